"""Generate tests/golden/framequeue.npz from the UNMODIFIED reference FrameQueue
(dsp/frame/FrameQueue.cc + Indexing.h compiled into oracle/_ref by oracle/Makefile).
Run in the build container only:   python tests/golden/make_golden_framequeue.py
Own seed, so the other golden files (make_golden.py) are not regenerated."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import pyoracle as po  # noqa: E402

po.build_ref()
R = po.Ref(strict=True)
rng = np.random.default_rng(20261105)
d = {}
cases = []
for L in (0, 1, 3, 4, 40, 333):
    x = rng.uniform(-1, 1, L).astype(np.float32)
    d[f"x_{L}"] = x
    for N, H in ((8, 4), (32, 16), (32, 24), (16, 16), (31, 15), (256, 64)):
        for center in (0, 1):
            for mode in (0, 1, 2):
                d[f"f_{L}_{N}_{H}_{center}_{mode}"] = R.framequeue(x, N, H, bool(center), mode)
                cases.append((L, N, H, center, mode))
d["cases"] = np.array(cases, np.int64)
# the pattern the reference's own test prints (tests/frame_queue_test.cc:321-364)
d["pattern_1234_8_4_reflect"] = R.framequeue(np.array([1, 2, 3, 4], np.float32), 8, 4, True, 1)
p = os.path.join(HERE, "framequeue.npz")
np.savez_compressed(p, **d)
print(f"framequeue.npz: {os.path.getsize(p)/1024:.1f} KiB, {len(cases)} cases")
