#!/usr/bin/env python
"""profiles/summarize.py -- turn the scratch artefacts a gpurun call brought back (gpurun_out/)
into the small tracked files under profiles/.

    python profiles/summarize.py ncu   gpurun_out/prof.ncu-rep  profiles/r1_ncu_c2.csv
    python profiles/summarize.py sass  profiles/r1_sass         stft_fwd_kernelILi2048ELi2E ...

`ncu`  : one row per profiled launch with the counters DESIGN.md / bench.py quote.
`sass` : per kernel, the instruction-only SASS listing (encodings stripped) plus an opcode
         histogram; proves UBLKCP (TMA 1-D bulk copy), FFMA2/FMUL2/FADD2 (packed fp32), SHFL.
"""
import collections
import csv
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "crlot-dsp_b200", "libcrlot_cuda.so")

METRICS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "lts__t_sector_hit_rate.pct",
    "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second",
    # round 2: which pipe bounds the kernel, tensor-memory traffic, top stall reasons (warps per issue-active cycle)
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__sass_inst_executed_op_tmem_ldt.sum", "smsp__sass_inst_executed_op_tmem_stt.sum",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
]


def do_ncu(rep, out):
    if rep.endswith(".csv"):       # already exported on the GPU box (`ncu -i x.ncu-rep --page raw --csv`)
        raw = open(rep).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    head, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(head)}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{m} [{units[idx[m]]}]" for m in METRICS if m in idx])
        for r in data:
            name = re.sub(r"void |unnamed>::|\(.*", "", r[idx["Kernel Name"]])
            w.writerow([name] + [r[idx[m]] for m in METRICS if m in idx])
    print("wrote", out)


def do_sass(outdir, patterns):
    os.makedirs(outdir, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    blocks = re.split(r"\n\s*Function : ", txt)
    for b in blocks[1:]:
        name = b.split("\n", 1)[0].strip()
        if not any(p in name for p in patterns):
            continue
        ins, hist = [], collections.Counter()
        for ln in b.splitlines():
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*", ln)
            if m:
                ins.append(f"/*{m.group(1)}*/ {m.group(2)};")
                op = m.group(2).split()
                op = op[1] if op[0].startswith("@") else op[0]
                hist[op.split(".")[0]] += 1
        short = re.sub(r"_ZN5crlot\d+_GLOBAL__N__[0-9a-f]+_\d+_\w+?_cu_[0-9a-f]+\d\d", "", name)
        short = re.sub(r"[^A-Za-z0-9_]", "_", short)[:80]
        with open(os.path.join(outdir, short + ".sass"), "w") as f:
            f.write(f"// {name}\n// {len(ins)} instructions; opcode histogram:\n")
            for op, n in hist.most_common():
                f.write(f"//   {op:12s} {n}\n")
            f.write("\n".join(ins) + "\n")
        print("wrote", short, len(ins), "instructions")


if __name__ == "__main__":
    if sys.argv[1] == "ncu":
        do_ncu(sys.argv[2], sys.argv[3])
    elif sys.argv[1] == "sass":
        do_sass(sys.argv[2], sys.argv[3:])
    else:
        sys.exit(__doc__)
