#!/usr/bin/env python
"""bench.py -- STFT -> iSTFT round-trip throughput of the hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4] [--impl reference]

One "step" = one pass of the three kernels (K1 frame+window+rFFT, K2 irFFT+synthesis window,
K3 gather overlap-add) over one batch of synthetic input already resident in HBM.  Workloads
(BASELINE.json configs):
  c2 (default)  per rank: 1 channel x 2^30 samples, frame 2048 hop 512 Hann      [configs[1]]
                N>1: N independent channels, one per rank, no communication (weak scaling)
  c3            256 channels x 2^24 samples, frame 1024 hop 256, channels sharded over the
                ranks (strong scaling; one rank processes its share in batches of 32 channels)
  c4            one signal of N x 2^30 samples, frame 4096 hop 1024, time-sharded with a
                (frame-hop)-sample halo exchanged between neighbour ranks over NCCL (weak)
  c5            4096 concurrent streams, frame 256 hop 64 Hamming, fed in 64-sample chunks through
                the device-resident rings (crlot_stream_*); a step = 512 ticks; streams sharded (strong)
Rank 0 prints ONE JSON line.  `--impl reference` times the reference's CPU path instead
(oracle/_ref built from the reference sources when present, else the plain-C oracle port).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c2": dict(nfft=2048, hop=512, channels=1, length=1 << 30,
               name="c2: 1ch x 2^30 fp32 samples per GPU, frame 2048 hop 512 Hann (BASELINE configs[1])"),
    "c3": dict(nfft=1024, hop=256, channels=256, length=1 << 24,
               name="c3: 256ch x 2^24 samples, frame 1024 hop 256 Hann, channel-sharded (BASELINE configs[2])"),
    "c4": dict(nfft=4096, hop=1024, channels=1, length=1 << 30,
               name="c4: single signal of n_gpus x 2^30 samples, frame 4096 hop 1024 Hann, time-sharded, NCCL halo (BASELINE configs[3])"),
    "c5": dict(nfft=256, hop=64, channels=4096, length=64 * 512, chunk=64, window="hamming",
               name="c5: 4096 concurrent streams, frame 256 hop 64 Hamming, 64-sample chunks, device ring-buffer ingest, "
                    "streams sharded over the ranks (BASELINE configs[4])"),
}
SEED = 0xC0FFEE

# stdout carries exactly ONE JSON line.  Libraries write to file descriptor 1 behind Python's back (NCCL prints
# its "NCCL version ..." banner there when NCCL_DEBUG is set on the box), so fd 1 is pointed at stderr for the
# whole run and the line goes to the saved descriptor.
_JSON_FD = None


def claim_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


# ------------------------------------------------------------------------------------------------
# CPU arms (the only places bench.py touches oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_roundtrip_throughput(nfft, hop, per_thread_len, threads, reps=1):
    """All-thread reference-CPU throughput in samples/s on a bounded sample.  Each thread runs
    `reps` independent mono round trips of per_thread_len samples (the reference is
    single-threaded; independent channels are the only parallelism it admits)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    orc = po.Oracle()
    if po.Ref.available():
        ref = po.Ref()
        w = orc.window(po.HANN, nfft)
        xs = [orc.synth_uniform(SEED, t, 0, per_thread_len) for t in range(threads)]
        bar = threading.Barrier(threads + 1)
        t_end = [0.0] * threads

        def work(i):
            bar.wait()
            for _ in range(reps):
                ref.roundtrip(xs[i], nfft, hop, w, po.DROP, taps=False)  # ctypes drops the GIL
            t_end[i] = time.perf_counter()
        th = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
        for t in th:
            t.start()
        bar.wait()
        t0 = time.perf_counter()
        for t in th:
            t.join()
        dt = max(t_end) - t0
        kind = "reference"
    else:
        dt = sum(orc.bench_roundtrip_mt(threads, per_thread_len, nfft, hop, po.HANN, SEED)[0] for _ in range(reps))
        kind = "port"
    return threads * per_thread_len * reps / dt, kind, dt


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per = 1 << 22
    reps = 8                                   # ~1-2 s of work per step on 16 threads
    vals = []
    kind = "port"
    for i in range(args.warmup + args.steps):
        v, kind, dt = cpu_roundtrip_throughput(wl["nfft"], wl["hop"], per, threads, reps)
        if i >= args.warmup:
            vals.append((v, dt))
    v = float(np.mean([a for a, _ in vals]))
    sample = f"{threads} threads x {reps} x 2^22 samples (independent mono round trips) per step, frame {wl['nfft']} hop {wl['hop']}"
    line = {
        "impl": "reference", "metric": "stft_istft_roundtrip_throughput", "value": v / 1e6, "unit": "Msamples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": float(np.mean([b for _, b in vals])) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "note": "reference CPU path on host cores; Highway kernels replaced by the "
                   "reference's scalar kernels auto-vectorised by GCC (Highway not vendored)"},
        "cpu_baseline": {"value": v / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class Clocks:
    """SM clock / power / throttle reasons sampled every ~4 ms DURING the timed region through NVML
    (the nvidia-smi CLI takes longer to start than a 10-step region lasts); nvidia-smi is the fallback."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}

    def __init__(self, index):
        self.index, self.p, self.th, self.stop_flag = index, None, None, False
        self.sm, self.pw, self.reasons, self.mx = [], [], set(), None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _poll(self):
        nv = self.nvml
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.pw.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for name, bit in self.BITS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.004)

    def start(self):
        if self.nvml:
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                       "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if self.nvml:
            self.stop_flag = True
            self.th.join(timeout=2)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                    "power_w_max": max(self.pw) if self.pw else None, "samples": len(self.sm),
                    "reasons": sorted(self.reasons), "source": "nvml, sampled every ~4 ms inside the timed region"}
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except Exception:
            out = ""
        sm, mx, reasons, pw = [], [], set(), []
        for ln in out.splitlines():
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvidia-smi -lms 20"}


# ------------------------------------------------------------------------------------------------
# GPU arm, streaming workload (c5)
# ------------------------------------------------------------------------------------------------
def run_c5(args, wl, rank, world, local, torch, dist, dsp):
    dev = torch.device("cuda", local)
    N, H, chunk = wl["nfft"], wl["hop"], wl["chunk"]
    assert wl["channels"] % world == 0
    S = wl["channels"] // world
    T = wl["length"] // chunk                                    # ticks per step
    w = dsp.window_make(dsp.HAMMING, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP, device=local)
    st = dsp.Stream(plan, S, max_chunk=chunk)
    x = dsp.synth_uniform(S, T * chunk, seed=SEED, channel0=rank * S, device=local)
    out = torch.empty((S, (T + 1) * chunk), dtype=torch.float32, device=dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        st.run(x, chunk, n_ticks=T, out=out)
    sync_all()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    t_beg, t_end = ev(), ev()
    sync_all()
    t_beg.record()
    for _ in range(args.steps):
        st.run(x, chunk, n_ticks=T, out=out)
    t_end.record()
    sync_all()
    clk = clocks.stop() if rank == 0 else None
    tmax = torch.tensor([t_beg.elapsed_time(t_end)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    total = S * world * T * chunk
    value = total * args.steps / (ms * 1e-3)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    bps = dsp.bytes_per_sample(N, H)
    gbps = bps * (value / world) / 1e9
    roofline = {"bound": "hbm", "kernel": "one tick = stream_tick1_kernel (ingest + window/rFFT + irFFT + OLA ring + produce in ONE launch; 32 ticks per CUDA graph)",
                "achieved": gbps, "peak": peak, "unit": "GB/s", "frac": gbps / peak, "traffic": None, "peak_source": peak_src,
                "note": "latency/launch-bound by construction: a tick moves 19 MB per GPU and its working set is "
                        "L2-resident; the figure to read is tick_latency_us",
                "path": {"bytes_per_sample": bps, "achieved_GBps_per_gpu": gbps, "frac_of_measured": gbps / peak}}

    e2e = None
    if not args.no_e2e:
        ticks = min(T, 256)
        hin = torch.empty((S, ticks * chunk), dtype=torch.float32).pin_memory()
        hin.copy_(x[:, :ticks * chunk])
        hout = torch.empty((S, chunk), dtype=torch.float32).pin_memory()
        hin_np, hout_np = hin.numpy(), hout.numpy()
        for t in range(8):
            st.tick_host(hin_np[:, t * chunk:(t + 1) * chunk], hout_np)
        sync_all()
        lat = []
        t0 = time.perf_counter()
        for t in range(ticks):
            a = time.perf_counter()
            st.tick_host(hin_np[:, t * chunk:(t + 1) * chunk], hout_np)      # returns when h_out is complete
            lat.append(time.perf_counter() - a)
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": S * world * ticks * chunk / float(dt.item()) / 1e6, "unit": "Msamples/s",
               "h2d_bytes_per_step": S * chunk * 4, "d2h_bytes_per_step": S * H * (chunk // H) * 4, "steps": ticks,
               "api": "crlot_stream_tick_host: per tick H2D of the chunk, one fused tick kernel, D2H of the hop, synchronous",
               "tick_latency_us": {"median": float(np.median(lat)) * 1e6, "p99": float(np.percentile(lat, 99)) * 1e6}}

    cpu = None
    if rank == 0 and not args.no_cpu and world == 1:
        threads = os.cpu_count() or 1
        per, reps = 1 << 22, 8
        v, kind, dt = cpu_roundtrip_throughput(N, H, per, threads, reps)
        cpu = {"value": v / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind,
               "sample": f"{threads} threads x {reps} x 2^22 samples, independent mono round trips, frame {N} hop {H} Hann "
                         f"({dt:.1f} s); Highway kernels replaced by the reference's scalar kernels (Highway not vendored)"}
    if rank == 0:
        line = {
            "metric": "stft_istft_roundtrip_throughput", "value": value / 1e6, "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["name"], "frame": N, "hop": H, "streams_per_rank": S, "chunk": chunk,
                       "ticks_per_step": T, "tick_latency_us": ms / args.steps / T * 1e3,
                       "cache": "working set (rings 29 MB + chunk) is L2-resident by nature of the workload; inputs of a "
                                "step (512 MB) stream from HBM",
                       "generator": "counter hash uniform[-1,1), seed 0xC0FFEE"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": (1 if chunk == H else 5) * T * args.steps, "clocks": clk,
        }
        emit(line)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def bind_to_gpu_numa_node(torch, local):
    """One process per GPU: run this process on the CPUs of the NUMA node the GPU hangs off, so that
    the page-locked host buffers of the e2e leg (first touch) and the threads that feed the copy
    engines sit next to the GPU's PCIe root.  Returns a description for the JSON line, or None when
    the topology cannot be read (then nothing is changed)."""
    try:
        p = torch.cuda.get_device_properties(local)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"gpu": bdf, "numa_node": node, "cpus": len(cpus)}
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--length-log2", type=int, default=None, help="override samples per channel (debug)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    claim_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    wl = dict(WORKLOADS[args.workload])
    if args.length_log2:
        wl["length"] = 1 << args.length_log2
        wl["name"] += f" [length overridden to 2^{args.length_log2}]"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, wl, rank, world)
        return

    import torch
    import torch.distributed as dist
    import crlot_dsp_b200 as dsp
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (the product has no CPU path)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(torch, local)      # pinned host buffers land on the GPU's own NUMA node
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"

    if args.workload == "c5":
        run_c5(args, wl, rank, world, local, torch, dist, dsp)
        if world > 1:
            dist.destroy_process_group()
        return

    N, H = wl["nfft"], wl["hop"]
    w = dsp.window_make(dsp.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP, device=local)
    halo = N - H

    # ---- lay the workload out on this rank ------------------------------------------------------
    if args.workload == "c2":
        chans, L, batch = 1, wl["length"], 1
        x = dsp.synth_uniform(1, L, seed=SEED, channel0=rank, device=local)
        local_samples = L
        scaling = "weak"
    elif args.workload == "c3":
        assert wl["channels"] % world == 0
        chans, L = wl["channels"] // world, wl["length"]
        batch = min(32, chans)                      # workspace batch: 32 ch = 21.5 GB of spectra+frames
        x = dsp.synth_uniform(chans, L, seed=SEED, channel0=rank * chans, device=local)
        local_samples = chans * L
        scaling = "strong"
    else:  # c4: rank r owns the output/time range [r*S, (r+1)*S); halos from BOTH neighbours (sharding.py)
        from crlot_dsp_b200 import sharding
        chans, batch = 1, 1
        S = wl["length"]
        sh = sharding.make_time_shard(S * world, N, H, rank, world)
        left, right = sh.left_halo, sh.right_halo
        xbuf = torch.zeros(left + S + right, dtype=torch.float32, device=dev)
        dsp.synth_uniform(1, S, seed=SEED, channel0=0, i0=rank * S, device=local, out=xbuf[left:left + S].unsqueeze(0))
        local_samples = S
        scaling = "weak"

    def halo_exchange():
        """NCCL send/recv of the (frame-hop)-sample halos between neighbouring ranks."""
        sharding.halo_exchange(xbuf, sh)

    # workspaces
    if args.workload == "c4":
        Ltot = left + S + right
        F, n_out = sh.n_frames, sh.n_out
        spec = torch.empty((1, F, plan.bins), dtype=torch.complex64, device=dev)
        frames = torch.empty((1, F, N), dtype=torch.float32, device=dev)
        out = torch.empty((1, n_out), dtype=torch.float32, device=dev)
    else:
        F = plan.n_frames(L)
        spec = torch.empty((batch, F, plan.bins), dtype=torch.complex64, device=dev)
        frames = torch.empty((batch, F, N), dtype=torch.float32, device=dev)
        out = torch.empty((chans, plan.out_len(L)), dtype=torch.float32, device=dev)

    ev = lambda: torch.cuda.Event(enable_timing=True)
    kt = {"k1": [], "k2": [], "k3": []}

    def step(record=False):
        launches = 0
        if args.workload == "c4":
            if world > 1:
                halo_exchange()
            e = [ev() for _ in range(4)] if record else None
            if record: e[0].record()
            plan.stft(xbuf.unsqueeze(0), length=Ltot, n_frames=F, out=spec)
            if record: e[1].record()
            plan.istft_frames(spec, fuse_synthesis_window=True, out=frames)
            if record: e[2].record()
            plan.ola(frames, prewindowed=True, t_offset=sh.t0, frame_t0=sh.frame_t0, n_out=n_out, out=out)
            if record: e[3].record()
            if record:
                kt["_ev"].append(e)
            return 3
        for c0 in range(0, chans, batch):
            xb = x[c0:c0 + batch]
            e = [ev() for _ in range(4)] if record else None
            if record: e[0].record()
            plan.stft(xb, out=spec)
            if record: e[1].record()
            plan.istft_frames(spec, fuse_synthesis_window=True, out=frames)
            if record: e[2].record()
            plan.ola(frames, prewindowed=True, out=out[c0:c0 + batch])
            if record: e[3].record()
            if record:
                kt["_ev"].append(e)
            launches += 3
        return launches

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    sync_all()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    kt["_ev"] = []
    t_beg, t_end = ev(), ev()
    sync_all()
    t_beg.record()
    launches = 0
    for _ in range(args.steps):
        launches += step(record=True)
    t_end.record()
    sync_all()
    clk = clocks.stop() if rank == 0 else None
    ms = t_beg.elapsed_time(t_end)
    for e in kt["_ev"]:
        kt["k1"].append(e[0].elapsed_time(e[1])); kt["k2"].append(e[1].elapsed_time(e[2])); kt["k3"].append(e[2].elapsed_time(e[3]))
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    total_samples = local_samples * world
    value = total_samples * args.steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel (rank 0's launches) --------------------------------------
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
    bins = N // 2 + 1
    frames_per_launch = F * (batch if args.workload != "c4" else 1)
    samples_per_launch = frames_per_launch * H             # new samples consumed per launch
    alg = {"k1": 4.0 * samples_per_launch + 8.0 * bins * frames_per_launch,
           "k2": (8.0 * bins + 4.0 * N) * frames_per_launch,
           "k3": 4.0 * N * frames_per_launch + 4.0 * samples_per_launch}
    kavg = {k: float(np.mean(kt[k])) for k in ("k1", "k2", "k3")}
    dom = max(kavg, key=kavg.get)
    names = {"k1": "stft_fwd_kernel (K1 frame+window+rFFT)", "k2": "istft_inv_kernel (K2 irFFT+synthesis window)",
             "k3": "ola_gather_kernel (K3 gather overlap-add)"}
    achieved = alg[dom] / (kavg[dom] * 1e-3) / 1e9
    per_kernel = {k: {"ms": kavg[k], "alg_GB": alg[k] / 1e9, "GBps": alg[k] / (kavg[k] * 1e-3) / 1e9,
                      "frac": alg[k] / (kavg[k] * 1e-3) / 1e9 / peak} for k in kavg}
    path_gbps = dsp.bytes_per_sample(N, H) * (value / world) / 1e9
    roofline = {"bound": "hbm", "kernel": names[dom], "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": peak_src, "per_kernel": per_kernel,
                "path": {"bytes_per_sample": dsp.bytes_per_sample(N, H), "achieved_GBps_per_gpu": path_gbps,
                         "frac_of_measured": path_gbps / peak, "frac_of_8TBps_nominal": path_gbps / 8000.0}}

    # measured DRAM traffic of the dominant kernel, from the committed ncu --set full capture
    # (profiles/traffic.json: bytes per frame, read+write, per kernel and frame size)
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_path):
        tj = json.load(open(traffic_path))
        ent = tj.get("kernels", {}).get(f"{dom}_n{N}_h{H}")
        if ent:
            roofline["traffic"] = ent["dram_bytes_per_frame"] * frames_per_launch
            roofline["traffic_source"] = f"profiles/traffic.json ({ent['source']}), bytes/frame x frames per launch"

    # ---- the same round trip with kernels 2+3 fused (iFFT frames never materialised) --------------
    fused = None
    c4_fused_ok = args.workload == "c4" and plan.istft_ola_supported(t_offset=sh.t0, frame_t0=sh.frame_t0, out=out)
    if (args.workload != "c4" and plan.istft_ola_supported()) or c4_fused_ok:
        def fstep(record=None):
            if args.workload == "c4":
                if world > 1:
                    halo_exchange()
                e = [ev() for _ in range(3)]
                e[0].record()
                plan.stft(xbuf.unsqueeze(0), length=Ltot, n_frames=F, out=spec)
                e[1].record()
                plan.istft_ola(spec, t_offset=sh.t0, frame_t0=sh.frame_t0, n_out=n_out, out=out)
                e[2].record()
                if record is not None:
                    record.append(e)
                return
            for c0 in range(0, chans, batch):
                xb = x[c0:c0 + batch]
                if record is not None:
                    e = [ev() for _ in range(3)]
                    e[0].record()
                plan.stft(xb, out=spec)
                if record is not None:
                    e[1].record()
                plan.istft_ola(spec, out=out[c0:c0 + batch])
                if record is not None:
                    e[2].record()
                    record.append(e)
        for _ in range(args.warmup):
            fstep()
        sync_all()
        rec = []
        f_beg, f_end = ev(), ev()
        f_beg.record()
        for _ in range(args.steps):
            fstep(rec)
        f_end.record()
        sync_all()
        fms = torch.tensor([f_beg.elapsed_time(f_end)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(fms, op=dist.ReduceOp.MAX)
        fms = float(fms.item())
        fk1 = float(np.mean([e[0].elapsed_time(e[1]) for e in rec]))
        fk23 = float(np.mean([e[1].elapsed_time(e[2]) for e in rec]))
        alg23 = (8.0 * bins + 4.0 * H) * frames_per_launch
        fbps = 8.0 + 16.0 * bins / H
        fval = total_samples * args.steps / (fms * 1e-3)
        fused = {"value": fval / 1e6, "unit": "Msamples/s", "ms_per_step": fms / args.steps,
                 "kernels": "stft_fwd_kernel (K1) + istft_ola_kernel (K2+K3 fused, bit-identical output)",
                 "k1_ms": fk1, "k23_ms": fk23, "k23_alg_GB": alg23 / 1e9, "k23_GBps": alg23 / (fk23 * 1e-3) / 1e9,
                 "k23_frac": alg23 / (fk23 * 1e-3) / 1e9 / peak, "bytes_per_sample": fbps,
                 "path_GBps_per_gpu": fbps * (fval / world) / 1e9, "path_frac_of_measured": fbps * (fval / world) / 1e9 / peak,
                 "note": "the fused path moves 8+16*bins/hop bytes per sample instead of 8+16*bins/hop+8*N/hop; `value` "
                         "above stays on the three-kernel path that BASELINE.json's north_star names"}

    # ---- e2e: host buffers through the C-ABI host entry point, copies inside the timed region -----
    e2e = None
    if not args.no_e2e:
        Le = wl["length"] if args.workload != "c3" else wl["length"]
        ce = 1 if args.workload != "c3" else min(chans, 4)
        nbytes = ce * Le * 4
        hin = torch.empty((ce, Le), dtype=torch.float32).pin_memory()
        hin.copy_(x[:ce, :Le] if args.workload != "c4" else xbuf[left:left + Le].unsqueeze(0))
        n_out_e = plan.out_len(Le)
        hout = torch.empty((ce, n_out_e), dtype=torch.float32).pin_memory()
        hin_np, hout_np = hin.numpy(), hout.numpy()
        plan.roundtrip_host(hin_np, out=hout_np)                 # warm-up (allocates the slab workspace)
        sync_all()
        n_e2e = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            plan.roundtrip_host(hin_np, out=hout_np)             # synchronous: returns when h_out is complete
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": ce * Le * world * n_e2e / float(dt.item()) / 1e6, "unit": "Msamples/s",
               "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": ce * n_out_e * 4, "steps": n_e2e,
               "api": "crlot_roundtrip_host (pinned host buffers, slabbed H2D/compute/D2H pipeline; K1 + fused K2+K3 per slab)",
               "numa_binding": numa,
               "workload": f"{ce}ch x 2^{int(np.log2(Le))} samples per rank",
               "ceiling": "PCIe: 45.8 GB/s per direction with H2D and D2H running together (tools/pcie_probe.py) = 11.4 Gsamples/s per GPU"}
        del hin, hout

    # ---- reconstruction quality of the round trip just measured (untimed, north_star: "the round-trip
    # reconstruction SNR must be reported").  vs the input: what the reference's own OLA normalisation
    # gives (sum-of-window envelope, not sum of squares: ~7.3 dB for Hann at 75 % overlap);
    # vs the reference: the CPU reference/oracle on the first 2^18 samples of the same input (checker only).
    step()
    torch.cuda.synchronize()
    if args.workload == "c4":       # out[0, j] is output sample t0 + j = rank*S + j, xbuf[left + j] the same input sample
        hi = min(S, n_out) - N
        xin, yout = xbuf[left + N:left + hi], out[0, N:hi]
    else:
        c_last = ((chans - 1) // batch) * batch
        xin, yout = x[c_last, N:L - N], out[c_last, N:L - N]
    num = den = 0.0
    for o in range(0, xin.numel(), 1 << 26):
        a_, b_ = xin[o:o + (1 << 26)].double(), yout[o:o + (1 << 26)].double()
        num += float((a_ * a_).sum().item()); den += float(((a_ - b_) ** 2).sum().item())
    recon = {"snr_db_vs_input": 10.0 * math.log10(num / den) if den > 0 else float("inf"),
             "samples": int(xin.numel()),
             "note": "the reference normalises by the window SUM envelope (norm_builder.cc), so its own round trip is "
                     "a scaled copy of the input (~7.3 dB for Hann at 75 % overlap); parity is judged against the reference"}
    if rank == 0 and not args.no_cpu and world == 1:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pyoracle as po
        seg = (1 << 18) + N
        xs = (xbuf[left:left + seg] if args.workload == "c4" else x[c_last, :seg]).cpu().numpy()
        orc = po.Ref() if po.Ref.available() else po.Oracle()
        want = orc.roundtrip(xs, N, H, po.Oracle().window(po.HANN, N), po.DROP, taps=False)["out"][N:1 << 18].astype(np.float64)
        got = (out[0] if args.workload == "c4" else out[c_last])[N:1 << 18].cpu().numpy().astype(np.float64)
        err = float(np.sum((got - want) ** 2))
        recon["vs_reference"] = {"rel_l2": math.sqrt(err / float(np.sum(want ** 2))),
                                 "snr_db": 10.0 * math.log10(float(np.sum(want ** 2)) / err) if err > 0 else float("inf"),
                                 "samples": int(want.size), "checker": "reference" if po.Ref.available() else "port",
                                 "bar": "1e-5 relative L2 (north_star)"}

    cpu = None
    if rank == 0 and not args.no_cpu and world == 1:
        threads = os.cpu_count() or 1
        per, reps = 1 << 24, 24
        v, kind, dt = cpu_roundtrip_throughput(N, H, per, threads, reps)
        cpu = {"value": v / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind,
               "sample": f"{threads} threads x {reps} x 2^24 samples, independent mono round trips, frame {N} hop {H} ({dt:.1f} s); "
                         "Highway kernels replaced by the reference's scalar kernels (Highway not vendored)"}

    if rank == 0:
        line = {
            "metric": "stft_istft_roundtrip_throughput", "value": value / 1e6, "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["name"], "frame": N, "hop": H, "channels_per_rank": chans,
                       "samples_per_rank": local_samples, "frames_per_launch": frames_per_launch,
                       "cache": "inputs larger than L2 (every buffer >> 126 MB), no flush needed",
                       "generator": "counter hash uniform[-1,1), seed 0xC0FFEE"},
            "roofline": roofline, "fused_k23": fused, "reconstruction": recon, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clk,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
