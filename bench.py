#!/usr/bin/env python
"""bench.py -- ITA int8 multi-head-attention throughput on B200 (frames/s, int8 TOP/s), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A *step* is one pass of the hot path (one fused-kernel launch through ``ita_mha_forward``) over one batch of synthetic
depth-frame token tiles.  Workload = BASELINE.json configs[1]: the ITAViTLSTM attention layer at the exported dims
(S=128, E=64, P=192, H=1), batch 4096 frames per GPU, seeded synthetic int8 inputs/weights.

* ``value``        frames/s, whole job (all ranks), inputs resident in HBM, CUDA-event timed, max over ranks
* ``e2e``          same metric through the host-buffer C-ABI entry (``ita_mha_forward_host``): pinned host input ->
                   H2D -> kernel -> D2H, every step
* ``roofline``     int8 tensor-core roofline of the fused kernel (algorithmic ops / measured launch time)
* ``cpu_baseline`` the CPU oracle (oracle/ita_ref.c, OpenMP) timed on this box's host cores on a bounded sample
* ``--impl reference``  times the reference path on the host cores.  The reference's own implementation is Python /
  torch.ao quantized modules plus an absent golden model and cannot travel to the GPU box, so this arm runs the
  oracle port (kind "port"), with all host threads.

L2 hygiene: 8 input/output buffer sets (8 x 2 x 33.5 MB = 537 MB > 126 MB L2) are cycled, so no step finds its
input in L2.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(S=128, E=64, P=192, H=1)
WORKLOAD = "ITAViTLSTM attention layer at exported ONNX dims (S=128,E=64,P=192,H=1), int8, batch 4096 frames/GPU"
NBUF = 8


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.lines, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # samples under load = the upper half of the observed clocks (idle samples before/after the region are dropped)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        load = sorted(sm)[len(sm) // 2:]
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def host_threads() -> int:
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1, so the OpenMP default is not it)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_baseline(params, seconds: float = 10.0, native: bool = True):
    """Oracle port timed on this box's host cores; bounded sample (~`seconds` of CPU work)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ita_ref as R
    import ita_ref_c as RC
    lib = RC.load(native=native)
    chunk = 512
    X = R.make_input(chunk, params.S, params.E, seed=3)
    nt = host_threads()
    lib.mha(params, X[:32], intermediates=False, nthreads=nt)          # warm-up (thread pool, page faults)
    n, t0 = 0, time.perf_counter()
    while True:
        lib.mha(params, X, intermediates=False, nthreads=nt)
        n += chunk
        dt = time.perf_counter() - t0
        if dt >= seconds:
            break
    return {"value": n / dt, "unit": "frames/s", "cores": nt, "kind": "port",
            "sample": f"{n} frames of the same workload in {dt:.1f} s, oracle/ita_ref.c -O3 -march=native OpenMP over frames"}


def run_reference(args):
    """--impl reference: the oracle port on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ita_ref as R
    import ita_ref_c as RC
    lib = RC.load(native=True)
    p = R.make_synthetic(**CFG, seed=0)
    sample = 1024                                           # frames per step (bounded sample of the 4096-frame batch)
    X = R.make_input(sample, p.S, p.E, seed=3)
    nt = host_threads()
    for _ in range(max(args.warmup, 1)):
        lib.mha(p, X[:128], intermediates=False, nthreads=nt)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lib.mha(p, X, intermediates=False, nthreads=nt)
    dt = time.perf_counter() - t0
    fps = sample * args.steps / dt
    ops = R.ops_per_frame(**CFG)
    line = {"impl": "reference", "metric": "ITA int8 MHA frames/s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "s8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample_frames_per_step": sample},
            "int8_tops": fps * ops / 1e12,
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": nt, "kind": "port",
                             "sample": f"{sample} frames/step x {args.steps} steps, oracle/ita_ref.c OpenMP, all host threads"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)
    return 0


_JSON_FD = None


def emit(line: dict) -> None:
    """The one JSON line, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # stdout carries exactly one JSON line: everything else that libraries print there (NCCL's "NCCL version ..." banner
    # at communicator creation, for one) is sent to stderr by pointing fd 1 at fd 2 for the duration of the run.
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4096, help="frames per GPU per step")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    import ita_b200 as ita
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ita_ref as R     # synthetic workload generator + cpu_baseline only; never on the measured path

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the host-CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = max(args.warmup, 3)
    K = args.steps
    B = args.batch

    p = R.make_synthetic(**CFG, seed=0)
    params = ita.ITAQuantParams(S=p.S, E=p.E, P=p.P, H=p.H, Wq=p.Wq, Wk=p.Wk, Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv,
                                Bo=p.Bo, mult=p.mult, shift=p.shift, add=p.add, eff=p.eff, round_mode=p.round_mode,
                                softmax_mode=p.softmax_mode, head_mode=p.head_mode)
    eng = ita.ITAEngine(params, dev)
    if not eng.fused_supported:
        raise SystemExit("fused tcgen05 path unavailable: " + eng.fused_unsupported_reason())
    lib = ita._lib.load()

    # ---- synthetic inputs: NBUF distinct batches (per rank), resident in HBM
    xs_host = [torch.from_numpy(R.make_input(B, p.S, p.E, seed=100 * rank + i)).pin_memory() for i in range(NBUF)]
    xs = [x.to(dev) for x in xs_host]
    ys = [torch.empty_like(x) for x in xs]
    ys_host = [torch.empty_like(x).pin_memory() for x in xs_host[:2]]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (value) + per-launch kernel time (roofline)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    for i in range(W):                                   # the W untimed warm-up steps, right before the timed region
        eng.forward(xs[i % NBUF], out=ys[i % NBUF], path=ita.PATH_FUSED)
    eng.sync_check()
    barrier()
    n0 = lib.ita_kernel_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ev[0].record()
    for i in range(K):
        eng.forward(xs[i % NBUF], out=ys[i % NBUF], path=ita.PATH_FUSED)
        ev[i + 1].record()
    torch.cuda.synchronize()
    total_ms = ev[0].elapsed_time(ev[K])
    launches = lib.ita_kernel_launch_count() - n0
    barrier()
    eng.sync_check()
    kern_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(K)]
    total_ms_max = ita.max_over_ranks(total_ms, dev)

    # ---- end to end through the host-buffer entry point (pinned host in, pinned host out)
    # Two handles on two streams through the asynchronous entry (ita_mha_forward_host_async): every step still copies its
    # own input H2D and its own result D2H, but the H2D copies of step i+1 overlap the D2H copies of step i (PCIe is full
    # duplex), which is how a serving loop drives the op.  Results are valid (and checked) after sync_check.
    eng2 = ita.ITAEngine(params, dev)
    engs, streams = (eng, eng2), (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
    for i in range(2):
        engs[i].forward_host(xs_host[i], out=ys_host[i], stream=streams[i], sync=False)
    for i in range(2):
        engs[i].sync_check(streams[i])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    Ke = max(4, min(K, 20))
    cur = torch.cuda.current_stream(dev)
    e0.record()
    for st in streams:
        st.wait_stream(cur)
    for i in range(Ke):
        engs[i % 2].forward_host(xs_host[i % NBUF], out=ys_host[i % 2], stream=streams[i % 2], sync=False)
    for st in streams:
        cur.wait_stream(st)
    e1.record()
    torch.cuda.synchronize()
    for i in range(2):
        engs[i].sync_check(streams[i])
    e2e_ms = ita.max_over_ranks(e0.elapsed_time(e1), dev)
    barrier()
    clocks = sampler.stop() if sampler else None

    # ---- a parity spot-check on the bench data itself (outside the timed regions): CUDA result == oracle, 64 frames
    parity = None
    if rank == 0:
        ref = R.mha_forward(p, xs_host[0][:64].numpy())["Y"]
        got = ys[0][:64].cpu().numpy() if K >= 1 else None
        parity = bool(got is not None and np.array_equal(ref, got))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    ops = R.ops_per_frame(**CFG)
    fps = world * B * K / (total_ms_max * 1e-3)
    e2e_fps = world * B * Ke / (e2e_ms * 1e-3)
    kms = statistics.mean(kern_ms)
    achieved = B * ops / (kms * 1e-3) / 1e12
    peaks = _peaks()
    meas_i8 = None
    try:
        meas_i8 = ita.microbench(0, local)[0]
    except Exception:
        pass
    bf16_burst = peaks.get("bf16_tflops")
    peak = meas_i8 if meas_i8 else (2 * bf16_burst if bf16_burst else 2 * 1590.0)
    peak_src = ("measured here: ita_microbench kind::i8 M128 N256 cta_group::1 issue loop" if meas_i8 else
                ("2 x MEASURED_PEAKS.json bf16_tflops (no int8 figure recorded)" if bf16_burst else "2 x fallback 1.59 PFLOP/s"))
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:
        pass
    line = {
        "metric": "ITA int8 MHA frames/s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "s8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_gpu_per_step": B, "parallelism": f"batch-shard x{world}, no collective",
                   "l2": f"{NBUF} rotating input/output buffer sets ({NBUF * 2 * B * p.S * p.E / 1e6:.0f} MB) > 126 MB L2",
                   "rounding": "ita_half_up", "softmax": "repo_global"},
        "int8_tops": fps * ops / 1e12,
        "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": B * p.S * p.E, "d2h_bytes_per_step": B * p.S * p.E,
                "steps": Ke, "api": "ITAEngine.forward_host(sync=False) -> ita_mha_forward_host_async, pinned host buffers, two handles on two streams"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TOP/s", "frac": achieved / peak,
                     "traffic": traffic, "kernel": "k_ita_fused<64,192>", "kernel_ms": kms, "ops_per_launch": B * ops,
                     "peak_source": peak_src,
                     "frac_of_2x_measured_bf16": (achieved / (2 * bf16_burst)) if bf16_burst else None,
                     "frac_of_nominal_4500": achieved / 4500.0,
                     "hbm_gbs": B * 2 * p.S * p.E / (kms * 1e-3) / 1e9,
                     "hbm_frac_of_measured": (B * 2 * p.S * p.E / (kms * 1e-3) / 1e9 / peaks["hbm_gbs"]) if peaks.get("hbm_gbs") else None},
        "clocks": clocks,
        "parity_spot_check": parity,
    }
    if world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_baseline(p, args.cpu_seconds)
        except Exception as e:   # the baseline is a reported figure, never the product path
            line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e}"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
