"""BASELINE config 5: batch sweep 1 ... 2^20 frames of the C2 attention layer (S=128, E=64, P=192, H=1, reference-pinned
arithmetic) on N GPUs (one process per GPU, batch-sharded, no collective), next to the host-CPU oracle port at the same batch.

    python tools/sweep.py [--config c2] [--max-log2 20] [--out profiles/sweep.json]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/sweep.py ...

Per batch B (frames per GPU; weak scaling: every rank runs B frames):
  * latency_us    one launch on resident data, CUDA events, median of the repetitions (B=1 is what the deployed batch-1 loop
                  samples/inference_udp_FPGA_custom_dispatch/main.cpp:150-231 would see)
  * device_fps    N * B / (max-over-ranks time of `reps` back-to-back launches over rotating buffers)
  * e2e_fps       the same through ITAEngine.forward_host (pinned host in, H2D, kernel, D2H, pinned host out), B <= 2^17
  * cpu_fps       oracle port (oracle/ita_ref.c, all host threads) on min(B, 2048) frames, rank 0 only
Writes one JSON document (rank 0) and prints a table.
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2")
    ap.add_argument("--max-log2", type=int, default=20)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    import torch.distributed as dist
    import ita_b200 as ita
    import ita_ref as R
    import bench

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = bench.CONFIGS[args.config]
    S, E = cfg["S"], cfg["E"]
    p, pdesc = bench.make_params(args.config, 2)
    eng = ita.ITAEngine(p, dev)
    assert eng.fused_supported, eng.fused_unsupported_reason()
    po = bench.to_oracle(R, p)
    ops = R.ops_per_frame(cfg["S"], cfg["E"], cfg["P"], cfg["H"])
    fb = S * E

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rows = []
    rc = None
    if rank == 0 and not args.no_cpu:
        import ita_ref_c as RCm
        rc = RCm.load(native=True)
    for lg in range(0, args.max_log2 + 1):
        B = 1 << lg
        # rotate over enough buffers to exceed the 126 MB L2 (at most 1 GiB per direction)
        nbuf = max(2, min(64, (160 << 20) // (B * fb) + 1))
        while nbuf > 2 and nbuf * B * fb > (1 << 30):
            nbuf -= 1
        base = torch.from_numpy(R.make_input(min(B, 4096), S, E, seed=7 + rank))
        reps_t = (B + base.shape[0] - 1) // base.shape[0]
        xs = [base.repeat(reps_t, 1, 1)[:B].contiguous().to(dev).roll(i, 0) for i in range(nbuf)]
        ys = [torch.empty_like(x) for x in xs]
        for i in range(3):
            eng.forward(xs[i % nbuf], out=ys[i % nbuf], path=ita.PATH_FUSED)
        eng.sync_check()
        # latency: single launches, one event pair each
        lat = []
        for i in range(15):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            eng.forward(xs[i % nbuf], out=ys[i % nbuf], path=ita.PATH_FUSED)
            e1.record()
            e1.synchronize()
            lat.append(e0.elapsed_time(e1) * 1e3)
        # throughput: back-to-back launches, max over ranks
        reps = max(4, min(200, int(2e7 // max(B, 1)) or 4))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            eng.forward(xs[i % nbuf], out=ys[i % nbuf], path=ita.PATH_FUSED)
        e1.record()
        torch.cuda.synchronize()
        ms = ita.max_over_ranks(e0.elapsed_time(e1), dev)
        barrier()
        eng.sync_check()
        row = {"batch_per_gpu": B, "latency_us": statistics.median(lat), "device_fps": world * B * reps / (ms * 1e-3),
               "device_tops": world * B * reps * ops / (ms * 1e-3) / 1e12, "reps": reps, "buffers": nbuf}
        # parity spot check on this batch (rank 0, up to 8 frames)
        if rank == 0:
            n = min(B, 8)
            row["parity"] = bool(np.array_equal(R.mha_forward(po, xs[(reps - 1) % nbuf][:n].cpu().numpy())["Y"],
                                                ys[(reps - 1) % nbuf][:n].cpu().numpy()))
        # end to end
        if lg <= 17:
            xh = [x.cpu().pin_memory() for x in xs[:2]]
            yh = [torch.empty_like(x).pin_memory() for x in xh]
            for i in range(2):
                eng.forward_host(xh[i], out=yh[i], path=ita.PATH_FUSED)
            er = max(4, min(50, int(4e6 // max(B, 1)) or 4))
            barrier()
            t0 = time.perf_counter()
            for i in range(er):
                eng.forward_host(xh[i % 2], out=yh[i % 2], path=ita.PATH_FUSED)
            torch.cuda.synchronize()
            dt = ita.max_over_ranks((time.perf_counter() - t0) * 1e3, dev)
            barrier()
            row["e2e_fps"] = world * B * er / (dt * 1e-3)
            row["e2e_latency_us"] = dt * 1e3 / er
            del xh, yh
        if rc is not None:
            n = min(B, 2048)
            X = xs[0][:n].cpu().numpy()
            rc.mha(po, X[: min(n, 8)], intermediates=False, nthreads=bench.host_threads())
            t0 = time.perf_counter()
            k = 0
            while True:
                rc.mha(po, X, intermediates=False, nthreads=bench.host_threads())
                k += 1
                if time.perf_counter() - t0 > 0.4 or k >= 200:
                    break
            row["cpu_fps"] = n * k / (time.perf_counter() - t0)
            row["cpu_threads"] = bench.host_threads()
        rows.append(row)
        del xs, ys
        torch.cuda.empty_cache()
    if rank == 0:
        doc = {"config": args.config, "workload": cfg["workload"], "n_gpus": world, "rounding": "torch_rne_f32", "params": pdesc,
               "kernels": eng.describe(), "build_id": bench.build_id(), "rows": rows,
               "note": "device_fps / e2e_fps are whole-job figures (N GPUs x batch_per_gpu frames per launch); latency_us is one launch on rank 0"}
        if args.out:
            with open(args.out, "w") as f:
                json.dump(doc, f, indent=1)
        print(f"# sweep {args.config} on {world} GPU(s): {eng.describe().split(';')[0]}")
        print("batch/GPU | latency us | device M frames/s | int8 TOP/s | e2e M frames/s | e2e latency us | CPU port k frames/s | parity")
        for r in rows:
            print(f"{r['batch_per_gpu']:9d} | {r['latency_us']:10.1f} | {r['device_fps'] / 1e6:17.3f} | {r['device_tops']:10.1f} | "
                  f"{(r.get('e2e_fps') or 0) / 1e6:14.3f} | {r.get('e2e_latency_us', 0):14.1f} | {(r.get('cpu_fps') or 0) / 1e3:19.2f} | {r.get('parity')}")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
