"""Per-rank pinned-memory PCIe bandwidth (H2D, D2H, both at once) with 1..N ranks copying at the same time -- the measurement
behind the end-to-end scaling figure: is the N-GPU e2e rate a platform ceiling or a placement problem?

    python tools/pcie_probe.py                                  # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py [--bind]

Under torchrun every rank drives its own GPU; the ranks start each phase together (barrier) and report GB/s per rank; rank 0
prints the table and, with --out, writes JSON.  --bind pins the process to the CPUs NVML reports as local to its GPU before the
pinned buffers are allocated (first-touch NUMA placement)."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import bench

ap = argparse.ArgumentParser()
ap.add_argument("--bind", action="store_true")
ap.add_argument("--out", default="")
ap.add_argument("--mb", type=int, default=64)
args = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
aff = bench.bind_to_gpu_cpus(local) if args.bind else "not bound (--bind to pin to the GPU's local CPUs)"
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = args.mb << 20
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
h1.fill_(1); h2.fill_(2)
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, active, reps=12):
    """GB/s of `fn` on this rank while the ranks in `active` run it together (others idle)."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if rank not in active:
        if world > 1:
            dist.barrier()
        return 0.0
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    if world > 1:
        dist.barrier()
    return n / dt / 1e9


def both():
    with torch.cuda.stream(s1):
        d1.copy_(h1, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


rows = []
counts = sorted({c for c in (1, 2, 4, 8, world) if c <= world})
for c in counts:
    active = set(range(c))
    r = {"ranks_active": c,
         "h2d": timed(lambda: d1.copy_(h1, non_blocking=True), active),
         "d2h": timed(lambda: h2.copy_(d2, non_blocking=True), active),
         "both": timed(both, active)}
    if world > 1:
        t = torch.tensor([r["h2d"], r["d2h"], r["both"]], device="cuda")
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        per = [x.tolist() for x in g][:c]
    else:
        per = [[r["h2d"], r["d2h"], r["both"]]]
    rows.append({"ranks_active": c, "per_rank_h2d": [p[0] for p in per], "per_rank_d2h": [p[1] for p in per],
                 "per_rank_both_each_way": [p[2] for p in per]})
affs = [aff]
if world > 1:
    affs = [None] * world
    dist.all_gather_object(affs, aff)
if rank == 0:
    print(f"# pinned-memory PCIe bandwidth, {args.mb} MiB copies, {world} rank(s); affinity: {affs[0]}")
    print("ranks active | H2D GB/s per rank (sum) | D2H GB/s per rank (sum) | both directions, each way per rank (sum)")
    for r in rows:
        f = lambda v: f"{min(v):.1f}-{max(v):.1f} ({sum(v):.0f})"
        print(f"{r['ranks_active']:12d} | {f(r['per_rank_h2d']):23s} | {f(r['per_rank_d2h']):23s} | {f(r['per_rank_both_each_way'])}")
    if args.out:
        json.dump({"world": world, "mib": args.mb, "affinity": affs, "rows": rows}, open(args.out, "w"), indent=1)
if world > 1:
    dist.destroy_process_group()
