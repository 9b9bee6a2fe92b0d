"""world_size-2 gloo test of the multi-rank plumbing (CPU): batch sharding, per-rank compute on the owned shard, the
final result gather and the max-over-ranks timing reduction.  On CPU the per-shard compute is the oracle (tests may use
it); on the GPU box the same plumbing runs with the CUDA engine (tests/test_gpu_parity.py, bench.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, out_dir):
    import sys
    for p in (ROOT, os.path.join(ROOT, "oracle")):
        sys.path.insert(0, p)
    import ita_b200 as ita
    import ita_ref as R
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = R.make_synthetic(64, 64, 64, 1, seed=5)
    X = R.make_input(total, 64, 64, seed=6)                  # every rank can regenerate the global batch from the seed
    b, e = ita.shard_range(total, rank, world)
    y_local = torch.from_numpy(R.mha_forward(p, X[b:e])["Y"]) if e > b else torch.zeros((0, 64, 64), dtype=torch.int8)
    y_all = ita.gather_frames(y_local, total)
    t = ita.max_over_ranks(10.0 + rank)
    if rank == 0:
        want = R.mha_forward(p, X)["Y"]
        np.save(os.path.join(out_dir, "ok.npy"), np.array([np.array_equal(y_all.numpy(), want), t == 10.0 + world - 1]))
    dist.barrier()
    dist.destroy_process_group()


def _run(total, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, total, str(tmp_path)), nprocs=2, join=True)
    ok = np.load(os.path.join(str(tmp_path), "ok.npy"))
    assert ok.all(), ok


def test_two_rank_shard_gather_even(tmp_path):
    _run(6, tmp_path)


def test_two_rank_shard_gather_ragged_and_tiny(tmp_path):
    _run(5, tmp_path)      # ragged: 3 + 2
    _run(1, tmp_path)      # fewer frames than ranks: rank 1 owns an empty shard
