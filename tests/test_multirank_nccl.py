"""Two ranks, two GPUs, NCCL: the batch sharding of tests/test_multirank_gloo.py with the CUDA engine as the per-shard compute.
Each rank runs the tcgen05 kernels on the frames it owns, the shards are gathered over NCCL, and EVERY rank compares its own
shard and the gathered batch with the oracle (skipped on a box with fewer than two GPUs; `gpurun --gpus 2`)."""
import os
import socket

import numpy as np
import pytest

from conftest import ROOT

torch = pytest.importorskip("torch")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2,
                                                  reason="needs two CUDA devices")]


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
    import torch.distributed as dist
    import ita_b200 as ita
    import ita_ref as R
    from conftest import to_engine_params
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ok = []
    for dims in ((128, 64, 192, 1), (256, 128, 64, 4)):                   # BASELINE configs[1] and configs[2]
        S, E, P, H = dims
        p = R.make_synthetic(S, E, P, H, seed=5, head_mode=1 if H > 1 else 0)
        X = R.make_input(total, S, E, seed=6)                              # every rank regenerates the global batch from the seed
        want = R.mha_forward(p, X)["Y"]
        eng = ita.ITAEngine(to_engine_params(ita, p), dev)
        assert eng.fused_supported, eng.fused_unsupported_reason()
        b, e = ita.shard_range(total, rank, world)
        y_local = eng.forward(torch.from_numpy(X[b:e]).to(dev))
        eng.sync_check()
        y_all = ita.gather_frames(y_local, total)
        ok += [np.array_equal(y_local.cpu().numpy(), want[b:e]), np.array_equal(y_all.cpu().numpy(), want)]
        eng.close()
    ok.append(ita.max_over_ranks(10.0 + rank, dev) == 10.0 + world - 1)
    np.save(os.path.join(out_dir, f"ok{rank}.npy"), np.array(ok))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 1])        # ragged (4 + 3); fewer frames than ranks (rank 1 owns an empty shard)
def test_two_gpu_shard_gather_with_the_cuda_engine(tmp_path, total):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), total, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        ok = np.load(os.path.join(str(tmp_path), f"ok{r}.npy"))
        assert ok.all(), (r, ok)
