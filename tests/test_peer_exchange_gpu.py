"""The per-step loss exchange over peer memory (espnet_b200/csrc/peer.cu) on two or more GPUs of one box: the totals
written by b200_peer_sum_exchange equal an NCCL all-reduce of the same pairs, step after step, and are bitwise
identical on every rank.  Needs >= 2 GPUs (skipped on a single-GPU box; the CPU-side logic of StepLossExchange is
covered by tests/test_sharding_gloo.py)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    from espnet_b200.sharding import StepLossExchange

    ex = StepLossExchange(dev, n_utt=64 + rank, transport="peer")
    assert ex.transport == "peer"
    g = torch.Generator().manual_seed(5)
    vals = torch.rand(40, world, generator=g) * 1000.0          # every rank knows every rank's sums
    ok = True
    for step in range(40):
        mean = ex.step(vals[step, rank].to(dev))
        ref = torch.tensor([float(vals[step, rank]), 64.0 + rank], device=dev)
        dist.all_reduce(ref)
        got = ex.buf[0:2].clone()
        # sums in rank order, exactly
        s = torch.zeros((), dtype=torch.float32)
        for r in range(world):
            s = s + vals[step, r]
        ok = ok and float(got[0]) == float(s) and float(got[1]) == float(sum(64 + r for r in range(world)))
        ok = ok and abs(float(ref[0]) - float(got[0])) <= 1e-3 and float(ref[1]) == float(got[1])
        ok = ok and abs(float(mean) - float(got[0]) / float(got[1])) < 1e-6
        gathered = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(gathered, got)
        ok = ok and all(torch.equal(gathered[0], t) for t in gathered)   # bitwise the same on every rank
    ok = ok and not ex.peer.failed()
    ok = ok and float(ex.scale()) == pytest.approx((64 + rank) * world / sum(64 + r for r in range(world)))
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_peer_written_sums_equal_allreduce():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs on one box")
    import torch.multiprocessing as mp

    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, 29611, out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}
