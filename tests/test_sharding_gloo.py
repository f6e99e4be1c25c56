"""The N>1 path on CPU: two gloo ranks shard a batch by utterance, each computes the nll of its
own block, and one 2-float all-reduce gives the global mean loss.  The per-rank
nll here comes from the reference formula on CPU (no GPU in this test); what is under test is
the sharding and the reduction."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import _cases


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _nll(x, ilen, olen, ys):
    tg = torch.cat([ys[i, : olen[i]] for i in range(x.shape[0])])
    lp = x.double().transpose(0, 1).log_softmax(2)
    return torch.nn.functional.ctc_loss(lp, tg, ilen, olen, reduction="none")


def _worker(rank, world, port, n_utt, q):
    from espnet_b200.sharding import allreduce_loss, shard_bounds

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c = _cases.make_loss_case(n_utt, 40, 16, 8, seed=31)
    lo, hi = shard_bounds(n_utt, rank, world)
    nll = _nll(c["x"][lo:hi], c["ilen"][lo:hi], c["olen"][lo:hi], c["ys"][lo:hi])
    mean, n = allreduce_loss(nll.sum().float(), hi - lo)
    q.put((rank, lo, hi, float(mean), float(n)))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_utt", [6, 7])
def test_two_ranks_reproduce_the_global_mean(n_utt):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_utt, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    c = _cases.make_loss_case(n_utt, 40, 16, 8, seed=31)
    ref = float(_nll(c["x"], c["ilen"], c["olen"], c["ys"]).sum()) / n_utt
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == n_utt   # blocks tile the batch
    for r in res:
        np.testing.assert_allclose(r[3], ref, rtol=1e-6)
        assert r[4] == n_utt


def test_shard_bounds_cover_everything():
    from espnet_b200.sharding import shard_bounds

    for n in (0, 1, 7, 64, 512):
        for world in (1, 2, 3, 8):
            edges = [shard_bounds(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[k][1] == edges[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


def test_allreduce_without_process_group_is_local():
    from espnet_b200.sharding import allreduce_loss

    mean, n = allreduce_loss(torch.tensor(12.0), 4)
    assert float(mean) == 3.0 and float(n) == 4.0


def _ledger_worker(rank, world, port, q):
    from espnet_b200.sharding import LossLedger

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_utt = 3 + rank                                   # ragged blocks: 3 and 4 utterances
    led = LossLedger(4, n_utt, torch.device("cpu"))
    out = []
    for rnd in range(2):                               # the ledger is reusable after reduce()
        for i in range(4):
            led.slot(4 * rnd + i).fill_(float((rank + 1) * (i + 1) + rnd))
        means, n = led.reduce()
        out.append((means.tolist(), float(n)))
    q.put((rank, out))
    dist.destroy_process_group()


def test_ledger_exchanges_an_interval_of_steps_in_one_allreduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ledger_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, out in res:
        for rnd, (means, n) in enumerate(out):
            assert n == 7.0
            want = [((i + 1) + 2 * (i + 1) + 2 * rnd) / 7.0 for i in range(4)]
            np.testing.assert_allclose(means, want, rtol=1e-6)


def test_ledger_without_process_group_is_local():
    from espnet_b200.sharding import LossLedger

    led = LossLedger(3, 5, torch.device("cpu"))
    for i in range(3):
        led.slot(i).fill_(10.0 * (i + 1))
    means, n = led.reduce()
    assert float(n) == 5.0 and means.tolist() == [2.0, 4.0, 6.0]
    assert led.slot(4).data_ptr() == led.slot(1).data_ptr()
    with pytest.raises(ValueError):
        LossLedger(0, 1, torch.device("cpu"))


def _worker_step(rank, world, port, q):
    """StepLossExchange with the all-reduce transport: what trainer.py:386-397 does every iteration."""
    from espnet_b200.sharding import StepLossExchange

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_utt = 3 + rank                       # unequal shards: the backward scale is not 1
    ex = StepLossExchange(torch.device("cpu"), n_utt=n_utt)
    out = []
    for step in range(3):
        local = torch.tensor(10.0 * (step + 1) + rank)
        mean = ex.step(local)
        out.append((float(mean), float(ex.buf[1]), float(ex.scale())))
    q.put((rank, ex.transport, out))
    dist.destroy_process_group()


def test_step_loss_exchange_two_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_step, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, transport, out in res:
        assert transport == "allreduce"
        for step, (mean, n, scale) in enumerate(out):
            assert n == 7.0
            assert mean == pytest.approx((20.0 * (step + 1) + 1.0) / 7.0)
            # loss /= weight; loss *= world_size  ==  local mean * (n_local * world / n_global)
            assert scale == pytest.approx((3 + rank) * 2 / 7.0)


def test_step_loss_exchange_single_process():
    from espnet_b200.sharding import StepLossExchange

    ex = StepLossExchange(torch.device("cpu"), n_utt=4)
    assert ex.transport == "none"
    assert float(ex.step(torch.tensor(8.0))) == 2.0 and float(ex.scale()) == 1.0
