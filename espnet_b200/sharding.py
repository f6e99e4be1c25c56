"""Utterance sharding of the CTC loss across the GPUs of one box.

The path shards by utterance with no data-path collective: every rank runs the
kernels on its own contiguous block of utterances and the only message is one all-reduce of
two floats, (sum of nll, number of utterances), exactly what the reference's
``recursive_average`` does with (loss * weight, weight)
(``espnet2/torch_utils/recursive_op.py:24-53``).
"""
import torch
import torch.distributed as dist


def shard_bounds(n_utt, rank, world):
    """Contiguous block [lo, hi) of rank `rank`; blocks differ by at most one utterance."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside [0, {world})")
    base, rem = divmod(n_utt, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_loss(nll_sum, n_utt, group=None, out=None):
    """Global (mean loss, number of utterances) from this rank's (sum of nll, utterance count).

    nll_sum: 0-dim or 1-element tensor on the rank's device (NCCL: cuda, gloo: cpu).  One
    all-reduce(SUM) of a 2-element fp32 buffer; without an initialised process group the
    local values are returned unchanged.
    """
    buf = out if out is not None else torch.empty(2, dtype=torch.float32, device=nll_sum.device)
    buf[0:1].copy_(nll_sum.reshape(1))
    buf[1] = float(n_utt)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf[0] / buf[1], buf[1]


class LossLedger:
    """Per-step loss sums of this rank, exchanged once per `interval` steps.

    The kernels write each step's sum of nll straight into ``slot(i)`` (a 1-element view of
    one device buffer), so a step launches nothing beyond its own kernels; ``reduce()`` is ONE
    all-reduce(SUM) of the interval's sums plus the utterance count and returns the global mean
    loss of every step of the interval.  The reference's trainer exchanges the same pair every
    iteration (``espnet2/train/trainer.py:384-392`` -> ``recursive_average``) and only reads it
    at its log interval (``trainer.py:485-486``); batching the exchange to that interval
    keeps a 2-float collective (launch + NVLink latency, ~90 us measured at N=2) out of a
    ~225 us step.
    """

    def __init__(self, interval, n_utt, device, group=None):
        if interval < 1:
            raise ValueError("interval must be >= 1")
        self.interval, self.group = int(interval), group
        self.buf = torch.zeros(self.interval + 1, dtype=torch.float32, device=device)
        self.buf[self.interval] = float(n_utt)
        self.n_utt = float(n_utt)

    def slot(self, step):
        i = step % self.interval
        return self.buf[i:i + 1]

    def reduce(self):
        """(mean loss per step of the interval (interval,), global utterance count); resets the count slot."""
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(self.buf, op=dist.ReduceOp.SUM, group=self.group)
        n = self.buf[self.interval].clone()
        means = self.buf[: self.interval] / n
        self.buf[self.interval] = self.n_utt
        return means, n


class StepLossExchange:
    """Per-step exchange of (sum of nll, number of utterances) between the ranks, before backward.

    What ``espnet2/train/trainer.py:384-397`` does every iteration: ``stats, weight =
    recursive_average(stats, weight, distributed)`` (``recursive_op.py:46-53``: one all-reduce for the loss
    statistic, one for the weight), then ``loss /= weight; loss *= world_size``.  ``step(local_sum)``
    returns the global mean loss; ``scale()`` the factor B_local * world / sum(B) the local mean loss is
    multiplied by before ``backward`` (exactly 1 when every rank holds the same number of utterances, which
    keeps the one-pass gradient written during forward final).

    Transport: `peer` (symmetric memory: every rank's kernel stores its pair straight into the slot ring of
    all peers over NVLink and waits for theirs -- no NCCL launch on the step path) when
    ``espnet_b200.peer_exchange`` can set it up, else one NCCL/gloo all-reduce of a 2-float buffer.
    """

    def __init__(self, device, n_utt, group=None, transport="auto"):
        self.device, self.group = torch.device(device), group
        self.n_utt = float(n_utt)
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.buf = torch.zeros(4, dtype=torch.float32, device=self.device)  # total sum, total count, mean, scale
        self.peer = None
        self.transport = "none" if self.world == 1 else "allreduce"
        if self.world > 1 and transport in ("auto", "peer") and self.device.type == "cuda":
            try:
                from .peer_exchange import PeerSlots

                self.peer = PeerSlots(self.device, group)
                self.transport = "peer"
            except Exception as e:  # no symmetric memory on this box / build: the collective still works
                if transport == "peer":
                    raise
                self.why_not_peer = f"{type(e).__name__}: {e}"[:200]

    def step(self, local_sum, sum_scale=1.0):
        """local_sum * sum_scale: this rank's sum of nll (0-dim or 1-element device tensor x host number, e.g. the
        module's mean loss and B).  Returns the global mean loss (0-dim view, no kernel on the peer transport); the
        totals are left in ``self.buf[0:2]``."""
        if self.peer is not None:
            ls = local_sum if local_sum.dtype == torch.float32 else local_sum.float()
            self.peer.exchange(ls, sum_scale, self.n_utt, self.buf)
            return self.buf[2]
        self.buf[0:1].copy_(local_sum.reshape(1))
        if sum_scale != 1.0:
            self.buf[0:1].mul_(sum_scale)
        self.buf[1] = self.n_utt
        if self.world > 1:
            dist.all_reduce(self.buf[0:2], op=dist.ReduceOp.SUM, group=self.group)
        self.buf[2] = self.buf[0] / self.buf[1]
        self.buf[3] = (self.n_utt * self.world) / self.buf[1]
        return self.buf[2]

    def scale(self):
        """n_utt * world / (global utterance count) as a 0-dim device tensor (no host synchronisation): pass it as
        ``loss.backward(gradient=exchange.scale())``."""
        return self.buf[3]
