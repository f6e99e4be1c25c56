"""Peer-memory transport of ``sharding.StepLossExchange``: one kernel launch per step, no NCCL on the step path.

Every rank owns a slot ring in symmetric memory (``torch.distributed._symmetric_memory``: allocated here, peers'
addresses exchanged once at set-up); ``b200_peer_sum_exchange`` stores this rank's (sum of nll, utterance count) into
every peer's ring over NVLink and waits for the peers' entries in its own (``espnet_b200/csrc/peer.cu``).
Replaces the two all-reduces of ``espnet2/torch_utils/recursive_op.py:24,50`` for this message.
"""
import ctypes

import torch
import torch.distributed as dist

from . import _lib


class PeerSlots(object):
    def __init__(self, device, group=None):
        import torch.distributed._symmetric_memory as symm

        self.device = torch.device(device)
        group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        lib = _lib.load()
        n = lib.b200_peer_exchange_ring_floats(self.world)
        if n == 0:
            raise RuntimeError(f"peer exchange serves 1..16 ranks, got {self.world}")
        self.ring = symm.empty(n, dtype=torch.float32, device=self.device)
        self.ring.zero_()
        self.handle = symm.rendezvous(self.ring, group)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        if len(ptrs) != self.world or any(p == 0 for p in ptrs):
            raise RuntimeError("symmetric memory rendezvous returned no peer pointers")
        self.ptrs = (ctypes.c_uint64 * self.world)(*ptrs)
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._status_ptr = self.status.data_ptr()
        self._fn = lib.b200_peer_sum_exchange
        self.seq = 0
        torch.cuda.synchronize(self.device)
        dist.barrier(group)   # every ring is zeroed before anyone stores into it

    def exchange(self, local_sum, sum_scale, local_count, pair):
        """local_sum * sum_scale: this rank's sum (fp32 1-element device tensor x host number); local_count: its
        utterance count (host number); pair: fp32 (4,) device tensor that receives [total sum, total count, global
        mean, backward scale].  One kernel launch, nothing else."""
        self.seq += 1
        if self.seq >= 2 ** 31 - 1:
            raise RuntimeError("peer exchange sequence number exhausted")
        # (lean on purpose: this runs on every training step of every rank -- no device guard when the device is
        #  already current, raw stream handle)
        dev = self.device
        pp = pair.data_ptr()
        stream = torch._C._cuda_getCurrentRawStream(dev.index)
        if torch.cuda.current_device() == dev.index:
            rc = self._fn(self.ptrs, self.world, self.rank, self.seq, local_sum.data_ptr(), float(sum_scale),
                          float(local_count), pp, self._status_ptr, stream)
        else:
            with torch.cuda.device(dev):
                rc = self._fn(self.ptrs, self.world, self.rank, self.seq, local_sum.data_ptr(), float(sum_scale),
                              float(local_count), pp, self._status_ptr, stream)
        if rc:
            _lib.check(rc, "b200_peer_sum_exchange")
        return pair

    def failed(self):
        """True if a peer did not arrive within the kernel's time-out in any exchange so far (synchronises)."""
        return bool(int(self.status.item()))
