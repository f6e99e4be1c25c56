"""Developer check: what each piece the plugin call adds to the pre-planned C-ABI call costs ON THE DEVICE
(C2; run on the GPU box).  Each variant is timed with CUDA events over 300 back-to-back steps."""
import os
import sys

import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from espnet_b200 import _lib  # noqa: E402
from espnet_b200 import ctc_loss as cl  # noqa: E402
from espnet_b200.ctc import CTC  # noqa: E402

cfg = bench.CONFIGS["c2"]
batch = bench.make_batch(cfg, 1236)
pool = bench.label_pool(cfg, 1236, batch["ilen"])
x = batch["x"].cuda()
xt = x.transpose(0, 1)
T, B, V = xt.shape
ilen = batch["ilen"]
lib = _lib.load()
dev = x.device
plans = [cl._Plan(xt, t, None, ilen, o, cfg["L"], 0) for (t, o) in pool]
grads = [torch.empty_like(x).transpose(0, 1) for _ in range(3)]
nll = torch.empty(B, dtype=torch.float32, device=dev)
loss = torch.empty((), dtype=torch.float32, device=dev)
one = torch.ones((), dtype=torch.float32, device=dev)
st = torch.cuda.current_stream().cuda_stream
code = _lib.dtype_code(x.dtype)


def raw(i, g, fixed=None):
    plan = plans[(i if fixed is None else fixed) % len(plans)]
    rc = lib.b200_ctc_loss_fwd_bwd(*plan.common_args(), None, 1.0 / B, nll.data_ptr(), loss.data_ptr(), 1.0 / B,
                                   g.data_ptr(), g.stride(0), g.stride(1), plan.ws_ptr, plan.ws_bytes, st)
    assert rc == 0


def stage(i):
    t, o = pool[i % len(pool)]
    plan = plans[i % len(plans)]
    rc = lib.b200_ctc_stage_labels_pinned(t.data_ptr(), t.numel(), None, ilen.data_ptr(), o.data_ptr(), B,
                                          plan.targets_ptr, st)
    assert rc == 0


def rescale(i, g):
    plan = plans[i % len(plans)]
    rc = lib.b200_ctc_grad_rescale(g.data_ptr(), code, T, B, V, g.stride(0), g.stride(1), one.data_ptr(), 1, 1.0,
                                   plan.ilen_ptr, st)
    assert rc == 0


ctc = CTC(V, 8, ctc_type="b200")
xg = x.clone().requires_grad_()


def plugin(i):
    t, o = pool[i % len(pool)]
    xg.grad = None
    ctc.loss_fn(xg.transpose(0, 1), t, ilen, o).backward()


def timed(name, fn, n=300):
    for i in range(10):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:46s} {1e3 * e0.elapsed_time(e1) / n:8.1f} us/step", flush=True)


for rep in range(2):
    for k in range(len(plans)):
        timed(f"raw, label batch {k} only", lambda i: raw(i, grads[0], fixed=k), n=100)
    timed("raw, one gradient buffer", lambda i: raw(i, grads[0]))
    timed("raw, gradient buffers rotate (3)", lambda i: raw(i, grads[i % 3]))
    timed("raw + label staging copy", lambda i: (stage(i), raw(i, grads[0])))
    timed("raw + grad_rescale", lambda i: (raw(i, grads[0]), rescale(i, grads[0])))
    timed("raw + staging + rescale", lambda i: (stage(i), raw(i, grads[0]), rescale(i, grads[0])))
    timed("raw + staging + rescale, rotating gradient", lambda i: (stage(i), raw(i, grads[i % 3]), rescale(i, grads[i % 3])))
    timed("plugin call (loss_fn + backward)", plugin)
