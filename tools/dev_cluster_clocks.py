"""Developer run (library built with B200_NVCC_EXTRA=-DB200_SWEEP_CLOCKS): SM-clock stamps of the cluster sweep's chunks
64..127 for eight warps of the longest C5 utterance."""
import ctypes, sys, torch
import numpy as np
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL, _lib
from tests import _cases
B, T, V, L = 16, 4000, 5000, 1000
c = _cases.make_loss_case(B, T, V, L, seed=1236, dtype=torch.bfloat16, full=False)
xt = c["x"].cuda().transpose(0, 1)
plan = CL._Plan(xt, c["ys"].cuda().view(-1), torch.arange(B) * L, c["ilen"], c["olen"], L, 0)
grad = CL._empty_like_strided(xt)
for _ in range(3):
    CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
torch.cuda.synchronize()
lib = _lib.load()
buf = np.zeros(2 * 8 * 4 * 64, dtype=np.int64)
lib.b200_debug_sweep_clocks.argtypes = [ctypes.c_void_p]
print("rc", lib.b200_debug_sweep_clocks(buf.ctypes.data))
k = buf.reshape(2, 8, 4, 64)
names = [0, 1, 7, 8, 9, 16, 24, 31]
for d in range(2):
    print("direction", d)
    for w in range(8):
        a = k[d, w]
        if a[0, 1] == 0:
            continue
        cs = list(range(1, 62))
        tot = np.array([a[0, x + 1] - a[0, x] for x in cs])
        beg = np.array([a[1, x] - a[0, x] for x in cs])
        pol = np.array([a[2, x] - a[1, x] for x in cs])
        frm = np.array([a[3, x] - a[2, x] for x in cs])
        rest = np.array([a[0, x + 1] - a[3, x] for x in cs])
        print(f" warp {names[w]:2d}: chunk {tot.mean():.0f} (min {tot.min()}, max {tot.max()}) = loads+chunk_begin {beg.mean():.0f} "
              f"(even {beg[1::2].mean():.0f} / odd {beg[0::2].mean():.0f}), boundary poll {pol.mean():.0f}, 8 frames {frm.mean():.0f}, "
              f"rest {rest.mean():.0f};  lag to warp 0 at chunk 96: {int(a[0, 32] - k[d, 0, 0, 32])}")
