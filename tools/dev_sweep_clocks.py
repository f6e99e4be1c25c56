"""Developer run (library built with B200_NVCC_EXTRA=-DB200_SWEEP_CLOCKS): per-warp SM-clock stamps of the
staged sweep's full chunks for the longest utterance of the C2 batch."""
import ctypes, sys, torch
import numpy as np
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL, _lib
from tests import _cases
B, T, V, L = 64, 500, 5000, 150
c = _cases.make_loss_case(B, T, V, L, seed=1236, dtype=torch.bfloat16, full=False)
xt = c["x"].cuda().transpose(0, 1)
plan = CL._Plan(xt, c["ys"].cuda().view(-1), torch.arange(B) * L, c["ilen"], c["olen"], L, 0)
grad = CL._empty_like_strided(xt)
for _ in range(4):
    CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
torch.cuda.synchronize()
lib = _lib.load()
buf = np.zeros(2 * 8 * 4 * 64, dtype=np.int64)
lib.b200_debug_sweep_clocks.argtypes = [ctypes.c_void_p]
print("rc", lib.b200_debug_sweep_clocks(buf.ctypes.data))
k = buf.reshape(2, 8, 4, 64)
order = int(torch.argsort(c["ilen"] * 1000 + c["olen"], descending=True)[0])
print("ilen/olen max", int(c["ilen"].max()), int(c["olen"].max()))
for d in range(2):
    t0 = k[d, 0, 0, 1]
    print(f"direction {d}")
    for w in range(8):
        if k[d, w, 0, 1] == 0:
            continue
        a = k[d, w]
        cs = [c_ for c_ in range(1, 62) if a[0, c_] and a[0, c_ + 1]]
        tot = np.array([a[0, c_ + 1] - a[0, c_] for c_ in cs])
        wait = np.array([a[1, c_] - a[0, c_] for c_ in cs])
        beg = np.array([a[2, c_] - a[1, c_] for c_ in cs])
        pol = np.array([a[3, c_] - a[2, c_] for c_ in cs])
        frm = np.array([a[0, c_ + 1] - a[3, c_] for c_ in cs])
        print(f" warp {w}: start of chunk 1 at +{a[0,1]-t0}; per chunk: total {tot.mean():.0f} (min {tot.min()}, max {tot.max()}) = "
              f"lp wait+lds {wait.mean():.0f}, chunk_begin {beg.mean():.0f} (even {beg[1::2].mean():.0f}/odd {beg[0::2].mean():.0f}), "
              f"boundary poll {pol.mean():.0f}, 8 frames {frm.mean():.0f}")
        if w in (0, 4):
            print("   chunk totals:", tot[:40].tolist())
            print("   lag to warp0 at chunk c:", [int(a[0, c_] - k[d, 0, 0, c_]) for c_ in cs[:40:4]])
print("timeline (direction 0), cycles relative to warp 0 chunk 28 start; per warp stamps 0..3")
base = k[0, 0, 0, 28]
for c_ in range(28, 34):
    for w in range(8):
        if k[0, w, 0, c_]:
            print(f"  chunk {c_} warp {w}: " + " ".join(f"{int(k[0, w, j, c_] - base):7d}" for j in range(4)))
