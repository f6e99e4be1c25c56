"""Developer run (library built with B200_NVCC_EXTRA=-DB200_TIMELINE): global-timer stamps of the kernels of one C2
call: when the stream kernel starts / ends, when the sweeps start / end (per utterance), when the patch kernel's warps
see their first row and finish their last."""
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
lib = _lib.load()
lib.b200_debug_timeline.argtypes = [ctypes.c_void_p, ctypes.c_int]
for rep in range(6):
    torch.cuda.synchronize()
    lib.b200_debug_timeline(None, 1)
    CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
    torch.cuda.synchronize()
    buf = np.zeros(1024, dtype=np.uint64)
    lib.b200_debug_timeline(buf.ctypes.data, 0)
    t0 = int(buf[0])
    rel = lambda v: (int(v) - t0) / 1e3
    ends = np.array([[rel(buf[64 + 2 * b + d]) for d in range(2)] for b in range(B)])
    utt_ready = ends.max(1)
    print(f"rep {rep}: stream {rel(buf[0]):.1f}..{rel(buf[1]):.1f} us | sweeps start {rel(buf[2]):.1f}, last end {rel(buf[3]):.1f} "
          f"(utterances ready: first {utt_ready.min():.1f}, median {np.median(utt_ready):.1f}, last {utt_ready.max():.1f}) | "
          f"patch kernel resident {rel(buf[6]):.1f}, first row {rel(buf[4]):.1f}, last warp done {rel(buf[5]):.1f}")
