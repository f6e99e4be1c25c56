"""Developer run: C2-shaped fwd+bwd a few times (CUDA-event timing; also the target of ncu launch lists)."""
import sys, torch
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL
from tests import _cases
from espnet_b200 import _lib
import os
if os.environ.get("B200_LIN"):
    _lib.load().b200_ctc_debug_force_log_sweep(0)
B, T, V, L = (16, 4000, 5000, 1000) if os.environ.get("B200_C5") else (64, 500, 5000, 150)
dt = torch.bfloat16 if len(sys.argv) < 2 or sys.argv[1] == "bf16" else torch.float32
full = len(sys.argv) > 2 and sys.argv[2] == "full"
n = int(sys.argv[3]) if len(sys.argv) > 3 else 20
c = _cases.make_loss_case(B, T, V, L, seed=1236, dtype=dt, full=full)
xt = c["x"].cuda().transpose(0, 1)
plan = CL._Plan(xt, c["ys"].cuda().view(-1), torch.arange(B) * L, c["ilen"], c["olen"], L, 0)
grad = CL._empty_like_strided(xt)
def step():
    return CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
for _ in range(3):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    nll, ls, _ = step()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / n
print(f"fused fwd+bwd {us:.1f} us/step  {B / us * 1e6:.0f} utt/s  loss {float(ls) / B:.4f}")
