"""Developer run: fused fwd+bwd at an arbitrary shape (CUDA-event timing; also a target for ncu launch lists).

    python tools/dev_shape.py B T V L [bf16|f32] [ragged|full] [n]      e.g. C5: 16 4000 5000 1000 bf16 ragged 5
"""
import sys, torch
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL
from tests import _cases
B, T, V, L = (int(a) for a in sys.argv[1:5])
dt = torch.float32 if len(sys.argv) > 5 and sys.argv[5] == "f32" else torch.bfloat16
full = len(sys.argv) > 6 and sys.argv[6] == "full"
n = int(sys.argv[7]) if len(sys.argv) > 7 else 10
c = _cases.make_loss_case(B, T, V, L, seed=1239, dtype=dt, full=full)
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
es = c["x"].element_size()
gb = (int(c["ilen"].sum()) * V * es + B * T * V * es) / 1e9
print(f"B={B} T={T} V={V} L={L} {dt}: fused fwd+bwd {us:.1f} us/step  {B / us * 1e6:.0f} utt/s  "
      f"{gb / (us * 1e-6):.0f} GB/s algorithmic  loss {float(ls) / B:.4f}")
