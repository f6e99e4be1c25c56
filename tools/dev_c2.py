"""Developer run: C2-shaped fwd+bwd a few times (for ncu launch lists)."""
import sys, torch
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL
from tools.dev_check_loss import make
B, T, V, L = 64, 500, 5000, 150
dt = torch.bfloat16 if len(sys.argv) < 2 or sys.argv[1] == "bf16" else torch.float32
full = len(sys.argv) > 2 and sys.argv[2] == "full"
x, ilen, olen, ys = make(B, T, V, L, dt, full=full)
xt = x.cuda().transpose(0, 1)
plan = CL._Plan(xt, ys.cuda().view(-1), torch.arange(B) * L, ilen, olen, L, 0)
grad = CL._empty_like_strided(xt)
for _ in range(4):
    CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
torch.cuda.synchronize()
print("ok")
