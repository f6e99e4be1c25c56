"""Developer run: time the forward (rowstats etc.) under debug variants."""
import os, sys, torch
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL, _lib
from tools.dev_check_loss import make
B, T, V, L = 64, 500, 5000, 150
x, ilen, olen, ys = make(B, T, V, L, torch.bfloat16)
xt = x.cuda().transpose(0, 1)
plan = CL._Plan(xt, ys.cuda().view(-1), torch.arange(B) * L, ilen, olen, L, 0)
nll = torch.empty(B, device="cuda")
lib = _lib.load()
def fwd():
    lib.b200_ctc_loss_forward(*plan.common_args(), nll.data_ptr(), None, plan.ws_ptr, plan.ws_bytes, torch.cuda.current_stream().cuda_stream)
for v in sys.argv[1:]:
    os.environ["B200_DEBUG_VARIANT"] = v
    for _ in range(3): fwd()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10): fwd()
    e1.record(); torch.cuda.synchronize()
    print("variant", v, "forward us", e0.elapsed_time(e1) * 100)
