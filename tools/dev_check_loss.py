"""Developer check (GPU): b200 CTC loss+grad vs torch.nn.CTCLoss, with per-call timing."""
import sys, time
import torch
sys.path.insert(0, ".")
from espnet_b200 import ctc_loss as CL


def make(B, T, V, L, dtype, seed=0, full=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, T, V, generator=g).to(dtype)
    ilen = torch.randint(int(0.6 * T), T + 1, (B,), generator=g); ilen[0] = T
    if full: ilen[:] = T
    olen = torch.randint(max(L // 2, 1), L + 1, (B,), generator=g)
    ys = torch.randint(1, V, (B, L), generator=g)
    if L >= 4: ys[0, 1] = ys[0, 0]; ys[0, 3] = ys[0, 0]
    return x, ilen, olen, ys


def ref(x_btv, ilen, olen, ys, dt=torch.float64):
    B, T, V = x_btv.shape
    xr = x_btv.to(dt).transpose(0, 1).detach().requires_grad_()
    tg = torch.cat([ys[i, :olen[i]] for i in range(B)]).to(x_btv.device)
    lp = xr.log_softmax(2)
    nll = torch.nn.functional.ctc_loss(lp, tg, ilen, olen, reduction="none")
    (nll.sum() / B).backward()
    return nll.detach(), xr.grad.detach()


def run(B, T, V, L, dtype, full=False, iters=5):
    x, ilen, olen, ys = make(B, T, V, L, dtype, full=full)
    xd = x.cuda()
    off = torch.arange(B) * L
    xt = xd.transpose(0, 1)
    nll, ls, grad = CL.ctc_loss_fwd_bwd(xt, ys.cuda().view(-1), off, ilen, olen, L, grad_scale=1.0 / B)
    torch.cuda.synchronize()
    rn, rg = ref(xd, ilen, olen, ys)
    rel = ((nll.double() - rn).abs() / rn.abs()).max().item()
    gd = (grad.double() - rg).abs().max().item() / rg.abs().max().item()
    rn32, rg32 = ref(xd, ilen, olen, ys, torch.float32)
    gd32 = (rg32.double() - rg).abs().max().item() / rg.abs().max().item()
    print(f"B={B} T={T} V={V} L={L} {dtype}: nll rel {rel:.2e} grad normrel {gd:.2e} (torch fp32 vs fp64: {gd32:.2e}) loss_sum {ls.item():.4f} vs {rn.sum().item():.4f}")
    plan = CL._Plan(xt, ys.cuda().view(-1), off, ilen, olen, L, 0)
    for _ in range(2): CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(iters): CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0 / B, grad_out=grad, plan=plan)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    byt = 2 * B * T * V * x.element_size()
    print(f"   fwd+bwd {ms*1e3:.1f} us  -> {B/ms*1e3:.0f} utt/s, algorithmic {byt/ms/1e6:.0f} GB/s")


if __name__ == "__main__":
    run(3, 17, 7, 4, torch.float32)
    run(8, 200, 30, 40, torch.float32)
    run(4, 50, 64, 10, torch.float32)
    run(4, 50, 64, 10, torch.bfloat16)
    run(16, 300, 1000, 80, torch.float32)
    run(64, 500, 5000, 150, torch.bfloat16)
    run(64, 500, 5000, 150, torch.bfloat16, full=True)
    run(64, 500, 5000, 150, torch.float32)
    run(4, 2000, 5000, 1000, torch.bfloat16, iters=2)
