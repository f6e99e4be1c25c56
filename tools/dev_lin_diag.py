"""Developer check: where do the linear-domain and the log-domain sweeps disagree? (run on the GPU box)"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from espnet_b200 import _lib, ctc_loss as CL
from tests import _cases

kind = sys.argv[1] if len(sys.argv) > 1 else "level"
c = _cases.make_loss_case_kind(kind, 64, 500, 5000, 150, seed=1236, dtype=torch.bfloat16)
B, L = 64, c["ys"].shape[1]
xt = c["x"].cuda().transpose(0, 1)
lib = _lib.load()
out = {}
for name, flag in (("log", 1), ("lin", 0)):
    lib.b200_ctc_debug_force_log_sweep(flag)
    plan = CL._Plan(xt, c["ys"].cuda().reshape(-1), torch.arange(B) * L, c["ilen"], c["olen"], L, 0)
    nll, _, g = CL.ctc_loss_fwd_bwd(xt, None, None, None, None, L, grad_scale=1.0, plan=plan)
    torch.cuda.synchronize()
    out[name] = (nll.cpu().numpy(), g.transpose(0, 1).float().cpu().numpy(), plan.status())
print("status log/lin", out["log"][2], out["lin"][2])
d = np.abs(out["lin"][1] - out["log"][1])
print("max diff", d.max(), "nll diff", np.abs(out["lin"][0] - out["log"][0]).max())
bad = np.argwhere(d > 0.05)
print("entries off by > 0.05:", len(bad))
ub = sorted(set(int(b) for b in bad[:, 0]))
for b in ub[:6]:
    rows = bad[bad[:, 0] == b]
    ts = sorted(set(int(t) for t in rows[:, 1]))
    Tb, Lb = int(c["ilen"][b]), int(c["olen"][b])
    ys = c["ys"][b, :Lb].tolist()
    print(f"utt {b}: ilen {Tb} olen {Lb} half {Tb // 2}; bad frames {len(ts)}: {ts[:12]} ... {ts[-4:]}")
    for t in ts[:3] + ts[-2:]:
        cols = rows[rows[:, 1] == t][:, 2].tolist()
        pos = [[k for k, y in enumerate(ys) if y == v] if v != 0 else ["blank"] for v in cols]
        print(f"   t={t}: columns {cols} label positions {pos} lin {out['lin'][1][b, t, cols].tolist()} log {out['log'][1][b, t, cols].tolist()}")
