"""Developer run: b200_beam_prune at the C3 shape (32 utterances x beam 10 x V=5000) against the eager formula."""
import sys, torch
sys.path.insert(0, ".")
from espnet_b200.recog_batch import beam_prune, LOGZERO

batch, beam, V = 32, 10, 5000
g = torch.Generator(device="cuda").manual_seed(1)
logp = torch.log_softmax(torch.randn(batch * beam, V, generator=g, device="cuda"), -1)
ctc = torch.log_softmax(torch.randn(batch * beam, V, generator=g, device="cuda"), -1) * 3
vs = torch.randn(batch, beam, generator=g, device="cuda") - 20


def eager():
    local = 0.7 * logp
    local[:, 0] = LOGZERO
    local = (local + 0.3 * ctc).view(batch, beam, V)
    eos_v = local[:, :, V - 1] + vs
    acc = vs.view(batch, beam, 1) + local
    acc[:, :, V - 1] = LOGZERO
    return torch.topk(acc.view(batch, -1), beam, 1)


def timeit(fn, n=200):
    for _ in range(5):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


s, ids, *_ = beam_prune(logp, ctc, vs, V - 1, 0.7, 0.3, False)
rs, rids = eager()
print("equal", torch.equal(s, rs), torch.equal(ids, rids))
print("b200_beam_prune us", timeit(lambda: beam_prune(logp, ctc, vs, V - 1, 0.7, 0.3, False)))
print("eager formula us", timeit(eager))
