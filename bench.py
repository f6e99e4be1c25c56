#!/usr/bin/env python
"""bench.py -- CTC loss fwd+bwd utterances/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c2|c5]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload: BASELINE.json configs[1] (C2: B=64 per GPU, T=500, V=5000, L<=150, bf16 logits, synthetic, SURVEY.md
8d input recipe); `--config c5` is configs[4] (B=16, T=4000, L<=1000).  With N GPUs every rank takes its own
utterances (weak scaling; N=8 at C2 is config C4, global B=512) with no data-path collective.

What a step is.  `value` times the PLUGIN CALL a trainer makes -- ``CTC(ctc_type="b200").loss_fn(th_pred,
th_target, th_ilen, th_olen)`` followed by ``loss.backward()`` (espnet2/asr/ctc.py:46-52 + autograd) -- on logits
resident in HBM, with a fresh label batch every step: the label host->device copy, the workspace and gradient
allocations and the launch arguments are all built inside the timed region.  `e2e` is the same call with the
LOGITS in pinned host memory (uploaded every step) and the loss read back.  `raw_abi` repeats the measurement
through the C ABI with everything prepared ahead (round 1's number), to show what the Python layer costs.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    "c2": dict(B=64, T=500, V=5000, L=150, name="C2: LibriSpeech-shaped CTC loss fwd+bwd, B=64 per GPU, T=500, V=5000 BPE, "
                                                 "L<=150, bf16 logits"),
    "c5": dict(B=16, T=4000, V=5000, L=1000, name="C5: long-form CTC loss fwd+bwd, B=16 per GPU, T=4000, V=5000, L<=1000, "
                                                  "bf16 logits"),
}
LABEL_POOL = 8        # distinct label batches cycled through the timed steps
NOMINAL_HBM_GBS = 8000.0


def make_batch(cfg, seed, full_length=False, dtype=torch.bfloat16):
    from tests import _cases

    return _cases.make_loss_case(cfg["B"], cfg["T"], cfg["V"], cfg["L"], seed=seed, dtype=dtype, full=full_length)


def label_pool(cfg, seed, ilen):
    """LABEL_POOL label batches for one logits batch: (concatenated targets, olen) on the host, as the data loader
    hands them to CTC.forward every step (labels drawn as in SURVEY 8d; every utterance stays feasible)."""
    from tests import _cases

    pool = []
    for k in range(LABEL_POOL):
        c = _cases.make_loss_case(cfg["B"], cfg["T"], cfg["V"], cfg["L"], seed=seed + 1000 * (k + 1))
        olen, ys = c["olen"].clone(), c["ys"]
        for b in range(cfg["B"]):
            lb = int(olen[b])
            rep = int((ys[b, 1:lb] == ys[b, : lb - 1]).sum()) if lb > 1 else 0
            if int(ilen[b]) < lb + rep:
                olen[b] = max(1, int(ilen[b]) // 2)
        tg = torch.cat([ys[b, : int(olen[b])] for b in range(cfg["B"])])
        pool.append((tg, olen))
    return pool


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [s.strip() for s in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            self._stop_evt.wait(0.02)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(self.samples[0][1]) if self.samples and self.samples[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.samples)}


def _gpu_cpu_affinity(world):
    """{gpu index: set of CPU cores next to it} for GPUs 0..world-1, from NVML (`nvmlDeviceGetCpuAffinity`: what the
    "CPU Affinity" column of `nvidia-smi topo -m` prints).  Eight ranks starting `nvidia-smi topo -m` at once took more
    than 10 s each on the 8-GPU box; the NVML query is a few milliseconds.  Falls back to parsing that command."""
    try:
        import pynvml

        pynvml.nvmlInit()
        try:
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            order = [int(v) for v in vis.split(",")] if vis and all(v.strip().isdigit() for v in vis.split(",")) else None
            ncpu = os.cpu_count() or 1
            words = (max(ncpu, 1) + 63) // 64
            aff = {}
            for r in range(world):
                h = pynvml.nvmlDeviceGetHandleByIndex(order[r] if order else r)
                mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
                aff[r] = {64 * w + bit for w, m in enumerate(mask) for bit in range(64) if (int(m) >> bit) & 1}
            return aff, "nvml"
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        pass
    out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=90).stdout
    lines = [ln for ln in out.splitlines() if ln.strip()]
    hdr = next(ln for ln in lines if "CPU Affinity" in ln)
    k = [c.strip() for c in hdr.split("\t")].index("CPU Affinity")
    aff = {}
    for ln in lines:
        f = [c.strip() for c in ln.split("\t")]
        if f and f[0].startswith("GPU") and f[0][3:].isdigit() and len(f) > k:
            cores = set()
            for part in f[k].split(","):
                lo, _, hi = part.partition("-")
                cores.update(range(int(lo), int(hi or lo) + 1))
            aff[int(f[0][3:])] = cores
    return aff, "nvidia-smi topo -m"


def pin_to_gpu_numa_node(local_rank, world):
    """Pins this rank (and therefore the first touch of its pinned host buffers) to the CPU cores next to its GPU.
    With N ranks uploading 257 MB per step each, unpinned processes put their staging buffers wherever the scheduler
    ran them first: round 1 measured 52 GB/s of H2D per GPU at N=1 falling to 23 GB/s at N=8; and with one loss
    exchange per step every rank waits for the slowest host.  GPUs that share one affinity set split its cores evenly
    between their ranks.  Returns a description for the JSON line; never raises."""
    try:
        aff, source = _gpu_cpu_affinity(world)
        allowed = os.sched_getaffinity(0)
        mine = aff[local_rank] & allowed
        if not mine:
            return {"pinned": False, "why": "no usable cores next to GPU %d" % local_rank, "source": source}
        sharers = [r for r in range(world) if aff.get(r) == aff[local_rank]]
        if len(sharers) > 1:   # several GPUs on one node: every rank gets its own slice of the node's cores
            lst = sorted(mine)
            per = max(1, len(lst) // len(sharers))
            k = sharers.index(local_rank)
            mine = set(lst[k * per:(k + 1) * per]) or mine
        os.sched_setaffinity(0, mine)
        lst = sorted(mine)
        return {"pinned": True, "cores": len(mine), "first_core": lst[0], "last_core": lst[-1],
                "gpus_sharing_the_node": len(sharers), "source": source}
    except Exception as e:
        return {"pinned": False, "why": f"{type(e).__name__}: {e}"[:120]}


# ----------------------------------------------------------------------------------------------- reference legs
def reference_ctc_module(V):
    """The reference's own espnet2 CTC class (oracle/_ref, staged from /root/reference) when it travelled to this
    box, else None (the callers then restate CTC.loss_fn: log_softmax + torch.nn.CTCLoss(sum) / B)."""
    try:
        from oracle import stage_ref

        if stage_ref.activate():
            from espnet2.asr.ctc import CTC as RefCTC

            return RefCTC(V, 8, ctc_type="builtin")
    except Exception:
        pass
    return None


def reference_loss_fn(V):
    mod = reference_ctc_module(V)
    if mod is not None:
        return mod.loss_fn, "reference"
    crit = torch.nn.CTCLoss(reduction="sum")
    return (lambda xt, tg, il, ol: crit(xt.log_softmax(2), tg, il, ol) / xt.size(1)), "port"


def cpu_reference_pass(batch, n_utt, threads, loss_fn):
    """The reference's CPU path for this metric: CTC.loss_fn (espnet2/asr/ctc.py:46-52), forward and backward."""
    torch.set_num_threads(threads)
    x = batch["x"][:n_utt].float()
    ilen, olen, ys = batch["ilen"][:n_utt], batch["olen"][:n_utt], batch["ys"][:n_utt]
    tg = torch.cat([ys[i, : olen[i]] for i in range(n_utt)])
    xt = x.transpose(0, 1).detach().requires_grad_()
    t0 = time.perf_counter()
    loss = loss_fn(xt, tg, ilen, olen)
    loss.backward()
    return time.perf_counter() - t0, float(loss.detach())


def gpu_reference_pass(batch, dev, steps=10):
    """SURVEY 8d "reference GPU baseline on the same box": CTC.loss_fn (espnet2/asr/ctc.py:46-52) on a CUDA device
    -- .float() of the bf16 logits (autocast's promotion), log_softmax(2), torch.nn.CTCLoss(sum) / B, backward to the
    fp32 logits -- timed with CUDA events."""
    x = batch["x"].to(dev)
    ilen, olen = batch["ilen"].to(dev), batch["olen"].to(dev)
    tg = torch.cat([batch["ys"][i, : batch["olen"][i]] for i in range(x.size(0))]).to(dev)
    loss_fn, kind = reference_loss_fn(x.size(2))
    xt = x.transpose(0, 1)

    def one():
        xf = xt.float().requires_grad_()
        loss = loss_fn(xf, tg, ilen, olen)
        loss.backward()
        return loss

    for _ in range(3):
        one()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = one()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"value": x.size(0) / (ms * 1e-3), "unit": "utterances/s", "ms_per_step": ms, "steps": steps,
            "loss": float(loss.detach()), "kind": kind,
            "what": "the reference's espnet2 CTC.loss_fn (builtin) on the same GPU: bf16 logits .float() -> log_softmax(2) "
                    "-> torch.nn.CTCLoss(sum)/B -> backward (espnet2/asr/ctc.py:46-52 as it runs under autocast)"}


def bench_c5_leg(dev, peak, steps=10):
    """BASELINE config 5 (long form: B=16, T=4000, V=5000, L<=1000, bf16) inside the default line: the plugin call
    CTC(ctc_type="b200").loss_fn + loss.backward() on device-resident logits, and the reference's call on the same
    GPU.  (`python bench.py --config c5` runs the full protocol on this configuration.)"""
    from espnet_b200.ctc import CTC

    cfg = CONFIGS["c5"]
    batch = make_batch(cfg, 1239)
    tg = torch.cat([batch["ys"][b, : int(batch["olen"][b])] for b in range(cfg["B"])])
    x = batch["x"].to(dev).requires_grad_()
    ctc = CTC(cfg["V"], 8, ctc_type="b200")

    def one():
        x.grad = None
        loss = ctc.loss_fn(x.transpose(0, 1), tg, batch["ilen"], batch["olen"])
        loss.backward()
        return loss

    for _ in range(5):
        one()
    torch.cuda.synchronize()
    # median of three timed blocks: a 2.7 GB working set right after the other legs freed theirs can see one block
    # pay for the caching allocator's cudaMalloc / cudaFree (observed once: 3.2 ms instead of 1.1 ms per step)
    blocks = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = one()
        e1.record()
        torch.cuda.synchronize()
        blocks.append(e0.elapsed_time(e1) / steps)
    ms = sorted(blocks)[1]
    nbytes = 2 * cfg["B"] * cfg["T"] * cfg["V"] * 2       # logits read once + gradient written once
    lattice = 2 * 4 * int(batch["ilen"].sum()) * (2 * cfg["L"] + 8)   # alpha + beta rows, written once and read once
    gbs = nbytes / (ms * 1e-3) / 1e9
    out = {"workload": cfg["name"], "value": cfg["B"] / (ms * 1e-3), "unit": "utterances/s", "ms_per_step": ms, "steps": steps,
           "ms_per_step_blocks": blocks,
           "loss": float(loss.detach()), "algorithmic_bytes": nbytes, "lattice_bytes_alpha_plus_beta": lattice,
           "roofline_step": {"achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                             "frac_with_lattice_traffic": (nbytes + 2 * lattice) / (ms * 1e-3) / 1e9 / peak},
           "note": "lattice sweeps: 4-CTA thread-block clusters per (utterance, direction), hand-off through distributed "
                   "shared memory"}
    x.grad = None
    del x
    torch.cuda.empty_cache()
    try:
        out["reference_gpu"] = gpu_reference_pass(batch, dev, steps=3)
        out["reference_gpu"].pop("what", None)
    except Exception as e:  # reported baseline only
        out["reference_gpu"] = {"error": str(e)[:200]}
    torch.cuda.empty_cache()
    return out


def cpu_oracle_pass(batch, n_utt):
    """The C restatement (oracle/ctc_oracle.c), one utterance per OpenMP thread."""
    import numpy as np

    from oracle import ctc_oracle as O

    x = np.ascontiguousarray(batch["x"][:n_utt].float().numpy().transpose(1, 0, 2))
    ys = batch["ys"][:n_utt].numpy()
    t0 = time.perf_counter()
    O.ctc_loss(x, ys.reshape(-1), np.arange(n_utt) * ys.shape[1], batch["ilen"][:n_utt].numpy(),
               batch["olen"][:n_utt].numpy(), grad_scale=1.0 / n_utt)
    return time.perf_counter() - t0, O.num_threads()


def cpu_c1_reference(threads):
    """BASELINE config 1, the reference's own CPU-runnable case: B=8, T=200, V=30, L<=40, fp32, ctc_type=builtin."""
    from tests import _cases

    c = _cases.make_loss_case(8, 200, 30, 40, seed=1235)
    loss_fn, kind = reference_loss_fn(30)
    cpu_reference_pass(c, 8, threads, loss_fn)
    n, t = 0, 0.0
    while t < 1.0 and n < 200:
        dt, _ = cpu_reference_pass(c, 8, threads, loss_fn)
        t += dt
        n += 1
    return {"value": 8 * n / t, "unit": "utterances/s", "cores": threads, "kind": kind, "ms_per_step": 1e3 * t / n,
            "workload": "C1: B=8, T=200, V=30, L<=40, fp32, CPU (BASELINE.json configs[0])"}


def run_reference(args, rank):
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    batch = make_batch(cfg, 1236, args.full_length)
    loss_fn, kind = reference_loss_fn(cfg["V"])
    n_utt = min(16, cfg["B"])  # bounded sample of the batch per step
    cpu_reference_pass(batch, n_utt, cores, loss_fn)   # (cold: thread pool, first touch of the buffers)
    t16 = min(cpu_reference_pass(batch, n_utt, cores, loss_fn)[0] for _ in range(2))
    # keep the whole run within about two and a half minutes whatever --steps is (the reference parallelises over the
    # utterances of a batch: a smaller sample per step costs it throughput, so it is only shrunk when it has to be)
    n_utt = max(1, min(n_utt, int(n_utt * 150.0 / max(args.steps * t16, 1e-9))))
    for _ in range(args.warmup):
        cpu_reference_pass(batch, n_utt, cores, loss_fn)
    t = 0.0
    for _ in range(args.steps):
        dt, _ = cpu_reference_pass(batch, n_utt, cores, loss_fn)
        t += dt
    value = n_utt * args.steps / t
    sample = f"{n_utt} of the {cfg['B']} {args.config.upper()} utterances per step, fp32 copy of the bf16 logits, torch CPU"
    what = ("the reference's own espnet2 CTC.loss_fn (builtin), staged unmodified under oracle/_ref" if kind == "reference"
            else "espnet2/asr/ctc.py:46-52 restated: log_softmax + torch.nn.CTCLoss(sum)/B (oracle/_ref is not staged on this box)")
    line = {"impl": "reference", "metric": "ctc_loss_fwd_bwd_utterances_per_sec", "value": value,
            "unit": "utterances/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, 1),
            "cpu_baseline": {"value": value, "unit": "utterances/s", "cores": cores, "kind": kind,
                             "sample": sample + " (" + what + " + backward)"},
            "e2e": {"value": value, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    cfg = CONFIGS[args.config]
    return {"workload": cfg["name"] + (", all utterances full length" if args.full_length else ", ilen~U[0.6T,T] (SURVEY 8d)"),
            "global_batch": cfg["B"] * world, "frames": cfg["T"], "vocab": cfg["V"], "max_labels": cfg["L"],
            "call": "CTC(ctc_type='b200').loss_fn(th_pred, th_target, th_ilen, th_olen) + loss.backward(), fresh labels every step",
            "sharding": (f"utterances x{world}, no data-path collective; per-step loss sums exchanged through "
                         "peer-written slots (espnet_b200/sharding.py)") if world > 1 else "single GPU",
            "l2_policy": "inputs larger than L2 (logits + gradient per step: 640 MB at C2 vs 126 MB L2)"}


# ----------------------------------------------------------------------------------------------- prefix scoring
def bench_decode_loop(dev, margin=0, pre_beam=None):
    """The whole multi-utterance search loop at BASELINE config 3 (B=32 utterances x beam 10, T=500, V=5000,
    ctc_weight 0.3): espnet_b200.recog_batch.recognize_beam_batch = CTC prefix scorer + top-k + state gathers +
    ended-hypothesis bookkeeping (decoders.py:776-960), driven by a toy attention decoder (embedding -> linear ->
    log-softmax, attention weights = a window that moves with the output position).  Reported as output positions
    per second of the WHOLE loop; the toy decoder's own kernels are inside the number."""
    from espnet_b200.recog_batch import DecoderStep, recognize_beam_batch
    from tests import _cases

    B, H, T_, V_, L = 32, 10, 500, 5000, 24
    c = _cases.make_prefix_case(B, T_, V_, H, seed=1238, nsteps=1, kind="trained", min_ratio=0.6, L=L)
    lpz = c["lp"].to(dev)
    g = torch.Generator().manual_seed(1239)
    emb = (torch.randn(V_, 64, generator=g) * 0.3).to(dev)
    proj = (torch.randn(64, V_, generator=g) * 0.3).to(dev)
    frames = torch.arange(T_, device=dev, dtype=torch.float32).view(1, -1)

    class Toy(DecoderStep):
        def start(self, n_bb):
            return 0

        def step(self, vy, pos):
            logp = torch.log_softmax(emb[vy] @ proj, dim=1)
            centre = min(8.0 + 14.0 * pos, T_ - 1.0)
            att = torch.softmax(-0.05 * (frames - centre) ** 2, dim=1).expand(vy.numel(), T_)
            return logp, att, pos + 1

        def select(self, pos, vidx):
            return pos

    def run():
        stats = {}
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = recognize_beam_batch(Toy(), lpz, c["xlens"], H, V_ - 1, V_ - 1, ctc_weight=0.3, nbest=1, ctc_margin=margin,
                                   pre_beam=pre_beam, stats=stats)
        torch.cuda.synchronize()
        return stats["steps"], time.perf_counter() - t0, out

    run()                      # warm-up (allocator, first launches)
    runs = sorted((run() for _ in range(3)), key=lambda r: r[1] / r[0])
    n, dt, out = runs[1]       # median of three
    return {"metric": "decode_loop_positions_per_sec", "value": n / dt, "unit": "output positions/s", "us_per_position": 1e6 * dt / n,
            "positions": n, "best_len_utt0": len(out[0][0]["yseq"]),
            "config": {"workload": f"C3 search loop: B=32 x beam 10, T=500, V=5000, ctc_weight 0.3, margin={margin}, "
                                   + (f"pre-beam {pre_beam}" if pre_beam else "all tokens scored")
                                   + ", toy attention decoder, one device->host copy per position"}}


def bench_prefix(dev, peak, steps=20, snum=0):
    """Second half of the metric: CTC prefix-score steps/s at BASELINE config 3 (B=32 utterances x beam 10, T=500,
    V=5000, fp32 log-probs).  snum=0: full-vocabulary scoring as recognize_beam_batch does on CUDA; snum=15:
    scoring_ids = top-15 per hypothesis (1.5 x beam), the pre-beam path BatchBeamSearch takes
    (batch_beam_search.py:228-238).  One step = __call__ + index_select_state for all 320 hypotheses; hypotheses
    advance by a seeded random walk."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200

    B, H, T_, V_ = 32, 10, 500, 5000
    eos = V_ - 1
    g = torch.Generator(device=dev).manual_seed(1237)
    lp = torch.log_softmax(torch.randn(B, T_, V_, generator=g, device=dev), dim=-1)
    gc = torch.Generator().manual_seed(1237)
    xlens = torch.randint(300, T_ + 1, (B,), generator=gc)
    xlens[0] = T_
    nwalk = steps + 3
    hyp = torch.randint(0, H, (nwalk, B, H), generator=gc)
    if snum:
        ids_all = torch.stack([torch.stack([torch.randperm(V_ - 2, generator=gc)[:snum] + 1 for _ in range(B * H)])
                               for _ in range(nwalk)])                                  # (nwalk, n_bh, snum)
        pick = torch.randint(0, snum, (nwalk, B, H), generator=gc)
        tok = ids_all.view(nwalk, B, H, snum)[torch.arange(nwalk).view(-1, 1, 1), torch.arange(B).view(1, -1, 1), hyp, pick]
        ids_dev = ids_all.to(dev)
    else:
        tok = torch.randint(1, V_ - 1, (nwalk, B, H), generator=gc)
        ids_dev = [None] * nwalk
    best = (hyp * V_ + tok).to(dev)
    parents = (hyp + torch.arange(B).view(1, -1, 1) * H).view(nwalk, -1).tolist()
    toks = tok.view(nwalk, -1).tolist()

    # The prefixes of every step are inputs of the walk: they are laid out before the timed region, as
    # (n_bh, len) device tensors for the b200 scorer and as token lists for the reference.
    ys, y = [], [[eos] for _ in range(B * H)]
    for i in range(nwalk):
        ys.append(y)
        y = [y[parents[i][k]] + [toks[i][k]] for k in range(B * H)]
    y_dev = [torch.tensor(yi, dtype=torch.int64, device=dev) for yi in ys]

    host_us = []

    def walk(make_impl, do_step, do_select, n):
        impl = make_impl()
        state = None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for i in range(n + 2):
            if i == 2:  # steps 0 and 1 (warm-up: first label, first step with a previous state) are not timed
                torch.cuda.synchronize()
                e0.record()
                h0 = time.perf_counter()
            st = do_step(impl, i, state)
            state = do_select(impl, st, best[i])
        e1.record()
        host_us.append((time.perf_counter() - h0) * 1e6 / n)   # time the host needed to ISSUE a step
        torch.cuda.synchronize()
        return n / (e0.elapsed_time(e1) * 1e-3)

    ours = walk(lambda: CTCPrefixScoreB200(lp.clone(), xlens, 0, eos),
                lambda m, i, s: m(y_dev[i], s, ids_dev[i])[1], lambda m, st, b: m.index_select_state(st, b), steps)
    if snum:
        # gathered log-prob columns (snum per hypothesis, whole 32-byte sectors) + states + the full-size score rows
        alg = 32 * B * H * snum * T_ + 4 * (2 * B * H * T_ * 2 + B * H * V_)
    else:
        alg = 4 * (B * T_ * V_ + 2 * B * H * T_ * 2 + B * H * V_)
    gbs = alg * ours / 1e9
    out = {"metric": "ctc_prefix_score_steps_per_sec", "value": ours, "unit": "steps/s", "us_per_step": 1e6 / ours,
           "config": {"workload": "C3: B=32 utterances x beam 10 (n_bh=320), T=500, V=5000, fp32 log-probs, "
                                  + ("scoring_ids=top-15 (pre-beam)" if snum else "scoring_ids=None") + ", margin=0",
                      "steps": steps},
           "algorithmic_bytes_per_step": alg, "host_issue_us_per_step": host_us[0],
           "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                        "frac_of_nominal_8TBs": gbs / NOMINAL_HBM_GBS,
                        "what": "whole step (all kernels of __call__ + index_select_state) against its algorithmic bytes"}}
    if snum:
        # the candidates' sums read a (B, V, T) column copy of x that the scorer builds at its first call with
        # scoring_ids (inside the untimed first step, once per batch of utterances): its cost, and the step without it
        plain = walk(lambda: CTCPrefixScoreB200(lp.clone(), xlens, 0, eos, columns=False),
                     lambda m, i, s: m(y_dev[i], s, ids_dev[i])[1], lambda m, st, b: m.index_select_state(st, b), steps)
        m = CTCPrefixScoreB200(lp.clone(), xlens, 0, eos)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        m._column_copy()
        e1.record()
        torch.cuda.synchronize()
        out["column_copy"] = {"one_time_us": e0.elapsed_time(e1) * 1e3, "bytes": 2 * 4 * B * T_ * V_,
                              "us_per_step_without_it": 1e6 / plain,
                              "what": "x (B,T,V) -> (B,V,T) once per batch of utterances, so that a candidate's frames are "
                                      "contiguous (b200_ctc_prefix_columns / b200_ctc_prefix_score_columns)"}
        del m
    ref_kind = None
    try:
        from oracle import stage_ref

        if stage_ref.activate():
            from espnet.nets.ctc_prefix_score import CTCPrefixScoreTH

            ref = walk(lambda: CTCPrefixScoreTH(lp.clone(), xlens, 0, eos),
                       lambda m, i, s: m(ys[i], s, ids_dev[i])[1],
                       lambda m, st, b: m.index_select_state(st, b), 3)
            ref_kind = "reference"
            what = "the reference's own CTCPrefixScoreTH (oracle/_ref, unmodified) on the same GPU, 3 steps"
        else:
            from oracle.prefix_th_port import PrefixScoreEager

            if snum:
                raise RuntimeError("the eager port covers the full-vocabulary path only")
            y_last = [[yk[-1] for yk in yi] for yi in ys]
            ref = walk(lambda: PrefixScoreEager(lp.clone(), xlens, 0, eos),
                       lambda m, i, s: m.step(y_last[i], i, s)[1], lambda m, st, b: m.select(st, b), 3)
            ref_kind = "port"
            what = "eager-PyTorch restatement of CTCPrefixScoreTH (oracle/prefix_th_port.py) on the same GPU, 3 steps"
        out["reference_gpu"] = {"value": ref, "unit": "steps/s", "kind": ref_kind, "what": what}
        out["speedup_vs_reference_gpu"] = ours / ref
    except Exception as e:  # e.g. out of memory on a shared box
        out["reference_gpu"] = {"error": str(e)[:200]}
    torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------------------------- the b200 arm
def run_b200(args, rank, world, local_rank):
    import torch.distributed as dist

    from espnet_b200 import _lib
    from espnet_b200 import ctc_loss as CL
    from espnet_b200.ctc import CTC
    from espnet_b200.sharding import StepLossExchange

    cfg = CONFIGS[args.config]
    Bn, T, V, LMAX = cfg["B"], cfg["T"], cfg["V"], cfg["L"]
    affinity = pin_to_gpu_numa_node(local_rank, world)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import datetime

        # NCCL prints its version banner on stdout when the communicator comes up (NCCL_DEBUG=VERSION on some
        # boxes): keep stdout for the ONE JSON line by pointing fd 1 at stderr until the first collective is done
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    lib = _lib.load()
    batch = make_batch(cfg, 1236 + rank, args.full_length)
    pool = label_pool(cfg, 1236 + rank, batch["ilen"])
    ilen_host = batch["ilen"]
    x_host = batch["x"].pin_memory()                      # (B,T,V) bf16, what ctc_lo produces
    x_dev = torch.empty_like(x_host, device=dev)
    x_dev.copy_(x_host)
    x_dev.requires_grad_()                                # the logits are the leaf the gradient is asked for
    # the trainer's upstream gradient is known (1 here; ctc_weight / accum_grad x loss scale in training): the
    # gradient written during forward is final, backward only verifies that on the device
    ctc = CTC(V, 8, ctc_type="b200", grad_scale_hint=1.0)
    exchange = StepLossExchange(dev, n_utt=Bn) if world > 1 else None
    nstep = [0]

    def step(x_leaf=x_dev):
        """One trainer-side CTC step: the plugin call with this step's labels, backward, loss exchange."""
        i = nstep[0]
        tg, olen = pool[i % LABEL_POOL]
        x_leaf.grad = None                                 # optimizer.zero_grad(set_to_none=True)
        loss = ctc.loss_fn(x_leaf.transpose(0, 1), tg, ilen_host, olen)
        nstep[0] = i + 1
        if exchange is None:
            loss.backward()
            return loss.detach()
        # trainer.py:384-397: (loss * weight).sum(); all-reduce of the statistic and of the weight; loss /= weight;
        # loss *= world_size -- every step, before backward
        mean = exchange.step(loss.detach(), float(Bn))
        loss.backward(gradient=exchange.scale())
        return mean

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    # ---- device-resident timing of the plugin call
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    host_issue_s = time.perf_counter() - t0
    sync_all()
    ms_local = e0.elapsed_time(e1)
    print(f"[bench] rank {rank}: {1e3 * ms_local / args.steps:.1f} us/step (host issued a step every "
          f"{1e6 * host_issue_s / args.steps:.1f} us), {int(batch['ilen'].sum())} frames", file=sys.stderr, flush=True)
    ms = max_over_ranks(ms_local)
    loss_value = float(loss.item())                        # mean nll per utterance
    value = world * Bn * args.steps / (ms * 1e-3)

    # ---- end to end: the same plugin call with the LOGITS in pinned host memory.  Every step uploads its logits
    #      (ctc_loss.upload_valid_frames: only the frames t < ilen[b] cross PCIe, the kernels never read the padding),
    #      runs loss_fn + backward and reads the loss back.  Two device buffers: the upload of step i + 1 runs on a
    #      copy stream while step i computes.
    n_e2e = min(args.steps, 100)                           # ~5 ms per step at C2: PCIe bound
    ilen_list = batch["ilen"].tolist()
    x_dev_b = torch.empty_like(x_host, device=dev)
    x_dev_b.copy_(x_host)
    x_dev_b.requires_grad_()
    bufs = [x_dev, x_dev_b]
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    h2d = [0]

    def upload(k):
        with torch.cuda.stream(copy_stream), torch.no_grad():
            copy_stream.wait_event(consumed[k])            # the step that last read this buffer is done
            h2d[0] = CL.upload_valid_frames(x_host, bufs[k], ilen_list)
            copied[k].record(copy_stream)

    def step_on(k):
        main_stream.wait_event(copied[k])
        out = step(bufs[k])
        consumed[k].record(main_stream)
        return out

    for k in (0, 1):
        consumed[k].record(main_stream)
    for w in range(2):                                     # warm up both buffers
        upload(w % 2)
        step_on(w % 2)
    sync_all()
    t0 = time.perf_counter()
    e0.record()
    upload(0)
    for i in range(n_e2e):
        if i + 1 < n_e2e:
            upload((i + 1) % 2)
        ls = step_on(i % 2)
        _ = ls.cpu()                                       # 4-byte device->host read, synchronises
    e1.record()
    sync_all()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    wall_e2e = time.perf_counter() - t0
    clocks = sampler.stop() if sampler else None
    label_bytes = 8 * (int(pool[0][0].numel()) + 3 * Bn)
    e2e = {"value": world * Bn * n_e2e / (ms_e2e * 1e-3), "unit": "utterances/s",
           "h2d_bytes_per_step": (h2d[0] + label_bytes) * world, "d2h_bytes_per_step": 4 * world, "steps": n_e2e,
           "ms_per_step": ms_e2e / n_e2e, "wall_ms_per_step": 1e3 * wall_e2e / n_e2e,
           "h2d_gbs_per_gpu": h2d[0] * n_e2e / (ms_e2e * 1e-3) / 1e9,
           "note": "CTC.loss_fn + backward with pinned host logits -> device (PCIe; valid frames only, the kernels never "
                   "read padded frames) and host labels, loss read back, every step; upload of step i+1 overlaps step i"}
    x_dev_b.grad = None
    del x_dev_b, bufs

    # ---- the same work through the raw C ABI with everything prepared ahead (labels on the device, workspace and
    #      gradient allocated once): what the kernels alone take
    xt = x_dev.detach().transpose(0, 1)
    tg0, olen0 = pool[0]
    off0 = torch.cumsum(olen0, 0) - olen0
    plan = CL._Plan(xt, tg0, off0, batch["ilen"], olen0, LMAX, 0)
    x_dev.grad = None
    grad = CL._empty_like_strided(xt)
    nll_buf = torch.empty(Bn, dtype=torch.float32, device=dev)
    ls_buf = torch.empty(1, dtype=torch.float32, device=dev)
    n_raw = min(args.steps, 300)
    for _ in range(3):
        CL.ctc_loss_fwd_bwd(xt, None, None, None, None, LMAX, grad_scale=1.0 / Bn, grad_out=grad, plan=plan,
                            nll_out=nll_buf, loss_sum_out=ls_buf)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n_raw):
        CL.ctc_loss_fwd_bwd(xt, None, None, None, None, LMAX, grad_scale=1.0 / Bn, grad_out=grad, plan=plan,
                            nll_out=nll_buf, loss_sum_out=ls_buf)
    e1.record()
    torch.cuda.synchronize()
    raw_ms = max_over_ranks(e0.elapsed_time(e1) / n_raw)
    raw_abi = {"value": world * Bn / (raw_ms * 1e-3), "unit": "utterances/s", "ms_per_step": raw_ms, "steps": n_raw,
               "what": "b200_ctc_loss_fwd_bwd through ctypes with a pre-built plan (labels resident, no allocation)"}

    # ---- dominant kernel alone: the one pass over the logits (row statistics + w*softmax written to the
    #      gradient), through the measurement entry point b200_ctc_loss_stream_pass
    st = torch.cuda.current_stream().cuda_stream

    def stream_pass():
        rc = lib.b200_ctc_loss_stream_pass(*plan.common_args(), None, 1.0 / Bn, grad.data_ptr(), grad.stride(0),
                                           grad.stride(1), plan.ws_ptr, plan.ws_bytes, st)
        _lib.check(rc, "b200_ctc_loss_stream_pass")

    n_k = min(args.steps, 200)
    for _ in range(3):
        stream_pass()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n_k):
        stream_pass()
    e1.record()
    torch.cuda.synchronize()
    k_ms = e0.elapsed_time(e1) / n_k
    es = x_host.element_size()
    k_bytes = int(batch["ilen"].sum()) * V * es + Bn * T * V * es   # logits read (valid frames) + grad write
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["stream_kernel_dram_bytes"]
    except Exception:
        pass
    achieved = k_bytes / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "ctc_stream_kernel alone (b200_ctc_loss_stream_pass)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "kernel_ms": k_ms, "algorithmic_bytes": k_bytes, "launches_timed": n_k,
                "frac_of_nominal_8TBs": achieved / NOMINAL_HBM_GBS,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)"}
    step_bytes = 2 * Bn * T * V * es
    step_gbs = step_bytes * args.steps / (ms_local * 1e-3) / 1e9
    raw_gbs = step_bytes / (raw_ms * 1e-3) / 1e9
    roofline_step = {"algorithmic_bytes_per_gpu": step_bytes, "achieved": step_gbs, "frac": step_gbs / peak,
                     "frac_of_nominal_8TBs": step_gbs / NOMINAL_HBM_GBS, "frac_raw_abi": raw_gbs / peak,
                     "note": "whole step (all kernels of the plugin call) against logits read once + gradient written once "
                             "(SURVEY 8d); frac_raw_abi is the same for the pre-planned C-ABI call"}
    del grad, plan

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    cpu_c1 = None
    if world == 1:
        cores = os.cpu_count() or 1
        n_utt = min(16, Bn)
        loss_fn, kind = reference_loss_fn(V)
        cpu_reference_pass(batch, min(4, Bn), cores, loss_fn)
        t_ref, _ = cpu_reference_pass(batch, n_utt, cores, loss_fn)
        try:
            from oracle import ctc_oracle as O

            O.build()
            t_or, thr = cpu_oracle_pass(batch, n_utt)
            cpu = {"value": n_utt / t_or, "unit": "utterances/s", "cores": thr, "kind": "port",
                   "sample": f"{n_utt} of the {Bn} {args.config.upper()} utterances, oracle/ctc_oracle.c (double, OpenMP)",
                   "reference_cpu": {"value": n_utt / t_ref, "cores": cores, "kind": kind,
                                     "what": "espnet2 CTC.loss_fn (builtin) fwd+bwd on the host cores"}}
        except Exception as e:  # the oracle is only a reported baseline
            cpu = {"value": n_utt / t_ref, "unit": "utterances/s", "cores": cores, "kind": kind,
                   "sample": f"{n_utt} {args.config.upper()} utterances, espnet2 CTC.loss_fn on CPU (oracle build failed: {e})"}
        try:
            cpu_c1 = cpu_c1_reference(cores)
        except Exception as e:
            cpu_c1 = {"error": str(e)[:200]}
    prefix = prefix15 = decode_loop = c5 = None
    ref_gpu = None
    if world == 1:
        x_dev.grad = None
        torch.cuda.empty_cache()
        try:
            ref_gpu = gpu_reference_pass(batch, dev)
            ref_gpu["speedup_device_resident"] = value / ref_gpu["value"]
        except Exception as e:  # reported baseline only
            ref_gpu = {"error": str(e)[:200]}
        torch.cuda.empty_cache()
    if world == 1 and not args.no_prefix and args.config == "c2":
        del x_dev
        torch.cuda.empty_cache()
        prefix = bench_prefix(dev, peak)
        prefix15 = bench_prefix(dev, peak, snum=15)
        try:
            c5 = bench_c5_leg(dev, peak)
        except Exception as e:  # a reported extra, never the reason the line is missing
            c5 = {"error": str(e)[:200]}
        try:
            decode_loop = [bench_decode_loop(dev), bench_decode_loop(dev, margin=40), bench_decode_loop(dev, pre_beam=15)]
        except Exception as e:  # a reported extra, never the reason the line is missing
            decode_loop = {"error": str(e)[:200]}
    kernels = ["ctc_label_fetch_kernel (labels from the pinned staging buffer)", "ctc_stream_kernel",
               "ctc_prep_kernel (under the stream kernel)", "ctc_sweep kernel (alpha+beta)",
               "ctc_occ_kernel<PATCH>", "ctc_grad_rescale_kernel (backward: verifies the upstream scale)"]
    line = {"metric": "ctc_loss_fwd_bwd_utterances_per_sec", "value": value, "unit": "utterances/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 logits, f32 lattice",
            "data": "synthetic", "config": workload_config(args, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": (len(kernels) + (1 if world > 1 else 0)) * args.steps,
            "kernels_per_step": kernels + (["peer_sum_exchange_kernel (N > 1)"] if world > 1 else []),
            "host_issue_us_per_step": 1e6 * host_issue_s / args.steps, "cpu_affinity": affinity,
            "roofline": roofline, "roofline_step": roofline_step, "raw_abi": raw_abi,
            "cpu_baseline": cpu, "cpu_reference_c1": cpu_c1, "reference_gpu": ref_gpu, "loss": loss_value,
            "prefix": prefix, "prefix_pre_beam": prefix15, "decode_loop": decode_loop, "c5": c5}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--no-prefix", action="store_true", help="skip the prefix-scoring (C3) measurement")
    ap.add_argument("--full-length", action="store_true", help="every utterance uses all T frames")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
