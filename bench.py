#!/usr/bin/env python
"""bench.py -- CTC loss fwd+bwd utterances/s on B200 (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload: BASELINE.json configs[1] (C2: B=64 per GPU, T=500, V=5000, L<=150, bf16 logits,
synthetic, SURVEY.md 8d input recipe).  With N GPUs every rank takes its own 64 utterances
(weak scaling; N=8 is config C4, global B=512) with no data-path collective; the per-step
loss sums are exchanged in one NCCL all-reduce per LEDGER_STEPS steps (sharding.LossLedger), inside
the timed region.  A "step" is one fused log_softmax + CTC forward + backward over the rank's batch.
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

B_PER_GPU, T, V, LMAX = 64, 500, 5000, 150
LEDGER_STEPS = 50     # N > 1: per-step loss sums are exchanged in one all-reduce per 50 steps (sharding.LossLedger)
KERNELS_PER_STEP = 4  # prep, stream (one pass over the logits), sweep (alpha+beta), patch


def make_batch(seed, full_length=False, dtype=torch.bfloat16):
    from tests import _cases

    return _cases.make_loss_case(B_PER_GPU, T, V, LMAX, seed=seed, dtype=dtype, full=full_length)


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


def cpu_reference_pass(batch, n_utt, threads):
    """The reference's CPU path for this metric: CTC.loss_fn (espnet2/asr/ctc.py:46-52) =
    log_softmax(2) + torch.nn.CTCLoss(sum) / B, forward and backward, on the host cores."""
    torch.set_num_threads(threads)
    x = batch["x"][:n_utt].float()
    ilen, olen, ys = batch["ilen"][:n_utt], batch["olen"][:n_utt], batch["ys"][:n_utt]
    tg = torch.cat([ys[i, : olen[i]] for i in range(n_utt)])
    crit = torch.nn.CTCLoss(reduction="sum")
    xt = x.transpose(0, 1).detach().requires_grad_()
    t0 = time.perf_counter()
    loss = crit(xt.log_softmax(2), tg, ilen, olen) / xt.size(1)
    loss.backward()
    return time.perf_counter() - t0, float(loss.detach())


def gpu_reference_pass(batch, dev, steps=10):
    """SURVEY 8d "reference GPU baseline on the same box": what CTC.loss_fn (espnet2/asr/ctc.py:46-52) executes
    on a CUDA device -- .float() of the bf16 logits (autocast's promotion), log_softmax(2),
    torch.nn.CTCLoss(sum) / B, backward to the fp32 logits -- timed with CUDA events.  Stock PyTorch only."""
    x = batch["x"].to(dev)
    ilen, olen = batch["ilen"].to(dev), batch["olen"].to(dev)
    tg = torch.cat([batch["ys"][i, : batch["olen"][i]] for i in range(x.size(0))]).to(dev)
    crit = torch.nn.CTCLoss(reduction="sum")
    xt = x.transpose(0, 1)

    def one():
        xf = xt.float().requires_grad_()
        loss = crit(xf.log_softmax(2), tg, ilen, olen) / xf.size(1)
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
            "loss": float(loss.detach()), "kind": "stock torch on the same GPU",
            "what": "bf16 logits .float() -> log_softmax(2) -> torch.nn.CTCLoss(sum)/B -> backward "
                    "(espnet2/asr/ctc.py:46-52 as it runs under autocast)"}


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


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    batch = make_batch(1236, args.full_length)
    n_utt = 16  # bounded sample of the C2 batch per step
    t16, _ = cpu_reference_pass(batch, n_utt, cores)
    # keep the whole run within about two minutes whatever --steps is
    n_utt = max(1, min(16, int(16 * 120.0 / max(args.steps * t16, 1e-9))))
    for _ in range(args.warmup):
        cpu_reference_pass(batch, n_utt, cores)
    t = 0.0
    for _ in range(args.steps):
        dt, _ = cpu_reference_pass(batch, n_utt, cores)
        t += dt
    value = n_utt * args.steps / t
    sample = f"{n_utt} of the {B_PER_GPU} C2 utterances per step, fp32 copy of the bf16 logits, torch CPU"
    line = {"impl": "reference", "metric": "ctc_loss_fwd_bwd_utterances_per_sec", "value": value,
            "unit": "utterances/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, 1),
            "cpu_baseline": {"value": value, "unit": "utterances/s", "cores": cores, "kind": "port",
                             "sample": sample + " (espnet2/asr/ctc.py:46-52 restated: log_softmax + "
                                       "torch.nn.CTCLoss(sum)/B + backward; the reference is Python and "
                                       "cannot travel to the box)"},
            "e2e": {"value": value, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {"workload": "C2: LibriSpeech-shaped CTC loss fwd+bwd, B=64 per GPU, T=500, V=5000 BPE, L<=150, "
                        "bf16 logits" + (", all utterances full length" if args.full_length else
                                         ", ilen~U[0.6T,T] (SURVEY 8d)"),
            "global_batch": B_PER_GPU * world, "frames": T, "vocab": V, "max_labels": LMAX,
            "sharding": f"utterances x{world}, no data-path collective; loss sums all-reduced once per {LEDGER_STEPS} steps" if world > 1 else "single GPU",
            "l2_policy": "inputs larger than L2 (320 MB logits + 320 MB gradient per step vs 126 MB L2)"}


def bench_prefix(dev, steps=20):
    """Second half of the metric: CTC prefix-score steps/s at BASELINE config 3
    (B=32 utterances x beam 10, T=500, V=5000, fp32 log-probs, full-vocabulary scoring as
    recognize_beam_batch does on CUDA).  One step = __call__ + index_select_state for all 320
    hypotheses; hypotheses advance by a seeded random walk."""
    from espnet_b200.ctc_prefix_score import CTCPrefixScoreB200
    from oracle.prefix_th_port import PrefixScoreEager

    B, H, T_, V_ = 32, 10, 500, 5000
    eos = V_ - 1
    g = torch.Generator(device=dev).manual_seed(1237)
    lp = torch.log_softmax(torch.randn(B, T_, V_, generator=g, device=dev), dim=-1)
    gc = torch.Generator().manual_seed(1237)
    xlens = torch.randint(300, T_ + 1, (B,), generator=gc)
    xlens[0] = T_
    nwalk = steps + 3
    hyp = torch.randint(0, H, (nwalk, B, H), generator=gc)
    tok = torch.randint(1, V_ - 1, (nwalk, B, H), generator=gc)
    best = (hyp * V_ + tok).to(dev)
    parents = (hyp + torch.arange(B).view(1, -1, 1) * H).view(nwalk, -1).tolist()
    toks = tok.view(nwalk, -1).tolist()

    # The prefixes of every step are inputs of the walk: they are laid out before the timed region, as
    # (n_bh, len) device tensors for the b200 scorer and as last-token lists for the eager port.
    ys, y = [], [[eos] for _ in range(B * H)]
    for i in range(nwalk):
        ys.append(y)
        y = [y[parents[i][k]] + [toks[i][k]] for k in range(B * H)]
    y_dev = [torch.tensor(yi, dtype=torch.int64, device=dev) for yi in ys]
    y_last = [[yk[-1] for yk in yi] for yi in ys]

    def walk(make_impl, do_step, do_select, n):
        impl = make_impl()
        state = None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for i in range(n + 2):
            if i == 2:  # steps 0 and 1 (warm-up: first label, first step with a previous state) are not timed
                torch.cuda.synchronize()
                e0.record()
            st = do_step(impl, i, state)
            state = do_select(impl, st, best[i])
        e1.record()
        torch.cuda.synchronize()
        return n / (e0.elapsed_time(e1) * 1e-3)

    ours = walk(lambda: CTCPrefixScoreB200(lp.clone(), xlens, 0, eos),
                lambda m, i, s: m(y_dev[i], s)[1], lambda m, st, b: m.index_select_state(st, b), steps)
    out = {"metric": "ctc_prefix_score_steps_per_sec", "value": ours, "unit": "steps/s",
           "config": {"workload": "C3: B=32 utterances x beam 10 (n_bh=320), T=500, V=5000, fp32 log-probs, "
                                  "scoring_ids=None, margin=0", "steps": steps},
           "algorithmic_bytes_per_step": 4 * (B * T_ * V_ + 2 * B * H * T_ * 2 + B * H * V_)}
    try:
        ref = walk(lambda: PrefixScoreEager(lp.clone(), xlens, 0, eos),
                   lambda m, i, s: m.step(y_last[i], i, s)[1],
                   lambda m, st, b: m.select(st, b), 3)
        out["reference_port_gpu"] = {"value": ref, "unit": "steps/s", "kind": "port",
                                     "what": "eager-PyTorch restatement of CTCPrefixScoreTH (oracle/prefix_th_port.py) "
                                             "on the same GPU, 3 steps"}
        out["speedup_vs_reference_port_gpu"] = ours / ref
    except Exception as e:  # e.g. out of memory on a shared box
        out["reference_port_gpu"] = {"error": str(e)[:200]}
    return out


def run_b200(args, rank, world, local_rank):
    import torch.distributed as dist

    from espnet_b200 import _lib
    from espnet_b200 import ctc_loss as CL

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
    batch = make_batch(1236 + rank, args.full_length)
    x_host = batch["x"].pin_memory()                      # (B,T,V) bf16, what ctc_lo produces
    x_dev = torch.empty_like(x_host, device=dev)
    x_dev.copy_(x_host)
    xt = x_dev.transpose(0, 1)                            # the (T,B,V) view CTC.forward hands to loss_fn
    ys = batch["ys"].to(dev).reshape(-1)
    off = torch.arange(B_PER_GPU) * batch["ys"].shape[1]
    plan = CL._Plan(xt, ys, off, batch["ilen"], batch["olen"], LMAX, 0)
    grad = CL._empty_like_strided(xt)
    nll_buf = torch.empty(B_PER_GPU, dtype=torch.float32, device=dev)
    from espnet_b200.sharding import LossLedger

    # Per-step loss sums land in a ledger on the device; with N > 1 ranks the ledger is exchanged
    # in ONE all-reduce per LEDGER_STEPS steps (the reference's log interval), inside the timed region.
    ledger = LossLedger(LEDGER_STEPS, B_PER_GPU, dev)
    gscale = 1.0 / (B_PER_GPU * world)
    nstep = [0]
    last = [None]

    def step():
        i = nstep[0]
        nll, ls, _ = CL.ctc_loss_fwd_bwd(xt, None, None, None, None, LMAX, grad_scale=gscale, grad_out=grad,
                                         plan=plan, nll_out=nll_buf, loss_sum_out=ledger.slot(i))
        nstep[0] = i + 1
        if world > 1 and (i + 1) % LEDGER_STEPS == 0:
            last[0] = ledger.reduce()[0][-1:]              # global mean loss of the newest step
            return last[0]
        return ls

    def join():
        """Flush a partly filled ledger (the tail of a run whose length is not a multiple of the interval)."""
        if world > 1 and nstep[0] % LEDGER_STEPS:
            k = nstep[0] % LEDGER_STEPS
            last[0] = ledger.reduce()[0][k - 1:k]
        return last[0]

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
    if world > 1:
        ledger.reduce()                                     # the exchange itself is warmed up too
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    # ---- device-resident timing
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    if world > 1:
        loss = join()
    e1.record()
    sync_all()
    ms_local = e0.elapsed_time(e1)
    print(f"[bench] rank {rank}: {1e3 * ms_local / args.steps:.1f} us/step, "
          f"{int(batch['ilen'].sum())} frames", file=sys.stderr, flush=True)
    ms = max_over_ranks(ms_local)
    loss_value = float(loss.item()) / (B_PER_GPU if world == 1 else 1)   # mean nll per utterance
    value = world * B_PER_GPU * args.steps / (ms * 1e-3)

    # ---- end to end through the public call with HOST buffers (pinned): every step uploads its logits
    #      (ctc_loss.upload_valid_frames: only the frames t < ilen[b] cross PCIe, the kernels never read the
    #      padding), runs the fused fwd+bwd and reads the loss back.  Two device buffers: the upload of step
    #      i + 1 runs on a copy stream while step i computes.
    n_e2e = min(args.steps, 100)                           # ~5 ms per step: PCIe bound
    ilen_list = batch["ilen"].tolist()
    x_dev_b = torch.empty_like(x_dev)
    x_dev_b.copy_(x_dev)
    bufs = [x_dev, x_dev_b]
    plans = [plan, CL._Plan(x_dev_b.transpose(0, 1), ys, off, batch["ilen"], batch["olen"], LMAX, 0)]
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream(dev)
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    h2d = [0]

    def upload(k):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[k])            # the step that last read this buffer is done
            h2d[0] = CL.upload_valid_frames(x_host, bufs[k], ilen_list)
            copied[k].record(copy_stream)

    def step_on(k):
        i = nstep[0]
        main_stream.wait_event(copied[k])
        _, ls, _ = CL.ctc_loss_fwd_bwd(bufs[k].transpose(0, 1), None, None, None, None, LMAX, grad_scale=gscale,
                                       grad_out=grad, plan=plans[k], nll_out=nll_buf, loss_sum_out=ledger.slot(i))
        consumed[k].record(main_stream)
        nstep[0] = i + 1
        if world > 1 and (i + 1) % LEDGER_STEPS == 0:
            return ledger.reduce()[0][-1:]
        return ls

    for k in (0, 1):
        consumed[k].record(main_stream)
    for w in range(2):                                     # warm up both buffers / plans
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
    e2e = {"value": world * B_PER_GPU * n_e2e / (ms_e2e * 1e-3), "unit": "utterances/s",
           "h2d_bytes_per_step": h2d[0] * world, "d2h_bytes_per_step": 4 * world, "steps": n_e2e,
           "ms_per_step": ms_e2e / n_e2e, "wall_ms_per_step": 1e3 * wall_e2e / n_e2e,
           "h2d_gbs_per_gpu": h2d[0] * n_e2e / (ms_e2e * 1e-3) / 1e9,
           "note": "pinned host logits -> device (PCIe; valid frames only, the kernels never read padded frames), "
                   "fused fwd+bwd, loss read back, every step; upload of step i+1 overlaps step i (two device buffers)"}
    del x_dev_b, plans

    # ---- dominant kernel alone: the one pass over the logits (row statistics + w*softmax written to the
    #      gradient), through the measurement entry point b200_ctc_loss_stream_pass
    st = torch.cuda.current_stream().cuda_stream

    def stream_pass():
        rc = lib.b200_ctc_loss_stream_pass(*plan.common_args(), None, gscale, grad.data_ptr(), grad.stride(0),
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
    k_bytes = int(batch["ilen"].sum()) * V * es + B_PER_GPU * T * V * es   # logits read (valid frames) + grad write
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
    roofline = {"bound": "hbm", "kernel": "ctc_stream_kernel<bf16,2> alone (b200_ctc_loss_stream_pass)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "kernel_ms": k_ms, "algorithmic_bytes": k_bytes, "launches_timed": n_k,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)"}
    step_bytes = 2 * B_PER_GPU * T * V * es
    step_gbs = step_bytes * args.steps / (ms * 1e-3) / 1e9
    roofline_step = {"algorithmic_bytes_per_gpu": step_bytes, "achieved": step_gbs, "frac": step_gbs / peak,
                     "note": "whole step (all kernels) against logits read once + gradient written once"}

    if rank != 0:
        return
    cpu = None
    if world == 1:
        cores = os.cpu_count() or 1
        n_utt = 16
        cpu_reference_pass(batch, 4, cores)
        t_ref, _ = cpu_reference_pass(batch, n_utt, cores)
        try:
            from oracle import ctc_oracle as O

            O.build()
            t_or, thr = cpu_oracle_pass(batch, n_utt)
            cpu = {"value": n_utt / t_or, "unit": "utterances/s", "cores": thr, "kind": "port",
                   "sample": f"{n_utt} of the {B_PER_GPU} C2 utterances, oracle/ctc_oracle.c (double, OpenMP)",
                   "torch_cpu": {"value": n_utt / t_ref, "cores": cores,
                                 "what": "log_softmax + torch.nn.CTCLoss fwd+bwd (the reference's CPU path)"}}
        except Exception as e:  # the oracle is only a reported baseline
            cpu = {"value": n_utt / t_ref, "unit": "utterances/s", "cores": cores, "kind": "port",
                   "sample": f"{n_utt} C2 utterances, torch CPU ctc_loss (oracle build failed: {e})"}
    prefix = None
    ref_gpu = None
    if world == 1:
        del grad
        torch.cuda.empty_cache()
        try:
            ref_gpu = gpu_reference_pass(batch, dev)
            ref_gpu["speedup_device_resident"] = value / ref_gpu["value"]
        except Exception as e:  # reported baseline only
            ref_gpu = {"error": str(e)[:200]}
        torch.cuda.empty_cache()
    if world == 1 and not args.no_prefix:
        prefix = bench_prefix(dev)
    line = {"metric": "ctc_loss_fwd_bwd_utterances_per_sec", "value": value, "unit": "utterances/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 logits, f32 lattice",
            "data": "synthetic", "config": workload_config(args, world), "clocks": clocks, "e2e": e2e,
            "gpu_launches": KERNELS_PER_STEP * args.steps,
            "kernels_per_step": ["ctc_stream_kernel", "ctc_prep_kernel (under the stream kernel)", "ctc_sweep_staged_kernel", "ctc_occ_kernel<PATCH>"], "roofline": roofline, "roofline_step": roofline_step,
            "cpu_baseline": cpu, "reference_gpu": ref_gpu, "loss": loss_value, "prefix": prefix}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
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
