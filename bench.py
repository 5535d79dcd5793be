#!/usr/bin/env python
"""bench.py - DeepSpeech2 training hot path (fwd + bwd + CTC) on N B200s of one node.

Contract: `python bench.py --gpus N --steps K --warmup W [--impl reference]`; for N>1 launched under
torch.distributed.run, one rank per GPU.  Rank 0 prints ONE JSON line.

workload  : the reference's dummy=1 training step (train.sh config: B=32/GPU, 7 bi-RNN layers, H=1760,
            32 conv filters) on its longest bucket T=1500 frames (15 s utterances, 480 audio-s / GPU / step),
            'stacked' wiring (the intended topology), bf16 tensor-core path.
value     : audio-seconds trained per second, whole job, inputs resident in HBM.
e2e       : same metric through the public API (deepSpeech_NCHW.inference/loss) with a fresh host batch
            every step: pinned-host -> device copy of the features + device -> host read of the loss inside
            the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "audio_sec_per_sec_trained_fwd_bwd_ctc"
UNIT = "audio-s/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32, help="utterances per GPU (train.sh: 32)")
    ap.add_argument("--frames", type=int, default=1500, help="input frames per utterance (dummy bucket)")
    ap.add_argument("--hidden", type=int, default=1760)
    ap.add_argument("--layers", type=int, default=7)
    ap.add_argument("--cell", default="relu_clip", choices=["relu_clip", "gru"])
    ap.add_argument("--fp32", action="store_true", help="fp32 CUDA-core path instead of bf16 tensor cores")
    ap.add_argument("--no-extras", action="store_true", help="skip CTC sweep / kernel roofline / cpu baseline")
    return ap.parse_args()


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler(object):
    """nvidia-smi in loop mode (-lms 100) for the duration of the `with` block; summary() gives the median SM clock
    and any throttle reason that was ever active."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            time.sleep(0.35)                       # let the first samples arrive before the timed region starts
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc is None:
            return
        try:
            time.sleep(0.15)
            self.proc.terminate()
            out, _ = self.proc.communicate(timeout=10)
            for line in out.strip().splitlines():
                f = [x.strip() for x in line.split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
        except Exception:
            pass

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        isnum = lambda x: x.replace(".", "", 1).isdigit()
        mhz = sorted(float(s[0]) for s in self.samples if isnum(s[0]))
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": float(self.samples[0][1]) if isnum(self.samples[0][1]) else None,
                "reasons": reasons, "samples": len(self.samples)}


# ----------------------------------------------------------------------------- FLOP model
def step_flops(B, T, H, L, nf=32, cell="relu_clip"):
    T1 = (T - 20) // 2 + 1
    T2 = (T1 - 10) // 2 + 1
    G = 3 if cell == "gru" else 1
    conv = 2.0 * B * nf * T1 * 79 * 100 + 2.0 * B * nf * T2 * 75 * 1600
    rnn = 0.0
    for l in range(L):
        I = 75 * nf if l == 0 else H
        rnn += 2 * 2.0 * T2 * B * G * H * (I + H)
    fc = 2.0 * T2 * B * H * 29
    return 3.0 * (conv + rnn + fc), 3.0 * rnn, T2


def _pick_cpu_bucket(timer, batch, n_steps, budget_s):
    """Largest dummy bucket whose n_steps steps fit budget_s seconds.  Step time grows faster than linearly with the
    bucket length on some hosts (measured: 1.1 s at T=100, 7.6 s at T=400, 131 s at T=1500 on 8 cores), so the buckets
    are probed in increasing order with one real step each, and a longer one is only tried when a quadratic
    extrapolation of the last measurement still fits.  Parameter construction is not part of any probe."""
    t0 = time.perf_counter()
    timer.step(batch, 0)                               # warm-up (allocator, thread pool)
    chosen, T_c, t_c = 0, 100, timer.step(batch, 0)[0]
    for b, T in ((1, 200), (3, 400), (7, 800), (14, 1500)):
        pred = t_c * (T / float(T_c)) ** 2
        spent = time.perf_counter() - t0
        if pred * (n_steps + 1) > budget_s or spent + pred > budget_s:
            break
        t = timer.step(batch, b)[0]
        if t * n_steps > budget_s:
            break
        chosen, T_c, t_c = b, T, t
    return chosen, T_c, t_c


def reference_arm(args, rank):
    """The reference's own TF-CPU path is unavailable (no tensorflow, py2 source): time the oracle's fp32
    torch-CPU port of the same step on this box's host cores (kind 'port'), on the bench workload's own bucket when the
    K + W steps fit ~4 minutes, else on the largest bucket that does (stated in the line)."""
    if rank != 0:
        return
    from oracle import ds2_cpu_baseline as cb
    threads = os.cpu_count() or 1
    timer = cb.StepTimer(threads, args.hidden, args.layers)
    bucket, frames, est100 = _pick_cpu_bucket(timer, args.batch, args.steps + args.warmup, 240.0)
    r = timer.run(args.batch, bucket, args.steps, args.warmup)
    same = frames == args.frames
    sample = "dummy bucket T=%d frames x B=%d (%.0f audio-s per step), %d warm-up + %d timed steps, fp32 torch-CPU port%s" % (
        frames, args.batch, r["audio_s_per_step"], args.warmup, args.steps,
        "" if same else " (the T=%d bucket does not fit the time budget on this host: one step of this bucket takes %.1f s)" % (args.frames, est100))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "DS2 dummy-data training step (fwd+bwd+CTC), train.sh config: B=%d/GPU, %d-layer bi-RNN(clipped ReLU) H=%d, "
                                   "32 conv filters, bucket T=%d frames, stacked wiring, no optimizer; CPU sample: %s"
                                   % (args.batch, args.layers, args.hidden, frames, sample),
                       "global_batch": args.batch, "frames": frames, "same_bucket_as_gpu_arm": same},
            "step_ms": {"min": r["ms_min"], "median": r["ms_median"], "max": r["ms_max"]},
            "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference TF-1.x CPU path unavailable on this box (no tensorflow; py2-only source): oracle port timed instead"}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # the all-reduce kernels run next to the latency-bound cooperative recurrence: keep them to a few CTAs so that they
        # neither delay the placement of its clusters nor crowd its L2 traffic (NVLink is far from saturated by 356 MB a step)
        os.environ.setdefault("NCCL_MAX_CTAS", "8")
        dist.init_process_group("nccl", device_id=dev)

    from deepspeech_b200 import _lib, deepSpeech_NCHW as ds, ops
    from deepspeech_b200.deepSpeech_dummy import DummyData, utt_lengths
    from deepspeech_b200.dp import GradAllReducer

    B, T, H, L = args.batch, args.frames, args.hidden, args.layers
    params = SimpleNamespace(use_fp16=not args.fp32, num_filters=32, batch_size=B, num_hidden=H, num_rnn_layers=L, engine="mkl",
                             rnn_type="bi-dir", wiring="stacked", cell=args.cell, seed=1234)
    model = ds.model(params)
    reducer = GradAllReducer(model.parameters(), prescaled=True) if world > 1 else None     # loss / world is back-propagated
    bucket = utt_lengths.index(T) if T in utt_lengths else None
    data = DummyData(B, seed=1234 + rank)
    stress_pool = []
    stress_rng = np.random.default_rng(4321 + rank)          # non-bucket length (BASELINE config 4(ii): T2 = 1500 recurrent steps
    if bucket is None and T < utt_lengths[-1]:               # = 6034 frames): standard-normal features, bucket-14 labels
        raise SystemExit("--frames must be one of the dummy buckets %s or longer than the longest" % utt_lengths)

    pinned = [torch.empty((B, T, 161), dtype=torch.float32).pin_memory() for _ in range(2)]
    feats_dev = torch.empty((B, T, 161), dtype=torch.float32, device=dev)

    def host_batch(slot):
        if bucket is None:
            _, idx, vals, shape, _ = data.inputs(len(utt_lengths) - 1)
            # like the reference's generator (src/deepSpeech_dummy.py:32-52): a pool drawn once, a different slice per step
            if not stress_pool:
                stress_pool.append(stress_rng.standard_normal((B * T + 4096, 161), dtype=np.float32))
            off = int(stress_rng.integers(0, 4096))
            pinned[slot].copy_(torch.from_numpy(stress_pool[0][off:off + B * T].reshape(B, T, 161)))
            return (idx, vals, shape), np.full(B, T, dtype=np.int32)
        feats, idx, vals, shape, lens = data.inputs(bucket)
        pinned[slot].copy_(torch.from_numpy(feats))
        return (idx, vals, shape), lens

    def train_step(sparse, lens, feats=None):
        model.zero_grad(set_to_none=True)
        logits = ds.inference(None, feats_dev if feats is None else feats, lens, params)
        loss = ds.loss(logits, sparse, lens)
        if reducer is not None:
            (loss * reducer.loss_scale).backward()
            reducer.finish()
        else:
            loss.backward()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- resident-input timing (value)
    sparse, lens = host_batch(0)
    feats_dev.copy_(pinned[0], non_blocking=True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        for _ in range(max(args.warmup, 3)):
            loss = train_step(sparse, lens)
        barrier()
        launches0 = _lib.lib.ds2_launch_count()
        barrier()
        ev0.record()
        for _ in range(args.steps):
            loss = train_step(sparse, lens)
        ev1.record()
        barrier()
    launches = _lib.lib.ds2_launch_count() - launches0
    ms_step = max_over_ranks(ev0.elapsed_time(ev1) / args.steps)
    final_loss = float(loss.item())
    audio_s_step = world * B * T / 100.0
    value = audio_s_step / (ms_step * 1e-3)

    # ---------------- end-to-end timing (host batch -> H2D -> step -> loss D2H), double-buffered host prep
    # Every step's batch is copied host -> device INSIDE the timed region (the first one included); the copy of batch
    # i+1 runs on a copy stream into the other device buffer while step i computes (the prefetch any input pipeline does).
    dev_bufs = [feats_dev, torch.empty_like(feats_dev)]
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream()
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    nxt = host_batch(0)
    barrier()
    t_e0, t_e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_e0.record()
    wall0 = time.perf_counter()
    with torch.cuda.stream(copy_stream):
        dev_bufs[0].copy_(pinned[0], non_blocking=True)
        copied[0].record()
    for i in range(args.steps):
        cur, slot = nxt, i % 2
        main_stream.wait_event(copied[slot])
        loss = train_step(*cur, feats=dev_bufs[slot])
        consumed[slot].record()
        if i + 1 < args.steps:
            nslot = (i + 1) % 2
            nxt = host_batch(nslot)                          # host work overlaps with the GPU step
            with torch.cuda.stream(copy_stream):
                if i >= 1:
                    copy_stream.wait_event(consumed[nslot])  # step i-1 has finished reading that buffer
                dev_bufs[nslot].copy_(pinned[nslot], non_blocking=True)
                copied[nslot].record()
        lv = loss.item()                                     # device -> host read of the result
    t_e1.record()
    barrier()
    e2e_ms = max_over_ranks(max(t_e0.elapsed_time(t_e1), (time.perf_counter() - wall0) * 1e3) / args.steps)
    e2e_value = audio_s_step / (e2e_ms * 1e-3)

    total_flops, rnn_flops, T2 = step_flops(B, T, H, L, 32, args.cell)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    # the roofline kernel is timed ALONE (a loop of launches of that one kernel, CUDA events): the burst figure applies
    peak_tf = float(peaks.get("bf16_tflops", 1600.0))
    peak_src = "measured (MEASURED_PEAKS.json, burst: the kernel is timed alone)" if peaks else "fallback (B200_PROFILING.md)"
    peak_hbm = float(peaks.get("hbm_gbs", 6650.0))

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.fp32 else "bf16", "data": "synthetic",
        "config": {"workload": "DS2 dummy-data training step (fwd+bwd+CTC), train.sh config: B=%d/GPU, %d-layer bi-%s H=%d, 32 conv filters, "
                               "bucket T=%d frames (T2=%d RNN steps), stacked wiring, no optimizer (metric is fwd+bwd+CTC)"
                               % (B, L, "GRU" if args.cell == "gru" else "RNN(clipped ReLU)", H, T, T2),
                   "global_batch": B * world, "frames": T, "parallelism": "dp%d" % world,
                   "l2": "per-step working set (weights %.0f MB fp32 + activations > 1 GB) exceeds the 126 MB L2; no explicit flush"
                         % (sum(p.numel() for p in model.parameters()) * 4 / 1e6)},
        "clocks": clk.summary(),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(B * T * 161 * 4), "d2h_bytes_per_step": 4,
                "ms_per_step": e2e_ms},
        "gpu_launches": int(launches),
        "dp": ({"all_reduces": reducer.launched, "gradients_copied_after_reduce": reducer.copied} if reducer is not None else None),
        "loss": final_loss,
        "step_tflops": total_flops / (ms_step * 1e-3) / 1e12 / 1.0,
        # whole step against the SUSTAINED bf16 peak (the step is long; the burst figure applies to a kernel timed alone)
        "roofline_step": {"bound": "tensor", "achieved": total_flops / (ms_step * 1e-3) / 1e12, "unit": "TFLOP/s",
                          "peak": float(peaks.get("bf16_tflops_sustained", 1377.0)),
                          "frac": total_flops / (ms_step * 1e-3) / 1e12 / float(peaks.get("bf16_tflops_sustained", 1377.0)),
                          "flops_per_step": total_flops},
        "model_flops_per_step": total_flops,
    }

    if rank == 0 and not args.no_extras:
        try:
            line.update(extras(args, dev, model, ops, _lib, B, T2, H, L, peak_tf, peak_hbm, peak_src, ms_step))
        except Exception as e:                                # never lose the headline line
            line["extras_error"] = repr(e)[:300]
    if world > 1:
        dist.barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _profiled_traffic(dom_gemm, T2, B, H):
    if dom_gemm or (T2, B, H) != (366, 32, 1760):
        return None
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "r2_rnn2_ncu.json")))["dram_bytes_per_launch"])
    except Exception:
        return None


def _time_kernel(fn, iters=10, warm=3):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def extras(args, dev, model, ops, _lib, B, T2, H, L, peak_tf, peak_hbm, peak_src, ms_step):
    """Kernel-level numbers next to the headline: recurrent-stack TFLOP/s, the dominant kernel's roofline
    (measured live with CUDA events on the launching stream), the CTC sweep (config 2) and the CPU baseline."""
    import torch
    out = {}
    x = torch.randn((T2, B, H), device=dev).to(torch.bfloat16 if not args.fp32 else torch.float32)
    lens = torch.full((B,), T2, dtype=torch.int32, device=dev)
    lyr = 1 if L > 1 else 0
    W, U, Bv = model.rnn_W[lyr].detach(), model.rnn_U[lyr].detach(), model.rnn_B[lyr].detach()
    I = W.shape[1]
    if args.cell == "relu_clip" and I == H:
        cdt = ops.rnn_path_dtype(B, H, 2, not args.fp32)
        Wg, Ug = ops.cast(W, cdt), ops.cast(U, cdt)
        xg = ops.cast(x, cdt)
        # (1) batched input projection GEMM
        ms_gemm = _time_kernel(lambda: ops.gemm_nt(xg.view(T2 * B, H), Wg))
        gemm_flops = 2.0 * T2 * B * H * 2 * H
        # (2) the recurrence (persistent kernel on the bf16 path)
        wx = ops.gemm_nt(xg.view(T2 * B, H), Wg)
        hs = torch.empty((2, T2, B, H), dtype=cdt, device=dev)
        wsb = _lib.lib.ds2_rnn_workspace_bytes(T2, B, H, 2, ops._dt(hs))
        ws = ops._workspace(wsb, dev, "rnn") if wsb else None

        def rec():
            ops.check(_lib.lib.ds2_rnn_relu_fwd(ops._p(wx), ops.SBN_SCALE, ops._p(Ug), ops._p(Bv), ops._p(lens), ops._p(hs), ops._dt(hs),
                                                T2, B, H, 2, 2, 20.0, ops._p(ws), wsb, ops._stream()), "rnn_fwd")
        ms_rec = _time_kernel(rec, iters=5, warm=2)
        rec_flops = 2 * 2.0 * T2 * B * H * H
        out["rnn_layer_fwd"] = {"gemm_ms": ms_gemm, "gemm_tflops": gemm_flops / ms_gemm / 1e9, "recurrence_ms": ms_rec,
                                "recurrence_tflops": rec_flops / ms_rec / 1e9, "us_per_timestep": ms_rec * 1e3 / T2,
                                "path": "persistent tcgen05 (bf16)" if cdt == torch.bfloat16 else "per-step CUDA-core kernels (fp32)"}
        dom_gemm = ms_gemm >= ms_rec
        ach = (gemm_flops / ms_gemm if dom_gemm else rec_flops / ms_rec) / 1e9
        out["roofline"] = {"kernel": "gemm_tcgen05_kernel (W.x, M=%d N=%d K=%d)" % (T2 * B, 2 * H, H) if dom_gemm
                           else "recurrent forward kernel (T=%d steps, B=%d, H=%d, 2 directions)" % (T2, B, H),
                           "bound": "tensor", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach / peak_tf,
                           # dram__bytes_read+write of one launch of this kernel, read from the committed ncu --set full
                           # extract of the same shape (profiles/r2_rnn2_ncu.json); null for any other shape / kernel
                           "traffic": _profiled_traffic(dom_gemm, T2, B, H), "traffic_source": "profiles/r2_rnn2_ncu.json (extract: r2_rnn2_final_ncu.txt)",
                           "algorithmic_bytes": 4.0 * T2 * B * 2 * H + 2.0 * 2 * H * H + 2.0 * 2 * T2 * B * H,
                           "peak_source": peak_src,
                           "note": "latency-bound T-step dependency chain (4 chains x %d steps); stage cycles in profiles/r2_rnn_stage_cycles_v3.txt" % T2}
    total_flops, rnn_flops, _ = step_flops(B, args.frames, H, L, 32, args.cell)
    out["rnn_tflops_in_step"] = {"note": "all RNN-stack GEMM FLOPs (fwd+bwd) / whole step time (lower bound)",
                                 "value": rnn_flops / (ms_step * 1e-3) / 1e12,
                                 "frac_of_sustained_peak": rnn_flops / (ms_step * 1e-3) / 1e12 / 1377.0}
    # ---- config 2: standalone CTC sweep, B=64 (+ B=512 bandwidth-regime point)
    sweep = []
    for (Tc, Lc, Bc) in [(200, 50, 64), (400, 100, 64), (800, 200, 64), (1600, 400, 64), (400, 100, 512)]:
        rng = np.random.default_rng(7)
        xl = torch.tensor(rng.standard_normal((Tc, Bc, 29)), dtype=torch.float32, device=dev)
        rngl = np.random.default_rng(11)
        labs = []
        for b in range(Bc):                                   # adjacent repeats collapsed, as the reference call does
            r = rngl.integers(0, 28, size=Lc)
            labs.append([int(v) for i, v in enumerate(r) if i == 0 or v != r[i - 1]])
        Lm = max(len(r) for r in labs)
        lab = np.zeros((Bc, Lm), np.int32)
        for b, r in enumerate(labs):
            lab[b, :len(r)] = r
        ll = np.asarray([len(r) for r in labs], np.int32)
        lab_d = torch.from_numpy(lab).to(dev)
        ll_d = torch.from_numpy(ll).to(dev)
        sl_d = torch.full((Bc,), Tc, dtype=torch.int32, device=dev)
        loss = torch.empty(Bc, device=dev)
        grad = torch.empty_like(xl)
        wsb = _lib.lib.ds2_ctc_workspace_bytes(Tc, Bc, 29, Lm)
        ws = ops._workspace(wsb, dev, "ctc")

        def run():
            ops.check(_lib.lib.ds2_ctc_loss_fwd_bwd(ops._p(xl), Tc, Bc, 29, ops._p(lab_d), Lm, ops._p(ll_d), ops._p(sl_d), None,
                                                    ops._p(loss), ops._p(grad), ops._p(ws), wsb, ops._stream()), "ctc")
        ms = _time_kernel(run, iters=10, warm=3)
        byts = 2.0 * Tc * Bc * 29 * 4
        sweep.append({"T": Tc, "L": Lc, "B": Bc, "ms": ms, "algorithmic_GBps": byts / ms / 1e6, "frac_of_hbm": byts / ms / 1e6 / peak_hbm,
                      "ns_per_lattice_step": ms * 1e6 / Tc})
    out["ctc_sweep"] = sweep
    out["ctc_peak_source"] = "measured (MEASURED_PEAKS.json, hbm copy)" if "measured" in peak_src else peak_src
    # ---- the step after the hot path (NOT part of the metric): fused Adam + EMA over all parameters
    try:
        from deepspeech_b200.optim import AdamEMA
        ps = [q for q in model.parameters() if q.requires_grad and q.grad is not None]
        opt = AdamEMA(ps, 1e-5)
        n_par = sum(q.numel() for q in ps)
        ms_opt = _time_kernel(lambda: opt.step())
        out["optimizer_step"] = {"ms": ms_opt, "params": n_par, "algorithmic_GBps": 36.0 * n_par / ms_opt / 1e6,
                                 "frac_of_hbm": 36.0 * n_par / ms_opt / 1e6 / peak_hbm,
                                 "note": "Adam + EMA, one launch, 20 B read + 16 B written per parameter; not in the timed step"}
    except Exception as e:  # noqa: BLE001 - reported, never fatal for the headline
        out["optimizer_step"] = {"error": str(e)}
    # ---- CPU baseline on this box's host cores (bounded sample: 3 warm-up + 5 timed steps on the largest bucket that
    #      keeps them within ~45 s)
    if args.gpus == 1:
        from oracle import ds2_cpu_baseline as cb
        threads = os.cpu_count() or 1
        timer = cb.StepTimer(threads, H, L)
        bucket, frames, _ = _pick_cpu_bucket(timer, B, 8, 45.0)
        r = timer.run(B, bucket, 5, 3)
        out["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": "dummy bucket T=%d frames x B=%d (%.0f audio-s/step), 3 warm-up + 5 timed steps, fp32 torch-CPU "
                                         "port of the same step (reference TF path unavailable on this box)" % (frames, B, r["audio_s_per_step"]),
                               "ms_per_step": r["ms_per_step"], "step_ms": {"min": r["ms_min"], "median": r["ms_median"], "max": r["ms_max"]}}
        if threads > 8:
            # the reference's own setting: OMP_NUM_THREADS=8 / intra_op=8 (src/setenvs.py:22-23, src/deepSpeech_train.py:518-523)
            t8 = cb.StepTimer(8, H, L)
            t8.step(B, bucket)
            r8 = t8.run(B, bucket, 2, 0)
            out["cpu_baseline"]["at_8_threads"] = {"value": r8["value"], "ms_per_step": r8["ms_per_step"]}
    return out


if __name__ == "__main__":
    main()
