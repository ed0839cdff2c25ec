#!/usr/bin/env python
"""bench.py -- train tokens/s of the fused B200 training step on the configuration BASELINE.json's
metric is quoted on: GPT-2 124M (12 x 12 x 768, vocab 50257, block_size 1024), bf16 compute, synthetic
tokens, data parallel over N GPUs (weak scaling: per-GPU batch fixed).

    python bench.py --gpus N --steps K --warmup W            # our arm  (N > 1: launched by torchrun)
    python bench.py --impl reference --steps K --warmup W    # the reference's arithmetic on the host cores
    python bench.py --model gpt2-medium | gpt2-xl --accum 4 | gpt-mini | gpt-nano   # the other BASELINE.json configs
                                                             # (default --model gpt2 = config 3, the headline)

One "step" = forward, backward (gradient all-reduce overlapped when N > 1), global-norm clip + AdamW:
exactly GPTTrainer._run_batch (mingpt/trainer.py:118-133).  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch

GPT2 = dict(n_layer=12, n_head=12, n_embed=768, vocab_size=50257, block_size=1024)
NOMINAL_BF16 = 2.25e15
# BASELINE.json configs (SURVEY.md 8, table of shapes): per-GPU batch, sequence length, vocabulary, accumulation
MODELS = {
    "gpt-nano": dict(dims=dict(n_layer=3, n_head=3, n_embed=48), batch=64, seq=128, vocab=65, accum=1, char=True, cfg=1),
    "gpt-mini": dict(dims=dict(n_layer=6, n_head=6, n_embed=192), batch=64, seq=256, vocab=65, accum=1, char=True, cfg=2),
    "gpt2": dict(dims=dict(n_layer=12, n_head=12, n_embed=768), batch=32, seq=1024, vocab=50257, accum=1, char=False, cfg=3),
    "gpt2-medium": dict(dims=dict(n_layer=24, n_head=16, n_embed=1024), batch=32, seq=1024, vocab=50257, accum=1, char=False, cfg=4),
    "gpt2-xl": dict(dims=dict(n_layer=48, n_head=25, n_embed=1600), batch=8, seq=1024, vocab=50257, accum=4, char=False, cfg=5),
}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(bf16_burst=p.get("bf16_tflops", 1590.0), bf16_sustained=p.get("bf16_tflops_sustained", 1400.0),
                    hbm=p.get("hbm_gbs", 6650.0), source="MEASURED_PEAKS.json")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def make_tokens(B, T, V, seed):
    g = torch.Generator().manual_seed(seed)
    tok = torch.randint(0, V, (B, T + 1), generator=g)
    return tok[:, :-1].contiguous(), tok[:, 1:].contiguous()


def make_corpus(n_bytes, vocab, seed):
    """seeded synthetic byte corpus over a `vocab`-symbol alphabet (configs 1-2 are char-LMs "via CharDataset")"""
    g = torch.Generator().manual_seed(seed)
    alphabet = torch.arange(33, 33 + vocab, dtype=torch.uint8)
    return bytes(alphabet[torch.randint(0, vocab, (n_bytes,), generator=g)].tolist())


# ----------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle (a port of the reference's arithmetic) on the host cores
# ----------------------------------------------------------------------------------------------------
def cpu_oracle_tokens_per_s(variant: str, steps: int, warmup: int, budget_s: float, B: int = 1, T: int = 1024, dims=None):
    from oracle import gpt_oracle as go
    import mingpt_distributed_b200 as m
    dims = dims or GPT2
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = m.GPT(m.GPTConfig(model_type=None, embed_drop=0.0, resid_drop=0.0, attn_drop=0.0, variant=variant, **dims), verbose=False)
    tr = go.OracleTrainer(model.state_dict(), dims["n_head"], variant)
    del model
    times = []
    t_start = time.time()
    for i in range(warmup + steps):
        x, y = make_tokens(B, T, dims["vocab_size"], 1234 + i)
        t0 = time.time()
        tr.run_batch(x, y)
        dt = time.time() - t0
        if i >= warmup:
            times.append(dt)
        if time.time() - t_start > budget_s and len(times) >= 1:
            break
    ms = 1000.0 * sum(times) / len(times)
    return B * T / (ms / 1000.0), ms, cores, len(times), f"{len(times)} steps of B={B} x T={T} tokens, fp32, torch {torch.__version__} CPU, {cores} threads"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    budget = 150.0
    spec = MODELS[args.model]
    dims = dict(vocab_size=spec["vocab"], block_size=spec["seq"], **spec["dims"])
    Bc = spec["batch"] if spec["char"] else 1         # the small char-LM configs fit the host at their full batch
    val, ms, cores, n, sample = cpu_oracle_tokens_per_s("reference", args.steps, min(args.warmup, 1), budget, B=Bc, T=spec["seq"], dims=dims)
    line = dict(impl="reference", metric="train_tokens_per_s", value=val, unit="tokens/s", n_gpus=args.gpus, steps=n,
                warmup=min(args.warmup, 1), ms_per_step=ms, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic",
                config=dict(workload=f"{args.model} train step (L{dims['n_layer']} H{dims['n_head']} C{dims['n_embed']} V{dims['vocab_size']} T{dims['block_size']})",
                            global_batch=Bc, seq_len=spec["seq"],
                            parallelism="cpu", note="the reference's as-written graph (oracle port of mingpt/model.py + trainer.py:118-133; "
                                                    "the reference itself is CUDA+NCCL-only and absent from this box)"),
                cpu_baseline=dict(value=val, unit="tokens/s", cores=cores, kind="port", sample=sample),
                e2e=dict(value=val, unit="tokens/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--model", default="gpt2", choices=list(MODELS), help="BASELINE.json config; gpt2 (config 3) is the headline")
    ap.add_argument("--batch", type=int, default=0, help="sequences per GPU per micro-step (0 = the config's)")
    ap.add_argument("--seq", type=int, default=0)
    ap.add_argument("--accum", type=int, default=0, help="gradient-accumulation micro-steps per optimizer step (0 = the config's)")
    ap.add_argument("--variant", default="mingpt")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=25.0)
    ap.add_argument("--profile-gemm", action="store_true", default=True)
    ap.add_argument("--quick", action="store_true", help="timed loop only (for runs under ncu): no e2e, roofline or cpu legs")
    ap.add_argument("--dropout", type=float, default=0.0, help="embed/resid/attn dropout (reference default 0.1); headline is 0")
    ap.add_argument("--no-graph", action="store_true", help="issue every kernel from Python instead of replaying a CUDA graph")
    ap.add_argument("--opt", action="append", default=[], help="library option key=value (mgb_set_option), repeatable")
    ap.add_argument("--grad-wire", default="fp32", choices=["fp32", "bf16"], help="dtype of the gradient all-reduce (N > 1)")
    ap.add_argument("--comm-sms", type=int, default=0, help="SMs left to NCCL while all-reduces are in flight (N > 1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return

    # libraries (NCCL's version banner, ...) may write to fd 1: keep stdout clean for the ONE JSON line
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    import torch.distributed as dist
    import mingpt_distributed_b200 as m
    from mingpt_distributed_b200 import _lib, ops
    opts = dict(kv.split("=") for kv in args.opt)
    for key in ("pdl",):                           # experiment knobs: MGB_OPT_PDL=0 python bench.py ...
        val = os.environ.get("MGB_OPT_" + key.upper())
        if val is not None:
            opts.setdefault(key, val)
    for key, val in opts.items():
        _lib.set_option(key, int(val))

    spec = MODELS[args.model]
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch or spec["batch"]
    T = args.seq or spec["seq"]
    K = args.accum or spec["accum"]
    V = spec["vocab"]
    dims = dict(vocab_size=V, block_size=T, **spec["dims"])
    headline = args.model == "gpt2" and B == 32 and T == 1024 and K == 1 and args.variant == "mingpt"

    torch.manual_seed(0)
    cfg = m.GPTConfig(model_type=None, embed_drop=args.dropout, resid_drop=args.dropout, attn_drop=args.dropout,
                      variant=args.variant, **dims)
    model = m.GPT(cfg, verbose=False)
    opt = m.create_optimizer(model, m.OptimizerConfig())
    model = model.to(dev)
    dp = None
    if world > 1:
        dp = m.DataParallel(model, grad_wire_dtype=torch.bfloat16 if args.grad_wire == "bf16" else None, comm_sms=args.comm_sms)

    nbatches = 4
    loader = None
    if spec["char"]:
        # configs 1-2 are char-LMs "via CharDataset": a seeded synthetic byte corpus, encoded once, resident on the
        # device; a batch is cut out of it by the window-gather kernel from host-chosen window starts
        ds = m.CharDataset(m.DataConfig(path=None, block_size=T, train_split=0.9, truncate=1.0), data=make_corpus(1 << 20, V, 7 + rank))
        codes = ds.device_codes(dev)
        g = torch.Generator().manual_seed(1234 + rank)
        starts_host = [torch.randint(0, len(ds), (K * B,), generator=g).pin_memory() for _ in range(nbatches)]
        err = torch.zeros(1, dtype=torch.int32, device=dev)

        def cut(starts_dev):
            x = torch.empty(K * B, T, dtype=torch.int64, device=dev); y = torch.empty_like(x)
            ops.token_windows(codes, starts_dev, T, x, y, err)
            return x, y
        resident = [cut(s.to(dev)) for s in starts_host]
        h2d_bytes = K * B * 8
    else:
        host = [make_tokens(K * B, T, V, 1234 + rank + 100 * i) for i in range(nbatches)]
        pinned = [(x.pin_memory(), y.pin_memory()) for x, y in host]
        resident = [(x.to(dev), y.to(dev)) for x, y in host]
        h2d_bytes = 2 * K * B * T * 8

    stepper = m.GraphedTrainStep(model, opt, 1.0, dp, use_graph=not args.no_graph, accum=K)

    def step_resident(i):
        x, y = resident[i % nbatches]
        return stepper(x, y)

    def step_e2e(i):
        if spec["char"]:                     # pinned host window starts -> device -> gather kernel -> step
            x, y = cut(starts_host[i % nbatches].to(dev, non_blocking=True))
        else:
            x, y = pinned[i % nbatches]      # pinned host tokens -> static device buffers (H2D inside the step)
        return stepper(x, y).item()          # device -> host read of the loss every step

    def step_eager(i):
        x, y = resident[i % nbatches]
        if K == 1:
            return m.train_step(model, opt, x, y, 1.0, dp, zero_grad=True)
        for p in model.parameters():
            p.grad = None
        for k in range(K):
            last = k == K - 1
            m.train_step(model, opt, x[k * B:(k + 1) * B], y[k * B:(k + 1) * B], 1.0, dp, accumulate=not last, loss_div=float(K))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        n0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        n = _lib.launch_count() - n0
        if stepper.use_graph:     # CUDA-graph replays: the kernels were recorded once, at capture
            n += stepper.launches_per_replay((K * B, T)) * steps
        return ms, n

    # W warm-up steps, then more of them until the part has been under load for 1.5 s: it is power-capped and sheds clock
    # over the first second, so a timed leg that starts earlier runs up to 3 % ahead of the legs behind it
    import time
    for i in range(args.warmup):
        step_resident(i)
    torch.cuda.synchronize()
    t_one = time.time()
    step_resident(args.warmup)          # one more, timed on the host: by now it is a graph replay
    torch.cuda.synchronize()
    t_one = max(time.time() - t_one, 1e-4)
    extra_warm = min(200, int(1.5 / t_one) + 1)
    if world > 1:                       # every rank must run the same number of (collective) steps
        t = torch.tensor([extra_warm], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        extra_warm = int(t.item())
    for i in range(extra_warm):
        step_resident(args.warmup + 1 + i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_total, launches = timed(step_resident, args.steps)
    if args.quick:
        ms_e2e = ms_total
        args.profile_gemm, args.no_cpu_baseline = False, True
    else:
        for i in range(2):
            step_e2e(i)
        ms_e2e, _ = timed(step_e2e, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    tokens_per_step = world * K * B * T
    value = tokens_per_step / (ms_step / 1000.0)
    e2e_value = tokens_per_step / (ms_e2e / args.steps / 1000.0)

    # roofline of the dominant kernel (the tcgen05 GEMM): every launch of two extra steps bracketed by CUDA events
    roof = None
    peaks = measured_peaks()
    if args.profile_gemm:
        records = []
        orig = ops.gemm

        def timed_gemm(a, b, out, **kw):
            M, N = out.shape
            Kd = a.shape[0] if kw.get("a_mn") else a.shape[1]
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = orig(a, b, out, **kw)
            e.record()
            records.append((2.0 * M * N * Kd, s, e))
            return r

        ops.gemm = timed_gemm
        try:
            step_eager(0)
            records.clear()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            t0.record()
            for i in range(2):
                step_eager(i)
            t1.record()
            torch.cuda.synchronize()
        finally:
            ops.gemm = orig
        # DRAM traffic per launch from the committed ncu --set full capture of every GEMM shape of the step
        # (tools/gemm_ncu.py -> profiles/*_gemm_shapes_ncu.json), weighted by how often each shape runs
        traffic, traffic_alg, tsrc = None, None, None
        for name in ("r2_gemm_shapes_ncu.json", "r1_e_gemm_shapes_ncu.json"):
            tpath = os.path.join(ROOT, "profiles", name)
            if os.path.exists(tpath) and headline:
                shapes = json.load(open(tpath))
                weights = [2 if sh["name"].startswith("head") else GPT2["n_layer"] for sh in shapes]
                traffic = sum(w * sh["dram_bytes"] for w, sh in zip(weights, shapes)) / sum(weights)
                traffic_alg = sum(w * sh["algorithmic_bytes"] for w, sh in zip(weights, shapes)) / sum(weights)
                tsrc = f"profiles/{name} (ncu --set full, dram read+write, mean per launch)"
                break
        flops = sum(r[0] for r in records)
        gemm_ms = sum(r[1].elapsed_time(r[2]) for r in records)
        step_ms_instr = t0.elapsed_time(t1)
        achieved = flops / (gemm_ms / 1000.0) / 1e12
        roof = dict(bound="tensor", kernel="mgb::gemm::gemm_pair_kernel (all %d launches of a step)" % (len(records) // 2),
                    achieved=achieved, peak=peaks["bf16_sustained"], unit="TFLOP/s", frac=achieved / peaks["bf16_sustained"],
                    peak_source=peaks["source"] + " bf16_tflops_sustained (kernel timed inside a long step)",
                    frac_of_nominal=achieved * 1e12 / NOMINAL_BF16, traffic=traffic, traffic_algorithmic=traffic_alg,
                    traffic_source=tsrc,
                    share_of_step=gemm_ms / step_ms_instr, flops_per_launch_avg=flops / max(len(records), 1),
                    ms_per_launch_avg=gemm_ms / max(len(records), 1))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        Bc = B if spec["char"] else 1
        v_ref, ms_ref, cores, n, sample = cpu_oracle_tokens_per_s("reference", 2, 1, args.cpu_budget, B=Bc, T=T, dims=dims)
        cpu = dict(value=v_ref, unit="tokens/s", cores=cores, kind="port",
                   sample="reference's as-written graph via the oracle port: " + sample)

    if rank == 0:
        f_tok = m.flops_per_token(dims["n_layer"], dims["n_embed"], V, T, args.variant)
        line = dict(metric="train_tokens_per_s", value=value, unit="tokens/s", n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=ms_step, higher_is_better=True, scaling="weak", vs_baseline=None,
                    dtype="bf16", data="synthetic",
                    config=dict(workload=f"{args.model} train step (BASELINE.json config {spec['cfg']}: L{dims['n_layer']} H{dims['n_head']} "
                                         f"C{dims['n_embed']} V{V} T{T}), variant={args.variant}, "
                                         "fp32 master weights + AdamW state, bf16 GEMM/attention operands"
                                         + (", tokens cut from a device-resident synthetic byte corpus (CharDataset)" if spec["char"] else ""),
                                global_batch=world * K * B, micro_batch_per_gpu=B, accumulation_steps=K, seq_len=T,
                                parallelism=f"dp{world}", dropout=args.dropout, options=opts,
                                grad_allreduce=(None if world == 1 else f"{args.grad_wire} SUM per >=25 MiB bucket, comm_sms={args.comm_sms}"),
                                launch=("eager (one Python call per kernel)" if not stepper.use_graph else
                                        "CUDA graph replay of the whole step" if dp is None else
                                        "CUDA graph segments cut at gradient-bucket boundaries, NCCL all-reduces launched between them"),
                                warmup_extra=f"{extra_warm + 1} more untimed steps: load held for 1.5 s before the timed legs (power-capped clocks settle)",
                                l2="per-step working set (activations + parameters / optimizer state, GBs) is far larger than the 126 MB L2"
                                   if not spec["char"] else "inputs rotate over 4 batches; working set of the small char-LM configs partly fits the 126 MB L2"),
                    mfu=dict(flops_per_token=f_tok, of_nominal_2p25pf=value * f_tok / (world * NOMINAL_BF16),
                             of_measured_sustained=value * f_tok / (world * peaks["bf16_sustained"] * 1e12)),
                    clocks=clocks,
                    e2e=dict(value=e2e_value, unit="tokens/s", h2d_bytes_per_step=h2d_bytes, d2h_bytes_per_step=4,
                             ms_per_step=ms_e2e / args.steps),
                    gpu_launches=launches, roofline=roof, cpu_baseline=cpu)
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
