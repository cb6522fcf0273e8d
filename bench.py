#!/usr/bin/env python
"""Headline benchmark: decode tokens/s (and prefill tokens/s) of the layer-sharded pipeline on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one decode step of the whole pipeline over the batch (one new token per sequence).
N=1 : BASELINE.json configs[1]  -- TinyLlama-1.1B bf16, batch 1, 512-token prompt, one stage on one B200.
N>1 : BASELINE.json configs[2]  -- Llama-3-8B bf16 layer-sharded over N stages (one rank per GPU), batch 1 per
      micro-batch (--batch B for 2..16), N micro-batches in flight, hidden-state hand-off over NCCL send/recv (NVLink).
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle pipeline (the reference itself emits no
tokens) on the host cores for the same metric / config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "distributed-llms_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "decode_tokens_per_sec"
UNIT = "tokens/s"


# ------------------------------------------------------------------------------------------------------
# helpers
# ------------------------------------------------------------------------------------------------------

def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    return 6650.0, 1590.0, "fallback"  # B200_PROFILING.md fallback figures


def stage_bytes_per_step(hf: dict, n_layers: int, has_head: bool, batch: int, ctx: float, esz: int = 2) -> float:
    """Algorithmic HBM bytes of one decode step of one stage: every weight of the stage once,
    the K and V of every cached token once (+ lm_head on the last stage).  Norm weights, biases and
    activations (< 0.1 %) are left out."""
    d = hf.get("hidden_size", hf.get("n_embd"))
    H = hf.get("num_attention_heads", hf.get("n_head"))
    KV = hf.get("num_key_value_heads", H)
    hd = hf.get("head_dim") or d // H
    ffn = hf.get("intermediate_size", hf.get("n_inner") or 4 * d)
    g = 2 if hf["model_type"] == "gpt2" else 3
    w = d * H * hd + 2 * d * KV * hd + H * hd * d + g * d * ffn
    kv = batch * ctx * 2 * KV * hd
    total = n_layers * (w + kv) * esz
    if has_head:
        total += hf["vocab_size"] * d * esz
    return float(total)


def stage_prefill_flops(hf: dict, n_layers: int, has_head: bool, batch: int, T: int) -> float:
    """Algorithmic flops of one prefill of `batch` prompts of T tokens through one stage: the dense
    projections, causal attention (half of T^2), and the lm_head for the last position of every prompt."""
    d = hf.get("hidden_size", hf.get("n_embd"))
    H = hf.get("num_attention_heads", hf.get("n_head"))
    KV = hf.get("num_key_value_heads", H)
    hd = hf.get("head_dim") or d // H
    ffn = hf.get("intermediate_size", hf.get("n_inner") or 4 * d)
    g = 2 if hf["model_type"] == "gpt2" else 3
    w = d * H * hd + 2 * d * KV * hd + H * hd * d + g * d * ffn
    per_layer = 2.0 * T * w + 4.0 * T * T * H * hd / 2.0
    total = batch * n_layers * per_layer
    if has_head:
        total += batch * 2.0 * hf["vocab_size"] * d
    return total


def prefill_roofline(flops: float, ms: float) -> dict:
    """Prefill is bound by the bf16 tensor cores: achieved TFLOP/s against the measured sustained cuBLAS figure."""
    _, tf_peak, kind = load_peaks()
    ach = flops / (ms * 1e-3) / 1e12
    return {"bound": "tensor", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach / tf_peak, "peak_kind": kind}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.path = gpu_index, None, f"/tmp/dllm_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for n, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.remove(self.path)
        except Exception:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_name(model: str, batch: int, prompt_len: int, stages: int, n_layers: int, n_micro: int = 0) -> str:
    """The same string in both arms' `config.workload`."""
    if stages == 1:
        return f"{model} decode, batch {batch}, {prompt_len}-token prompt, 1 stage"
    return (f"{model} decode, {n_layers} layers over {stages} stages, {n_micro or stages} micro-batches of {batch} in flight, "
            f"{prompt_len}-token prompts")


def make_config(model: str, batch: int, prompt_len: int, stages: int, n_layers: int, n_micro: int = 0) -> dict:
    """`config` of the JSON line -- byte-identical in our arm and in the reference arm for the same workload.
    Arm-specific details (layer partition, hand-off kind, ...) go under `detail`, not here."""
    return {"workload": workload_name(model, batch, prompt_len, stages, n_layers, n_micro), "model": model, "batch": batch,
            "micro_batches": (n_micro or stages) if stages > 1 else 1, "prompt_len": prompt_len, "stages": stages,
            "weights": "random-init (HF default init, seeded)",
            "l2": "inputs larger than L2: every decode step streams all the weights of every stage (>= 1.7 GB per stage) "
                  "from memory"}


WORKLOADS = {
    # name: (model, batch per micro-batch, prompt length)
    "tinyllama-1.1b": ("tinyllama-1.1b", 1, 512),
    "llama3-8b": ("llama3-8b", 1, 512),
    "gpt2-small": ("gpt2-small", 1, 16),
    "qwen2.5-14b": ("qwen2.5-14b", 1, 2048),   # BASELINE configs[3]: 4 stages, micro-batched prefill of 4 x 2048 tokens
    "llama3-70b": ("llama3-70b", 32, 4096),    # BASELINE configs[4]: 8 stages, batch 32, 4k context
}


# ------------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle pipeline on the host cores
# ------------------------------------------------------------------------------------------------------

def cpu_pipeline_tokens_per_sec(model: str, batch: int, prompt_len: int, steps: int, warmup: int,
                                n_stages: int = 1, budget_s: float = 150.0):
    """Time `steps` greedy decode steps of the CPU oracle pipeline (fp32 math on seeded random-init weights, all host
    threads) after a `prompt_len`-token prefill.  Returns (tokens/s, description, cores, ms per step, steps done).

    Every timed step runs ALL the model's layers and the lm_head once -- nothing is extrapolated.  Models whose fp32
    weights exceed a bounded host sample (8B and up: 32 GB+) hold `distinct` (4) seeded layers and run them
    n_layers / 4 times in sequence, each pass with its own KV cache: the same flops and the same bytes per step as
    the full depth (4 layers are 3.5 GB+, far beyond any cache), at an eighth of the memory and start-up time."""
    from oracle.shard_oracle import (NAMED_CONFIGS, OracleShard, embed_keys, head_keys, make_prompts, make_state_dict,
                                     pipeline_step)
    cfg = NAMED_CONFIGS[model]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    L = cfg.n_layers
    distinct = L
    if L * (cfg.hidden * cfg.hidden * 2 + 3 * cfg.hidden * cfg.ffn) * 4 > 24e9:
        distinct = 4
    max_ctx = prompt_len + warmup + steps + 8
    note = ""
    if distinct == L:
        from oracle.shard_oracle import build_pipeline
        sd = make_state_dict(cfg, 1234, torch.float32)
        stages = build_pipeline(cfg, sd, min(n_stages, L), torch.float32, max_context=max_ctx)
    else:
        sub = type(cfg)(**{**cfg.__dict__, "n_layers": distinct})
        sd = make_state_dict(sub, 1234, torch.float32)
        body = {k: v for k, v in sd.items() if k.startswith("layers.") or k.startswith("h.")}
        n_pass = (L + distinct - 1) // distinct
        stages = []
        for i in range(n_pass):  # pass i: the same `distinct` layers' tensors (shared storage), its own KV cache
            part = dict(body)
            if i == 0:
                part.update({k: sd[k] for k in embed_keys(sub)})
            if i == n_pass - 1:
                part.update({k: sd[k] for k in head_keys(sub)})
            stages.append(OracleShard(i, part, sub, torch.float32, max_ctx))
        note = (f"; {L} layer executions per step over {distinct} distinct seeded layers' weights (the 32 GB+ fp32 model is "
                f"not materialised), nothing extrapolated")
    prompts = make_prompts(cfg.vocab, batch, prompt_len, seed=7)
    with torch.no_grad():
        out = pipeline_step(stages, {"input_ids": prompts})
        tok = out["next_token"]
        for _ in range(warmup):
            tok = pipeline_step(stages, {"input_ids": tok[:, None]})["next_token"]
        t0 = time.perf_counter()
        done = 0
        for _ in range(steps):
            tok = pipeline_step(stages, {"input_ids": tok[:, None]})["next_token"]
            done += 1
            if time.perf_counter() - t0 > budget_s:
                break
        dt = time.perf_counter() - t0
    tps = batch * done / dt
    desc = (f"{model} fp32 CPU oracle pipeline ({n_stages} stage(s) run one after another on the shared host cores), batch {batch}, "
            f"{prompt_len}-token prompt, {done} decode steps timed after {warmup} warm-up{note}")
    return tps, desc, cores, dt / done * 1e3, done


def run_reference(args) -> int:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    model, batch, prompt = WORKLOADS[args.workload or ("tinyllama-1.1b" if args.gpus == 1 else "llama3-8b")]
    steps = args.steps if args.steps > 0 else 16
    warm = args.warmup if args.warmup >= 0 else 3
    if args.batch:
        batch = args.batch
    from oracle.shard_oracle import NAMED_CONFIGS as ORACLE_CONFIGS
    n_stages = max(1, args.gpus)
    # N micro-batches of `batch` rows go through the N CPU stages one after another (the host cores are shared by
    # every stage, so micro-batches cannot overlap): a "step" of the N-stage workload is one decode step of all N
    # micro-batches = N passes of `batch` rows through all the layers.  One pass is timed per sample step (bounded
    # sample); tokens/s is the same either way: batch tokens per pass.
    tps, desc, cores, ms, done = cpu_pipeline_tokens_per_sec(model, batch, prompt, steps, warm, n_stages=n_stages)
    n_layers = ORACLE_CONFIGS[model].n_layers
    line = {
        "impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
        "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": make_config(model, batch, prompt, n_stages, n_layers),
        "cpu_baseline": {"value": tps, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "detail": {"sample_step": "one micro-batch of `batch` rows through every layer and the lm_head (measured wall time; "
                                  "`ms_per_step` x `steps` is the timed region)",
                   "ms_per_step_all_micro_batches": ms * n_stages},
        "note": "the reference as shipped emits no tokens (BASELINE.md 1); this is its CPU restatement (oracle/)",
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------------
# our arm, one GPU
# ------------------------------------------------------------------------------------------------------

def run_single(args) -> int:
    from src.engine import cabi
    from src.model import random_init
    from src.worker.node import ModelShard

    model, batch, prompt_len = WORKLOADS[args.workload or "tinyllama-1.1b"]
    if args.prompt_len:
        prompt_len = args.prompt_len
    if args.batch:
        batch = args.batch
    K = args.steps if args.steps > 0 else 256
    W = args.warmup if args.warmup >= 0 else 8
    W = max(W, 3)
    hf = random_init.NAMED_CONFIGS[model]
    L = hf.get("num_hidden_layers", hf.get("n_layer"))
    torch.cuda.set_device(0)
    dev = "cuda:0"
    params = random_init.random_stage_parameters(hf, range(L), True, True, 1234, torch.bfloat16, device=dev)
    max_ctx = prompt_len + 2 * (W + K) + 16
    shard = ModelShard(0, params, config=hf, max_rows=max(min(prompt_len, 2048) * batch, 128), max_seqs=batch, max_context=max_ctx, page_size=64)
    shard.to_device()
    ex = shard.executor
    g = torch.Generator().manual_seed(7)
    prompts = torch.randint(0, hf["vocab_size"], (batch, prompt_len), generator=g)
    hbm_peak, _, peak_kind = load_peaks()

    # ---- prefill (warm once, then timed) ----
    out = shard.compute({"input_ids": prompts})
    torch.cuda.synchronize()
    shard.reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = shard.compute({"input_ids": prompts})
    e1.record()
    torch.cuda.synchronize()
    prefill_ms = e0.elapsed_time(e1)

    # ---- token check: the device-resident loop and the token-by-token host path must generate the same ids from the
    # same prefill (a timing of a kernel that computes something else is worthless); the bench exits 3 otherwise ----
    V = 8
    first_tok = out["next_token"]
    loop_ids = shard.decode_loop(batch, V).cpu()
    shard.reset()
    nt = shard.compute({"input_ids": prompts})["next_token"]
    step_ids = []
    for _ in range(V):
        nt = shard.compute({"input_ids": nt[:, None]})["next_token"]
        step_ids.append(nt.cpu())
    tokens_checked = bool(torch.equal(loop_ids.to(torch.int64), torch.stack(step_ids, 1).to(torch.int64)))
    shard.reset()
    out = shard.compute({"input_ids": prompts})
    assert torch.equal(out["next_token"].cpu(), first_tok.cpu())
    torch.cuda.synchronize()

    # ---- decode, device-resident (value) ----
    shard.decode_loop(batch, W)
    torch.cuda.synchronize()
    launches0 = cabi.kernel_launch_count()
    sampler = ClockSampler(0)
    sampler.start()
    torch.cuda.synchronize()
    e0.record()
    toks = shard.decode_loop(batch, K)
    e1.record()
    torch.cuda.synchronize()
    dec_ms = e0.elapsed_time(e1)
    launches = cabi.kernel_launch_count() - launches0
    value = batch * K / (dec_ms * 1e-3)
    ctx_mid = prompt_len + W + K / 2.0

    # ---- decode, end to end through the C ABI with host buffers (e2e) ----
    seqs = list(range(batch))
    cur = prompt_len + W + K  # tokens cached per sequence so far
    ids_h = toks[:, -1].to(torch.int32).cpu().pin_memory()
    pos_h = torch.full((batch,), cur, dtype=torch.int32).pin_memory()
    seq_h = torch.arange(batch, dtype=torch.int32).pin_memory()
    nxt_h = torch.zeros(batch, dtype=torch.int32).pin_memory()
    for s in seqs:
        ex.reserve(s, cur + W + K + 1)
    stream = torch.cuda.current_stream()

    def host_step():
        ex.step_host(ids_h, pos_h, seq_h, next_token_out=nxt_h)
        stream.synchronize()  # the host needs the token before it can issue the next step
        ids_h.copy_(nxt_h)
        pos_h.add_(1)

    for _ in range(W):
        host_step()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        host_step()
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    e2e_ms = e0.elapsed_time(e1)
    e2e = batch * K / (e2e_ms * 1e-3)

    # ---- roofline of the decode step (one graph replay = one launch of the whole stage) ----
    algo = stage_bytes_per_step(hf, L, True, batch, ctx_mid)
    step_s = dec_ms * 1e-3 / K
    achieved = algo / step_s / 1e9
    traffic = None
    try:  # DRAM bytes per step from the committed ncu --set full capture of this workload (profiles/)
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(model)
        if t and batch == t.get("batch", WORKLOADS[model][1]) and not args.prompt_len:
            traffic = t["dram_bytes_per_step"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                "traffic": traffic, "peak_kind": peak_kind,
                "kernel": ("decode_megakernel: one persistent launch per decode loop; unit = one decode step of the stage"
                           if launches <= 4 else
                           f"one decode step = a CUDA graph of {int(launches) // K} kernels (swap-AB tcgen05 GEMMs, grouped-head "
                           "attention, row norms): more rows than the persistent kernel takes at this width"),
                "algorithmic_bytes_per_launch": algo}

    # ---- CPU baseline on the host cores (bounded sample) ----
    cpu = None
    if not args.no_cpu_baseline:
        tps, desc, cores, _, _ = cpu_pipeline_tokens_per_sec(model, batch, prompt_len, 8, 1, 1, budget_s=40.0)
        cpu = {"value": tps, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": dec_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": make_config(model, batch, prompt_len, 1, L),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 3 * 4 * batch, "d2h_bytes_per_step": 4 * batch,
                "ms_per_step": e2e_ms / K,
                "transfer": "ids / positions / slots are read from, and the sampled id written to, pinned host buffers every step "
                            "(by the kernel, in place: dllm_shard_forward adds no copy nodes for pinned int32 buffers); the host "
                            "synchronises the stream and feeds the id back"},
        "gpu_launches": int(launches),
        "tokens_checked": tokens_checked,
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "prefill": {"tokens_per_sec": batch * prompt_len / (prefill_ms * 1e-3), "ms": prefill_ms,
                    "roofline": prefill_roofline(stage_prefill_flops(hf, L, True, batch, prompt_len), prefill_ms),
                    "note": "persistent tcgen05/TMA GEMMs for the projections (rotary + KV append in the q|k|v epilogue, K-split tails), TMA + tcgen05 flash-style causal attention over the paged cache"},
    }
    print(json.dumps(line))
    return 0 if tokens_checked else 3


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=-1)
    ap.add_argument("--warmup", type=int, default=-1)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=[None, *WORKLOADS])
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--micro", type=int, default=0, help="micro-batches in flight (N > 1; default: one per stage)")
    ap.add_argument("--prompt-len", type=int, default=0, help="override the workload's prompt length (single GPU)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device: the shard forward has no CPU path"}))
        return 1
    if args.gpus <= 1:
        return run_single(args)
    from bench_pipeline import run_pipeline  # multi-GPU arm
    return run_pipeline(args)


if __name__ == "__main__":
    sys.exit(main())
