"""Multi-GPU arm of bench.py: Llama-3-8B layer-sharded over N stages (one rank per GPU, torchrun), decode with N
micro-batches in flight, hidden-state hand-off over NCCL send/recv (NVLink), token ring last -> first stage.
Weak scaling: every GPU always runs all its layers once per micro-batch step; the number of micro-batches (and
so of concurrent sequences) grows with N."""
from __future__ import annotations

import json
import os

import torch
import torch.distributed as dist

from bench import METRIC, UNIT, WORKLOADS, ClockSampler, load_peaks, stage_bytes_per_step, workload_name


def run_pipeline(args) -> int:
    from src.engine import cabi
    from src.engine.pipeline import ShardStepper, StagePipeline
    from src.model import random_init
    from src.model.shard_manager import contiguous_partition
    from src.network.handoff import StageLink, init_stage_group
    from src.worker.node import ModelShard

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world = init_stage_group("nccl")
    dev = torch.device("cuda", local)
    model, _, prompt_len = WORKLOADS[args.workload or "llama3-8b"]
    batch = args.batch or 1
    n_micro = world
    K = args.steps if args.steps > 0 else 64
    W = max(args.warmup if args.warmup >= 0 else 4, 3)
    hf = random_init.NAMED_CONFIGS[model]
    L = hf["num_hidden_layers"]
    layers = contiguous_partition(L, world)[rank]
    first, last = rank == 0, rank == world - 1
    params = random_init.random_stage_parameters(hf, layers, first, last, 1234, torch.bfloat16, device=str(dev))
    max_ctx = prompt_len + 2 * (W + K) + 16
    shard = ModelShard(rank, params, config=hf, max_rows=max(128, batch * prompt_len), max_seqs=n_micro * batch, max_context=max_ctx,
                       page_size=64)
    shard.to_device()
    ex = shard.executor
    link = StageLink(rank, world)
    d = hf["hidden_size"]
    hbm_peak, _, peak_kind = load_peaks()

    # ---- prefill: the micro-batches follow each other through the stages (stage s works on micro-batch m while
    # stage s+1 works on m-1); the first tokens return to stage 0 after the last hand-off has been queued ----
    def prefill_all(seed: int):
        g = torch.Generator().manual_seed(seed)
        toks, keep = [], []
        for m in range(n_micro):
            prompts = torch.randint(0, hf["vocab_size"], (batch, prompt_len), generator=g)
            seqs = torch.arange(m * batch, (m + 1) * batch, dtype=torch.int32)
            pos = torch.arange(prompt_len, dtype=torch.int32)[None, :].expand(batch, prompt_len)
            if first:
                out = shard.compute({"input_ids": prompts, "positions": pos, "seq_ids": seqs})
            else:
                tile = torch.empty(batch, prompt_len, d, dtype=torch.bfloat16, device=dev)
                dist.recv(tile, src=rank - 1)
                out = shard.compute({"hidden_states": tile, "positions": pos, "seq_ids": seqs})
            if not last:
                hid = out["hidden_states"].contiguous()
                keep.append(hid)
                dist.send(hid, dst=rank + 1)
            toks.append(out["next_token"].to(torch.int32) if last else torch.zeros(batch, dtype=torch.int32, device=dev))
        if world > 1:
            for m in range(n_micro):
                if last:
                    dist.send(toks[m], dst=0)
                elif first:
                    dist.recv(toks[m], src=world - 1)
        return toks

    prefill_all(3)  # untimed: NCCL point-to-point communicators, tensor maps and module loading happen here
    torch.cuda.synchronize()
    shard.reset()
    t_pre0, t_pre1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    t_pre0.record()
    first_tokens = prefill_all(7)
    t_pre1.record()
    torch.cuda.synchronize()
    pre = torch.tensor([t_pre0.elapsed_time(t_pre1)], device=dev)
    dist.all_reduce(pre, op=dist.ReduceOp.MAX)
    prefill_ms = float(pre)

    stepper = ShardStepper(shard, batch, n_micro)
    for m in range(n_micro):
        stepper.set_positions(m, torch.full((batch,), prompt_len, dtype=torch.int32), reserve_to=max_ctx)
    pipe = StagePipeline(link, stepper, batch, n_micro, d, torch.bfloat16, dev)

    def timed(n_steps, start_tokens, record=False):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        got = pipe.decode(start_tokens, n_steps, record=record)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)  # the slowest rank defines the step time
        return float(ms), got

    # warm-up, then the timed region (device-resident: tokens go round the ring without touching the host)
    _, got = timed(W, first_tokens if first else None, record=True)
    nxt = [got[m, :, -1].contiguous() for m in range(n_micro)] if first else None
    launches0 = cabi.kernel_launch_count()
    sampler = ClockSampler(local)
    if first:
        sampler.start()
    dec_ms, got = timed(K, nxt, record=True)
    launches = torch.tensor([cabi.kernel_launch_count() - launches0], device=dev, dtype=torch.int64)
    dist.all_reduce(launches)
    value = n_micro * batch * K / (dec_ms * 1e-3)

    # end to end: the first stage takes every step's token ids from pinned host memory and the returned tokens
    # are read back to the host before the next step (what a master-driven loop does)
    nxt = [got[m, :, -1].contiguous() for m in range(n_micro)] if first else None
    host_in = [t.cpu().pin_memory() for t in nxt] if first else None
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cur_in = [h.to(dev, non_blocking=True) for h in host_in] if first else None
    pipe.decode(cur_in, K, record=True, via_host=True)
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(ms)
    clocks = sampler.stop() if first else None

    # per-stage roofline: every stage runs n_micro * K micro-batch steps in the timed region
    ctx_mid = prompt_len + W + K / 2.0
    my_bytes = stage_bytes_per_step(hf, len(layers), last, batch, ctx_mid)
    fr = torch.zeros(world, device=dev)
    fr[rank] = my_bytes * n_micro * K / (dec_ms * 1e-3) / 1e9 / hbm_peak
    dist.all_reduce(fr)
    if first:
        worst = int(torch.argmin(fr))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dec_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(model, batch, prompt_len, world, L), "model": model, "batch": batch,
                       "micro_batches": n_micro, "prompt_len": prompt_len, "stages": world,
                       "weights": "random-init (HF default init, seed 1234)", "handoff": "NCCL send/recv, bf16 [B,d] tile",
                       "l2": "inputs larger than L2: every stage streams >= 1.7 GB of weights per micro-batch step"},
            "e2e": {"value": n_micro * batch * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 4 * batch * n_micro,
                    "d2h_bytes_per_step": 4 * batch * n_micro, "ms_per_step": e2e_ms / K},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": float(fr[worst]) * hbm_peak, "peak": hbm_peak, "unit": "GB/s",
                         "frac": float(fr[worst]), "traffic": None, "peak_kind": peak_kind,
                         "kernel": f"decode step of stage {worst} (slowest stage; persistent kernel per micro-batch step)",
                         "per_stage_frac": [round(float(x), 4) for x in fr]},
            "cpu_baseline": None,
            "prefill": {"tokens_per_sec": n_micro * batch * prompt_len / (prefill_ms * 1e-3), "ms": prefill_ms,
                        "note": "micro-batches follow each other through the stages; tcgen05 GEMMs + flash-style attention per stage"},
        }
        print(json.dumps(line))
    dist.barrier()
    dist.destroy_process_group()
    return 0
