"""Multi-GPU arm of bench.py: Llama-3-8B layer-sharded over N stages (one rank per GPU, torchrun), decode with N
micro-batches in flight.  `value`: the fused hand-off -- ONE persistent kernel per stage for the whole timed loop,
hidden rows and token ids move through peer mailboxes over NVLink (dllm_shard_decode_pipelined).  `e2e` and
`nccl_handoff`: the launch-per-micro-batch-step schedule with NCCL send/recv and the token ring last -> first stage.
Weak scaling: every GPU always runs all its layers once per micro-batch step; the number of micro-batches (and
so of concurrent sequences) grows with N."""
from __future__ import annotations

import json
import os

import torch
import torch.distributed as dist

from bench import (METRIC, UNIT, WORKLOADS, ClockSampler, load_peaks, prefill_roofline, stage_bytes_per_step,
                   stage_prefill_flops, workload_name)


def run_pipeline(args) -> int:
    from src.engine import cabi
    from src.engine.pipeline import ShardStepper, StagePipeline
    from src.model import random_init
    from src.model.shard_manager import balanced_partition
    from src.network.handoff import StageLink, exchange_mailboxes, init_stage_group
    from src.worker.node import ModelShard

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    rank, world = init_stage_group("nccl")
    dev = torch.device("cuda", local)
    model, wl_batch, prompt_len = WORKLOADS[args.workload or "llama3-8b"]
    if args.prompt_len:
        prompt_len = args.prompt_len
    batch = args.batch or (wl_batch if model == "llama3-70b" else 1)
    fused = batch <= 8  # the persistent kernel (and with it the fused hand-off) takes up to 8 rows at these widths
    n_micro = args.micro if args.micro >= world else world
    K = args.steps if args.steps > 0 else 64
    W = max(args.warmup if args.warmup >= 0 else 4, 3)
    hf = random_init.NAMED_CONFIGS[model]
    L = hf["num_hidden_layers"]
    # the last stage also streams the lm_head: give it fewer layers (head bytes / bytes of one layer)
    per_layer = stage_bytes_per_step(hf, 1, False, 1, 0.0)
    head_cost = (stage_bytes_per_step(hf, 0, True, 1, 0.0)) / per_layer
    partition = balanced_partition(L, world, head_cost)
    layers = partition[rank]
    first, last = rank == 0, rank == world - 1
    params = random_init.random_stage_parameters(hf, layers, first, last, 1234, torch.bfloat16, device=str(dev))
    max_ctx = prompt_len + 3 * (W + K) + 16
    shard = ModelShard(rank, params, config=hf, max_rows=max(128, min(batch * prompt_len, 4096)), max_seqs=n_micro * batch,
                       max_context=max_ctx, page_size=64)
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

    big = n_micro * batch * prompt_len > 65536
    if big:  # a second full prefill would cost minutes: warm the links and the kernels with a short prompt instead
        full_len, prompt_len = prompt_len, 64
        prefill_all(3)
        prompt_len = full_len
    else:
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

    # ---- fused hand-off: one launch per stage for W warm-up steps, one for the K timed steps ----
    inbox = outbox = None
    launches0 = cabi.kernel_launch_count()
    sampler = ClockSampler(local)
    done = 0  # decode steps so far
    fused_last = first_tokens if first else None
    dec_ms = None
    if fused:
        inbox, outbox = exchange_mailboxes(rank, world, local, n_micro, batch, d)
        if first:
            for m in range(n_micro):
                inbox.post_tokens(m, first_tokens[m], ticket=1)
        seq_all = torch.arange(n_micro * batch, dtype=torch.int32).reshape(n_micro, batch)
        pos0 = torch.full((n_micro, batch), prompt_len, dtype=torch.int32)
        ex.decode_pipelined(inbox, outbox, n_micro, batch, W, 0, pos0, seq_all)
        torch.cuda.synchronize()
        dist.barrier()
        launches0 = cabi.kernel_launch_count()
        if first:
            sampler.start()
        torch.cuda.synchronize()
        dist.barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        ex.decode_pipelined(inbox, outbox, n_micro, batch, K, W, pos0 + W, seq_all)
        f1.record()
        torch.cuda.synchronize()
        dist.barrier()
        fms = torch.tensor([f0.elapsed_time(f1)], device=dev)
        dist.all_reduce(fms, op=dist.ReduceOp.MAX)  # the slowest rank defines the step time
        dec_ms = float(fms)
        fused_last = [inbox.read_tokens(m)[0].to(dev) for m in range(n_micro)] if first else None
        done = W + K
    launches = torch.tensor([cabi.kernel_launch_count() - launches0], device=dev, dtype=torch.int64)

    stepper = ShardStepper(shard, batch, n_micro)
    for m in range(n_micro):
        stepper.set_positions(m, torch.full((batch,), prompt_len + done, dtype=torch.int32), reserve_to=max_ctx)
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

    # the same loop with a launch per micro-batch step and NCCL send/recv (device-resident: tokens go round the ring
    # without touching the host) -- reported beside the fused number
    _, got = timed(W, fused_last, record=True)
    nxt = [got[m, :, -1].contiguous() for m in range(n_micro)] if first else None
    if not fused:  # more rows than the persistent kernel takes: this schedule is the measured one
        launches0 = cabi.kernel_launch_count()
        if first:
            sampler.start()
    nccl_ms, got = timed(K, nxt, record=True)
    if not fused:
        launches = torch.tensor([cabi.kernel_launch_count() - launches0], device=dev, dtype=torch.int64)
        dec_ms = nccl_ms
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
            "config": {"workload": workload_name(model, batch, prompt_len, world, L, n_micro), "model": model, "batch": batch,
                       "micro_batches": n_micro, "prompt_len": prompt_len, "stages": world,
                       "layers_per_stage": [len(r) for r in partition],
                       "weights": "random-init (HF default init, seed 1234)",
                       "handoff": ("fused: peer stores from the down-projection epilogue into the next stage's mailbox + tickets"
                                   if fused else "NCCL send/recv per micro-batch step (batch above the persistent kernel's 8 rows)"),
                       "l2": "inputs larger than L2: every stage streams >= 1.7 GB of weights per micro-batch step"},
            "e2e": {"value": n_micro * batch * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 4 * batch * n_micro,
                    "d2h_bytes_per_step": 4 * batch * n_micro, "ms_per_step": e2e_ms / K},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": float(fr[worst]) * hbm_peak, "peak": hbm_peak, "unit": "GB/s",
                         "frac": float(fr[worst]), "traffic": None, "peak_kind": peak_kind,
                         "kernel": (f"decode_megakernel of stage {worst} (slowest stage): one persistent launch for the timed loop, "
                                    "unit = one micro-batch step of the stage") if fused else
                                   (f"micro-batch step of stage {worst} (slowest stage): CUDA graph of swap-AB tcgen05 GEMMs + "
                                    "grouped-head decode attention"),
                         "per_stage_frac": [round(float(x), 4) for x in fr]},
            "cpu_baseline": None,
            "note": "N=1 of this bench is BASELINE configs[1] (TinyLlama-1.1B); N >= 2 is configs[2] (this model). "
                    "For the same model on one B200 run `bench.py --workload " + (args.workload or "llama3-8b") + "` "
                    "(Llama-3-8B: 296 tok/s, profiles/r1d_summary.md)",
            "nccl_handoff": {"value": n_micro * batch * K / (nccl_ms * 1e-3), "unit": UNIT, "ms_per_step": nccl_ms / K,
                             "note": "same schedule, one launch per micro-batch step, NCCL send/recv + token ring"},
            "prefill": {"tokens_per_sec": n_micro * batch * prompt_len / (prefill_ms * 1e-3), "ms": prefill_ms,
                        # the busiest stage's flops for all micro-batches over the whole (pipelined) prefill time:
                        # includes the fill and drain bubbles of the (stages + micro-batches - 1)-slot schedule
                        "roofline": prefill_roofline(n_micro * stage_prefill_flops(hf, max(len(r) for r in partition), False,
                                                                                  batch, prompt_len), prefill_ms),
                        "note": "micro-batches follow each other through the stages; tcgen05 GEMMs + flash-style attention per stage"},
        }
        print(json.dumps(line))
    dist.barrier()
    if outbox is not None and outbox is not inbox:
        outbox.close()
    dist.barrier()
    if inbox is not None:
        inbox.close()
    dist.destroy_process_group()
    return 0
