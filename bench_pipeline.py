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

from bench import (METRIC, UNIT, WORKLOADS, ClockSampler, load_peaks, make_config, prefill_roofline, stage_bytes_per_step,
                   stage_prefill_flops)


def run_pipeline(args) -> int:
    from src.engine import cabi
    from src.engine.pipeline import ShardStepper, StagePipeline
    from src.model import random_init
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
    fused = batch <= 64  # the mailbox hand-off: inside the persistent kernel up to 8 rows at these widths, as wait / send kernels
    persistent = batch <= 8  # around the step's CUDA graph above that
    n_micro = args.micro if args.micro >= world else world
    K = args.steps if args.steps > 0 else 64
    W = max(args.warmup if args.warmup >= 0 else 4, 3)
    hf = random_init.NAMED_CONFIGS[model]
    L = hf["num_hidden_layers"]
    # Stages are contiguous runs of BLOCKS (a layer's attention block, its MLP block): a boundary may fall inside a layer --
    # the residual row is the hand-off either way -- and the last stage, which also streams the lm_head, gets fewer of
    # them.  Costs = bytes one decode step streams.
    d_, H_, KV_ = hf["hidden_size"], hf["num_attention_heads"], hf.get("num_key_value_heads", hf["num_attention_heads"])
    hd_ = hf.get("head_dim") or d_ // H_
    attn_cost = 2.0 * (2 * d_ * H_ * hd_ + 2 * d_ * KV_ * hd_)
    mlp_cost = 2.0 * 3 * d_ * hf["intermediate_size"]
    head_cost = stage_bytes_per_step(hf, 0, True, 1, 0.0)
    from src.model.shard_manager import block_partition, decode_block_times
    # balanced on estimated decode-step TIME per block (streamed bytes + the serial latency of the block's phases), not bytes
    t_attn, t_mlp, t_head, t_first = decode_block_times(hf, batch, prompt_len)
    stages_blocks = block_partition(L, world, t_attn, t_mlp, t_head, t_first)
    blocks = stages_blocks[rank]
    block_w = {"full": 1.0, "attn": attn_cost / (attn_cost + mlp_cost), "mlp": mlp_cost / (attn_cost + mlp_cost)}
    layers_eq = [sum(block_w[b] for _, b in st) for st in stages_blocks]   # stage sizes in layer equivalents
    first, last = rank == 0, rank == world - 1
    params = random_init.random_stage_parameters(hf, blocks, first, last, 1234, torch.bfloat16, device=str(dev))
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

    # ---- token check (multi-GPU parity inside the bench): from a fresh prefill, V decode steps through the fused
    # hand-off and through the NCCL schedule must give the same ids for every micro-batch, and micro-batch 0 must
    # equal the SAME model run as one stage on rank 0 alone.  A mismatch makes the bench exit non-zero. ----
    V = 8
    checked = {"fused_vs_nccl": None, "vs_single_stage": None, "steps": V}

    def fresh_state():
        shard.reset()
        toks = prefill_all(7)
        torch.cuda.synchronize()
        dist.barrier()
        return toks

    seq_all = torch.arange(n_micro * batch, dtype=torch.int32).reshape(n_micro, batch)
    fused_ids = None
    if fused:
        toks0 = fresh_state()
        vin, vout = exchange_mailboxes(rank, world, local, n_micro, batch, d)
        if first:
            for m in range(n_micro):
                vin.post_tokens(m, toks0[m], ticket=1)
        dist.barrier()
        fed = ex.decode_pipelined(vin, vout, n_micro, batch, V, 0, torch.full((n_micro, batch), prompt_len, dtype=torch.int32), seq_all)
        torch.cuda.synchronize()
        dist.barrier()
        if first:  # [n_micro, batch, V + 1]: the seed, then the ids generated by steps 0 .. V-1
            tail = torch.stack([vin.read_tokens(m)[0] for m in range(n_micro)])  # generated by the last step
            fused_ids = torch.cat([fed.cpu().permute(1, 2, 0), tail[:, :, None]], dim=2)
        dist.barrier()
        if vout is not vin:
            vout.close()
        dist.barrier()
        vin.close()
    toks0 = fresh_state()
    for m in range(n_micro):
        stepper.set_positions(m, torch.full((batch,), prompt_len, dtype=torch.int32), reserve_to=max_ctx)
    got = pipe.decode(toks0 if first else None, V, record=True)
    torch.cuda.synchronize()
    dist.barrier()
    ok = True
    single_tps = None
    if first:
        nccl_ids = torch.cat([torch.stack(toks0).cpu()[:, :, None], got.cpu()], dim=2)  # [n_micro, batch, V + 1]
        if fused_ids is not None:
            checked["fused_vs_nccl"] = bool(torch.equal(fused_ids.to(torch.int64), nccl_ids.to(torch.int64)))
            ok &= checked["fused_vs_nccl"]
        # the same model as ONE stage on this GPU: the ids of micro-batch 0, and the single-GPU tokens/s of this very
        # model (the driver's scaling table divides by the N=1 bench, which times a different, smaller model)
        whole_bytes = stage_bytes_per_step(hf, L, True, 1, 0.0)
        if 2.2 * whole_bytes < 100e9:
            whole = ModelShard(100, random_init.random_stage_parameters(hf, range(L), True, True, 1234, torch.bfloat16, device=str(dev)),
                               config=hf, max_rows=max(128, min(batch * prompt_len, 4096)), max_seqs=batch,
                               max_context=max_ctx, page_size=64)
            whole.to_device()
            g = torch.Generator().manual_seed(7)
            prompts0 = torch.randint(0, hf["vocab_size"], (batch, prompt_len), generator=g)  # micro-batch 0 of prefill_all(7)
            t0 = whole.compute({"input_ids": prompts0})["next_token"].cpu()
            loop = whole.decode_loop(batch, V).cpu()
            single_ids = torch.cat([t0[:, None], loop], dim=1)
            checked["vs_single_stage"] = bool(torch.equal(single_ids, nccl_ids[0].to(torch.int64)))
            ok &= checked["vs_single_stage"]
            whole.decode_loop(batch, W)
            torch.cuda.synchronize()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            whole.decode_loop(batch, K)
            s1.record()
            torch.cuda.synchronize()
            single_tps = batch * K / (s0.elapsed_time(s1) * 1e-3)
            del whole
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    ok = bool(int(flag))

    # per-stage roofline: every stage runs n_micro * K micro-batch steps in the timed region
    ctx_mid = prompt_len + W + K / 2.0
    n_attn = sum(1 for _, b in blocks if b != "mlp")   # layers whose KV this stage reads
    my_bytes = (stage_bytes_per_step(hf, 1, False, 1, 0.0) * layers_eq[rank] + (head_cost if last else 0.0)
                + n_attn * batch * ctx_mid * 2 * KV_ * hd_ * 2)
    fr = torch.zeros(world, device=dev)
    fr[rank] = my_bytes * n_micro * K / (dec_ms * 1e-3) / 1e9 / hbm_peak
    dist.all_reduce(fr)
    if first:
        worst = int(torch.argmin(fr))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dec_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": make_config(model, batch, prompt_len, world, L, n_micro),
            "detail": {"layers_per_stage": [round(x, 2) for x in layers_eq],
                       "stage_blocks": [f"{st[0][0]}{'' if st[0][1] == 'full' else ':' + st[0][1]}..{st[-1][0]}{'' if st[-1][1] == 'full' else ':' + st[-1][1]}" for st in stages_blocks],
                       "handoff": ("fused: peer stores from the down-projection epilogue into the next stage's mailbox + tickets"
                                   if persistent else
                                   "fused: per micro-batch step a wait kernel, the step's CUDA graph and a send kernel (peer stores into the next "
                                   "stage's mailbox + tickets), all enqueued up front; no NCCL, no host between steps" if fused else
                                   "NCCL send/recv per micro-batch step")},
            "tokens_checked": ok, "token_check": checked,
            "same_model_n1": ({"value": single_tps, "unit": UNIT, "scaling_efficiency": value / (world * single_tps),
                               "note": "the same model as one stage on one B200, same batch; efficiency = value / (N x this)"}
                              if single_tps else None),
            "e2e": {"value": n_micro * batch * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 4 * batch * n_micro,
                    "d2h_bytes_per_step": 4 * batch * n_micro, "ms_per_step": e2e_ms / K},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": float(fr[worst]) * hbm_peak, "peak": hbm_peak, "unit": "GB/s",
                         "frac": float(fr[worst]), "traffic": None, "peak_kind": peak_kind,
                         "kernel": (f"decode_megakernel of stage {worst} (the stage with the lowest fraction: every stage takes the same step "
                                    "time in the steady state, this one moves the fewest bytes): one persistent launch for the timed loop, "
                                    "unit = one micro-batch step of the stage") if persistent else
                                   (f"micro-batch step of stage {worst} (the stage with the lowest fraction): CUDA graph of swap-AB tcgen05 GEMMs + "
                                    "grouped-head decode attention"),
                         "per_stage_frac": [round(float(x), 4) for x in fr]},
            "cpu_baseline": None,
            "note": "N=1 of this bench is BASELINE configs[1] (TinyLlama-1.1B); N >= 2 is configs[2] (this model): "
                    "`same_model_n1` is this model on one B200, measured in this run on rank 0",
            "nccl_handoff": {"value": n_micro * batch * K / (nccl_ms * 1e-3), "unit": UNIT, "ms_per_step": nccl_ms / K,
                             "note": "same schedule, one launch per micro-batch step, NCCL send/recv + token ring"},
            "prefill": {"tokens_per_sec": n_micro * batch * prompt_len / (prefill_ms * 1e-3), "ms": prefill_ms,
                        # the busiest stage's flops for all micro-batches over the whole (pipelined) prefill time:
                        # includes the fill and drain bubbles of the (stages + micro-batches - 1)-slot schedule
                        "roofline": prefill_roofline(n_micro * stage_prefill_flops(hf, 1, False, batch, prompt_len) * max(layers_eq), prefill_ms),
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
    return 0 if ok else 3
