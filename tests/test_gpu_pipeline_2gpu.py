"""GPU, two devices: the pipelined 2-stage decode over NCCL must give exactly the tokens of one stage on one GPU
(same kernels, same order; the hand-off copies bits).  Skipped on a single-GPU box."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

HF = dict(model_type="llama", hidden_size=256, num_hidden_layers=4, num_attention_heads=4, num_key_value_heads=2,
          intermediate_size=512, vocab_size=500, max_position_embeddings=128, rms_norm_eps=1e-5, rope_theta=10000.0)
B, N_MICRO, PROMPT, STEPS = 2, 2, 9, 6


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _prompts(B=B):
    g = torch.Generator().manual_seed(3)
    return [torch.randint(0, HF["vocab_size"], (B, PROMPT), generator=g) for _ in range(N_MICRO)]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from src.engine.pipeline import ShardStepper, StagePipeline
    from src.model import random_init
    from src.model.shard_manager import contiguous_partition
    from src.network.handoff import StageLink
    from src.worker.node import ModelShard
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    dev = torch.device("cuda", rank)
    layers = contiguous_partition(HF["num_hidden_layers"], world)[rank]
    params = random_init.random_stage_parameters(HF, layers, rank == 0, rank == world - 1, 1234, torch.bfloat16, init="full")
    shard = ModelShard(rank, params, config=HF, max_rows=32, max_seqs=N_MICRO * B, max_context=64, page_size=16)
    shard.to_device()
    link = StageLink(rank, world)
    first_tokens = []
    for m, prompts in enumerate(_prompts()):
        seqs = torch.arange(m * B, (m + 1) * B, dtype=torch.int32)
        if rank == 0:
            out = shard.compute({"input_ids": prompts, "seq_ids": seqs})
            dist.send(out["hidden_states"].contiguous(), dst=1)
            tok = torch.zeros(B, dtype=torch.int32, device=dev)
            dist.recv(tok, src=1)
        else:
            tile = torch.empty(B, PROMPT, HF["hidden_size"], dtype=torch.bfloat16, device=dev)
            dist.recv(tile, src=0)
            pos = torch.arange(PROMPT, dtype=torch.int32)[None].expand(B, PROMPT)
            out = shard.compute({"hidden_states": tile, "positions": pos, "seq_ids": seqs})
            tok = out["next_token"].to(torch.int32)
            dist.send(tok, dst=0)
        first_tokens.append(tok)
    stepper = ShardStepper(shard, B, N_MICRO)
    for m in range(N_MICRO):
        stepper.set_positions(m, torch.full((B,), PROMPT, dtype=torch.int32), reserve_to=64)
    pipe = StagePipeline(link, stepper, B, N_MICRO, HF["hidden_size"], torch.bfloat16, dev)
    got = pipe.decode(first_tokens if rank == 0 else None, STEPS, record=True)
    torch.cuda.synchronize()
    if rank == 0:
        q.put((torch.stack(first_tokens).cpu(), got.cpu()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_stage_pipeline_equals_single_stage():
    from src.model import random_init
    from src.worker.node import ModelShard
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    first, got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    params = random_init.random_stage_parameters(HF, range(HF["num_hidden_layers"]), True, True, 1234, torch.bfloat16, init="full")
    one = ModelShard(0, params, config=HF, max_rows=32, max_seqs=N_MICRO * B, max_context=64, page_size=16)
    one.to_device()
    for m, prompts in enumerate(_prompts()):
        seqs = torch.arange(m * B, (m + 1) * B, dtype=torch.int32)
        tok = one.compute({"input_ids": prompts, "seq_ids": seqs})["next_token"].cpu()
        assert torch.equal(tok.to(torch.int32), first[m])
        for t in range(STEPS):
            tok = one.compute({"input_ids": tok[:, None], "seq_ids": seqs})["next_token"].cpu()
            assert torch.equal(tok.to(torch.int32), got[m, :, t]), (m, t)


def _fused_worker(rank, world, port, q, B=B, megakernel="1"):
    """Same model and prompts, hand-off through peer mailboxes (IPC handles exchanged over the process group): fused into
    the persistent kernels -- one launch per stage for all the steps -- or, with the persistent kernel off / more rows
    than it takes, as wait / send kernels around every step's CUDA graph."""
    os.environ["DLLM_MEGAKERNEL"] = megakernel
    import torch.distributed as dist
    from src.model import random_init
    from src.model.shard_manager import contiguous_partition
    from src.network.handoff import exchange_mailboxes
    from src.worker.node import ModelShard
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    dev = torch.device("cuda", rank)
    layers = contiguous_partition(HF["num_hidden_layers"], world)[rank]
    params = random_init.random_stage_parameters(HF, layers, rank == 0, rank == world - 1, 1234, torch.bfloat16, init="full")
    shard = ModelShard(rank, params, config=HF, max_rows=64, max_seqs=N_MICRO * B, max_context=64, page_size=16)
    shard.to_device()
    first_tokens = []
    for m, prompts in enumerate(_prompts(B)):
        seqs = torch.arange(m * B, (m + 1) * B, dtype=torch.int32)
        if rank == 0:
            out = shard.compute({"input_ids": prompts, "seq_ids": seqs})
            dist.send(out["hidden_states"].contiguous(), dst=1)
            tok = torch.zeros(B, dtype=torch.int32, device=dev)
            dist.recv(tok, src=1)
        else:
            tile = torch.empty(B, PROMPT, HF["hidden_size"], dtype=torch.bfloat16, device=dev)
            dist.recv(tile, src=0)
            pos = torch.arange(PROMPT, dtype=torch.int32)[None].expand(B, PROMPT)
            out = shard.compute({"hidden_states": tile, "positions": pos, "seq_ids": seqs})
            tok = out["next_token"].to(torch.int32)
            dist.send(tok, dst=0)
        first_tokens.append(tok)
    inbox, outbox = exchange_mailboxes(rank, world, rank, N_MICRO, B, HF["hidden_size"])
    if rank == 0:
        for m in range(N_MICRO):
            inbox.post_tokens(m, first_tokens[m], ticket=1)
    dist.barrier()
    pos = torch.full((N_MICRO, B), PROMPT, dtype=torch.int32)
    seq = torch.arange(N_MICRO * B, dtype=torch.int32).reshape(N_MICRO, B)
    fed = shard.executor.decode_pipelined(inbox, outbox, N_MICRO, B, STEPS, 0, pos, seq)
    torch.cuda.synchronize()
    dist.barrier()  # every stage has finished: the ids of the final step are in stage 0's inbox
    if rank == 0:
        last = torch.stack([inbox.read_tokens(m)[0] for m in range(N_MICRO)])       # [n_micro, B]
        got = torch.cat([fed.cpu().permute(1, 2, 0)[:, :, 1:], last[:, :, None]], 2)  # generated ids [n_micro, B, STEPS]
        q.put((torch.stack(first_tokens).cpu(), got))
    dist.barrier()
    if outbox is not inbox:
        outbox.close()
    dist.barrier()
    inbox.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("B,megakernel", [(2, "1"), (2, "0"), (12, "0")])
def test_fused_handoff_two_stages_equals_single_stage(B, megakernel, monkeypatch):
    from src.model import random_init
    from src.worker.node import ModelShard
    monkeypatch.setenv("DLLM_MEGAKERNEL", megakernel)   # the single-stage run below takes the same kernels as the stages
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_fused_worker, args=(r, 2, port, q, B, megakernel)) for r in range(2)]
    for p in procs:
        p.start()
    first, got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    params = random_init.random_stage_parameters(HF, range(HF["num_hidden_layers"]), True, True, 1234, torch.bfloat16, init="full")
    one = ModelShard(0, params, config=HF, max_rows=64, max_seqs=N_MICRO * B, max_context=64, page_size=16)
    one.to_device()
    for m, prompts in enumerate(_prompts(B)):
        seqs = torch.arange(m * B, (m + 1) * B, dtype=torch.int32)
        tok = one.compute({"input_ids": prompts, "seq_ids": seqs})["next_token"].cpu()
        assert torch.equal(tok.to(torch.int32), first[m])
        for t in range(STEPS):
            tok = one.compute({"input_ids": tok[:, None], "seq_ids": seqs})["next_token"].cpu()
            assert torch.equal(tok.to(torch.int32), got[m, :, t]), (m, t)
