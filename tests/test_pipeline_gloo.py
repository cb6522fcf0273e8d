"""CPU, world_size 2 over gloo: the N>1 host logic (static micro-batch schedule, hidden-state hand-off, token ring).
The stage compute is a stub (the product's compute needs a GPU; this checks who talks to whom and in what order)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from src.engine.pipeline import StagePipeline
from src.network.handoff import StageLink


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _stub(rank, world, hidden):
    """Deterministic stand-in for a stage: first stage 'embeds' a token, last stage emits token = f(hidden)."""
    def compute(m, x):
        if rank == 0:
            h = x.float()[:, None].repeat(1, hidden) + 1.0          # embed(token)
        else:
            h = x.float() * 2.0 + (m + 1)                            # a layer that depends on the micro-batch
        if rank == world - 1:
            return (h.sum(dim=1).to(torch.int64) % 997).to(torch.int32)  # "argmax"
        return h
    return compute


def _expected(world, hidden, first, n_micro, n_steps):
    out = torch.zeros(n_micro, first[0].numel(), n_steps, dtype=torch.int32)
    for m in range(n_micro):
        tok = first[m]
        for t in range(n_steps):
            x = tok
            for r in range(world):
                x = _stub(r, world, hidden)(m, x)
            tok = x
            out[m, :, t] = tok
    return out


def _worker(rank, world, port, hidden, n_micro, n_steps, q, via_host=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        link = StageLink(rank, world)
        first = [torch.tensor([3 + m, 11 + m], dtype=torch.int32) for m in range(n_micro)]
        pipe = StagePipeline(link, _stub(rank, world, hidden), batch=2, n_micro=n_micro, hidden=hidden,
                             act_dtype=torch.float32, device="cpu")
        got = pipe.decode(first if rank == 0 else None, n_steps, record=True, via_host=via_host)
        if rank == 0:
            q.put(got)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_micro,n_steps,via_host", [(2, 5, False), (3, 4, False), (2, 4, True)])
def test_two_stage_schedule_and_token_ring(n_micro, n_steps, via_host):
    """`via_host`: the master-driven loop (every token id staged through host memory) follows the same schedule."""
    world, hidden = 2, 8
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, hidden, n_micro, n_steps, q, via_host)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    first = [torch.tensor([3 + m, 11 + m], dtype=torch.int32) for m in range(n_micro)]
    assert torch.equal(got, _expected(world, hidden, first, n_micro, n_steps))


def test_stage_link_roles():
    a, b, c = (StageLink(r, 3, make_ring_group=False) for r in range(3))
    assert (a.prev, a.next, a.is_first, a.is_last) == (None, 1, True, False)
    assert (b.prev, b.next, b.is_first, b.is_last) == (0, 2, False, False)
    assert (c.prev, c.next, c.is_first, c.is_last) == (1, None, False, True)
