"""Micro-batch pipelining of decode across layer-sharded stages (one process per GPU).

The reference's master broadcasts one request to every worker and never chains them
(reference src/master/node.py:256-269).  Here every rank runs the same static schedule:

    for step in range(n_steps):
        for m in range(n_micro):                 # micro-batches keep all stages busy (n_micro >= n_stages)
            first stage : [token ids of (m, step) from the last stage]  -> embed + layers -> hidden to stage 1
            middle stage: hidden from stage s-1                          -> layers         -> hidden to stage s+1
            last stage  : hidden from stage S-2                          -> layers + head  -> token ids to stage 0

All transfers are NCCL send/recv on device tensors (NVLink), stream-ordered: the host never waits for a token.
`compute` is injected so that the schedule can be exercised over gloo on CPUs with a stub (tests); the product
passes `ShardStepper`, which drives `dllm_shard_forward` on pre-allocated device buffers.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch

from ..network.handoff import StageLink


class ShardStepper:
    """Allocation-free decode step of one stage on device buffers (wraps ShardExecutor.step_host)."""

    def __init__(self, shard, batch: int, n_micro: int):
        self.ex = shard.executor
        ex = self.ex
        dev = ex.device
        self.batch, self.n_micro = batch, n_micro
        self.seq = [torch.arange(m * batch, (m + 1) * batch, dtype=torch.int32, device=dev) for m in range(n_micro)]
        self.pos = [torch.zeros(batch, dtype=torch.int32, device=dev) for _ in range(n_micro)]
        self.hidden_out = [torch.empty(batch, ex.geom.hidden, dtype=ex.dtype, device=dev) for _ in range(n_micro)]
        self.tokens = [torch.zeros(batch, dtype=torch.int32, device=dev) for _ in range(n_micro)]

    def set_positions(self, m: int, positions: torch.Tensor, reserve_to: int) -> None:
        self.pos[m].copy_(positions.to(torch.int32))
        for s in self.seq[m].tolist():
            self.ex.reserve(s, reserve_to)

    def __call__(self, m: int, x: torch.Tensor) -> torch.Tensor:
        """x: int32 token ids [B] (first stage) or hidden [B, d].  Returns hidden [B, d] or int32 tokens [B]."""
        ex = self.ex
        if ex.has_head:
            ex.step_host(x, self.pos[m], self.seq[m], next_token_out=self.tokens[m])
            out = self.tokens[m]
        else:
            ex.step_host(x, self.pos[m], self.seq[m], hidden_out=self.hidden_out[m])
            out = self.hidden_out[m]
        self.pos[m].add_(1)
        return out


class StagePipeline:
    """The static decode schedule of one rank."""

    def __init__(self, link: StageLink, compute: Callable[[int, torch.Tensor], torch.Tensor], batch: int,
                 n_micro: int, hidden: int, act_dtype: torch.dtype, device):
        if n_micro < 1:
            raise ValueError("need at least one micro-batch")
        self.link, self.compute = link, compute
        self.batch, self.n_micro = batch, n_micro
        self.hidden_in = [torch.empty(batch, hidden, dtype=act_dtype, device=device) for _ in range(n_micro)]
        self.tok_in = [torch.zeros(batch, dtype=torch.int32, device=device) for _ in range(n_micro)]
        self.tok_host = None  # pinned staging of the host-driven loop (first stage, allocated on first use)
        self.tok_dev = [torch.zeros(batch, dtype=torch.int32, device=device) for _ in range(n_micro)]

    def _via_host(self, m: int, x: torch.Tensor) -> torch.Tensor:
        """Token ids of micro-batch m: device -> pinned host (synchronous read, what a master does with every result)
        -> device again as the next step's input.  Only micro-batch m waits; the other micro-batches keep the
        stages busy meanwhile."""
        if self.tok_host is None:
            pin = torch.cuda.is_available()
            self.tok_host = [torch.zeros(self.batch, dtype=torch.int32, pin_memory=pin) for _ in range(self.n_micro)]
        self.tok_host[m].copy_(x)
        self.tok_dev[m].copy_(self.tok_host[m], non_blocking=True)
        return self.tok_dev[m]

    def decode(self, first_tokens: Optional[List[torch.Tensor]], n_steps: int,
               record: bool = False, via_host: bool = False) -> Optional[torch.Tensor]:
        """`first_tokens[m]`: int32 [B] for step 0 (first stage only).  With `record`, the first stage returns the
        int32 tokens [n_micro, B, n_steps] that came back around the ring (one extra ring transfer per step).
        With `via_host`, every token id the first stage receives is read back to pinned host memory and fed to the
        next step from there (the master-driven loop): same schedule, one host round trip per micro-batch step."""
        link = self.link
        log = [[] for _ in range(self.n_micro)] if (record and link.is_first) else None
        pending = [None] * self.n_micro  # the send of micro-batch m's previous output (its buffer is reused)
        for step in range(n_steps):
            for m in range(self.n_micro):
                if pending[m] is not None:
                    pending[m].wait()
                    pending[m] = None
                if link.is_first:
                    if step == 0:
                        x = first_tokens[m]
                    else:
                        x = link.recv_tokens(self.tok_in[m]) if link.world > 1 else self._loop_tok[m]
                        if log is not None:
                            log[m].append(x.clone())
                    if via_host:
                        x = self._via_host(m, x)
                else:
                    x = link.recv_hidden(self.hidden_in[m])
                out = self.compute(m, x)
                if link.is_last:
                    if link.world > 1:
                        pending[m] = link.send_tokens(out)
                    else:
                        self._loop_tok = getattr(self, "_loop_tok", [None] * self.n_micro)
                        self._loop_tok[m] = out.clone()
                else:
                    pending[m] = link.send_hidden(out)
        for w in pending:
            if w is not None:
                w.wait()
        # drain: the tokens of the final step are still travelling to the first stage
        if link.is_first:
            for m in range(self.n_micro):
                x = link.recv_tokens(self.tok_in[m]) if link.world > 1 else self._loop_tok[m]
                if via_host:
                    x = self._via_host(m, x)
                if log is not None:
                    log[m].append(x.clone())
        if log is not None:
            return torch.stack([torch.stack(l, dim=-1) for l in log])
        return None
