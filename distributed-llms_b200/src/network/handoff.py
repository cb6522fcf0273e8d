"""Stage-to-stage hand-off of the hidden-state tile: the data plane that replaces pickled tensors over
`MessageProtocol.send_message / receive_message` (reference src/network/protocol.py:38-129; reference
src/worker/node.py:224,236 pickles whole tensor dicts) inside one NVLink box.

One process per GPU; adjacent pipeline stages exchange the bf16 `[B, T, d]` tile with NCCL send/recv over NVLink
(8 KiB ... 20 MiB) and the last stage returns `[B]` int32 token ids to the first.  KV never
moves; positions and sequence ids are implied by the schedule both sides run.  On a machine without GPUs the
same code runs over gloo with CPU tensors (tests of the host logic).
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist


def init_stage_group(backend: Optional[str] = None) -> tuple[int, int]:
    """Join the process group torchrun set up (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT).  Returns (rank, world)."""
    if not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group(backend=backend, rank=int(os.environ.get("RANK", "0")),
                                world_size=int(os.environ.get("WORLD_SIZE", "1")))
    return dist.get_rank(), dist.get_world_size()


class StageLink:
    """Point-to-point links of one pipeline stage: previous stage, next stage, and the token ring
    (last stage -> first stage).

    Sends are asynchronous (`isend`): a stage never blocks in a send while its neighbour is itself blocked in a
    send to it -- with blocking sends the ring (stage 0 -> 1 -> ... -> S-1 -> 0) deadlocks as soon as more than
    one micro-batch is in flight.  The token ring uses its own process group (its own NCCL communicator and
    stream), so that a token waiting for stage 0 never holds up the hidden-state traffic behind it."""

    def __init__(self, rank: int, world: int, make_ring_group: bool = True):
        self.rank, self.world = rank, world
        self.prev = rank - 1 if rank > 0 else None
        self.next = rank + 1 if rank < world - 1 else None
        self.is_first, self.is_last = rank == 0, rank == world - 1
        self.ring_group = None
        if make_ring_group and world > 1 and dist.is_initialized():
            self.ring_group = dist.new_group(ranks=list(range(world)))  # collective: every rank constructs a link

    # hidden-state tile ------------------------------------------------------------------------------
    def send_hidden(self, tile: torch.Tensor):
        """Returns the pending work (wait on it before overwriting `tile`), or None."""
        return dist.isend(tile, dst=self.next) if self.next is not None else None

    def recv_hidden(self, buf: torch.Tensor) -> torch.Tensor:
        if self.prev is not None:
            dist.recv(buf, src=self.prev)
        return buf

    # token ring -------------------------------------------------------------------------------------
    def send_tokens(self, tokens: torch.Tensor):
        """Last stage -> first stage (no-op for a single-stage pipeline)."""
        if self.is_last and self.world > 1:
            return dist.isend(tokens, dst=0, group=self.ring_group)
        return None

    def recv_tokens(self, buf: torch.Tensor) -> torch.Tensor:
        if self.is_first and self.world > 1:
            dist.recv(buf, src=self.world - 1, group=self.ring_group)
        return buf


# ---- fused hand-off: mailboxes in peer memory ----------------------------------------------------------------

class Mailbox:
    """A stage's inbox (owner side) or a producer's view of another stage's inbox (`dllm_mailbox_*`,
    include/dllm_b200.h).  Per slot (micro-batch): a ticket, `rows` token ids, `rows x hidden` bf16 rows."""

    def __init__(self, handle, device: int, n_slots: int, rows: int, hidden: int, owner: bool):
        self.handle, self.device, self.n_slots, self.rows, self.hidden, self.owner = handle, device, n_slots, rows, hidden, owner

    @classmethod
    def create(cls, device: int, n_slots: int, rows: int, hidden: int) -> "Mailbox":
        import ctypes as C
        from ..engine import cabi
        h = C.c_void_p()
        cabi.check(cabi.lib().dllm_mailbox_create(device, n_slots, rows, hidden, C.byref(h)))
        box = cls(h, device, n_slots, rows, hidden, True)
        _LOCAL_MAILBOXES[box.export_handle()] = box
        return box

    def export_handle(self) -> bytes:
        import ctypes as C
        from ..engine import cabi
        buf = C.create_string_buffer(64)
        cabi.check(cabi.lib().dllm_mailbox_export(self.handle, buf))
        return buf.raw

    @classmethod
    def open_peer(cls, device: int, ipc: bytes, n_slots: int, rows: int, hidden: int) -> "Mailbox":
        """Another process's mailbox, mapped on `device` (cudaIpcOpenMemHandle)."""
        import ctypes as C
        from ..engine import cabi
        h = C.c_void_p()
        cabi.check(cabi.lib().dllm_mailbox_import(device, C.create_string_buffer(ipc, 64), n_slots, rows, hidden, C.byref(h)))
        return cls(h, device, n_slots, rows, hidden, False)

    def view_from(self, device: int) -> "Mailbox":
        """Same process, other GPU."""
        import ctypes as C
        from ..engine import cabi
        h = C.c_void_p()
        cabi.check(cabi.lib().dllm_mailbox_view(self.handle, device, C.byref(h)))
        return Mailbox(h, device, self.n_slots, self.rows, self.hidden, False)

    def post_tokens(self, slot: int, tokens: torch.Tensor, ticket: int) -> None:
        from ..engine import cabi
        t = tokens.to(torch.int32).contiguous()
        st = torch.cuda.current_stream(self.device).cuda_stream
        cabi.check(cabi.lib().dllm_mailbox_post_tokens(self.handle, slot, t.data_ptr(), ticket, st))

    def read_tokens(self, slot: int) -> tuple[torch.Tensor, int]:
        import ctypes as C
        from ..engine import cabi
        out = torch.empty(self.rows, dtype=torch.int32)
        ticket = C.c_uint32(0)
        st = torch.cuda.current_stream(self.device).cuda_stream
        cabi.check(cabi.lib().dllm_mailbox_read_tokens(self.handle, slot, out.data_ptr(), C.byref(ticket), st))
        return out, int(ticket.value)

    def put_rows(self, slot: int, rows: torch.Tensor) -> None:
        """Producer side: copy a [n, hidden] bf16 device tile into `slot` of this (usually another stage's) mailbox."""
        from ..engine import cabi
        t = rows.reshape(-1, self.hidden).contiguous()
        st = torch.cuda.current_stream(t.device).cuda_stream
        cabi.check(cabi.lib().dllm_mailbox_put_rows(self.handle, slot, t.data_ptr(), int(t.shape[0]), st))

    def get_rows(self, slot: int, n_rows: int, dtype=torch.bfloat16) -> torch.Tensor:
        """Owner side: the first `n_rows` rows of `slot` as a fresh [n_rows, hidden] device tensor."""
        from ..engine import cabi
        out = torch.empty(n_rows, self.hidden, dtype=dtype, device=torch.device("cuda", self.device))
        st = torch.cuda.current_stream(out.device).cuda_stream
        cabi.check(cabi.lib().dllm_mailbox_get_rows(self.handle, slot, out.data_ptr(), n_rows, st))
        return out

    def close(self) -> None:
        if self.handle is not None:
            from ..engine import cabi
            if self.owner:
                for k in [k for k, v in _LOCAL_MAILBOXES.items() if v is self]:
                    del _LOCAL_MAILBOXES[k]
            cabi.lib().dllm_mailbox_destroy(self.handle)
            self.handle = None


# Mailboxes created in this process, by IPC handle: a worker in the same process (threads, tests) cannot open its
# neighbour's handle through CUDA IPC and takes a peer view of the object instead.
_LOCAL_MAILBOXES: dict = {}


def connect_mailbox(device: int, info: dict) -> Mailbox:
    """The producer's side of `info` = {"ipc", "pid", "n_slots", "rows", "hidden"} (what a MAILBOX_OPEN reply carries)."""
    local = _LOCAL_MAILBOXES.get(info["ipc"]) if info.get("pid") == os.getpid() else None
    if local is not None:
        return local if local.device == device else local.view_from(device)
    return Mailbox.open_peer(device, info["ipc"], info["n_slots"], info["rows"], info["hidden"])


def exchange_mailboxes(rank: int, world: int, device: int, n_slots: int, rows: int, hidden: int):
    """Every stage creates its inbox and opens the next stage's (the first stage's for the last stage).  Collective:
    the 64-byte IPC handles travel through `all_gather` of the existing process group.  Returns (inbox, outbox)."""
    inbox = Mailbox.create(device, n_slots, rows, hidden)
    if world == 1:
        return inbox, inbox
    dev = torch.device("cuda", device)
    mine = torch.frombuffer(bytearray(inbox.export_handle()), dtype=torch.uint8).to(dev)
    everyone = [torch.empty(64, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(everyone, mine)
    nxt = (rank + 1) % world
    outbox = Mailbox.open_peer(device, bytes(everyone[nxt].cpu().numpy().tobytes()), n_slots, rows, hidden)
    return inbox, outbox
