"""Host side of one pipeline stage: turns an HF-keyed parameter dict into a `dllm_shard` handle and
drives `dllm_shard_forward` / `dllm_shard_decode_loop` (include/dllm_b200.h).

This is what the body of the reference's `ModelShard.compute` (reference src/worker/node.py:24-32, a
per-key matmul placeholder) becomes: a real layer-shard forward with a resident paged KV cache.  torch
supplies device memory and the current stream; all arithmetic happens in libdllm_b200.so.
"""
from __future__ import annotations

import ctypes as C
import json
import threading
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import torch

from . import cabi

CONFIG_KEY = "__config__"  # uint8 tensor holding the HF config.json bytes, stored next to the weights


@dataclass
class StageGeometry:
    """What a stage needs from an HF ``config.json`` (GPT-2, Llama, Qwen2)."""
    model_type: str
    hidden: int
    n_layers: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    ffn: int
    vocab: int
    max_pos: int
    norm_eps: float
    rope_theta: float
    tie_head: bool

    @staticmethod
    def from_hf(c: dict) -> "StageGeometry":
        mt = c.get("model_type")
        if mt == "gpt2":
            d = c.get("n_embd", c.get("hidden_size"))
            h = c.get("n_head", c.get("num_attention_heads"))
            return StageGeometry("gpt2", d, c.get("n_layer", c.get("num_hidden_layers")), h, h, d // h,
                                 c.get("n_inner") or 4 * d, c["vocab_size"],
                                 c.get("n_positions", c.get("max_position_embeddings", 1024)),
                                 c.get("layer_norm_epsilon", 1e-5), 0.0, True)
        if mt in ("llama", "qwen2"):
            d, h = c["hidden_size"], c["num_attention_heads"]
            theta = c.get("rope_theta")
            if theta is None:
                theta = (c.get("rope_parameters") or {}).get("rope_theta", 10000.0)
            scaling = c.get("rope_scaling") or {}
            if scaling and scaling.get("rope_type", scaling.get("type", "default")) != "default":
                raise ValueError("rope scaling other than 'default' is not supported")
            return StageGeometry(mt, d, c["num_hidden_layers"], h, c.get("num_key_value_heads", h),
                                 c.get("head_dim") or d // h, c["intermediate_size"], c["vocab_size"],
                                 c.get("max_position_embeddings", 4096), c.get("rms_norm_eps", 1e-6),
                                 float(theta), bool(c.get("tie_word_embeddings", False)))
        raise ValueError(f"unsupported model_type {mt!r} (supported: gpt2, llama, qwen2)")


def encode_config(cfg: dict) -> torch.Tensor:
    return torch.tensor(list(json.dumps(cfg).encode()), dtype=torch.uint8)


def decode_config(t: torch.Tensor) -> dict:
    return json.loads(bytes(t.cpu().tolist()).decode())


def rope_tables(head_dim: int, theta: float, n_pos: int):
    """fp32 cos/sin [n_pos, head_dim/2] as HF computes them (default rope:
    HF:llama/modeling_llama.py:106-135): inv_freq = 1/theta^(2i/hd), freqs = pos * inv_freq."""
    inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).float() / head_dim))
    freqs = torch.arange(n_pos, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return freqs.cos().contiguous(), freqs.sin().contiguous()


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class ShardExecutor:
    """One `dllm_shard` handle plus the device tensors it points into."""

    def __init__(self, parameters: Dict[str, torch.Tensor], hf_config: dict, device: torch.device,
                 max_rows: int = 256, max_seqs: int = 8, max_context: Optional[int] = None,
                 page_size: int = 64):
        if cabi.device_count() == 0 or not torch.cuda.is_available():
            raise cabi.DllmError("no CUDA device: the shard forward pass has no CPU path")
        self.geom = g = StageGeometry.from_hf(hf_config)
        self.device = torch.device(device)
        self.lock = threading.Lock()  # a handle is not re-entrant (include/dllm_b200.h)
        prefix = "h." if g.model_type == "gpt2" else "layers."
        self.layer_ids: List[int] = sorted({int(k.split(".")[1]) for k in parameters if k.startswith(prefix)})
        if self.layer_ids and self.layer_ids != list(range(self.layer_ids[0], self.layer_ids[-1] + 1)):
            raise ValueError(f"a shard must hold a contiguous layer range, got {self.layer_ids}")
        emb_key = "wte.weight" if g.model_type == "gpt2" else "embed_tokens.weight"
        self.has_embed = emb_key in parameters
        self.has_head = "lm_head.weight" in parameters
        floats = [v for k, v in parameters.items() if k != CONFIG_KEY and v.is_floating_point()]
        src_dtype = floats[0].dtype if floats else torch.bfloat16
        # the reference loads fp16 on CUDA (src/model/loader.py:13,22); the B200 path computes in bf16
        self.dtype = torch.float32 if src_dtype == torch.float32 else torch.bfloat16
        self.max_rows, self.max_seqs, self.page_size = int(max_rows), int(max_seqs), int(page_size)
        self.max_context = int(max_context or min(g.max_pos, 2048))
        if g.model_type == "gpt2":
            self.max_context = min(self.max_context, g.max_pos)

        self._keep: List[torch.Tensor] = []  # everything the handle points into

        def dev(t: torch.Tensor, transpose: bool = False) -> torch.Tensor:
            t = t.to(self.device, self.dtype)
            t = t.t().contiguous() if transpose else t.contiguous()
            self._keep.append(t)
            return t

        P = parameters
        # A stage boundary may fall inside a layer, between its attention block and its MLP block: the first layer of a
        # stage may then hold only the MLP block, the last one only the attention block (include/dllm_b200.h).
        attn_key = (lambda i: f"h.{i}.attn.c_attn.weight") if g.model_type == "gpt2" else (lambda i: f"layers.{i}.self_attn.q_proj.weight")
        mlp_key = (lambda i: f"h.{i}.mlp.c_proj.weight") if g.model_type == "gpt2" else (lambda i: f"layers.{i}.mlp.down_proj.weight")
        self.first_layer_mlp_only = bool(self.layer_ids) and attn_key(self.layer_ids[0]) not in P
        self.last_layer_attn_only = bool(self.layer_ids) and mlp_key(self.layer_ids[-1]) not in P
        for i in self.layer_ids:
            whole = attn_key(i) in P and mlp_key(i) in P
            if not whole and not ((i == self.layer_ids[0] and self.first_layer_mlp_only and mlp_key(i) in P) or
                                  (i == self.layer_ids[-1] and self.last_layer_attn_only and attn_key(i) in P)):
                raise ValueError(f"layer {i} is incomplete: only the first layer of a stage may lack its attention block "
                                 "and only the last one its MLP block")
        layers = (cabi.LayerWeights * max(1, len(self.layer_ids)))()
        for n, i in enumerate(self.layer_ids):
            lw = layers[n]
            has_attn = not (n == 0 and self.first_layer_mlp_only)
            has_mlp = not (n == len(self.layer_ids) - 1 and self.last_layer_attn_only)
            if g.model_type == "gpt2":
                p = f"h.{i}."
                d = g.hidden
                vals = {}
                if has_attn:
                    wqkv = dev(P[p + "attn.c_attn.weight"], transpose=True)  # Conv1D [in,out] -> [3d, d]
                    bqkv = dev(P[p + "attn.c_attn.bias"])
                    esz = wqkv.element_size()
                    vals.update(
                        norm1_w=_ptr(dev(P[p + "ln_1.weight"])), norm1_b=_ptr(dev(P[p + "ln_1.bias"])),
                        wq=wqkv.data_ptr(), wk=wqkv.data_ptr() + d * d * esz, wv=wqkv.data_ptr() + 2 * d * d * esz,
                        bq=bqkv.data_ptr(), bk=bqkv.data_ptr() + d * esz, bv=bqkv.data_ptr() + 2 * d * esz,
                        wo=_ptr(dev(P[p + "attn.c_proj.weight"], True)), bo=_ptr(dev(P[p + "attn.c_proj.bias"])))
                if has_mlp:
                    vals.update(
                        norm2_w=_ptr(dev(P[p + "ln_2.weight"])), norm2_b=_ptr(dev(P[p + "ln_2.bias"])),
                        w_gate=None, w_up=_ptr(dev(P[p + "mlp.c_fc.weight"], True)), b_up=_ptr(dev(P[p + "mlp.c_fc.bias"])),
                        w_down=_ptr(dev(P[p + "mlp.c_proj.weight"], True)), b_down=_ptr(dev(P[p + "mlp.c_proj.bias"])))
            else:
                p = f"layers.{i}."
                opt = lambda k: _ptr(dev(P[k])) if k in P else None  # noqa: E731
                vals = {}
                if has_attn:
                    vals.update(
                        norm1_w=_ptr(dev(P[p + "input_layernorm.weight"])), norm1_b=None,
                        wq=_ptr(dev(P[p + "self_attn.q_proj.weight"])), bq=opt(p + "self_attn.q_proj.bias"),
                        wk=_ptr(dev(P[p + "self_attn.k_proj.weight"])), bk=opt(p + "self_attn.k_proj.bias"),
                        wv=_ptr(dev(P[p + "self_attn.v_proj.weight"])), bv=opt(p + "self_attn.v_proj.bias"),
                        wo=_ptr(dev(P[p + "self_attn.o_proj.weight"])), bo=None)
                if has_mlp:
                    vals.update(
                        norm2_w=_ptr(dev(P[p + "post_attention_layernorm.weight"])), norm2_b=None,
                        w_gate=_ptr(dev(P[p + "mlp.gate_proj.weight"])),
                        w_up=_ptr(dev(P[p + "mlp.up_proj.weight"])), b_up=None,
                        w_down=_ptr(dev(P[p + "mlp.down_proj.weight"])), b_down=None)
            for k, v in vals.items():
                setattr(lw, k, v)

        io = cabi.IoWeights()
        if self.has_embed:
            io.wte = _ptr(dev(P[emb_key]))
            if g.model_type == "gpt2":
                io.wpe = _ptr(dev(P["wpe.weight"]))
        if self.has_head:
            if g.model_type == "gpt2":
                io.final_norm_w, io.final_norm_b = _ptr(dev(P["ln_f.weight"])), _ptr(dev(P["ln_f.bias"]))
            else:
                io.final_norm_w = _ptr(dev(P["norm.weight"]))
            io.lm_head = _ptr(dev(P["lm_head.weight"]))
        if g.model_type != "gpt2":
            cos, sin = rope_tables(g.head_dim, g.rope_theta, self.max_context)
            cos, sin = cos.to(self.device), sin.to(self.device)
            self._keep += [cos, sin]
            io.rope_cos, io.rope_sin = cos.data_ptr(), sin.data_ptr()

        cfg = cabi.ShardConfig(
            arch=cabi.ARCH_IDS[g.model_type], dtype=cabi.DTYPE_F32 if self.dtype == torch.float32 else cabi.DTYPE_BF16,
            hidden=g.hidden, n_heads=g.n_heads, n_kv_heads=g.n_kv_heads, head_dim=g.head_dim, ffn=g.ffn,
            vocab=g.vocab, n_layers=len(self.layer_ids), has_embed=int(self.has_embed), has_head=int(self.has_head),
            norm_eps=g.norm_eps, max_rows=self.max_rows, max_seqs=self.max_seqs, max_context=self.max_context,
            page_size=self.page_size, num_pages=0, first_layer_mlp_only=int(self.first_layer_mlp_only),
            last_layer_attn_only=int(self.last_layer_attn_only))
        handle = C.c_void_p()
        with torch.cuda.device(self.device):
            cabi.check(cabi.lib().dllm_shard_create(C.byref(cfg), layers, C.byref(io), C.byref(handle)))
        self._handle = handle
        self._seq_len: Dict[int, int] = {}

    # ---- life cycle ---------------------------------------------------------------------------------
    def close(self) -> None:
        h, self._handle = getattr(self, "_handle", None), None
        if h:
            cabi.lib().dllm_shard_destroy(h)
        self._keep = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def device_bytes(self) -> int:
        return int(cabi.lib().dllm_shard_device_bytes(self._handle))

    def reset(self, seq_ids: Optional[Sequence[int]] = None) -> None:
        """Drop the KV state of the given sequence slots (all when None)."""
        with self.lock:
            if seq_ids is None:
                cabi.check(cabi.lib().dllm_shard_kv_release(self._handle, -1))
                self._seq_len.clear()
            else:
                for s in seq_ids:
                    cabi.check(cabi.lib().dllm_shard_kv_release(self._handle, int(s)))
                    self._seq_len.pop(int(s), None)

    def seq_len(self, seq_id: int) -> int:
        return self._seq_len.get(int(seq_id), 0)

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def reserve(self, seq_id: int, n_tokens: int) -> None:
        cabi.check(cabi.lib().dllm_shard_kv_reserve(self._handle, int(seq_id), int(n_tokens), self._stream()))

    # ---- the hot path ----------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, positions: torch.Tensor, seq_ids: Sequence[int],
                want_logits: bool = True) -> Dict[str, torch.Tensor]:
        """x: token ids [B,T] (stage with embeddings) or hidden states [B,T,d]; positions int [B,T];
        seq_ids: B sequence slots.  Long prompts are fed in chunks of <= max_rows rows (chunked prefill)."""
        g = self.geom
        B, T = int(x.shape[0]), int(x.shape[1])
        if len(seq_ids) != B:
            raise ValueError("seq_ids must have one entry per batch row")
        pos_host = positions.to("cpu", torch.int32).reshape(B, T).contiguous()
        seq_host = torch.tensor([int(s) for s in seq_ids], dtype=torch.int32)
        if int(pos_host.min()) < 0 or int(pos_host.max()) >= self.max_context:
            raise ValueError(f"position out of range [0, {self.max_context})")
        if int(seq_host.min()) < 0 or int(seq_host.max()) >= self.max_seqs:
            raise ValueError(f"seq_id out of range [0, {self.max_seqs})")
        if B > self.max_rows:
            raise ValueError(f"batch {B} exceeds max_rows {self.max_rows}")
        if self.has_embed:
            xin = x.to(self.device, torch.int32).reshape(B, T)
        else:
            xin = x.to(self.device, self.dtype).reshape(B, T, g.hidden)
        hidden_out = None if self.has_head else torch.empty(B, T, g.hidden, dtype=self.dtype, device=self.device)
        logits = next_tok = None
        if self.has_head and want_logits:
            logits = torch.empty(B, g.vocab, dtype=torch.float32, device=self.device)
            next_tok = torch.empty(B, dtype=torch.int32, device=self.device)
        chunk_t = max(1, self.max_rows // B)
        L = cabi.lib()
        with self.lock, torch.cuda.device(self.device):
            st = self._stream()
            for b in range(B):
                need = int(pos_host[b].max()) + 1
                cabi.check(L.dllm_shard_kv_reserve(self._handle, int(seq_host[b]), need, st))
                self._seq_len[int(seq_host[b])] = max(self._seq_len.get(int(seq_host[b]), 0), need)
            for t0 in range(0, T, chunk_t):
                t1 = min(T, t0 + chunk_t)
                n = t1 - t0
                last = t1 == T
                xc = xin[:, t0:t1].contiguous()
                pc = pos_host[:, t0:t1].contiguous()
                sc = seq_host[:, None].expand(B, n).contiguous()
                hc = None
                if hidden_out is not None:
                    hc = hidden_out if n == T else torch.empty(B, n, g.hidden, dtype=self.dtype, device=self.device)
                lr = None
                emit = self.has_head and last and want_logits
                if emit and n > 1:  # logits of the last position of every row (n == 1: every row, the default)
                    lr = (torch.arange(B, dtype=torch.int32) * n + (n - 1)).contiguous()
                cabi.check(L.dllm_shard_forward(
                    self._handle, xc.data_ptr(), pc.data_ptr(), sc.data_ptr(), B * n, _ptr(hc),
                    _ptr(lr), B if lr is not None else (0 if emit or not self.has_head else -1),
                    _ptr(logits) if emit else None, _ptr(next_tok) if emit else None, st))
                if hc is not None and hc is not hidden_out:
                    hidden_out[:, t0:t1] = hc
        self._last_seqs = [int(v) for v in seq_host]
        if self.has_head:
            out = {}
            if logits is not None:
                out["logits"] = logits
                out["next_token"] = next_tok.to(torch.int64)
            return out
        return {"hidden_states": hidden_out}

    def step_host(self, x: torch.Tensor, positions: torch.Tensor, seq_ids: torch.Tensor,
                  hidden_out: Optional[torch.Tensor] = None, next_token_out: Optional[torch.Tensor] = None,
                  logits_out: Optional[torch.Tensor] = None) -> None:
        """One `dllm_shard_forward` call on caller-owned buffers, host (pinned) or device: x = int32 ids [R]
        or hidden [R,d]; positions / seq_ids int32 [R].  Nothing is allocated; the copies in and out are
        part of the call (this is the end-to-end path a master's decode loop drives).  KV pages must have
        been reserved (`reserve`)."""
        n = int(positions.numel())
        with self.lock, torch.cuda.device(self.device):
            cabi.check(cabi.lib().dllm_shard_forward(
                self._handle, x.data_ptr(), positions.data_ptr(), seq_ids.data_ptr(), n, _ptr(hidden_out),
                None, 0, _ptr(logits_out), _ptr(next_token_out), self._stream()))

    def decode_loop(self, batch: int, n_steps: int) -> torch.Tensor:
        """`n_steps` greedy steps on the device for the `batch` sequences of the last forward() call
        (single-stage models only).  Returns int64 token ids [batch, n_steps]."""
        if n_steps <= 0:
            return torch.empty(batch, 0, dtype=torch.int64, device=self.device)
        toks = torch.empty(n_steps, batch, dtype=torch.int32, device=self.device)
        with self.lock, torch.cuda.device(self.device):
            st = self._stream()
            if batch != len(self._last_seqs):
                raise ValueError("decode_loop continues the batch of the last forward() call")
            for s in self._last_seqs:
                if self._seq_len[s] + n_steps > self.max_context:
                    raise ValueError("decode_loop would run past max_context")
            for s in self._last_seqs:
                cabi.check(cabi.lib().dllm_shard_kv_reserve(self._handle, s, self._seq_len[s] + n_steps, st))
                self._seq_len[s] += n_steps
            cabi.check(cabi.lib().dllm_shard_decode_loop(self._handle, batch, n_steps, toks.data_ptr(), st))
        return toks.t().to(torch.int64)

    def check_decode_pipelined(self, inbox, outbox, n_micro: int, rows: int, n_steps: int, positions: torch.Tensor,
                               seq_ids: torch.Tensor) -> None:
        """Pre-flight of `decode_pipelined`: raises what the launch would raise (shape the persistent kernel does not
        take, mailbox mismatch, context limit, KV pool exhausted) -- and reserves the KV pages -- without launching."""
        pos = positions.to("cpu", torch.int32).reshape(n_micro * rows)
        seq = seq_ids.to("cpu", torch.int32).reshape(n_micro * rows)
        if int(pos.max()) + n_steps > self.max_context:
            raise ValueError("decode_pipelined would run past max_context")
        with self.lock, torch.cuda.device(self.device):
            cabi.check(cabi.lib().dllm_shard_decode_pipelined_check(self._handle, inbox.handle, outbox.handle, n_micro, rows, n_steps))
            st = self._stream()
            for s_, p_ in zip(seq.tolist(), pos.tolist()):
                cabi.check(cabi.lib().dllm_shard_kv_reserve(self._handle, s_, p_ + n_steps, st))

    def pipeline_status(self) -> None:
        """Waits for the pipelined loops enqueued so far and raises `DllmError` if one of them was abandoned because a
        hand-off ticket did not arrive in time (`dllm_shard_pipeline_status`); the context stays usable either way."""
        with self.lock, torch.cuda.device(self.device):
            cabi.check(cabi.lib().dllm_shard_pipeline_status(self._handle, self._stream()))

    def decode_pipelined(self, inbox, outbox, n_micro: int, rows: int, n_steps: int, step_base: int,
                         positions: torch.Tensor, seq_ids: torch.Tensor) -> Optional[torch.Tensor]:
        """`n_steps` decode steps of `n_micro` micro-batches in ONE persistent launch with the hand-off fused into
        the kernel (`dllm_shard_decode_pipelined`; mailboxes from `src/network/handoff.py`).  `positions` /
        `seq_ids`: int32 [n_micro, rows] -- position of the token fed at the first step, KV slot of each row.
        Asynchronous on the current stream.  The first stage returns the int32 ids it fed, [n_steps, n_micro, rows]."""
        pos = positions.to("cpu", torch.int32).reshape(n_micro * rows).contiguous()
        seq = seq_ids.to("cpu", torch.int32).reshape(n_micro * rows).contiguous()
        if int(pos.max()) + n_steps > self.max_context:
            raise ValueError("decode_pipelined would run past max_context")
        fed = torch.empty(n_steps, n_micro, rows, dtype=torch.int32, device=self.device) if self.has_embed else None
        with self.lock, torch.cuda.device(self.device):
            st = self._stream()
            for s_, p_ in zip(seq.tolist(), pos.tolist()):
                cabi.check(cabi.lib().dllm_shard_kv_reserve(self._handle, s_, p_ + n_steps, st))
                self._seq_len[s_] = max(self._seq_len.get(s_, 0), p_ + n_steps)
            cabi.check(cabi.lib().dllm_shard_decode_pipelined(
                self._handle, inbox.handle, outbox.handle, n_micro, rows, n_steps, step_base, pos.data_ptr(),
                seq.data_ptr(), fed.data_ptr() if fed is not None else None, st))
        return fed

    _last_seqs: List[int] = []
