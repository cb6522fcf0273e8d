"""Seeded random-init checkpoints of the BASELINE.json model list (no network here: there are no real
checkpoints to load).  Per-tensor seeds derive from (base_seed, key), so a stage can create only its own
layers -- a 70B stage never materialises the other 70 layers.  HF default initialisation: normal(0, 0.02)
matrices, unit norm weights, zero biases."""
from __future__ import annotations

import zlib
from typing import Dict, Optional

import torch

NAMED_CONFIGS: Dict[str, dict] = {
    "gpt2-small": dict(model_type="gpt2", n_embd=768, n_layer=12, n_head=12, n_inner=3072, vocab_size=50257,
                       n_positions=1024, layer_norm_epsilon=1e-5, tie_word_embeddings=True),
    "tinyllama-1.1b": dict(model_type="llama", hidden_size=2048, num_hidden_layers=22, num_attention_heads=32,
                           num_key_value_heads=4, intermediate_size=5632, vocab_size=32000,
                           max_position_embeddings=2048, rms_norm_eps=1e-5, rope_theta=10000.0),
    "llama3-8b": dict(model_type="llama", hidden_size=4096, num_hidden_layers=32, num_attention_heads=32,
                      num_key_value_heads=8, intermediate_size=14336, vocab_size=128256,
                      max_position_embeddings=8192, rms_norm_eps=1e-5, rope_theta=500000.0),
    "qwen2.5-14b": dict(model_type="qwen2", hidden_size=5120, num_hidden_layers=48, num_attention_heads=40,
                        num_key_value_heads=8, intermediate_size=13824, vocab_size=152064,
                        max_position_embeddings=32768, rms_norm_eps=1e-6, rope_theta=1000000.0),
    "llama3-70b": dict(model_type="llama", hidden_size=8192, num_hidden_layers=80, num_attention_heads=64,
                       num_key_value_heads=8, intermediate_size=28672, vocab_size=128256,
                       max_position_embeddings=8192, rms_norm_eps=1e-5, rope_theta=500000.0),
}


def key_seed(base_seed: int, key: str) -> int:
    return (zlib.crc32(key.encode()) ^ (base_seed * 0x9E3779B1)) & 0x7FFFFFFF


def layer_shapes(cfg: dict, i: int) -> Dict[str, tuple]:
    if cfg["model_type"] == "gpt2":
        d, f, p = cfg["n_embd"], cfg.get("n_inner") or 4 * cfg["n_embd"], f"h.{i}."
        return {p + "ln_1.weight": ("norm", (d,)), p + "ln_1.bias": ("bias", (d,)),
                p + "attn.c_attn.weight": ("w", (d, 3 * d)), p + "attn.c_attn.bias": ("bias", (3 * d,)),
                p + "attn.c_proj.weight": ("w", (d, d)), p + "attn.c_proj.bias": ("bias", (d,)),
                p + "ln_2.weight": ("norm", (d,)), p + "ln_2.bias": ("bias", (d,)),
                p + "mlp.c_fc.weight": ("w", (d, f)), p + "mlp.c_fc.bias": ("bias", (f,)),
                p + "mlp.c_proj.weight": ("w", (f, d)), p + "mlp.c_proj.bias": ("bias", (d,))}
    d, f, H = cfg["hidden_size"], cfg["intermediate_size"], cfg["num_attention_heads"]
    KV = cfg.get("num_key_value_heads", H)
    hd = cfg.get("head_dim") or d // H
    p = f"layers.{i}."
    out = {p + "input_layernorm.weight": ("norm", (d,)),
           p + "self_attn.q_proj.weight": ("w", (H * hd, d)), p + "self_attn.k_proj.weight": ("w", (KV * hd, d)),
           p + "self_attn.v_proj.weight": ("w", (KV * hd, d)), p + "self_attn.o_proj.weight": ("w", (d, H * hd)),
           p + "post_attention_layernorm.weight": ("norm", (d,)),
           p + "mlp.gate_proj.weight": ("w", (f, d)), p + "mlp.up_proj.weight": ("w", (f, d)),
           p + "mlp.down_proj.weight": ("w", (d, f))}
    if cfg["model_type"] == "qwen2":
        out.update({p + "self_attn.q_proj.bias": ("bias", (H * hd,)), p + "self_attn.k_proj.bias": ("bias", (KV * hd,)),
                    p + "self_attn.v_proj.bias": ("bias", (KV * hd,))})
    return out


def end_shapes(cfg: dict, first: bool, last: bool) -> Dict[str, tuple]:
    out: Dict[str, tuple] = {}
    if cfg["model_type"] == "gpt2":
        d = cfg["n_embd"]
        if first:
            out.update({"wte.weight": ("w", (cfg["vocab_size"], d)), "wpe.weight": ("w", (cfg["n_positions"], d))})
        if last:
            out.update({"ln_f.weight": ("norm", (d,)), "ln_f.bias": ("bias", (d,))})
    else:
        d = cfg["hidden_size"]
        if first:
            out["embed_tokens.weight"] = ("w", (cfg["vocab_size"], d))
        if last:
            out["norm.weight"] = ("norm", (d,))
            if not cfg.get("tie_word_embeddings", False):
                out["lm_head.weight"] = ("w", (cfg["vocab_size"], d))
    return out


def block_of(key: str) -> str:
    """"attn" or "mlp": which block of its decoder layer a tensor belongs to (norms go with the block they feed)."""
    k = key.split(".", 2)[2]
    return "attn" if k.startswith(("ln_1", "attn", "input_layernorm", "self_attn")) else "mlp"


def random_stage_parameters(cfg: dict, layers, first: bool, last: bool, base_seed: int = 1234,
                            dtype: torch.dtype = torch.bfloat16, device: Optional[str] = None,
                            init: str = "hf") -> Dict[str, torch.Tensor]:
    """Parameters of one stage.  `layers`: layer indices, or (layer, block) pairs with block in "full" / "attn" /
    "mlp" for a stage that starts or ends inside a layer.  With device=None tensors are drawn on the CPU from per-key
    seeded generators (bit-reproducible: what parity tests use).  With a CUDA device they are drawn on that
    device (fast; used for throughput runs on 8B-70B models, where values only need the right scale)."""
    shapes: Dict[str, tuple] = dict(end_shapes(cfg, first, False))
    for item in layers:
        i, block = item if isinstance(item, tuple) else (item, "full")
        shapes.update({k: v for k, v in layer_shapes(cfg, i).items() if block == "full" or block_of(k) == block})
    shapes.update(end_shapes(cfg, False, last))
    out: Dict[str, torch.Tensor] = {}
    on_gpu = device is not None and str(device).startswith("cuda")
    for k, (kind, shape) in shapes.items():
        if on_gpu:
            g = torch.Generator(device=device).manual_seed(key_seed(base_seed, k))
            if kind == "w":
                t = torch.empty(shape, dtype=dtype, device=device).normal_(0.0, 0.02, generator=g)
            elif kind == "norm":
                t = torch.ones(shape, dtype=dtype, device=device)
            else:
                t = torch.zeros(shape, dtype=dtype, device=device)
        else:
            g = torch.Generator().manual_seed(key_seed(base_seed, k))
            if kind == "w":
                t = torch.randn(shape, generator=g, dtype=torch.float32) * 0.02
            elif kind == "norm":
                t = torch.ones(shape, dtype=torch.float32)
                if init == "full":
                    t = t + 0.1 * torch.randn(shape, generator=g, dtype=torch.float32)
            else:
                t = torch.zeros(shape, dtype=torch.float32)
                if init == "full":
                    t = 0.02 * torch.randn(shape, generator=g, dtype=torch.float32)
            t = t.to(dtype)
        out[k] = t
    if last and "lm_head.weight" not in out:  # tied head
        emb = "wte.weight" if cfg["model_type"] == "gpt2" else "embed_tokens.weight"
        if emb in out:
            out["lm_head.weight"] = out[emb]
        else:
            kind, shape = end_shapes(cfg, True, False)[emb]
            out["lm_head.weight"] = random_stage_parameters(cfg, range(0), True, False, base_seed, dtype, device, init)[emb]
    return out
