"""Pin the CPU oracle against HF ``transformers`` random-init models (test infrastructure).

The reference computes no transformer; the algorithm lives in the third-party
``transformers`` package its loader imports (/root/reference/src/model/loader.py:10,19; unpinned,
5.5.0 in this image).  This script loads the oracle's seeded state dict into the stock HF classes
(eager attention) and checks, for GPT-2, Llama (GQA) and Qwen2 (q/k/v bias):

* fp32: last-position logits of prefill and of every decode step within 1e-4 relative, and greedy
  tokens identical to ``generate(do_sample=False)``, with the model split into 1, 2 and 3 shards;
* bf16: logits within 1e-2 relative of HF's bf16 forward, greedy tokens compared (reported).

Run:  python oracle/validate_against_hf.py        (CPU, ~1 min)
"""
from __future__ import annotations

import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.shard_oracle import (ShardModelConfig, build_pipeline, greedy_generate,  # noqa: E402
                                 make_prompts, make_state_dict)


def hf_model(cfg: ShardModelConfig, sd, dtype):
    import transformers as tf
    if cfg.model_type == "gpt2":
        hc = tf.GPT2Config(vocab_size=cfg.vocab, n_positions=cfg.max_pos, n_embd=cfg.hidden,
                           n_layer=cfg.n_layers, n_head=cfg.n_heads, n_inner=cfg.ffn,
                           layer_norm_epsilon=cfg.norm_eps, activation_function="gelu_new",
                           resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
        hc._attn_implementation = "eager"
        m = tf.GPT2LMHeadModel(hc)
        full = {"transformer." + k: v for k, v in sd.items()}
        full["lm_head.weight"] = sd["wte.weight"]
    else:
        kw = dict(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.ffn,
                  num_hidden_layers=cfg.n_layers, num_attention_heads=cfg.n_heads,
                  num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim,
                  max_position_embeddings=cfg.max_pos, rms_norm_eps=cfg.norm_eps,
                  rope_theta=cfg.rope_theta, tie_word_embeddings=False, attention_dropout=0.0)
        if cfg.model_type == "llama":
            hc = tf.LlamaConfig(**kw)
            hc._attn_implementation = "eager"
            m = tf.LlamaForCausalLM(hc)
        else:
            hc = tf.Qwen2Config(**kw)
            hc._attn_implementation = "eager"
            m = tf.Qwen2ForCausalLM(hc)
        full = {("model." + k if k != "lm_head.weight" else k): v for k, v in sd.items()}
    missing, unexpected = m.load_state_dict(full, strict=False)
    assert not unexpected, unexpected
    assert all("rotary" in k or "attn.bias" in k or "masked_bias" in k for k in missing), missing
    return m.to(dtype).eval()


def hf_greedy(m, prompts, new_tokens):
    """Manual greedy loop with DynamicCache -> tokens [B,n] and last-position logits [B,n,V]."""
    toks, logits = [], []
    with torch.no_grad():
        out = m(input_ids=prompts, use_cache=True)
        past = out.past_key_values
        lg = out.logits[:, -1, :]
        for _ in range(new_tokens):
            logits.append(lg.float())
            t = torch.argmax(lg, dim=-1)
            toks.append(t)
            out = m(input_ids=t[:, None], past_key_values=past, use_cache=True)
            past = out.past_key_values
            lg = out.logits[:, -1, :]
    return torch.stack(toks, 1), torch.stack(logits, 1)


CASES = {
    "gpt2": ShardModelConfig("gpt2", 64, 4, 4, 4, 16, 256, 211, 64, 1e-5, tie_head=True),
    "llama": ShardModelConfig("llama", 128, 4, 4, 2, 32, 352, 257, 96, 1e-5, 10000.0),
    "qwen2": ShardModelConfig("qwen2", 128, 3, 8, 2, 16, 192, 300, 96, 1e-6, 1000000.0),
}


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main() -> int:
    torch.manual_seed(0)
    ok = True
    for name, cfg in CASES.items():
        sd32 = make_state_dict(cfg, 1234, torch.float32, init="full")
        prompts = make_prompts(cfg.vocab, 2, 9, seed=7)
        n_new = 12
        m = hf_model(cfg, sd32, torch.float32)
        ref_t, ref_l = hf_greedy(m, prompts, n_new)
        gen = m.generate(prompts, max_new_tokens=n_new, do_sample=False, pad_token_id=0)[:, prompts.shape[1]:]
        assert torch.equal(gen, ref_t), "manual HF loop != generate()"
        for S in (1, 2, 3):
            stages = build_pipeline(cfg, sd32, S, torch.float32)
            ids, lg, mg = greedy_generate(stages, prompts, n_new, record=True)
            e = rel_err(lg, ref_l)
            same = torch.equal(ids, ref_t)
            print(f"[fp32] {name:6s} shards={S}: logits rel err {e:.2e}  tokens identical={same} "
                  f"min top2 margin {float(mg.min()):.2e}")
            ok &= same and e < 1e-4
        sd16 = {k: v.to(torch.bfloat16) for k, v in sd32.items()}
        m16 = hf_model(cfg, sd16, torch.bfloat16)
        ref_t16, ref_l16 = hf_greedy(m16, prompts, n_new)
        stages = build_pipeline(cfg, sd16, 2, torch.bfloat16)
        ids, lg, mg = greedy_generate(stages, prompts, n_new, record=True)
        e = rel_err(lg, ref_l16)
        nsame = int((ids == ref_t16).sum())
        print(f"[bf16] {name:6s} shards=2: logits rel err {e:.2e}  tokens equal {nsame}/{ids.numel()} "
              f"min top2 margin {float(mg.min()):.2e}")
        ok &= e <= 1e-2
    print("ORACLE PINNED AGAINST HF" if ok else "ORACLE MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
