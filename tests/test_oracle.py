"""CPU: the oracle against its pins (golden vectors from HF transformers and from the reference itself)."""
import os

import pytest
import torch

from oracle.shard_oracle import (NAMED_CONFIGS, OracleShard, ShardModelConfig, build_pipeline,
                                 contiguous_partition, greedy_generate, legacy_compute, make_state_dict,
                                 pipeline_step, split_state_dict)

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    return torch.load(os.path.join(GOLD, name), weights_only=True)


def test_placeholder_matches_reference_golden():
    """legacy_compute restates reference src/worker/node.py:28-32; golden = the reference's own output."""
    g = _load("reference_placeholder.pt")
    out = legacy_compute(g["params"], g["inputs"])
    assert set(out) == set(g["outputs"]) == {"weight", "other"}
    for k in out:
        assert torch.equal(out[k], g["outputs"][k])


@pytest.mark.parametrize("name", ["gpt2", "llama", "qwen2"])
@pytest.mark.parametrize("shards", [1, 2, 3])
def test_oracle_fp32_matches_hf_golden(name, shards):
    g = _load(f"hf_{name}.pt")
    cfg = ShardModelConfig.from_hf_dict(g["config"])
    sd = make_state_dict(cfg, 1234, torch.float32, init="full")
    stages = build_pipeline(cfg, sd, shards, torch.float32)
    ids, lg, _ = greedy_generate(stages, g["prompts"], g["tokens"].shape[1], record=True)
    assert torch.equal(ids, g["tokens"])  # identical greedy tokens
    rel = float((lg - g["logits"]).abs().max() / g["logits"].abs().max())
    assert rel < 1e-4, rel  # north-star fp32 tolerance


@pytest.mark.parametrize("name", ["gpt2", "llama", "qwen2"])
@pytest.mark.parametrize("shards", [1, 2])
def test_oracle_bf16_matches_hf_bf16_golden(name, shards):
    """The bf16 oracle (fp32 math, bf16 rounding at the HF op boundaries) against HF's own bf16 forward on the same
    bf16 weights: identical greedy tokens and logits within the north star's 1e-2 relative, step by step."""
    g = _load(f"hf_{name}_bf16.pt")
    cfg = ShardModelConfig.from_hf_dict(g["config"])
    sd = {k: v.to(torch.bfloat16) for k, v in make_state_dict(cfg, 1234, torch.float32, init="full").items()}
    stages = build_pipeline(cfg, sd, shards, torch.bfloat16)
    ids, lg, _ = greedy_generate(stages, g["prompts"], g["tokens"].shape[1], record=True)
    assert torch.equal(ids, g["tokens"])
    for t in range(lg.shape[1]):
        rel = float((lg[:, t] - g["logits"][:, t]).abs().max() / g["logits"][:, t].abs().max())
        assert rel <= 1e-2, (t, rel)


@pytest.mark.parametrize("name", ["gpt2", "llama", "qwen2"])
@pytest.mark.parametrize("tag,dt", [("fp32", torch.float32), ("bf16", torch.bfloat16)])
def test_oracle_regression_pins(name, tag, dt):
    g = _load(f"oracle_{name}_{tag}.pt")
    cfg = ShardModelConfig.from_hf_dict(g["config"])
    sd = {k: v.to(dt) for k, v in make_state_dict(cfg, 1234, torch.float32, init="full").items()}
    stages = build_pipeline(cfg, sd, 2, dt)
    ids, lg, mg = greedy_generate(stages, g["prompts"], g["tokens"].shape[1], record=True)
    assert torch.equal(ids, g["tokens"])
    assert torch.allclose(lg, g["logits"], rtol=0, atol=1e-5 * float(g["logits"].abs().max()))


def test_contiguous_partition():
    assert [list(r) for r in contiguous_partition(4, 2)] == [[0, 1], [2, 3]]
    assert [len(r) for r in contiguous_partition(32, 8)] == [4] * 8
    assert [len(r) for r in contiguous_partition(22, 4)] == [6, 5, 6, 5]
    for L, S in [(12, 2), (22, 3), (80, 8), (5, 5)]:
        parts = contiguous_partition(L, S)
        assert [i for r in parts for i in r] == list(range(L))  # every layer once, in order


def test_split_keeps_embeddings_and_head():
    cfg = ShardModelConfig("gpt2", 64, 4, 4, 4, 16, 256, 97, 32, 1e-5, tie_head=True)
    parts = split_state_dict(cfg, make_state_dict(cfg), 2)
    assert "wte.weight" in parts[0] and "wpe.weight" in parts[0]
    assert "ln_f.weight" in parts[1] and "lm_head.weight" in parts[1]
    assert any(k.startswith("h.0.") for k in parts[0]) and any(k.startswith("h.3.") for k in parts[1])
    assert not any(k.startswith("h.2.") for k in parts[0])


def test_decode_equals_prefill_positions():
    """KV-cache consistency: feeding a prompt token by token gives the logits of one prefill call."""
    cfg = ShardModelConfig("llama", 64, 2, 4, 2, 16, 128, 101, 64, 1e-5)
    sd = make_state_dict(cfg, 5, torch.float32, init="full")
    prompts = torch.randint(0, cfg.vocab, (2, 7), generator=torch.Generator().manual_seed(1))
    a = build_pipeline(cfg, sd, 2)
    full = pipeline_step(a, {"input_ids": prompts})["logits"]
    b = build_pipeline(cfg, sd, 2)
    for t in range(prompts.shape[1]):
        step = pipeline_step(b, {"input_ids": prompts[:, t:t + 1]})["logits"]
    assert torch.allclose(full, step, atol=1e-5)


def test_ragged_batch_rows_are_independent():
    """Rows of different context length in one call (explicit positions / seq ids)."""
    cfg = ShardModelConfig("llama", 64, 2, 4, 2, 16, 128, 101, 64, 1e-5)
    sd = make_state_dict(cfg, 5, torch.float32, init="full")
    g = torch.Generator().manual_seed(3)
    p0, p1 = torch.randint(0, 101, (1, 5), generator=g), torch.randint(0, 101, (1, 9), generator=g)
    st = OracleShard(0, sd, cfg)
    st.compute({"input_ids": p0, "seq_ids": torch.tensor([0])})
    st.compute({"input_ids": p1, "seq_ids": torch.tensor([1])})
    nxt = torch.tensor([[3], [8]])
    both = st.compute({"input_ids": nxt, "seq_ids": torch.tensor([0, 1])})["logits"]
    solo = OracleShard(0, sd, cfg)
    solo.compute({"input_ids": p1, "seq_ids": torch.tensor([0])})
    alone = solo.compute({"input_ids": nxt[1:], "seq_ids": torch.tensor([0])})["logits"]
    assert torch.allclose(both[1], alone[0], atol=1e-5)


def test_named_configs_match_survey_table():
    c = NAMED_CONFIGS["llama3-8b"]
    per_layer = c.hidden * c.n_heads * c.head_dim * 2 + 2 * c.hidden * c.n_kv_heads * c.head_dim + 3 * c.hidden * c.ffn
    assert per_layer == 218_103_808  # 218.1 M params per layer
    t = NAMED_CONFIGS["tinyllama-1.1b"]
    assert t.hidden * t.n_heads * t.head_dim * 2 + 2 * t.hidden * t.n_kv_heads * t.head_dim + 3 * t.hidden * t.ffn == 44_040_192
