"""GPU: the CUDA path (through the C ABI) against the CPU oracle and the golden fixtures.

Tolerances are the north star's: fp32 logits within 1e-4 relative, bf16 within 1e-2 relative (relative to the
largest |logit| of the step), greedy token ids identical.  Integer outputs (token ids) are compared exactly.
"""
import os

import pytest
import torch

from oracle.shard_oracle import (ShardModelConfig, build_pipeline, greedy_generate, make_prompts, make_state_dict,
                                 split_state_dict)
from src.engine import cabi
from src.model import random_init
from src.worker.node import ModelShard, WorkerNode

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = {torch.float32: 1e-4, torch.bfloat16: 1e-2}


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


# Every bf16 comparison of this file lands here, so that the suite can print (and bound) how often the strict 1e-2
# is exceeded and how many greedy ids differ, instead of passing silently: (relative error, error in bf16 quanta).
BF16_LOG = []
BF16_FLIPS = []   # (oracle top-2 margin / scale) of every first differing id


def bf16_logits_close(a, b, log=True):
    """North-star bf16 tolerance: 1e-2 relative to the largest |logit|.  HF (and the oracle) return logits rounded to
    bf16; on the small random-init test models max|logit| < 1, where ONE output quantum is 0.4-0.8 % of the scale, so
    two correctly rounded results whose exact values differ by just over one quantum land two quanta apart (up to 1.6 %).
    Pairs up to three quanta apart are accepted and counted: `test_bf16_parity_summary` prints the counts and fails
    when they grow.  The BASELINE-size tests (test_gpu_full_size.py) have no quantum floor: they measure the oracle's
    own sensitivity to summation order in the same run and hold the CUDA path to max(1e-2, twice that)."""
    a, b = a.float().cpu(), b.float().cpu()
    scale = float(b.abs().max())
    ulp = 2.0 ** (torch.floor(torch.log2(torch.tensor(scale))).item() - 7)
    err = float((a - b).abs().max())
    if log:   # (a test that re-evaluates one case through several kernel variants logs it once: the summary counts cases)
        BF16_LOG.append((err / scale, err / ulp))
    return err <= max(1e-2 * scale, 3.0 * ulp)


def check_bf16_generation(ids, lg, o_ids, o_lg, o_mg, log=True):
    """Free-running bf16 generation against the oracle's: logits are comparable while the token histories agree;
    a first differing id must sit on an oracle top-2 margin inside bf16 resolution of the logit scale (2^-7), and is
    counted."""
    same_prefix = (ids == o_ids).long().cumprod(1).bool()
    for t in range(ids.shape[1]):
        rows = same_prefix[:, t - 1] if t else torch.ones(ids.shape[0], dtype=torch.bool)
        if rows.any():
            assert bf16_logits_close(lg[rows, t], o_lg[rows, t], log), t
    for b in range(ids.shape[0]):
        if not same_prefix[b].all():
            t = int((~same_prefix[b]).float().argmax())
            rel_margin = float(o_mg[b, t]) / float(o_lg[b, t].abs().max())
            if log:
                BF16_FLIPS.append(rel_margin)
            assert rel_margin <= 2 ** -7, (b, t, rel_margin)


def stream():
    return torch.cuda.current_stream().cuda_stream


# ---- single operators -------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("rows,n_out,k", [(1, 64, 256), (1, 257, 2048), (3, 50, 768), (5, 33, 5632), (1, 2, 8), (16, 96, 512)])
def test_gemv_matches_fp32_reference(dtype, rows, n_out, k):
    g = torch.Generator().manual_seed(rows * 1000 + n_out)
    x = torch.randn(rows, k, generator=g).to(dtype)
    w = (torch.randn(n_out, k, generator=g) * 0.05).to(dtype)
    b = torch.randn(n_out, generator=g).to(dtype)
    y0 = torch.randn(rows, n_out, generator=g).to(dtype)
    ref = x.float() @ w.float().t() + b.float()
    ref_acc = ref.to(dtype).float() + y0.float()
    xd, wd, bd = x.cuda(), w.cuda(), b.cuda()
    code = cabi.DTYPE_F32 if dtype == torch.float32 else cabi.DTYPE_BF16
    y = torch.empty(rows, n_out, dtype=dtype, device="cuda")
    cabi.check(cabi.lib().dllm_op_gemv(code, xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), y.data_ptr(), rows, n_out, k, 0, stream()))
    ya = y0.cuda()
    cabi.check(cabi.lib().dllm_op_gemv(code, xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), ya.data_ptr(), rows, n_out, k, 1, stream()))
    torch.cuda.synchronize()
    tol = 2e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(y, ref) < tol
    assert rel_err(ya, ref_acc) < tol


@pytest.mark.parametrize("m,n_out,k", [(128, 128, 64), (512, 256, 2048), (37, 130, 704), (200, 2560, 256), (1, 64, 128),
                                       (300, 5632, 2048), (1300, 4096, 320), (2500, 2048, 192), (16, 4096, 4096), (33, 1024, 2048),
                                       (64, 640, 1088), (7, 50, 72)])
def test_prefill_gemm_matches_fp32_reference(m, n_out, k, monkeypatch):
    """The tcgen05 / TMA GEMM (bf16 in, fp32 accumulate in tensor memory) against a plain fp32 matmul, with ragged
    tiles in every dimension, bias, and the in-place residual form.  Shapes with more than 148 tiles of 128 x 256 take
    the 256-wide tiles; m <= 64 takes the swap-AB kernel that splits K over a cluster (16 x 4096 x 4096: 8 ways)."""
    monkeypatch.setenv("DLLM_GEMM_CG2", "1")   # CTA pairs (cta_group::2) wherever the shape allows, not only from 1024 rows on
    g = torch.Generator().manual_seed(m * 7 + n_out)
    x = torch.randn(m, k, generator=g).to(torch.bfloat16)
    w = (torch.randn(n_out, k, generator=g) * 0.05).to(torch.bfloat16)
    b = torch.randn(n_out, generator=g).to(torch.bfloat16)
    y0 = torch.randn(m, n_out, generator=g).to(torch.bfloat16)
    ref = (x.float() @ w.float().t() + b.float()).to(torch.bfloat16).float()
    ref_acc = (ref + y0.float()).to(torch.bfloat16).float()
    xd, wd, bd = x.cuda(), w.cuda(), b.cuda()
    y = torch.full((m, n_out), float("nan"), dtype=torch.bfloat16, device="cuda")
    cabi.check(cabi.lib().dllm_op_gemm(xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), y.data_ptr(), m, n_out, k, 0, stream()))
    ya = y0.cuda()
    cabi.check(cabi.lib().dllm_op_gemm(xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), ya.data_ptr(), m, n_out, k, 1, stream()))
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all()
    assert rel_err(y, ref) < 1e-2 and rel_err(ya, ref_acc) < 1e-2
    # the decode GEMV computes the same product: the two paths must agree to bf16 rounding
    yv = torch.empty(m, n_out, dtype=torch.bfloat16, device="cuda")
    cabi.check(cabi.lib().dllm_op_gemv(cabi.DTYPE_BF16, xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), yv.data_ptr(), m, n_out, k, 0, stream()))
    torch.cuda.synchronize()
    assert rel_err(y, yv) < 1e-2


def test_gemv_rejects_bad_k():
    x = torch.zeros(1, 12, device="cuda", dtype=torch.bfloat16)
    rc = cabi.lib().dllm_op_gemv(0, x.data_ptr(), x.data_ptr(), None, x.data_ptr(), 1, 1, 12, 0, stream())
    assert rc != 0 and b"multiple" in cabi.lib().dllm_last_error()


def test_handoff_peer_copy_moves_the_tile_bit_for_bit():
    """`dllm_handoff_peer_copy` (the hand-off tile as one stream-ordered device-to-device copy): same device here, the
    peer-to-peer form over NVLink when a second GPU is present."""
    L = cabi.lib()
    g = torch.Generator().manual_seed(1)
    tile = torch.randn(2, 37, 512, generator=g).to(torch.bfloat16).cuda()
    dst = torch.zeros_like(tile)
    cabi.check(L.dllm_handoff_peer_copy(dst.data_ptr(), 0, tile.data_ptr(), 0, tile.numel() * 2, stream()))
    torch.cuda.synchronize()
    assert torch.equal(dst, tile)
    assert L.dllm_handoff_peer_copy(None, 0, tile.data_ptr(), 0, 16, stream()) != 0      # null destination is refused
    if torch.cuda.device_count() >= 2:
        far = torch.zeros(tile.shape, dtype=torch.bfloat16, device="cuda:1")
        cabi.check(L.dllm_handoff_peer_copy(far.data_ptr(), 1, tile.data_ptr(), 0, tile.numel() * 2, stream()))
        torch.cuda.synchronize()
        assert torch.equal(far.cpu(), tile.cpu())


def _paged_kv(lens, KV, hd, page, dtype, gen):
    """Random K/V for sequences of the given lengths, scattered over shuffled pages."""
    pps = (max(lens) + page - 1) // page
    n_pages = pps * len(lens) + 3
    perm = torch.randperm(n_pages, generator=gen)
    kpool = torch.zeros(n_pages, KV, page, hd)
    vpool = torch.zeros(n_pages, KV, page, hd)
    bt = torch.full((len(lens), pps), -1, dtype=torch.int32)
    ks, vs, nxt = [], [], 0
    for s, n in enumerate(lens):
        k = torch.randn(n, KV, hd, generator=gen).to(dtype).float()
        v = torch.randn(n, KV, hd, generator=gen).to(dtype).float()
        ks.append(k); vs.append(v)
        for p in range((n + page - 1) // page):
            pg = int(perm[nxt]); nxt += 1
            bt[s, p] = pg
            t0, t1 = p * page, min(n, (p + 1) * page)
            kpool[pg, :, : t1 - t0] = k[t0:t1].transpose(0, 1)
            vpool[pg, :, : t1 - t0] = v[t0:t1].transpose(0, 1)
    return kpool.to(dtype), vpool.to(dtype), bt, ks, vs, pps


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("H,KV,hd,splits,lens", [(4, 2, 64, 3, [5, 70]), (8, 1, 128, 8, [200, 1, 33]), (4, 4, 32, 1, [17]),
                                                 (32, 4, 64, 16, [512]), (5, 1, 64, 64, [40, 41]),
                                                 (10, 2, 128, 2, [65, 129, 64]), (16, 1, 64, 4, [300, 63]),
                                                 (7, 1, 32, 1, [1, 64, 127]), (40, 8, 128, 3, [257])])
def test_decode_attention_matches_reference(dtype, H, KV, hd, splits, lens):
    gen = torch.Generator().manual_seed(H * 100 + hd)
    page = 16
    kpool, vpool, bt, ks, vs, pps = _paged_kv(lens, KV, hd, page, dtype, gen)
    R = len(lens)
    q = torch.randn(R, H, hd, generator=gen).to(dtype)
    pos = torch.tensor([n - 1 for n in lens], dtype=torch.int32)
    seq = torch.arange(R, dtype=torch.int32)
    ref = torch.empty(R, H * hd)
    G = H // KV
    rnd = (lambda t: t.to(dtype).float())
    for r in range(R):
        K = ks[r].repeat_interleave(G, dim=1)
        V = vs[r].repeat_interleave(G, dim=1)
        sc = rnd(rnd(torch.einsum("hd,chd->hc", q[r].float(), K)) * hd ** -0.5)
        ref[r] = torch.einsum("hc,chd->hd", torch.softmax(sc, -1), V).reshape(-1)
    code = cabi.DTYPE_F32 if dtype == torch.float32 else cabi.DTYPE_BF16
    L = cabi.lib()
    scratch = torch.zeros(L.dllm_op_decode_attention_scratch_bytes(R, H, hd, splits), dtype=torch.uint8, device="cuda")
    out = torch.empty(R, H * hd, dtype=dtype, device="cuda")
    args = [t.cuda() for t in (q, kpool, vpool, bt, pos, seq)]
    for _ in range(2):  # second call checks that the counters were left at zero
        out.zero_()
        cabi.check(L.dllm_op_decode_attention(code, args[0].data_ptr(), args[1].data_ptr(), args[2].data_ptr(), args[3].data_ptr(),
                                              pps, page, args[4].data_ptr(), args[5].data_ptr(), R, H, KV, hd, splits,
                                              scratch.data_ptr(), out.data_ptr(), stream()))
        torch.cuda.synchronize()
        assert rel_err(out, ref) < (2e-5 if dtype == torch.float32 else 1e-2)


def test_placeholder_matches_reference_golden():
    """The reference's shipped `compute` (per-key matmul), through ModelShard on the GPU."""
    g = torch.load(os.path.join(GOLD, "reference_placeholder.pt"), weights_only=True)
    shard = ModelShard(1, {k: v.clone() for k, v in g["params"].items()})
    shard.to_device()
    out = shard.compute(g["inputs"])
    assert set(out) == set(g["outputs"])
    for k in out:
        assert out[k].is_cuda and out[k].shape == g["outputs"][k].shape
        assert rel_err(out[k], g["outputs"][k]) < 1e-5


def test_reference_worker_tests_on_gpu():
    """reference tests/worker/test_node.py:26-30, 87-112: shapes and the two-shard merge."""
    shard = ModelShard(1, {"weight": torch.randn(3, 3)})
    r = shard.compute({"weight": torch.randn(1, 3)})
    assert "weight" in r and r["weight"].shape == (1, 3)
    w = WorkerNode()
    w.shards[1] = ModelShard(1, {"weight1": torch.randn(3, 3)})
    w.shards[2] = ModelShard(2, {"weight2": torch.randn(3, 3)})
    res = w.schedule_computation({"weight1": torch.randn(1, 3), "weight2": torch.randn(1, 3)}, [1, 2])
    assert res["weight1"].shape == (1, 3) and res["weight2"].shape == (1, 3)


# ---- the layer-shard pipeline ---------------------------------------------------------------------------------
def gpu_stages(cfg: ShardModelConfig, sd, n_shards, dtype, **limits):
    hf = cfg.to_hf_dict()
    stages = []
    for i, part in enumerate(split_state_dict(cfg, {k: v.to(dtype) for k, v in sd.items()}, n_shards)):
        st = ModelShard(i, dict(part), config=hf, **limits)
        st.to_device()
        stages.append(st)
    return stages


def gpu_step(stages, inputs):
    x = inputs
    for st in stages:
        x = st.compute(x)
    return x


def gpu_generate(stages, prompts, n_new):
    for st in stages:
        st.reset()
    out = gpu_step(stages, {"input_ids": prompts})
    toks, logits = [out["next_token"].cpu()], [out["logits"].cpu()]
    for _ in range(n_new - 1):
        out = gpu_step(stages, {"input_ids": toks[-1][:, None]})
        toks.append(out["next_token"].cpu())
        logits.append(out["logits"].cpu())
    return torch.stack(toks, 1), torch.stack(logits, 1)


@pytest.mark.parametrize("name", ["gpt2", "llama", "qwen2"])
@pytest.mark.parametrize("tag,dtype", [("fp32", torch.float32), ("bf16", torch.bfloat16)])
@pytest.mark.parametrize("shards", [1, 2, 3])
def test_pipeline_matches_golden(name, tag, dtype, shards):
    g = torch.load(os.path.join(GOLD, f"oracle_{name}_{tag}.pt"), weights_only=True)
    cfg = ShardModelConfig.from_hf_dict(g["config"])
    sd = make_state_dict(cfg, 1234, torch.float32, init="full")
    stages = gpu_stages(cfg, sd, shards, dtype, max_rows=64, max_seqs=4, max_context=cfg.max_pos, page_size=8)
    ids, lg = gpu_generate(stages, g["prompts"], g["tokens"].shape[1])
    if dtype == torch.float32:
        err = max(rel_err(lg[:, t], g["logits"][:, t]) for t in range(lg.shape[1]))
        assert err < TOL[dtype], err
    else:
        check_bf16_generation(ids, lg, g["tokens"], g["logits"], g["margins"])
    if dtype == torch.float32:
        assert torch.equal(ids, g["tokens"])  # bit-identical greedy tokens
        hf = torch.load(os.path.join(GOLD, f"hf_{name}.pt"), weights_only=True)
        assert torch.equal(ids, hf["tokens"]) and rel_err(lg, hf["logits"]) < 1e-4  # and against HF itself


@pytest.mark.parametrize("name", ["llama", "gpt2"])
def test_boundary_hidden_state_matches_golden(name):
    """The hand-off tile itself (stage 0 -> stage 1), fp32."""
    g = torch.load(os.path.join(GOLD, f"oracle_{name}_fp32.pt"), weights_only=True)
    cfg = ShardModelConfig.from_hf_dict(g["config"])
    sd = make_state_dict(cfg, 1234, torch.float32, init="full")
    st = gpu_stages(cfg, sd, 2, torch.float32, max_rows=64, max_seqs=4, max_context=cfg.max_pos, page_size=8)[0]
    out = st.compute({"input_ids": g["prompts"]})
    assert out["hidden_states"].shape == g["boundary_hidden"].shape
    assert rel_err(out["hidden_states"], g["boundary_hidden"]) < 1e-4
    assert out["positions"].tolist() == [list(range(g["prompts"].shape[1]))] * g["prompts"].shape[0]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_seeded_pipeline_against_live_oracle(dtype):
    """Same seeded inputs through the oracle (CPU) and the CUDA path, GQA llama with 128-wide heads, 32 tokens."""
    cfg = ShardModelConfig("llama", 512, 3, 4, 2, 128, 1024, 1000, 256, 1e-5, 500000.0)
    sd = make_state_dict(cfg, 99, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, 3, 21, seed=5)
    sdd = {k: v.to(dtype) for k, v in sd.items()}
    o_ids, o_lg, o_mg = greedy_generate(build_pipeline(cfg, sdd, 2, dtype), prompts, 32, record=True)
    stages = gpu_stages(cfg, sd, 2, dtype, max_rows=64, max_seqs=4, max_context=128, page_size=16)
    ids, lg = gpu_generate(stages, prompts, 32)
    if dtype == torch.float32:
        assert torch.equal(ids, o_ids)
        assert rel_err(lg, o_lg) < 1e-4
    else:
        check_bf16_generation(ids, lg, o_ids, o_lg, o_mg)


def test_decode_loop_equals_stepwise():
    """Device-resident greedy loop == one forward call per token (token ids exact)."""
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, 2, 10, seed=2)
    for dtype in (torch.float32, torch.bfloat16):
        (st,) = gpu_stages(cfg, sd, 1, dtype, max_rows=32, max_seqs=2, max_context=128, page_size=16)
        ids, _ = gpu_generate([st], prompts, 20)
        st.reset()
        first = st.compute({"input_ids": prompts})["next_token"].cpu()
        rest = st.decode_loop(2, 19).cpu()
        assert torch.equal(torch.cat([first[:, None], rest], 1), ids)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_decode_loop_twice_equals_one_long_loop(dtype):
    """decode_loop(a) followed by decode_loop(b) == decode_loop(a + b), bit for bit, with prompt + a landing on a KV
    page boundary (the second call must start at the pending token's position, not one past it)."""
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, 2, 10, seed=2)
    (st,) = gpu_stages(cfg, sd, 1, dtype, max_rows=32, max_seqs=2, max_context=128, page_size=16)
    st.compute({"input_ids": prompts})
    want = st.decode_loop(2, 6 + 11).cpu()
    st.reset()
    st.compute({"input_ids": prompts})
    a = st.decode_loop(2, 6).cpu()    # 10 prompt tokens + 6 = 16 = one full page
    b = st.decode_loop(2, 11).cpu()
    assert torch.equal(torch.cat([a, b], 1), want)
    c = st.decode_loop(2, 1).cpu()    # and a third call keeps going from where the second stopped
    st.reset()
    st.compute({"input_ids": prompts})
    assert torch.equal(st.decode_loop(2, 18).cpu()[:, -1:], c)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_prefill_single_step_then_decode_loop(dtype):
    """prefill (logit rows T-1, 2T-1, ...) -> one single-token forward -> decode_loop: the loop must gather its
    positions and slots from the single-token call, not from the prefill's stale logit rows."""
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, 2, 10, seed=2)
    (st,) = gpu_stages(cfg, sd, 1, dtype, max_rows=32, max_seqs=2, max_context=128, page_size=16)
    ids, _ = gpu_generate([st], prompts, 12)
    st.reset()
    first = st.compute({"input_ids": prompts})["next_token"].cpu()
    second = st.compute({"input_ids": first[:, None]})["next_token"].cpu()
    rest = st.decode_loop(2, 10).cpu()
    assert torch.equal(torch.cat([first[:, None], second[:, None], rest], 1), ids)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("arch", ["llama", "gpt2"])
def test_stage_boundary_inside_a_layer(dtype, arch):
    """A stage may end after a layer's attention block and the next one start with that layer's MLP block (the residual
    row is the hand-off either way): prefill, single steps and the oracle's tokens must not notice.  Against the oracle
    (fp32: ids identical, logits 1e-4) and, bit for bit, against the same model as one stage."""
    from src.model.random_init import block_of
    if arch == "llama":
        cfg = ShardModelConfig("llama", 256, 4, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    else:
        cfg = ShardModelConfig("gpt2", 128, 4, 4, 4, 32, 512, 211, 96, 1e-5, 10000.0, tie_head=True)
    sd = {k: v.to(dtype) for k, v in make_state_dict(cfg, 3, torch.float32, init="full").items()}
    (whole,) = split_state_dict(cfg, sd, 1)
    prefix = "h." if arch == "gpt2" else "layers."
    layer = lambda k: int(k.split(".")[1]) if k.startswith(prefix) else None
    # stage A: embedding, layers 0-1, attention block of layer 2; stage B: MLP block of layer 2, layer 3, final norm, head
    a = {k: v for k, v in whole.items() if (layer(k) is not None and (layer(k) < 2 or (layer(k) == 2 and block_of(k) == "attn")))
         or k.startswith(("wte", "wpe", "embed_tokens"))}
    b = {k: v for k, v in whole.items() if k not in a or k == "lm_head.weight"}
    if arch == "gpt2":
        b["lm_head.weight"] = whole["lm_head.weight"]
        a.pop("lm_head.weight", None)
    limits = dict(max_rows=32, max_seqs=2, max_context=64, page_size=16)
    hf = cfg.to_hf_dict()
    one = ModelShard(0, dict(whole), config=hf, **limits); one.to_device()
    sa = ModelShard(0, a, config=hf, **limits); sa.to_device()
    sb = ModelShard(1, b, config=hf, **limits); sb.to_device()
    assert sa.executor.last_layer_attn_only and sb.executor.first_layer_mlp_only
    prompts = make_prompts(cfg.vocab, 2, 9, seed=4)
    ids1, lg1 = gpu_generate([one], prompts, 8)
    ids2, lg2 = gpu_generate([sa, sb], prompts, 8)
    assert torch.equal(ids1, ids2) and torch.equal(lg1, lg2)
    if dtype == torch.float32:
        o_ids, o_lg, _ = greedy_generate(build_pipeline(cfg, sd, 1, dtype), prompts, 8, record=True)
        assert torch.equal(ids2, o_ids) and rel_err(lg2, o_lg) < 1e-4


def test_single_steps_on_pinned_host_buffers_read_in_place():
    """`dllm_shard_forward` on pinned host int32 buffers (what a master's token loop passes): the persistent kernel reads
    ids / positions / slots in place and writes the sampled id straight to host memory.  Same ids as the same steps on
    device tensors (copy path) and as the device loop; a decode loop right after such a step continues correctly (the
    handle's device copies were refreshed by the kernel)."""
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, 2, 10, seed=2)
    (st,) = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=32, max_seqs=2, max_context=128, page_size=16)
    ex = st.executor
    first = st.compute({"input_ids": prompts})["next_token"].cpu()
    want = st.decode_loop(2, 9).cpu()                      # 9 further ids from the device loop
    st.reset()
    st.compute({"input_ids": prompts})
    for s_ in range(2):
        ex.reserve(s_, 10 + 9 + 1)
    ids_h = first.to(torch.int32).pin_memory()
    pos_h = torch.full((2,), 10, dtype=torch.int32).pin_memory()
    seq_h = torch.arange(2, dtype=torch.int32).pin_memory()
    nxt_h = torch.zeros(2, dtype=torch.int32).pin_memory()
    got = []
    for _ in range(5):
        ex.step_host(ids_h, pos_h, seq_h, next_token_out=nxt_h)
        torch.cuda.current_stream().synchronize()
        got.append(nxt_h.clone())
        ids_h.copy_(nxt_h); pos_h.add_(1)
    assert torch.equal(torch.stack(got, 1).to(torch.int64), want[:, :5])
    ex._seq_len = {0: 15, 1: 15}
    rest = st.decode_loop(2, 4).cpu()                      # the loop picks up from the in-place steps
    assert torch.equal(rest, want[:, 5:])


def test_chunked_prefill_equals_single_call():
    """max_rows smaller than the prompt -> chunked prefill; same logits (fp32 tolerance) and token."""
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, 2, 37, seed=2)
    big = gpu_stages(cfg, sd, 2, torch.float32, max_rows=128, max_seqs=2, max_context=128, page_size=16)
    small = gpu_stages(cfg, sd, 2, torch.float32, max_rows=10, max_seqs=2, max_context=128, page_size=16)
    a, b = gpu_step(big, {"input_ids": prompts}), gpu_step(small, {"input_ids": prompts})
    assert torch.equal(a["next_token"], b["next_token"]) and rel_err(a["logits"], b["logits"]) < 1e-5


def test_ragged_batch_and_slot_reuse():
    """Rows with different context lengths in one call; releasing a slot and reusing its pages."""
    cfg = ShardModelConfig("qwen2", 256, 2, 8, 2, 32, 384, 300, 128, 1e-6, 1000000.0)
    sd = make_state_dict(cfg, 8, torch.float32, init="full")
    from oracle.shard_oracle import OracleShard
    oracle = OracleShard(0, sd, cfg, max_context=64)
    (st,) = gpu_stages(cfg, sd, 1, torch.float32, max_rows=32, max_seqs=3, max_context=64, page_size=8)
    g = torch.Generator().manual_seed(4)
    p0, p1 = torch.randint(0, 300, (1, 5), generator=g), torch.randint(0, 300, (1, 19), generator=g)
    for s, p in ((0, p0), (2, p1)):
        st.compute({"input_ids": p, "seq_ids": torch.tensor([s])})
        oracle.compute({"input_ids": p, "seq_ids": torch.tensor([s])})
    nxt = torch.tensor([[3], [8]])
    got = st.compute({"input_ids": nxt, "seq_ids": torch.tensor([0, 2])})
    want = oracle.compute({"input_ids": nxt, "seq_ids": torch.tensor([0, 2])})
    assert torch.equal(got["next_token"].cpu(), want["next_token"]) and rel_err(got["logits"], want["logits"]) < 1e-4
    free_before = cabi.lib().dllm_shard_kv_free_pages(st.executor._handle)
    st.reset([2]); oracle.reset([2])
    assert cabi.lib().dllm_shard_kv_free_pages(st.executor._handle) == free_before + 3  # 20 tokens / 8 per page
    got = st.compute({"input_ids": p0, "seq_ids": torch.tensor([2])})
    want = oracle.compute({"input_ids": p0, "seq_ids": torch.tensor([2])})
    assert rel_err(got["logits"], want["logits"]) < 1e-4


def test_limits_are_enforced():
    cfg = ShardModelConfig("llama", 256, 1, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32)
    (st,) = gpu_stages(cfg, sd, 1, torch.float32, max_rows=8, max_seqs=1, max_context=16, page_size=8)
    with pytest.raises(ValueError):
        st.compute({"input_ids": torch.zeros(1, 17, dtype=torch.long)})  # past max_context
    with pytest.raises(ValueError):
        st.compute({"input_ids": torch.zeros(1, 1, dtype=torch.long), "seq_ids": torch.tensor([1])})  # no such slot
    with pytest.raises(KeyError):
        gpu_stages(cfg, sd, 2, torch.float32, max_rows=8, max_seqs=1, max_context=16, page_size=8)[1].compute(
            {"input_ids": torch.zeros(1, 1, dtype=torch.long)})  # stage 1 needs hidden_states


# ---- full-size properties (BASELINE configs[1]: TinyLlama-1.1B bf16, 512-token prompt) ---------------------------
def test_tinyllama_full_size_properties():
    """Size-independent properties at the bench configuration (the oracle comparison at this size is
    tests/test_gpu_full_size.py):
    (1) splitting the model into 2 stages does not change a single bit of the hand-off or the tokens,
    (2) the device-resident loop reproduces step-by-step decoding exactly,
    (3) a 512-token chunked prefill and token-by-token prefill agree within the bf16 tolerance."""
    hf = random_init.NAMED_CONFIGS["tinyllama-1.1b"]
    L = hf["num_hidden_layers"]
    params = random_init.random_stage_parameters(hf, range(L), True, True, 1234, torch.bfloat16, device="cuda")
    limits = dict(max_rows=128, max_seqs=1, max_context=640, page_size=64)
    one = ModelShard(0, dict(params), config=hf, **limits); one.to_device()
    keys0 = {k for k in params if k.startswith("embed") or (k.startswith("layers.") and int(k.split(".")[1]) < L // 2)}
    s0 = ModelShard(0, {k: params[k] for k in keys0}, config=hf, **limits); s0.to_device()
    s1 = ModelShard(1, {k: v for k, v in params.items() if k not in keys0}, config=hf, **limits); s1.to_device()
    prompt = make_prompts(hf["vocab_size"], 1, 512, seed=7)
    a = one.compute({"input_ids": prompt})
    b = s1.compute(s0.compute({"input_ids": prompt}))
    assert torch.equal(a["logits"], b["logits"]) and torch.equal(a["next_token"], b["next_token"])
    loop = one.decode_loop(1, 31).cpu()
    toks = [b["next_token"].cpu()]
    for _ in range(31):
        toks.append(s1.compute(s0.compute({"input_ids": toks[-1][:, None]}))["next_token"].cpu())
    assert torch.equal(torch.stack(toks, 1)[:, 1:], loop)
    one.reset()
    for t in range(64):
        step = one.compute({"input_ids": prompt[:, t:t + 1]})
    one2 = ModelShard(0, dict(params), config=hf, **limits); one2.to_device()
    chunk = one2.compute({"input_ids": prompt[:, :64]})
    # two bf16 evaluation orders (different split-K and row tiling), each within 1e-2 of exact arithmetic
    assert rel_err(step["logits"], chunk["logits"]) < 3e-2


@pytest.mark.parametrize("batch", [1, 3, 8])
def test_streamed_rows_variant_matches_oracle(batch, monkeypatch):
    """The persistent kernel's second variant (down-projection rows streamed from global memory with the weights,
    what Llama-3-8B at batch 2-8 uses) forced on a small model: bf16 logits against the live oracle."""
    monkeypatch.setenv("DLLM_MK_XG", "1")
    cfg = ShardModelConfig("llama", 512, 3, 4, 2, 128, 1024, 1000, 256, 1e-5, 500000.0)
    sd = make_state_dict(cfg, 99, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, batch, 21, seed=5)
    sdd = {k: v.to(torch.bfloat16) for k, v in sd.items()}
    o_ids, o_lg, o_mg = greedy_generate(build_pipeline(cfg, sdd, 1, torch.bfloat16), prompts, 12, record=True)
    stages = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=64, max_seqs=8, max_context=128, page_size=16)
    ids, lg = gpu_generate(stages, prompts, 12)
    check_bf16_generation(ids, lg, o_ids, o_lg, o_mg)
    # and the device-resident loop through the same variant reproduces the stepwise tokens
    stages[0].reset()
    first = stages[0].compute({"input_ids": prompts})["next_token"].cpu()
    rest = stages[0].decode_loop(batch, 11).cpu()
    assert torch.equal(torch.cat([first[:, None], rest], 1), ids)


@pytest.mark.parametrize("arch", ["llama", "gpt2"])
@pytest.mark.parametrize("n_micro,rows,persistent", [(1, 1, True), (3, 2, True), (2, 8, True),
                                                     (2, 2, False), (3, 12, False), (2, 24, False)])
def test_fused_handoff_single_stage_equals_decode_loop(n_micro, rows, persistent, arch, monkeypatch):
    """One stage, its own mailbox as in- and outbox: the token ids go round through the mailbox tickets -- inside the
    persistent kernel, or (shapes it does not take: here it is switched off) through the wait / send kernels around every
    step's CUDA graph.  Every micro-batch must reproduce, bit for bit, what the classic device loop generates."""
    from src.network.handoff import Mailbox
    if persistent and os.environ.get("DLLM_MEGAKERNEL") == "0":
        pytest.skip("the persistent kernel is switched off")
    if not persistent:
        monkeypatch.setenv("DLLM_MEGAKERNEL", "0")
    if arch == "llama":
        cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    else:
        cfg = ShardModelConfig("gpt2", 128, 3, 4, 4, 32, 512, 211, 96, 1e-5, 10000.0, tie_head=True)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    n_steps, T = 7, 10
    (st,) = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=64, max_seqs=n_micro * rows, max_context=64, page_size=16)
    prompts = [make_prompts(cfg.vocab, rows, T, seed=20 + m) for m in range(n_micro)]
    want, seeds = [], []
    for m in range(n_micro):
        seqs = torch.arange(m * rows, (m + 1) * rows, dtype=torch.int32)
        first = st.compute({"input_ids": prompts[m], "seq_ids": seqs})["next_token"].cpu()
        want.append(torch.cat([first[:, None], st.decode_loop(rows, n_steps).cpu()], 1))  # [rows, 1 + n_steps]
    st.reset()
    for m in range(n_micro):
        seqs = torch.arange(m * rows, (m + 1) * rows, dtype=torch.int32)
        seeds.append(st.compute({"input_ids": prompts[m], "seq_ids": seqs})["next_token"].to(torch.int32))
    box = Mailbox.create(0, n_micro, rows, cfg.hidden)
    try:
        for m in range(n_micro):
            box.post_tokens(m, seeds[m], ticket=1)
        pos = torch.full((n_micro, rows), T, dtype=torch.int32)
        seq = torch.arange(n_micro * rows, dtype=torch.int32).reshape(n_micro, rows)
        # two launches: tickets and positions carry over
        fed_a = st.executor.decode_pipelined(box, box, n_micro, rows, 4, 0, pos, seq)
        fed_b = st.executor.decode_pipelined(box, box, n_micro, rows, n_steps - 4, 4, pos + 4, seq)
        torch.cuda.synchronize()
        fed = torch.cat([fed_a, fed_b]).cpu()  # [n_steps, n_micro, rows]: seed, then the ids generated one step earlier
        for m in range(n_micro):
            last, ticket = box.read_tokens(m)
            assert ticket == n_steps + 1
            got = torch.cat([fed[:, m, :].t(), last[:, None]], 1).to(torch.int64)
            assert torch.equal(got, want[m]), m
    finally:
        box.close()


@pytest.mark.parametrize("batch", [1, 3])
def test_time_stamped_kernel_variant_generates_the_same_ids(batch, tmp_path, monkeypatch):
    """DLLM_MK_TRACE selects the persistent kernel's variants with per-phase time stamps compiled in (their own template
    instantiations, generic in everything else): same ids as the specialised kernels, and the trace file is written."""
    if os.environ.get("DLLM_MEGAKERNEL") == "0":
        pytest.skip("the persistent kernel is switched off")
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 5, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, batch, 9, seed=batch)
    ids = {}
    for traced in (False, True):
        if traced:
            monkeypatch.setenv("DLLM_MK_TRACE", str(tmp_path / "trace.csv"))
        (st,) = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=64, max_seqs=batch, max_context=64, page_size=16)
        first = st.compute({"input_ids": prompts})["next_token"].cpu()
        ids[traced] = torch.cat([first[:, None], st.decode_loop(batch, 6).cpu()], 1)
    assert torch.equal(ids[True], ids[False])
    lines = open(tmp_path / "trace.csv").read().splitlines()
    assert lines[0].startswith("step,phase,kind") and len(lines) > 6


@pytest.mark.parametrize("persistent", [True, False])
def test_pipelined_loop_gives_up_on_a_missing_ticket_and_the_stage_stays_usable(persistent, monkeypatch):
    """A stage whose input never arrives (nobody posted the seed ids) must not hang or trap: after the time-out its loop is
    abandoned, `pipeline_status` reports it once, and the very same handle then runs the loop correctly -- in the
    persistent kernel (every CTA leaves, the barrier words end up clean) and in the per-op loop (wait / send kernels)."""
    from src.network.handoff import Mailbox
    if persistent and os.environ.get("DLLM_MEGAKERNEL") == "0":
        pytest.skip("the persistent kernel is switched off")
    if not persistent:
        monkeypatch.setenv("DLLM_MEGAKERNEL", "0")
    monkeypatch.setenv("DLLM_TICKET_TIMEOUT_MS", "300")
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 3, torch.float32, init="full")
    rows, n_steps, T = 2, 5, 10
    (st,) = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=64, max_seqs=rows, max_context=64, page_size=16)
    prompts = make_prompts(cfg.vocab, rows, T, seed=20)
    first = st.compute({"input_ids": prompts})["next_token"].cpu()
    want = torch.cat([first[:, None], st.decode_loop(rows, n_steps).cpu()], 1)
    st.reset()
    seed = st.compute({"input_ids": prompts})["next_token"].to(torch.int32)
    box = Mailbox.create(0, 1, rows, cfg.hidden)
    try:
        pos = torch.full((1, rows), T, dtype=torch.int32)
        seq = torch.arange(rows, dtype=torch.int32).reshape(1, rows)
        st.executor.decode_pipelined(box, box, 1, rows, n_steps, 0, pos, seq)   # no ids were posted: ticket 1 never comes
        with pytest.raises(cabi.DllmError, match="abandoned"):
            st.executor.pipeline_status()
        st.executor.pipeline_status()          # reported once, cleared
        # the handle, its barrier words and the CUDA context are intact: the same loop with its input
        st.reset()
        assert torch.equal(st.compute({"input_ids": prompts})["next_token"].to(torch.int32), seed)
        box.post_tokens(0, seed, ticket=1)
        fed = st.executor.decode_pipelined(box, box, 1, rows, n_steps, 0, pos, seq)
        st.executor.pipeline_status()
        last, ticket = box.read_tokens(0)
        assert ticket == n_steps + 1
        got = torch.cat([fed.cpu()[:, 0, :].t(), last[:, None]], 1).to(torch.int64)
        assert torch.equal(got, want)
    finally:
        box.close()


@pytest.mark.parametrize("qt128", ["0", "1"])
@pytest.mark.parametrize("batch,prompt_len", [(1, 1300), (2, 2100)])
def test_long_context_decode_matches_oracle(batch, prompt_len, qt128, monkeypatch):
    """Contexts beyond 64 tokens x 16 splits: every attention unit of the persistent kernel walks several K/V tiles
    (online softmax across tiles, merge across splits); chunked flash-style prefill before it.  bf16 vs the oracle."""
    monkeypatch.setenv("DLLM_PREFILL_QT128", qt128)  # both query-tile sizes of the prefill attention kernel
    cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 512, 500, 2304, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 11, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, batch, prompt_len, seed=9)
    sdd = {k: v.to(torch.bfloat16) for k, v in sd.items()}
    n_new = 6
    o_ids, o_lg, o_mg = greedy_generate(build_pipeline(cfg, sdd, 1, torch.bfloat16, max_context=prompt_len + 16), prompts, n_new, record=True)
    stages = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=1024, max_seqs=batch, max_context=prompt_len + 16, page_size=64)
    ids, lg = gpu_generate(stages, prompts, n_new)
    check_bf16_generation(ids, lg, o_ids, o_lg, o_mg)


@pytest.mark.parametrize("hd,H,KV", [(64, 8, 2), (128, 4, 1)])
@pytest.mark.parametrize("batch,T,max_rows", [(1, 130, 256), (2, 300, 1024), (1, 700, 256)])
def test_tcgen05_prefill_attention_against_oracle_and_mma_sync_kernel(hd, H, KV, batch, T, max_rows, monkeypatch):
    """The TMA + tcgen05 prefill attention (64-token pages, head_dim 64 / 128) on ragged tile edges -- prompts that end
    inside a 128-row query tile and inside a KV page, several sequences, a prompt fed in chunks (queries that start
    deep inside the context) -- against the bf16 oracle, and against the mma.sync kernel it replaces."""
    cfg = ShardModelConfig("llama", H * hd, 2, H, KV, hd, 512, 500, 1024, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 21, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, batch, T, seed=T)
    sdd = {k: v.to(torch.bfloat16) for k, v in sd.items()}
    o_ids, o_lg, o_mg = greedy_generate(build_pipeline(cfg, sdd, 1, torch.bfloat16, max_context=T + 8), prompts, 2, record=True)
    outs = {}
    for tc in ("1", "0"):
        monkeypatch.setenv("DLLM_PREFILL_ATTN_TC", tc)
        stages = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=max_rows, max_seqs=batch, max_context=T + 8, page_size=64)
        ids, lg = gpu_generate(stages, prompts, 2)
        check_bf16_generation(ids, lg, o_ids, o_lg, o_mg)
        outs[tc] = lg
    assert bf16_logits_close(outs["1"][:, 0], outs["0"][:, 0])


@pytest.mark.parametrize("arch,hd,H,KV", [("llama", 64, 8, 2), ("llama", 128, 4, 1), ("qwen2", 64, 6, 2), ("gpt2", 64, 4, 4)])
@pytest.mark.parametrize("batch,T", [(1, 130), (3, 200)])
def test_prefill_qkv_epilogue_with_rotary_and_kv_append_is_bit_identical_to_the_separate_pass(arch, hd, H, KV, batch, T, monkeypatch):
    """Prefill K1 with rotary and the paged-KV append inside the projection's epilogue (GEMM_EPI_QKV) against the
    projection followed by the separate rotary + append kernel: same roundings in the same order, so the logits of the
    prefill AND of the decode steps that read the appended K/V back are the same bits; both against the bf16 oracle.
    Covers rotary (llama), q/k/v biases (qwen2), no rotary + biases (gpt2), head_dim 64 / 128, tiles that end past M."""
    cfg = ShardModelConfig(arch, H * hd, 2, H, KV, hd, 512, 500, 512, 1e-5, 10000.0, tie_head=arch == "gpt2")
    sd = make_state_dict(cfg, 31, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, batch, T, seed=T + hd)
    sdd = {k: v.to(torch.bfloat16) for k, v in sd.items()}
    o_ids, o_lg, o_mg = greedy_generate(build_pipeline(cfg, sdd, 1, torch.bfloat16, max_context=T + 8), prompts, 3, record=True)
    outs = {}
    for fused in ("1", "0"):
        monkeypatch.setenv("DLLM_QKV_FUSED", fused)
        stages = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=1024, max_seqs=batch, max_context=T + 8, page_size=64)
        ids, lg = gpu_generate(stages, prompts, 3)
        check_bf16_generation(ids, lg, o_ids, o_lg, o_mg)
        outs[fused] = (ids, lg)
    assert torch.equal(outs["1"][0], outs["0"][0])
    assert torch.equal(outs["1"][1], outs["0"][1])


@pytest.mark.parametrize("model", ["llama3-8b", "qwen2.5-14b", "llama3-70b"])
def test_full_width_layers_persistent_kernel_vs_per_op_path(model, monkeypatch):
    """Two layers at the real widths of the BASELINE configs (hidden 4096 / 5120 / 8192, GQA 4 / 5 / 8, head_dim 128,
    full vocabulary): the persistent kernel (several rounds of panels per CTA, the streamed-rows variant at batch 4-8)
    against the independent per-op CUDA path -- a cross-check between two implementations.  The comparison with the
    oracle at these widths is tests/test_gpu_full_size.py::test_full_width_decode_against_oracle."""
    hf = dict(random_init.NAMED_CONFIGS[model])
    hf["num_hidden_layers"] = 2
    params = random_init.random_stage_parameters(hf, range(2), True, True, 77, torch.bfloat16, device="cuda")
    limits = dict(max_rows=512, max_seqs=16, max_context=128, page_size=64)
    shards = {}
    for mk in ("1", "0"):
        monkeypatch.setenv("DLLM_MEGAKERNEL", mk)
        shards[mk] = ModelShard(0, dict(params), config=hf, **limits)
        shards[mk].to_device()
    for batch in (1, 4, 8, 16):  # 16: the swap-AB GEMMs + grouped-head attention on one side, per-op kernels on the other
        prompts = make_prompts(hf["vocab_size"], batch, 24, seed=batch)
        outs = {}
        for mk, sh in shards.items():
            sh.reset()
            first = sh.compute({"input_ids": prompts})
            step = sh.compute({"input_ids": first["next_token"][:, None]})
            outs[mk] = (first["logits"].cpu(), step["logits"].cpu())
        assert torch.equal(outs["1"][0], outs["0"][0])               # same prefill kernels on both sides
        assert bf16_logits_close(outs["1"][1], outs["0"][1]), batch  # decode step: persistent kernel vs per-op kernels


def test_prefill_tile_widths_agree_bit_for_bit(monkeypatch):
    """A 4096-token prefill at Llama-3-8B width takes the 256-wide GEMM tiles (q | k | v as row segments of one
    launch) on CTA pairs (cta_group::2, 256-row tiles); with the pairs switched off one CTA per 128-row tile, with the
    wide tiles switched off 128-wide tiles.  All walk K in the same order, so the logits must be identical -- the tile
    shape is a scheduling choice, not a numerical one.  (The same prefill against the oracle:
    tests/test_gpu_full_size.py::test_long_prefill_against_oracle.)"""
    hf = dict(random_init.NAMED_CONFIGS["llama3-8b"])
    hf["num_hidden_layers"] = 2
    params = random_init.random_stage_parameters(hf, range(2), True, True, 5, torch.bfloat16, device="cuda")
    prompt = make_prompts(hf["vocab_size"], 1, 4096, seed=3)
    outs = []
    monkeypatch.setenv("DLLM_GEMM_SPLIT", "0")  # whole tiles only: a K-split tail sums in another order (its own test below)
    for bn256, qt128, cg2 in (("1", None, "1"), ("1", None, "0"), ("0", None, "1"), ("1", "0", "1")):
        monkeypatch.setenv("DLLM_GEMM_BN256", bn256)
        monkeypatch.setenv("DLLM_GEMM_CG2", cg2)   # CTA pairs on 256-row tiles / one CTA per 128-row tile: same K order
        if qt128 is not None:  # the same for the attention's query tile: 128 rows by default at this size, 64 forced
            monkeypatch.setenv("DLLM_PREFILL_QT128", qt128)
        sh = ModelShard(0, dict(params), config=hf, max_rows=4096, max_seqs=1, max_context=4160, page_size=64)
        sh.to_device()
        outs.append(sh.compute({"input_ids": prompt})["logits"].cpu())
        del sh
    assert torch.isfinite(outs[0]).all() and all(torch.equal(outs[0], o) for o in outs[1:])


@pytest.mark.parametrize("batch,T", [(1, 512), (1, 300), (2, 330)])
def test_prefill_gemm_split_tail_against_oracle_and_whole_tiles(batch, T, monkeypatch):
    """Mid-size prompts are all tail: fewer output tiles than SMs.  The persistent GEMM then cuts the tiles of the last
    wave along K, parks the fp32 slices in the handle's scratch area and lets the last slice to arrive sum them in slice
    order and run the epilogue (bias / residual / SwiGLU / rotary + KV append).  Against the bf16 oracle, against the
    same prefill with whole tiles (another summation order: close, not identical), and twice in a row (the sum does not
    depend on which slice arrives last: identical bits, tickets back at zero)."""
    cfg = ShardModelConfig("llama", 1024, 2, 16, 4, 64, 2816, 500, 1024, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 41, torch.float32, init="full")
    prompts = make_prompts(cfg.vocab, batch, T, seed=T)
    sdd = {k: v.to(torch.bfloat16) for k, v in sd.items()}
    o_ids, o_lg, o_mg = greedy_generate(build_pipeline(cfg, sdd, 1, torch.bfloat16, max_context=T + 8), prompts, 2, record=True)
    for pairs in ("1", "0"):   # split tiles of CTA pairs (each CTA's half is a tile of its own in the scratch area) / of single CTAs
        monkeypatch.setenv("DLLM_GEMM_CG2", pairs)
        outs = {}
        for split in ("1", "0", "1"):
            monkeypatch.setenv("DLLM_GEMM_SPLIT", split)
            stages = gpu_stages(cfg, sd, 1, torch.bfloat16, max_rows=1024, max_seqs=batch, max_context=T + 8, page_size=64)
            ids, lg = gpu_generate(stages, prompts, 2)
            check_bf16_generation(ids, lg, o_ids, o_lg, o_mg, log=pairs == "0" and split not in outs)
            if split == "1" and split in outs:
                assert torch.equal(outs[split], lg)
            outs[split] = lg
        assert bf16_logits_close(outs["1"][:, 0], outs["0"][:, 0], log=False)


def test_bf16_parity_summary():
    """Runs after every bf16 comparison of this file: print how strict the parity actually was, and fail when it
    degrades.  (Collected last in this file; run alone it has nothing to summarise.)"""
    if not BF16_LOG:
        pytest.skip("no bf16 comparison ran in this process")
    n = len(BF16_LOG)
    over = [e for e, _ in BF16_LOG if e > 1e-2]
    worst_rel = max(e for e, _ in BF16_LOG)
    worst_q = max(q for _, q in BF16_LOG)
    mean_rel = sum(e for e, _ in BF16_LOG) / n
    print(f"\n[parity] bf16, small models: {n} logit comparisons, mean rel err {mean_rel:.2e}, worst {worst_rel:.2e} "
          f"({worst_q:.1f} output quanta); {len(over)} above the strict 1e-2 (all within three quanta of max|logit| < 1); "
          f"{len(BF16_FLIPS)} first differing ids, oracle margin / scale there: {[round(m, 5) for m in BF16_FLIPS]}")
    assert worst_q <= 3.0
    assert len(over) <= max(2, n // 20), (len(over), n)   # a handful of two-quanta pairs, not a trend
