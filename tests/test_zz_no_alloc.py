"""GPU: the allocation-free attention entry point against the ordinary one.

`dllm_op_decode_attention_static` is for callers that cannot provide a scratch buffer because they cannot
allocate device memory where they stand -- a process whose own pool owns the whole GPU, or a stage thread that
must not stall its neighbours with a synchronising cudaMalloc.  It keeps its scratch in a fixed area inside the
library (4 MiB per GPU), shared by all calls and never cleared between them by anyone but the call itself.

This file is collected last on purpose: if it fails, a `-x` run has already executed every other test.
The checks below call it twice in a row with different inputs (the second call sees the first one's leftovers
in the shared area) and compare each result with the ordinary entry point and with a plain fp32 reference,
through the C ABI, within the tolerances the other attention tests use.  It does not depend on any other test having run.
"""
import pytest
import torch

from src.engine import cabi

pytestmark = pytest.mark.gpu


def _problem(seed, lens, H, KV, hd, page, dtype):
    """Random paged K/V for sequences of the given lengths, plus one query row per sequence."""
    g = torch.Generator().manual_seed(seed)
    pps = (max(lens) + page - 1) // page
    n_pages = pps * len(lens) + 2
    perm = torch.randperm(n_pages, generator=g)
    kpool = torch.randn(n_pages, KV, page, hd, generator=g).to(dtype)
    vpool = torch.randn(n_pages, KV, page, hd, generator=g).to(dtype)
    bt = torch.full((len(lens), pps), -1, dtype=torch.int32)
    nxt = 0
    for s, n in enumerate(lens):
        for p in range((n + page - 1) // page):
            bt[s, p] = int(perm[nxt]); nxt += 1
    q = torch.randn(len(lens), H, hd, generator=g).to(dtype)
    pos = torch.tensor([n - 1 for n in lens], dtype=torch.int32)
    seq = torch.arange(len(lens), dtype=torch.int32)
    return [t.cuda() for t in (q, kpool, vpool, bt, pos, seq)], pps


def _run_both(dtype, lens, H, KV, hd, splits, seed, page=16):
    (q, kpool, vpool, bt, pos, seq), pps = _problem(seed, lens, H, KV, hd, page, dtype)
    L = cabi.lib()
    R = len(lens)
    code = cabi.DTYPE_F32 if dtype == torch.float32 else cabi.DTYPE_BF16
    st = torch.cuda.current_stream().cuda_stream
    want = torch.empty(R, H * hd, dtype=dtype, device="cuda")
    scratch = torch.zeros(L.dllm_op_decode_attention_scratch_bytes(R, H, hd, splits), dtype=torch.uint8, device="cuda")
    cabi.check(L.dllm_op_decode_attention(code, q.data_ptr(), kpool.data_ptr(), vpool.data_ptr(), bt.data_ptr(), pps, page,
                                          pos.data_ptr(), seq.data_ptr(), R, H, KV, hd, splits, scratch.data_ptr(),
                                          want.data_ptr(), st))
    got = torch.full((R, H * hd), float("nan"), dtype=dtype, device="cuda")
    cabi.check(L.dllm_op_decode_attention_static(code, q.data_ptr(), kpool.data_ptr(), vpool.data_ptr(), bt.data_ptr(), pps,
                                                 page, pos.data_ptr(), seq.data_ptr(), R, H, KV, hd, splits,
                                                 got.data_ptr(), st))
    torch.cuda.synchronize()
    return got.float().cpu(), want.float().cpu(), _fp32_reference(q, kpool, vpool, bt, pos, H, KV, hd, page, dtype)


def _fp32_reference(q, kpool, vpool, bt, pos, H, KV, hd, page, dtype):
    """Plain fp32 attention over the gathered pages, rounded where the kernels round (scores before and after scaling):
    the reference the ordinary entry point is held to in tests/test_gpu_parity.py."""
    rnd = (lambda t: t.to(dtype).float())
    q, kpool, vpool, bt, pos = (t.cpu() for t in (q, kpool, vpool, bt, pos))
    out = torch.empty(q.shape[0], H * hd)
    for r in range(q.shape[0]):
        n = int(pos[r]) + 1
        pages = bt[r, : (n + page - 1) // page].long()
        K = kpool[pages].float().permute(0, 2, 1, 3).reshape(-1, KV, hd)[:n].repeat_interleave(H // KV, dim=1)
        V = vpool[pages].float().permute(0, 2, 1, 3).reshape(-1, KV, hd)[:n].repeat_interleave(H // KV, dim=1)
        sc = rnd(rnd(torch.einsum("hd,chd->hc", q[r].float(), K)) * hd ** -0.5)
        out[r] = torch.einsum("hc,chd->hd", torch.softmax(sc, -1), V).reshape(-1)
    return out


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_matches_ordinary_entry_point_twice_in_a_row(dtype):
    tol = 2e-5 if dtype == torch.float32 else 1e-2
    # two different problems back to back: the second runs on the first one's leftovers in the shared area
    for seed, lens, H, KV, hd, splits in [(1, [70, 5, 200], 8, 2, 64, 6), (2, [33, 129], 4, 4, 128, 3)]:
        got, want, ref = _run_both(dtype, lens, H, KV, hd, splits, seed)
        assert torch.isfinite(got).all()
        err = float((got - want).abs().max() / want.abs().max())
        assert err < tol, (seed, err)
        # and against the same fp32 reference the ordinary entry point is checked with
        err_ref = float((got - ref).abs().max() / ref.abs().max())
        assert err_ref < tol, (seed, err_ref)
    # and the very same problem again gives the very same bits
    a, _, _ = _run_both(dtype, [70, 5, 200], 8, 2, 64, 6, seed=1)
    b, _, _ = _run_both(dtype, [70, 5, 200], 8, 2, 64, 6, seed=1)
    assert torch.equal(a, b)


def test_static_scratch_rejects_shapes_beyond_its_area():
    L = cabi.lib()
    x = torch.zeros(16, device="cuda")
    i = torch.zeros(16, dtype=torch.int32, device="cuda")
    # 64 rows x 64 heads x 64 splits x 128 dims of fp32 partials = far more than 4 MiB
    rc = L.dllm_op_decode_attention_static(cabi.DTYPE_BF16, x.data_ptr(), x.data_ptr(), x.data_ptr(), i.data_ptr(), 1, 16,
                                           i.data_ptr(), i.data_ptr(), 64, 64, 8, 128, 64, x.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream)
    msg = L.dllm_last_error().decode()
    assert rc != 0 and "built-in area" in msg and str(4 * 1024 * 1024) in msg
