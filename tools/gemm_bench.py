"""Throughput of the tcgen05 prefill GEMM (dllm_op_gemm) on prefill shapes.  Usage: python tools/gemm_bench.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llms_b200")]
import torch  # noqa: E402

from src.engine import cabi  # noqa: E402

SHAPES = [  # (name, M, N, K)
    ("tinyllama qkv  M=512 ", 512, 2560, 2048), ("tinyllama o    M=512 ", 512, 2048, 2048),
    ("tinyllama down M=512 ", 512, 2048, 5632),
    ("llama8b qkv    M=512 ", 512, 6144, 4096), ("llama8b o      M=512 ", 512, 4096, 4096),
    ("llama8b down   M=512 ", 512, 4096, 14336), ("llama8b qkv    M=4096", 4096, 6144, 4096),
    ("llama8b o      M=4096", 4096, 4096, 4096), ("llama8b down   M=4096", 4096, 4096, 14336),
    ("qwen14b qkv    M=8192", 8192, 7168, 5120), ("qwen14b o      M=8192", 8192, 5120, 5120),
    ("qwen14b down   M=8192", 8192, 5120, 13824), ("llama70b o     M=4096", 4096, 8192, 8192),
]
st = torch.cuda.current_stream().cuda_stream
for name, m, n, k in SHAPES:
    x = torch.randn(m, k, device="cuda").to(torch.bfloat16)
    w = (torch.randn(n, k, device="cuda") * 0.02).to(torch.bfloat16)
    y = torch.empty(m, n, dtype=torch.bfloat16, device="cuda")
    for _ in range(3):
        cabi.check(cabi.lib().dllm_op_gemm(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), m, n, k, 0, st))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        cabi.check(cabi.lib().dllm_op_gemm(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), m, n, k, 0, st))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    ref = torch.matmul(x, w.t())
    for _ in range(3):
        torch.matmul(x, w.t(), out=ref)
    e0.record()
    for _ in range(reps):
        torch.matmul(x, w.t(), out=ref)
    e1.record()
    torch.cuda.synchronize()
    ms_ref = e0.elapsed_time(e1) / reps
    tf = 2.0 * m * n * k / (ms * 1e-3) / 1e12
    print(f"{name}  {ms:8.3f} ms  {tf:7.1f} TFLOP/s   (cuBLAS {2.0 * m * n * k / (ms_ref * 1e-3) / 1e12:7.1f})")
