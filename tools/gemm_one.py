"""One GEMM shape through dllm_op_gemm, a few launches (for ncu).  Usage: python tools/gemm_one.py M N K [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llms_b200")]
import torch
from src.engine import cabi
m, n, k = (int(v) for v in sys.argv[1:4])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
x = torch.randn(m, k, device="cuda").to(torch.bfloat16)
w = (torch.randn(n, k, device="cuda") * 0.05).to(torch.bfloat16)
y = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    cabi.check(cabi.lib().dllm_op_gemm(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), m, n, k, 0, st))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    cabi.check(cabi.lib().dllm_op_gemm(x.data_ptr(), w.data_ptr(), None, y.data_ptr(), m, n, k, 0, st))
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"M={m} N={n} K={k}: {ms:.3f} ms  {2.0*m*n*k/ms/1e9:.1f} TFLOP/s")
