"""Times a long chunked prefill several times and samples SM clocks / power meanwhile (debug aid)."""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llms_b200")]
import torch
from src.model import random_init
from src.worker.node import ModelShard

model, T, rows = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
hf = random_init.NAMED_CONFIGS[model]
L = hf["num_hidden_layers"]
params = random_init.random_stage_parameters(hf, range(L), True, True, 1234, torch.bfloat16, device="cuda:0")
shard = ModelShard(0, params, config=hf, max_rows=rows, max_seqs=1, max_context=T + 16, page_size=64)
shard.to_device()
prompts = torch.randint(0, hf["vocab_size"], (1, T))
shard.compute({"input_ids": prompts}); torch.cuda.synchronize(); shard.reset()
smi = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader", "-lms", "50"],
                       stdout=subprocess.PIPE, text=True)
for i in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(); shard.compute({"input_ids": prompts}); e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"prefill {T} tokens (chunks of {rows}): device {e0.elapsed_time(e1):.1f} ms, host enqueue {1e3*(t1-t0):.1f} ms, wall {1e3*(t2-t0):.1f} ms")
    shard.reset()
smi.terminate()
lines = smi.stdout.read().strip().splitlines()
print("clock samples:", lines[:: max(1, len(lines) // 12)])
