"""Decode tokens/s through the reference-facing API (MasterNode.run_inference with one WorkerNode per GPU):
the master's per-token RUN_INFERENCE relay against its fused decode loop.  Usage: python tools/master_bench.py [model] [new_tokens]"""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llms_b200")]
import torch  # noqa: E402

from src.master.node import MasterNode  # noqa: E402
from src.worker.node import WorkerNode  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "tinyllama-1.1b"
n_new = int(sys.argv[2]) if len(sys.argv) > 2 else 65
n_gpus = torch.cuda.device_count()
master = MasterNode(host="127.0.0.1", port=0)
master.start()
workers = []
for dev in range(n_gpus):
    w = WorkerNode(host="127.0.0.1", port=0, master_address=f"127.0.0.1:{master.port}", device=dev)
    w.start()
    workers.append(w)
    while len(master.workers) < dev + 1:
        time.sleep(0.02)
with tempfile.TemporaryDirectory() as tmp:
    t0 = time.time()
    master.initialize_model(f"random:{model}", num_shards=n_gpus, cache_dir=tmp)
    master.assign_shards()
    master.distribute_shards(timeout=600)
    print(f"{model}: {n_gpus} stage(s) loaded in {time.time() - t0:.1f} s")
    prompt = torch.randint(0, 1000, (1, 512), generator=torch.Generator().manual_seed(1))
    for fused in (True, False, True):
        master.use_fused_handoff = fused
        master.run_inference(prompt, timeout=600, max_new_tokens=1)  # prefill only: subtract it below
        t0 = time.time(); master.run_inference(prompt, timeout=600, max_new_tokens=1); t_pre = time.time() - t0
        t0 = time.time(); out = master.run_inference(prompt, timeout=600, max_new_tokens=n_new); t_all = time.time() - t0
        print(f"decode path {master.last_decode_path:8s}: {(n_new - 1) / (t_all - t_pre):8.1f} tok/s "
              f"({(t_all - t_pre) / (n_new - 1) * 1e3:.2f} ms/token; prefill round trip {t_pre * 1e3:.1f} ms)")
    # prefill hand-off: hidden-state tiles worker -> worker over NVLink (mailbox tile slots) against the pickled relay
    # through the master, 2048-token prompt (the Qwen2.5-14B tile of BASELINE configs[3] is 20 MiB per hop)
    long_prompt = torch.randint(0, 1000, (1, 2048), generator=torch.Generator().manual_seed(2))
    for direct in (True, False, True):
        master.use_direct_tiles = direct
        master.run_inference(long_prompt, timeout=600, max_new_tokens=1)
        t0 = time.time()
        for _ in range(3):
            master.run_inference(long_prompt, timeout=600, max_new_tokens=1)
        print(f"prefill hand-off {master.last_prefill_path:6s}: {(time.time() - t0) / 3 * 1e3:8.1f} ms per 2048-token prefill round trip")
for w in workers:
    w.stop()
master.stop()
