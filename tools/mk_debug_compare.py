"""Debug aid: one decode step of a small stage through the persistent kernel and through the per-op path
(DLLM_MEGAKERNEL=0), printing where the hand-off tiles differ.  Usage: python tools/mk_debug_compare.py [arch]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llms_b200")]


def child(out_path, arch):
    import torch
    from oracle.shard_oracle import ShardModelConfig, make_prompts, make_state_dict, split_state_dict
    from src.worker.node import ModelShard
    if arch == "gpt2":
        cfg = ShardModelConfig("gpt2", 128, 2, 4, 4, 32, 512, 211, 96, 1e-5, 10000.0)
    else:
        cfg = ShardModelConfig("llama", 256, 2, 4, 2, 64, 704, 257, 128, 1e-5, 10000.0)
    sd = make_state_dict(cfg, 1234, torch.float32, init="full")
    parts = split_state_dict(cfg, {k: v.to(torch.bfloat16) for k, v in sd.items()}, 2)
    st = ModelShard(0, dict(parts[0]), config=cfg.to_hf_dict(), max_rows=64, max_seqs=4, max_context=64, page_size=8)
    st.to_device()
    prompts = make_prompts(cfg.vocab, 3, 9, seed=1)
    st.compute({"input_ids": prompts})
    out = st.compute({"input_ids": prompts[:, :1]})
    torch.save(out["hidden_states"].float().cpu(), out_path)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(sys.argv[2], sys.argv[3])
        sys.exit(0)
    import torch
    arch = sys.argv[1] if len(sys.argv) > 1 else "llama"
    outs = []
    for mk in ("0", "1"):
        path = f"/tmp/mk_debug_{mk}.pt"
        subprocess.run([sys.executable, __file__, "--child", path, arch], check=True, env={**os.environ, "DLLM_MEGAKERNEL": mk})
        outs.append(torch.load(path))
    a, b = outs
    a, b = a.reshape(-1, a.shape[-1]), b.reshape(-1, b.shape[-1])
    print("shape", tuple(a.shape), "max |per-op|", float(a.abs().max()), "nan in persistent:", int(torch.isnan(b).sum()))
    bad = (~torch.isclose(a, b, rtol=2e-2, atol=2e-2)) | torch.isnan(b)
    print("mismatching elements:", int(bad.sum()), "of", bad.numel())
    for r in range(a.shape[0]):
        cols = bad[r].nonzero().flatten().tolist()
        print(f"row {r}: {len(cols)} bad; first cols {cols[:24]}")
        if cols:
            c0 = cols[0]
            print("   per-op    ", [round(float(v), 3) for v in a[r, c0:c0 + 8]])
            print("   persistent", [round(float(v), 3) for v in b[r, c0:c0 + 8]])
