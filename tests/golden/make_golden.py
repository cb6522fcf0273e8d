"""Generate the golden fixtures under tests/golden/ (run in the build container, where /root/reference
and transformers are present; the GPU box has neither).

    python tests/golden/make_golden.py

Writes
  reference_placeholder.pt : outputs of the UNMODIFIED reference `ModelShard.compute`
                             (/root/reference/src/worker/node.py:24-32), imported from /root/reference,
                             on seeded inputs -- pins the placeholder path against the reference itself.
  hf_<model>.pt            : HF transformers 5.5.0 (eager attention, fp32) on the oracle's seeded
                             random-init weights: prompts, greedy tokens, last-position logits per step.
  hf_<model>_bf16.pt       : the same HF classes in bf16 (weights cast to bf16, HF's own CPU bf16 kernels): greedy
                             tokens and last-position logits per step -- pins the bf16 oracle's rounding points.
  oracle_<model>_<dtype>.pt: the oracle's own outputs (stage-boundary hidden states of the prefill, logits,
                             tokens, top-2 margins) for 2 shards, fp32 and bf16 -- regression pins.
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.shard_oracle import (ShardModelConfig, build_pipeline, greedy_generate, make_prompts,  # noqa: E402
                                 make_state_dict, pipeline_step)
from oracle.validate_against_hf import hf_greedy, hf_model  # noqa: E402

# GPU-runnable tiny geometries (head_dim in {32, 64}; every dim a multiple of 8)
CASES = {
    "gpt2": ShardModelConfig("gpt2", 128, 4, 4, 4, 32, 512, 211, 96, 1e-5, tie_head=True),
    "llama": ShardModelConfig("llama", 256, 4, 4, 2, 64, 704, 257, 128, 1e-5, 10000.0),
    "qwen2": ShardModelConfig("qwen2", 256, 3, 8, 2, 32, 384, 300, 128, 1e-6, 1000000.0),
}
N_NEW, PROMPT_LEN, BATCH = 12, 9, 2


def reference_placeholder():
    sys.path.insert(0, "/root/reference")
    from src.worker.node import ModelShard as RefShard  # the reference, unmodified
    g = torch.Generator().manual_seed(11)
    params = {"weight": torch.randn(48, 40, generator=g), "other": torch.randn(16, 24, generator=g)}
    inputs = {"weight": torch.randn(3, 48, generator=g), "other": torch.randn(1, 16, generator=g),
              "absent": torch.randn(2, 2, generator=g)}
    shard = RefShard(1, {k: v.clone() for k, v in params.items()})
    shard.device = "cpu"
    out = shard.compute(inputs)
    sys.path.pop(0)
    for m in [m for m in sys.modules if m == "src" or m.startswith("src.")]:
        del sys.modules[m]
    torch.save({"params": params, "inputs": inputs, "outputs": out}, os.path.join(HERE, "reference_placeholder.pt"))
    print("reference_placeholder.pt:", {k: tuple(v.shape) for k, v in out.items()})


def main():
    reference_placeholder()
    for name, cfg in CASES.items():
        sd = make_state_dict(cfg, 1234, torch.float32, init="full")
        prompts = make_prompts(cfg.vocab, BATCH, PROMPT_LEN, seed=7)
        m = hf_model(cfg, sd, torch.float32)
        hf_t, hf_l = hf_greedy(m, prompts, N_NEW)
        torch.save({"config": cfg.to_hf_dict(), "prompts": prompts, "tokens": hf_t, "logits": hf_l},
                   os.path.join(HERE, f"hf_{name}.pt"))
        m16 = hf_model(cfg, {k: v.to(torch.bfloat16) for k, v in sd.items()}, torch.bfloat16)
        hf_t16, hf_l16 = hf_greedy(m16, prompts, N_NEW)
        torch.save({"config": cfg.to_hf_dict(), "prompts": prompts, "tokens": hf_t16, "logits": hf_l16},
                   os.path.join(HERE, f"hf_{name}_bf16.pt"))
        for dt, tag in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
            sdd = {k: v.to(dt) for k, v in sd.items()}
            stages = build_pipeline(cfg, sdd, 2, dt)
            for st in stages:
                st.reset()
            boundary = stages[0].compute({"input_ids": prompts})["hidden_states"]
            ids, lg, mg = greedy_generate(stages, prompts, N_NEW, record=True)
            torch.save({"config": cfg.to_hf_dict(), "prompts": prompts, "tokens": ids, "logits": lg, "margins": mg,
                        "boundary_hidden": boundary},
                       os.path.join(HERE, f"oracle_{name}_{tag}.pt"))
            if dt == torch.float32:
                err = float((lg - hf_l).abs().max() / hf_l.abs().max())
                assert torch.equal(ids, hf_t) and err < 1e-4, (name, err)
                print(f"{name}: oracle fp32 == HF (tokens identical, logits rel err {err:.1e}); "
                      f"min margin {float(mg.min()):.2e}")
            else:
                same = (ids == hf_t16).long().cumprod(1).bool()
                err16 = max(float((lg[:, t][same[:, max(t - 1, 0)] | (t == 0)] - hf_l16[:, t][same[:, max(t - 1, 0)] | (t == 0)]).abs().max()
                                  / hf_l16[:, t].abs().max()) for t in range(N_NEW) if bool((same[:, max(t - 1, 0)] | (t == 0)).any()))
                print(f"{name}: bf16 min margin {float(mg.min()):.2e}; tokens equal to fp32 run: "
                      f"{int((ids == hf_t).sum())}/{ids.numel()}; vs HF bf16: tokens equal {int((ids == hf_t16).sum())}/{ids.numel()}, "
                      f"logits rel err {err16:.2e} while the histories agree")


if __name__ == "__main__":
    main()
