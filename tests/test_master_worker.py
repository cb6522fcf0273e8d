"""Master <-> workers over localhost ZeroMQ: registration, placement, shard shipping, heartbeat eviction (CPU), and --
on a GPU -- BASELINE config[0] in miniature: a model split into 2 layer shards on a master + 2 localhost workers,
greedy decode, tokens identical to the CPU oracle."""
import json
import os
import time

import pytest
import torch

from src.master.node import GreedySampler, MasterNode
from src.model import random_init
from src.worker.node import WorkerNode

zmq = pytest.importorskip("zmq")

CFG = dict(model_type="llama", hidden_size=128, num_hidden_layers=4, num_attention_heads=4, num_key_value_heads=2,
           intermediate_size=256, vocab_size=300, max_position_embeddings=64, rms_norm_eps=1e-5, rope_theta=10000.0)


def _model_dir(tmp_path, dtype):
    d = str(tmp_path / "model")
    os.makedirs(d, exist_ok=True)
    sd = random_init.random_stage_parameters(CFG, range(4), True, True, 1234, dtype, init="full")
    torch.save({("model." + k if k != "lm_head.weight" else k): v for k, v in sd.items()}, os.path.join(d, "pytorch_model.bin"))
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(CFG, f)
    return d, sd


def _cluster(n_workers):
    master = MasterNode(host="127.0.0.1", port=0)
    master.start()
    workers = []
    for i in range(n_workers):
        w = WorkerNode(host="127.0.0.1", port=0, master_address=f"127.0.0.1:{master.port}")
        w.start()
        workers.append(w)
        deadline = time.time() + 10
        while len(master.workers) < i + 1 and time.time() < deadline:  # register in a known order
            time.sleep(0.02)
    assert len(master.workers) == n_workers
    return master, workers


def _shutdown(master, workers):
    for w in workers:
        w.stop()
    master.stop()


def test_register_assign_distribute_and_errors(tmp_path):
    d, _ = _model_dir(tmp_path, torch.float32)
    master, workers = _cluster(2)
    try:
        assert all("has_gpu" in info for info in master.workers.values())  # REGISTER carried the capabilities
        shard_dir = master.initialize_model(d, num_shards=2)
        assert sorted(f for f in os.listdir(shard_dir) if f.endswith(".pt")) == ["shard_0.pt", "shard_1.pt"]
        assign = master.assign_shards()
        assert sorted(assign.values()) == [[0], [1]]
        master.distribute_shards(timeout=60)          # LOAD_SHARD -> SHARD_LOADED round trips over ZeroMQ
        assert all(len(w.shards) == 1 for w in workers)
        assert workers[0].shards[0].config == CFG and "embed_tokens.weight" in workers[0].shards[0].parameters
        assert "lm_head.weight" in workers[1].shards[1].parameters
        if not torch.cuda.is_available():
            # without CUDA the stage cannot run: the worker reports it and the master raises (no silent CPU path)
            with pytest.raises(RuntimeError, match="no CPU"):
                master.run_inference([1, 2, 3], timeout=20, max_new_tokens=2)
    finally:
        _shutdown(master, workers)


def test_workers_load_their_stage_from_the_model_directory(tmp_path):
    """`distribute_shards(by_path=True)`: the master sends a path, every worker reads only its own stage -- from
    `model.safetensors` when there is one, else from the shard file the master wrote."""
    from safetensors.torch import save_file
    d, sd = _model_dir(tmp_path, torch.float32)
    master, workers = _cluster(2)
    try:
        master.initialize_model(d, num_shards=2)      # pytorch_model.bin only: workers fall back to shards/shard_i.pt
        master.assign_shards()
        master.distribute_shards(timeout=60, by_path=True)
        assert all(len(w.shards) == 1 for w in workers)
        assert "embed_tokens.weight" in workers[0].shards[0].parameters and "lm_head.weight" in workers[1].shards[1].parameters
        for w in workers:
            for sid in list(w.shards):
                w.unload_shard(sid)
        master._loaded_shards.clear()
        full = torch.load(os.path.join(d, "pytorch_model.bin"), weights_only=True)
        save_file({k: v.contiguous() for k, v in full.items()}, os.path.join(d, "model.safetensors"))
        os.remove(os.path.join(d, "pytorch_model.bin"))
        master.distribute_shards(timeout=60, by_path=True)   # now straight from model.safetensors, stage by stage
        got = workers[1].shards[1].parameters
        assert torch.equal(got["lm_head.weight"].float().cpu(), sd["lm_head.weight"].float())
        assert not any(k.startswith("layers.0.") for k in got) and any(k.startswith("layers.3.") for k in got)
    finally:
        _shutdown(master, workers)


def test_no_workers_and_heartbeat_eviction():
    master = MasterNode(host="127.0.0.1", port=0)
    master.start()
    try:
        master.shard_manager = type("M", (), {"num_shards": 2})()
        with pytest.raises(ValueError):
            master.assign_shards()
        master.workers["w"] = {"identity": b"w", "last_heartbeat": time.time() - 100, "status": "connected"}
        master.shard_assignments = {"w": [0]}
        assert master.evict_stale_workers() == ["w"] and not master.workers and not master.shard_assignments
    finally:
        master.stop()


def test_fused_decode_is_only_tried_when_every_stage_has_its_own_gpu():
    """Eligibility of the master's fused loop is decided from what the workers registered -- before any command goes out."""
    m = MasterNode(host="127.0.0.1", port=0)
    tok = torch.zeros(1, 2, dtype=torch.int64)

    def cluster(places, shards=None):
        m.workers = {f"w{i}": {"identity": b"x", "pid": p, "device_index": d} for i, (p, d) in enumerate(places)}
        m.shard_assignments = shards or {f"w{i}": [i] for i in range(len(places))}

    cluster([(10, 0), (10, 0)])                       # two stages of one process on one GPU: their kernels would wait for each other
    assert m._fused_decode("t", tok, 4) is None
    cluster([(10, 0), (None, None)])                  # a worker that did not say where it runs
    assert m._fused_decode("t", tok, 4) is None
    cluster([(10, 0), (11, 0)], {"w0": [0, 1], "w1": [2]})   # a worker with two stages
    assert m._fused_decode("t", tok, 4) is None
    cluster([(10, 0), (10, 1)])
    assert m._fused_decode("t", torch.zeros(1, 65, dtype=torch.int64), 4) is None   # more rows than a mailbox-fed step takes
    assert m._fused_decode("t", tok, 0) is None


def test_sampler():
    logits = torch.tensor([[0.1, 2.0, 2.0, -1.0]])
    assert GreedySampler()(logits).tolist() == [1]                       # first index of the maximum
    s = GreedySampler(temperature=1.0, top_k=2, seed=1)
    assert all(int(s(logits)) in (1, 2) for _ in range(20))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_master_two_workers_greedy_decode_matches_oracle(tmp_path, dtype):
    from oracle.shard_oracle import ShardModelConfig, build_pipeline, greedy_generate
    d, sd = _model_dir(tmp_path, dtype)
    master, workers = _cluster(2)
    try:
        master.initialize_model(d, num_shards=2)
        master.assign_shards()
        master.distribute_shards(timeout=120)
        prompt = torch.randint(0, CFG["vocab_size"], (1, 11), generator=torch.Generator().manual_seed(5))
        got = master.run_inference(prompt, timeout=120, max_new_tokens=16)
        cfg = ShardModelConfig.from_hf_dict(CFG)
        want, _, margins = greedy_generate(build_pipeline(cfg, sd, 2, dtype), prompt, 16, record=True)
        got_ids = torch.tensor(got["token_ids"])
        if dtype == torch.float32:
            assert torch.equal(got_ids, want)
        else:
            same = (got_ids == want).long().cumprod(1).bool()
            assert same.all() or float(margins[0, int((~same[0]).float().argmax())]) < 2 ** -6
    finally:
        _shutdown(master, workers)


@pytest.mark.gpu
def test_schedule_computation_command_over_the_socket(tmp_path):
    """The SCHEDULE_COMPUTATION message (reference src/worker/node.py:240-261: RUN_INFERENCE with a `priority` header that
    the reference's constant sort key turns into a no-op) sent to a worker over ZeroMQ: same RESULT as RUN_INFERENCE and as
    the in-process `schedule_computation`, fp32, checked against the oracle."""
    import pickle
    from oracle.shard_oracle import ShardModelConfig, build_pipeline, pipeline_step
    d, sd = _model_dir(tmp_path, torch.float32)
    master, workers = _cluster(1)
    try:
        master.initialize_model(d, num_shards=2)
        master.assign_shards()                       # both shards on the one worker, in pipeline order
        master.distribute_shards(timeout=120)
        prompt = torch.randint(0, CFG["vocab_size"], (2, 7), generator=torch.Generator().manual_seed(8))
        worker = next(iter(master.workers.values()))
        replies = {}
        for cmd in ("SCHEDULE_COMPUTATION", "RUN_INFERENCE"):
            task = f"t_{cmd}"
            master._replies[task] = __import__("queue").Queue()
            master._send(worker, cmd, pickle.dumps({"input_ids": prompt}),
                         {"task_id": task, "shard_ids": [0, 1], "reset": True, "priority": 3})
            header, payload = master._replies[task].get(timeout=60)
            assert header["command"] == "RESULT" and not header.get("error"), header
            replies[cmd] = pickle.loads(payload)
        a, b = replies["SCHEDULE_COMPUTATION"], replies["RUN_INFERENCE"]
        assert torch.equal(a["next_token"], b["next_token"]) and torch.equal(a["logits"], b["logits"])
        workers[0].shards[0].reset(); workers[0].shards[1].reset()
        direct = workers[0].schedule_computation({"input_ids": prompt}, [0, 1])
        assert torch.equal(direct["logits"].cpu(), a["logits"])
        cfg = ShardModelConfig.from_hf_dict(CFG)
        want = pipeline_step(build_pipeline(cfg, sd, 2, torch.float32), {"input_ids": prompt})
        assert torch.equal(a["next_token"], want["next_token"])
        assert float((a["logits"] - want["logits"]).abs().max() / want["logits"].abs().max()) < 1e-4
    finally:
        _shutdown(master, workers)


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_master_uses_the_fused_handoff_when_every_stage_has_its_own_gpu(tmp_path):
    """Two workers on two GPUs: after the prefill the master opens a mailbox on each stage, connects them into a ring
    and starts one persistent decode loop per stage (no message per token).  Same tokens as the per-step loop."""
    from oracle.shard_oracle import ShardModelConfig, build_pipeline, greedy_generate
    d, sd = _model_dir(tmp_path, torch.bfloat16)
    master = MasterNode(host="127.0.0.1", port=0)
    master.start()
    workers = []
    try:
        for dev in range(2):
            w = WorkerNode(host="127.0.0.1", port=0, master_address=f"127.0.0.1:{master.port}", device=dev)
            w.start()
            workers.append(w)
            deadline = time.time() + 10
            while len(master.workers) < dev + 1 and time.time() < deadline:
                time.sleep(0.02)
        master.initialize_model(d, num_shards=2)
        master.assign_shards()
        master.distribute_shards(timeout=120)
        prompt = torch.randint(0, CFG["vocab_size"], (2, 11), generator=torch.Generator().manual_seed(5))
        got = torch.tensor(master.run_inference(prompt, timeout=180, max_new_tokens=16)["token_ids"])
        assert master.last_decode_path == "fused"
        master.use_fused_handoff = False
        again = torch.tensor(master.run_inference(prompt, timeout=180, max_new_tokens=16)["token_ids"])
        assert master.last_decode_path == "per-step"
        assert torch.equal(got, again)  # same kernels, the hand-off copies bits
        # prefill and the per-step relay moved their hidden-state tiles worker -> worker over NVLink (mailbox tile slots);
        # the same request with the tiles pickled through the master gives the same tokens
        assert master.last_prefill_path == "tiles"
        master.use_direct_tiles = False
        relayed = torch.tensor(master.run_inference(prompt, timeout=180, max_new_tokens=16)["token_ids"])
        assert master.last_prefill_path == "relay" and torch.equal(relayed, again)
        master.use_direct_tiles = True
        cfg = ShardModelConfig.from_hf_dict(CFG)
        want, _, margins = greedy_generate(build_pipeline(cfg, sd, 2, torch.bfloat16), prompt, 16, record=True)
        same = (got == want).long().cumprod(1).bool()
        for b in range(2):
            assert same[b].all() or float(margins[b, int((~same[b]).float().argmax())]) < 2 ** -6
        # several requests at once = micro-batches of one pipelined loop (prompts of different lengths)
        g = torch.Generator().manual_seed(9)
        prompts = [torch.randint(0, CFG["vocab_size"], (1, n), generator=g) for n in (7, 13, 10)]
        master.use_fused_handoff = True
        many = torch.tensor(master.run_many(prompts, timeout=180, max_new_tokens=12)["token_ids"])  # [3, 1, 12]
        assert master.last_decode_path == "fused"
        master.use_fused_handoff = False
        relay = torch.tensor(master.run_many(prompts, timeout=180, max_new_tokens=12)["token_ids"])
        assert master.last_decode_path == "per-step" and torch.equal(many, relay)
        for m, pr in enumerate(prompts):
            alone = torch.tensor(master.run_inference(pr, timeout=180, max_new_tokens=12)["token_ids"])
            assert torch.equal(alone, many[m])
    finally:
        _shutdown(master, workers)
