"""CPU: the reference-shaped host API (names, signatures, error behaviour) -- reads like the reference's own
tests (reference tests/worker/test_node.py, tests/model/test_shard_manager.py, tests/network/test_protocol.py)."""
import io
import json
import os
import pickle
import socket
import threading
from unittest.mock import MagicMock

import pytest
import torch

from src.engine import cabi
from src.engine.shard_exec import CONFIG_KEY, StageGeometry, decode_config, encode_config, rope_tables
from src.model import random_init
from src.model.shard_manager import ModelShardManager, contiguous_partition, partition_state_dict
from src.network.protocol import HEADER_SIZE, MessageProtocol
from src.worker.node import ModelShard, WorkerNode

NO_GPU = not torch.cuda.is_available()


# ---- ModelShard / WorkerNode surface (reference tests/worker/test_node.py) ----------------------------
class TestModelShard:
    def setup_method(self):
        self.shard = ModelShard(1, {"weight": torch.randn(3, 3)})

    def test_init(self):
        assert self.shard.shard_id == 1 and list(self.shard.parameters) == ["weight"]
        assert self.shard.device == ("cuda" if torch.cuda.is_available() else "cpu")

    def test_to_device(self):
        self.shard.to_device()
        for p in self.shard.parameters.values():
            assert p.device.type == self.shard.device

    @pytest.mark.skipif(not NO_GPU, reason="loud-failure check for machines without CUDA")
    def test_compute_without_gpu_fails_loudly(self):
        with pytest.raises(cabi.DllmError, match="no CPU fallback"):
            self.shard.compute({"weight": torch.randn(1, 3)})

    def test_compute_ignores_unknown_keys(self):
        assert self.shard.compute({"something_else": torch.randn(1, 3)}) == {}

    def test_config_travels_with_parameters(self):
        cfg = random_init.NAMED_CONFIGS["tinyllama-1.1b"]
        shard = ModelShard(0, {CONFIG_KEY: encode_config(cfg)})
        assert shard.config == cfg and decode_config(encode_config(cfg)) == cfg


class TestWorkerNode:
    def setup_method(self):
        self.worker = WorkerNode()

    def test_init(self):
        w = self.worker
        assert (w.host, w.port, w.master_address, w.shards, w.running) == ("0.0.0.0", 65433, None, {}, False)

    def test_start_does_not_block_and_stop(self):
        w = WorkerNode(host="127.0.0.1", port=0)
        w.start()
        assert w.running
        w.stop()
        assert not w.running

    def test_load_and_unload_shard(self):
        buf = io.BytesIO()
        torch.save({"weight": torch.randn(3, 3)}, buf)
        assert self.worker.load_shard(1, buf.getvalue()) is True
        assert 1 in self.worker.shards
        assert self.worker.load_shard(2, pickle.dumps({"weight": torch.randn(3, 3)})) is True  # the reference's codec
        assert self.worker.load_shard(3, b"not a shard") is False
        assert self.worker.unload_shard(1) is True and self.worker.unload_shard(1) is False
        assert 1 not in self.worker.shards

    def test_unknown_shard_raises_keyerror(self):
        """reference tests/worker/test_node.py:114-123 expects KeyError after an unload."""
        self.worker.shards[1] = ModelShard(1, {"weight": torch.randn(3, 3)})
        self.worker.unload_shard(1)
        with pytest.raises(KeyError):
            self.worker.schedule_computation({"weight": torch.randn(1, 3)}, [1])


# ---- shard files (reference tests/model/test_shard_manager.py) -------------------------------------------
def _write_model(tmp_path, cfg, sd):
    os.makedirs(tmp_path, exist_ok=True)
    torch.save(sd, os.path.join(tmp_path, "pytorch_model.bin"))
    with open(os.path.join(tmp_path, "config.json"), "w") as f:
        json.dump(cfg, f)


def test_shard_manager_reference_case(tmp_path):
    d = str(tmp_path / "m")
    _write_model(d, {"model_type": "test"}, {"weight1": torch.randn(3, 3), "weight2": torch.randn(3, 3)})
    mgr = ModelShardManager(d, num_shards=2)
    shard_dir = mgr.shard_model()
    files = [f for f in os.listdir(shard_dir) if f.startswith("shard_") and f.endswith(".pt")]
    assert len(files) == 2 and os.path.exists(os.path.join(shard_dir, "shard_info.json"))
    paths = mgr.get_shard_paths()
    assert len(paths) == 2 and all(p.startswith(d) for p in paths)
    sd, cfg = ModelShardManager.reconstruct_model(paths, os.path.join(shard_dir, "config.json"))
    assert "weight1" in sd and "weight2" in sd and cfg["model_type"] == "test"  # the reference loses both keys


def test_shard_manager_contiguous_and_complete(tmp_path):
    cfg = dict(model_type="llama", hidden_size=64, num_hidden_layers=5, num_attention_heads=2, num_key_value_heads=1,
               intermediate_size=96, vocab_size=50, max_position_embeddings=32, rms_norm_eps=1e-5, rope_theta=1e4)
    sd = random_init.random_stage_parameters(cfg, range(5), True, True, dtype=torch.float32)
    prefixed = {("model." + k if k != "lm_head.weight" else k): v for k, v in sd.items()}  # LlamaForCausalLM keys
    d = str(tmp_path / "llama")
    _write_model(d, cfg, prefixed)
    mgr = ModelShardManager(d, 3)
    mgr.shard_model()
    shards = [torch.load(p, weights_only=True) for p in mgr.get_shard_paths()]
    layers = [sorted({int(k.split(".")[1]) for k in s if k.startswith("layers.")}) for s in shards]
    assert layers == [[0, 1], [2, 3], [4]]  # contiguous (the reference would give [0,3],[1,4],[2])
    assert "embed_tokens.weight" in shards[0] and "norm.weight" in shards[2] and "lm_head.weight" in shards[2]
    assert all(decode_config(s[CONFIG_KEY]) == cfg for s in shards)
    full, _ = ModelShardManager.reconstruct_model(mgr.get_shard_paths(), os.path.join(mgr.shard_dir, "config.json"))
    assert set(full) == set(sd) and all(torch.equal(full[k], sd[k]) for k in sd)


def test_tied_head_is_materialised_on_last_shard():
    cfg = random_init.NAMED_CONFIGS["gpt2-small"] | {"n_embd": 32, "n_layer": 2, "n_head": 2, "n_inner": 64,
                                                     "vocab_size": 40, "n_positions": 16}
    sd = random_init.random_stage_parameters(cfg, range(2), True, True, dtype=torch.float32)
    sd.pop("lm_head.weight")
    parts = partition_state_dict({"transformer." + k: v for k, v in sd.items()}, 2)
    assert torch.equal(parts[1]["lm_head.weight"], sd["wte.weight"]) and "wte.weight" in parts[0]


def test_partition_matches_oracle_partition():
    from oracle.shard_oracle import contiguous_partition as oracle_partition
    for L, S in [(12, 2), (22, 4), (32, 8), (80, 8), (48, 4), (7, 3)]:
        assert contiguous_partition(L, S) == oracle_partition(L, S)


def test_random_init_matches_oracle_generator():
    from oracle.shard_oracle import ShardModelConfig, make_state_dict
    hf = dict(model_type="qwen2", hidden_size=64, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
              intermediate_size=96, vocab_size=50, max_position_embeddings=32, rms_norm_eps=1e-6, rope_theta=1e6)
    ours = random_init.random_stage_parameters(hf, range(2), True, True, 1234, torch.float32, init="full")
    ref = make_state_dict(ShardModelConfig.from_hf_dict(hf), 1234, torch.float32, init="full")
    assert set(ours) == set(ref) and all(torch.equal(ours[k], ref[k]) for k in ref)


def test_rope_tables_match_oracle():
    from oracle.shard_oracle import ShardModelConfig, rope_tables as oracle_rope
    cfg = ShardModelConfig("llama", 256, 1, 4, 2, 64, 64, 10, 128, 1e-5, 500000.0)
    c, s = rope_tables(64, 500000.0, 128)
    oc, os_ = oracle_rope(cfg, 128)
    assert torch.equal(c, oc) and torch.equal(s, os_)


def test_geometry_from_named_configs():
    g = StageGeometry.from_hf(random_init.NAMED_CONFIGS["llama3-8b"])
    assert (g.hidden, g.n_heads, g.n_kv_heads, g.head_dim, g.ffn, g.vocab) == (4096, 32, 8, 128, 14336, 128256)
    g = StageGeometry.from_hf(random_init.NAMED_CONFIGS["gpt2-small"])
    assert (g.model_type, g.head_dim, g.ffn, g.tie_head) == ("gpt2", 64, 3072, True)
    with pytest.raises(ValueError):
        StageGeometry.from_hf({"model_type": "opt"})


# ---- wire protocol (reference tests/network/test_protocol.py) ---------------------------------------------
class TestMessageProtocol:
    def test_send_on_class_like_the_reference_tests(self):
        sock = MagicMock()
        assert MessageProtocol.send_message(sock, "TEST_COMMAND") is True
        sock.sendall.assert_called()
        sock = MagicMock()
        assert MessageProtocol.send_message(sock, "TEST_COMMAND", payload=b"test_payload") is True
        assert sock.sendall.call_count == 2

    def test_tcp_round_trip(self):
        a, b = socket.socketpair()
        proto = MessageProtocol()
        payload = os.urandom(300_000)
        t = threading.Thread(target=lambda: proto.send_message(a, "RESULT", payload, {"task_id": "t1"}))
        t.start()
        header, got = proto.receive_message(b, timeout=10)
        t.join()
        assert header["command"] == "RESULT" and header["task_id"] == "t1" and got == payload
        MessageProtocol.send_message(a, "HEARTBEAT", metadata={"timestamp": 1.0})
        header, got = MessageProtocol.receive_message(b, timeout=10)
        assert header["command"] == "HEARTBEAT" and got is None
        a.close(); b.close()

    def test_header_length_field_width(self):
        sock = MagicMock()
        MessageProtocol.send_message(sock, "X")
        first = sock.sendall.call_args_list[0][0][0]
        assert int(first[:HEADER_SIZE].decode().strip()) == len(first) - HEADER_SIZE

    def test_timeout_bad_header_and_closed_connection(self):
        sock = MagicMock()
        sock.recv.side_effect = socket.timeout()
        with pytest.raises(TimeoutError):
            MessageProtocol.receive_message(sock)
        sock = MagicMock()
        sock.recv.side_effect = [b"invalid   ", b"header"]
        with pytest.raises(ValueError):
            MessageProtocol.receive_message(sock)
        sock = MagicMock()
        sock.recv.return_value = b""
        with pytest.raises(ConnectionError):
            MessageProtocol.receive_message(sock)

    def test_zmq_router_dealer_round_trip(self):
        zmq = pytest.importorskip("zmq")
        ctx = zmq.Context.instance()
        server, client = MessageProtocol(ctx), MessageProtocol(ctx)
        router = server.setup_zmq_socket(zmq.ROUTER, "tcp://127.0.0.1:*")  # binds (the reference never did)
        endpoint = router.getsockopt_string(zmq.LAST_ENDPOINT)
        dealer = client.setup_zmq_socket(zmq.DEALER, endpoint)
        assert client.send_message(dealer, "REGISTER", metadata={"capabilities": {"has_gpu": False}})
        header, payload = server.receive_message(router, timeout=5)
        assert header["command"] == "REGISTER" and payload is None and header["identity"]
        assert server.send_message(router, "LOAD_SHARD", b"abc", {"identity": header["identity"], "shard_id": 3})
        header, payload = client.receive_message(dealer, timeout=5)
        assert header["command"] == "LOAD_SHARD" and header["shard_id"] == 3 and payload == b"abc"
        router.close(); dealer.close()


def test_balanced_partition_accounts_for_the_head():
    from src.model.shard_manager import balanced_partition, contiguous_partition
    # Llama-3-8B: the lm_head streams as many bytes as 2.4 layers
    for stages, biggest, last in ((2, 17, 15), (4, 9, 6), (8, 5, 2)):
        parts = balanced_partition(32, stages, 2.4)
        assert max(len(r) for r in parts) == biggest and len(parts[-1]) == last and len(parts) == stages
        assert [i for r in parts for i in r] == list(range(32))  # contiguous, complete, in order
    assert balanced_partition(32, 1, 2.4) == contiguous_partition(32, 1)
    assert balanced_partition(12, 4, 0.0) == contiguous_partition(12, 4)
    assert [len(r) for r in balanced_partition(4, 4, 5.0)] == [1, 1, 1, 1]


def _tiny_checkpoint(tmp_path, model_type, sharded):
    """A small safetensors checkpoint with HF key prefixes (one file, or two files + index)."""
    import json as _json
    from safetensors.torch import save_file
    from src.model import random_init
    if model_type == "llama":
        cfg = dict(model_type="llama", hidden_size=64, num_hidden_layers=5, num_attention_heads=4, num_key_value_heads=2,
                   intermediate_size=128, vocab_size=50, max_position_embeddings=32, rms_norm_eps=1e-5, rope_theta=10000.0)
        sd = random_init.random_stage_parameters(cfg, range(5), True, True, 1, torch.float32, init="full")
        full = {("model." + k if k != "lm_head.weight" else k): v.contiguous() for k, v in sd.items()}
    else:
        cfg = dict(model_type="gpt2", n_embd=64, n_layer=4, n_head=4, n_inner=256, vocab_size=50, n_positions=32,
                   layer_norm_epsilon=1e-5, tie_word_embeddings=True)
        sd = random_init.random_stage_parameters(cfg, range(4), True, True, 1, torch.float32, init="full")
        sd.pop("lm_head.weight", None)  # tied: the checkpoint holds the embedding only
        full = {"transformer." + k: v.contiguous() for k, v in sd.items()}
    d = str(tmp_path / f"ckpt_{model_type}_{int(sharded)}")
    os.makedirs(d, exist_ok=True)
    if sharded:
        keys = sorted(full)
        half = len(keys) // 2
        files = {"model-00001-of-00002.safetensors": keys[:half], "model-00002-of-00002.safetensors": keys[half:]}
        for fname, ks in files.items():
            save_file({k: full[k] for k in ks}, os.path.join(d, fname))
        with open(os.path.join(d, "model.safetensors.index.json"), "w") as f:
            _json.dump({"weight_map": {k: fname for fname, ks in files.items() for k in ks}}, f)
    else:
        save_file(full, os.path.join(d, "model.safetensors"))
    with open(os.path.join(d, "config.json"), "w") as f:
        _json.dump(cfg, f)
    return d, cfg, full


@pytest.mark.parametrize("model_type", ["llama", "gpt2"])
@pytest.mark.parametrize("sharded", [False, True])
def test_stage_loader_reads_exactly_the_partition(tmp_path, model_type, sharded):
    """`load_stage_from_safetensors` gives each stage the tensors `partition_state_dict` gives it (one file or an
    index of files; a tied head is materialised on the last stage), without reading the other stages' tensors."""
    from src.engine.shard_exec import CONFIG_KEY, decode_config
    from src.model.shard_manager import load_stage_from_safetensors, partition_state_dict
    d, cfg, full = _tiny_checkpoint(tmp_path, model_type, sharded)
    for n in (1, 2, 3):
        want = partition_state_dict(full, n, tie_head=bool(cfg.get("tie_word_embeddings", False)))
        for i in range(n):
            got = load_stage_from_safetensors(d, i, n)
            assert decode_config(got.pop(CONFIG_KEY)) == cfg
            assert set(got) == set(want[i])
            assert all(torch.equal(got[k], want[i][k]) for k in got)
    bf = load_stage_from_safetensors(d, 0, 2, torch.bfloat16)
    assert all(v.dtype == torch.bfloat16 for k, v in bf.items() if k != CONFIG_KEY)
    with pytest.raises(FileNotFoundError):
        load_stage_from_safetensors(str(tmp_path), 0, 2)


@pytest.mark.parametrize("sharded", [False, True])
def test_shard_manager_reads_multi_file_checkpoints_and_can_skip_the_rewrite(tmp_path, sharded):
    """Checkpoints that come as an index of safetensors files (Llama-3-8B and larger) shard like single-file ones, and a
    master whose workers load by path plans the partition without reading a tensor or writing `shards/*.pt`."""
    from src.master.node import MasterNode
    from src.model.shard_manager import ModelShardManager, partition_state_dict
    d, cfg, full = _tiny_checkpoint(tmp_path, "llama", sharded)
    mgr = ModelShardManager(d, 2)
    state = mgr._load_state_dict()
    assert set(state) == set(full) and all(torch.equal(state[k], full[k]) for k in full)
    mgr.shard_model()
    want = partition_state_dict(full, 2, tie_head=False)
    for i, path in enumerate(mgr.get_shard_paths()):
        got = torch.load(path, map_location="cpu", weights_only=True)
        got.pop("__config__", None)
        assert set(got) == set(want[i])
    # by path: nothing read, nothing re-written
    d2, _, _ = _tiny_checkpoint(tmp_path / "again", "llama", sharded)
    m = MasterNode(host="127.0.0.1", port=0)
    out = m.initialize_model(d2, num_shards=2, by_path=True)
    assert os.path.exists(os.path.join(out, "config.json")) and not any(f.endswith(".pt") for f in os.listdir(out))
    assert m.shard_manager.num_shards == 2 and len(m.shard_manager.get_shard_paths()) == 2


def test_block_partition_cuts_inside_layers_and_balances_the_heaviest_stage():
    """Stages are contiguous runs of attention | MLP blocks: every block exactly once and in order, a boundary may fall
    inside a layer, the lm_head weighs on the last stage and the embedding phase on the first, and no other contiguous
    split has a cheaper heaviest stage (checked by brute force on a small case)."""
    import itertools
    from src.model import random_init
    from src.model.shard_manager import block_partition, decode_block_times
    a, m, h, f = decode_block_times(random_init.NAMED_CONFIGS["llama3-8b"], 1, 512)
    assert 0.3 < a / (a + m) < 0.4 and h > a + m          # by time an attention block is a third of a layer, not 0.19
    for n_layers, n_stages in ((32, 8), (32, 4), (6, 4), (5, 3)):
        stages = block_partition(n_layers, n_stages, a, m, h, f)
        flat = []
        for st in stages:
            for layer, blk in st:
                flat += [(layer, "attn"), (layer, "mlp")] if blk == "full" else [(layer, blk)]
        assert flat == [(i, b) for i in range(n_layers) for b in ("attn", "mlp")]
        assert all(st for st in stages)
    w = {"attn": a, "mlp": m}
    units = [(i, b) for i in range(5) for b in ("attn", "mlp")]

    def heaviest(bounds):
        cuts = [0, *bounds, len(units)]
        t = [sum(w[b] for _, b in units[x:y]) for x, y in zip(cuts[:-1], cuts[1:])]
        t[0] += f
        t[-1] += h
        return max(t)
    best = min(heaviest(c) for c in itertools.combinations(range(1, len(units)), 2))
    got = block_partition(5, 3, a, m, h, f)
    sizes = [sum(2 if blk == "full" else 1 for _, blk in st) for st in got]
    assert abs(heaviest([sizes[0], sizes[0] + sizes[1]]) - best) < 1e-9
    assert any(blk != "full" for st in block_partition(32, 8, a, m, h, f) for _, blk in st)   # it does cut inside layers

