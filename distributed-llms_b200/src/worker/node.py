"""Worker side of the drop-in: `ModelShard` and `WorkerNode` with the reference's names and signatures
(reference src/worker/node.py:13-32, 34-302), the body of `compute` replaced by the B200 shard forward.

Kept from the reference: class / method names, constructor arguments and attributes, the dict-in /
dict-out `compute` contract, the message names, `load_shard -> bool`, `unload_shard -> bool`.
Fixed behind the names (each noted where it happens):
  * `compute` runs a real layer-shard forward when the shard carries a model config; without one it
    keeps the reference's per-key matmul placeholder, on the GPU (`dllm_op_legacy_matmul`);
  * `schedule_computation` raises KeyError for an unknown shard id, which is what the reference's own
    test expects (reference tests/worker/test_node.py:114-123), and *chains* shards in pipeline order
    instead of merging their outputs;
  * `start()` no longer blocks forever unless asked to (`block=True` keeps the old behaviour).
"""
from __future__ import annotations

import io
import os
import pickle
import socket
import threading
import time
from typing import Any, Dict, List, Optional

import torch

try:  # control plane only; the data path never touches zmq
    import zmq
except Exception:  # pragma: no cover
    zmq = None

from ..engine import cabi
from ..engine.shard_exec import CONFIG_KEY, ShardExecutor, decode_config
from ..network.protocol import MessageProtocol, OwnedSocket

META_KEYS = ("positions", "seq_ids")


class ModelShard:
    """One pipeline stage: a contiguous range of decoder layers (+ embedding on the first stage,
    + final norm / lm_head / greedy sampler on the last), resident on one B200.

    Reference: src/worker/node.py:13-32.  `parameters` is the HF-keyed dict a shard file holds
    (`h.N.*` / `layers.N.*`, plus `wte/wpe/embed_tokens`, `ln_f/norm`, `lm_head`); the model config
    travels with it under ``"__config__"`` (uint8 tensor of config.json bytes) or as `config=`.
    """

    def __init__(self, shard_id: int, parameters: Dict[str, torch.Tensor], config: Optional[dict] = None,
                 **limits: Any):
        self.shard_id = shard_id
        self.parameters = parameters
        self.device = "cuda" if torch.cuda.is_available() else "cpu"  # reference :17
        if config is None and CONFIG_KEY in parameters:
            config = decode_config(parameters[CONFIG_KEY])
        self.config = config
        self.limits = limits  # max_rows / max_seqs / max_context / page_size
        self._exec: Optional[ShardExecutor] = None
        self._lock = threading.Lock()

    # -- reference :19-22 -----------------------------------------------------------------------------
    def to_device(self):
        """Move parameters to the device and, for a transformer shard, build the executor (KV cache,
        block table, CUDA graphs).  On a machine without CUDA this only mirrors the reference."""
        for key in self.parameters:
            if key != CONFIG_KEY:
                self.parameters[key] = self.parameters[key].to(self.device)
        if self.config is not None and self.device == "cuda":
            self._exec = ShardExecutor(self.parameters, self.config, torch.device("cuda", torch.cuda.current_device()),
                                       **self.limits)
            # the executor may have re-laid-out tensors (GPT-2 Conv1D transposes, fp16 -> bf16);
            # `parameters` keeps the HF-keyed view for reconstruct / inspection

    @property
    def executor(self) -> ShardExecutor:
        if self._exec is None:
            if self.config is None:
                raise RuntimeError("this shard carries no model config (placeholder shard)")
            if not torch.cuda.is_available():
                raise cabi.DllmError("ModelShard.compute needs a CUDA device: there is no CPU fallback")
            self.to_device()
        return self._exec

    @property
    def has_head(self) -> bool:
        return "lm_head.weight" in self.parameters

    def reset(self, seq_ids=None) -> None:
        if self._exec is not None:
            self._exec.reset(seq_ids)

    # -- reference :24-32 -----------------------------------------------------------------------------
    def compute(self, inputs: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """Run this stage.

        inputs : ``input_ids`` int [B,T] (first stage) or ``hidden_states`` [B,T,d];
                 optional ``positions`` int [B,T] (default: continue each sequence),
                 optional ``seq_ids`` int [B] (default 0..B-1).
        returns: ``hidden_states`` [B,T,d] + ``positions`` + ``seq_ids`` (the hand-off to the next
                 stage), or on the last stage ``logits`` fp32 [B,V] (last position of each row) and
                 ``next_token`` int64 [B] (greedy).
        A dict without those keys takes the reference's placeholder path (per-key matmul)."""
        if self.config is None or not ("input_ids" in inputs or "hidden_states" in inputs):
            return self._legacy_compute(inputs)
        ex = self.executor
        x = inputs["input_ids"] if (ex.has_embed and "input_ids" in inputs) else inputs.get("hidden_states")
        if x is None:
            raise KeyError("this stage needs 'hidden_states' (it does not own the embedding)")
        if x.dim() == 1:
            x = x[None, :]
        B, T = int(x.shape[0]), int(x.shape[1])
        seq_ids = inputs.get("seq_ids")
        seq_list = [int(s) for s in seq_ids] if seq_ids is not None else list(range(B))
        if "positions" in inputs:
            pos = inputs["positions"].reshape(B, T)
        else:
            pos = torch.stack([torch.arange(ex.seq_len(s), ex.seq_len(s) + T) for s in seq_list])
        out = ex.forward(x, pos, seq_list)
        if "hidden_states" in out:
            out["positions"] = pos.to(torch.int32)
            out["seq_ids"] = torch.tensor(seq_list, dtype=torch.int32)
        return out

    def decode_loop(self, batch: int, n_steps: int) -> torch.Tensor:
        """Device-resident greedy loop for a single-stage model (first and last stage in one)."""
        return self.executor.decode_loop(batch, n_steps)

    def _legacy_compute(self, inputs: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """`torch.matmul(inputs[key], parameters[key])` for every shared key (reference :28-32), run by
        the library's fp32 kernel; no CUDA device -> error (no CPU path)."""
        results: Dict[str, torch.Tensor] = {}
        L = None
        for key in self.parameters:
            if key == CONFIG_KEY or key not in inputs:
                continue
            if L is None:
                L = cabi.lib()
                if cabi.device_count() == 0 or not torch.cuda.is_available():
                    raise cabi.DllmError("ModelShard.compute needs a CUDA device: there is no CPU fallback")
            w = self.parameters[key].to("cuda", torch.float32).contiguous()
            a = inputs[key].to("cuda", torch.float32).contiguous()
            if w.dim() != 2 or a.shape[-1] != w.shape[0]:
                raise ValueError(f"shape mismatch for {key!r}: {tuple(a.shape)} @ {tuple(w.shape)}")
            a2 = a.reshape(-1, w.shape[0])
            out = torch.empty(a2.shape[0], w.shape[1], dtype=torch.float32, device="cuda")
            cabi.check(L.dllm_op_legacy_matmul(a2.data_ptr(), w.data_ptr(), out.data_ptr(), a2.shape[0], w.shape[0],
                                               w.shape[1], torch.cuda.current_stream().cuda_stream))
            results[key] = out.reshape(*a.shape[:-1], w.shape[1])
        return results


def _load_parameters(shard_data: bytes) -> Dict[str, torch.Tensor]:
    """One codec for shard bytes.  The reference's master sends `torch.save` bytes and its worker
    `pickle.loads` them; accept both, torch.save first."""
    try:
        return torch.load(io.BytesIO(shard_data), map_location="cpu", weights_only=True)
    except Exception:
        return pickle.loads(shard_data)


class WorkerNode:
    """Reference src/worker/node.py:34-302."""

    def __init__(self, host: str = "0.0.0.0", port: int = 65433, master_address: Optional[str] = None,
                 device: Optional[int] = None):
        self.host = host
        self.port = port
        self.master_address = master_address
        self.device = device  # CUDA device index of this worker (None: the process's current device)
        self._inbox = None    # fused hand-off: this stage's mailbox and the view of the next stage's
        self._outbox = None
        self._tile_in = None  # prefill / relay hand-off: a one-slot mailbox that holds a whole [B*T, d] tile, and the next stage's
        self._tile_out = None
        self.shards: Dict[int, ModelShard] = {}
        self.running = False
        self.server_socket = None
        self.master_socket = None
        self.zmq_context = zmq.Context.instance() if zmq is not None else None
        self._shards_lock = threading.Lock()  # reference mutates self.shards from several threads unlocked
        self._protocol: Optional[MessageProtocol] = None
        self._owned: Optional[OwnedSocket] = None  # the DEALER socket to the master: owned by the _handle_master thread

    # -- life cycle (reference :45-99) ------------------------------------------------------------------
    def start(self, block: bool = False):
        self.server_socket = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.server_socket.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        bound = False
        for _ in range(5):  # reference :51-65: port-bump retry
            try:
                self.server_socket.bind((self.host, self.port))
                bound = True
                break
            except OSError as e:
                if "Address already in use" in str(e):
                    self.port += 1
                    continue
                raise
        if not bound:
            raise RuntimeError(f"Could not bind to any port in range {self.port - 5}-{self.port - 1}")
        self.server_socket.listen(5)
        self.running = True
        print(f"Worker listening on {self.host}:{self.port}")
        if self.master_address:
            self._connect_to_master()
        self.accept_thread = threading.Thread(target=self._accept_connections, daemon=True)
        self.accept_thread.start()
        self.heartbeat_thread = threading.Thread(target=self._send_heartbeat, daemon=True)
        self.heartbeat_thread.start()
        if block:  # the reference always blocked here
            try:
                while self.running:
                    time.sleep(1)
            except KeyboardInterrupt:
                self.stop()

    def stop(self):
        self.running = False
        if self.server_socket:
            self.server_socket.close()
        t = getattr(self, "_master_thread", None)
        if t is not None and t is not threading.current_thread():
            t.join(timeout=3.0)   # that thread owns the master socket and closes it on its way out

    # -- control plane (reference :101-187) ---------------------------------------------------------------
    def _connect_to_master(self, retries: int = 3):
        """DEALER socket to the master's bound ROUTER (the reference connects a REQ to a ROUTER that
        never binds)."""
        for attempt in range(retries):
            try:
                host, port = self.master_address.split(":")
                self._protocol = MessageProtocol(zmq_context=self.zmq_context)
                self.master_socket = self._protocol.setup_zmq_socket(zmq.DEALER, f"tcp://{host}:{int(port)}")
                capabilities = {
                    "has_gpu": torch.cuda.is_available(),
                    "device": "cuda" if torch.cuda.is_available() else "cpu",
                    "memory": torch.cuda.get_device_properties(self.device or 0).total_memory if torch.cuda.is_available() else 0,
                    "pid": os.getpid(),
                    "device_index": self._device_index() if torch.cuda.is_available() else -1,
                }
                self._owned = OwnedSocket(self.zmq_context, self.master_socket)
                self._owned.send("REGISTER", metadata={"capabilities": capabilities})  # on the wire once the owner polls
                self._master_thread = threading.Thread(target=self._handle_master, daemon=True)
                self._master_thread.start()
                print(f"Successfully connected to master at {self.master_address}")
                return
            except Exception as e:  # bounded retry instead of the reference's unbounded recursion (:130-136)
                print(f"Error connecting to master: {e}")
                if not self.running or attempt == retries - 1:
                    return
                time.sleep(1)

    def _accept_connections(self):
        while self.running:
            try:
                client_socket, address = self.server_socket.accept()
                threading.Thread(target=self._handle_client, args=(client_socket, address), daemon=True).start()
            except Exception as e:
                if self.running:
                    print(f"Error accepting connection: {e}")

    def _handle_client(self, client_socket, address):
        protocol = MessageProtocol()  # raw-TCP framing
        try:
            while self.running:
                header, payload = protocol.receive_message(client_socket)
                if not header:
                    break
                self._handle_command(client_socket, header.get("command", ""), header, payload, protocol)
        except Exception as e:
            print(f"Error handling client {address}: {e}")
        finally:
            client_socket.close()

    def _handle_master(self):
        try:
            while self.running:
                try:
                    header, payload = self._owned.receive(timeout=0.5)
                except TimeoutError:
                    continue
                if not header:
                    continue
                self._handle_command(self.master_socket, header.get("command", ""), header, payload, self._protocol)
        except Exception as e:
            if self.running:
                print(f"Error handling master connection: {e}")
        finally:
            owned, self._owned, self.master_socket = self._owned, None, None
            if owned is not None:
                owned.close()

    # -- shard-execution entry point (reference :189-261) ---------------------------------------------------
    def _device_index(self) -> int:
        return self.device if self.device is not None else torch.cuda.current_device()

    def _handle_command(self, sock, command, header, payload, protocol: Optional[MessageProtocol] = None):
        if self.device is not None and torch.cuda.is_available():
            with torch.cuda.device(self.device):  # the current device is per thread
                return self._dispatch_command(sock, command, header, payload, protocol)
        return self._dispatch_command(sock, command, header, payload, protocol)

    # fused hand-off (include/dllm_b200.h "mailboxes"): the master opens a mailbox on every stage, tells every stage
    # where the next one's is, seeds stage 0 and then starts ONE persistent decode loop per stage
    def _mailbox_command(self, command: str, header: dict, payload: bytes) -> dict:
        from ..network.handoff import Mailbox, connect_mailbox
        args = pickle.loads(payload) if payload else {}
        dev = self._device_index()
        if command == "TILE_OPEN":   # this stage's inbox for whole hidden-state tiles (prefill, per-step relay)
            hidden = self.shards[header["shard_id"]].executor.geom.hidden
            if self._tile_in is not None:
                self._tile_in.close()
            self._tile_in = Mailbox.create(dev, 1, int(header["rows"]), hidden)
            return {"ipc": self._tile_in.export_handle(), "pid": os.getpid(), "device_index": dev, "n_slots": 1,
                    "rows": self._tile_in.rows, "hidden": hidden}
        if command == "TILE_CONNECT":
            if self._tile_out is not None and self._tile_out is not self._tile_in:
                self._tile_out.close()
            self._tile_out = connect_mailbox(dev, args)
            return {}
        if command == "TILE_CLOSE":
            for box in (self._tile_out, self._tile_in):
                if box is not None:
                    box.close()
            self._tile_in = self._tile_out = None
            return {}
        if command == "MAILBOX_OPEN":
            shard = self.shards[header["shard_id"]]
            hidden = shard.executor.geom.hidden
            self._close_mailboxes()
            self._inbox = Mailbox.create(dev, int(header["n_slots"]), int(header["rows"]), hidden)
            return {"ipc": self._inbox.export_handle(), "pid": os.getpid(), "device_index": dev,
                    "n_slots": self._inbox.n_slots, "rows": self._inbox.rows, "hidden": hidden}
        if self._inbox is None:
            raise RuntimeError("no mailbox is open on this worker")
        if command == "MAILBOX_CONNECT":
            mine = args["ipc"] == self._inbox.export_handle() and args.get("pid") == os.getpid()
            self._outbox = self._inbox if mine else connect_mailbox(dev, args)
            return {}
        if command == "MAILBOX_POST":
            for slot, toks in enumerate(args["tokens"]):
                self._inbox.post_tokens(slot, torch.as_tensor(toks, dtype=torch.int32), int(args["ticket"]))
            return {}
        if command in ("PREPARE_DECODE_LOOP", "RUN_DECODE_LOOP"):
            if self._outbox is None:
                raise RuntimeError("the mailbox of the next stage is not connected")
            ex = self.shards[header["shard_id"]].executor
            n_micro, rows, n_steps = int(header["n_micro"]), int(header["rows"]), int(header["n_steps"])
            seq = torch.as_tensor(args["seq_ids"], dtype=torch.int32).reshape(n_micro, rows)
            pos = args.get("positions")
            if pos is None:  # continue every sequence where its cache ends
                pos = torch.tensor([ex.seq_len(int(s)) for s in seq.reshape(-1)], dtype=torch.int32)
            pos = torch.as_tensor(pos, dtype=torch.int32).reshape(n_micro, rows)
            if command == "PREPARE_DECODE_LOOP":
                # everything that can refuse the loop is checked (and the KV pages reserved) before ANY stage starts:
                # a stage failing after its neighbours launched would leave them waiting for tickets that never come
                ex.check_decode_pipelined(self._inbox, self._outbox, n_micro, rows, n_steps, pos, seq)
                return {"positions": pos}
            fed = ex.decode_pipelined(self._inbox, self._outbox, n_micro, rows, n_steps, int(header.get("step_base", 0)), pos, seq)
            ex.pipeline_status()  # waits for the loop; raises if a ticket never arrived (the RESULT then carries the error)
            return {"fed": fed.cpu() if fed is not None else None}
        if command == "MAILBOX_READ":
            got = [self._inbox.read_tokens(slot) for slot in range(self._inbox.n_slots)]
            return {"tokens": torch.stack([g[0] for g in got]), "tickets": [g[1] for g in got]}
        if command == "MAILBOX_CLOSE":
            self._close_mailboxes()
            return {}
        raise ValueError(f"unknown mailbox command {command}")

    def _close_mailboxes(self) -> None:
        if self._outbox is not None and self._outbox is not self._inbox:
            self._outbox.close()
        if self._inbox is not None:
            self._inbox.close()
        self._inbox = self._outbox = None

    def _dispatch_command(self, sock, command, header, payload, protocol: Optional[MessageProtocol] = None):
        protocol = protocol or MessageProtocol()
        if sock is not None and sock is self.master_socket and self._owned is not None:
            class _ViaOwner:  # replies to the master leave through the socket's owner (OwnedSocket)
                def send_message(_, _sock, command, payload=None, metadata=None, owned=self._owned):
                    return owned.send(command, payload, metadata)
            protocol = _ViaOwner()
        if command.startswith(("MAILBOX_", "TILE_")) or command in ("RUN_DECODE_LOOP", "PREPARE_DECODE_LOOP"):
            task_id = header.get("task_id")
            md = {"task_id": task_id, "stage": header.get("stage")}
            try:
                out = self._mailbox_command(command, header, payload)
                protocol.send_message(sock, "RESULT", payload=pickle.dumps(out), metadata=md)
            except Exception as e:
                md["error"] = f"{type(e).__name__}: {e}"
                protocol.send_message(sock, "RESULT", payload=pickle.dumps({}), metadata=md)
        elif command == "LOAD_SHARD":
            shard_id = header.get("shard_id")
            ok = self.load_shard(shard_id, payload)
            protocol.send_message(sock, "SHARD_LOADED", metadata={"shard_id": shard_id, "ok": ok})
        elif command == "LOAD_SHARD_FROM_PATH":  # same box / shared file system: read the stage, do not receive it
            shard_id = header.get("shard_id")
            ok = self.load_shard_from_path(shard_id, header.get("model_path"), int(header.get("num_shards", 1)),
                                           header.get("shard_file"))
            protocol.send_message(sock, "SHARD_LOADED", metadata={"shard_id": shard_id, "ok": ok})
        elif command == "UNLOAD_SHARD":
            shard_id = header.get("shard_id")
            self.unload_shard(shard_id)
            protocol.send_message(sock, "SHARD_UNLOADED", metadata={"shard_id": shard_id})
        elif command in ("RUN_INFERENCE", "SCHEDULE_COMPUTATION"):
            task_id = header.get("task_id")
            shard_ids = header.get("shard_ids", [])
            try:
                inputs = pickle.loads(payload)
                if header.get("in_tile"):  # the hidden states are already in this GPU's tile inbox (put there over NVLink)
                    B, T = header["in_tile"]
                    d = self._tile_in.hidden
                    inputs["hidden_states"] = self._tile_in.get_rows(0, int(B) * int(T)).reshape(int(B), int(T), d)
                if header.get("reset"):
                    for sid in shard_ids:
                        if sid in self.shards:
                            self.shards[sid].reset()
                # `priority` (reference :244-251) is a constant sort key there, i.e. a no-op; kept as such
                results = self.schedule_computation(inputs, [s for s in shard_ids if s in self.shards])
                if header.get("out_tile") and "hidden_states" in results:
                    # hand-off over NVLink: the tile goes straight into the next stage's inbox; only its shape travels back
                    h = results.pop("hidden_states")
                    self._tile_out.put_rows(0, h)
                    torch.cuda.current_stream(h.device).synchronize()  # the master releases the next stage on this reply
                    results["tile_shape"] = [int(h.shape[0]), int(h.shape[1])]
                results = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in results.items()}
                protocol.send_message(sock, "RESULT", payload=pickle.dumps(results), metadata={"task_id": task_id})
            except Exception as e:  # the master's scheduler is waiting for this RESULT: tell it what went wrong
                protocol.send_message(sock, "RESULT", payload=pickle.dumps({}),
                                      metadata={"task_id": task_id, "error": f"{type(e).__name__}: {e}"})

    def _send_heartbeat(self):
        while self.running:
            try:
                if self._owned is not None:
                    self._owned.send("HEARTBEAT", metadata={"timestamp": time.time()})
                time.sleep(5)
            except Exception as e:
                print(f"Error sending heartbeat: {e}")
                time.sleep(1)

    # -- direct API (reference :278-302) ---------------------------------------------------------------------
    def load_shard(self, shard_id: int, shard_data: bytes) -> bool:
        try:
            parameters = _load_parameters(shard_data)
            shard = ModelShard(shard_id, parameters)
            shard.to_device()
            with self._shards_lock:
                self.shards[shard_id] = shard
            return True
        except Exception as e:
            print(f"Error loading shard {shard_id}: {e}")
            return False

    def load_shard_from_path(self, shard_id: int, model_path: str, num_shards: int, shard_file: Optional[str] = None) -> bool:
        """Load stage `shard_id` of `num_shards` from disk: only its tensors out of `model.safetensors` (or the files of
        the index), else the shard file the master wrote.  The reference ships every shard as one message
        (src/master/node.py:213-225), 7-17 GB per stage at the BASELINE sizes."""
        from ..model.shard_manager import load_stage_from_safetensors
        try:
            dtype = torch.bfloat16 if torch.cuda.is_available() else None
            try:
                parameters = load_stage_from_safetensors(model_path, shard_id, num_shards, dtype)
            except FileNotFoundError:
                if not shard_file or not os.path.exists(shard_file):
                    raise
                parameters = torch.load(shard_file, map_location="cpu", weights_only=True, mmap=True)
            shard = ModelShard(shard_id, parameters)
            shard.to_device()
            with self._shards_lock:
                self.shards[shard_id] = shard
            return True
        except Exception as e:
            print(f"Error loading shard {shard_id} from {model_path}: {e}")
            return False

    def unload_shard(self, shard_id: int) -> bool:
        with self._shards_lock:
            if shard_id in self.shards:
                del self.shards[shard_id]
                return True
        return False

    def schedule_computation(self, inputs: Dict[str, torch.Tensor], shard_ids: List[int]) -> Dict[str, torch.Tensor]:
        """Run the given shards in pipeline order, feeding each one's hand-off into the next.  Placeholder
        shards (no model config) keep the reference's merge-by-`dict.update` behaviour.  Unknown shard id
        -> KeyError (reference tests/worker/test_node.py:114-123; the shipped code returned {})."""
        with self._shards_lock:
            shards = [self.shards[s] for s in shard_ids]  # KeyError on a missing id
        results: Dict[str, torch.Tensor] = {}
        x = inputs
        for shard in shards:
            with shard._lock:  # serialise per shard: several handler threads may call in
                out = shard.compute(x)
            if shard.config is None:
                results.update(out)
            else:
                results = out
                x = out
        return results
