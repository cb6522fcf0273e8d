"""Master side of the drop-in: `MasterNode` with the reference's surface (reference src/master/node.py:14-277) --
`start / stop / initialize_model / assign_shards / distribute_shards / run_inference` -- plus what the reference's
master lacks for the hot path to produce tokens:

  * a *scheduler* that drives the stages in pipeline order (stage 0 only gets the request; each worker chains its
    own shards; the hand-off of one worker becomes the input of the next) instead of broadcasting the same
    payload to every worker (reference :256-269);
  * a decode loop (prefill, then `max_new_tokens` steps) and a *sampler* (greedy argmax; optional temperature /
    top-k) -- the reference returns the first RESULT's raw bytes (:117-138, :208-211);
  * a blocking wait per task instead of polling a shared queue every 100 ms (:127-138).

This is the control-plane route: hidden states between workers travel through the master over ZeroMQ, which is what
the reference's topology allows (a star).  Inside one NVLink box the data plane bypasses it: stages exchange
tiles directly (src/network/handoff.py, src/engine/pipeline.py).
"""
from __future__ import annotations

import json
import os
import pickle
import threading
import time
from queue import Empty, Queue
from typing import Any, Dict, List, Optional, Sequence, Union

import torch

try:
    import zmq
except Exception:  # pragma: no cover
    zmq = None

from ..model.shard_manager import ModelShardManager
from ..network.protocol import MessageProtocol, OwnedSocket


class GreedySampler:
    """The sampler the reference's master never had.  Greedy = first index of the maximum (torch.argmax)."""

    def __init__(self, temperature: float = 0.0, top_k: int = 0, seed: int = 0):
        self.temperature, self.top_k = temperature, top_k
        self.gen = torch.Generator().manual_seed(seed)

    def __call__(self, logits: torch.Tensor) -> torch.Tensor:
        if self.temperature <= 0.0:
            return torch.argmax(logits, dim=-1)
        l = logits.float().cpu() / self.temperature
        if self.top_k > 0:
            kth = torch.topk(l, min(self.top_k, l.shape[-1]), dim=-1).values[..., -1:]
            l = l.masked_fill(l < kth, float("-inf"))
        return torch.multinomial(torch.softmax(l, dim=-1), 1, generator=self.gen).squeeze(-1)


class MasterNode:
    def __init__(self, host: str = "0.0.0.0", port: int = 65432):
        self.host = host
        self.port = port
        self.workers: Dict[str, Dict[str, Any]] = {}  # worker_id -> info (identity, capabilities, last_heartbeat, ...)
        self.model_path = None
        self.tokenizer = None
        self.shard_manager: Optional[ModelShardManager] = None
        self.shard_assignments: Dict[str, List[int]] = {}
        self.running = False
        self.server_socket = None
        self.task_queue: Queue = Queue()
        self.result_queue: Queue = Queue()            # kept for API parity; results are delivered per task below
        self.zmq_context = zmq.Context.instance() if zmq is not None else None
        self.sampler = GreedySampler()
        self._protocol: Optional[MessageProtocol] = None
        self._owned: Optional[OwnedSocket] = None     # the ROUTER socket: owned by the accept thread, see OwnedSocket
        self._workers_lock = threading.Lock()
        self._replies: Dict[str, Queue] = {}          # task_id -> queue of (header, payload) from workers
        self._finished: Dict[str, Queue] = {}         # task_id -> final result for run_inference()
        self._loaded = threading.Condition()
        self._loaded_shards: set = set()
        self.heartbeat_timeout = 30.0
        self.use_fused_handoff = True                 # decode through peer mailboxes when every stage has its own GPU
        self.use_direct_tiles = True                  # prefill / relay: hidden-state tiles go worker -> worker over NVLink
        self._tile_rows = 0                           # capacity of the tile inboxes opened on the workers (0 = none)
        self._tile_workers: List[str] = []
        self.last_prefill_path = None                 # "tiles" | "relay": how the last prompt's hidden states travelled
        self.last_decode_path = None                  # "fused" | "per-step": how the last request was decoded

    # ---- life cycle (reference :29-52) -------------------------------------------------------------------
    def start(self):
        self._protocol = MessageProtocol(zmq_context=self.zmq_context)
        self.server_socket = self._protocol.setup_zmq_socket(zmq.ROUTER, f"tcp://{self.host}:{self.port}")  # binds
        if self.port == 0:
            self.port = int(self.server_socket.getsockopt_string(zmq.LAST_ENDPOINT).rsplit(":", 1)[1])
        self._owned = OwnedSocket(self.zmq_context, self.server_socket)
        self.running = True
        print(f"Master node listening on {self.host}:{self.port}")
        self.accept_thread = threading.Thread(target=self._accept_connections, daemon=True)
        self.accept_thread.start()
        self.task_thread = threading.Thread(target=self._process_tasks, daemon=True)
        self.task_thread.start()

    def stop(self):
        self.running = False
        t = getattr(self, "accept_thread", None)
        if t is not None and t is not threading.current_thread():
            t.join(timeout=2.0)   # the accept thread owns the socket: it must be out of its poll before the close
        if self._owned is not None:
            self._owned.close()
            self._owned = None
        self.server_socket = None

    # ---- model preparation (reference :54-82) ---------------------------------------------------------------
    def initialize_model(self, model_id: str, num_shards: int, cache_dir: str = "./models", by_path: bool = False):
        """`model_id`: a local directory (config.json + pytorch_model.bin / model.safetensors / a safetensors index),
        `random:<name>` for a seeded random-init checkpoint of a known architecture (no network needed), or a Hub id
        (needs a network).  `by_path=True`: the workers will read their own stages from the checkpoint
        (`distribute_shards(by_path=True)`), so nothing is loaded or re-written here."""
        if model_id.startswith("random:"):
            self.model_path = self._materialise_random(model_id[len("random:"):], cache_dir)
        elif os.path.exists(model_id):
            self.model_path = model_id
        else:
            from ..model.downloader import download_model
            print(f"Downloading model {model_id}...")
            self.model_path = download_model(model_id, cache_dir=cache_dir)
        try:  # the tokenizer is optional: prompts may be given as token ids
            from transformers import AutoTokenizer
            self.tokenizer = AutoTokenizer.from_pretrained(self.model_path)
        except Exception:
            self.tokenizer = None
        print(f"Sharding model into {num_shards} parts...")
        self.shard_manager = ModelShardManager(self.model_path, num_shards)
        if by_path and any(os.path.exists(os.path.join(self.model_path, n))
                           for n in ("model.safetensors", "model.safetensors.index.json")):
            return self.shard_manager.plan_only()
        return self.shard_manager.shard_model()

    @staticmethod
    def _materialise_random(name: str, cache_dir: str, seed: int = 1234, dtype: torch.dtype = torch.bfloat16) -> str:
        from ..model import random_init
        cfg = random_init.NAMED_CONFIGS[name]
        path = os.path.join(cache_dir, f"random-{name}")
        os.makedirs(path, exist_ok=True)
        if not os.path.exists(os.path.join(path, "pytorch_model.bin")):
            L = cfg.get("num_hidden_layers", cfg.get("n_layer"))
            sd = random_init.random_stage_parameters(cfg, range(L), True, True, seed, dtype)
            if cfg.get("tie_word_embeddings", cfg["model_type"] == "gpt2"):
                sd.pop("lm_head.weight", None)
            torch.save(sd, os.path.join(path, "pytorch_model.bin"))
            with open(os.path.join(path, "config.json"), "w") as f:
                json.dump(cfg, f)
        return path

    # ---- placement (reference :84-115) -------------------------------------------------------------------------
    def assign_shards(self):
        """Shard i -> worker i (round-robin when there are more shards than workers; consecutive shards of one worker
        stay in pipeline order)."""
        with self._workers_lock:
            workers = list(self.workers.keys())
        if not workers:
            raise ValueError("No workers connected")
        num_shards = self.shard_manager.num_shards
        self.shard_assignments = {}
        for i in range(num_shards):
            self.shard_assignments.setdefault(workers[i % len(workers)], []).append(i)
        return self.shard_assignments

    def distribute_shards(self, wait: bool = True, timeout: float = 600.0, by_path: bool = False):
        """Give every worker its shards.  `by_path`: the workers can read the model directory themselves (same box or a
        shared file system) -- they are told where it is and load only their stage; otherwise the shard files travel
        as messages like in the reference (:106-115)."""
        shard_paths = self.shard_manager.get_shard_paths()
        wanted = set()
        for worker_id, shard_ids in self.shard_assignments.items():
            for shard_id in shard_ids:
                if by_path:
                    self._send(self.workers[worker_id], "LOAD_SHARD_FROM_PATH", None,
                               {"shard_id": shard_id, "model_path": os.path.abspath(self.model_path),
                                "num_shards": self.shard_manager.num_shards, "shard_file": os.path.abspath(shard_paths[shard_id])})
                else:
                    self._send_shard(self.workers[worker_id], shard_id, shard_paths[shard_id])
                wanted.add(shard_id)
        if wait:
            deadline = time.time() + timeout
            with self._loaded:
                while not wanted <= self._loaded_shards:
                    left = deadline - time.time()
                    if left <= 0:
                        raise TimeoutError(f"workers did not confirm shards {sorted(wanted - self._loaded_shards)}")
                    self._loaded.wait(left)

    def _send_shard(self, worker, shard_id, shard_path):
        print(f"Sending shard {shard_id} to worker")
        with open(shard_path, "rb") as f:
            shard_data = f.read()
        self._send(worker, "LOAD_SHARD", shard_data, {"shard_id": shard_id})

    def _send(self, worker, command, payload=None, metadata=None) -> bool:
        md = dict(metadata or {})
        md["identity"] = worker["identity"]
        return self._owned.send(command, payload, md)  # any thread: the accept thread puts it on the wire

    # ---- inference (reference :117-138, :227-277) -----------------------------------------------------------------
    def run_inference(self, input_text: Union[str, Sequence[int], torch.Tensor], timeout: int = 60,
                      max_new_tokens: int = 32):
        """Queue a request and wait for it.  Returns {"token_ids": [...], "text": str | None}."""
        task_id = f"task_{int(time.time() * 1000)}_{threading.get_ident() & 0xffff}"
        done: Queue = Queue()
        self._finished[task_id] = done
        self.task_queue.put((task_id, input_text, max_new_tokens))
        try:
            result = done.get(timeout=timeout)
        except Empty:
            raise TimeoutError(f"Inference timed out after {timeout} seconds")
        finally:
            self._finished.pop(task_id, None)
        if isinstance(result, Exception):
            raise result
        return result

    def _stage_order(self) -> List[tuple]:
        """[(worker_id, [shard ids in pipeline order]), ...] sorted by the first shard a worker holds."""
        order = sorted(((min(s), w, sorted(s)) for w, s in self.shard_assignments.items() if s))
        flat = [i for _, _, s in order for i in s]
        if flat != sorted(flat):
            raise RuntimeError("shard assignment is not a pipeline: a worker holds non-consecutive shards")
        return [(w, s) for _, w, s in order]

    def _ensure_tiles(self, task_id: str, rows: int) -> bool:
        """Worker-to-worker hand-off of hidden-state tiles: every stage but the first opens an inbox for `rows` rows on its
        GPU and its predecessor maps it (CUDA IPC / peer access).  Needs one shard per worker and every worker on its
        own GPU, like the fused decode loop; otherwise (and on any refusal) the tiles keep travelling through the
        master as pickled RESULT payloads, which is what the reference's star topology does (src/master/node.py:256-269)."""
        if not self.use_direct_tiles:
            return False
        order = self._stage_order()
        ids = [w for w, _ in order]
        if len(order) < 2 or any(len(s) != 1 for _, s in order):
            return False
        places = [(self.workers[w].get("pid"), self.workers[w].get("device_index")) for w in ids]
        if any(p[0] is None or p[1] is None or p[1] < 0 for p in places) or len(set(places)) != len(places):
            return False
        if self._tile_rows >= rows and self._tile_workers == ids:
            return True
        try:
            infos = {}
            for st in range(1, len(order)):
                infos[st] = self._request(self.workers[ids[st]], task_id, "TILE_OPEN", None, {"shard_id": order[st][1][0], "rows": rows})
            for st in range(len(order) - 1):
                self._request(self.workers[ids[st]], task_id, "TILE_CONNECT", infos[st + 1], {})
        except (RuntimeError, TimeoutError) as e:
            print(f"direct tile hand-off unavailable ({e}); hidden states go through the master")
            self._drain_replies(task_id)
            self._tile_rows, self._tile_workers = 0, []
            self.use_direct_tiles = False
            return False
        self._tile_rows, self._tile_workers = rows, ids
        return True

    def _pipeline_step(self, task_id: str, inputs: Dict[str, torch.Tensor], reset: bool, timeout: float) -> Dict[str, torch.Tensor]:
        """One pass through all stages.  The hidden states go from worker to worker over NVLink when the cluster allows
        it (`_ensure_tiles`); else each worker's RESULT (the hand-off) is forwarded to the next worker by the master."""
        x = inputs
        ids = inputs.get("input_ids")
        rows = int(ids.shape[0]) * int(ids.shape[1]) if ids is not None and ids.dim() == 2 else 0
        tiles = rows > 0 and self._ensure_tiles(task_id, rows)
        self.last_prefill_path = "tiles" if tiles else "relay"
        order = self._stage_order()
        for i, (worker_id, shard_ids) in enumerate(order):
            worker = self.workers[worker_id]
            md = {"task_id": task_id, "shard_ids": shard_ids, "reset": reset}
            if tiles:
                md["out_tile"] = i < len(order) - 1
                if i > 0:
                    md["in_tile"] = x.pop("tile_shape")
            self._send(worker, "RUN_INFERENCE", pickle.dumps(x), md)
            try:
                header, payload = self._replies[task_id].get(timeout=timeout)
            except Empty:
                raise TimeoutError(f"worker {worker_id} did not answer")
            if header.get("error"):
                raise RuntimeError(f"worker {worker_id}: {header['error']}")
            x = pickle.loads(payload)
        return x

    # ---- fused hand-off: one persistent decode loop per stage, no message per token ---------------------------------
    def _request(self, worker, task_id: str, command: str, payload=None, metadata=None, timeout: float = 60.0):
        """Send a command and wait for its RESULT.  Returns the unpickled payload."""
        md = dict(metadata or {})
        md["task_id"] = task_id
        self._send(worker, command, pickle.dumps(payload) if payload is not None else None, md)
        try:
            header, body = self._replies[task_id].get(timeout=timeout)
        except Empty:
            raise TimeoutError(f"no answer to {command}")
        if header.get("error"):
            raise RuntimeError(header["error"])
        return pickle.loads(body) if body else {}

    def _drain_replies(self, task_id: str) -> None:
        """Drop answers still queued for this task (a late RESULT must not be taken for the next request's)."""
        q = self._replies.get(task_id)
        while q is not None and not q.empty():
            try:
                q.get_nowait()
            except Empty:
                break

    def _fused_decode(self, task_id: str, first_tokens: torch.Tensor, n_steps: int, timeout: float = 300.0) -> Optional[torch.Tensor]:
        """Greedy decode of `n_steps` further tokens for the micro-batches whose prefill just ran (`first_tokens`:
        [n_micro, rows], sequence slots m * rows + r), with the hand-off fused into the stages' kernels (workers:
        MAILBOX_* / RUN_DECODE_LOOP).  Needs one shard per worker and every worker on its own GPU (two persistent
        kernels that wait for each other cannot share a device).  Returns int64 [n_micro, rows, n_steps] -- the tokens
        after `first_tokens` -- or None when this cluster / batch cannot take the fused path."""
        order = self._stage_order()
        if first_tokens.dim() == 1:
            first_tokens = first_tokens[None, :]
        n_micro, rows = int(first_tokens.shape[0]), int(first_tokens.shape[1])
        if n_steps < 1 or rows > 64 or any(len(s) != 1 for _, s in order):   # the stages' pre-flight decides about the shape itself
            return None
        places = [(self.workers[w].get("pid"), self.workers[w].get("device_index")) for w, _ in order]
        if any(p[0] is None or p[1] is None or p[1] < 0 for p in places) or len(set(places)) != len(places):
            return None
        workers = [self.workers[w] for w, _ in order]
        shard_ids = [s[0] for _, s in order]
        opened = []
        try:
            infos = []
            for st, (wk, sid) in enumerate(zip(workers, shard_ids)):
                infos.append(self._request(wk, task_id, "MAILBOX_OPEN", None, {"shard_id": sid, "n_slots": n_micro, "rows": rows, "stage": st}))
                opened.append(wk)
            for st, wk in enumerate(workers):
                self._request(wk, task_id, "MAILBOX_CONNECT", infos[(st + 1) % len(workers)], {"stage": st})
            self._request(workers[0], task_id, "MAILBOX_POST", {"tokens": first_tokens.to(torch.int32), "ticket": 1}, {"stage": 0})
            seq = torch.arange(n_micro * rows, dtype=torch.int32)
            loop = {"n_micro": n_micro, "rows": rows, "n_steps": n_steps, "step_base": 0}
            # pre-flight: every stage checks shape, mailboxes, context limit and reserves its KV pages.  Nothing is
            # launched unless all accept -- a refusal here raises RuntimeError and the caller falls back to the relay.
            positions = []
            for st, (wk, sid) in enumerate(zip(workers, shard_ids)):
                ans = self._request(wk, task_id, "PREPARE_DECODE_LOOP", {"seq_ids": seq, "positions": None},
                                    {"shard_id": sid, "stage": st, **loop})
                positions.append(ans["positions"])
            for st, (wk, sid) in enumerate(zip(workers, shard_ids)):  # every stage starts its loop; they finish together
                self._send(wk, "RUN_DECODE_LOOP", pickle.dumps({"seq_ids": seq, "positions": positions[st]}),
                           {"task_id": task_id, "shard_id": sid, "stage": st, **loop})
            fed, errors = None, []
            for _ in workers:  # wait for every stage, also when one of them failed: no reply may be left in the queue
                try:
                    header, body = self._replies[task_id].get(timeout=timeout)
                except Empty:
                    # a launched loop that never finishes: its kernels trap after 20 s without a ticket, which costs the
                    # worker its CUDA context -- that worker has to be restarted; do not fall back onto it
                    raise TimeoutError("a stage did not finish its decode loop (restart its worker)")
                if header.get("error"):
                    errors.append(header["error"])
                elif header.get("stage") == 0:
                    fed = pickle.loads(body)["fed"]
            if errors:
                raise RuntimeError("; ".join(errors))
            last = self._request(workers[0], task_id, "MAILBOX_READ", None, {"stage": 0})["tokens"]
            # fed[i] = ids fed at step i (fed[0] = first_tokens); the ids generated by the last step are in the mailbox
            gen = torch.cat([fed[1:], last.reshape(1, n_micro, rows)], dim=0)  # [n_steps, n_micro, rows]
            return gen.permute(1, 2, 0).to(torch.int64)
        finally:
            for wk in opened:
                try:
                    self._request(wk, task_id, "MAILBOX_CLOSE", None, {}, timeout=10.0)
                except Exception:
                    pass

    def _decode_micro_batches(self, task_id: str, prompts: List[torch.Tensor], max_new: int) -> Dict[str, Any]:
        """Several requests at once: prompt m ([rows, T_m] token ids, the same `rows` for all) becomes micro-batch m, its
        rows use the sequence slots m * rows ... -- the stages work on different micro-batches at the same time."""
        prompts = [p[None, :] if p.dim() == 1 else p for p in (torch.as_tensor(p, dtype=torch.int64) for p in prompts)]
        rows = int(prompts[0].shape[0])
        if any(int(p.shape[0]) != rows for p in prompts):
            raise ValueError("every micro-batch needs the same number of rows")
        M = len(prompts)
        seqs = [torch.arange(m * rows, (m + 1) * rows, dtype=torch.int32) for m in range(M)]
        first = []
        for m, ids in enumerate(prompts):  # prefill, one micro-batch after the other through the stages
            out = self._pipeline_step(task_id, {"input_ids": ids, "seq_ids": seqs[m]}, reset=(m == 0), timeout=60.0)
            first.append(out["next_token"].reshape(-1).to(torch.int64))
        tokens = torch.stack(first)[:, :, None]  # [M, rows, 1]
        fused = None
        if self.use_fused_handoff and max_new > 1:
            try:
                fused = self._fused_decode(task_id, torch.stack(first), max_new - 1)
            except RuntimeError as e:
                print(f"fused decode unavailable ({e}); falling back to the per-step loop")
                self._drain_replies(task_id)
        self.last_decode_path = "fused" if fused is not None else "per-step"
        if fused is not None:
            tokens = torch.cat([tokens, fused], dim=2)
        else:
            cur = first
            for _ in range(max_new - 1):
                nxt = []
                for m in range(M):
                    out = self._pipeline_step(task_id, {"input_ids": cur[m].reshape(-1, 1), "seq_ids": seqs[m]}, reset=False, timeout=60.0)
                    nxt.append(out["next_token"].reshape(-1).to(torch.int64))
                cur = nxt
                tokens = torch.cat([tokens, torch.stack(cur)[:, :, None]], dim=2)
        return {"token_ids": tokens.tolist(), "text": None}

    def run_many(self, prompts: Sequence[torch.Tensor], timeout: int = 120, max_new_tokens: int = 32):
        """Decode several prompts as micro-batches of one pipeline pass.  Returns {"token_ids": [n_prompts][rows][new]}."""
        return self.run_inference([torch.as_tensor(p) for p in prompts], timeout=timeout, max_new_tokens=max_new_tokens)

    def _process_tasks(self):
        while self.running:
            try:
                task_id, request, max_new = self.task_queue.get(block=True, timeout=0.2)
            except Empty:
                for w in self.evict_stale_workers():   # idle: workers whose heartbeat stopped leave the pipeline
                    print(f"Worker {w} evicted: no heartbeat for {self.heartbeat_timeout:.0f} s")
                continue
            done = self._finished.get(task_id)
            self._replies[task_id] = Queue()
            try:
                if not self.shard_assignments:
                    raise RuntimeError("no shard assignments: call assign_shards() / distribute_shards() first")
                if isinstance(request, (list, tuple)) and request and isinstance(request[0], torch.Tensor):
                    result = self._decode_micro_batches(task_id, list(request), max_new)  # prompts = micro-batches
                else:
                    if isinstance(request, str):
                        if self.tokenizer is None:
                            raise ValueError("no tokenizer is available: pass the prompt as token ids")
                        ids = self.tokenizer(request, return_tensors="pt")["input_ids"].to(torch.int64)
                    else:
                        ids = torch.as_tensor(request, dtype=torch.int64)
                        if ids.dim() == 1:
                            ids = ids[None, :]
                    out = self._pipeline_step(task_id, {"input_ids": ids}, reset=True, timeout=60.0)   # prefill
                    tokens = [self.sampler(out["logits"]) if "logits" in out else out["next_token"]]
                    fused = None
                    if self.use_fused_handoff and max_new > 1 and self.sampler.temperature <= 0.0:
                        try:
                            fused = self._fused_decode(task_id, tokens[0].reshape(1, -1), max_new - 1)
                            fused = fused[0] if fused is not None else None
                        except RuntimeError as e:  # e.g. a shape the persistent kernel does not take: fall back below
                            print(f"fused decode unavailable ({e}); falling back to the per-step loop")
                            self._drain_replies(task_id)
                    self.last_decode_path = "fused" if fused is not None else "per-step"
                    if fused is not None:
                        tokens.extend(fused[:, i] for i in range(fused.shape[1]))
                    for _ in range(0 if fused is not None else max_new - 1):                             # decode
                        out = self._pipeline_step(task_id, {"input_ids": tokens[-1].reshape(-1, 1).to(torch.int64)},
                                                  reset=False, timeout=60.0)
                        tokens.append(self.sampler(out["logits"]) if "logits" in out else out["next_token"])
                    token_ids = torch.stack([t.reshape(-1).cpu() for t in tokens], dim=1)
                    text = None
                    if self.tokenizer is not None:
                        text = self.tokenizer.batch_decode(token_ids.tolist())
                    result = {"token_ids": token_ids.tolist(), "text": text}
            except Exception as e:  # delivered to the caller instead of being printed and dropped (reference :240-242)
                result = e
            finally:
                self._replies.pop(task_id, None)
            if done is not None:
                done.put(result)
            self.result_queue.put((task_id, result))

    # ---- worker messages (reference :140-211) ------------------------------------------------------------------------
    def _accept_connections(self):
        while self.running:
            try:
                header, payload = self._owned.receive(timeout=0.2)
            except TimeoutError:
                continue
            except Exception as e:
                if self.running:
                    print(f"Error accepting connection: {e}")
                continue
            if not header:
                continue
            identity = header.get("identity", b"")
            worker_id = identity.hex() or "worker"
            with self._workers_lock:
                if worker_id not in self.workers:
                    print(f"New connection from {worker_id}")
                    self.workers[worker_id] = {"identity": identity, "connected_time": time.time(),
                                               "last_heartbeat": time.time(), "status": "connected"}
            self._handle_worker_command(worker_id, header.get("command", ""), header, payload)

    def _handle_worker_command(self, worker_id, command, header, payload):
        if command == "REGISTER":
            capabilities = header.get("capabilities", {})
            self.workers[worker_id].update(capabilities)
            print(f"Worker {worker_id} registered with capabilities: {capabilities}")
        elif command == "HEARTBEAT":
            self.workers[worker_id]["last_heartbeat"] = time.time()
        elif command == "SHARD_LOADED":
            shard_id = header.get("shard_id")
            print(f"Worker {worker_id} loaded shard {shard_id}")
            with self._loaded:
                if header.get("ok", True):
                    self._loaded_shards.add(shard_id)
                self._loaded.notify_all()
        elif command == "RESULT":
            task_id = header.get("task_id")
            q = self._replies.get(task_id)
            if q is not None:
                q.put((header, payload))

    def evict_stale_workers(self) -> List[str]:
        """Drop workers whose heartbeat is older than `heartbeat_timeout` (the reference records heartbeats and never
        looks at them, :199-201)."""
        now = time.time()
        with self._workers_lock:
            stale = [w for w, info in self.workers.items() if now - info["last_heartbeat"] > self.heartbeat_timeout]
            for w in stale:
                del self.workers[w]
                self.shard_assignments.pop(w, None)
        return stale
