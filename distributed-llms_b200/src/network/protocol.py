"""Control-plane wire protocol, same names and framings as the reference (src/network/protocol.py:8-176):
raw TCP  = "<header_len:<10>" + pickle(header) + payload;   ZeroMQ = multipart [identity?, pickle(header), payload].

Only the control plane (REGISTER / LOAD_SHARD / HEARTBEAT / small RESULTs) uses this.  The data plane --
the hidden-state tile between stages -- goes over NVLink (src/network/handoff.py), not through pickle.

Repairs behind the reference's names:
  * ROUTER sockets `bind()` (the reference only binds PUB/PUSH/REP, so its master never listened);
  * `receive_message` always returns `(header, payload)`; a ROUTER's peer identity is `header["identity"]`;
  * `send_message` / `receive_message` work on an instance *and* on the class, which is how the
    reference's master and its tests call them (src/master/node.py:174,220; tests/network/test_protocol.py:12);
  * the TCP payload is received into one pre-sized buffer instead of 4 KiB `bytes +=` chunks.
"""
from __future__ import annotations

import pickle
import socket
import threading
from typing import Any, Dict, Optional, Tuple

try:
    import zmq
except Exception:  # pragma: no cover
    zmq = None

HEADER_SIZE = 10  # width of the header-length field (reference :8)


class _hybridmethod:
    """Method usable as `obj.m(...)` and as `Class.m(...)` (then `self` is None)."""

    def __init__(self, fn):
        self.fn = fn
        self.__doc__ = fn.__doc__

    def __get__(self, obj, cls):
        fn = self.fn

        def bound(*a, **k):
            return fn(obj, *a, **k)
        bound.__doc__ = fn.__doc__
        bound.__name__ = fn.__name__
        return bound


class MessageProtocol:
    MESSAGE_TYPES = {
        "REGISTER": "Worker registration",
        "LOAD_SHARD": "Transfer model shard",
        "UNLOAD_SHARD": "Drop a model shard",
        "RUN_INFERENCE": "Execute computation",
        "SCHEDULE_COMPUTATION": "Execute computation with a priority",
        "RESULT": "Return computation results",
        "SHARD_LOADED": "Shard load confirmation",
        "SHARD_UNLOADED": "Shard unload confirmation",
        "HEARTBEAT": "Health check",
        "SHARD_REQUEST": "Request specific shard",
        "TASK_ASSIGN": "Assign computation task",
    }

    def __init__(self, zmq_context=None):
        self.zmq_context = zmq_context
        self.zmq_socket = None
        self.use_zmq = False

    def setup_zmq_socket(self, socket_type, address=None):
        if zmq is None:
            raise RuntimeError("pyzmq is not installed")
        if self.zmq_context is None:
            self.zmq_context = zmq.Context.instance()
        sock = self.zmq_context.socket(socket_type)
        sock.setsockopt(zmq.LINGER, 0)
        if address:
            if socket_type in (zmq.PUB, zmq.PUSH, zmq.REP, zmq.ROUTER):
                sock.bind(address)
            else:
                sock.connect(address)
        self.zmq_socket = sock
        self.use_zmq = True
        return sock

    # ---- send ---------------------------------------------------------------------------------------
    @_hybridmethod
    def send_message(self, sock, command: str, payload: Optional[bytes] = None,
                     metadata: Optional[Dict[str, Any]] = None) -> bool:
        if zmq is not None and isinstance(sock, zmq.Socket):
            return MessageProtocol._send_zmq(sock, command, payload, metadata)
        try:
            header = {"command": command}
            if metadata:
                header.update(metadata)
            if payload is not None:
                header["payload_size"] = len(payload)
            header_bytes = pickle.dumps(header)
            sock.sendall(f"{len(header_bytes):<{HEADER_SIZE}}".encode() + header_bytes)
            if payload is not None:
                sock.sendall(payload)
            return True
        except socket.timeout:
            print("Timeout while sending message")
            return False
        except Exception as e:
            print(f"Error sending message: {e}")
            return False

    # ---- receive ------------------------------------------------------------------------------------
    @_hybridmethod
    def receive_message(self, sock, timeout: int = 60) -> Tuple[Dict[str, Any], Optional[bytes]]:
        if zmq is not None and isinstance(sock, zmq.Socket):
            return MessageProtocol._receive_zmq(sock, timeout)
        try:
            sock.settimeout(timeout)
            raw = MessageProtocol._recv_exact(sock, HEADER_SIZE, allow_eof=True)
            if raw is None:
                raise ConnectionError("Connection closed")
            header_len = int(raw.decode().strip())  # ValueError on a malformed length field
            header_bytes = MessageProtocol._recv_exact(sock, header_len)
            header = pickle.loads(header_bytes)
            payload = None
            if "payload_size" in header:
                payload = MessageProtocol._recv_exact(sock, header["payload_size"])
            return header, payload
        except socket.timeout:
            raise TimeoutError("Timeout while receiving message")

    @staticmethod
    def _recv_exact(sock, n: int, allow_eof: bool = False) -> Optional[bytes]:
        buf = bytearray()
        while len(buf) < n:
            chunk = sock.recv(n - len(buf))
            if not chunk:
                if allow_eof and not buf:
                    return None
                raise ConnectionError("Connection closed while receiving")
            buf += chunk
        return bytes(buf)

    # ---- ZeroMQ framing -------------------------------------------------------------------------------
    @staticmethod
    def _zmq_frames(sock_type, command, payload, metadata) -> list:
        header = {"command": command}
        identity = None
        if metadata:
            header.update(metadata)
            identity = header.pop("identity", None)
        if payload is not None:
            header["payload_size"] = len(payload)
        frames = [pickle.dumps(header), payload if payload is not None else b""]
        if sock_type == zmq.ROUTER:
            if identity is None:
                raise ValueError("ROUTER send needs metadata['identity']")
            frames.insert(0, identity)
        return frames

    @staticmethod
    def _send_zmq(sock, command, payload, metadata) -> bool:
        try:
            sock.send_multipart(MessageProtocol._zmq_frames(sock.type, command, payload, metadata))
            return True
        except Exception as e:
            print(f"Error sending ZMQ message: {e}")
            return False

    @staticmethod
    def _receive_zmq(sock, timeout) -> Tuple[Dict[str, Any], Optional[bytes]]:
        if not sock.poll(int(timeout * 1000)):
            raise TimeoutError("Timeout while receiving ZMQ message")
        return MessageProtocol._parse_zmq(sock.type, sock.recv_multipart())

    @staticmethod
    def _parse_zmq(sock_type, parts) -> Tuple[Dict[str, Any], Optional[bytes]]:
        identity = None
        if sock_type == zmq.ROUTER:
            identity, parts = parts[0], parts[1:]
        if not parts:
            return {}, None
        header = pickle.loads(parts[0])
        if identity is not None:
            header["identity"] = identity
        payload = parts[1] if len(parts) > 1 and "payload_size" in header else None
        return header, payload

    # the reference's private names, kept as aliases
    def _send_zmq_message(self, command, payload=None, metadata=None) -> bool:
        return MessageProtocol._send_zmq(self.zmq_socket, command, payload, metadata)

    def _receive_zmq_message(self, timeout: int = 60):
        return MessageProtocol._receive_zmq(self.zmq_socket, timeout)


class OwnedSocket:
    """A ZeroMQ socket with ONE owning thread.  libzmq sockets are not thread-safe, and both nodes have several threads
    that want to talk on the same socket (the reference shares its sockets between threads unguarded: heartbeat,
    handler and task threads, src/worker/node.py:124-126,263-276, src/master/node.py:39-46).  Here the thread that
    receives owns the socket; every other thread hands its frames to it through a thread-local inproc PUSH socket, and
    the owner forwards them between polls -- it blocks on both at once, so a send is on the wire as soon as the owner
    wakes up, and nothing is ever sent and received on the socket by two threads at a time."""

    def __init__(self, context, sock):
        self.context, self.sock = context, sock
        self.endpoint = f"inproc://dllm-outbox-{id(self):x}"
        self._pull = context.socket(zmq.PULL)
        self._pull.setsockopt(zmq.LINGER, 0)
        self._pull.bind(self.endpoint)
        self._poller = zmq.Poller()
        self._poller.register(self.sock, zmq.POLLIN)
        self._poller.register(self._pull, zmq.POLLIN)
        self._local = threading.local()
        self._closed = False

    def send(self, command: str, payload: Optional[bytes] = None, metadata: Optional[Dict[str, Any]] = None) -> bool:
        """From any thread."""
        if self._closed:
            return False
        try:
            frames = MessageProtocol._zmq_frames(self.sock.type, command, payload, metadata)
            push = getattr(self._local, "push", None)
            if push is None:
                push = self.context.socket(zmq.PUSH)
                push.setsockopt(zmq.LINGER, 0)
                push.connect(self.endpoint)
                self._local.push = push
            push.send_multipart(frames)
            return True
        except Exception as e:
            print(f"Error sending ZMQ message: {e}")
            return False

    def _forward_outgoing(self) -> None:
        while True:
            try:
                frames = self._pull.recv_multipart(zmq.NOBLOCK)
            except zmq.Again:
                return
            self.sock.send_multipart(frames)

    def receive(self, timeout: float = 60) -> Tuple[Dict[str, Any], Optional[bytes]]:
        """Owning thread only.  Forwards whatever other threads queued, returns the next incoming message or raises
        TimeoutError."""
        import time
        deadline = time.monotonic() + timeout
        while True:
            self._forward_outgoing()
            left = deadline - time.monotonic()
            if left <= 0:
                raise TimeoutError("Timeout while receiving ZMQ message")
            events = dict(self._poller.poll(int(left * 1000) + 1))
            if self._pull in events:
                self._forward_outgoing()
            if self.sock in events:
                return MessageProtocol._parse_zmq(self.sock.type, self.sock.recv_multipart())

    def close(self) -> None:
        """Owning thread (or after it has stopped)."""
        self._closed = True
        try:
            self._forward_outgoing()
        except Exception:
            pass
        self._pull.close()
        self.sock.close()
