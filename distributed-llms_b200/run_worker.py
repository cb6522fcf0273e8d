#!/usr/bin/env python
"""Entry script of a worker (reference run_worker.py:12-23): connects to the master and serves its shards."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from src.worker.node import WorkerNode  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--master", default="localhost:65432")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=65433)
    ap.add_argument("--gpu", type=int, default=None, help="CUDA device of this worker (one worker per GPU)")
    a = ap.parse_args()
    # the device is a property of the worker, not of the main thread: commands are handled on other threads
    worker = WorkerNode(host=a.host, port=a.port, master_address=a.master, device=a.gpu)
    print(f"Worker node initialized and connecting to master at {a.master}")
    try:
        worker.start(block=True)
    except KeyboardInterrupt:
        pass
    finally:
        worker.stop()
        print("Worker node stopped")


if __name__ == "__main__":
    main()
