#!/usr/bin/env python
"""Entry script with the reference's commands (reference run_master.py:26-48: assign / distribute / inference / exit).
The model defaults to a seeded random-init GPT-2 small (`random:gpt2-small`): there is no network here."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from src.master.node import MasterNode  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=65432)
    ap.add_argument("--model", default="random:gpt2-small", help="local dir, random:<name> or a Hub id")
    ap.add_argument("--shards", type=int, default=2)
    ap.add_argument("--cache-dir", default="./models")
    ap.add_argument("--by-path", action="store_true",
                    help="workers read their stage from the model directory (same box / shared file system) instead of receiving it")
    a = ap.parse_args()
    master = MasterNode(host=a.host, port=a.port)
    master.start()
    print("Initializing and sharding model...")
    print(f"Model initialized and sharded at: {master.initialize_model(a.model, num_shards=a.shards, cache_dir=a.cache_dir, by_path=a.by_path)}")
    print("\nWaiting for worker connections...\nStart workers in separate terminals with: python run_worker.py")
    try:
        while True:
            cmd = input("\nEnter command (assign/distribute/inference/exit): ").strip()
            if cmd == "assign":
                print(f"Shard assignments: {master.assign_shards()}")
            elif cmd == "distribute":
                master.distribute_shards(by_path=a.by_path)
                print("Shards distributed to workers")
            elif cmd == "inference":
                text = input("Enter text (or space-separated token ids) for inference: ")
                req = [int(t) for t in text.split()] if all(t.lstrip("-").isdigit() for t in text.split()) and text.strip() else text
                print(f"Inference result: {master.run_inference(req)}")
            elif cmd == "exit":
                break
            else:
                print("Unknown command")
    except (KeyboardInterrupt, EOFError):
        pass
    finally:
        master.stop()
        print("Master node stopped")


if __name__ == "__main__":
    main()
