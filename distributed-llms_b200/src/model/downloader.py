"""Reference src/model/downloader.py:4-6, unchanged in meaning (needs a network; out of the hot path)."""
from huggingface_hub import snapshot_download


def download_model(model_id, cache_dir="./models"):
    return snapshot_download(repo_id=model_id, cache_dir=cache_dir)
