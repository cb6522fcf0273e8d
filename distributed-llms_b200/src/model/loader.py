"""Reference src/model/loader.py:5-25 (weight acquisition; out of the hot path).  Same call surface;
bf16 instead of fp16 on CUDA, because the B200 path computes in bf16."""
import torch
from transformers import AutoModel, AutoTokenizer


def load_model(model_id, device_map="auto"):
    tokenizer = AutoTokenizer.from_pretrained(model_id)
    model = AutoModel.from_pretrained(
        model_id, device_map=device_map,
        torch_dtype=torch.bfloat16 if torch.cuda.is_available() else torch.float32)
    return model, tokenizer
