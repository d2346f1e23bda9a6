#!/usr/bin/env python
"""Write a synthetic HuggingFace model directory (config.json + a WordLevel tokenizer) for a
Qwen3 size, so the reference's `LLM(path)` can be constructed offline.
No weights are written: the e2e harness (tools/e2e/sitecustomize.py) seeds them in place."""
import json
import os
import sys

SIZES = {  # public HF configs (SURVEY.md section 8 header)
    "0.6b": dict(hidden_size=1024, num_hidden_layers=28, num_attention_heads=16, num_key_value_heads=8, intermediate_size=3072,
                 tie_word_embeddings=True),
    "8b": dict(hidden_size=4096, num_hidden_layers=36, num_attention_heads=32, num_key_value_heads=8, intermediate_size=12288,
               tie_word_embeddings=False),
    "14b": dict(hidden_size=5120, num_hidden_layers=40, num_attention_heads=40, num_key_value_heads=8, intermediate_size=17408,
                tie_word_embeddings=False),
    "32b": dict(hidden_size=5120, num_hidden_layers=64, num_attention_heads=64, num_key_value_heads=8, intermediate_size=25600,
                tie_word_embeddings=False),
}


def make(size: str, out_dir: str, num_layers: int | None = None) -> str:
    os.makedirs(out_dir, exist_ok=True)
    cfg = dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", attention_bias=False, attention_dropout=0.0,
               bos_token_id=151643, eos_token_id=151645, head_dim=128, hidden_act="silu", initializer_range=0.02,
               max_position_embeddings=40960, max_window_layers=28, rms_norm_eps=1e-6, rope_scaling=None,
               rope_theta=1000000, sliding_window=None, torch_dtype="bfloat16", use_cache=True,
               use_sliding_window=False, vocab_size=151936)
    cfg.update(SIZES[size])
    if num_layers:
        cfg["num_hidden_layers"] = num_layers
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(cfg, f, indent=1)
    from tokenizers import Tokenizer, models, pre_tokenizers
    vocab = {f"t{i}": i for i in range(1000)}
    vocab["<unk>"] = 1000
    tok = Tokenizer(models.WordLevel(vocab=vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.Whitespace()
    tok.save(os.path.join(out_dir, "tokenizer.json"))
    with open(os.path.join(out_dir, "tokenizer_config.json"), "w") as f:
        json.dump({"tokenizer_class": "PreTrainedTokenizerFast", "unk_token": "<unk>", "eos_token": "t999",
                   "model_max_length": 40960}, f)
    return out_dir


if __name__ == "__main__":
    print(make(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else None))
