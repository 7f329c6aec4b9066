"""Generation entrypoint - same flags and importable function as reference
``scripts/generate_text.py:7-60`` (``--model_path --input_text --max_new_tokens``;
``generate_text(model_path, input_text, max_new_tokens=100, device='cuda') -> str``).

Additions: ``--input_ids "1,2,3"`` / ``generate_ids`` for machines without the tokenizer files
(tiktoken downloads its BPE table on first use - SURVEY D14), ``--device``, ``--temperature``,
``--top_k``, ``--seed``.  Checkpoints written by the reference (including ``_orig_mod.`` /
``module.`` prefixed keys, defect D11) load unchanged.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from config.config import default_config as config  # noqa: E402
from src.models.transformer import Transformer  # noqa: E402


def _load_model(model_path: str, device: str) -> Transformer:
    checkpoint = torch.load(model_path, map_location="cpu", weights_only=False)
    # The reference builds the model from config/config.py (generate_text.py:12-20), so a checkpoint of another shape
    # needs the config edited first.  Checkpoints written by this framework's Trainer also carry the shape they were
    # trained with ("trainer_state" -> "config"): prefer it, fall back to the config file for reference checkpoints.
    shape = dict(config)
    saved = (checkpoint.get("trainer_state") or {}).get("config") or {}
    for k in ("n_head", "n_embed", "context_length", "vocab_size", "n_blocks"):
        if k in saved:
            shape[k] = saved[k]
    model = Transformer(n_head=shape["n_head"], n_embed=shape["n_embed"],
                        context_length=shape["context_length"], vocab_size=shape["vocab_size"],
                        N_BLOCKS=shape["n_blocks"])
    model.load_state_dict(checkpoint["model_state_dict"])
    return model.eval().to(device)


def _encoder():
    try:
        import tiktoken
        return tiktoken.get_encoding("r50k_base")
    except Exception as e:  # no package or no network for the BPE table
        raise RuntimeError("the r50k_base tokenizer is not available offline; use --input_ids / generate_ids "
                           f"instead ({type(e).__name__}: {e})")


def generate_ids(model_path: str, input_ids, max_new_tokens: int = 100, device: str = "cuda", *,
                 temperature: float = 1.0, top_k=None, seed=None):
    """Token ids in, token ids out (offline mode).  ``input_ids``: list[int] or list[list[int]]."""
    if device.startswith("cuda") and not torch.cuda.is_available():
        device = "cpu"
    model = _load_model(model_path, device)
    ids = torch.tensor(input_ids, dtype=torch.long, device=device)
    if ids.dim() == 1:
        ids = ids.unsqueeze(0)
    gen = None
    if seed is not None:
        gen = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        out = model.generate(ids, max_new_tokens=max_new_tokens, temperature=temperature, top_k=top_k,
                             generator=gen)
    return out.tolist()


def generate_text(model_path: str, input_text: str, max_new_tokens: int = 100, device: str = "cuda") -> str:
    enc = _encoder()
    start_ids = enc.encode_ordinary(input_text)
    tokens = generate_ids(model_path, start_ids, max_new_tokens, device)[0]
    return enc.decode(tokens)


def main() -> None:
    parser = argparse.ArgumentParser(description="Generate text using a pre-trained Transformer model.")
    parser.add_argument("--model_path", type=str, help="Path to the saved model checkpoint.")
    parser.add_argument("--input_text", type=str, help="The initial text to start generation from.")
    parser.add_argument("--max_new_tokens", type=int, default=100, help="Maximum number of new tokens to generate.")
    parser.add_argument("--input_ids", type=str, default=None, help="Comma-separated token ids (offline mode).")
    parser.add_argument("--device", type=str, default="cuda")
    parser.add_argument("--temperature", type=float, default=1.0)
    parser.add_argument("--top_k", type=int, default=None)
    parser.add_argument("--seed", type=int, default=None)
    args = parser.parse_args()
    if args.input_ids is not None:
        ids = [int(t) for t in args.input_ids.split(",") if t.strip()]
        out = generate_ids(args.model_path, ids, args.max_new_tokens, args.device, temperature=args.temperature,
                           top_k=args.top_k, seed=args.seed)
        print("Generated ids:\n" + ",".join(map(str, out[0])))
        return
    generated = generate_text(args.model_path, args.input_text, args.max_new_tokens, args.device)
    print(f"Generated text:\n{generated}")


if __name__ == "__main__":
    main()
