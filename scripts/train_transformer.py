"""Training entrypoint - same launch contract as reference ``scripts/train_transformer.py``:

    python scripts/train_transformer.py                      # single GPU (or CPU if device=cpu)
    torchrun --nproc_per_node=8 scripts/train_transformer.py # one process per GPU

driven by ``config/config.py`` and the ``RANK/LOCAL_RANK/WORLD_SIZE`` environment (reference
:15-20).  Additions: ``key=value`` overrides on the command line (``preset=gpt2-small
synthetic_data=true t_train_steps=100``) so benchmark shapes do not need file edits.  There is no
``TORCH_COMPILE`` switch: the CUDA path is hand-written kernels, not a tracing compiler.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from config.config import apply_overrides, default_config as config, model_args, validate  # noqa: E402
from data_loader.data_loader import get_batch_iterator  # noqa: E402
from pretraining_llm_b200.parallel.bootstrap import init_distributed  # noqa: E402
from pretraining_llm_b200.runtime.trainer import Trainer  # noqa: E402
from src.models.transformer import Transformer  # noqa: E402


def main(argv=None):
    apply_overrides(config, argv if argv is not None else sys.argv[1:])
    validate(config)
    ctx = init_distributed(config)
    device = ctx.device
    torch.manual_seed(config["seed"])            # same seed on every rank => identical initial weights
    model = Transformer(**model_args(config)).to(device)
    if ctx.master:
        print(f"model: {model.num_parameters():,} parameters on {device}, world size {ctx.world}")
    data_train = None if config["synthetic_data"] else config["train_path"]
    data_val = None if config["synthetic_data"] else config["val_path"]
    kw = dict(seed=config["seed"], vocab_size=config["vocab_size"])
    train_iterator = get_batch_iterator(data_train, config["t_batch_size"], config["context_length"], device,
                                        ctx.ddp, ctx.rank, ctx.world, **kw)
    val_iterator = get_batch_iterator(data_val, config["t_batch_size"], config["context_length"], device,
                                      ctx.ddp, ctx.rank, ctx.world, seed=config["seed"] + 1,
                                      vocab_size=config["vocab_size"])
    trainer = Trainer(model, train_iterator, val_iterator, None, config=config, dist_ctx=ctx)
    trainer.train()
    ctx.shutdown()


if __name__ == "__main__":
    main()
