

def make_data_parallel(eng, ctx, mode: str = "ddp", vocab_parallel: bool = False):
    """Attach the multi-GPU machinery to an Engine: ``ddp`` = replicated weights, fused all-reduce + AdamW
    (parallel/dp.py); ``zero`` = parameters / gradients / optimizer state sharded, all-gather fused into the GEMMs
    (parallel/zero.py)."""
    if mode == "zero":
        if vocab_parallel:
            raise ValueError("vocab_parallel goes with dp_mode='ddp' (the sharded mode shards the LM head already)")
        from .zero import ZeroDataParallel
        dp = ZeroDataParallel(eng, ctx)
    elif mode == "ddp":
        from .dp import DataParallel
        dp = DataParallel(eng, ctx, mode="ddp", vocab_parallel=vocab_parallel)
    else:
        raise ValueError(f"dp_mode must be 'ddp' or 'zero', got {mode!r}")
    eng.dp = dp
    return dp
