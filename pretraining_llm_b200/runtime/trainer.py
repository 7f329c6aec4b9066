"""Training runtime: ``Trainer`` with the reference's surface and schedule.

Mirrors reference ``scripts/train_transformer.py:35-109`` (``Trainer(model, train_iterator,
val_iterator, optimizer)``, ``_get_lr``, ``_evaluate``, ``_train_step``, ``train``) with the defects
of SURVEY 2.5 repaired: evaluation runs on every rank and is averaged (D7), gradients are
synchronised every step with a correct ``grad_accum_steps`` option (D8), checkpoints carry clean
keys (D11), the output directory is created, and a run can be resumed.

On a CUDA model the step is ``Engine.train_step`` (hand-written kernels, fused AdamW, CUDA graph);
on CPU it is plain PyTorch with ``torch.optim.AdamW`` - same numbers, same checkpoint format.
"""
from __future__ import annotations

import json
import os
import time
from typing import Iterator, Optional

import torch

from ..config import default_config
from ..parallel import bootstrap


def get_lr(it: int, cfg: dict) -> float:
    """Reference schedule (train_transformer.py:43-49): linear warm-up over the first 10% of
    ``t_train_steps`` (so step 0 has lr 0), then constant ``t_lr``.  With
    ``lr_schedule == 'decay'`` the otherwise unused reference keys ``t_lr_decay_step`` /
    ``t_lr_decayed`` are honoured: from that step on the rate is ``t_lr_decayed``."""
    warm = cfg["t_train_steps"] * 0.1
    if it < warm:
        return cfg["t_lr"] * it / warm
    if cfg.get("lr_schedule", "reference") == "decay" and it >= cfg["t_lr_decay_step"]:
        return cfg["t_lr_decayed"]
    return cfg["t_lr"]


class Trainer:
    def __init__(self, model, train_iterator: Iterator, val_iterator: Optional[Iterator], optimizer=None,
                 config: Optional[dict] = None, dist_ctx: Optional[bootstrap.DistContext] = None):
        self.config = dict(default_config) if config is None else config
        self.model = model
        self.train_iterator = train_iterator
        self.val_iterator = val_iterator
        self.dist = dist_ctx or bootstrap.DistContext()
        # The hand-written kernels compute in bf16 (fp32 statistics / masters).  dtype="float32" and "float16" are the
        # reference's other two branches (train_transformer.py:41,69,92-93): they run its recipe in plain PyTorch -
        # fp32 without autocast, fp16 under autocast with a GradScaler - instead of being silently computed in bf16.
        dtype = self.config.get("dtype", "bfloat16")
        if dtype != "bfloat16" and hasattr(model, "use_native"):
            model.use_native = False
        self.native = bool(getattr(model, "_native_ok", lambda: False)())
        self.optimizer = optimizer
        if not self.native and optimizer is None:
            self.optimizer = torch.optim.AdamW(model.parameters(), lr=self.config["t_lr"])
        on_cuda = next(model.parameters()).is_cuda
        self.scaler = torch.amp.GradScaler("cuda", enabled=(dtype == "float16" and on_cuda and not self.native))
        self.step = 0
        self.tokens_per_step = None
        self.log_file = open(self.config["log_jsonl"], "a") if self.config.get("log_jsonl") and self.dist.master else None
        if self.native and self.dist.world > 1:
            from ..parallel import make_data_parallel
            make_data_parallel(model.engine, self.dist, self.config.get("dp_mode", "ddp"),
                               bool(self.config.get("vocab_parallel", False)))

    # reference-compatible private names
    def _get_lr(self, it: int) -> float:
        return get_lr(it, self.config)

    def _autocast(self):
        dev = "cuda" if next(self.model.parameters()).is_cuda else "cpu"
        dtype = self.config["dtype"]
        if dtype == "float16" and dev == "cuda":
            return torch.autocast(device_type=dev, dtype=torch.float16)
        return torch.autocast(device_type=dev, dtype=torch.bfloat16, enabled=dtype == "bfloat16")

    @torch.no_grad()
    def _evaluate(self) -> float:
        """Mean validation loss over ``t_eval_iters`` batches, on all ranks (averaged)."""
        if self.val_iterator is None:
            return float("nan")
        self.model.eval()
        if self.native and self.model.engine.dp is not None:
            self.model.engine.dp.gather_weights()      # sharded / vocab-parallel weights -> complete local copy
        n = self.config["t_eval_iters"]
        total = torch.zeros((), device=next(self.model.parameters()).device)
        for _ in range(n):
            X, Y = next(self.val_iterator)
            if self.native:
                _, _, loss = self.model.engine.forward(X.contiguous(), Y.contiguous(), train=False)
            else:
                with self._autocast():
                    _, loss = self.model(X, Y)
            total += loss.float()
        self.model.train()
        mean = total / n
        mean = self.dist.all_reduce_mean(mean)
        return float(mean.item())

    def _train_step(self, xb, yb, lr: float):
        accum = self.config.get("grad_accum_steps", 1)
        if self.native:
            eng = self.model.engine
            if accum == 1:
                return eng.train_step(xb, yb, lr, use_graph=self.config.get("use_cuda_graph", True))
            loss_acc = None
            # the fused reduce+AdamW kernels of the last micro-step read the staged scalars while backward is still
            # running: stage them BEFORE the loop (same order as Engine.train_step)
            eng.set_hyper(lr)
            if eng.dp is not None:
                eng.dp.begin_step()
            for k in range(accum):
                if k:
                    xb, yb = next(self.train_iterator)
                w, _, loss = eng.forward(xb.contiguous(), yb.contiguous(), train=True)
                eng.backward(w, grad_scale=1.0 / accum, accumulate=k > 0, dp_sync=(k == accum - 1))
                loss_acc = loss.clone() if loss_acc is None else loss_acc + loss
            eng.optimizer_step()
            return loss_acc / accum
        for g in self.optimizer.param_groups:
            g["lr"] = lr
        loss_acc = None
        for k in range(accum):
            if k:
                xb, yb = next(self.train_iterator)
            with self._autocast():
                _, loss = self.model(xb, yb)
            self.scaler.scale(loss / accum).backward()
            loss_acc = loss.detach() if loss_acc is None else loss_acc + loss.detach()
        self.dist.all_reduce_grads(self.model)
        self.scaler.step(self.optimizer)          # plain optimizer.step() unless fp16 scaling is on
        self.scaler.update()
        self.optimizer.zero_grad(set_to_none=True)
        return loss_acc / accum

    # ------------------------------------------------------------------ checkpoints
    def state_for_checkpoint(self) -> dict:
        """{'model_state_dict', 'optimizer_state_dict'} in the reference's layout
        (train_transformer.py:106-109) plus 'trainer_state' for resuming."""
        model = self.model
        if self.native and model.engine.dp is not None:
            # each rank holds current fp32 masters only for the slices it owns: complete them (and the moments)
            # BEFORE the state_dict is cloned out of the flat buffer
            model.engine.dp.check_errors()
            model.engine.dp.gather_full_state()
        sd = model.state_dict()
        lr = self._get_lr(max(self.step - 1, 0))
        if self.native:
            eng = model.engine
            eng.ensure_opt_state()
            osd = model.layout.to_reference_optimizer_state(eng.exp_avg, eng.exp_avg_sq, eng.step_count, lr,
                                                            eng.betas, eng.eps, eng.weight_decay)
            if eng.dp is not None:
                eng.release_full_opt_state()      # the sharded dp.m / dp.v are the live copies
                if hasattr(eng.dp, "release_full"):
                    sd = {k: v.cpu() for k, v in sd.items()}
                    eng.dp.release_full()         # sharded mode: drop the materialised full parameters again
        else:
            osd = self._reference_optimizer_state()
        return {"model_state_dict": sd, "optimizer_state_dict": osd,
                "trainer_state": {"step": self.step, "config": {k: v for k, v in self.config.items()
                                                                 if isinstance(v, (int, float, str, bool))}}}

    def _reference_optimizer_state(self) -> dict:
        """torch optimizer over packed params -> reference per-head layout."""
        lay = self.model.layout
        flat_m = torch.zeros_like(self.model.flat_params)
        flat_v = torch.zeros_like(self.model.flat_params)
        step = 0
        for name, p in self.model.named_parameters():
            st = self.optimizer.state.get(p)
            if st:
                lay.view(flat_m, name).copy_(st["exp_avg"])
                lay.view(flat_v, name).copy_(st["exp_avg_sq"])
                step = int(float(st["step"]))
        g = self.optimizer.param_groups[0]
        return lay.to_reference_optimizer_state(flat_m, flat_v, step, g["lr"], g["betas"], g["eps"], g["weight_decay"])

    def save_checkpoint(self, path: str) -> None:
        state = self.state_for_checkpoint()      # collective when sharded: every rank calls it
        if self.dist.master:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            tmp = path + ".tmp"
            torch.save(state, tmp)
            os.replace(tmp, path)

    def load_checkpoint(self, path: str) -> None:
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        sharded = self.native and self.model.engine.dp is not None and hasattr(self.model.engine.dp, "materialize")
        if sharded:
            self.model.engine.dp.materialize()     # collective: full buffers to load into, re-sharded below
        self.model.load_state_dict(ckpt["model_state_dict"])
        osd = ckpt.get("optimizer_state_dict")
        lay = self.model.layout
        if osd and osd.get("state"):
            if self.native:
                eng = self.model.engine
                eng.ensure_opt_state()
                m, v = torch.zeros(lay.total), torch.zeros(lay.total)
                step, _ = lay.from_reference_optimizer_state(osd, m, v)
                eng.exp_avg.copy_(m)
                eng.exp_avg_sq.copy_(v)
                eng.step_count = step
                if eng.dp is not None:
                    eng.dp.state_loaded()
            else:
                m, v = torch.zeros(lay.total), torch.zeros(lay.total)
                step, _ = lay.from_reference_optimizer_state(osd, m, v)
                for name, p in self.model.named_parameters():
                    self.optimizer.state[p] = {"step": torch.tensor(float(step)),
                                               "exp_avg": lay.view(m, name).clone().to(p.device),
                                               "exp_avg_sq": lay.view(v, name).clone().to(p.device)}
        if sharded and self.model.engine.dp.full_flat is not None:
            self.model.engine.dp.state_loaded()    # no optimizer state in the file: re-shard the weights alone
        if self.native and self.dist.world > 1:
            # peers read each other's memory (weight shards, embedding rows): nobody starts a step before every rank
            # has finished loading
            torch.cuda.synchronize()
            self.dist.barrier()
        ts = ckpt.get("trainer_state")
        if ts:
            self.step = int(ts["step"])

    def _mfu(self, tokens_per_s: float, T: int) -> float:
        """Model-FLOP utilisation of the whole job against the measured bf16 tensor peak (utils/flops.py)."""
        from ..utils.flops import mfu
        m = self.model
        return mfu(tokens_per_s, m.N_BLOCKS, m.n_embed, m.vocab_size, T, self.dist.world)

    # ------------------------------------------------------------------ main loop
    def train(self) -> None:
        cfg = self.config
        self.model.train()
        if cfg.get("resume_from"):
            self.load_checkpoint(cfg["resume_from"])
        cuda = next(self.model.parameters()).is_cuda
        t0 = time.time()
        ev0 = ev1 = None
        if cuda:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        steps_in_window = 0
        while self.step < cfg["t_train_steps"]:
            step = self.step
            lr = self._get_lr(step)
            xb, yb = next(self.train_iterator)
            if self.tokens_per_step is None:
                self.tokens_per_step = xb.numel() * cfg.get("grad_accum_steps", 1) * self.dist.world
            loss = self._train_step(xb, yb, lr)
            steps_in_window += 1
            self.step += 1
            if step % cfg["t_eval_steps"] == 0:
                if cuda:
                    ev1.record()
                    ev1.synchronize()
                    dt_dev = ev0.elapsed_time(ev1) / 1e3
                    dt_dev = float(self.dist.all_reduce_max(torch.tensor(dt_dev, device=loss.device)).item())
                else:
                    dt_dev = time.time() - t0
                t1 = time.time()
                if self.native and self.model.engine.dp is not None:
                    self.model.engine.dp.check_errors()     # a peer time-out must stop the run, not train on garbage
                val_loss = self._evaluate()
                train_loss = float(self.dist.all_reduce_mean(loss.detach().float()).item())
                toks = self.tokens_per_step * steps_in_window / max(dt_dev, 1e-9)
                util = self._mfu(toks, xb.shape[-1]) if cuda else float("nan")
                if self.dist.master:
                    # first five fields: the reference's line (train_transformer.py:101)
                    print(f"Step {step}: Train Loss={train_loss:.4f}, Val Loss={val_loss:.4f}, LR={lr:.2e}, "
                          f"Time={(t1 - t0) * 1000:.2f}ms, tokens/s={toks:,.0f}"
                          + (f", MFU={util * 100:.1f}%" if cuda else ""), flush=True)
                    if self.log_file:
                        self.log_file.write(json.dumps({"step": step, "train_loss": train_loss, "val_loss": val_loss,
                                                        "lr": lr, "tokens_per_s": toks, "device_s": dt_dev,
                                                        "mfu": util}) + "\n")
                        self.log_file.flush()
                t0 = time.time()
                steps_in_window = 0
                if cuda:
                    ev0.record()
            if cfg.get("save_every", 0) and self.step % cfg["save_every"] == 0:
                self.save_checkpoint(cfg["t_out_path"])
        if self.dist.master:
            print(f"Training finished. Saving model to {cfg['t_out_path']}")
        self.save_checkpoint(cfg["t_out_path"])
