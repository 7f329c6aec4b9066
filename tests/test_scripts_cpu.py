"""CPU tier: the four entrypoints of the reference (scripts/*.py) run end to end on a tiny local setup."""
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV_TINY = dict(B200_CFG_VOCAB_SIZE="300", B200_CFG_CONTEXT_LENGTH="32", B200_CFG_N_EMBED="64", B200_CFG_N_HEAD="4",
                B200_CFG_N_BLOCKS="2", B200_CFG_T_BATCH_SIZE="4", B200_CFG_DEVICE="cpu", B200_CFG_DTYPE="float32")


def _run(args, env_extra, cwd=ROOT, timeout=600):
    env = dict(os.environ, **ENV_TINY, **env_extra)
    env.pop("RANK", None)
    return subprocess.run([sys.executable] + args, capture_output=True, text=True, env=env, cwd=cwd, timeout=timeout)


def test_preprocess_train_generate_pipeline(tmp_path):
    """text files -> data_preprocess.py (byte tokenizer) -> train_transformer.py on the .bin files ->
    generate_text.py --input_ids on the checkpoint.  The reference's scripts die on missing config keys
    before reaching any of these stages."""
    txt = tmp_path / "corpus.txt"
    txt.write_text("\n\n".join("document %d says hello to the world %s" % (i, "abc " * (5 + i % 7)) for i in range(400)))
    train_bin, val_bin, ckpt = tmp_path / "train.bin", tmp_path / "val.bin", tmp_path / "out" / "model.pt"
    paths = dict(B200_CFG_TRAIN_PATH=str(train_bin), B200_CFG_VAL_PATH=str(val_bin), B200_CFG_T_OUT_PATH=str(ckpt))
    r = _run(["scripts/data_preprocess.py", "--text_files", str(txt), "--tokenizer", "bytes", "--val_fraction", "0.05"], paths)
    assert r.returncode == 0, r.stderr[-2000:]
    tr = np.memmap(train_bin, dtype=np.uint16, mode="r")
    va = np.memmap(val_bin, dtype=np.uint16, mode="r")
    assert len(tr) > 5000 and len(va) > 100 and int(tr.max()) <= 256          # bytes + EOT(256)
    assert (np.asarray(tr) == 256).sum() >= 300                              # one EOT per document

    r = _run(["scripts/train_transformer.py", "t_train_steps=6", "t_eval_steps=3", "t_eval_iters=2", "t_lr=0.01"], paths)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-2500:]
    assert "Step 0: Train Loss=" in r.stdout and "Step 3: Train Loss=" in r.stdout
    assert ckpt.exists()
    ck = torch.load(ckpt, weights_only=False)
    assert {"model_state_dict", "optimizer_state_dict"} <= set(ck)

    r = _run(["scripts/generate_text.py", "--model_path", str(ckpt), "--input_ids", "72,101,108,108,111",
              "--max_new_tokens", "7", "--device", "cpu", "--seed", "3"], paths)
    assert r.returncode == 0, r.stderr[-2500:]
    ids = [int(t) for t in r.stdout.strip().splitlines()[-1].split(",")]
    assert len(ids) == 12 and ids[:5] == [72, 101, 108, 108, 111] and max(ids) < 300


def test_generate_text_without_tokenizer_reports_cleanly(tmp_path):
    from pretraining_llm_b200.models import Transformer
    torch.manual_seed(0)
    m = Transformer(4, 64, 32, 300, N_BLOCKS=2)
    p = tmp_path / "m.pt"
    torch.save({"model_state_dict": m.state_dict()}, p)
    r = _run(["scripts/generate_text.py", "--model_path", str(p), "--input_text", "hello", "--max_new_tokens", "3",
              "--device", "cpu"], {})
    # offline box: tiktoken (or its BPE table) is not available -> a clear message, not a stack of network errors
    assert r.returncode == 0 or "tokenizer is not available offline" in (r.stderr + r.stdout)


def test_data_download_offline_message():
    r = _run(["scripts/data_download.py"], {})
    assert r.returncode == 0
    assert "could not download" in r.stdout or "documents" in r.stdout


def test_bench_prints_exactly_one_json_line_on_stdout():
    """The driver parses bench.py's stdout: one JSON line, whatever libraries print (they are routed to stderr)."""
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d.get("impl") == "reference"          # on a CPU box: {"impl": "reference", "unavailable": ...}


def test_gemm_stats_entries_contract():
    from pretraining_llm_b200 import ops
    assert ops.gemm_stats_entries(50304) == 2 * 197        # 256-column tiles x 2 epilogue groups
    assert ops.gemm_stats_entries(256) == 2 and ops.gemm_stats_entries(257) == 4
    assert ops.gemm_stats_entries(1000, bn=64) == 16       # 64-column tiles: one group
