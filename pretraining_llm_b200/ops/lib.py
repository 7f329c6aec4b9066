"""Build + load the native library (``libb200llm.so``) and expose its C ABI through ctypes.

The library is plain CUDA C++ with ``extern "C"`` entry points declared in
``include/b200llm.h``; every entry point takes raw device pointers and a ``cudaStream_t``, so
Python passes ``tensor.data_ptr()`` and ``torch.cuda.current_stream().cuda_stream`` and the
launches are capturable in CUDA graphs.  It is built in-tree with
``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (``build()``), so the ``.so``
travels with a repo snapshot and shows up in ``/proc/<pid>/maps`` of any process that used it.

There is no fallback: on a CUDA device, ops raise if the library is missing.
"""
from __future__ import annotations

import ctypes
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import Optional

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
OUT_DIR = Path(__file__).resolve().parent
SO_PATH = OUT_DIR / "libb200llm.so"
OBJ_DIR = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC",
          "-Xcompiler", "-Wall", "--expt-relaxed-constexpr", "-I", str(INCLUDE), "-I", str(CSRC)]

_lib: Optional[ctypes.CDLL] = None


def sources():
    return sorted(CSRC.rglob("*.cu"))


def _digest(path: Path) -> str:
    h = hashlib.sha1()
    h.update(path.read_bytes())
    for hdr in sorted(list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list(INCLUDE.glob("*.h"))):
        h.update(hdr.read_bytes())
    h.update(" ".join(ARCH_FLAGS + COMMON).encode())
    return h.hexdigest()[:16]


def _compile_one(src: Path, verbose: bool) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dig = _digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj
    cmd = [NVCC, *ARCH_FLAGS, *COMMON, "-Xptxas", "-v", "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = OBJ_DIR / (src.stem + ".ptxas.log")
    log.write_text(r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"nvcc failed for {src}")
    if verbose:
        print(f"[build] {src.relative_to(ROOT)} ok")
    stamp.write_text(dig)
    return obj


def build(force: bool = False, verbose: bool = True) -> Path:
    """Compile every ``csrc/**/*.cu`` for sm_100a and link ``libb200llm.so`` in-tree."""
    srcs = sources()
    if not srcs:
        raise RuntimeError("no CUDA sources found under csrc/")
    if force:
        for f in OBJ_DIR.glob("*.sha"):
            f.unlink()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose), srcs))
    newest = max(o.stat().st_mtime for o in objs)
    if force or not SO_PATH.exists() or SO_PATH.stat().st_mtime < newest:
        cmd = [NVCC, *ARCH_FLAGS, "-shared", "-o", str(SO_PATH), *map(str, objs), "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
        if verbose:
            print(f"[build] linked {SO_PATH.relative_to(ROOT)}")
    return SO_PATH


def available() -> bool:
    """True when the native library can be used: CUDA device present and the .so built."""
    try:
        import torch
        if not torch.cuda.is_available():
            return False
    except Exception:
        return False
    if not SO_PATH.exists():
        raise RuntimeError(
            f"{SO_PATH} is missing on a CUDA machine: run `python -c 'import __graft_entry__ as g; "
            "g.build()'` (there is no PyTorch fallback for the CUDA hot path)")
    return True


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not SO_PATH.exists():
            raise RuntimeError(f"{SO_PATH} not built; run pretraining_llm_b200.ops.lib.build()")
        _lib = ctypes.CDLL(str(SO_PATH), mode=ctypes.RTLD_GLOBAL)
        _declare(_lib)
    return _lib


# ---------------------------------------------------------------------------------------
# C ABI: signatures are parsed from include/b200llm.h (single source of truth)
# ---------------------------------------------------------------------------------------
_CTYPES = {
    "int": ctypes.c_int, "float": ctypes.c_float, "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong, "cudaStream_t": ctypes.c_void_p, "size_t": ctypes.c_size_t,
}


def _parse_header():
    import re
    text = (INCLUDE / "b200llm.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    sigs = {}
    for m in re.finditer(r"^\s*(const char\*|long long|int)\s+(b200_\w+)\s*\(([^;]*)\)\s*;", text, flags=re.M):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                    continue
                ty = a.replace("const ", "").rsplit(" ", 1)[0].strip()
                argtypes.append(_CTYPES[ty])
        sigs[name] = (argtypes, ctypes.c_char_p if "char" in ret else (ctypes.c_longlong if "long" in ret else ctypes.c_int))
    return sigs


def _declare(lib: ctypes.CDLL) -> None:
    for name, (argtypes, restype) in _parse_header().items():
        if not hasattr(lib, name):
            raise RuntimeError(f"{name} is declared in include/b200llm.h but missing from {SO_PATH.name}")
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().b200_last_error()
        raise RuntimeError(f"native call failed ({what}): rc={rc} {msg.decode() if msg else ''}")


if __name__ == "__main__":
    build(force="--force" in sys.argv)
