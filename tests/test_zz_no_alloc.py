"""The allocation-free LayerNorm-backward entry point, called straight through the C ABI.

``b200_ln_bwd_static`` is for hosts that cannot give the kernel a scratch buffer - the device
memory all belongs to their own arena, or cudaMalloc's device-wide synchronisation is not
acceptable in their loop - so the scratch is a fixed array inside the library.  This test drives it
with ctypes (no Python wrapper in between), twice in a row with different inputs because the array
keeps the previous call's contents, and checks every output against the regular entry point fed
with an ordinary caller-allocated scratch buffer.

Collected last on purpose: if it fails under ``-x`` every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

BF16, F32 = torch.bfloat16, torch.float32


def _lib():
    from pretraining_llm_b200.ops import lib
    assert torch.cuda.is_available() and lib.available()
    return lib.load()           # a ctypes.CDLL: the calls below are the C ABI itself


def _p(t):
    return ctypes.c_void_p(None if t is None else t.data_ptr())


def _run(L, fn_static, M, d, seed, with_colsum):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(M, d, device="cuda", generator=g)
    dy = torch.randn(M, d, device="cuda", generator=g).to(BF16)
    dx_in = torch.randn(M, d, device="cuda", generator=g)
    gamma = torch.randn(d, device="cuda", generator=g)
    mean = x.mean(-1).contiguous()
    rstd = (x.var(-1, unbiased=False) + 1e-5).rsqrt().contiguous()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for static in (False, True):
        dx = torch.empty(M, d, device="cuda")
        dxb = torch.empty(M, d, device="cuda", dtype=BF16)
        dgamma, dbeta, cs = (torch.full((d,), float("nan"), device="cuda") for _ in range(3))
        csp = _p(cs) if with_colsum else _p(None)
        if static:
            rc = fn_static(_p(dy), _p(x), _p(mean), _p(rstd), _p(gamma), _p(dx_in), _p(dx), _p(dxb), _p(dgamma),
                           _p(dbeta), csp, M, d, 0, st)
        else:
            nparts = L.b200_ln_bwd_nparts(M, d)
            partial = torch.empty(nparts * 3 * d, device="cuda")
            rc = L.b200_ln_bwd(_p(dy), _p(x), _p(mean), _p(rstd), _p(gamma), _p(dx_in), _p(dx), _p(dxb), _p(partial),
                               _p(dgamma), _p(dbeta), csp, M, d, 0, st)
        assert rc == 0, L.b200_last_error()
        outs.append((dx, dxb.float(), dgamma, dbeta, cs if with_colsum else torch.zeros(d, device="cuda")))
    torch.cuda.synchronize()
    for a, b in zip(*outs):
        assert torch.isfinite(b).all()
        rel = ((a - b).norm() / (a.norm() + 1e-12)).item()
        assert rel < 1e-4, rel            # the tolerance the LayerNorm kernel tests use


def test_ln_bwd_static_matches_regular_entry_point_twice():
    L = _lib()
    fn = L.b200_ln_bwd_static
    assert L.b200_ln_bwd_static_capacity() == 296 * 3 * 2048
    # two calls in a row on the shared array, different shapes and inputs
    _run(L, fn, M=777, d=768, seed=1, with_colsum=True)
    _run(L, fn, M=4096, d=2048, seed=2, with_colsum=False)
    _run(L, fn, M=333, d=256, seed=3, with_colsum=True)


def test_ln_bwd_static_refuses_oversized_requests():
    L = _lib()
    M, d = 4096, 4096                       # 296 rows x 2 x 4096 floats > the built-in array
    z = torch.zeros(8, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.b200_ln_bwd_static(_p(z), _p(z), _p(z), _p(z), _p(z), _p(None), _p(z), _p(None), _p(z), _p(z), _p(None),
                              M, d, 0, st)
    assert rc != 0
    msg = L.b200_last_error().decode()
    assert "built-in array" in msg and "4096" in msg
    torch.cuda.synchronize()
