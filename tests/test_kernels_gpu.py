"""GPU tier: every native kernel against a plain PyTorch fp32 reference of the same op."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BF16, F32 = torch.bfloat16, torch.float32


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tier needs a CUDA device"
    from pretraining_llm_b200 import ops as _ops
    from pretraining_llm_b200.ops import lib
    assert lib.available()
    lib.load()
    return _ops


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _rand(*shape, scale=1.0, dtype=BF16):
    return (torch.randn(*shape, device="cuda", dtype=F32) * scale).to(dtype)


# ------------------------------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 128), (384, 768, 1024), (200, 328, 72),
                                   (1024, 2304, 768), (32, 3072, 768)])
@pytest.mark.parametrize("bn", [0, 64, 128, 256])
def test_gemm_tn_bf16(ops, M, N, K, bn):
    torch.manual_seed(M + N + K)
    a, b = _rand(M, K), _rand(N, K)
    c = ops.gemm(a, b, bn=bn)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    assert _rel(c, ref) < 6e-3, _rel(c, ref)


@pytest.mark.parametrize("bn", [128, 256])
@pytest.mark.parametrize("major", ["tn", "nt", "tt"])
def test_gemm_2cta(ops, bn, major):
    """cta_group::2 kernel: all operand majors, tails in M / N / K, bf16 and fp32 epilogues."""
    torch.manual_seed(11)
    M, N, K = 1000, 712, 328
    if major == "tn":
        a, b = _rand(M, K), _rand(N, K)
        bias = torch.randn(N, device="cuda")
        c = ops.gemm(a, b, bias=bias, relu=True, bn=bn, two_cta=True)
        ref = F.relu(a.float() @ b.float().t() + bias)
        assert _rel(c, ref) < 6e-3, _rel(c, ref)
    elif major == "nt":
        a, b = _rand(M, K), _rand(K, N)
        c = ops.gemm(a, b, b_mn=True, bn=bn, two_cta=True)
        assert _rel(c, a.float() @ b.float()) < 6e-3
    else:
        a, b = _rand(K, M), _rand(K, N)
        out = torch.zeros(M, N, device="cuda")
        ops.gemm(a, b, out=out, a_mn=True, b_mn=True, f32_mode=2, split_k=2, bn=bn, two_cta=True)
        assert _rel(out, a.float().t() @ b.float()) < 2e-3
    big_a, big_b = _rand(2048, 768), _rand(3072, 768)
    c = ops.gemm(big_a, big_b, bn=bn, two_cta=True)
    assert _rel(c, big_a.float() @ big_b.float().t()) < 6e-3


@pytest.mark.parametrize("M", [1, 3, 32, 48, 128])
def test_gemm_skinny(ops, M):
    """Decode-sized weight-streaming GEMM: automatic and forced cluster split-K (DSMEM reduction)."""
    torch.manual_seed(12 + M)
    for (N, K) in ((6144, 2048), (2048, 8192), (1000, 328)):
        x, w = _rand(M, K), _rand(N, K, scale=0.05)
        bias = torch.randn(N, device="cuda")
        ref = F.relu(x.float() @ w.float().t() + bias)
        for sk in (0, 1, 2, 4, 8):
            out = ops.gemm_skinny(x, w, bias=bias, relu=True, split_k=sk)
            assert _rel(out, ref) < 6e-3, (N, K, sk, _rel(out, ref))


def test_gemm_epilogues(ops):
    torch.manual_seed(1)
    M, N, K = 512, 1024, 256
    a, b = _rand(M, K), _rand(N, K, scale=0.1)
    bias = torch.randn(N, device="cuda")
    c = ops.gemm(a, b, bias=bias, relu=True)
    ref = F.relu(a.float() @ b.float().t() + bias)
    assert _rel(c, ref) < 6e-3
    act = _rand(M, N)
    act = torch.where(act > 0, act, torch.zeros_like(act))
    c2 = ops.gemm(a, b, mask=act, alpha=0.5)
    ref2 = 0.5 * (a.float() @ b.float().t()) * (act > 0)
    assert _rel(c2, ref2) < 6e-3


@pytest.mark.parametrize("two_cta", [False, True])
@pytest.mark.parametrize("shape", [(512, 1024, 256), (300, 200, 128), (640, 3072, 768)])
def test_gemm_relu_sign_bits(ops, two_cta, shape):
    """Forward GEMM (+bias, ReLU) writes one sign bit per output; the dgrad GEMM masks with them."""
    torch.manual_seed(4)
    M, N, K = shape
    a, b = _rand(M, K), _rand(N, K, scale=0.1)
    bias = torch.randn(N, device="cuda")
    bits = torch.zeros(M, (N + 63) // 64, device="cuda", dtype=torch.int64)
    act = ops.gemm(a, b, bias=bias, relu=True, relu_bits_out=bits, two_cta=two_cta)
    # bit j of word w of row m  <->  act[m, 64 w + j] > 0
    cols = torch.arange(N, device="cuda")
    got = ((bits[:, cols // 64] >> (cols % 64)) & 1).bool()
    assert torch.equal(got, act > 0)
    # unused high bits of the last word stay clear
    if N % 64:
        assert int((bits[:, -1] >> (N % 64)).abs().max()) == 0
    # dgrad: dH[M, N] = (dOut[M, P] @ W2[P, N]) * (act > 0)
    P = 256
    dout, w2 = _rand(M, P), _rand(P, N, scale=0.1)
    dh = ops.gemm(dout, w2, b_mn=True, relu_bits=bits, two_cta=two_cta)
    ref = (dout.float() @ w2.float()) * (act > 0)
    assert _rel(dh, ref) < 6e-3
    assert torch.equal(dh == 0, ~(act > 0) | (dh == 0))


@pytest.mark.parametrize("bn", [64, 128, 256])
def test_gemm_nt_dgrad(ops, bn):
    """dX[M,K] = dY[M,N] @ W[N,K]: B operand MN-major."""
    torch.manual_seed(2)
    M, N, K = 384, 640, 512
    dy, w = _rand(M, N), _rand(N, K, scale=0.1)
    dx = ops.gemm(dy, w, b_mn=True, bn=bn)
    ref = dy.float() @ w.float()
    assert _rel(dx, ref) < 6e-3, _rel(dx, ref)


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_gemm_tt_wgrad(ops, bn, mode):
    """dW[N,K] = dY[M,N]^T @ X[M,K]: both operands MN-major, fp32 output."""
    torch.manual_seed(3)
    M, N, K = 1024, 384, 320
    dy, x = _rand(M, N), _rand(M, K)
    base = torch.randn(N, K, device="cuda") if mode else torch.zeros(N, K, device="cuda")
    out = base.clone()
    ops.gemm(dy, x, out=out, a_mn=True, b_mn=True, f32_mode=mode, bn=bn, split_k=(3 if mode == 2 else 0))
    ref = dy.float().t() @ x.float() + (base if mode else 0)
    assert _rel(out, ref) < 2e-3, _rel(out, ref)


def test_gemm_stats(ops):
    """Softmax statistics from the LM-head epilogue reproduce logsumexp of the bf16 logits."""
    torch.manual_seed(4)
    M, V, K = 300, 1000, 128
    h, w = _rand(M, K), _rand(V, K, scale=0.2)
    bias = torch.randn(V, device="cuda") * 0.1
    Vp = (V + 7) // 8 * 8
    logits = torch.empty(M, Vp, device="cuda", dtype=BF16)
    nt = 2 * ((V + 255) // 256)          # one entry per 256-column tile and epilogue warp group
    stats = torch.zeros(nt, M, 2, device="cuda")
    ops.gemm(h, w, out=logits[:, :V], bias=bias, stats=stats, bn=256)
    lse = torch.empty(M, device="cuda")
    rows = torch.empty(M, device="cuda")
    tot = torch.zeros(1, device="cuda")
    tgt = torch.randint(0, V, (M,), device="cuda")
    ops.ce_finalize(stats, nt, logits, tgt, lse, rows, tot, V)
    ref_lse = torch.logsumexp(logits[:, :V].float(), dim=-1)
    assert torch.allclose(lse, ref_lse, atol=2e-3, rtol=1e-4)
    ref_loss = F.cross_entropy(logits[:, :V].float(), tgt, reduction="none")
    assert torch.allclose(rows, ref_loss, atol=3e-3, rtol=1e-4)
    assert abs(tot.item() - ref_loss.sum().item()) < 1e-2 * M
    # gradient in place
    ref_g = (torch.softmax(logits[:, :V].float(), -1) - F.one_hot(tgt, V)) / M
    dbias = torch.zeros(V, device="cuda")
    ops.ce_grad(logits[:, :V], lse, tgt, dbias, V, 1.0 / M)
    assert _rel(logits[:, :V], ref_g) < 1e-2
    assert _rel(dbias, ref_g.sum(0)) < 2e-2


# ------------------------------------------------------------------------------------------ LN / embed / AdamW
@pytest.mark.parametrize("d", [256, 768, 1024, 2048])
def test_layernorm_fwd_bwd(ops, d):
    torch.manual_seed(5)
    M = 777
    x = torch.randn(M, d, device="cuda")
    br = _rand(M, d)
    g = torch.randn(d, device="cuda")
    b = torch.randn(d, device="cuda")
    x_ref = x + br.float()
    xr = x_ref.clone().requires_grad_(True)
    gr, brr = g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y_ref = F.layer_norm(xr, (d,), gr, brr, 1e-5)
    xk = x.clone()
    y, mean, rstd = ops.ln_fwd(xk, g, b, branch=br)           # in place
    xo = torch.empty_like(x)
    y2, _, _ = ops.ln_fwd(x, g, b, branch=br, x_out=xo)      # out of place
    assert torch.equal(xo, xk) and torch.equal(y2, y)
    assert torch.allclose(xk, x_ref, atol=1e-6)
    assert _rel(y, y_ref) < 5e-3
    assert torch.allclose(mean, x_ref.mean(-1), atol=1e-5)
    dy = _rand(M, d)
    dx_in = torch.randn(M, d, device="cuda")
    y_ref.backward(dy.float())
    dgamma, dbeta, cs = (torch.zeros(d, device="cuda") for _ in range(3))
    dxb = torch.empty(M, d, device="cuda", dtype=BF16)
    dx = ops.ln_bwd(dy, xk, mean, rstd, g, dgamma, dbeta, dx_in=dx_in, dx_bf16=dxb, colsum_out=cs)
    assert _rel(dx, xr.grad + dx_in) < 1e-4
    assert _rel(dxb, dx) < 5e-3
    assert _rel(dgamma, gr.grad) < 1e-4 and _rel(dbeta, brr.grad) < 1e-4
    assert _rel(cs, dx_in.sum(0)) < 1e-4
    # no residual-gradient input, no colsum
    dx2 = ops.ln_bwd(dy, xk, mean, rstd, g, dgamma, dbeta)
    assert _rel(dx2, xr.grad) < 1e-4


def test_embed(ops):
    torch.manual_seed(6)
    B, T, d, V = 3, 17, 256, 500
    idx = torch.randint(0, V, (B, T), device="cuda")
    wte, wpe = torch.randn(V, d, device="cuda"), torch.randn(32, d, device="cuda")
    x = ops.embed_fwd(idx, wte, wpe)
    ref = wte[idx] + wpe[:T][None]
    assert torch.equal(x.view(B, T, d), ref)
    dx = torch.randn(B * T, d, device="cuda")
    dwte, dwpe = torch.zeros_like(wte), torch.zeros_like(wpe)
    ops.embed_bwd(idx, dx, dwte, dwpe[:T])
    ref_wte = torch.zeros_like(wte).index_add_(0, idx.view(-1), dx)
    assert torch.allclose(dwte, ref_wte, atol=1e-5)
    assert torch.allclose(dwpe[:T], dx.view(B, T, d).sum(0), atol=1e-5)


def test_adamw_matches_torch(ops):
    torch.manual_seed(7)
    n = 100003 // 4 * 4
    p = torch.randn(n, device="cuda")
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref_p], lr=5e-4)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    pb = torch.empty(n, device="cuda", dtype=BF16)
    hyper = torch.empty(8, device="cuda")
    for step in range(1, 11):
        g = torch.randn(n, device="cuda")
        ref_p.grad = g.clone()
        opt.step()
        hyper.copy_(torch.tensor(ops.adamw_hyper(5e-4, step)))
        ops.adamw(p, g, m, v, pb, hyper)
    assert torch.allclose(p, ref_p.data, rtol=1e-5, atol=1e-7)
    st = opt.state[ref_p]
    assert torch.allclose(m, st["exp_avg"], rtol=1e-5, atol=1e-6)
    assert torch.allclose(v, st["exp_avg_sq"], rtol=1e-5, atol=1e-7)
    assert torch.equal(pb, p.to(BF16))


def test_colsum_cast(ops):
    a = _rand(1000, 328)
    assert _rel(ops.colsum(a), a.float().sum(0)) < 1e-5
    x = torch.randn(4096 + 4, device="cuda")
    assert torch.equal(ops.cast_bf16(x), x.to(BF16))


# ------------------------------------------------------------------------------------------ attention
def _ref_attn(qkv, B, T, H, hs):
    q, k, v = qkv.float().view(B, T, 3, H, hs).permute(2, 0, 3, 1, 4)
    o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return o.transpose(1, 2).reshape(B * T, H * hs)


@pytest.mark.parametrize("impl", ["tc", "mma"])
@pytest.mark.parametrize("hs", [64, 128])
@pytest.mark.parametrize("T", [1, 17, 128, 200, 1024])
def test_flash_attention_fwd_bwd(ops, hs, T, impl):
    torch.manual_seed(8)
    B, H = 2, 3
    qkv = _rand(B * T, 3 * H * hs)
    out, lse = ops.attn_fwd(qkv, B, T, H, hs, impl=impl)
    qr = qkv.float().requires_grad_(True)
    ref = _ref_attn(qr, B, T, H, hs)
    assert _rel(out, ref) < 8e-3, _rel(out, ref)
    q, k = qkv.float().view(B, T, 3, H, hs).permute(2, 0, 3, 1, 4)[:2]
    s = (q @ k.transpose(-1, -2)) / math.sqrt(hs)
    s = s.masked_fill(torch.triu(torch.ones(T, T, device="cuda", dtype=torch.bool), 1), float("-inf"))
    assert torch.allclose(lse, torch.logsumexp(s, -1), atol=2e-3, rtol=1e-3)
    dout = _rand(B * T, H * hs)
    ref.backward(dout.float())
    dqkv = ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs, impl=impl)
    g = qr.grad.view(B * T, 3, H * hs)
    mine = dqkv.float().view(B * T, 3, H * hs)
    floor = 1e-3 * dout.float().norm()   # T == 1: dq is exactly 0 in exact arithmetic
    for i, name in enumerate("qkv"):
        err = (mine[:, i] - g[:, i]).norm() / (g[:, i].norm() + floor)
        assert err < 2e-2, (name, err.item())


def test_ln_bwd_two_step_equals_one_step(ops):
    """ln_bwd_partial x 2 + ln_bwd_finalize2 (one reduce launch for the two LayerNorms of a block) == two ln_bwd calls."""
    torch.manual_seed(31)
    M, d = 1000, 768
    outs = []
    for two_step in (False, True):
        torch.manual_seed(32)
        res = []
        parts = [torch.empty(ops.ln_bwd_nparts(M, d) * 3 * d, device="cuda") for _ in range(2)]
        grads = []
        for k, cs in enumerate((True, False)):
            x = torch.randn(M, d, device="cuda"); dy = _rand(M, d); g = torch.randn(d, device="cuda")
            mean = x.mean(-1); rstd = (x.var(-1, unbiased=False) + 1e-5).rsqrt()
            dxi = torch.randn(M, d, device="cuda")
            dg, db, col = torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda")
            dxo, dxb = torch.empty(M, d, device="cuda"), torch.empty(M, d, device="cuda", dtype=BF16)
            if two_step:
                ops.ln_bwd_partial(dy, x, mean, rstd, g, parts[k], dx_in=dxi, dx_out=dxo, dx_bf16=dxb, with_colsum=cs)
            else:
                ops.ln_bwd(dy, x, mean, rstd, g, dg, db, dx_in=dxi, dx_out=dxo, dx_bf16=dxb,
                           colsum_out=col if cs else None, partial=parts[k])
            grads.append((dg, db, col if cs else None))
            res += [dxo, dxb.float()]
        if two_step:
            ops.ln_bwd_finalize2(M, d, parts[0], *grads[0], parts[1], *grads[1])
        for dg, db, col in grads:
            res += [dg, db] + ([col] if col is not None else [])
        outs.append(res)
    for a, b in zip(*outs):
        assert torch.equal(a, b)


@pytest.mark.parametrize("tiles", ["tc1", "tc2"])
@pytest.mark.parametrize("hs", [64, 128])
@pytest.mark.parametrize("T", [100, 384, 640, 1024, 2048])
def test_flash_attention_fwd_both_layouts(ops, hs, T, tiles):
    """One and two query tiles per CTA (the default picks by head size): same output and log-sum-exp.  T = 100 / 384 / 640
    give an odd number of 128-row tiles, i.e. a last pair whose second tile is empty or partial; B*H*pairs > #SMs at
    T = 1024 makes the persistent CTAs walk several work items."""
    torch.manual_seed(21)
    B, H = (25, 6) if T == 1024 else ((3, 4) if T == 2048 else (2, 3))
    qkv = _rand(B * T, 3 * H * hs)
    out, lse = ops.attn_fwd(qkv, B, T, H, hs, impl=tiles)
    ref = _ref_attn(qkv.float(), B, T, H, hs)
    assert _rel(out, ref) < 8e-3, _rel(out, ref)
    out2, lse2 = ops.attn_fwd(qkv, B, T, H, hs, impl="mma")
    assert torch.allclose(lse, lse2, atol=2e-3, rtol=1e-3)


@pytest.mark.parametrize("hs", [64, 128])
def test_decode_attention(ops, hs):
    torch.manual_seed(9)
    B, H, Tmax, L = 3, 4, 64, 37
    hist = _rand(B * L, 3 * H * hs)
    kc = torch.zeros(B, H, Tmax, hs, device="cuda", dtype=BF16)
    vc = torch.zeros_like(kc)
    ops.kv_append(hist, kc, vc, B, L, H, hs, Tmax, pos0=0)
    k_ref = hist.view(B, L, 3, H, hs)[:, :, 1].permute(0, 2, 1, 3)
    assert torch.equal(kc[:, :, :L], k_ref)
    qkv = hist.view(B, L, -1)[:, -1].contiguous()
    out = torch.empty(B, H * hs, device="cuda", dtype=BF16)
    pos = torch.tensor([L - 1], device="cuda", dtype=torch.int32)
    ops.attn_decode(qkv, kc, vc, out, B, H, hs, Tmax, pos_dev=pos)
    ref = _ref_attn(hist, B, L, H, hs).view(B, L, -1)[:, -1]
    assert _rel(out, ref) < 8e-3
    # fused append: the cache only holds L-1 positions, the kernel adds the last one itself
    kc2, vc2 = kc.clone(), vc.clone()
    kc2[:, :, L - 1] = 0
    vc2[:, :, L - 1] = 0
    out2 = torch.empty_like(out)
    ops.attn_decode(qkv, kc2, vc2, out2, B, H, hs, Tmax, pos_dev=pos, append=True)
    assert _rel(out2, ref) < 8e-3 and torch.equal(kc2, kc) and torch.equal(vc2, vc)


def test_sampler_distribution(ops):
    torch.manual_seed(10)
    V, B = 50, 4096
    row = torch.randn(V, device="cuda") * 2
    logits = row.to(BF16)[None].repeat(B, 1).contiguous()
    nxt = torch.empty(B, device="cuda", dtype=torch.int64)
    ops.sample(logits, nxt, V, seed=123, step=5)
    assert int(nxt.min()) >= 0 and int(nxt.max()) < V
    p = torch.softmax(logits[0].float(), -1)
    counts = torch.bincount(nxt, minlength=V).float()
    chi2 = (((counts - B * p) ** 2) / (B * p + 1e-9))[B * p > 5].sum().item()
    assert chi2 < 3 * V, chi2  # dof ~ V: generous bound, catches any systematic error
    nxt2 = torch.empty_like(nxt)
    ops.sample(logits, nxt2, V, seed=123, step=5)
    assert torch.equal(nxt, nxt2)
