"""GPU tier, second batch: the shapes the benchmarks run (long sequences, many work items per persistent CTA, the
real ragged vocabulary), the all-gather GEMM with emulated peers, the sharded data-parallel machinery with one rank,
native gradient accumulation, the autograd-path fixes and a loss overlay against the reference module."""
import ctypes
import math
import os
import sys

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

BF16, F32 = torch.bfloat16, torch.float32
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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


def _ref_attn(qkv_f32, B, T, H, hs):
    q, k, v = qkv_f32.view(B, T, 3, H, hs).permute(2, 0, 3, 1, 4)
    o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    return o.permute(0, 2, 1, 3).reshape(B * T, H * hs)


# ------------------------------------------------------------------------------------------ attention at benchmark scale
@pytest.mark.parametrize("hs,T,B,H", [(128, 2048, 2, 20), (64, 1024, 4, 12), (64, 2048, 2, 12)])
def test_attention_bwd_many_items_per_cta(ops, hs, T, B, H):
    """Persistent backward with every CTA walking >= 3 work items (B*H*T/128 > 2*148), T up to the 1.3B shape's 2048:
    the double-buffered multi-item path against the fp32 oracle."""
    assert B * H * (T // 128) > 2 * 148
    torch.manual_seed(hs + T)
    qkv = _rand(B * T, 3 * H * hs)
    out, lse = ops.attn_fwd(qkv, B, T, H, hs)
    qr = qkv.float().requires_grad_(True)
    ref = _ref_attn(qr, B, T, H, hs)
    assert _rel(out, ref) < 8e-3
    dout = _rand(B * T, H * hs)
    ref.backward(dout.float())
    dqkv = ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs)
    g = qr.grad.view(B * T, 3, H * hs)
    mine = dqkv.float().view(B * T, 3, H * hs)
    for i, name in enumerate("qkv"):
        err = ((mine[:, i] - g[:, i]).norm() / g[:, i].norm()).item()
        assert err < 2e-2, (name, err)


# ------------------------------------------------------------------------------------------ LM head at the real vocabulary
def test_lm_head_stats_and_grad_at_real_vocab(ops):
    """V = 50304 = 196.5 tiles of 256 columns: the ragged last tile, targets on tile edges, M = 4096."""
    torch.manual_seed(5)
    M, V, K = 4096, 50304, 768
    h, w = _rand(M, K), _rand(V, K, scale=0.05)
    bias = torch.randn(V, device="cuda") * 0.1
    logits = torch.empty(M, V, device="cuda", dtype=BF16)
    nt = ops.gemm_stats_entries(V)
    stats = torch.zeros(nt, M, 2, device="cuda")
    ops.gemm(h, w, out=logits, bias=bias, stats=stats, bn=256)
    lse, rows, tot = torch.empty(M, device="cuda"), torch.empty(M, device="cuda"), torch.zeros(1, device="cuda")
    tgt = torch.randint(0, V, (M,), device="cuda")
    tgt[:8] = torch.tensor([0, V - 1, V - 2, 255, 256, 50175, 50176, 50303], device="cuda")   # tile edges
    ops.ce_finalize(stats, nt, logits, tgt, lse, rows, tot, V)
    lf = logits.float()
    assert torch.allclose(lse, torch.logsumexp(lf, -1), atol=2e-3, rtol=1e-4)
    ref_loss = F.cross_entropy(lf, tgt, reduction="none")
    assert torch.allclose(rows, ref_loss, atol=3e-3, rtol=1e-4)
    ref_g = (torch.softmax(lf, -1) - F.one_hot(tgt, V)) / M
    dbias = torch.zeros(V, device="cuda")
    ops.ce_grad(logits, lse, tgt, dbias, V, 1.0 / M)
    assert _rel(logits, ref_g) < 1e-2
    assert _rel(dbias, ref_g.sum(0)) < 2e-2


def test_sampler_distribution_at_real_vocab(ops):
    """Chi-square of 262144 draws over V = 50304 (categories with an expected count < 8 pooled into one bin)."""
    torch.manual_seed(10)
    V, B = 50304, 262144
    row = (torch.randn(V, device="cuda") * 2.5).to(BF16)
    logits = row[None].expand(B, V).contiguous()
    nxt = torch.empty(B, device="cuda", dtype=torch.int64)
    ops.sample(logits, nxt, V, seed=7, step=3)
    assert int(nxt.min()) >= 0 and int(nxt.max()) < V
    p = torch.softmax(row.float(), -1)
    counts = torch.bincount(nxt, minlength=V).double()
    exp = (B * p).double()
    big = exp >= 8
    o = torch.cat([counts[big], counts[~big].sum()[None]])
    e = torch.cat([exp[big], exp[~big].sum()[None]])
    chi2 = (((o - e) ** 2) / e).sum().item()
    dof = o.numel() - 1
    assert chi2 < dof + 5 * math.sqrt(2 * dof), (chi2, dof)     # 5 sigma of the chi-square distribution


# ------------------------------------------------------------------------------------------ all-gather GEMM, emulated peers
def _gemm_ag(lib, a, bw, out, b_mn, world, rank, gen, err, jobs, wait=None, bias=None, relu=False):
    L = lib.load()
    if b_mn:
        K, N = bw.shape
    else:
        N, K = bw.shape
    nojob = (None, None, 0, 0, None, None)
    j0 = jobs[0] if len(jobs) > 0 else nojob
    j1 = jobs[1] if len(jobs) > 1 else nojob
    wf, wc, wo = wait if wait else (None, 0, 0)
    rc = L.b200_gemm_ag(a.data_ptr(), bw.data_ptr(), out.data_ptr(), a.shape[0], N, K, a.stride(0), bw.stride(0),
                        out.stride(0), int(b_mn), None if bias is None else bias.data_ptr(), int(relu), None, None,
                        wf, wc, wo, world, rank, gen.data_ptr(), err.data_ptr(), len(jobs), *j0, *j1,
                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.b200_last_error()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_gemm_ag_with_emulated_peers(ops, world):
    """b200_gemm_ag on ONE GPU: the "peers" are separate buffers on the same device, one chunk each.  Forward form:
    the launch gathers its own B operand (tiles wait per chunk) plus a second layer into another window; backward
    form (B [K, N]): the operand is complete, the launch only carries a prefetch job.  Results vs torch.matmul, the
    windows vs the concatenated chunks."""
    from pretraining_llm_b200.ops import lib
    torch.manual_seed(world)
    M, N, K = 1024, 1536, 768
    rank = world - 1
    layer = _rand(N * K + 3 * 128, scale=0.1)               # the "layer": W [N, K] at offset 128, plus small slots
    w_off = 128
    total = layer.numel()
    chunk = ((total + world - 1) // world + 127) // 128 * 128
    def shardify(t):          # one padded chunk per "rank", each its own allocation like a peer's shard memory
        padded = torch.cat([t, torch.zeros(world * chunk - t.numel(), device="cuda", dtype=BF16)])
        return [padded[r * chunk:(r + 1) * chunk].clone() for r in range(world)]

    shards = shardify(layer)
    src = (ctypes.c_void_p * world)(*[s.data_ptr() for s in shards])
    other = _rand(total, scale=0.1)
    oshards = shardify(other)
    osrc = (ctypes.c_void_p * world)(*[s.data_ptr() for s in oshards])
    win0 = torch.zeros(world * chunk, device="cuda", dtype=BF16)
    win1 = torch.zeros(world * chunk, device="cuda", dtype=BF16)
    flags = torch.zeros(2, world, device="cuda", dtype=torch.int32)
    counters = torch.zeros(2, world, device="cuda", dtype=torch.int32)
    gen = torch.ones(1, device="cuda", dtype=torch.int32)
    err = torch.zeros(1, device="cuda", dtype=torch.int32)
    a = _rand(M, K)
    bias = torch.randn(N, device="cuda")
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    bw = win0[w_off:w_off + N * K].view(N, K)
    jobs = [(src, win0.data_ptr(), chunk, total, flags[0].data_ptr(), counters[0].data_ptr()),
            (osrc, win1.data_ptr(), chunk, total, flags[1].data_ptr(), counters[1].data_ptr())]
    _gemm_ag(lib, a, bw, out, False, world, rank, gen, err, jobs, wait=(flags[0].data_ptr(), chunk, w_off), bias=bias, relu=True)
    torch.cuda.synchronize()
    assert int(err.item()) == 0
    assert torch.equal(win0[:total], layer) and torch.equal(win1[:total], other)
    assert torch.equal(flags, torch.ones_like(flags)) and int(counters.abs().sum()) == 0
    W = layer[w_off:w_off + N * K].view(N, K)
    ref = F.relu(a.float() @ W.float().t() + bias)
    assert _rel(out, ref) < 6e-3, _rel(out, ref)
    # backward form: dx = dy @ W (B operand [K_gemm = N, N_gemm = K]) carrying one prefetch job into a fresh window
    gen.fill_(2)
    win2 = torch.zeros(world * chunk, device="cuda", dtype=BF16)
    dy = _rand(M, N)
    dx = torch.empty(M, K, device="cuda", dtype=BF16)
    _gemm_ag(lib, dy, bw, dx, True, world, rank, gen, err,
             [(osrc, win2.data_ptr(), chunk, total, flags[1].data_ptr(), counters[1].data_ptr())])
    torch.cuda.synchronize()
    assert torch.equal(win2[:total], other) and int(flags[1].min()) == 2
    assert _rel(dx, dy.float() @ W.float()) < 6e-3


# ------------------------------------------------------------------------------------------ sharded mode with one rank
def test_zero_mode_single_rank_matches_plain_engine():
    """The whole sharded data-parallel path (all-gather jobs, rolling windows, fused reduce + AdamW on shards,
    embedding rows from the master shard, CUDA graph) with world = 1 against the plain engine: same kernels for the
    math, so losses and weights agree to fp32 rounding."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import zero_check1
    zero_check1.run(dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000), 3, 4, 128, 5, 1e-3, False)
    zero_check1.run(dict(n_head=2, n_embed=256, context_length=256, vocab_size=1000), 4, 2, 256, 5, 1e-3, True)
    # gradient accumulation: every micro-step reduce-scattered into the owners' accumulators, AdamW on the last
    zero_check1.run_accum(dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000), 3, 4, 128, 3, 1e-3, 3)


def test_zero_plan_memory_claim():
    """Per-rank parameter / gradient / optimizer bytes of the 1.3B shape on 8 ranks stay under 6 GB."""
    from pretraining_llm_b200.models.layout import ParamLayout
    from pretraining_llm_b200.parallel.zero import ZeroPlan
    plan = ZeroPlan(ParamLayout(16, 2048, 2048, 50304, 24), 8)
    total = sum(plan.bytes_per_rank().values())
    assert total < 6e9, total


# ------------------------------------------------------------------------------------------ trainer paths
def _batches(n, B, T, V, seed):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        xy = torch.randint(0, V, (B, T + 1), generator=g)
        out.append((xy[:, :-1].contiguous().cuda(), xy[:, 1:].contiguous().cuda()))
    return out


def test_native_grad_accumulation_equals_big_batch():
    """Trainer's native micro-step loop (forward/backward with accumulate, one fused AdamW) vs the same tokens as one
    batch through Engine.train_step."""
    from pretraining_llm_b200.config import default_config
    from pretraining_llm_b200.models import Transformer
    from pretraining_llm_b200.runtime.trainer import Trainer
    cfg = dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000)
    torch.manual_seed(3)
    a = Transformer(**cfg, N_BLOCKS=2).cuda()
    b = Transformer(**cfg, N_BLOCKS=2).cuda()
    b.load_state_dict(a.state_dict())
    data = _batches(8, 4, 128, 1000, seed=1)
    conf = dict(default_config, grad_accum_steps=2, t_train_steps=4, dtype="bfloat16")
    it = iter(data)
    tr = Trainer(a, it, None, config=conf)
    for s in range(4):
        xb, yb = next(it)
        la = tr._train_step(xb, yb, 1e-3)
        x2 = torch.cat([data[2 * s][0], data[2 * s + 1][0]])
        y2 = torch.cat([data[2 * s][1], data[2 * s + 1][1]])
        lb = b.engine.train_step(x2, y2, 1e-3, use_graph=False)
        assert abs(float(la) - float(lb)) < 3e-3, (s, float(la), float(lb))
    dw = (a.engine.flat - b.engine.flat).abs().mean().item()
    assert dw < 5e-3 * 1e-3 * 4, dw


def test_autograd_path_accumulates_and_returns_copies():
    """module(x, y)[1].backward() twice without zero_grad adds the two gradients (p.grad must not alias the engine's
    gradient buffer), and logits handed out by forward survive the next call."""
    from pretraining_llm_b200.models import Transformer
    cfg = dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000)
    torch.manual_seed(4)
    m = Transformer(**cfg, N_BLOCKS=2).cuda()
    (x1, y1), (x2, y2) = _batches(2, 2, 128, 1000, seed=2)
    _, l1 = m(x1, y1)
    l1.backward()
    g1 = m.lm_head.weight.grad.clone()
    m.zero_grad(set_to_none=True)
    _, l2 = m(x2, y2)
    l2.backward()
    g2 = m.lm_head.weight.grad.clone()
    m.zero_grad(set_to_none=True)
    m(x1, y1)[1].backward()
    m(x2, y2)[1].backward()
    assert _rel(m.lm_head.weight.grad, g1 + g2) < 1e-3
    with torch.no_grad():
        la, _ = m(x1)
        keep = la.clone()
        m(x2)
    assert torch.equal(la, keep)


def test_loss_overlay_vs_reference_module(ref_models):
    """20 steps of a GPT-2-small-shaped stack (d=768, 12 heads, T=1024, V=50304; 4 layers to keep it short) from the
    reference's weights on the same batches: the reference module under bf16 autocast + torch AdamW vs the engine."""
    from pretraining_llm_b200.models import Transformer
    H, d, ctx, V, L, B = 12, 768, 1024, 50304, 4, 4
    torch.manual_seed(1234)
    ref = ref_models.Transformer(H, d, ctx, V, L).cuda()
    ours = Transformer(H, d, ctx, V, N_BLOCKS=L).cuda()
    ours.load_state_dict(ref.state_dict())
    opt = torch.optim.AdamW(ref.parameters(), lr=5e-4)
    worst = 0.0
    first = last = None
    for step in range(20):
        g = torch.Generator().manual_seed(500 + step)
        x0 = torch.randint(0, 1000, (B, 1), generator=g)
        toks = [x0]
        for _ in range(ctx):
            noise = torch.rand(B, 1, generator=g) < 0.1
            toks.append(torch.where(noise, torch.randint(0, 1000, (B, 1), generator=g), (toks[-1] * 7 + 13) % 1000))
        seq = torch.cat(toks, 1)
        x, y = seq[:, :-1].contiguous().cuda(), seq[:, 1:].contiguous().cuda()
        lr = 5e-4 * min(1.0, (step + 1) / 10)
        for gp in opt.param_groups:
            gp["lr"] = lr
        with torch.autocast(device_type="cuda", dtype=torch.bfloat16):
            _, loss = ref(x, y)
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        lo = float(ours.engine.train_step(x, y, lr))
        worst = max(worst, abs(lo - float(loss)))
        first = first if first is not None else lo
        last = lo
    assert worst < 0.05, worst
    assert last < first - 1.0, (first, last)       # the curve actually moved


def test_embed_bwd_sorted_is_deterministic_and_matches_index_add(ops):
    """The sort-by-token embedding backward against torch index_add_, bit-identical across two runs, with the
    accumulate form and heavy token repetition (Zipf-like ids)."""
    torch.manual_seed(12)
    B, T, d, V = 4, 256, 768, 5000
    idx = (torch.rand(B, T, device="cuda") ** 4 * V).long().clamp_(0, V - 1)     # many repeats of small ids
    dx = torch.randn(B * T, d, device="cuda")
    order = torch.sort(idx.view(-1), stable=True).indices
    outs = []
    for _ in range(2):
        dwte = torch.zeros(V, d, device="cuda")
        dwpe = torch.zeros(T, d, device="cuda")
        ops.embed_bwd_sorted(idx, order, dx, dwte, dwpe)
        outs.append((dwte, dwpe))
    torch.cuda.synchronize()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    ref = torch.zeros(V, d, device="cuda", dtype=torch.float64).index_add_(0, idx.view(-1), dx.double())
    assert _rel(outs[0][0], ref) < 1e-6
    assert _rel(outs[0][1], dx.view(B, T, d).sum(0)) < 1e-6
    ops.embed_bwd_sorted(idx, order, dx, outs[0][0], outs[0][1], accumulate=True)
    assert _rel(outs[0][0], 2 * ref) < 1e-6


@pytest.mark.parametrize("hs", [64, 128])
@pytest.mark.parametrize("splits", [1, 3, 8])
@pytest.mark.parametrize("L", [1, 37, 700])
def test_decode_attention_split_kv(ops, hs, splits, L):
    """Split-KV decode attention (coalesced rows, partials merged by the last CTA) against the fp32 oracle, with and
    without the fused KV append, twice in a row on the same scratch / tickets."""
    torch.manual_seed(L + splits)
    B, H, Tmax = 3, 4, 1024
    hist = _rand(B * L, 3 * H * hs)
    kc = torch.zeros(B, H, Tmax, hs, device="cuda", dtype=BF16)
    vc = torch.zeros_like(kc)
    ops.kv_append(hist, kc, vc, B, L, H, hs, Tmax, pos0=0)
    qkv = hist.view(B, L, -1)[:, -1].contiguous()
    pos = torch.tensor([L - 1], device="cuda", dtype=torch.int32)
    scratch = torch.empty(B * H * splits * (hs + 2), device="cuda")
    tickets = torch.zeros(B * H, device="cuda", dtype=torch.int32)
    ref = _ref_attn(hist.float(), B, L, H, hs).view(B, L, -1)[:, -1]
    for rep in range(2):
        out = torch.empty(B, H * hs, device="cuda", dtype=BF16)
        ops.attn_decode_split(qkv, kc, vc, out, B, H, hs, Tmax, scratch, tickets, splits, pos_dev=pos)
        assert _rel(out, ref) < 8e-3, (rep, _rel(out, ref))
        assert int(tickets.abs().sum()) == 0
    kc2, vc2 = kc.clone(), vc.clone()
    kc2[:, :, L - 1] = 0
    vc2[:, :, L - 1] = 0
    out2 = torch.empty_like(out)
    ops.attn_decode_split(qkv, kc2, vc2, out2, B, H, hs, Tmax, scratch, tickets, splits, pos_dev=pos, append=True)
    assert _rel(out2, ref) < 8e-3 and torch.equal(kc2, kc) and torch.equal(vc2, vc)


def test_deterministic_mode_gives_bit_identical_gradients():
    """B200_DETERMINISTIC=1 (read at import / first launch, hence the subprocess): two backward passes from the same
    weights on the same batch produce bit-identical gradient buffers - embedding scatter, bias column sums included."""
    import subprocess
    code = (
        "import torch, sys; sys.path.insert(0, %r)\n"
        "from pretraining_llm_b200.models import Transformer\n"
        "torch.manual_seed(0)\n"
        "m = Transformer(n_head=4, n_embed=256, context_length=256, vocab_size=5000, N_BLOCKS=3).cuda()\n"
        "e = m.engine\n"
        "x = (torch.rand(8, 256, device='cuda') ** 3 * 5000).long(); y = torch.randint(0, 5000, (8, 256), device='cuda')\n"
        "gs = []\n"
        "for _ in range(3):\n"
        "    w, _, loss = e.forward(x, y, train=True); e.backward(w); torch.cuda.synchronize(); gs.append(e.grad.clone())\n"
        "assert torch.equal(gs[0], gs[1]) and torch.equal(gs[1], gs[2]), 'gradients differ between runs'\n"
        "assert float(gs[0].abs().sum()) > 0\n"
        "print('DETERMINISTIC_OK')\n") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, B200_DETERMINISTIC="1"))
    assert r.returncode == 0 and "DETERMINISTIC_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_forward_embedding_runs_on_the_native_kernels():
    """Transformer.forward_embedding on a CUDA model (no grad) takes the engine's kernels and matches the PyTorch
    statement of the same blocks; with gradients enabled it stays on the differentiable PyTorch path."""
    from pretraining_llm_b200 import ops as _ops
    from pretraining_llm_b200.models import Transformer
    torch.manual_seed(6)
    cfg = dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000)
    m = Transformer(**cfg, N_BLOCKS=3).cuda()
    o = Transformer(**cfg, N_BLOCKS=3).cuda()
    o.load_state_dict(m.state_dict())
    o.use_native = False
    x = torch.randint(0, 1000, (2, 100), device="cuda")
    n0 = _ops.LAUNCHES
    with torch.no_grad():
        hid, res = m.forward_embedding(x)
        hid_o, res_o = o.forward_embedding(x)
    assert _ops.LAUNCHES > n0                                   # the native path really ran
    assert hid.shape == hid_o.shape == (2, 100, 1024) and res.shape == res_o.shape == (2, 100, 256)
    assert _rel(res, res_o) < 2e-2 and _rel(hid, hid_o) < 3e-2
    n1 = _ops.LAUNCHES
    hid_g, _ = m.forward_embedding(x)                           # grad mode: PyTorch path, differentiable
    assert _ops.LAUNCHES == n1 and hid_g.requires_grad


@pytest.mark.parametrize("shape", [(1024, 3072, 768), (1000, 712, 328), (16384, 3072, 768)])
def test_gemm_dgrad_with_fused_column_sums(ops, shape):
    """dgrad GEMM (A K-major, B [K, N]) whose epilogue also accumulates the column sums of its bf16 output (the bias
    gradient), with the ReLU sign-bit mask, against torch: the GEMM result, a fresh sum and an accumulated one."""
    torch.manual_seed(13)
    M, N, K = shape
    a, b = _rand(M, K), _rand(K, N, scale=0.1)
    bits = torch.randint(-2 ** 62, 2 ** 62, (M, (N + 63) // 64), device="cuda", dtype=torch.int64)
    keep = ((bits.unsqueeze(-1) >> torch.arange(64, device="cuda")) & 1).bool().view(M, -1)[:, :N]
    ref = (a.float() @ b.float()) * keep
    cs = torch.zeros(N, device="cuda")
    out = ops.gemm(a, b, b_mn=True, relu_bits=bits, colsum_out=cs, two_cta=True)
    assert _rel(out, ref) < 6e-3
    want = out.float().sum(0)
    assert _rel(cs, want) < 1e-4, _rel(cs, want)
    ops.gemm(a, b, out=out, b_mn=True, relu_bits=bits, colsum_out=cs, two_cta=True)       # accumulates
    assert _rel(cs, 2 * want) < 1e-4
