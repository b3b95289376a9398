"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op
(SURVEY §4.3 "kernel correctness" tier).  Run on a B200 with ``pytest -m gpu``."""
import math

import pytest
import torch
import torch.nn.functional as F

from distributed_training_b200.ops import native

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bf16(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16).to(DEV)


def test_library_loaded_on_gpu_box():
    assert native.available(), \
        "libdtb200.so must be present on a GPU machine (python -m distributed_training_b200._native)"
    assert native.lib().dtb_version() >= 100


# ------------------------------------------------------------------------ GEMM
def _gemm_ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()          # logical [M,K]
    B = b.float().t() if b_mn else b.float()          # logical [N,K]
    return A @ B.t()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("cta_pair,tile_n", [(1, 128), (1, 256), (2, 256)])
@pytest.mark.parametrize("m,n,k", [(256, 512, 256), (384, 320, 200), (1024, 768, 1024)])
def test_gemm_layouts_and_tiles(a_mn, b_mn, cta_pair, tile_n, m, n, k):
    a = _bf16(*((k, m) if a_mn else (m, k)), seed=1)
    b = _bf16(*((k, n) if b_mn else (n, k)), seed=2)
    out = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    native.gemm(a, b, out, m=m, n=n, k=k, a_mn=a_mn, b_mn=b_mn, cta_pair=cta_pair, tile_n=tile_n)
    ref = _gemm_ref(a, b, a_mn, b_mn)
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * math.sqrt(k) + 0.1 * ref.abs().max().item() / 32, f"max err {err}"
    assert torch.allclose(out.float(), ref, atol=0.25, rtol=2e-2)


def _gemm_ref_bf16_inputs(a, b, a_mn, b_mn):
    """fp32 oracle computed on the GPU in chunks (the north-star shapes are too large for one fp32 matmul on a
    shared box to be cheap on the host; TF32 is disabled so the oracle is true fp32)."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        return _gemm_ref(a, b, a_mn, b_mn)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


# The shapes the benchmarks actually run ("on the north-star shapes"): the wide MLP's 8192-class GEMMs in
# all four operand layouts, GPT-2 XL's fc forward (K = 1600: 25 k-blocks, N tail), its fc-proj wgrad (K = 4096 rows),
# the short-K family of GPT-2 small, and the LM head over the padded vocabulary at K = 768.
NORTH_STAR = [
    # m, n, k, a_mn, b_mn
    (4096, 8192, 8192, False, False),      # wide MLP forward
    (4096, 8192, 8192, False, True),       # dgrad
    (8192, 8192, 4096, True, True),        # wgrad
    (8192, 8192, 8192, True, False),       # the fourth layout at the 8192 class
    (4096, 6400, 1600, False, False),      # GPT-2 XL fc forward
    (4096, 1600, 6400, False, True),       # GPT-2 XL fc dgrad
    (6400, 1600, 4096, True, True),        # GPT-2 XL fc wgrad
    (8192, 2304, 768, False, False),       # GPT-2 small qkv forward
    (8192, 50304, 768, False, False),      # LM head forward
    (8192, 768, 50304, False, True),       # LM head dgrad
]


@pytest.mark.parametrize("m,n,k,a_mn,b_mn", NORTH_STAR)
def test_gemm_north_star_shapes(m, n, k, a_mn, b_mn):
    a = _bf16(*((k, m) if a_mn else (m, k)), seed=21)
    b = _bf16(*((k, n) if b_mn else (n, k)), scale=1.0 / math.sqrt(k), seed=22)
    out = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    native.gemm(a, b, out, m=m, n=n, k=k, a_mn=a_mn, b_mn=b_mn)
    ref = _gemm_ref_bf16_inputs(a, b, a_mn, b_mn)
    torch.cuda.synchronize()
    # outputs are O(1); bf16 rounding of the result dominates (2^-8 relative), accumulation is fp32 in TMEM
    assert torch.allclose(out.float(), ref, atol=2e-2, rtol=1e-2), (out.float() - ref).abs().max().item()
    del ref


def test_gemm_bias_gelu_epilogue_and_preactivation():
    m, n, k = 512, 1024, 512
    x, w = _bf16(m, k, seed=3), _bf16(n, k, scale=0.05, seed=4)
    bias = _bf16(n, seed=5)
    h = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    z = torch.empty_like(h)
    native.gemm(x, w, h, m=m, n=n, k=k, bias=bias, act=1, d2=z)
    zr = x.float() @ w.float().t() + bias.float()
    assert torch.allclose(z.float(), zr, atol=0.06, rtol=2e-2)
    assert torch.allclose(h.float(), F.gelu(zr, approximate="tanh"), atol=0.06, rtol=2e-2)


def test_gemm_dgelu_and_axpby_epilogues_fp32_out():
    m, n, k = 256, 512, 384
    dy, w = _bf16(m, k, seed=6), _bf16(k, n, scale=0.05, seed=7)       # dgrad: B stored [K,N]
    zprev = _bf16(m, n, seed=8)
    out = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    native.gemm(dy, w, out, m=m, n=n, k=k, b_mn=True, aux=zprev, act=2)
    zf = zprev.float().requires_grad_(True)
    (F.gelu(zf, approximate="tanh")).backward(dy.float() @ w.float())
    assert torch.allclose(out.float(), zf.grad, atol=0.06, rtol=2e-2)
    # accumulate into an fp32 buffer: D = 0.5*acc + 2*aux
    acc = torch.randn(m, n, device=DEV)
    res = torch.empty_like(acc)
    native.gemm(dy, w, res, m=m, n=n, k=k, b_mn=True, aux=acc, alpha=0.5, beta=2.0)
    assert torch.allclose(res, 0.5 * (dy.float() @ w.float()) + 2 * acc, atol=0.05, rtol=1e-2)


@pytest.mark.parametrize("m,n,k,cta_pair", [(512, 1024, 4096, 0), (300, 520, 2048, 0), (4096, 2048, 4096, 2),
                                              (640, 768, 256, 1)])
def test_gemm_colsum_in_the_epilogue(m, n, k, cta_pair):
    """dgrad * gelu'(z) with the bias gradient (column sums of the bf16 output) added from the epilogue."""
    dy, w = _bf16(m, k, seed=31), _bf16(k, n, scale=1.0 / math.sqrt(k), seed=32)
    zprev = _bf16(m, n, seed=33)
    out = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    cs = torch.zeros(n, device=DEV)
    for rep in range(2):                       # the second call must ADD to the first
        native.gemm(dy, w, out, m=m, n=n, k=k, b_mn=True, aux=zprev, act=2, colsum=cs, cta_pair=cta_pair)
    ref_out = torch.empty_like(out)
    native.gemm(dy, w, ref_out, m=m, n=n, k=k, b_mn=True, aux=zprev, act=2, cta_pair=cta_pair, split_k=1)
    assert torch.equal(out, ref_out)          # same kernel, same tile walk: the option must not change D
    want = 2 * out.float().sum(0)
    assert torch.allclose(cs, want, rtol=1e-4, atol=1e-3 * math.sqrt(m)), (cs - want).abs().max().item()
    grad = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    native.take_f32_to_bf16(cs, grad)
    assert torch.allclose(grad.float(), want, rtol=2 ** -7, atol=1e-2) and float(cs.abs().max()) == 0.0


def test_gemm_batched_strided_heads():
    # attention-style: per (batch, head) problem carved out of a packed [B, T, 3C] tensor
    B, T, H, D = 2, 256, 3, 64
    C3 = 3 * H * D
    qkv = _bf16(B, T, C3, seed=9)
    s = torch.empty(B, H, T, T, dtype=torch.bfloat16, device=DEV)
    q, kk = qkv, qkv[:, :, H * D:]
    native.gemm(q, kk, s, m=T, n=T, k=D, lda=C3, ldb=C3, ldd=T, batch=(H, B), a_strides=(D, T * C3),
                b_strides=(D, T * C3), d_strides=(T * T, H * T * T), alpha=0.125)
    qh = qkv[:, :, :H * D].view(B, T, H, D).permute(0, 2, 1, 3).float()
    kh = qkv[:, :, H * D:2 * H * D].view(B, T, H, D).permute(0, 2, 1, 3).float()
    ref = 0.125 * qh @ kh.transpose(-1, -2)
    assert torch.allclose(s.float(), ref, atol=0.1, rtol=2e-2)


def test_gemm_vocab_edge_tile():
    m, n, k = 192, 50304 // 8, 128         # N not a multiple of the tile, M not a multiple of 128
    a, b = _bf16(m, k, seed=10), _bf16(n, k, seed=11)
    out = torch.full((m, n), float("nan"), dtype=torch.bfloat16, device=DEV)
    native.gemm(a, b, out, m=m, n=n, k=k)
    assert torch.allclose(out.float(), a.float() @ b.float().t(), atol=0.2, rtol=2e-2)


def test_linear_small_matches_torch():
    x = torch.randn(32, 20, device=DEV)
    lin = torch.nn.Linear(20, 1).to(DEV)
    y = torch.empty(32, 1, device=DEV)
    native.linear_small_fwd(x, lin.weight.detach(), lin.bias.detach(), y)
    assert torch.allclose(y, lin(x), atol=1e-5)
    dy = torch.randn(32, 1, device=DEV)
    dx, dw, db = torch.empty_like(x), torch.empty(1, 20, device=DEV), torch.empty(1, device=DEV)
    native.linear_small_bwd(x, lin.weight.detach(), dy, dx, dw, db)
    xr = x.clone().requires_grad_(True)
    lin(xr).backward(dy)
    assert torch.allclose(dx, xr.grad, atol=1e-5) and torch.allclose(dw, lin.weight.grad, atol=1e-4)
    assert torch.allclose(db, lin.bias.grad, atol=1e-4)


# ------------------------------------------------------------------- elementwise
@pytest.mark.parametrize("n", [8, 1000, 4096 * 33 + 5])
def test_gelu_add_cast(n):
    x, dy = _bf16(n, seed=1), _bf16(n, seed=2)
    y = torch.empty_like(x)
    native.gelu_fwd(x, y)
    assert torch.allclose(y.float(), F.gelu(x.float(), approximate="tanh"), atol=2e-2, rtol=2e-2)
    dx = torch.empty_like(x)
    native.gelu_bwd(x, dy, dx)
    xr = x.float().requires_grad_(True)
    F.gelu(xr, approximate="tanh").backward(dy.float())
    assert torch.allclose(dx.float(), xr.grad, atol=3e-2, rtol=2e-2)
    s = torch.empty_like(x)
    native.add(x, dy, s)
    assert torch.equal(s, (x.float() + dy.float()).to(torch.bfloat16))
    f = torch.randn(n, device=DEV)
    b = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    native.cast_f32_to_bf16(f, b)
    assert torch.equal(b, f.to(torch.bfloat16))
    back = torch.empty(n, device=DEV)
    native.cast_bf16_to_f32(b, back)
    assert torch.equal(back, b.float())


@pytest.mark.parametrize("rows,cols", [(4096, 8192), (100, 72), (33, 1600)])
def test_colsum_and_sumsq_and_mean(rows, cols):
    x = _bf16(rows, cols, seed=3)
    out = torch.empty(cols, dtype=torch.float32, device=DEV)
    native.colsum(x, out)
    assert torch.allclose(out, x.float().sum(0), atol=1e-2 * math.sqrt(rows), rtol=1e-3)
    outb = torch.empty(cols, dtype=torch.bfloat16, device=DEV)
    native.colsum(x, outb)
    assert torch.allclose(outb.float(), x.float().sum(0), atol=0.5, rtol=2e-2)
    acc = torch.zeros((), device=DEV)
    native.sumsq(x, acc)
    assert torch.allclose(acc, (x.float() ** 2).sum(), rtol=1e-3)
    m = torch.empty((), device=DEV)
    native.mean_f32(out, m)
    assert torch.allclose(m, out.mean(), rtol=1e-4, atol=1e-4)


# --------------------------------------------------------------------- layernorm
@pytest.mark.parametrize("rows,cols", [(2048, 768), (512, 1600), (77, 1024), (64, 4096)])
@pytest.mark.parametrize("with_res", [False, True])
def test_layernorm_fwd(rows, cols, with_res):
    x, r = _bf16(rows, cols, seed=1), _bf16(rows, cols, seed=2)
    g, b = _bf16(cols, seed=3) * 0.1 + 1, _bf16(cols, seed=4) * 0.1
    y, xres = torch.empty_like(x), torch.empty_like(x)
    mean, rstd = torch.empty(rows, device=DEV), torch.empty(rows, device=DEV)
    native.layernorm_fwd(x, r if with_res else None, g, b, y, xres if with_res else None, mean, rstd)
    xin = (x.float() + r.float()).to(torch.bfloat16).float() if with_res else x.float()
    ref = F.layer_norm(xin, (cols,), g.float(), b.float(), 1e-5)
    assert torch.allclose(y.float(), ref, atol=3e-2, rtol=2e-2)
    assert torch.allclose(mean, xin.mean(-1), atol=1e-3)
    assert torch.allclose(rstd, 1 / torch.sqrt(xin.var(-1, unbiased=False) + 1e-5), rtol=1e-3)
    if with_res:
        assert torch.equal(xres, xin.to(torch.bfloat16))


@pytest.mark.parametrize("rows,cols", [(2048, 768), (300, 1600), (8192, 768), (4096, 1600), (513, 1024), (64, 2048),
                                       (100, 72), (3, 512)])
def test_layernorm_bwd(rows, cols):
    x, dy, dres = _bf16(rows, cols, seed=1), _bf16(rows, cols, seed=2), _bf16(rows, cols, seed=5)
    g, b = _bf16(cols, seed=3) * 0.1 + 1, _bf16(cols, seed=4) * 0.1
    y = torch.empty_like(x)
    mean, rstd = torch.empty(rows, device=DEV), torch.empty(rows, device=DEV)
    native.layernorm_fwd(x, None, g, b, y, None, mean, rstd)
    dx = torch.empty_like(x)
    dgamma, dbeta = torch.empty(cols, device=DEV), torch.empty(cols, device=DEV)
    scratch = torch.empty(native.layernorm_bwd_scratch_floats(rows, cols), device=DEV)
    native.layernorm_bwd(dy, x, g, mean, rstd, dres, dx, dgamma, dbeta, scratch)
    xr, gr, br = x.float().requires_grad_(True), g.float().requires_grad_(True), b.float().requires_grad_(True)
    F.layer_norm(xr, (cols,), gr, br, 1e-5).backward(dy.float())
    assert torch.allclose(dx.float(), xr.grad + dres.float(), atol=6e-2, rtol=3e-2)
    assert torch.allclose(dgamma, gr.grad, atol=2e-2 * math.sqrt(rows), rtol=1e-2)
    assert torch.allclose(dbeta, br.grad, atol=2e-2 * math.sqrt(rows), rtol=1e-2)


def test_layernorm_bwd_bf16_param_grads_match_fp32_path():
    rows, cols = 1024, 768
    x, dy = _bf16(rows, cols, seed=1), _bf16(rows, cols, seed=2)
    g, b = _bf16(cols, seed=3) * 0.1 + 1, _bf16(cols, seed=4) * 0.1
    y = torch.empty_like(x)
    mean, rstd = torch.empty(rows, device=DEV), torch.empty(rows, device=DEV)
    native.layernorm_fwd(x, None, g, b, y, None, mean, rstd)
    scratch = torch.empty(native.layernorm_bwd_scratch_floats(rows, cols), device=DEV)
    dx32, dg32, db32 = torch.empty_like(x), torch.empty(cols, device=DEV), torch.empty(cols, device=DEV)
    native.layernorm_bwd(dy, x, g, mean, rstd, None, dx32, dg32, db32, scratch)
    dx16 = torch.empty_like(x)
    dg16 = torch.empty(cols, device=DEV, dtype=torch.bfloat16)
    db16 = torch.empty(cols, device=DEV, dtype=torch.bfloat16)
    native.layernorm_bwd_bf16grads(dy, x, g, mean, rstd, None, dx16, dg16, db16, scratch)
    assert torch.equal(dx16, dx32)
    assert torch.equal(dg16, dg32.to(torch.bfloat16)) and torch.equal(db16, db32.to(torch.bfloat16))
    native.layernorm_bwd_bf16grads(dy, x, g, mean, rstd, None, dx16, dg16, db16, scratch, accumulate=True)
    assert torch.allclose(dg16.float(), 2 * dg32, rtol=2e-2, atol=1e-2)


# ------------------------------------------------------------------------ losses
@pytest.mark.parametrize("rows,vocab,ld", [(256, 8192, 8192), (64, 50257, 50304), (33, 1000, 1000)])
def test_cross_entropy_fused(rows, vocab, ld):
    logits = _bf16(rows, ld, scale=2.0, seed=1)
    targets = torch.randint(0, vocab, (rows,), device=DEV)
    logits[0, targets[0]] = 14.0          # a confident row: p_t -> 1, the gradient p_t - 1 must not lose its digits
    logits[1, targets[1]] = 9.0
    targets[2] = -100                      # ignore_index: zero loss, zero gradient
    ref_in = logits[:, :vocab].float().requires_grad_(True)
    ref_loss = F.cross_entropy(ref_in, targets, reduction="none")
    (ref_loss.sum() / rows).backward()
    row_loss = torch.empty(rows, device=DEV)
    work = logits.clone()
    native.cross_entropy_fwd_bwd(work, targets, row_loss, vocab, 1.0 / rows)
    assert torch.allclose(row_loss, ref_loss, atol=2e-3, rtol=1e-3)
    assert torch.allclose(work[:, :vocab].float(), ref_in.grad, atol=2e-3 / rows + 1e-5, rtol=2e-2)
    for r in (0, 1):                       # relative accuracy of the target column itself
        got, want = float(work[r, targets[r]]), float(ref_in.grad[r, targets[r]])
        assert abs(got - want) <= 1e-2 * abs(want) + 1e-9, (r, got, want)
    assert float(work[2].float().abs().max()) == 0.0 and float(row_loss[2]) == 0.0
    if ld > vocab:
        assert float(work[:, vocab:].float().abs().max()) == 0.0
    # loss only (no gradient): the logits must be left untouched
    work2 = logits.clone()
    native.cross_entropy_fwd_bwd(work2, targets, row_loss, vocab, 1.0 / rows, write_grad=False)
    assert torch.equal(work2, logits) and torch.allclose(row_loss, ref_loss, atol=2e-3, rtol=1e-3)


def test_soft_cross_entropy_and_mse():
    x = torch.randn(32, 1, device=DEV)
    t = torch.rand(32, 1, device=DEV)
    row, d = torch.empty(32, device=DEV), torch.empty_like(x)
    native.soft_cross_entropy_fwd_bwd(x, t, row, d, 1.0 / 32)
    assert float(row.abs().max()) == 0.0 and float(d.abs().max()) == 0.0       # reference quirk Q1
    x = torch.randn(16, 10, device=DEV)
    t = torch.softmax(torch.randn(16, 10, device=DEV), -1)
    row, d = torch.empty(16, device=DEV), torch.empty_like(x)
    native.soft_cross_entropy_fwd_bwd(x, t, row, d, 1.0 / 16)
    xr = x.clone().requires_grad_(True)
    F.cross_entropy(xr, t).backward()
    assert torch.allclose(row.mean(), F.cross_entropy(x, t), atol=1e-5)
    assert torch.allclose(d, xr.grad, atol=1e-6)
    p, tt = torch.randn(1000, device=DEV), torch.randn(1000, device=DEV)
    s, dp = torch.empty((), device=DEV), torch.empty_like(p)
    native.mse_fwd_bwd(p, tt, s, dp, 1.0 / 1000)
    assert torch.allclose(s / 1000, F.mse_loss(p, tt), rtol=1e-5)
    assert torch.allclose(dp, 2 * (p - tt) / 1000, atol=1e-7)


# --------------------------------------------------------------------- embedding
def test_embedding_fwd_bwd():
    B, T, C, V = 4, 64, 768, 1000
    wte, wpe = _bf16(V, C, seed=1), _bf16(T, C, seed=2)
    tok = torch.randint(0, V, (B, T), device=DEV)
    out = torch.empty(B, T, C, dtype=torch.bfloat16, device=DEV)
    native.embedding_fwd(tok, wte, wpe, out)
    assert torch.equal(out, (wte[tok] + wpe[None]).to(torch.bfloat16))
    dout = _bf16(B, T, C, seed=3)
    dwte, dwpe = torch.zeros_like(wte), torch.empty_like(wpe)
    native.embedding_bwd(dout, tok, dwte, dwpe)
    ref = torch.zeros(V, C, device=DEV).index_add_(0, tok.view(-1), dout.float().view(-1, C))
    assert torch.allclose(dwte.float(), ref, atol=0.15, rtol=3e-2)
    assert torch.allclose(dwpe.float(), dout.float().sum(0), atol=0.05, rtol=2e-2)


# -------------------------------------------------------------------- optimizers
@pytest.mark.parametrize("pdtype,gdtype,master", [(torch.float32, torch.float32, False),
                                                   (torch.bfloat16, torch.bfloat16, False),
                                                   (torch.bfloat16, torch.bfloat16, True)])
def test_fused_sgd_and_adamw(pdtype, gdtype, master):
    from distributed_training_b200.optim import FusedAdamW, FusedSGD

    shapes = [(1000,), (33, 7), (512, 512), (5,)]
    for kind in ("sgd", "adamw"):
        torch.manual_seed(0)
        ps = [torch.randn(s, device=DEV).to(pdtype).requires_grad_(True) for s in shapes]
        ref = [p.detach().float().clone().requires_grad_(True) for p in ps]
        if kind == "sgd":
            opt = FusedSGD(ps, lr=0.1, momentum=0.9, weight_decay=0.01, master_weights=master)
            ropt = torch.optim.SGD(ref, lr=0.1, momentum=0.9, weight_decay=0.01)
        else:
            opt = FusedAdamW(ps, lr=1e-2, weight_decay=0.1, master_weights=master)
            ropt = torch.optim.AdamW(ref, lr=1e-2, weight_decay=0.1)
        for step in range(3):
            for p, r in zip(ps, ref):
                g = torch.randn(p.shape, device=DEV)
                p.grad = g.to(gdtype)
                r.grad = p.grad.float()
            opt.step()
            ropt.step()
        tol = 1e-5 if pdtype == torch.float32 else (2e-2 if not master else 1e-2)
        for p, r in zip(ps, ref):
            assert torch.allclose(p.detach().float(), r.detach(), atol=tol, rtol=tol), (kind, pdtype, master)
        if master and pdtype != torch.float32:
            for p, r in zip(ps, ref):
                assert torch.allclose(opt.state[p]["master"], r.detach(), atol=1e-5, rtol=1e-5)


# ----------------------------------------------------------------- model-level
def test_native_mlp_step_matches_autograd():
    from distributed_training_b200.models.mlp import MLP

    torch.manual_seed(0)
    model = MLP(256, 512, 384, num_layers=3, dtype=torch.bfloat16).to(DEV)
    assert model.native_ready()
    x = _bf16(192, 256, seed=1)
    y = torch.randint(0, 384, (192,), device=DEV)
    loss = model.forward_backward(x, y)
    native_grads = [p.grad.float().clone() for p in model.parameters()]
    for p in model.parameters():
        p.grad = None
    ref = MLP(256, 512, 384, num_layers=3, dtype=torch.float32).to(DEV)
    ref.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
    ref_loss = F.cross_entropy(ref(x.float()), y)
    ref_loss.backward()
    assert abs(float(loss) - float(ref_loss.detach())) < 3e-2
    for g, p in zip(native_grads, ref.parameters()):
        scale = p.grad.abs().max().item() + 1e-6
        assert (g - p.grad).abs().max().item() <= 0.06 * scale + 1e-4


def test_native_mlp_two_stream_backward_matches_one_stream(monkeypatch):
    """Weight gradients on the side stream (tails of consecutive GEMMs overlap) must give what the one-stream order
    gives: bit-identical weight gradients and loss, bias gradients equal up to the order of the fp32 reductions."""
    from distributed_training_b200.models import mlp as mlp_mod

    torch.manual_seed(3)
    model = mlp_mod.MLP(2048, 2048, 2048, num_layers=5, dtype=torch.bfloat16).to(DEV)
    x = _bf16(2048, 2048, seed=41)
    y = torch.randint(0, 2048, (2048,), device=DEV)
    results = []
    for flag in (False, True, True):
        monkeypatch.setattr(mlp_mod, "_ASYNC_WGRAD", flag)
        for p in model.parameters():
            p.grad = None
        loss = model.forward_backward(x, y)
        torch.cuda.synchronize()
        results.append((float(loss), [p.grad.clone() for p in model.parameters()]))
    for loss, grads in results[1:]:
        assert loss == results[0][0]
        for p, g, g0 in zip(model.parameters(), grads, results[0][1]):
            if p.dim() == 2:
                assert torch.equal(g, g0)
            else:
                assert torch.allclose(g.float(), g0.float(), rtol=2 ** -6, atol=1e-3)


def test_launch_counter_counts():
    native.reset_launch_count()
    x = _bf16(1024, seed=1)
    native.gelu_fwd(x, torch.empty_like(x))
    native.gelu_fwd(x, torch.empty_like(x))
    assert native.launch_count() == 2


# --------------------------------------------------------------------- attention
def _attn_ref(qkv, heads):
    B, T, C3 = qkv.shape
    C = C3 // 3
    q, k, v = qkv.float().view(B, T, 3, heads, C // heads).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(C // heads)
    mask = torch.ones(T, T, device=qkv.device, dtype=torch.bool).tril()
    s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, -1)
    o = torch.softmax(s, -1) @ v
    return o.transpose(1, 2).reshape(B, T, C), lse


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (2, 256, 3), (2, 1024, 12), (4, 1024, 25)])
def test_attention_fwd(B, T, H):
    qkv = _bf16(B, T, 3 * H * 64, seed=1)
    out = torch.empty(B, T, H * 64, dtype=torch.bfloat16, device=DEV)
    lse = torch.empty(B, H, T, device=DEV)
    native.attention_fwd(qkv, out, lse, H)
    ref, ref_lse = _attn_ref(qkv, H)
    assert torch.allclose(lse, ref_lse, atol=2e-2, rtol=1e-3)
    assert torch.allclose(out.float(), ref, atol=3e-2, rtol=3e-2)


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (2, 384, 2), (1, 1024, 4), (4, 1024, 25), (8, 1024, 12)])
def test_attention_bwd(B, T, H):
    qkv = _bf16(B, T, 3 * H * 64, seed=2)
    dout = _bf16(B, T, H * 64, seed=3)
    out = torch.empty(B, T, H * 64, dtype=torch.bfloat16, device=DEV)
    lse = torch.empty(B, H, T, device=DEV)
    native.attention_fwd(qkv, out, lse, H)
    dqkv = torch.empty_like(qkv)
    scratch = torch.empty(native.attention_bwd_scratch_floats(B, T, H), device=DEV)
    native.attention_bwd(qkv, out, dout, lse, dqkv, scratch, H)
    qr = qkv.float().requires_grad_(True)
    ref, _ = _attn_ref(qr, H)
    ref.backward(dout.float())
    scale = qr.grad.abs().max().item()
    err = (dqkv.float() - qr.grad).abs().max().item()
    assert err <= 0.03 * scale + 2e-2, f"max err {err} (grad scale {scale})"


def test_single_gpu_fused_sgd_in_wgrad_epilogue_matches_reference():
    """World size 1: the wgrad GEMM epilogue applies the SGD update; weights must track an fp32 autograd run."""
    from distributed_training_b200.models.mlp import MLP
    from distributed_training_b200.parallel.ddp import DDPStrategy
    from distributed_training_b200.parallel.strategy import OptimizerSpec

    torch.manual_seed(0)
    model = MLP(256, 512, 384, num_layers=3, dtype=torch.bfloat16).to(DEV)
    ref = MLP(256, 512, 384, num_layers=3, dtype=torch.float32).to(DEV)
    ref.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
    strategy = DDPStrategy(0, True)
    wrapped = strategy.prepare_model(model, 0)
    opt = strategy.build_optimizer(wrapped, OptimizerSpec("sgd", lr=0.05, weight_decay=0.01))
    wrapped.attach_optimizer(opt)
    assert wrapped._fused_sgd_hyper() is not None and wrapped.comm.world_size == 1
    ropt = torch.optim.SGD(ref.parameters(), lr=0.05, weight_decay=0.01)
    for step in range(4):
        x, y = _bf16(192, 256, seed=step), torch.randint(0, 384, (192,), device=DEV)
        wrapped.forward_backward(x, y)
        opt.step()
        ropt.zero_grad()
        F.cross_entropy(ref(x.float()), y).backward()
        ropt.step()
    for (k, v), (_, rv) in zip(model.state_dict().items(), ref.state_dict().items()):
        assert torch.allclose(v.float(), rv, atol=3e-2, rtol=3e-2), (k, (v.float() - rv).abs().max())
    strategy.comm.close()


def test_native_gpt2_step_matches_autograd():
    from distributed_training_b200.models.gpt2 import GPT2, GPT2Config

    torch.manual_seed(0)
    cfg = GPT2Config(vocab_size=1000, n_positions=256, n_embd=128, n_layer=2, n_head=2)
    model = GPT2(cfg, dtype=torch.bfloat16).to(DEV)
    assert model.native_ready()
    ref = GPT2(cfg, dtype=torch.float32).to(DEV)
    ref.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
    tok = torch.randint(0, 1000, (2, 256), device=DEV)
    tgt = torch.randint(0, 1000, (2, 256), device=DEV)
    loss = model.forward_backward(tok, tgt)
    ref_loss = ref.loss(tok, tgt)
    ref_loss.backward()
    assert abs(float(loss) - float(ref_loss.detach())) < 5e-2, (float(loss), float(ref_loss))
    bad = []
    for (name, p), (_, rp) in zip(model.named_parameters(), ref.named_parameters()):
        scale = rp.grad.abs().max().item() + 1e-6
        err = (p.grad.float() - rp.grad).abs().max().item()
        if err > 0.08 * scale + 2e-4:
            bad.append((name, err, scale))
    assert not bad, bad[:6]


@pytest.mark.parametrize("m,n,k,cta_pair,tile_n", [(512, 1024, 512, 1, 128), (1024, 2048, 768, 2, 256),
                                                    (384, 1600, 256, 1, 256)])
def test_allgather_gemm_single_rank(m, n, k, cta_pair, tile_n):
    """World size 1: the gather warps copy the (only) shard into the full matrix inside the GEMM kernel."""
    x, shard = _bf16(m, k, seed=1), _bf16(n, k, scale=0.05, seed=2)
    bias = _bf16(n, seed=3)
    full = torch.full((n, k), float("nan"), dtype=torch.bfloat16, device=DEV)
    out = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    flags = torch.ones(native.gemm_allgather_flag_ints(n), dtype=torch.int32, device=DEV) * 7   # call must zero them
    native.gemm_allgather(x, full, out, [shard.data_ptr()], flags, 0, n, m=m, n=n, k=k, bias=bias,
                          cta_pair=cta_pair, tile_n=tile_n)
    torch.cuda.synchronize()
    assert torch.equal(full, shard)
    ref = x.float() @ shard.float().t() + bias.float()
    assert torch.allclose(out.float(), ref, atol=0.08, rtol=2e-2)


def test_functional_autograd_wrappers_match_torch():
    from distributed_training_b200.ops import functional as NF

    # bf16 Linear + GELU + LayerNorm + CE through autograd
    torch.manual_seed(0)
    x = _bf16(96, 256, seed=1).requires_grad_(True)
    lin = NF.NativeLinear(256, 512).to(DEV).to(torch.bfloat16)
    ln = torch.nn.LayerNorm(512).to(DEV).to(torch.bfloat16)
    tgt = torch.randint(0, 512, (96,), device=DEV)
    loss = NF.cross_entropy(NF.layer_norm(NF.gelu(lin(x)), ln.weight, ln.bias), tgt)
    loss.backward()
    xr = x.detach().float().requires_grad_(True)
    w, b = lin.weight.detach().float().requires_grad_(True), lin.bias.detach().float().requires_grad_(True)
    g, be = ln.weight.detach().float().requires_grad_(True), ln.bias.detach().float().requires_grad_(True)
    ref = F.cross_entropy(F.layer_norm(F.gelu(F.linear(xr, w, b), approximate="tanh"), (512,), g, be), tgt)
    ref.backward()
    assert abs(float(loss) - float(ref.detach())) < 3e-2
    for got, want in ((x.grad, xr.grad), (lin.weight.grad, w.grad), (lin.bias.grad, b.grad), (ln.weight.grad, g.grad)):
        scale = want.abs().max().item() + 1e-6
        assert (got.float() - want).abs().max().item() <= 0.08 * scale + 1e-4
    # the reference's own toy model: fp32 Linear(20, 1) + MSE
    toy = NF.NativeLinear(20, 1).to(DEV)
    xs, ys = torch.randn(32, 20, device=DEV), torch.randn(32, 1, device=DEV)
    l1 = NF.mse_loss(toy(xs), ys)
    l1.backward()
    wr, br = toy.weight.detach().clone().requires_grad_(True), toy.bias.detach().clone().requires_grad_(True)
    l2 = F.mse_loss(F.linear(xs, wr, br), ys)
    l2.backward()
    assert torch.allclose(l1, l2, atol=1e-6)
    assert torch.allclose(toy.weight.grad, wr.grad, atol=1e-5) and torch.allclose(toy.bias.grad, br.grad, atol=1e-5)
    # soft-label CE, the reference trainer's objective (identically zero for one logit)
    l3 = NF.cross_entropy(toy(xs), torch.rand(32, 1, device=DEV))
    assert float(l3) == 0.0


@pytest.mark.parametrize("split", [0, 2, 5])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, True)])
def test_gemm_split_k(split, a_mn, b_mn):
    """Few output tiles, long K (transformer wgrad shape): partial sums through the workspace + reduce kernel."""
    m, n, k = 768, 512, 4096
    a = _bf16(*((k, m) if a_mn else (m, k)), scale=0.5, seed=1)
    b = _bf16(*((k, n) if b_mn else (n, k)), scale=0.5, seed=2)
    bias, res = _bf16(n, seed=3), _bf16(m, n, seed=4)
    out = torch.empty(m, n, dtype=torch.bfloat16, device=DEV)
    native.gemm(a, b, out, m=m, n=n, k=k, a_mn=a_mn, b_mn=b_mn, bias=bias, aux=res, beta=0.5, alpha=0.25, split_k=split)
    ref = 0.25 * _gemm_ref(a, b, a_mn, b_mn) + bias.float() + 0.5 * res.float()
    assert torch.allclose(out.float(), ref, atol=0.15, rtol=2e-2), (out.float() - ref).abs().max()
    # split 1 (disabled) must agree with the split result to bf16 rounding
    one = torch.empty_like(out)
    native.gemm(a, b, one, m=m, n=n, k=k, a_mn=a_mn, b_mn=b_mn, bias=bias, aux=res, beta=0.5, alpha=0.25, split_k=1)
    assert torch.allclose(out.float(), one.float(), atol=0.13, rtol=2e-2)


# ------------------------------------------------------------- loss curves
def test_loss_curve_wide_mlp_north_star_shape_matches_torch_baseline():
    """BASELINE.json config 2 at full size (8 x 8192, batch 4096, bf16, SGD): the native step and the plain-torch
    model of the reference arm start from the same weights, see the same batches and must trace the same loss curve."""
    from baseline.models import wide_mlp_torch
    from distributed_training_b200.models.mlp import wide_mlp
    from distributed_training_b200.optim import FusedSGD

    torch.manual_seed(0)
    ours = wide_mlp().to(DEV)
    ref = wide_mlp_torch().to(DEV)
    linears = [m for m in ref if isinstance(m, torch.nn.Linear)]
    with torch.no_grad():
        for mine, theirs in zip(ours.layers, linears):
            theirs.weight.copy_(mine.weight)
            theirs.bias.copy_(mine.bias)
    lr = 0.05
    opt = FusedSGD(ours.parameters(), lr=lr)
    ropt = torch.optim.SGD(ref.parameters(), lr=lr)
    g = torch.Generator().manual_seed(3)
    curve, rcurve = [], []
    for step in range(6):
        x = torch.randn(4096, 8192, generator=g).to(torch.bfloat16).to(DEV)
        y = torch.randint(0, 8192, (4096,), generator=g).to(DEV)
        curve.append(float(ours.forward_backward(x, y)))
        opt.step()
        ropt.zero_grad(set_to_none=True)
        rl = F.cross_entropy(ref(x).float(), y)
        rl.backward()
        ropt.step()
        rcurve.append(float(rl))
    assert all(abs(a - b) < 2e-2 for a, b in zip(curve, rcurve)), (curve, rcurve)
    # the weights moved (bf16 SGD at this lr) and moved the same way
    w0, r0 = ours.layers[-1].weight.float(), linears[-1].weight.float()
    assert float((w0 - r0).abs().max()) < 2e-2 * float(r0.abs().max()) + 1e-4


def test_loss_curve_gpt2_native_matches_eager_fp32():
    """A 4-layer GPT-2 of GPT-2 small's width at seq 1024: native step + fused AdamW (fp32 master weights) against
    eager fp32 autograd + torch AdamW over the same batches."""
    from distributed_training_b200.models.gpt2 import GPT2, GPT2Config
    from distributed_training_b200.optim import FusedAdamW

    torch.manual_seed(0)
    cfg = GPT2Config(n_layer=4)
    model = GPT2(cfg, dtype=torch.bfloat16).to(DEV)
    ref = GPT2(cfg, dtype=torch.float32).to(DEV)
    ref.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
    opt = FusedAdamW(model.parameters(), lr=3e-4, weight_decay=0.1, master_weights=True)
    ropt = torch.optim.AdamW(ref.parameters(), lr=3e-4, weight_decay=0.1)
    g = torch.Generator().manual_seed(11)
    batches = [torch.randint(0, cfg.vocab_size, (4, 1025), generator=g).to(DEV) for _ in range(2)]
    curve, rcurve = [], []
    for step in range(6):
        t = batches[step % 2]
        tok, tgt = t[:, :-1].contiguous(), t[:, 1:].contiguous()
        curve.append(float(model.forward_backward(tok, tgt)))
        opt.step()
        ropt.zero_grad(set_to_none=True)
        rl = ref.loss(tok, tgt)
        rl.backward()
        ropt.step()
        rcurve.append(float(rl))
    assert rcurve[-1] < rcurve[0] - 0.05, rcurve                     # the reference is learning (memorising 2 batches)
    assert all(abs(a - b) < 4e-2 for a, b in zip(curve, rcurve)), (curve, rcurve)
