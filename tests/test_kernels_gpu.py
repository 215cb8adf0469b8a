"""Numerics of every sm_100a kernel against a plain PyTorch fp32 expression of the same op (GPU only)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["2cta", "1cta"])
def K(request):
    from scaling_b200.ops import _lib, kernels

    assert _lib.available(), "native library must load on a GPU box"
    kernels.set_gemm_impl(request.param)
    yield kernels
    kernels.set_gemm_impl("2cta")


def _rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.mark.parametrize(
    "M,N,K_",
    [(128, 256, 64), (256, 512, 128), (4096, 4096, 4096), (1000, 776, 520), (4096, 6144, 4096), (72, 40, 136), (8, 8, 8)],
)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_layouts(K, M, N, K_, a_mn, b_mn):
    torch.manual_seed(0)
    dev = "cuda"
    a = torch.randn((K_, M) if a_mn else (M, K_), device=dev, dtype=torch.bfloat16)
    b = torch.randn((K_, N) if b_mn else (N, K_), device=dev, dtype=torch.bfloat16)
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    ref = A @ B
    out = K.gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    assert out.shape == (M, N)
    assert _rel_err(out, ref) < 1e-2, _rel_err(out, ref)
    out32 = K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    assert _rel_err(out32, ref) < 1e-4, _rel_err(out32, ref)


@pytest.mark.parametrize("M,N,K_", [(512, 512, 14336), (4096, 4096, 14336), (1000, 520, 28680), (2048, 1024, 32768)])
def test_gemm_split_k_tail(K, M, N, K_):
    """Shapes whose last wave of 256x256 tiles is partial: the tail tiles are split along K and reduced through the
    fp32 workspace (2-CTA kernel).  Every epilogue must see the full sum, and repeated launches must agree bitwise
    (the part order of the reduction is fixed, the arrival counters reset themselves)."""
    torch.manual_seed(5)
    a = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K_, device="cuda", dtype=torch.bfloat16) / math.sqrt(K_)
    ref = a.float() @ b.float().t()
    out = K.gemm(a, b, out_dtype=torch.float32)
    assert _rel_err(out, ref) < 1e-4, _rel_err(out, ref)
    for _ in range(3):
        assert torch.equal(K.gemm(a, b, out_dtype=torch.float32), out)
    assert _rel_err(K.gemm(a, b), ref) < 1e-2
    bt = b.t().contiguous()
    assert _rel_err(K.gemm(a, bt, b_mn=True, out_dtype=torch.float32), ref) < 1e-4
    at = a.t().contiguous()
    assert _rel_err(K.gemm(at, bt, a_mn=True, b_mn=True, out_dtype=torch.float32), ref) < 1e-4
    acc = torch.randn(M, N, device="cuda", dtype=torch.float32)
    want = acc + ref
    K.gemm(a, b, out=acc, accumulate=True)
    assert _rel_err(acc, want) < 1e-4
    bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    assert _rel_err(K.gemm(a, b, bias=bias), ref + bias.float()) < 1e-2
    if N % 16 == 0:
        F = N // 2
        h, pre = K.gemm_swiglu(a, b)
        g, u = ref[:, :F].bfloat16().float(), ref[:, F:].bfloat16().float()
        assert _rel_err(pre[:, :F], g) < 1e-2 and _rel_err(pre[:, F:], u) < 1e-2
        assert _rel_err(h, torch.nn.functional.silu(g) * u) < 2e-2


def test_gemm_accumulate_and_bias(K):
    torch.manual_seed(1)
    M, N, K_ = 512, 768, 320
    a = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K_, device="cuda", dtype=torch.bfloat16)
    bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    ref = a.float() @ b.float().t()
    acc = torch.randn(M, N, device="cuda", dtype=torch.float32)
    want = acc + ref
    K.gemm(a, b, out=acc, accumulate=True)
    assert _rel_err(acc, want) < 1e-4
    accb = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
    wantb = accb.float() + ref
    K.gemm(a, b, out=accb, accumulate=True)
    assert _rel_err(accb, wantb) < 1e-2
    out = K.gemm(a, b, bias=bias)
    assert _rel_err(out, ref + bias.float()) < 1e-2
    # strided rows (views into a wider buffer), as used for fused q|k|v outputs
    wide = torch.zeros(M, N + 64, device="cuda", dtype=torch.bfloat16)
    K.gemm(a, b, out=wide[:, :N])
    assert _rel_err(wide[:, :N], ref) < 1e-2 and float(wide[:, N:].abs().max()) == 0.0


@pytest.mark.parametrize("M,F,K_", [(256, 512, 256), (4096, 14336, 4096), (300, 136, 72)])
def test_gemm_swiglu(K, M, F, K_):
    torch.manual_seed(2)
    x = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(2 * F, K_, device="cuda", dtype=torch.bfloat16) / math.sqrt(K_)
    h, pre = K.gemm_swiglu(x, w)
    g = (x.float() @ w[:F].float().t()).bfloat16().float()
    u = (x.float() @ w[F:].float().t()).bfloat16().float()
    ref = torch.nn.functional.silu(g) * u
    assert _rel_err(pre[:, :F], g) < 1e-2 and _rel_err(pre[:, F:], u) < 1e-2
    assert _rel_err(h, ref) < 2e-2
    assert _rel_err(K.swiglu_fwd(pre), torch.nn.functional.silu(pre[:, :F].float()) * pre[:, F:].float()) < 1e-2


@pytest.mark.parametrize("rms", [True, False])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("rows,cols", [(64, 256), (1000, 4096), (7, 8192), (33, 12288)])
def test_norm_fwd_bwd(K, rms, dtype, rows, cols):
    torch.manual_seed(3)
    x = torch.randn(rows, cols, device="cuda", dtype=dtype)
    res = torch.randn(rows, cols, device="cuda", dtype=dtype)
    w = (1 + 0.1 * torch.randn(cols, device="cuda")).to(dtype)
    b = None if rms else (0.1 * torch.randn(cols, device="cuda")).to(dtype)
    dy = torch.randn(rows, cols, device="cuda", dtype=dtype)
    eps = 1e-5
    y, summed, mean, rstd = K.norm_fwd(x, w, b, eps, rms, residual=res)
    h = (x + res).float().requires_grad_(True)
    wf = w.float().requires_grad_(True)
    if rms:
        ref = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps) * wf
    else:
        ref = torch.nn.functional.layer_norm(h, (cols,), wf, b.float(), eps)
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    assert _rel_err(summed, x + res) < tol
    assert _rel_err(y, ref) < tol
    ref.backward(dy.float())
    dx, dw, db = K.norm_bwd(dy, summed, w, mean, rstd, rms, has_bias=not rms)
    assert _rel_err(dx, h.grad) < tol
    assert _rel_err(dw, wf.grad) < tol
    if not rms:
        assert _rel_err(db, dy.float().sum(0)) < tol
    # main-grad accumulation in fp32
    main = torch.ones(cols, device="cuda", dtype=torch.float32)
    K.norm_bwd(dy, summed, w, mean, rstd, rms, dw_out=main, accumulate=True, has_bias=False)
    assert _rel_err(main, 1 + wf.grad) < tol


def test_swiglu_bwd(K):
    torch.manual_seed(4)
    rows, F = 512, 1024
    pre = torch.randn(rows, 2 * F, device="cuda", dtype=torch.bfloat16)
    dh = torch.randn(rows, F, device="cuda", dtype=torch.bfloat16)
    p = pre.float().requires_grad_(True)
    h = torch.nn.functional.silu(p[:, :F]) * p[:, F:]
    h.backward(dh.float())
    dpre, hh = K.swiglu_bwd(pre, dh, recompute_h=True)
    assert _rel_err(dpre, p.grad) < 1e-2
    assert _rel_err(hh, h) < 1e-2


@pytest.mark.parametrize("silu", [True, False])
def test_activation(K, silu):
    x = torch.randn(1024, 512, device="cuda", dtype=torch.bfloat16)
    dy = torch.randn_like(x)
    xf = x.float().requires_grad_(True)
    f = torch.nn.functional.silu if silu else torch.nn.functional.gelu
    y = f(xf)
    y.backward(dy.float())
    assert _rel_err(K.act_fwd(x, silu), y) < 1e-2
    assert _rel_err(K.act_bwd(x, dy, silu), xf.grad) < 1e-2


@pytest.mark.parametrize("mode", [0, 1])
def test_rope(K, mode):
    torch.manual_seed(5)
    tokens, heads, hd, seq = 256, 4, 64, 128
    rot = hd if mode == 0 else 32
    x = torch.randn(tokens, heads, hd, device="cuda", dtype=torch.bfloat16)
    inv = 1.0 / (10000 ** (torch.arange(0, rot, 2, device="cuda").float() / rot))
    ang = torch.arange(seq, device="cuda").float()[:, None] * inv[None]
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    pos = (torch.arange(tokens, device="cuda") % seq).int()
    xf = x.float()
    c, s = cos[pos.long()][:, None, :], sin[pos.long()][:, None, :]
    ref = xf.clone()
    if mode == 0:
        a, b = xf[..., 0:rot:2], xf[..., 1:rot:2]
        ref[..., 0:rot:2] = a * c - b * s
        ref[..., 1:rot:2] = a * s + b * c
    else:
        a, b = xf[..., : rot // 2], xf[..., rot // 2 : rot]
        ref[..., : rot // 2] = a * c - b * s
        ref[..., rot // 2 : rot] = a * s + b * c
    y = K.rope_(x.clone(), cos, sin, pos, seq, mode, rot)
    assert _rel_err(y, ref) < 1e-2
    back = K.rope_(y.clone(), cos, sin, pos, seq, mode, rot, backward=True)
    assert _rel_err(back, x) < 2e-2
    y2 = K.rope_(x.clone(), cos, sin, None, seq, mode, rot)
    assert torch.equal(y, y2)


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("rot_dim,d", [(128, 128), (64, 128), (32, 64)])
def test_qkv_rope_split(K, mode, rot_dim, d):
    """Fused split + rotary against the plain rotation formulas, forward and backward, with explicit position ids."""
    from scaling_b200.ops import functional as OF

    torch.manual_seed(4)
    T, nh, nkv, seq = 96, 4, 2, 32
    y = torch.randn(T, (nh + 2 * nkv) * d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    pos = (torch.arange(T, device="cuda") % seq).flip(0).to(torch.int32).contiguous()
    half = rot_dim // 2
    inv = 1.0 / (10000 ** (torch.arange(0, half, device="cuda").float() / half))
    ang = torch.arange(seq, device="cuda").float()[:, None] * inv[None]
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()

    def ref_rot(x):  # x: [T, h, d] fp32
        c, s_ = cos[pos.long()][:, None], sin[pos.long()][:, None]
        r, rest = x[..., :rot_dim], x[..., rot_dim:]
        if mode == 0:
            a, b = r[..., 0::2], r[..., 1::2]
            out = torch.stack((a * c - b * s_, a * s_ + b * c), dim=-1).flatten(-2)
        else:
            a, b = r[..., :half], r[..., half:]
            out = torch.cat((a * c - b * s_, a * s_ + b * c), dim=-1)
        return torch.cat((out, rest), dim=-1)

    q, k, v = OF.qkv_rope_split(y, nh, nkv, d, cos, sin, pos, seq, mode, rot_dim)
    yf = y.detach().float().requires_grad_(True)
    qr = ref_rot(yf[:, :nh * d].view(T, nh, d))
    kr = ref_rot(yf[:, nh * d:(nh + nkv) * d].view(T, nkv, d))
    vr = yf[:, (nh + nkv) * d:].view(T, nkv, d)
    assert _rel_err(q, qr) < 1e-2 and _rel_err(k, kr) < 1e-2 and torch.equal(v.float(), vr.detach())
    gq, gk, gv = torch.randn_like(q), torch.randn_like(k), torch.randn_like(v)
    torch.autograd.backward((q, k, v), (gq, gk, gv))
    torch.autograd.backward((qr, kr, vr), (gq.float(), gk.float(), gv.float()))
    assert _rel_err(y.grad, yf.grad) < 1e-2


def test_bias_dropout_add(K):
    torch.manual_seed(6)
    x = torch.randn(512, 1024, device="cuda", dtype=torch.bfloat16)
    r = torch.randn_like(x)
    bias = torch.randn(1024, device="cuda", dtype=torch.bfloat16)
    out = K.bias_dropout_add(x, bias, r, 0.0, 1, 0)
    assert _rel_err(out, x.float() + bias.float() + r.float()) < 1e-2
    p = 0.25
    o1 = K.bias_dropout_add(x, None, None, p, 7, 3)
    o2 = K.bias_dropout_add(x, None, None, p, 7, 3)
    assert torch.equal(o1, o2)  # counter based: same (seed, offset) -> same mask
    kept = (o1 != 0).float().mean().item()
    assert abs(kept - (1 - p)) < 0.01
    mask = (o1 != 0)
    assert _rel_err(o1[mask], x.float()[mask] / (1 - p)) < 1e-2
    dx = K.dropout_bwd(torch.ones_like(x), p, 7, 3)
    assert torch.equal(dx != 0, mask)
    assert not torch.equal(o1, K.bias_dropout_add(x, None, None, p, 7, 4))


def test_embedding(K):
    torch.manual_seed(7)
    V, H, T = 1000, 256, 512
    w = torch.randn(V, H, device="cuda", dtype=torch.bfloat16)
    ids = torch.randint(0, 2 * V, (T,), device="cuda")
    out = K.embedding_fwd(ids, w, 0, V)
    inrange = ids < V
    ref = torch.zeros(T, H, device="cuda")
    ref[inrange] = w.float()[ids[inrange]]
    assert torch.equal(out.float(), ref)
    dout = torch.randn(T, H, device="cuda", dtype=torch.bfloat16)
    dW = torch.zeros(V, H, device="cuda", dtype=torch.float32)
    K.embedding_bwd(ids, dout, dW, 0, V)
    want = torch.zeros(V, H, device="cuda")
    want.index_add_(0, ids[inrange], dout.float()[inrange])
    assert _rel_err(dW, want) < 1e-5


@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
def test_embedding_backward_is_bit_reproducible_with_heavily_repeated_ids(gdtype):
    """Few distinct ids, hundreds of occurrences each (the regime where an atomicAdd scatter changes bits from run to
    run): the sorted kernel gives the same bits every time, accumulates onto existing gradients, honours the vocabulary
    range and matches the sequential fp32 sum in token order exactly (fp32 arena)."""
    from scaling_b200.ops import kernels as K

    torch.manual_seed(11)
    V, H, T = 64, 136, 4096
    ids = torch.randint(0, 24, (T,), device="cuda")  # ids 16..23 lie outside the local range [0, 16)
    dout = torch.randn(T, H, device="cuda", dtype=torch.bfloat16)
    base = torch.randn(V, H, device="cuda").to(gdtype)
    outs = []
    for _ in range(3):
        dW = base.clone()
        K.embedding_bwd(ids, dout, dW, 0, 16)
        outs.append(dW)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    assert torch.equal(outs[0][16:], base[16:]), "rows outside the range must stay untouched"
    want = torch.zeros(V, H, device="cuda")
    for r in range(16):  # sequential sum in token order, fp32
        rows = dout.float()[ids == r]
        acc = torch.zeros(H, device="cuda")
        for x in rows:
            acc = acc + x
        want[r] = acc
    if gdtype == torch.float32:
        assert torch.equal(outs[0][:16], base[:16] + want[:16])
    else:
        assert _rel_err(outs[0][:16], base[:16].float() + want[:16]) < 1e-2


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_cross_entropy(K, dtype):
    torch.manual_seed(8)
    T, V = 300, 32000
    logits = (3 * torch.randn(T, V, device="cuda")).to(dtype)
    targets = torch.randint(0, V, (T,), device="cuda")
    lf = logits.float().requires_grad_(True)
    loss_tok = torch.nn.functional.cross_entropy(lf, targets, reduction="none")
    # two "ranks": the two halves of the vocabulary
    halves = [logits[:, : V // 2].contiguous(), logits[:, V // 2 :].contiguous()]
    stats = [K.xent_stats(h, targets, i * (V // 2)) for i, h in enumerate(halves)]
    m = torch.maximum(stats[0][:, 0], stats[1][:, 0])
    s = sum(st[:, 1] * torch.exp(st[:, 0] - m) for st in stats)
    lse = m + torch.log(s)
    tl = stats[0][:, 2] + stats[1][:, 2]
    assert torch.allclose(lse - tl, loss_tok, atol=2e-3, rtol=1e-3)
    am = torch.where(stats[0][:, 0] >= stats[1][:, 0], stats[0][:, 3].view(torch.int32), stats[1][:, 3].view(torch.int32))
    assert torch.equal(am.long(), logits.float().argmax(-1))
    w = torch.rand(T, device="cuda")
    (loss_tok * w).sum().backward()
    for i, h in enumerate(halves):
        K.xent_grad_(h, targets, lse, w, None, i * (V // 2))
    got = torch.cat(halves, dim=1)
    assert _rel_err(got, lf.grad) < (2e-2 if dtype == torch.bfloat16 else 1e-4)


@pytest.mark.parametrize("gdtype", [torch.bfloat16, torch.float32])
def test_adamw_and_sumsq(K, gdtype):
    torch.manual_seed(9)
    n = 1_000_003
    p = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda").to(gdtype)
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    out = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    norm_sq = torch.zeros(1, device="cuda")
    bad = torch.zeros(1, device="cuda", dtype=torch.int32)
    clip = 1.0
    for step in range(1, 4):
        K.grad_sumsq(g, norm_sq, bad)
        assert abs(norm_sq.item() - g.float().pow(2).sum().item()) / norm_sq.item() < 1e-4
        coef = min(1.0, clip / (math.sqrt(norm_sq.item()) + 1e-6))
        ref_p.grad = g.float() * coef
        opt.step()
        K.adamw_step(p, m, v, g, out, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, step=step,
                     clip=clip, norm_sq=norm_sq)
        assert _rel_err(p, ref_p.data) < 1e-5
        assert _rel_err(out, ref_p.data) < 1e-2
    assert bad.item() == 0
    g[5] = float("inf")
    K.grad_sumsq(g, norm_sq, bad)
    assert bad.item() == 1
    before = p.clone()
    K.adamw_step(p, m, v, g, out, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, step=4, skip=bad)
    assert torch.equal(p, before)


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T,cu,heads,kv_heads", [
    (1024, [0, 1024], 4, 4),
    (1024, [0, 200, 512, 1024], 4, 2),
    (1000, [0, 37, 640, 1000], 8, 1),
    (4096, [0, 4096], 2, 1),
])
def test_attention_sm100(causal, T, cu, heads, kv_heads):
    from scaling_b200.ops.attention import attention_dense
    from scaling_b200.ops.attention_sm100 import AttentionFn

    torch.manual_seed(10)
    dev = "cuda"
    q = torch.randn(T, heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(T, kv_heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(T, kv_heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    cu_t = torch.tensor(cu, device=dev, dtype=torch.int32)
    scale = 1.0 / math.sqrt(128)
    out = AttentionFn.apply(q, k, v, cu_t, max(b - a for a, b in zip(cu[:-1], cu[1:])), causal, scale, None)
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ref = attention_dense(qf, kf, vf, cu_t, cu_t, causal, scale)
    assert _rel_err(out, ref) < 2e-2, _rel_err(out, ref)
    dout = torch.randn_like(out)
    out.backward(dout)
    ref.backward(dout.float())
    assert _rel_err(q.grad, qf.grad) < 3e-2, _rel_err(q.grad, qf.grad)
    assert _rel_err(k.grad, kf.grad) < 3e-2, _rel_err(k.grad, kf.grad)
    assert _rel_err(v.grad, vf.grad) < 3e-2, _rel_err(v.grad, vf.grad)


def _max_rel_per_row(out: torch.Tensor, ref: torch.Tensor) -> float:
    """Worst row: max-abs error of a row over that row's max-abs reference.  A few wrong rows in a tail tile barely
    move a Frobenius ratio over a 4096^2 output; they do move this."""
    out, ref = out.float(), ref.float()
    err = (out - ref).abs().amax(dim=-1)
    scale = ref.abs().amax(dim=-1).clamp_min(1e-6)
    return float((err / scale).max())


@pytest.mark.parametrize("M,N,K_", [(4096, 4096, 4096), (1000, 776, 520), (4096, 2048, 4096), (4096, 4096, 2048), (513, 257 * 8, 7168)])
def test_gemm_per_row_error(M, N, K_):
    from scaling_b200.ops import kernels as K

    torch.manual_seed(11)
    a = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K_, device="cuda", dtype=torch.bfloat16) / math.sqrt(K_)
    ref = a.float() @ b.float().t()
    assert _max_rel_per_row(K.gemm(a, b, out_dtype=torch.float32), ref) < 2e-4
    assert _max_rel_per_row(K.gemm(a, b), ref) < 2e-2
    assert _max_rel_per_row(K.gemm(a, b.t().contiguous(), b_mn=True), ref) < 2e-2


@pytest.mark.parametrize("M,N,K_,r", [(4096, 4096, 4096, 64), (1000, 776, 520, 16), (512, 1024, 4096, 192), (4096, 512, 2048, 8)])
@pytest.mark.parametrize("b_mn", [False, True])
def test_gemm_lowrank_slab(M, N, K_, r, b_mn):
    """D = A·Bᵀ + A2·B2ᵀ in one launch (LoRA as extra k-blocks) against the fp32 expression."""
    from scaling_b200.ops import kernels as K

    torch.manual_seed(3)
    a = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16)
    b = torch.randn((K_, N) if b_mn else (N, K_), device="cuda", dtype=torch.bfloat16) / math.sqrt(K_)
    a2 = torch.randn(M, r, device="cuda", dtype=torch.bfloat16)
    b2 = torch.randn((r, N) if b_mn else (N, r), device="cuda", dtype=torch.bfloat16) / math.sqrt(r)
    B = b.float() if b_mn else b.float().t()
    B2 = b2.float() if b_mn else b2.float().t()
    ref = a.float() @ B + a2.float() @ B2
    out = K.gemm_lowrank(a, b, a2, b2, b_mn=b_mn)
    assert _rel_err(out, ref) < 1e-2 and _max_rel_per_row(out, ref) < 3e-2, (_rel_err(out, ref), _max_rel_per_row(out, ref))
    # the slab really contributes: without it the result is far off
    assert _rel_err(K.gemm(a, b, b_mn=b_mn), ref) > 0.3


@pytest.mark.parametrize("kind", ["column", "row"])
def test_lora_linear_matches_autograd_of_the_plain_expression(kind):
    """Fused LoRA projection (forward and all four gradients) against torch autograd of y = x Wᵀ + s (x Aᵀ) Bᵀ."""
    from scaling_b200.ops import functional as OF

    torch.manual_seed(0)
    M, Kd, N, r, s = 1024, 1024, 768, 32, 0.5
    x = torch.randn(2, M // 2, Kd, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    W = (torch.randn(N, Kd, device="cuda") / math.sqrt(Kd)).bfloat16().requires_grad_(True)
    A = (torch.randn(r, Kd, device="cuda") / math.sqrt(Kd)).bfloat16().requires_grad_(True)
    B = (torch.randn(N, r, device="cuda") / math.sqrt(r)).bfloat16().requires_grad_(True)
    y = OF.lora_linear(x, W, A, B, s, kind)
    assert y is not None
    dy = torch.randn_like(y)
    y.backward(dy)
    got = [y.detach(), x.grad, W.grad, A.grad, B.grad]
    xf, Wf, Af, Bf = (t.detach().float().requires_grad_(True) for t in (x, W, A, B))
    yr = xf @ Wf.t() + s * (xf @ Af.t()) @ Bf.t()
    yr.backward(dy.float())
    for name, g, ref in zip(["y", "dx", "dW", "dA", "dB"], got, [yr.detach(), xf.grad, Wf.grad, Af.grad, Bf.grad]):
        assert g is not None and _rel_err(g, ref) < 2e-2, (name, _rel_err(g, ref))


def test_qkv_lora_as_one_low_rank_gemm_matches_three_separate_projections():
    """Frozen adjacent q|k|v weights + a LoRA pair on each: ONE fused GEMM each way (block-diagonal B) against torch
    autograd of the three plain expressions, output and every gradient that exists (x, A_i, B_i)."""
    from scaling_b200.core.optimizer.param_group import pack_frozen_parameters
    from scaling_b200.ops import functional as OF

    torch.manual_seed(3)
    M, Kd, r, s = 1024, 1024, 32, 2.0
    sizes = (1024, 256, 256)
    x = torch.randn(2, M // 2, Kd, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    Ws = [torch.nn.Parameter((torch.randn(n, Kd, device="cuda") / math.sqrt(Kd)).bfloat16(), requires_grad=False) for n in sizes]
    assert OF.fused_view(tuple(w.data for w in Ws)) is None  # separate allocations ...
    pack_frozen_parameters(Ws)
    assert OF.fused_view(tuple(w.data for w in Ws)) is not None  # ... adjacent after packing
    As = [(torch.randn(r, Kd, device="cuda") / math.sqrt(Kd)).bfloat16().requires_grad_(True) for _ in sizes]
    Bs = [(torch.randn(n, r, device="cuda") / math.sqrt(r)).bfloat16().requires_grad_(True) for n in sizes]
    y = OF.qkv_lora_linear(x, tuple(Ws), tuple(zip(As, Bs)), s)
    assert y is not None and y.shape == (2, M // 2, sum(sizes))
    dy = torch.randn_like(y)
    y.backward(dy)
    xf = x.detach().float().requires_grad_(True)
    Af = [a.detach().float().requires_grad_(True) for a in As]
    Bf = [b.detach().float().requires_grad_(True) for b in Bs]
    yr = torch.cat([xf @ w.float().t() + s * (xf @ a.t()) @ b.t() for w, a, b in zip(Ws, Af, Bf)], dim=-1)
    yr.backward(dy.float())
    assert _rel_err(y.detach(), yr.detach()) < 2e-2
    assert _rel_err(x.grad, xf.grad) < 2e-2
    for i in range(3):
        assert _rel_err(As[i].grad, Af[i].grad) < 2e-2, ("dA", i, _rel_err(As[i].grad, Af[i].grad))
        assert _rel_err(Bs[i].grad, Bf[i].grad) < 2e-2, ("dB", i, _rel_err(Bs[i].grad, Bf[i].grad))
    assert all(w.grad is None for w in Ws)


@pytest.mark.parametrize("M,N,K_", [(4096, 32000, 1024), (512, 1000, 256), (1000, 16000, 512), (256, 128, 64)])
def test_lm_head_gemm_with_cross_entropy_statistics_in_the_epilogue(M, N, K_):
    """Logits + per-tile CE statistics from ONE GEMM launch, merged by the combine / finalize kernels, against
    F.cross_entropy on the fp32 copy of the same bf16 logits; gradient through the in-place (softmax - onehot) kernel."""
    from scaling_b200.ops import functional as OF
    from scaling_b200.ops import kernels as K

    torch.manual_seed(1)
    x = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    w = (torch.randn(N, K_, device="cuda") * (2.0 / math.sqrt(K_))).bfloat16().requires_grad_(True)
    targets = torch.randint(0, N, (M,), device="cuda")
    logits_ref = K.gemm(x.detach(), w.detach())
    out = OF.linear_with_xent_stats(x, w)
    assert out is not None
    logits, partial = out
    assert torch.equal(logits.detach(), logits_ref), "the statistics epilogue must store the same bf16 logits"
    loss, argmax = OF.vocab_parallel_cross_entropy(logits, targets, partial=partial)
    lf = logits_ref.float()
    ref = torch.nn.functional.cross_entropy(lf, targets, reduction="none")
    assert float((loss - ref).abs().max()) < 2e-3, float((loss - ref).abs().max())
    # arg-max: equal, or a tie in value (bf16 logits tie often at large N)
    am_ref = lf.argmax(-1)
    bad = argmax != am_ref
    assert bool((lf.gather(1, argmax[:, None]) == lf.gather(1, am_ref[:, None]))[bad].all())
    weights = torch.rand(M, device="cuda")
    (loss * weights).sum().backward()
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().float().requires_grad_(True)
    (torch.nn.functional.cross_entropy(xr @ wr.t(), targets, reduction="none") * weights).sum().backward()
    assert _rel_err(x.grad, xr.grad) < 3e-2, _rel_err(x.grad, xr.grad)
    assert _rel_err(w.grad, wr.grad) < 3e-2, _rel_err(w.grad, wr.grad)



@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T,cu,heads,kv_heads,d,window", [
    (1024, [0, 1024], 4, 2, 128, 200),      # window inside one sequence, GQA
    (1000, [0, 300, 1000], 4, 4, 128, 64),  # packed documents, ragged tail, window smaller than a tile
    (777, [0, 100, 101, 777], 8, 1, 128, 300),  # MQA, T not a multiple of 64, a one-token document
    (512, [0, 512], 4, 2, 64, None),        # head dim 64 (zero-padded to the kernel's 128)
    (640, [0, 256, 640], 2, 2, 96, 100),    # head dim 96 + window
])
def test_attention_sm100_windows_and_small_heads(causal, T, cu, heads, kv_heads, d, window):
    """Sliding windows (as narrowed per-row key bounds), head dims below 128 and GQA/MQA on ragged packed batches,
    forward and all three gradients, against the dense fp32 expression."""
    from scaling_b200.ops.attention import attention, attention_dense
    from scaling_b200.ops.attention_sm100 import sm100_attention_supported

    torch.manual_seed(12)
    dev = "cuda"
    q = torch.randn(T, heads, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(T, kv_heads, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(T, kv_heads, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    cu_t = torch.tensor(cu, device=dev, dtype=torch.int32)
    scale = 1.0 / math.sqrt(d)
    assert sm100_attention_supported(q, k, v, cu_t, cu_t, causal, window, 0.0)
    out = attention(q, k, v, cu_t, max(b - a for a, b in zip(cu[:-1], cu[1:])), causal=causal, scale=scale, window=window)
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ref = attention_dense(qf, kf, vf, cu_t, cu_t, causal, scale, window)
    assert out.shape == ref.shape
    assert _rel_err(out, ref) < 2e-2 and _max_rel_per_row(out.flatten(1), ref.flatten(1)) < 6e-2, (
        _rel_err(out, ref), _max_rel_per_row(out.flatten(1), ref.flatten(1)))
    dout = torch.randn_like(out)
    out.backward(dout)
    ref.backward(dout.float())
    for name, g, r in (("dq", q.grad, qf.grad), ("dk", k.grad, kf.grad), ("dv", v.grad, vf.grad)):
        assert _rel_err(g, r) < 3e-2, (name, _rel_err(g, r))


@pytest.mark.parametrize("tq,tk,window", [(1, 700, None), (5, 300, None), (1, 129, 64), (130, 1000, None)])
def test_attention_sm100_kv_cache_decoding(tq, tk, window):
    """New query tokens against a longer key/value cache (right-aligned causal mask, forward only)."""
    from scaling_b200.ops.attention import attention, attention_dense

    torch.manual_seed(13)
    dev = "cuda"
    q = torch.randn(tq, 8, 128, device=dev, dtype=torch.bfloat16)
    k = torch.randn(tk, 2, 128, device=dev, dtype=torch.bfloat16)
    v = torch.randn(tk, 2, 128, device=dev, dtype=torch.bfloat16)
    cu_q = torch.tensor([0, tq], device=dev, dtype=torch.int32)
    cu_k = torch.tensor([0, tk], device=dev, dtype=torch.int32)
    scale = 1.0 / math.sqrt(128)
    with torch.no_grad():
        out = attention(q, k, v, cu_q, tq, causal=True, scale=scale, cu_seqlens_k=cu_k, max_seqlen_k=tk, window=window)
        ref = attention_dense(q.float(), k.float(), v.float(), cu_q, cu_k, True, scale, window)
    assert _rel_err(out, ref) < 2e-2, _rel_err(out, ref)


def _dropout_keep_mask(seed: int, p: float, heads: int, T: int, head_offset: int = 0) -> torch.Tensor:
    """[heads, T, T] bool: the kernel's stateless keep decision (csrc/attention_sm100.cu: make_drop / drop_keep)."""
    M = 0xFFFFFFFF
    s32 = ((((seed ^ (seed >> 32)) & M) * 0x9E3779B1) + 0x7F4A7C15) & M
    thr = max(1, int(p * 4294967296.0))
    q = torch.arange(T, dtype=torch.int64, device="cuda").view(1, T, 1)
    k = torch.arange(T, dtype=torch.int64, device="cuda").view(1, 1, T)
    h = (torch.arange(heads, dtype=torch.int64, device="cuda") + head_offset).view(heads, 1, 1)
    x = (q * 0x9E3779B1 + k * 0x85EBCA77 + h * 0xC2B2AE3D + s32) & M
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & M
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & M
    x = x ^ (x >> 16)
    return x >= thr


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T,cu,heads,kv_heads", [(512, [0, 512], 4, 2), (777, [0, 300, 777], 4, 1)])
def test_attention_sm100_dropout_matches_the_same_mask_in_fp32(causal, T, cu, heads, kv_heads):
    """Attention-probability dropout: forward, dK/dV blocks and dQ blocks regenerate one and the same mask.  The oracle
    applies that mask (rebuilt from the hash in Python) to the fp32 softmax; the keep rate must be about 1 - p."""
    from scaling_b200.ops.attention import dense_mask, repeat_kv
    from scaling_b200.ops.attention_sm100 import AttentionFn

    torch.manual_seed(14)
    dev, d, p, seed, hoff = "cuda", 128, 0.2, 0x1234ABCD5678, 3
    q = torch.randn(T, heads, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(T, kv_heads, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(T, kv_heads, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    cu_t = torch.tensor(cu, device=dev, dtype=torch.int32)
    scale = 1.0 / math.sqrt(d)
    out = AttentionFn.apply(q, k, v, cu_t, T, causal, scale, None, (p, seed, hoff))
    keep = _dropout_keep_mask(seed, p, heads, T, hoff)
    assert abs(float(keep.float().mean()) - (1 - p)) < 5e-3
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    kk, vv = repeat_kv(kf, heads // kv_heads), repeat_kv(vf, heads // kv_heads)
    scores = torch.einsum("qhd,khd->hqk", qf, kk) * scale
    probs = torch.softmax(scores.masked_fill(dense_mask(cu_t, T, cu_t, T, causal, None)[None], float("-inf")), dim=-1)
    ref = torch.einsum("hqk,khd->qhd", torch.nan_to_num(probs) * keep / (1 - p), vv)
    assert _rel_err(out, ref) < 2e-2, _rel_err(out, ref)
    dout = torch.randn_like(out)
    out.backward(dout)
    ref.backward(dout.float())
    for name, g, r in (("dq", q.grad, qf.grad), ("dk", k.grad, kf.grad), ("dv", v.grad, vf.grad)):
        assert _rel_err(g, r) < 3e-2, (name, _rel_err(g, r))
    # p = 0 goes through the same entry point and equals the plain kernel bit for bit
    a = AttentionFn.apply(q.detach(), k.detach(), v.detach(), cu_t, T, causal, scale, None, (0.0, seed, hoff))
    b = AttentionFn.apply(q.detach(), k.detach(), v.detach(), cu_t, T, causal, scale, None)
    assert torch.equal(a, b)


@pytest.mark.parametrize("name", ["gelu", "silu"])
@pytest.mark.parametrize("M,N,K_,bias", [(1024, 1024, 512, True), (1000, 520, 264, False), (4096, 1024, 4096, True)])
def test_linear_with_activation_in_the_epilogue(name, M, N, K_, bias):
    """act(x Wᵀ + b) from one GEMM launch (adapter / GELU-MLP up-projection) and its gradients vs torch autograd."""
    from scaling_b200.ops import functional as OF

    torch.manual_seed(2)
    x = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    w = (torch.randn(N, K_, device="cuda") / math.sqrt(K_)).bfloat16().requires_grad_(True)
    b = torch.randn(N, device="cuda").bfloat16().requires_grad_(True) if bias else None
    y = OF.linear_act(x, w, b, name)
    assert y is not None
    dy = torch.randn_like(y)
    y.backward(dy)
    xf, wf = x.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    bf = b.detach().float().requires_grad_(True) if bias else None
    act = torch.nn.functional.gelu if name == "gelu" else torch.nn.functional.silu
    ref = act(torch.nn.functional.linear(xf, wf, bf))
    ref.backward(dy.float())
    assert _rel_err(y, ref) < 2e-2 and _max_rel_per_row(y, ref) < 5e-2, (_rel_err(y, ref), _max_rel_per_row(y, ref))
    assert _rel_err(x.grad, xf.grad) < 3e-2 and _rel_err(w.grad, wf.grad) < 3e-2
    if bias:
        assert _rel_err(b.grad, bf.grad) < 3e-2


@pytest.mark.parametrize("M,nh,nkv,K_,rot", [(4096, 32, 8, 4096, 128), (1000, 4, 2, 520, 64), (512, 16, 4, 1024, 128)])
@pytest.mark.parametrize("with_pos", [False, True])
def test_qkv_gemm_with_rotary_in_the_epilogue(M, nh, nkv, K_, rot, with_pos):
    """q|k|v projection with the rotary embedding applied in the GEMM epilogue against GEMM followed by the rotary kernel
    (and against the fp32 expression): q and k columns rotated, v columns untouched."""
    from scaling_b200.ops import kernels as K

    torch.manual_seed(6)
    d, seq = 128, 256
    N = (nh + 2 * nkv) * d
    x = torch.randn(M, K_, device="cuda", dtype=torch.bfloat16)
    w = (torch.randn(N, K_, device="cuda") / math.sqrt(K_)).bfloat16()
    inv = 1.0 / (10000 ** (torch.arange(0, rot, 2, device="cuda").float() / rot))
    ang = torch.arange(seq, device="cuda").float()[:, None] * inv[None]
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    pos = torch.randint(0, seq, (M,), device="cuda", dtype=torch.int32) if with_pos else None
    y = K.gemm_rope(x, w, cos, sin, pos, seq, (nh + nkv) * d, d, rot)
    ref = (x.float() @ w.float().t()).view(M, nh + 2 * nkv, d)
    p = (pos.long() if with_pos else torch.arange(M, device="cuda") % seq)
    c, s = cos[p][:, None, :], sin[p][:, None, :]
    a, b = ref[:, :nh + nkv, 0:rot:2].clone(), ref[:, :nh + nkv, 1:rot:2].clone()
    ref[:, :nh + nkv, 0:rot:2] = a * c - b * s
    ref[:, :nh + nkv, 1:rot:2] = a * s + b * c
    ref = ref.view(M, N)
    assert _rel_err(y, ref) < 1e-2 and _max_rel_per_row(y, ref) < 3e-2, (_rel_err(y, ref), _max_rel_per_row(y, ref))
    # the unfused pair of kernels agrees as well
    y2 = K.gemm(x, w)
    K.rope_(y2.view(M, nh + 2 * nkv, d)[:, :nh + nkv], cos, sin, pos, seq, 0, rot)
    assert _rel_err(y, y2) < 1e-2
