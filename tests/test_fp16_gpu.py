"""fp16 on the sm_100a kernels: the linear-layer GEMMs (fprop / dgrad / wgrad), attention forward + backward and an fp16
training step with the dynamic loss scaler (reference acceptance grid: tests/core/test_training/test_training.py:149-235
trains fp16 + loss scaler).  Every kernel is compared against the fp32 PyTorch expression of the same op; fp16 has three
more mantissa bits than bf16, so the thresholds are TIGHTER than the bf16 tests' — a kernel that silently treated the
operands as bf16 fails them by orders of magnitude."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.mark.parametrize("M,N,K_", [(256, 128, 64), (4096, 4096, 4096), (1000, 776, 520), (512, 2048, 14336)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_fp16_gemm_layouts(M, N, K_, a_mn, b_mn):
    from scaling_b200.ops import kernels as K

    torch.manual_seed(0)
    a = torch.randn((K_, M) if a_mn else (M, K_), device="cuda", dtype=torch.float16)
    b = torch.randn((K_, N) if b_mn else (N, K_), device="cuda", dtype=torch.float16) / math.sqrt(K_)
    out = K.gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    assert out.dtype == torch.float16
    af = a.float().t() if a_mn else a.float()
    bf = b.float() if b_mn else b.float().t()
    ref = af @ bf
    assert _rel_err(out, ref) < 1e-3, _rel_err(out, ref)
    # per row: a few wrong rows in a tail tile must not hide in the Frobenius norm
    row = (out.float() - ref).norm(dim=1) / (ref.norm(dim=1) + 1e-6)
    assert float(row.max()) < 3e-3, float(row.max())


def test_fp16_wgrad_accumulates_into_an_fp32_arena_and_bias_is_added():
    from scaling_b200.ops import kernels as K

    torch.manual_seed(1)
    M, N, K_ = 2048, 512, 1024
    dy = torch.randn(M, N, device="cuda", dtype=torch.float16)
    x = torch.randn(M, K_, device="cuda", dtype=torch.float16)
    main = torch.randn(N, K_, device="cuda", dtype=torch.float32)
    ref = main + dy.float().t() @ x.float()
    K.gemm(dy, x, a_mn=True, b_mn=True, out=main, accumulate=True)
    assert _rel_err(main, ref) < 1e-5, _rel_err(main, ref)
    w = torch.randn(N, K_, device="cuda", dtype=torch.float16) / math.sqrt(K_)
    bias = torch.randn(N, device="cuda", dtype=torch.float16)
    y = K.gemm(x, w, bias=bias)
    assert _rel_err(y, x.float() @ w.float().t() + bias.float()) < 1e-3
    assert not K.gemm_f16_supported(128, 128, False, False) and not K.gemm_f16_supported(512, 512, True, False)


def test_fp16_linear_autograd_runs_the_own_gemm():
    from scaling_b200.ops import _lib
    from scaling_b200.ops import functional as OF

    torch.manual_seed(2)
    x = torch.randn(2, 512, 1024, device="cuda", dtype=torch.float16, requires_grad=True)
    w = (torch.randn(768, 1024, device="cuda") / 32).half().requires_grad_(True)
    b = torch.randn(768, device="cuda").half().requires_grad_(True)
    before = _lib.launch_count()
    y = OF.linear(x, w, b)
    dy = torch.randn_like(y)
    y.backward(dy)
    assert _lib.launch_count() - before >= 3, "fprop, dgrad and wgrad must be own launches"
    xf, wf, bf = (t.detach().float().requires_grad_(True) for t in (x, w, b))
    yr = torch.nn.functional.linear(xf, wf, bf)
    yr.backward(dy.float())
    for name, g, r in (("y", y, yr), ("dx", x.grad, xf.grad), ("dw", w.grad, wf.grad), ("db", b.grad, bf.grad)):
        assert _rel_err(g, r) < 2e-3, (name, _rel_err(g, r))


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("T,cu,heads,kv_heads,d", [
    (1024, [0, 200, 512, 1024], 4, 2, 128),
    (1000, [0, 37, 640, 1000], 8, 1, 128),
    (777, [0, 777], 4, 4, 64),
])
def test_fp16_attention_forward_and_backward(causal, T, cu, heads, kv_heads, d):
    from scaling_b200.ops.attention import attention, attention_dense

    torch.manual_seed(10)
    q = torch.randn(T, heads, d, device="cuda", dtype=torch.float16, requires_grad=True)
    k = torch.randn(T, kv_heads, d, device="cuda", dtype=torch.float16, requires_grad=True)
    v = torch.randn(T, kv_heads, d, device="cuda", dtype=torch.float16, requires_grad=True)
    cu_t = torch.tensor(cu, device="cuda", dtype=torch.int32)
    scale = 1.0 / math.sqrt(d)
    from scaling_b200.ops import _lib

    before = _lib.launch_count()
    out = attention(q, k, v, cu_t, max(b - a for a, b in zip(cu[:-1], cu[1:])), causal=causal, scale=scale)
    assert out.dtype == torch.float16 and _lib.launch_count() > before, "fp16 must run the sm_100a kernel"
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ref = attention_dense(qf, kf, vf, cu_t, cu_t, causal, scale)
    assert _rel_err(out, ref) < 3e-3, _rel_err(out, ref)
    dout = torch.randn_like(out)
    out.backward(dout)
    ref.backward(dout.float())
    for name, g, r in (("dq", q.grad, qf.grad), ("dk", k.grad, kf.grad), ("dv", v.grad, vf.grad)):
        assert _rel_err(g, r) < 5e-3, (name, _rel_err(g, r))


def _fp16_worker(rank: int, world_size: int, master_port: int, config_dict):
    from scaling_b200.core import LaunchConfig
    from scaling_b200.ops import _lib
    from scaling_b200.transformer.train import main

    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=config_dict)
    before = _lib.launch_count()
    metrics = main(launch, return_metrics=True)
    return [(m["training/loss"], m.get("training/loss_scale", m.get("training/current_loss_scale")))
            for m in (metrics or [])], _lib.launch_count() - before


def test_fp16_transformer_trains_with_the_dynamic_loss_scaler(tmp_path):
    """Tiny fp16 model, one GPU, through the public train entry: finite decreasing loss with the loss scaler on, and the
    step is made of own kernel launches (no silent library fallback for the projections / attention)."""
    from tests.transformer_utils import tiny_config, write_dataset
    from tests.utils import dist_launcher

    prefix = tmp_path / "d" / "tokens"
    write_dataset(prefix, n_docs=512, vocab=512)
    a = {"vocab_size": 512, "hidden_size": 512, "sequence_length": 256, "precision": "float16", "num_layers": 2,
         "masked_softmax": {"kernel": "flash_attention"}}
    cfg = tiny_config(prefix, iters=8, gas=2, arch=a, learning_rate_scheduler={"learning_rate": 2e-3},
                      optimizer={"zero": False, "loss_scaler": {"enable": True, "initial_scale": 4096.0, "window": 4}})
    rows, launches = dist_launcher(_fp16_worker, world_size=1, timeout=400, config_dict=cfg)[0]
    losses = [r[0] for r in rows]
    assert len(losses) == 8 and all(math.isfinite(x) for x in losses), losses
    assert losses[-1] < losses[0], losses
    # per micro-batch and layer: q|k|v (3), dense, gate, up, down projections fwd + 2 bwd each, attention fwd + bwd
    assert launches > 8 * 2 * 2 * 20, launches
