"""nn building blocks against plain-torch oracles, single process and TP=2 over gloo (reference: tests/core/test_nn/*,
tests/core/unit/nn/linear/test_utils.py; SURVEY §4.3 incl. the gradient-parity and non-vacuous LoRA gaps of §4.4)."""
import math

import pytest
import torch

from scaling_b200.core import (
    LayerNorm,
    LayerNormConfig,
    LoRaConfig,
    LoRAModuleType,
    MaskedSoftmax,
    MaskedSoftmaxConfig,
    NormType,
    ParallelSelfAttention,
    RelativePositionEmbeddingType,
    RMSNorm,
    RotaryConfig,
    RotaryEmbedding,
    RotaryEmbeddingComplex,
    Topology,
    TopologyConfig,
)
from scaling_b200.core.nn import cumulative_seq_lengths_to_dense_attention_mask, repeat_kv
from scaling_b200.ops.attention import attention_dense, dense_mask
from tests.utils import dist_launcher

CPU = torch.device("cpu")


# --------------------------------------------------------------------------------------------------- rotary
@pytest.mark.parametrize("cls", [RotaryEmbedding, RotaryEmbeddingComplex])
def test_rotary_position_invariants(cls):
    torch.manual_seed(0)
    rot = cls(RotaryConfig(dimensions=16, base=10000, max_seq_length=64), device=CPU)
    q, k = torch.randn(12, 2, 3, 16), torch.randn(12, 2, 3, 16)
    q0, k0 = rot(q, k)
    pos = torch.arange(12).unsqueeze(1).expand(12, 2)
    q1, k1 = rot(q, k, query_position_ids=pos, key_position_ids=pos)
    assert torch.equal(q0, q1) and torch.equal(k0, k1)  # explicit default positions: bit-equal
    # a sequence offset: rotating tokens 4.. of a longer sequence equals rotating them alone with offset positions
    q2, _ = rot(q[4:], k[4:], query_position_ids=pos[4:], key_position_ids=pos[4:])
    assert torch.equal(q2, q0[4:])
    # rotation preserves norms and q·k depends only on the relative distance
    assert torch.allclose(q0.norm(dim=-1), q.norm(dim=-1), atol=1e-5)
    qa, ka = rot(q[:1].expand(12, -1, -1, -1).contiguous(), k[:1].expand(12, -1, -1, -1).contiguous())
    d1 = (qa[5] * ka[3]).sum(-1)
    d2 = (qa[9] * ka[7]).sum(-1)
    assert torch.allclose(d1, d2, atol=1e-4)


def test_rotary_complex_equals_rotary_up_to_permutation():
    torch.manual_seed(1)
    cfg = RotaryConfig(dimensions=8, base=10000, max_seq_length=32)
    half, inter = RotaryEmbedding(cfg, device=CPU), RotaryEmbeddingComplex(cfg, device=CPU)
    x = torch.randn(6, 1, 2, 8)
    perm = torch.tensor([0, 4, 1, 5, 2, 6, 3, 7])  # interleaved pair i = (half i, half i + d/2)
    y_half, _ = half(x, x)
    y_int, _ = inter(x[..., perm], x[..., perm])
    assert torch.allclose(y_int, y_half[..., perm], atol=1e-5)


def test_partial_rotary_leaves_tail_untouched():
    rot = RotaryEmbedding(RotaryConfig(dimensions=4, max_seq_length=16), device=CPU)
    x = torch.randn(5, 1, 2, 8)
    y, _ = rot(x, x)
    assert torch.equal(y[..., 4:], x[..., 4:]) and not torch.equal(y[1:, ..., :4], x[1:, ..., :4])
    packed = rot.apply_packed(x.reshape(5, 2, 8), None, 5)
    assert torch.allclose(packed, y.reshape(5, 2, 8), atol=1e-6)


# ---------------------------------------------------------------------------------------------- norms / softmax
def test_norms_match_torch():
    torch.manual_seed(0)
    x = torch.randn(3, 5, 32)
    ln = LayerNorm(LayerNormConfig(layernorm_epsilon=1e-5), 32, device=CPU)
    with torch.no_grad():
        ln.weight.normal_()
        ln.get_bias().normal_()
    ref = torch.nn.functional.layer_norm(x, (32,), ln.weight, ln.get_bias(), 1e-5)
    assert torch.allclose(ln(x), ref, atol=1e-6)
    rms = RMSNorm(LayerNormConfig(layernorm_epsilon=1e-6), 32, device=CPU)
    with torch.no_grad():
        rms.weight.normal_()
    ref = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6) * rms.weight
    assert torch.allclose(rms(x), ref, atol=1e-6)
    y, summed = rms(x, residual=torch.ones_like(x))
    assert torch.allclose(summed, x + 1) and torch.allclose(y, rms(x + 1), atol=1e-6)
    named = dict(LayerNorm(LayerNormConfig(), 8, device=CPU, bitfit_bias_name="tune").named_parameters())
    assert set(named) == {"weight", "bias_tune"}


@pytest.mark.parametrize("fp32", [True, False])
def test_masked_softmax(fp32):
    torch.manual_seed(0)
    sm = MaskedSoftmax(MaskedSoftmaxConfig(softmax_in_fp32=fp32, scale=0.5))
    x = torch.randn(2, 4, 4).bfloat16()
    mask = torch.triu(torch.ones(4, 4, dtype=torch.bool), 1)[None]
    p = sm(x, mask)
    assert p.dtype == torch.bfloat16 and torch.all(p.masked_select(mask.expand_as(p)) == 0)
    ref = torch.softmax((x.float() * 0.5).masked_fill(mask, -10000.0), -1)
    assert torch.allclose(p.float(), ref, atol=2e-2)
    assert torch.allclose(p.float().sum(-1), torch.ones(2, 4), atol=2e-2)


# ---------------------------------------------------------------------------------------------- attention helpers
def test_dense_mask_packed_causal_window_and_cache():
    cu = torch.tensor([0, 3, 5], dtype=torch.int32)
    m = cumulative_seq_lengths_to_dense_attention_mask(cu, seq_length_per_batch_item=5, causal=True)
    allowed = ~m.reshape(5, 5) if m.dtype == torch.bool else m.reshape(5, 5) == 0
    expect = torch.tensor([[1, 0, 0, 0, 0], [1, 1, 0, 0, 0], [1, 1, 1, 0, 0], [0, 0, 0, 1, 0], [0, 0, 0, 1, 1]], dtype=torch.bool)
    assert torch.equal(allowed, expect)
    bidir = ~dense_mask(cu, 5, cu, 5, causal=False, window=None)
    assert torch.equal(bidir, torch.block_diag(torch.ones(3, 3), torch.ones(2, 2)).bool())
    win = ~dense_mask(torch.tensor([0, 6]), 6, torch.tensor([0, 6]), 6, causal=True, window=2)
    assert win.sum(-1).tolist() == [1, 2, 3, 3, 3, 3]
    # decoding with a KV cache: 2 new queries against 5 keys are right-aligned
    dec = ~dense_mask(torch.tensor([0, 2]), 2, torch.tensor([0, 5]), 5, causal=True, window=None)
    assert dec.tolist() == [[True] * 4 + [False], [True] * 5]


def test_repeat_kv_and_gqa_equivalence():
    torch.manual_seed(0)
    q, k, v = torch.randn(7, 4, 8), torch.randn(7, 2, 8), torch.randn(7, 2, 8)
    assert torch.equal(repeat_kv(k, 2)[:, [0, 2]], k) and torch.equal(repeat_kv(k, 2)[:, 0], repeat_kv(k, 2)[:, 1])
    cu = torch.tensor([0, 4, 7], dtype=torch.int32)
    out = attention_dense(q, k, v, cu, cu, True, 8 ** -0.5)
    ref = attention_dense(q, repeat_kv(k, 2), repeat_kv(v, 2), cu, cu, True, 8 ** -0.5)
    assert torch.allclose(out, ref, atol=1e-6)
    # packed sequences do not leak: the second sequence alone gives the same rows
    alone = attention_dense(q[4:], k[4:], v[4:], torch.tensor([0, 3]), torch.tensor([0, 3]), True, 8 ** -0.5)
    assert torch.allclose(out[4:], alone, atol=1e-6)


@pytest.mark.parametrize("kv_heads,local_heads", [(4, 0), (2, 0), (2, 2), (4, 4), (1, 1)])
def test_attention_module_local_and_global_heads(kv_heads, local_heads):
    torch.manual_seed(0)
    attn = ParallelSelfAttention(hidden_size=32, num_attention_heads=4, masked_softmax_config=MaskedSoftmaxConfig(),
                                 num_local_attention_heads=local_heads, local_attention_window_size=3 if local_heads else None,
                                 rotary_config=RotaryConfig(dimensions=8, max_seq_length=32),
                                 relative_position_embedding_type=RelativePositionEmbeddingType.ROTARY_COMPLEX, bias=False,
                                 device=CPU, qkv_in_one=False, num_kv_heads=kv_heads, norm_type=NormType.RMS)
    x = torch.randn(2, 10, 32, requires_grad=True)
    cu = torch.tensor([0, 6, 10, 20], dtype=torch.int32)
    y = attn(x, cumulative_seq_lengths=cu, position_ids=None)
    assert y.shape == x.shape
    y.sum().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all()
    # tokens of the first packed sequence are unaffected by what follows
    x2 = x.detach().clone()
    x2[0, 6:] += 1.0
    y2 = attn(x2, cumulative_seq_lengths=cu, position_ids=None)
    assert torch.allclose(y2[0, :6], y[0, :6].detach(), atol=1e-5)


# ------------------------------------------------------------------------------------------------------- LoRA
@pytest.mark.parametrize("qkv_in_one", [False, True])
def test_lora_merge_is_exact(qkv_in_one):
    """Merged weights reproduce the adapter forward (the reference's check passes with atol=1e6, i.e. nothing)."""
    torch.manual_seed(0)
    attn = ParallelSelfAttention(hidden_size=32, num_attention_heads=4, masked_softmax_config=MaskedSoftmaxConfig(),
                                 rotary_config=RotaryConfig(dimensions=8, max_seq_length=32), bias=False, device=CPU,
                                 qkv_in_one=qkv_in_one, lora_config=LoRaConfig(rank=4, alpha=8))
    for m in attn.lora_modules.values():  # B is zero-initialised: make the update non-trivial
        torch.nn.init.normal_(m.dense_out.weight, std=0.2)
    x = torch.randn(2, 6, 32)
    before_w = attn.dense.weight.detach().clone()
    with torch.no_grad():
        y_adapter = attn(x, cumulative_seq_lengths=None, position_ids=None)
        attn.merge_lora_weights()
        y_merged = attn(x, cumulative_seq_lengths=None, position_ids=None)
    assert not hasattr(attn, "lora_modules") and attn.lora_merged_state
    assert not torch.equal(attn.dense.weight, before_w)
    assert torch.allclose(y_adapter, y_merged, atol=1e-5), (y_adapter - y_merged).abs().max()


# ------------------------------------------------------------------------------------------- TP=2 parity (gloo)
def _tp_worker(rank: int, world_size: int, master_port: int):
    from scaling_b200.core import ColumnParallelLinear, ParallelMLP, ParallelSwiGLUMLP, RowParallelLinear, VocabParallelEmbedding
    from scaling_b200.core.nn.tp_ops import (
        all_concat,
        all_reduce,
        all_reduce_scatter_to_sequence_parallel,
        all_shard,
        copy_to_tensor_model_parallel_region,
        gather_from_sequence_parallel_region,
    )
    from scaling_b200.core.utils.param_merge import merge_parameter

    topo = Topology(TopologyConfig(model_parallel_size=2, pipe_parallel_size=1, data_parallel_size=1, global_rank=rank,
                                   micro_batch_size=1, gradient_accumulation_steps=1))
    topo.initialize_distributed("127.0.0.1", str(master_port))
    torch.manual_seed(7)  # same inputs on both ranks
    out = {}

    # conjugate operator pairs: forward semantics and gradients
    x = torch.randn(2, 4, 6, requires_grad=True)
    y = copy_to_tensor_model_parallel_region(x, topo)
    (y * (rank + 1)).sum().backward()
    assert torch.equal(y, x) and torch.allclose(x.grad, torch.full_like(x, 3.0))  # identity fwd, sum of grads bwd
    x = torch.randn(2, 4, 6, requires_grad=True)
    y = all_reduce(x * (rank + 1), topo)
    y.sum().backward()
    assert torch.allclose(y, x.detach() * 3) and torch.allclose(x.grad, torch.full_like(x, float(rank + 1)))
    full = torch.arange(24.0).reshape(2, 12).requires_grad_(True)
    shard = all_shard(full, dim=-1, topology=topo)
    assert torch.equal(shard, full.detach()[:, rank * 6:(rank + 1) * 6])
    back = all_concat(shard, dim=-1, topology=topo)
    assert torch.equal(back, full.detach())
    back.sum().backward()
    assert torch.allclose(full.grad, torch.ones_like(full))
    seq = torch.randn(2, 8, 4, requires_grad=True)
    part = all_reduce_scatter_to_sequence_parallel(seq * (rank + 1), topo)
    assert part.shape == (2, 4, 4) and torch.allclose(part, seq.detach()[:, rank * 4:(rank + 1) * 4] * 3)
    whole = gather_from_sequence_parallel_region(part, topo)
    assert torch.allclose(whole, seq.detach() * 3)
    whole.sum().backward()
    assert seq.grad is not None

    # column -> row pair equals the dense computation with the merged weights, forward AND gradients
    torch.manual_seed(11 + rank)  # different shards per rank
    col = ColumnParallelLinear(16, 24, bias=True, topology=topo, parallel_output=True)
    row = RowParallelLinear(24, 16, bias=True, topology=topo, parallel_input=True)
    with torch.no_grad():
        col.get_bias().normal_()
        row.get_bias().copy_(torch.linspace(-1, 1, 16))
    torch.manual_seed(3)
    x = torch.randn(5, 16, requires_grad=True)
    y = row(torch.tanh(col(x)))
    y.square().sum().backward()
    def merged(p):
        return merge_parameter(p.data, p.core_parameter_meta, topo)

    wc, bc = merged(col.weight), merged(col.get_bias())
    wr = merged(row.weight)
    xr = x.detach().clone().requires_grad_(True)
    wc_, wr_ = wc.clone().requires_grad_(True), wr.clone().requires_grad_(True)
    yr = torch.nn.functional.linear(torch.tanh(torch.nn.functional.linear(xr, wc_, bc)), wr_, row.get_bias().detach())
    yr.square().sum().backward()
    out["linear_fwd"] = (y - yr).abs().max().item()
    out["linear_dx"] = (x.grad - xr.grad).abs().max().item()
    out["linear_dwc"] = (col.weight.grad - wc_.grad[rank * 12:(rank + 1) * 12]).abs().max().item()
    out["linear_dwr"] = (row.weight.grad - wr_.grad[:, rank * 12:(rank + 1) * 12]).abs().max().item()
    gathered = ColumnParallelLinear(16, 24, bias=False, topology=topo, parallel_output=False)
    assert gathered(x.detach()).shape == (5, 24)

    # vocab-parallel embedding equals nn.Embedding with the merged table; finetunable ids mask the gradient
    emb = VocabParallelEmbedding(10, 8, topology=topo, finetunable_token_ids=[1, 7])
    ids = torch.tensor([[0, 1, 5, 7, 9, 9]])
    e = emb(ids)
    table = merged(emb.weight)
    out["emb_fwd"] = (e - torch.nn.functional.embedding(ids, table)).abs().max().item()
    e.sum().backward()
    nz = (emb.weight.grad.abs().sum(-1) > 0).nonzero().flatten() + emb.vocab_start_index
    out["emb_trainable_rows"] = nz.tolist()

    # MLPs run and agree between ranks (replicated output)
    for cls in (ParallelMLP, ParallelSwiGLUMLP):
        torch.manual_seed(21 + rank)
        mlp = cls(io_features=16, intermediate_feature_factor=2.0, bias=False, topology=topo)
        torch.manual_seed(5)
        h = mlp(torch.randn(3, 4, 16))
        both = [torch.empty_like(h) for _ in range(2)]
        torch.distributed.all_gather(both, h.detach(), group=topo.model_parallel_group)
        assert torch.equal(both[0], both[1])
    return out


def test_tensor_parallel_parity_gloo():
    res = dist_launcher(_tp_worker, 2)
    for r in (0, 1):
        o = res[r]
        assert o["linear_fwd"] < 1e-5 and o["linear_dx"] < 1e-4 and o["linear_dwc"] < 1e-4 and o["linear_dwr"] < 1e-4, o
        assert o["emb_fwd"] == 0.0
    assert res[0]["emb_trainable_rows"] == [1] and res[1]["emb_trainable_rows"] == [7]


# --------------------------------------------------------------------------------------------------- row-LoRA operands
def _row_lora_cache_worker(rank: int, world_size: int, master_port: int):
    import torch.distributed as dist

    from scaling_b200.ops import functional as OF

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{master_port}", rank=rank, world_size=world_size)
    try:
        torch.manual_seed(0)
        r, kl, n = 8, 6, 10
        a_full, b_full = torch.randn(r, kl * world_size), torch.randn(n, r)
        rl = r // world_size
        A = a_full[rank * rl:(rank + 1) * rl].clone()  # rank dimension is TP-split, all input columns
        B = b_full[:, rank * rl:(rank + 1) * rl].clone()
        tp = (world_size, rank, dist.group.WORLD)
        a1, b1 = OF._row_lora_operands(A, B, tp, kl)
        assert torch.equal(a1, a_full[:, rank * kl:(rank + 1) * kl]) and torch.equal(b1, b_full)
        a2, _ = OF._row_lora_operands(A, B, tp, kl)
        assert a2 is a1, "same weights within a step: the gathers run once"
        A.mul_(2.0)  # in-place change seen by autograd's version counter
        a3, _ = OF._row_lora_operands(A, B, tp, kl)
        assert a3 is not a1 and torch.equal(a3, 2.0 * a_full[:, rank * kl:(rank + 1) * kl])
        OF.bump_weight_epoch()  # what the optimizer's arena kernels announce
        a4, _ = OF._row_lora_operands(A, B, tp, kl)
        assert a4 is not a3
        # another tensor object at the same address must not hit the entry
        alias = A.view_as(A)
        a5, _ = OF._row_lora_operands(alias, B, tp, kl)
        assert a5 is not a4
        return True
    finally:
        dist.destroy_process_group()


def test_row_lora_operands_are_gathered_once_per_weight_version():
    assert all(dist_launcher(_row_lora_cache_worker, world_size=2).values())
