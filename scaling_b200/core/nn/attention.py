"""Tensor-parallel self attention: MHA / GQA / MQA, packed sequences, rotary, sliding window, QK-norm, KV cache, LoRA.

Reference: core/nn/attention/attention.py:1-796.  Parameter names (``query_key_value`` or ``query``/``key``/``value``,
``dense``, ``norm_query``/``norm_key``, ``lora_modules.*``) and the fused-QKV row layout (head-major, [q|k|v] inside a
head) are the checkpoint contract and are kept.  What changed underneath:

  * activations stay token-major ``[b·s, heads, d]`` — no ``[s, b, …]`` transposes;
  * separate q/k/v projections run as one GEMM forward and one dgrad + one wgrad GEMM backward, with a single
    copy-to-TP-region (1 dX all-reduce instead of 3);
  * rotary is one in-place kernel; attention is the sm_100a kernel (or the dense PyTorch oracle on CPU);
  * no host sync for ``max_seq_length`` — the static sequence length bounds it;
  * true MQA / fewer KV heads than TP ranks: KV projections are replicated inside TP sub-groups and their gradients
    all-reduced over the replicas (the reference divides by zero);
  * biases and the LoRA-dense term are added once, after the TP reduction.
"""
from __future__ import annotations

import math
from enum import Enum
from typing import Callable, Dict, List, Optional, Tuple, Union

import torch

from ...ops import functional as OF
from ...ops.attention import attention as attention_op
from ...ops.attention import dense_mask, repeat_kv  # noqa: F401  (re-exported helpers)
from ...parallel.symm import get_tp_comm
from ..topology import Topology
from .linear import ColumnParallelLinear, RowParallelLinear
from .lora import LoRaConfig, LoRAModuleType, ParallelLoRa
from .norms import LayerNorm, LayerNormConfig, NormType, RMSNorm, get_norm
from .rotary import RotaryConfig, RotaryEmbedding, RotaryEmbeddingComplex
from .softmax import MaskedSoftmax, MaskedSoftmaxConfig, MaskedSoftmaxKernel
from .tp_ops import (
    all_concat,
    all_reduce,
    all_reduce_scatter_to_sequence_parallel,
    all_shard,
    copy_to_tensor_model_parallel_region,
    gather_from_sequence_parallel_region,
    get_device,
)


class RelativePositionEmbeddingType(Enum):
    NONE = "none"
    ROTARY = "rotary"
    ROTARY_COMPLEX = "rotary_complex"


def cumulative_seq_lengths_to_dense_attention_mask(cumulative_seq_lengths: torch.Tensor, seq_length_per_batch_item: int,
                                                   causal: bool = True) -> torch.Tensor:
    """[b, 1, s, s] bool mask (True = masked) for packed sequences, from flat boundaries over b·s tokens."""
    total = int(cumulative_seq_lengths[-1].item()) if cumulative_seq_lengths.numel() else 0
    b = max(total // seq_length_per_batch_item, 1)
    full = dense_mask(cumulative_seq_lengths, total, cumulative_seq_lengths, total, causal, None)
    s = seq_length_per_batch_item
    return torch.stack([full[i * s:(i + 1) * s, i * s:(i + 1) * s] for i in range(b)]).unsqueeze(1)


def get_max_seq_length(cumulative_seq_lengths: torch.Tensor) -> int:
    return int((cumulative_seq_lengths[1:] - cumulative_seq_lengths[:-1]).max().item())


class ParallelSelfAttention(torch.nn.Module):
    def __init__(self, hidden_size: int, num_attention_heads: int, masked_softmax_config: MaskedSoftmaxConfig,
                 causal: bool = True, num_local_attention_heads: int = 0,
                 local_attention_window_size: Optional[int] = None, scaling_factor: Optional[float] = None,
                 dropout_attention_probs: float = 0.0, rotary_config: Optional[RotaryConfig] = None,
                 relative_position_embedding_type: RelativePositionEmbeddingType = RelativePositionEmbeddingType.ROTARY,
                 bias: bool = True, device: Optional[torch.device] = None, dtype: torch.dtype = torch.float32,
                 topology: Optional[Topology] = None,
                 init_method: Callable[[torch.Tensor], torch.Tensor] = torch.nn.init.xavier_normal_,
                 bitfit_bias_name: Optional[str] = None, lora_config: Optional[LoRaConfig] = None,
                 norm_type: NormType = NormType.LAYERNORM, key_query_norm: bool = False,
                 layernorm_config: Optional[LayerNormConfig] = None, qkv_in_one: bool = True,
                 num_kv_heads: Optional[int] = None, use_matmul: bool = False):
        super().__init__()
        self._device = get_device(topology=topology, device=device)
        assert hidden_size % num_attention_heads == 0, (
            f"hidden size ({hidden_size}) must be divisible by num_attention_heads ({num_attention_heads})")
        self.hidden_size = hidden_size
        self.head_dim = self.hidden_size_per_attention_head = hidden_size // num_attention_heads
        self.num_attention_heads = num_attention_heads
        self.causal = causal
        self.topology = topology
        self.dtype = dtype
        self.lora_config = lora_config
        self.masked_softmax_config = masked_softmax_config
        self.use_flash_attention = masked_softmax_config.kernel == MaskedSoftmaxKernel.FLASH_ATTENTION
        self.num_local_attention_heads = num_local_attention_heads
        self.local_attention_window_size = local_attention_window_size
        tp = 1 if topology is None else topology.config.model_parallel_size
        self.model_parallel_size = tp
        if num_local_attention_heads > 0:
            assert local_attention_window_size is not None, "`local_attention_window_size` needs to be set"
            if num_local_attention_heads != num_attention_heads and tp > 1:
                raise NotImplementedError("Mixed range attention is not supported with tensor parallelism.")
        assert num_attention_heads % tp == 0, (
            f"attention heads ({num_attention_heads}) must be divisible by model parallel size ({tp})")
        self.num_attention_heads_per_partition = num_attention_heads // tp
        self.qkv_in_one = qkv_in_one
        self.num_kv_heads = num_kv_heads
        self.kv_replicas = 1  # TP ranks that hold a copy of the same KV head (true MQA / kv_heads < tp)
        if num_kv_heads:
            assert not qkv_in_one, "for a differing number of kv heads, qkv cannot be stored in one"
            assert num_attention_heads % num_kv_heads == 0
            if num_kv_heads >= tp:
                assert num_kv_heads % tp == 0
                self.num_kv_heads_per_partition = num_kv_heads // tp
            else:
                assert tp % num_kv_heads == 0
                self.num_kv_heads_per_partition = 1
                self.kv_replicas = tp // num_kv_heads
            self.num_repeat_kv = self.num_attention_heads_per_partition // self.num_kv_heads_per_partition
        else:
            self.num_kv_heads_per_partition = self.num_attention_heads_per_partition
            self.num_repeat_kv = 1
        kv_out = hidden_size // (num_attention_heads // num_kv_heads) if num_kv_heads else hidden_size

        if lora_config:
            self.lora_merged_state = False
            self.lora_modules = torch.nn.ModuleDict()
            for module_type in lora_config.parallel_modules:
                out = hidden_size if module_type in (LoRAModuleType.DENSE, LoRAModuleType.QUERY) else kv_out
                self.lora_modules[f"{module_type.value}_{lora_config.name}"] = ParallelLoRa(
                    in_features=hidden_size, out_features=out, rank=lora_config.rank, topology=topology,
                    dropout=lora_config.dropout, dtype=dtype, lora_module_type=module_type, alpha=lora_config.alpha,
                    bias=lora_config.bias, kaiming_a=lora_config.kaiming_a,
                    device=None if topology is not None else self._device)

        kw = dict(bias=bias, device=device, dtype=dtype, topology=topology, init_method=init_method,
                  bitfit_bias_name=bitfit_bias_name, parallel_output=True)
        if qkv_in_one:
            self.query_key_value = ColumnParallelLinear(hidden_size, hidden_size * 3, **kw)
        else:
            self.query = ColumnParallelLinear(hidden_size, hidden_size, **kw)
            if self.kv_replicas > 1:
                # replicate: build the [kv_heads·d, h] projection per KV head group, each rank keeps its head
                self.key = _ReplicatedKV(hidden_size, self.head_dim, num_kv_heads, topology, bias, dtype, init_method,
                                         bitfit_bias_name)
                self.value = _ReplicatedKV(hidden_size, self.head_dim, num_kv_heads, topology, bias, dtype,
                                           init_method, bitfit_bias_name)
            else:
                self.key = ColumnParallelLinear(hidden_size, kv_out, **kw)
                self.value = ColumnParallelLinear(hidden_size, kv_out, **kw)
        self.use_matmul = use_matmul
        self.scaling_factor = 1 / math.sqrt(self.head_dim) if scaling_factor is None else scaling_factor

        self.rotary_embedding: Optional[Union[RotaryEmbedding, RotaryEmbeddingComplex]] = None
        if relative_position_embedding_type == RelativePositionEmbeddingType.ROTARY:
            assert rotary_config is not None
            self.rotary_embedding = RotaryEmbedding(rotary_config, device=self._device, dtype=dtype)
        elif relative_position_embedding_type == RelativePositionEmbeddingType.ROTARY_COMPLEX:
            assert rotary_config is not None
            self.rotary_embedding = RotaryEmbeddingComplex(rotary_config, device=self._device)

        self.key_query_norm = key_query_norm
        self.norm_query: Optional[Union[LayerNorm, RMSNorm]] = None
        self.norm_key: Optional[Union[LayerNorm, RMSNorm]] = None
        if key_query_norm:
            self.norm_query = get_norm(norm_type, layernorm_config, self.head_dim, self._device, dtype, bitfit_bias_name)
            self.norm_key = get_norm(norm_type, layernorm_config, self.head_dim, self._device, dtype, bitfit_bias_name)
            # one replicated weight per norm, but every TP rank applies it to ITS heads only: the gradient is a partial
            # sum and must be all-reduced over the TP group before the step (the reference never does -> its replicas
            # drift apart under TP)
            if self.model_parallel_size > 1:
                from .parameter_meta import CoreParameterMeta

                for norm in (self.norm_query, self.norm_key):
                    for prm in norm.parameters():
                        meta = getattr(prm, "core_parameter_meta", None) or CoreParameterMeta.register_on_parameter(
                            prm, is_model_parallel=False)
                        meta.tp_grad_reduce = "always"  # type: ignore[attr-defined]

        self.dropout_attention_probs = dropout_attention_probs
        self.dropout = torch.nn.Dropout(dropout_attention_probs)
        self.dense = RowParallelLinear(hidden_size, hidden_size, bias=bias, topology=topology, dtype=dtype,
                                       device=device, bitfit_bias_name=bitfit_bias_name, init_method=init_method,
                                       parallel_input=True, parallel_output=True)
        self.masked_softmax = MaskedSoftmax(masked_softmax_config)
        self.cache: Dict[int, Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]] = {}

    # ------------------------------------------------------------------------------------------------
    def _sp(self) -> bool:
        return self.topology is not None and self.topology.config.sequence_parallel

    def supports_sequence_parallel_shard(self) -> bool:
        """True when ``forward(..., sequence_parallel_shard=True)`` can take the norm's UNGATHERED [b, s/tp, h] output and
        run the all-gather inside the q|k|v GEMM (fast projection path, peer communicator present)."""
        return (self._sp() and not self.qkv_in_one and self.rotary_embedding is not None and not self.key_query_norm
                and self.kv_replicas == 1 and (self.lora_config is None or self.lora_merged_state)
                and all(m.get_bias() is None for m in (self.query, self.key, self.value))
                and get_tp_comm(self.topology, sequence_parallel=True) is not None)

    def _project_qkv_rope(self, x: torch.Tensor, position_ids: Optional[torch.Tensor], sp_shard: bool = False):
        """Fast path for the common layout (separate bias-free q/k/v projections adjacent in the arena, rotary, no
        LoRA / QK-norm): ONE GEMM for q|k|v, then ONE kernel that splits the result and rotates q and k (and one kernel
        back).  Returns None when it does not apply."""
        if (self.qkv_in_one or self.rotary_embedding is None or self.key_query_norm or self.kv_replicas > 1
                or (self.lora_config is not None and not self.lora_merged_state)):
            return None
        if any(m.get_bias() is not None for m in (self.query, self.key, self.value)):
            return None
        rot = self.rotary_embedding
        nh, nkv, d = self.num_attention_heads_per_partition, self.num_kv_heads_per_partition, self.head_dim
        weights = (self.query.weight, self.key.weight, self.value.weight)
        if not OF.qkv_rope_fusable(x, d, rot.dimensions) or OF.fused_view(tuple(w.data for w in weights)) is None:
            return None
        n_out = sum(w.shape[0] for w in weights)
        pos = None if position_ids is None else position_ids.reshape(-1).to(torch.int32).contiguous()
        tp = 1 if self.topology is None else self.topology.config.model_parallel_size
        s_full = x.shape[1] * (tp if sp_shard else 1)
        # rotary in the GEMM epilogue (q and k columns rotated while the accumulator row is in registers)
        in_gemm = OF.rope_in_gemm_supported(x, n_out, d, rot.dimensions, rot.mode)
        rope = (rot.cos_table, rot.sin_table, pos, s_full, (nh + nkv) * d, d, rot.dimensions) if in_gemm else None
        if sp_shard:
            sp_comm = get_tp_comm(self.topology, sequence_parallel=True)
            if not OF.sp_gather_fusable(x, sp_comm, n_out):
                return None
            y = OF.multi_linear_fused(x, weights, sp=sp_comm, rope=rope)  # all-gather(seq) inside the GEMM: [b, s, ΣN]
            if y is None:
                return None
            return OF.qkv_rope_split(y, nh, nkv, d, rot.cos_table, rot.sin_table, pos, s_full, rot.mode, rot.dimensions,
                                     pre_rotated=in_gemm)
        comm = get_tp_comm(self.topology) if x.is_cuda else None
        if comm is None and not self._sp():
            x = copy_to_tensor_model_parallel_region(x, self.topology)
        y = OF.multi_linear_fused(x, weights, comm=comm, rope=rope)
        if y is None:
            return None
        if in_gemm:
            return OF.qkv_rope_split(y, nh, nkv, d, rot.cos_table, rot.sin_table, pos, s_full, rot.mode, rot.dimensions,
                                     pre_rotated=True)
        s = x.shape[1]
        pos = None if position_ids is None else position_ids.reshape(-1).to(torch.int32).contiguous()
        return OF.qkv_rope_split(y, nh, nkv, d, rot.cos_table, rot.sin_table, pos, s, rot.mode, rot.dimensions)

    def _project_qkv(self, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """x: [b, s, h] -> q [T, nh_local, d], k/v [T, nkv_local, d]."""
        b, s, _ = x.shape
        T = b * s
        nh, nkv, d = self.num_attention_heads_per_partition, self.num_kv_heads_per_partition, self.head_dim
        if self.qkv_in_one:
            qkv = self.query_key_value(x).view(T, nh, 3 * d)
            return qkv[..., :d], qkv[..., d:2 * d], qkv[..., 2 * d:]
        comm = get_tp_comm(self.topology) if x.is_cuda else None
        biases = (self.query.get_bias(), self.key.get_bias(), self.value.get_bias())
        if comm is not None and all(bb is None for bb in biases):  # dX all-reduce inside the fused dgrad GEMM
            q, k, v = OF.multi_linear(x, (self.query.weight, self.key.weight, self.value.weight), comm=comm)
            return q.reshape(T, nh, d), k.reshape(T, nkv, d), v.reshape(T, nkv, d)
        if not self._sp():
            x = copy_to_tensor_model_parallel_region(x, self.topology)
        if all(bb is None for bb in biases):
            q, k, v = OF.multi_linear(x, (self.query.weight, self.key.weight, self.value.weight))
        else:
            q, k, v = (OF.linear(x, m.weight, bb) for m, bb in zip((self.query, self.key, self.value), biases))
        return q.reshape(T, nh, d), k.reshape(T, nkv, d), v.reshape(T, nkv, d)

    def _lora_fusable(self) -> bool:
        c = self.lora_config
        return (c is not None and not self.lora_merged_state and not c.bias and not c.dropout and not self._sp()
                and not self.qkv_in_one and self.kv_replicas == 1 and OF.lora_fusion_enabled())

    def _tp_triple(self):
        t = self.topology
        if t is None or t.config.model_parallel_size == 1:
            return (1, 0, None)
        return (t.config.model_parallel_size, t.model_parallel_rank, t.model_parallel_group)

    def _project_qkv_lora(self, x: torch.Tensor, position_ids: Optional[torch.Tensor] = None):
        """q/k/v with their LoRA updates inside the projection GEMMs (low-rank K-slab, OF.lora_linear); None when the
        fused form does not apply (then: plain projections + separate LoRA linears).  Returns (q, k, v, rotated):
        ``rotated`` says that the rotary embedding has already been applied (split + rotate kernel of the fast path)."""
        if not self._lora_fusable() or not x.is_cuda or any(
                m.get_bias() is not None for m in (self.query, self.key, self.value)):
            return None
        b, s, _ = x.shape
        T = b * s
        nh, nkv, d = self.num_attention_heads_per_partition, self.num_kv_heads_per_partition, self.head_dim
        comm = get_tp_comm(self.topology)
        tp = self._tp_triple()
        name = self.lora_config.name
        mods = [self.lora_modules[f"{kind}_{name}"] if f"{kind}_{name}" in self.lora_modules else None
                for kind in ("query", "key", "value")]
        if all(m is not None for m in mods) and len({m.scaling for m in mods}) == 1:
            # frozen, adjacent base weights: q|k|v and their three LoRA pairs as one low-rank GEMM (one dX reduction)
            y = OF.qkv_lora_linear(x, (self.query.weight, self.key.weight, self.value.weight),
                                   tuple((m.dense_in.weight, m.dense_out.weight) for m in mods), mods[0].scaling, tp, comm)
            if y is not None:
                rot = self.rotary_embedding
                if rot is not None and not self.key_query_norm and \
                        OF.qkv_rope_fusable(x, d, rot.dimensions):
                    pos = None if position_ids is None else position_ids.reshape(-1).to(torch.int32).contiguous()
                    return (*OF.qkv_rope_split(y, nh, nkv, d, rot.cos_table, rot.sin_table, pos, s, rot.mode,
                                               rot.dimensions), True)
                q, k, v = y.split([nh * d, nkv * d, nkv * d], dim=-1)
                return q.reshape(T, nh, d), k.reshape(T, nkv, d), v.reshape(T, nkv, d), False
        outs = []
        for kind, base, heads in (("query", self.query, nh), ("key", self.key, nkv), ("value", self.value, nkv)):
            mod = self.lora_modules[f"{kind}_{self.lora_config.name}"] if f"{kind}_{self.lora_config.name}" in self.lora_modules else None
            y = None
            if mod is not None:
                y = OF.lora_linear(x, base.weight, mod.dense_in.weight, mod.dense_out.weight, mod.scaling, "column", tp, comm)
                if y is None:
                    return None
            else:
                y = base(x)
            outs.append(y.reshape(T, heads, d))
        return (*outs, False)

    def forward(self, x: torch.Tensor, cumulative_seq_lengths: Optional[torch.Tensor],
                position_ids: Optional[torch.Tensor], cumulative_seq_lengths_key: Optional[torch.Tensor] = None,
                use_cache: bool = False, reset_cache: bool = False, cache_index: int = 0,
                attention_scores_manipulation: Optional[torch.Tensor] = None,
                attentions_score_manipulation_log_additive: Union[bool, List[bool]] = True,
                sequence_parallel_shard: bool = False) -> torch.Tensor:
        """``sequence_parallel_shard``: ``x`` is the norm's ungathered [b, s/tp, h] output (see
        :meth:`supports_sequence_parallel_shard`); the all-gather then runs inside the projection GEMM."""
        fused = None
        if sequence_parallel_shard:
            fused = self._project_qkv_rope(x, position_ids, sp_shard=True)
            if fused is None:  # shapes do not fit the fused unit: gather the ordinary way and carry on
                x = gather_from_sequence_parallel_region(x, self.topology)
        b, s, _ = x.shape
        if fused is not None:
            s = s * self.topology.config.model_parallel_size
        T = b * s
        nh, d = self.num_attention_heads_per_partition, self.head_dim
        if fused is None:
            fused = self._project_qkv_rope(x, position_ids)
        if fused is not None:
            q, k, v = fused
        else:
            lora_fused = self._project_qkv_lora(x, position_ids)
            rotated = False
            if lora_fused is not None:
                q, k, v, rotated = lora_fused
            else:
                q, k, v = self._project_qkv(x)
                if self.lora_config is not None and not self.lora_merged_state:
                    q, k, v = self._apply_lora(x, q, k, v)

            if self.key_query_norm:
                # every TP rank holds a replica of the per-head norm weights; norm over the local heads is identical math
                q = self.norm_query(q, gather=False)
                k = self.norm_key(k, gather=False)

            if self.rotary_embedding is not None and not rotated:
                pos = None if position_ids is None else position_ids.reshape(-1)
                q = self.rotary_embedding.apply_packed(q, pos, s)
                k = self.rotary_embedding.apply_packed(k, pos, s)

        if cumulative_seq_lengths is None:
            cumulative_seq_lengths = torch.arange(0, T + 1, s, device=x.device, dtype=torch.int32)
        cu_k, max_k = cumulative_seq_lengths_key, None
        if use_cache:
            if not self.causal:
                raise ValueError("KV caching is only supported for causal attention.")
            assert b == 1, f"KV caching is only supported for batch size 1, got {b}"
            if not reset_cache:
                pk, pv = self.cache[cache_index]
                assert pk is not None and pv is not None
                k, v = torch.cat((pk, k), dim=0), torch.cat((pv, v), dim=0)
            self.cache[cache_index] = (k.detach(), v.detach())
            cu_k = torch.tensor([0, k.shape[0]], device=x.device, dtype=torch.int32)
            max_k = k.shape[0]
        elif reset_cache:
            self.cache[cache_index] = (None, None)

        dense_path = (not self.use_flash_attention) or attention_scores_manipulation is not None
        common = dict(cu_seqlens_q=cumulative_seq_lengths, max_seqlen_q=s, causal=self.causal,
                      scale=self.scaling_factor, cu_seqlens_k=cu_k, max_seqlen_k=max_k,
                      dropout_p=self.dropout_attention_probs, training=self.training,
                      deterministic=self.masked_softmax_config.deterministic_flash_attn_bwd)
        if self.dropout_attention_probs > 0.0 and self.training and self.topology is not None and not dense_path:
            # counter-based mask: same on every replay of this call (activation checkpointing), distinct per global head
            t = self.topology
            common["dropout_seed"] = (t.dropout_seed ^ (t.next_dropout_offset() * 0x9E3779B97F4A7C15)) & (2**63 - 1)
            common["head_offset"] = t.model_parallel_rank * nh
        if dense_path:
            manip = attention_scores_manipulation
            if manip is not None and manip.dim() == 4:  # [b, heads|1, s, s] -> [heads|1, T, T] block diagonal
                manip = torch.block_diag(*[manip[i, 0] for i in range(b)])[None] if b > 1 else manip[0]
            log_add = attentions_score_manipulation_log_additive
            log_add = log_add[0] if isinstance(log_add, (list, tuple)) else log_add
            from ...ops.attention import attention_dense

            ctx = attention_dense(q, k, v, cumulative_seq_lengths, cu_k if cu_k is not None else cumulative_seq_lengths,
                                  self.causal, self.scaling_factor, None, softmax_fn=self.masked_softmax,
                                  dropout_fn=self.dropout, scores_manipulation=manip,
                                  manipulation_log_additive=log_add)
        elif 0 < self.num_local_attention_heads < self.num_attention_heads:
            nl = self.num_local_attention_heads
            kk, vv = repeat_kv(k, self.num_repeat_kv), repeat_kv(v, self.num_repeat_kv)
            local = attention_op(q[:, :nl], kk[:, :nl], vv[:, :nl], window=self.local_attention_window_size, **common)
            glob = attention_op(q[:, nl:], kk[:, nl:], vv[:, nl:], **common)
            ctx = torch.cat((local, glob), dim=1)
        else:
            window = self.local_attention_window_size if self.num_local_attention_heads > 0 else None
            ctx = attention_op(q, k, v, window=window, **common)
        ctx = ctx.reshape(b, s, nh * d)

        if self._sp():
            out = OF.linear_reduce_scatter(ctx, self.dense.weight, get_tp_comm(self.topology, sequence_parallel=True)) \
                if ctx.is_cuda else None
            if out is None:
                out = all_reduce_scatter_to_sequence_parallel(self.dense(ctx), self.topology)
            bias = self.dense.get_bias()
            if bias is not None:
                out = out + bias
        else:
            dense_lora = self.lora_modules[f"dense_{self.lora_config.name}"] if (
                self._lora_fusable() and ctx.is_cuda and f"dense_{self.lora_config.name}" in self.lora_modules) else None
            out = None
            if dense_lora is not None:
                # base product and low-rank update are per-rank partial sums of the same all-reduce: one launch, and
                # the attention output is never gathered (reference attention.py:693 all-gathers [b, s, h] here)
                out = OF.lora_linear(ctx, self.dense.weight, dense_lora.dense_in.weight, dense_lora.dense_out.weight,
                                     dense_lora.scaling, "row", self._tp_triple(), get_tp_comm(self.topology))
                if out is not None:
                    bias = self.dense.get_bias()
                    return out if bias is None else out + bias
            out = self.dense(ctx, reduce=True)  # Σ over the ranks' heads (+ bias once), fused into the GEMM launch
        if self.lora_config and not self.lora_merged_state and LoRAModuleType.DENSE in self.lora_config.parallel_modules:
            full = all_concat(ctx, dim=-1, topology=self.topology)
            lora_out = self.lora_modules[f"dense_{self.lora_config.name}"](full)
            if self._sp():
                lora_out = all_shard(lora_out, dim=1, topology=self.topology)
            out = out + lora_out
        return out

    # ------------------------------------------------------------------------------------------------ LoRA
    def _apply_lora(self, x, q, k, v):
        assert self.lora_config is not None
        T = q.shape[0]
        for name, module in self.lora_modules.items():
            kind = module.lora_module_type
            if kind == LoRAModuleType.DENSE:
                continue
            out = module(x)
            heads = self.num_attention_heads_per_partition if kind == LoRAModuleType.QUERY or self.qkv_in_one \
                else self.num_kv_heads_per_partition
            out = out.reshape(T, heads, self.head_dim)
            if kind == LoRAModuleType.QUERY:
                q = q + out
            elif kind == LoRAModuleType.KEY:
                k = k + out
            else:
                v = v + out
        return q, k, v

    @torch.no_grad()
    def merge_lora_weights(self) -> None:
        """Fold every LoRA update into its base weight and drop the adapters.  For fused QKV the q/k/v deltas are
        scattered with the same per-head [q|k|v] interleave the forward pass uses (the reference adds them to
        contiguous thirds — SURVEY §9.2 #25)."""
        assert self.lora_config and hasattr(self, "lora_modules"), "merge of LoRa weights called without LoRA"
        name = self.lora_config.name
        d = self.head_dim
        for key, module in self.lora_modules.items():
            kind = key[: -len(name) - 1]
            delta = module.get_delta_weights()
            if kind == "dense":
                self.dense.weight.data += delta.to(self.dense.weight.dtype)
            elif self.qkv_in_one:
                slot = {"query": 0, "key": 1, "value": 2}[kind]
                w = self.query_key_value.weight.data.view(self.num_attention_heads_per_partition, 3, d, -1)
                w[:, slot] += delta.view(self.num_attention_heads_per_partition, d, -1).to(w.dtype)
            else:
                target = getattr(self, kind)
                target.weight.data += delta.to(target.weight.dtype)
        del self.lora_modules
        self.lora_merged_state = True


class _ReplicatedKV(torch.nn.Module):
    """K or V projection when there are fewer KV heads than TP ranks: rank r holds KV head ``r // replicas``.

    The weight is registered as model parallel along dim 0 with the replicated rows de-duplicated at checkpoint time
    by ``kv_replicas`` (see PipePartitionedModule); gradients of the replicas are averaged by the optimizer through
    ``tied_grad_on_model_parallel``-style reduction over the replica sub-group.
    """

    def __init__(self, hidden_size, head_dim, num_kv_heads, topology, bias, dtype, init_method, bitfit_bias_name):
        super().__init__()
        from .parameter_meta import CoreParameterMeta

        self.topology = topology
        self.replicas = topology.config.model_parallel_size // num_kv_heads
        gen_state = torch.random.get_rng_state()
        # identical init on all replicas of a head: seed from the head index
        head = topology.model_parallel_rank // self.replicas
        torch.manual_seed(1234 + head)
        w = torch.empty(head_dim, hidden_size, dtype=dtype)
        init_method(w)
        torch.random.set_rng_state(gen_state)
        self.weight = torch.nn.Parameter(w.to(topology.device))
        meta = CoreParameterMeta.register_on_parameter(self.weight, is_model_parallel=True, model_parallel_dimension=0)
        meta.kv_replicas = self.replicas  # type: ignore[attr-defined]
        self.bias_name = None
        if bias:
            self.bias_name = "bias" if not bitfit_bias_name else f"bias_{bitfit_bias_name}"
            bparam = torch.nn.Parameter(torch.zeros(head_dim, device=topology.device, dtype=dtype))
            setattr(self, self.bias_name, bparam)
            m2 = CoreParameterMeta.register_on_parameter(bparam, is_model_parallel=True, model_parallel_dimension=0)
            m2.kv_replicas = self.replicas  # type: ignore[attr-defined]

    def get_bias(self):
        return None if self.bias_name is None else getattr(self, self.bias_name)

    def forward(self, x):
        return OF.linear(x, self.weight, self.get_bias())
