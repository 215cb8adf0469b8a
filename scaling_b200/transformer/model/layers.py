"""Transformer layers (reference: transformer/model/layers/{base,embedding,layer,layernorm,lm_head,lm_head_tied,
embedding_head}.py).  Class names and parameter names are part of the checkpoint file names / keys and are kept.

Layer math is the reference's pre-norm block — ``x + drop(attn(norm(x)))`` [+ adapter], ``x + drop(mlp(norm(x)))``
[+ adapter] — arranged so that the residual add in front of the second norm is fused into the norm kernel, the LM head
keeps its logits sharded over the vocabulary for the fused vocab-parallel cross-entropy while training, and dropout is
skipped outright when its probability is zero.
"""
from __future__ import annotations

import collections
from functools import partial
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import torch

from ...core import (
    BaseLayer,
    BaseLayerIO,
    ColumnParallelLinear,
    ParallelMLP,
    ParallelSelfAttention,
    ParallelSwiGLUMLP,
    RotaryConfig,
    VocabParallelEmbedding,
    get_norm,
)
from ...core.nn.tp_ops import all_concat, copy_to_tensor_model_parallel_region
from ...parallel.symm import get_tp_comm
from ...core.topology import Topology
from ...ops import functional as OF
from ..config import MLPType, TransformerArchitectureConfig
from ..data.inference_settings import InferenceSettings
from ..data.text_dataset_batch import TextDatasetBatch, _FieldTuple

InitFn = Callable[[torch.Tensor], torch.Tensor]


class TransformerLayerIO(_FieldTuple, BaseLayerIO):
    FIELDS = ("activations", "position_ids", "cumulative_seq_lengths", "cumulative_seq_lengths_padded", "loss_weights",
              "inference_settings", "embeddings", "embeddings_head", "attention_scores_manipulation")

    def __init__(self, activations: torch.Tensor, position_ids: torch.Tensor,
                 cumulative_seq_lengths_padded: Optional[torch.Tensor], cumulative_seq_lengths: Optional[torch.Tensor] = None,
                 loss_weights: Optional[torch.Tensor] = None, inference_settings: Optional[InferenceSettings] = None,
                 embeddings: Optional[torch.Tensor] = None, embeddings_head: Optional[torch.Tensor] = None,
                 attention_scores_manipulation: Optional[torch.Tensor] = None):
        self.activations = activations
        self.position_ids = position_ids
        self.cumulative_seq_lengths = cumulative_seq_lengths
        self.cumulative_seq_lengths_padded = cumulative_seq_lengths_padded
        self.loss_weights = loss_weights
        self.inference_settings = inference_settings
        self.embeddings = embeddings
        self.embeddings_head = embeddings_head
        self.attention_scores_manipulation = attention_scores_manipulation
        # set by the LM head while training: logits are this rank's vocabulary shard (start index, TP group)
        self.vocab_parallel_logits: Optional[Tuple[int, Any]] = None

    def derive(self, activations: torch.Tensor, **overrides: Any) -> "TransformerLayerIO":
        kw = dict(position_ids=self.position_ids, cumulative_seq_lengths=self.cumulative_seq_lengths,
                  cumulative_seq_lengths_padded=self.cumulative_seq_lengths_padded, loss_weights=self.loss_weights,
                  inference_settings=self.inference_settings, embeddings=self.embeddings)
        kw.update(overrides)
        return TransformerLayerIO(activations=activations, **kw)


class TransformerLayerBaseIO(BaseLayer[TransformerLayerIO, TransformerLayerIO, TransformerLayerIO]):
    @staticmethod
    def input_to_tuple(input: TransformerLayerIO) -> Tuple[Any, ...]:
        return input.as_tuple()

    @staticmethod
    def tuple_to_input(d: Tuple[Any, ...]) -> TransformerLayerIO:
        io = TransformerLayerIO.from_tuple(d)
        if io.cumulative_seq_lengths is None and io.cumulative_seq_lengths_padded is not None:
            from ..data.utils import remove_cumulative_seq_lengths_padding

            io.cumulative_seq_lengths = remove_cumulative_seq_lengths_padding(io.cumulative_seq_lengths_padded)
        return io

    @staticmethod
    def output_to_tuple(output: TransformerLayerIO) -> Tuple[Any, ...]:
        # the unpadded boundaries have a data-dependent length; only the padded copy travels between stages
        if output.cumulative_seq_lengths_padded is not None:
            output = output.derive(output.activations, cumulative_seq_lengths=None,
                                   attention_scores_manipulation=output.attention_scores_manipulation)
        return output.as_tuple()

    @staticmethod
    def tuple_to_last_stage_activation(d: Tuple[Any, ...]) -> TransformerLayerIO:
        return TransformerLayerBaseIO.tuple_to_input(d)


def _device(topology: Optional[Topology]) -> torch.device:
    if topology is not None:
        return topology.device
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def _dropout(x: torch.Tensor, p: float, training: bool, topology: Optional[Topology]) -> torch.Tensor:
    """Dropout whose mask is identical on every TP rank (drawn from the TP-constant RNG stream); free when p == 0."""
    if p <= 0.0 or not training:
        return x
    if topology is not None and OF.native_enabled(x) and x.dtype == torch.bfloat16 and x.numel() % 4 == 0:
        # counter-based (Philox) mask: no RNG state to fork, identical on all TP ranks and under recomputation
        return OF.bias_dropout_add(x, None, None, p, True, topology.dropout_seed, topology.next_dropout_offset(x.numel()))
    if topology is not None:
        with topology.model_parallel_constant_rng():
            return torch.nn.functional.dropout(x, p=p, training=True)
    return torch.nn.functional.dropout(x, p=p, training=True)


def _record_embedding(x: TransformerLayerIO, slot_layer: int, activations: torch.Tensor) -> None:
    s = x.inference_settings
    if s is not None and slot_layer in s.embedding_layers and x.embeddings is not None:
        x.embeddings[s.embedding_layers.index(slot_layer)] = activations


# ----------------------------------------------------------------------------------------------------------------
class EmbeddingInput(BaseLayer[TextDatasetBatch, TransformerLayerIO, TransformerLayerIO]):
    """Token embedding (+ optional image-encoder splice, soft-prompt prefix and attention-manipulation set-up)."""

    def __init__(self, architecture_config: TransformerArchitectureConfig, topology: Optional[Topology] = None,
                 init_method: InitFn = torch.nn.init.xavier_normal_):
        super().__init__()
        a = architecture_config
        self.architecture_config = a
        self.topology = topology
        self.embedding = VocabParallelEmbedding(a.vocab_size, a.hidden_size, topology=topology, dtype=a.precision.dtype,
                                                init_method=init_method, finetunable_token_ids=a.finetunable_token_ids,
                                                device=None if topology is not None else _device(None))
        self.dropout_p = a.dropout_embedding
        if a.image_encoder:
            from .image_encoder import ImageEncoder

            self.image_encoder = ImageEncoder(out_features=a.hidden_size, dropout_p=a.dropout_image_encoder,
                                              layernorm_config=a.layernorm, dtype=a.precision.dtype,
                                              device=_device(topology))
        self.softprompt_name: Optional[str] = None
        if a.softprompt_config is not None:
            self.softprompt_name = a.softprompt_config.name
            p = torch.nn.Parameter(torch.zeros(a.softprompt_config.n_tokens, a.hidden_size, dtype=a.precision.dtype,
                                               device=_device(topology)))
            init_method(p)
            setattr(self, f"softprompt_{self.softprompt_name}", p)
        self.cache: Dict[int, Optional[torch.Tensor]] = {}

    @staticmethod
    def estimate_parameter_count(architecture_config: TransformerArchitectureConfig, topology=None, **_: Any) -> int:
        tp = 1 if topology is None else topology.config.model_parallel_size
        return architecture_config.vocab_size * architecture_config.hidden_size // tp

    def forward(self, x: TextDatasetBatch) -> TransformerLayerIO:
        s = x.inference_settings
        use_cache, reset_cache, cache_index = (False, False, 0) if s is None else (s.use_cache, s.reset_cache, s.cache_index)
        if reset_cache:
            self.cache[cache_index] = None
        assert x.input_token_ids is not None
        h = self.embedding(x.input_token_ids)
        h = _dropout(h, self.dropout_p, self.training, self.topology)
        if x.input_images is not None:
            h = self._splice_images(x, h)
        if self.softprompt_name is not None:
            sp = getattr(self, f"softprompt_{self.softprompt_name}")
            h = torch.cat((sp.unsqueeze(0).expand(h.shape[0], -1, -1), h[:, sp.shape[0]:, :]), dim=1).contiguous()
        if s is not None and 0 in s.embedding_layers and x.embeddings is not None:
            x.embeddings[0] = h
        loss_weights = x.loss_weights
        manipulation = None
        if s is not None and s.inference_control_parameters is not None:
            manipulation, loss_weights = self._attention_manipulation(s, h, loss_weights, use_cache, reset_cache, cache_index)
        return TransformerLayerIO(activations=h, position_ids=x.position_ids, cumulative_seq_lengths=x.cumulative_seq_lengths,
                                  cumulative_seq_lengths_padded=x.cumulative_seq_lengths_padded,
                                  attention_scores_manipulation=manipulation, loss_weights=loss_weights,
                                  inference_settings=s, embeddings=x.embeddings)

    def _splice_images(self, x: TextDatasetBatch, h: torch.Tensor) -> torch.Tensor:
        rng = self.topology.model_parallel_constant_rng() if self.topology is not None else None
        if rng is not None:
            with rng:
                img = self.image_encoder(x.input_images)
        else:
            img = self.image_encoder(x.input_images)
        locations = x.input_image_locations
        if locations is None and not self.training:
            assert x.inference_settings is not None and x.inference_settings.input_image_locations is not None
            locations = x.inference_settings.input_image_locations
        if locations is not None:
            h = h.clone()
            locs = locations.tolist() if torch.is_tensor(locations) else locations
            assert len(locs) == img.shape[0]
            for emb, (b, start, end) in zip(img, locs):
                h[b, start:end] = emb
            return h
        # pre-training: one image per sample, placed at the beginning (cat keeps the graph differentiable)
        return torch.cat((img, h[:, img.shape[1]:, :]), dim=1).contiguous()

    # -- "AtMan" attention manipulation: per-token factors on the attention scores, optionally spread to tokens whose
    #    embeddings are cosine-similar to the controlled ones (reference embedding.py:166-278) ------------------------
    @staticmethod
    def get_control_factor_from_cosine_similarity(control_factor: float, cosine_similarity: float) -> float:
        if 0 <= cosine_similarity <= 1.0:
            return (1 - control_factor) * (1 - cosine_similarity) + control_factor
        return 1.0

    @staticmethod
    def get_similarity_matrix(a: torch.Tensor, b: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
        an = a / a.norm(dim=1, keepdim=True).clamp_min(eps)
        bn = b / b.norm(dim=1, keepdim=True).clamp_min(eps)
        return an @ bn.t()

    def get_embedding_similarity_matrix(self, embeddings: torch.Tensor) -> torch.Tensor:
        assert embeddings.ndim == 3
        with torch.no_grad():
            e = embeddings.float()
            return torch.stack([self.get_similarity_matrix(x, x) for x in e]).clip(-1, 1).cpu()

    def _attention_manipulation(self, s: InferenceSettings, h: torch.Tensor, loss_weights, use_cache, reset_cache, cache_index):
        params = s.inference_control_parameters
        assert len(params) == h.shape[0], "number of inference_control_parameters does not match batch size"
        b, n = h.shape[0], h.shape[1]
        m = torch.zeros(b, 1, n, n, device=h.device, dtype=h.dtype)
        for i, p in enumerate(params):
            if not p.control_log_additive:
                m[i] = 1.0
        sim = None
        if any(p.contextual_control_threshold is not None for p in params):
            ref = h
            if use_cache:
                if not reset_cache and self.cache.get(cache_index) is not None:
                    ref = torch.cat((self.cache[cache_index], ref), dim=1)
                self.cache[cache_index] = ref
            sim = self.get_embedding_similarity_matrix(ref)
        for i, p in enumerate(params):
            if p is None or p.controls is None or all(c.token_index == -1 for c in p.controls):
                continue
            factors: Dict[int, float] = collections.defaultdict(float)
            for c in p.controls:
                if c.token_index < 0:
                    continue
                factors[c.token_index] = c.factor
                if p.contextual_control_threshold is not None:
                    scores = sim[i][c.token_index]
                    for j in (scores >= p.contextual_control_threshold).nonzero().view(-1).tolist():
                        if j == c.token_index:
                            continue
                        f = self.get_control_factor_from_cosine_similarity(c.factor, float(scores[j]))
                        factors[j] = min(f, factors[j])
            for idx, f in factors.items():
                if reset_cache and loss_weights is not None:
                    loss_weights[i, idx] = loss_weights[i, idx] * f
                if p.control_log_additive:
                    m[i, :, :, idx] = -10000.0 if f == 0.0 else float(torch.log(torch.tensor(f)))
                else:
                    m[i, :, :, idx] = f
        return m, loss_weights

    @staticmethod
    def input_to_tuple(input: TextDatasetBatch) -> Tuple[Any, ...]:
        return input.as_tuple()

    @staticmethod
    def tuple_to_input(d: Tuple[Any, ...]) -> TextDatasetBatch:
        return TextDatasetBatch.from_tuple(d)

    @staticmethod
    def output_to_tuple(output: TransformerLayerIO) -> Tuple[Any, ...]:
        return TransformerLayerBaseIO.output_to_tuple(output)

    @staticmethod
    def tuple_to_last_stage_activation(d: Tuple[Any, ...]) -> TransformerLayerIO:
        return TransformerLayerBaseIO.tuple_to_input(d)


# ----------------------------------------------------------------------------------------------------------------
class TransformerLayer(TransformerLayerBaseIO):
    def __init__(self, architecture_config: TransformerArchitectureConfig, layer_index: int,
                 topology: Optional[Topology] = None, init_method: InitFn = torch.nn.init.xavier_normal_):
        super().__init__()
        a = architecture_config
        self.architecture_config = a
        self.topology = topology
        self.layer_index = layer_index
        bitfit = getattr(a.bitfit_bias_config, "name", None)
        dev = None if topology is not None else _device(None)
        norm_kw = dict(norm_type=a.norm_type, layernorm_config=a.layernorm, dimensions=a.hidden_size, device=dev,
                       dtype=a.precision.dtype, bitfit_bias_name=bitfit, topology=topology)
        self.input_layernorm = get_norm(**norm_kw)
        head_dim = a.hidden_size // a.num_attention_heads
        self.self_attention = ParallelSelfAttention(
            hidden_size=a.hidden_size, num_attention_heads=a.num_attention_heads,
            num_local_attention_heads=a.num_local_attention_heads, local_attention_window_size=a.local_attention_window_size,
            masked_softmax_config=a.masked_softmax, causal=a.causal, dropout_attention_probs=a.dropout_attention_probs,
            rotary_config=RotaryConfig(dimensions=int(a.rotary_percentage * head_dim), max_seq_length=a.sequence_length,
                                       base=a.rotary_embedding_base),
            relative_position_embedding_type=a.relative_position_embedding_type, bias=a.attention_bias, topology=topology,
            device=dev, dtype=a.precision.dtype, bitfit_bias_name=bitfit, init_method=init_method, lora_config=a.lora_config,
            norm_type=a.norm_type, key_query_norm=a.key_query_norm, layernorm_config=a.layernorm,
            qkv_in_one=a.attention_qkv_in_one, num_kv_heads=a.attention_num_kv_heads, use_matmul=a.attention_use_matmul)
        self.post_attention_layernorm = get_norm(**norm_kw)
        mlp_cls = ParallelMLP if a.mlp_type == MLPType.DEFAULT else ParallelSwiGLUMLP
        self.mlp = mlp_cls(io_features=a.hidden_size, intermediate_feature_factor=a.mlp_factor, bias=a.mlp_bias,
                           topology=topology, device=dev, dtype=a.precision.dtype, bitfit_bias_name=bitfit,
                           init_method=init_method)
        self.p_attn, self.p_mlp = a.dropout_after_attention, a.dropout_after_mlp
        if a.adapter_config is not None:
            self.load_adapter()

    @staticmethod
    def estimate_parameter_count(architecture_config: TransformerArchitectureConfig, topology=None, **_: Any) -> int:
        a = architecture_config
        tp = 1 if topology is None else topology.config.model_parallel_size
        h = a.hidden_size
        kv = h if not a.attention_num_kv_heads else h * a.attention_num_kv_heads // a.num_attention_heads
        f = int(h * a.mlp_factor)
        n = (h * h + 2 * h * kv + h * h + (3 if a.mlp_type == MLPType.SWIGLU else 2) * h * f) // tp
        return n + 2 * h

    def load_adapter(self) -> None:
        """Bias-free bottleneck MLPs after the attention and/or MLP block (reference layer.py:143-186)."""
        a, c = self.architecture_config, self.architecture_config.adapter_config
        make = lambda factor: ParallelMLP(io_features=a.hidden_size, intermediate_feature_factor=factor, bias=False,  # noqa: E731
                                          topology=self.topology, dtype=a.precision.dtype,
                                          device=None if self.topology is not None else _device(None),
                                          init_method=partial(torch.nn.init.normal_, mean=0.0, std=c.init_std))
        if c.attention_downsampling_factor is not None:
            self.attn_adapter_name = f"attn_adapter_{c.name}"
            setattr(self, self.attn_adapter_name, make(c.attention_downsampling_factor))
        if c.mlp_downsampling_factor is not None:
            self.mlp_adapter_name = f"mlp_adapter_{c.name}"
            setattr(self, self.mlp_adapter_name, make(c.mlp_downsampling_factor))

    def _adapter(self, h: torch.Tensor, name: str) -> torch.Tensor:
        """Adapters read the full sequence; under SP their input is gathered and their output re-sharded — the
        reference feeds them the sequence shard unchanged."""
        adapter = getattr(self, name)
        sp = self.topology is not None and self.topology.config.sequence_parallel
        if sp:
            from ...core.nn.tp_ops import gather_from_sequence_parallel_region

            h = gather_from_sequence_parallel_region(h, self.topology)
        return adapter(h)

    # The two halves of the block as separate calls (public in the reference, layer.py:189-239; its backward-compatibility
    # test drives them one by one).  ``forward`` below is the fused form of ``mlp_block(attention_block(h))``.
    def attention_block(self, hidden_state: torch.Tensor, cumulative_seq_lengths: Optional[torch.Tensor],
                        position_ids: Optional[torch.Tensor], use_cache: bool = False, reset_cache: bool = False,
                        cache_index: int = 0, attention_scores_manipulation: Optional[torch.Tensor] = None,
                        attentions_score_manipulation_log_additive: Any = True) -> torch.Tensor:
        a = self.self_attention(self.input_layernorm(hidden_state), cumulative_seq_lengths=cumulative_seq_lengths,
                                position_ids=position_ids, use_cache=use_cache, reset_cache=reset_cache,
                                cache_index=cache_index, attention_scores_manipulation=attention_scores_manipulation,
                                attentions_score_manipulation_log_additive=attentions_score_manipulation_log_additive)
        h = hidden_state + _dropout(a, self.p_attn, self.training, self.topology)
        if hasattr(self, "attn_adapter_name"):
            h = h + self._adapter(h, self.attn_adapter_name)
        return h

    def mlp_block(self, hidden_state: torch.Tensor) -> torch.Tensor:
        m = self.mlp(self.post_attention_layernorm(hidden_state))
        h = hidden_state + _dropout(m, self.p_mlp, self.training, self.topology)
        if hasattr(self, "mlp_adapter_name"):
            h = h + self._adapter(h, self.mlp_adapter_name)
        return h

    def forward(self, x: TransformerLayerIO) -> TransformerLayerIO:
        s = x.inference_settings
        use_cache, reset_cache, cache_index, log_add = (False, False, 0, True) if s is None else \
            (s.use_cache, s.reset_cache, s.cache_index, s.control_log_additive_batch)
        h = x.activations
        # sequence parallel on GPU: hand the norms' UNGATHERED shards to attention / MLP, whose first GEMM gathers them
        # over NVLink tile by tile (all-gather -> GEMM as one unit) instead of an all-gather in between
        shard_attn = h.is_cuda and getattr(self.self_attention, "supports_sequence_parallel_shard", lambda: False)()
        shard_mlp = h.is_cuda and getattr(self.mlp, "supports_sequence_parallel_shard", lambda: False)()
        extra = {"sequence_parallel_shard": True} if shard_attn else {}
        a = self.self_attention(self.input_layernorm(h, gather=not shard_attn),
                                cumulative_seq_lengths=x.cumulative_seq_lengths,
                                position_ids=x.position_ids, use_cache=use_cache, reset_cache=reset_cache,
                                cache_index=cache_index, attention_scores_manipulation=x.attention_scores_manipulation,
                                attentions_score_manipulation_log_additive=log_add, **extra)
        a = _dropout(a, self.p_attn, self.training, self.topology)
        if hasattr(self, "attn_adapter_name"):
            h = h + a
            h = h + self._adapter(h, self.attn_adapter_name)
            n = self.post_attention_layernorm(h, gather=not shard_mlp)
        else:
            n, h = self.post_attention_layernorm(a, residual=h, gather=not shard_mlp)  # fused: h <- h + a, n <- norm(h)
        m = self.mlp(n, sequence_parallel_shard=True) if shard_mlp else self.mlp(n)
        m = _dropout(m, self.p_mlp, self.training, self.topology)
        h = h + m
        if hasattr(self, "mlp_adapter_name"):
            h = h + self._adapter(h, self.mlp_adapter_name)
        _record_embedding(x, self.layer_index + 1, h)
        return x.derive(h, attention_scores_manipulation=x.attention_scores_manipulation)


class LayerNormWrapper(TransformerLayerBaseIO):
    def __init__(self, architecture_config: TransformerArchitectureConfig, layer_index: int, topology: Optional[Topology] = None):
        super().__init__()
        a = architecture_config
        self.topology = topology
        self.architecture_config = a
        self.layer_index = layer_index
        self.norm = get_norm(norm_type=a.norm_type, layernorm_config=a.layernorm, dimensions=a.hidden_size,
                             device=None if topology is not None else _device(None), dtype=a.precision.dtype,
                             bitfit_bias_name=getattr(a.bitfit_bias_config, "name", None), topology=topology)

    @staticmethod
    def estimate_parameter_count(architecture_config: TransformerArchitectureConfig, **_: Any) -> int:
        return architecture_config.hidden_size

    def forward(self, x: TransformerLayerIO) -> TransformerLayerIO:
        h = self.norm(x.activations)
        _record_embedding(x, self.layer_index + 1, h)
        return x.derive(h)


def _row_mask(token_ids: List[int], start: int, rows: int, like: torch.Tensor) -> torch.Tensor:
    mask = torch.zeros(rows, 1, dtype=like.dtype, device=like.device)
    for t in token_ids:
        if start <= t < start + rows:
            mask[t - start] = 1
    return mask


class _LMHeadBase(TransformerLayerBaseIO):
    topology: Optional[Topology]

    def _weight(self) -> torch.Tensor:
        raise NotImplementedError

    def forward(self, x: TransformerLayerIO) -> TransformerLayerIO:
        """Logits over this rank's vocabulary shard while training on the fused cross-entropy path; full logits
        (all-gathered like the reference, lm_head.py:56 / lm_head_tied.py:37-43) otherwise."""
        topo = self.topology
        tp = 1 if topo is None else topo.config.model_parallel_size
        h = x.activations
        w = self._weight()
        comm = get_tp_comm(topo) if h.is_cuda else None
        training = self.training and torch.is_grad_enabled()
        fused = None
        if training and (tp == 1 or comm is not None):
            # CE statistics in the GEMM epilogue: the loss never re-reads the logits
            fused = OF.linear_with_xent_stats(h, w, comm=comm)
        if fused is not None:
            logits, partial = fused
            out = x.derive(logits)
            start = 0 if topo is None else topo.model_parallel_rank * w.shape[0]
            out.vocab_parallel_logits = (start, None if tp == 1 else topo.model_parallel_group, partial, comm)
            return out
        if comm is not None:
            logits = OF.linear(h, w, None, comm=comm, reduce_input_grad=True)
        else:
            if tp > 1 and not topo.config.sequence_parallel:
                h = copy_to_tensor_model_parallel_region(h, topo)
            logits = OF.linear(h, w, None)
        out = x.derive(logits)
        keep_sharded = self.training and torch.is_grad_enabled() and OF.native_enabled(logits)
        if keep_sharded:
            start = 0 if topo is None else topo.model_parallel_rank * w.shape[0]
            out.vocab_parallel_logits = (start, None if tp == 1 else topo.model_parallel_group)
        elif tp > 1:
            out.activations = all_concat(logits, dim=-1, topology=topo)
        return out


class TransformerLMHead(_LMHeadBase):
    def __init__(self, architecture_config: TransformerArchitectureConfig, init_method: InitFn = torch.nn.init.xavier_normal_,
                 topology: Optional[Topology] = None):
        super().__init__()
        a = architecture_config
        self.topology = topology
        self.linear = ColumnParallelLinear(a.hidden_size, a.vocab_size, bias=False, topology=topology,
                                           device=None if topology is not None else _device(None),
                                           dtype=a.precision.dtype, init_method=init_method, parallel_output=True)
        if a.finetunable_token_ids:
            rows = self.linear.weight.shape[0]
            start = 0 if topology is None else topology.model_parallel_rank * rows
            mask = _row_mask(a.finetunable_token_ids, start, rows, self.linear.weight)
            self.linear.weight._sb_no_main_grad = True  # masked rows need the autograd hook below
            self.linear.weight.register_hook(lambda g: None if g is None else g * mask.to(g.dtype))

    @staticmethod
    def estimate_parameter_count(architecture_config: TransformerArchitectureConfig, topology=None, **_: Any) -> int:
        tp = 1 if topology is None else topology.config.model_parallel_size
        return architecture_config.vocab_size * architecture_config.hidden_size // tp

    def _weight(self) -> torch.Tensor:
        return self.linear.weight


class TransformerLMHeadTied(_LMHeadBase):
    def __init__(self, architecture_config: TransformerArchitectureConfig, init_method: InitFn = torch.nn.init.xavier_normal_,
                 topology: Optional[Topology] = None):
        super().__init__()
        a = architecture_config
        self.topology = topology
        self.embedding = VocabParallelEmbedding(a.vocab_size, a.hidden_size, topology=topology, dtype=a.precision.dtype,
                                                device=None if topology is not None else _device(None),
                                                init_method=init_method, finetunable_token_ids=a.finetunable_token_ids)

    @staticmethod
    def estimate_parameter_count(architecture_config: TransformerArchitectureConfig, topology=None, **_: Any) -> int:
        tp = 1 if topology is None else topology.config.model_parallel_size
        return architecture_config.vocab_size * architecture_config.hidden_size // tp

    def _weight(self) -> torch.Tensor:
        return self.embedding.weight


class TransformerEmbeddingHead(TransformerLayerBaseIO):
    """Position-weighted mean pooling (fp32) and an optional projection stack (reference embedding_head.py)."""

    def __init__(self, architecture_config: TransformerArchitectureConfig, topology: Optional[Topology] = None):
        super().__init__()
        a = architecture_config
        assert a.embedding_head_config is not None
        self.topology = topology
        self.embedding_head_config = a.embedding_head_config
        width = a.hidden_size
        for i, out in enumerate(self.embedding_head_config.proj_layers):
            setattr(self, f"embedding_head_proj_{self.embedding_head_config.name}_{i}",
                    torch.nn.Linear(width, out, bias=False, device=_device(topology), dtype=a.precision.dtype))
            width = out

    @staticmethod
    def weighted_mean_pooling(embeddings: torch.Tensor, loss_weights: torch.Tensor) -> torch.Tensor:
        e = embeddings.float()
        pos = torch.arange(1, e.shape[1] + 1, dtype=e.dtype, device=e.device).view(1, -1, 1)
        w = loss_weights.to(e.dtype).unsqueeze(-1) * pos
        den = w.sum(dim=1)
        if float(den.sum()) == 0.0:
            return torch.zeros_like(e[:, 0]).to(embeddings.dtype)
        return ((e * w).sum(dim=1) / den).to(embeddings.dtype)

    def apply_embedding_head_proj(self, embeddings: torch.Tensor) -> torch.Tensor:
        for i in range(len(self.embedding_head_config.proj_layers)):
            if i > 0:
                embeddings = torch.nn.functional.gelu(embeddings)
            embeddings = getattr(self, f"embedding_head_proj_{self.embedding_head_config.name}_{i}")(embeddings)
        return embeddings

    def forward(self, x: TransformerLayerIO) -> TransformerLayerIO:
        assert x.loss_weights is not None, "did not receive loss_weights for masking"
        pooled = self.weighted_mean_pooling(x.activations, x.loss_weights)
        if self.embedding_head_config.proj_layers:
            pooled = self.apply_embedding_head_proj(pooled)
        return x.derive(pooled)
