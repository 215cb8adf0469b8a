"""Next-token samplers over logits [batch, seq, vocab] → token id per batch row, taken at the last position
(reference: transformer/inference/sample.py:1-45)."""
from __future__ import annotations

import torch


def _last(logits: torch.Tensor) -> torch.Tensor:
    return logits[:, -1, :].float()


def sample_argmax(logits: torch.Tensor) -> torch.Tensor:
    return _last(logits).argmax(dim=-1)


def fast_multinomial(p: torch.Tensor) -> torch.Tensor:
    """Inverse-CDF draw: first index whose cumulative probability reaches a uniform sample."""
    u = torch.rand(p.shape[:-1], device=p.device, dtype=p.dtype).unsqueeze(-1)
    return (p.cumsum(-1) >= u).to(torch.uint8).argmax(-1)


def sample_temperature(logits: torch.Tensor, temperature: float = 1.0) -> torch.Tensor:
    return fast_multinomial(torch.softmax(_last(logits) / temperature, dim=-1))


def top_k_transform(logits: torch.Tensor, k: int = 10) -> torch.Tensor:
    """Everything below the k-th largest logit (per row) becomes -inf."""
    kth = torch.topk(logits, k, dim=-1).values[..., -1:]
    return logits.masked_fill(logits < kth, float("-inf"))


def top_p_transform(logits: torch.Tensor, threshold: float = 0.95) -> torch.Tensor:
    """Nucleus filtering per row: keep the smallest prefix of the sorted distribution whose mass exceeds
    ``threshold`` (the token that crosses it is kept)."""
    sorted_logits, order = torch.sort(logits, descending=True, dim=-1)
    mass = torch.softmax(sorted_logits.float(), dim=-1).cumsum(-1)
    drop_sorted = mass > threshold
    drop_sorted = torch.cat([torch.zeros_like(drop_sorted[..., :1]), drop_sorted[..., :-1]], dim=-1)
    drop = torch.zeros_like(drop_sorted).scatter(-1, order, drop_sorted)
    return logits.masked_fill(drop, float("-inf"))


def sample_top_k(logits: torch.Tensor, k: int = 10, temperature: float = 1.0) -> torch.Tensor:
    return sample_temperature(top_k_transform(logits, k), temperature)


def sample_top_p(logits: torch.Tensor, threshold: float = 0.95, temperature: float = 1.0) -> torch.Tensor:
    return sample_temperature(top_p_transform(logits, threshold), temperature)
