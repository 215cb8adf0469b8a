"""Packed-sequence helpers (reference: transformer/data/utils.py:1-108) — vectorised, no per-document Python loops on
device tensors.  Token id 0 is the end-of-document marker, as in the reference (§9.2 #19)."""
from __future__ import annotations

import torch

EOD_TOKEN = 0


def add_cumulative_seq_lengths_padding(cumulative_seq_lengths: torch.Tensor, pad_to: int, padding_value: int = -1) -> torch.Tensor:
    """Pad the boundary vector to a static length so that pipeline messages keep a fixed shape."""
    assert pad_to >= len(cumulative_seq_lengths)
    pad = torch.full((pad_to - len(cumulative_seq_lengths),), padding_value, dtype=torch.int32,
                     device=cumulative_seq_lengths.device)
    return torch.cat((cumulative_seq_lengths.to(torch.int32), pad))


def remove_cumulative_seq_lengths_padding(cumulative_seq_lengths: torch.Tensor, padding_value: int = -1) -> torch.Tensor:
    return cumulative_seq_lengths[cumulative_seq_lengths != padding_value]


def _segment_ends(input_token_ids: torch.Tensor, eod_token: int) -> torch.Tensor:
    """bool [b, s]: positions that close a sequence (an EOD token, or the last position of the row)."""
    ends = input_token_ids == eod_token
    ends[:, -1] = True
    return ends


def get_cumulative_seq_lengths(input_token_ids: torch.Tensor, reset_attention_mask: bool = True,
                               eod_token: int = EOD_TOKEN) -> torch.Tensor:
    """int32 boundaries over the flattened [b·s] tokens: 0, then the index after every sequence end."""
    b, s = input_token_ids.shape
    dev = input_token_ids.device
    if not reset_attention_mask:
        return torch.arange(0, b * s + 1, s, dtype=torch.int32, device=dev)
    flat_ends = _segment_ends(input_token_ids, eod_token).reshape(-1)
    idx = torch.nonzero(flat_ends, as_tuple=False).reshape(-1) + 1
    return torch.cat((torch.zeros(1, dtype=torch.int32, device=dev), idx.to(torch.int32)))


def get_position_ids(input_token_ids: torch.Tensor, reset_position_ids: bool = True, eod_token: int = EOD_TOKEN) -> torch.Tensor:
    """Positions that restart at 0 after every end-of-document token."""
    b, s = input_token_ids.shape
    dev = input_token_ids.device
    ar = torch.arange(s, dtype=torch.long, device=dev).unsqueeze(0).expand(b, s)
    if not reset_position_ids:
        return ar.contiguous()
    ends = input_token_ids == eod_token
    # start index of the segment each position belongs to = 1 + index of the latest EOD strictly before it
    start_after = torch.where(ends, ar + 1, torch.zeros_like(ar))
    start_after = torch.cat((torch.zeros(b, 1, dtype=torch.long, device=dev), start_after[:, :-1]), dim=1)
    seg_start = torch.cummax(start_after, dim=1).values
    return ar - seg_start
