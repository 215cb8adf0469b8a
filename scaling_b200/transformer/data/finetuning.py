"""Supervised-finetuning datasets (reference: transformer/data/finetuning_text_dataset.py:1-428,
finetuning_chat_dataset.py:1-365).

Two on-disk formats, one sample pipeline:

* :class:`FinetuningTextDataset` — prompt/completion pairs.  Either a JSON list of ``{"prompt", "completion"}``
  (``prompt`` may be a list mixing text pieces and ``*.jpg`` paths) tokenised on the fly, or a memory map written by
  :meth:`FinetuningTextDataset.convert_jsonl` whose documents are ``[len(prompt), *prompt, *completion]``.
  Loss is taken on the completion and on exactly one EOS after it.
* :class:`FinetuningChatDataset` — jsonl, one conversation per line: a list of ``{"type": "text"|"image",
  "content", "has_loss"}`` turns.  Loss follows ``has_loss``; no EOS is added (a warning says so once).

Both reduce a record to a :class:`_Sample` (token ids, per-token loss flag, image paths and their placeholder spans)
and share padding / soft-prompt shifting / collation / TP broadcast.  Every sample is one sequence of exactly
``sequence_length`` tokens padded with EOS (documents are never packed), an image occupies 144 placeholder tokens
that ``EmbeddingInput`` overwrites with the image encoder's output.
"""
from __future__ import annotations

import hashlib
import json
import random
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Optional, Tuple, Union

import numpy as np
import torch

from ...core import (
    BaseBlendedDataset,
    BaseDataset,
    BaseDatasetItem,
    MemoryMapDataset,
    MemoryMapDatasetBuilder,
    Topology,
    broadcast_data,
)
from ..model.image_encoder import IMAGE_ENCODER_TOKEN_COUNTS, IMAGE_SIZE, clip_transform
from .text_dataset_batch import TextDatasetBatch
from .utils import get_cumulative_seq_lengths

IMAGE_TRANSFORM_FN = clip_transform(IMAGE_SIZE)


class FinetuningTextDatasetItem(BaseDatasetItem):
    def __init__(self, input_token_ids: torch.Tensor, target_token_ids: torch.Tensor, cumulative_seq_lengths: torch.Tensor,
                 position_ids: torch.Tensor, loss_weights: torch.Tensor, input_images: Optional[List[torch.Tensor]] = None,
                 input_image_locations: Optional[List[Tuple[int, int]]] = None):
        self.input_token_ids = input_token_ids
        self.target_token_ids = target_token_ids
        self.cumulative_seq_lengths = cumulative_seq_lengths
        self.position_ids = position_ids
        self.loss_weights = loss_weights
        self.input_images = input_images
        self.input_image_locations = input_image_locations


@dataclass
class _Sample:
    """Unpadded token stream of one record: ``tokens[i]`` predicts ``tokens[i+1]``; ``loss[i]`` flags whether token i
    is a training target."""

    tokens: List[int]
    loss: List[int]
    image_paths: List[Path] = field(default_factory=list)
    image_spans: List[Tuple[int, int]] = field(default_factory=list)


def _vocab_digest(md5: "hashlib._Hash", tokenizer: Any) -> None:
    md5.update(json.dumps(tokenizer.tokenizer.get_vocab(), sort_keys=True, default=str).encode("utf-8"))


class _FinetuningBase(BaseDataset[FinetuningTextDatasetItem, TextDatasetBatch, TextDatasetBatch]):
    sequence_length: int
    softprompt_n_tokens: int
    data: List[Any]

    def _eos(self) -> int:
        return int(self.tokenizer.eos_token_id)  # type: ignore[attr-defined]

    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        if self.seed is not None and self.seed == seed:
            return
        if shuffle:
            random.Random(seed).shuffle(self.data)
        self.seed = seed

    def __len__(self) -> int:
        return len(self.data)

    # ---- sample → fixed-length item -----------------------------------------------------------------------
    def _finish(self, s: _Sample, loss_on_padding_eos: bool) -> FinetuningTextDatasetItem:
        L, eos, sp = self.sequence_length, self._eos(), self.softprompt_n_tokens
        tokens, loss = s.tokens, s.loss
        spans = s.image_spans
        if sp > 0:  # soft-prompt slots are prepended; they never carry loss and push the image spans right
            tokens = [0] * sp + tokens
            loss = [0] * sp + loss
            spans = [(a + sp, b + sp) for a, b in spans]
        n = len(tokens)
        ids = (tokens + [eos] * max(0, L + 1 - n))[:L + 1]
        tgt_flag = (loss + [0] * max(0, L + 1 - n))[:L + 1]
        if loss_on_padding_eos and n < L + 1:
            tgt_flag[n] = 1  # the first padding EOS is a real target: the model must learn to stop
        weights = torch.tensor(tgt_flag[1:], dtype=torch.float)  # weight i belongs to target i = token i+1
        images = [IMAGE_TRANSFORM_FN(_open_image(p)) for p in s.image_paths] if s.image_paths else None
        return FinetuningTextDatasetItem(
            input_token_ids=torch.tensor(ids[:-1], dtype=torch.long),
            target_token_ids=torch.tensor(ids[1:], dtype=torch.long),
            cumulative_seq_lengths=torch.tensor([0, L], dtype=torch.int32),
            position_ids=torch.arange(0, L),
            loss_weights=weights,
            input_images=images,
            input_image_locations=list(spans) if s.image_paths else None,
        )

    # ---- batch ----------------------------------------------------------------------------------------------
    def collate(self, batch: List[FinetuningTextDatasetItem]) -> TextDatasetBatch:
        input_token_ids = torch.stack([b.input_token_ids for b in batch])
        images: List[torch.Tensor] = []
        locations: List[List[int]] = []
        for row, item in enumerate(batch):
            imgs, spans = item.input_images or [], item.input_image_locations or []
            assert len(imgs) == len(spans)
            for img, (start, end) in zip(imgs, spans):
                images.append(img)
                locations.append([row, start, end])
        return TextDatasetBatch(
            input_token_ids=input_token_ids,
            target_token_ids=torch.stack([b.target_token_ids for b in batch]),
            cumulative_seq_lengths=get_cumulative_seq_lengths(input_token_ids, reset_attention_mask=False),
            position_ids=torch.stack([b.position_ids for b in batch]),
            loss_weights=torch.stack([b.loss_weights for b in batch]),
            input_images=torch.stack(images, dim=0) if images else None,
            input_image_locations=torch.tensor(locations, dtype=torch.long) if images else None,
        )

    @staticmethod
    def sync_batch_to_model_parallel(topology: Optional[Topology], batch: Optional[TextDatasetBatch]) -> TextDatasetBatch:
        """TP rank 0 holds the collated batch; one int64 and one fp32 broadcast bring it to the whole TP group
        (missing images travel as a ``[-1]`` sentinel so the tensor count is static)."""
        if topology is None:
            assert batch is not None
            return batch
        longs: Optional[List[Optional[torch.Tensor]]] = None
        floats: Optional[List[Optional[torch.Tensor]]] = None
        if topology.model_parallel_rank == 0:
            assert batch is not None
            batch.contiguous_()
            has_img = batch.input_images is not None
            longs = [batch.input_token_ids, batch.target_token_ids, batch.cumulative_seq_lengths_padded.to(torch.long),
                     batch.position_ids, batch.input_image_locations if has_img else torch.tensor([-1], dtype=torch.long)]
            floats = [batch.loss_weights, batch.input_images if has_img else torch.tensor([-1.0])]
        else:
            assert batch is None
        lt = broadcast_data(longs, torch.long, topology, n=5)
        ft = broadcast_data(floats, torch.float32, topology, n=2)
        no_img = bool((lt[4] == -1).all())
        return TextDatasetBatch(
            input_token_ids=lt[0], target_token_ids=lt[1], cumulative_seq_lengths_padded=lt[2].to(torch.int32),
            position_ids=lt[3], loss_weights=ft[0], input_images=None if no_img else ft[1],
            input_image_locations=None if no_img else lt[4],
        )


def _open_image(path: Path):
    from PIL import Image

    return Image.open(str(path))


# ---------------------------------------------------------------------------------------------------------------
class FinetuningTextDataset(_FinetuningBase):
    def __init__(self, data_prefix: Path, sequence_length: int, seed: int, softprompt_n_tokens: int, tokenizer: Any,
                 tokenizer_no_prefix_space: Any, memory_map_dataset: bool = False, shuffle: bool = True):
        self.data_prefix = Path(data_prefix)
        self.data_prefix_parent = self.data_prefix.parent
        self.sequence_length = sequence_length
        self.softprompt_n_tokens = softprompt_n_tokens
        self.tokenizer = tokenizer
        self.tokenizer_no_prefix_space = tokenizer_no_prefix_space
        self.memory_map_dataset = memory_map_dataset
        self.dataset: Optional[MemoryMapDataset] = None
        if memory_map_dataset:
            self.dataset = MemoryMapDataset(self.data_prefix)
            self.data = list(range(len(self.dataset)))
        else:
            with open(self.data_prefix, encoding="UTF-8") as f:
                self.data = json.load(f)
        super().__init__(seed=seed, shuffle=shuffle)

    def ident(self) -> str:
        md5 = hashlib.md5()
        md5.update(str(self.data_prefix).encode("utf-8"))
        if not self.memory_map_dataset:  # the memory map is already tokenised
            _vocab_digest(md5, self.tokenizer)
            _vocab_digest(md5, self.tokenizer_no_prefix_space)
        return f"{md5.hexdigest()}-seq-{self.sequence_length}"

    def _sample(self, index: int) -> _Sample:
        if self.memory_map_dataset:
            assert self.dataset is not None
            doc = np.asarray(self.dataset[self.data[index]]).tolist()
            n_prompt = int(doc[0])
            prompt, completion = doc[1:n_prompt + 1], doc[n_prompt + 1:]
            return _Sample(tokens=prompt + completion, loss=[0] * len(prompt) + [1] * len(completion))
        rec: Dict[str, Any] = self.data[index]
        prompt_ids: List[int] = []
        paths: List[Path] = []
        spans: List[Tuple[int, int]] = []
        pieces = rec["prompt"] if isinstance(rec["prompt"], list) else [rec["prompt"]]
        for i, piece in enumerate(pieces):
            assert isinstance(piece, str)
            if piece.endswith(".jpg") and len(piece) < 1024 and (self.data_prefix_parent / piece).is_file():
                spans.append((len(prompt_ids), len(prompt_ids) + IMAGE_ENCODER_TOKEN_COUNTS))
                paths.append(self.data_prefix_parent / piece)
                prompt_ids.extend([self._eos()] * IMAGE_ENCODER_TOKEN_COUNTS)
            else:  # only the very first piece gets the tokenizer's leading-space treatment
                prompt_ids.extend((self.tokenizer if i == 0 else self.tokenizer_no_prefix_space).encode(piece))
        completion_ids = self.tokenizer.encode(rec["completion"])
        return _Sample(tokens=prompt_ids + completion_ids, loss=[0] * len(prompt_ids) + [1] * len(completion_ids),
                       image_paths=paths, image_spans=spans)

    def __getitem__(self, index: int) -> FinetuningTextDatasetItem:
        return self._finish(self._sample(index), loss_on_padding_eos=True)

    @staticmethod
    def convert_jsonl(jsonl_file: Union[str, Path], tokenizer: Any, tokenizer_no_prefix_space: Any,
                      out_prefix_path: Union[str, Path]) -> None:
        """``{"prompt", "completion"}`` lines → memory map of ``[len(prompt), *prompt, *completion]`` documents."""
        with MemoryMapDatasetBuilder(Path(out_prefix_path)) as builder, open(jsonl_file, encoding="UTF-8") as f:
            for line in f:
                line = line.strip()
                if not line:
                    continue
                rec = json.loads(line)
                prompt = tokenizer.encode(rec["prompt"])
                completion = tokenizer_no_prefix_space.encode(rec["completion"])
                builder.add(np.asarray([len(prompt)] + prompt + completion))


class FinetuningChatDataset(_FinetuningBase):
    def __init__(self, data_path: Path, sequence_length: int, seed: int, softprompt_n_tokens: int, tokenizer: Any,
                 tokenizer_no_prefix_space: Any, shuffle: bool = True):
        self.data_path = Path(data_path)
        self.data_path_parent = self.data_path.parent
        self.sequence_length = sequence_length
        self.softprompt_n_tokens = softprompt_n_tokens
        self.tokenizer = tokenizer
        self.tokenizer_no_prefix_space = tokenizer_no_prefix_space
        with open(self.data_path, encoding="UTF-8") as f:
            conversations = [json.loads(line) for line in f.read().split("\n") if line != ""]
        self.data = [self._tokenize(turns) for turns in conversations]
        if any(self._eos() not in s.tokens for s in self.data):
            print("WARNING: No EOS Token detected. Other than the 'finetuning_text_dataset', the "
                  "'finetuning_chat_dataset' does not automatically add an EOS token to your data. "
                  "You have to add it yourself into your 'data.jsonl'.")
        super().__init__(seed=seed, shuffle=shuffle)

    def _tokenize(self, turns: List[Dict[str, Any]]) -> _Sample:
        s = _Sample(tokens=[], loss=[])
        first_text = True
        for turn in turns:
            kind, content = turn["type"], turn["content"]
            if kind == "text":
                ids = (self.tokenizer if first_text else self.tokenizer_no_prefix_space).encode(content)
                first_text = False
                s.tokens.extend(ids)
                s.loss.extend([int(bool(turn.get("has_loss", False)))] * len(ids))
            elif kind == "image":
                s.image_spans.append((len(s.tokens), len(s.tokens) + IMAGE_ENCODER_TOKEN_COUNTS))
                s.image_paths.append(self.data_path_parent / content)
                s.tokens.extend([self._eos()] * IMAGE_ENCODER_TOKEN_COUNTS)
                s.loss.extend([0] * IMAGE_ENCODER_TOKEN_COUNTS)
            else:
                raise NotImplementedError(f"Content type {kind} is not supported")
        return s

    def ident(self) -> str:
        md5 = hashlib.md5()
        md5.update(str(self.data_path).encode("utf-8"))
        _vocab_digest(md5, self.tokenizer)
        _vocab_digest(md5, self.tokenizer_no_prefix_space)
        return f"{md5.hexdigest()}-seq-{self.sequence_length}"

    def __getitem__(self, index: int) -> FinetuningTextDatasetItem:
        return self._finish(self.data[index], loss_on_padding_eos=False)


class FinetuningTextBlendedDataset(BaseBlendedDataset[FinetuningTextDatasetItem, TextDatasetBatch, TextDatasetBatch, FinetuningTextDataset]):
    @staticmethod
    def sync_batch_to_model_parallel(topology, batch):
        return _FinetuningBase.sync_batch_to_model_parallel(topology, batch)


class FinetuningChatBlendedDataset(BaseBlendedDataset[FinetuningTextDatasetItem, TextDatasetBatch, TextDatasetBatch, FinetuningChatDataset]):
    @staticmethod
    def sync_batch_to_model_parallel(topology, batch):
        return _FinetuningBase.sync_batch_to_model_parallel(topology, batch)
