"""Tokenizer wrapper around HF ``tokenizers`` (reference: transformer/tokenizer/tokenizer.py:1-103)."""
from __future__ import annotations

from pathlib import Path
from typing import List, Optional, Tuple, Union

DEFAULT_VOCAB = Path(__file__).parent / "alpha-001-128k.json"


class Tokenizer:
    def __init__(self, tokenizer: "object"):
        self.tokenizer = tokenizer
        self.eos_token_id = self._lookup_eos()

    def _lookup_eos(self) -> int:
        for tok in ("<|endoftext|>", "</s>", "<eos>", "[SEP]"):
            i = self.tokenizer.token_to_id(tok)
            if i is not None:
                return int(i)
        return 0

    @classmethod
    def from_file(cls, filename: Union[str, Path]) -> "Tokenizer":
        from tokenizers import Tokenizer as HFTokenizer

        return cls(HFTokenizer.from_file(str(filename)))

    @classmethod
    def from_str(cls, s: str) -> "Tokenizer":
        from tokenizers import Tokenizer as HFTokenizer

        return cls(HFTokenizer.from_str(s))

    @classmethod
    def default(cls) -> "Tokenizer":
        assert DEFAULT_VOCAB.is_file(), (f"default vocabulary {DEFAULT_VOCAB} is not shipped (large blob); pass a "
                                         f"vocab_file explicitly")
        return cls.from_file(DEFAULT_VOCAB)

    def __len__(self) -> int:
        return self.tokenizer.get_vocab_size()

    @property
    def vocab_size(self) -> int:
        return len(self)

    def encode(self, text: str) -> List[int]:
        return self.tokenizer.encode(text, add_special_tokens=False).ids

    def decode(self, token_ids: List[int], skip_special_tokens: bool = False) -> str:
        return self.tokenizer.decode(token_ids, skip_special_tokens=skip_special_tokens)


def load_tokenizers(vocab_file: Optional[Union[str, Path]] = None) -> Tuple[Tokenizer, Tokenizer]:
    """(tokenizer, tokenizer without the metaspace prefix space) — the second is used to continue text after an image
    or a role marker without inserting a spurious leading space."""
    import json

    path = Path(vocab_file) if vocab_file is not None else DEFAULT_VOCAB
    tok = Tokenizer.from_file(path)
    spec = json.loads(path.read_text(encoding="utf-8"))

    def strip_prefix_space(node: object) -> None:
        if isinstance(node, dict):
            if node.get("type") == "Metaspace":
                node["add_prefix_space"] = False
                if "prepend_scheme" in node:
                    node["prepend_scheme"] = "never"
            for v in node.values():
                strip_prefix_space(v)
        elif isinstance(node, list):
            for v in node:
                strip_prefix_space(v)

    strip_prefix_space(spec.get("pre_tokenizer"))
    return tok, Tokenizer.from_str(json.dumps(spec))
