"""Drop-in for the reference's optional Rust wheel ``blended-dataset-loop`` (requirements/base.txt:3), backed by this
framework's native C++ index builder (csrc/blend_index.cpp).

The reference calls ``blended_dataset_loop.sample(counts, stem)`` (core/data/blended_dataset.py:316-329) and then reads
``{stem}.bin`` / ``{stem}.meta.json`` / ``{stem}.input.json`` back (:357-371).  With this module on the import path the
unmodified reference can blend several datasets offline, and its blended-loader tests act as an oracle for the C++
routine.
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Sequence

import numpy as np


def sample(counts: Sequence[int], stem: str) -> None:
    """Writes the interleave index for ``counts`` samples per dataset: ``{stem}.bin`` (int64 rows
    ``[dataset index, index within dataset]``), ``{stem}.meta.json`` (``dtype``, ``shape``) and ``{stem}.input.json``."""
    from scaling_b200.core.data.blended import blend_index

    counts = np.asarray(counts, dtype=np.int64)
    index = blend_index(counts)
    Path(stem).parent.mkdir(parents=True, exist_ok=True)
    tmp = Path(str(stem) + ".bin.tmp")
    index.tofile(tmp)
    tmp.rename(str(stem) + ".bin")
    Path(str(stem) + ".input.json").write_text(json.dumps({"counts": counts.tolist()}))
    Path(str(stem) + ".meta.json").write_text(json.dumps({"dtype": "int64", "shape": list(index.shape)}))
