"""28x28 digit classification data (reference: examples/mlp_example/data.py).

MNIST is read through torchvision from ``root`` when it is already on disk (nothing is downloaded — training boxes have
no network); otherwise a deterministic synthetic set with the same shapes stands in: each class is a fixed random
28x28 template plus noise, which a small MLP separates in a few hundred steps.
"""
from __future__ import annotations

from pathlib import Path
from typing import List, Optional

import torch

from scaling_b200.core import BaseDataset, BaseDatasetBatch, BaseDatasetItem, Topology, broadcast_data


class MNISTDatasetItem(BaseDatasetItem):
    def __init__(self, input_: torch.Tensor, target: torch.Tensor):
        self.input = input_
        self.target = target


class MNISTDatasetBatch(BaseDatasetBatch):
    def __init__(self, inputs: Optional[torch.Tensor] = None, targets: Optional[torch.Tensor] = None):
        self.inputs = inputs
        self.targets = targets

    def only_inputs(self) -> "MNISTDatasetBatch":
        return MNISTDatasetBatch(inputs=self.inputs)

    def only_targets(self) -> "MNISTDatasetBatch":
        return MNISTDatasetBatch(targets=self.targets)


def _load_mnist(root: Path, train: bool):
    try:
        from torchvision import datasets

        ds = datasets.MNIST(root=str(root), train=train, download=False)
        return ds.data.float().div_(255.0).sub_(0.1307).div_(0.3081), ds.targets.long()
    except Exception:  # noqa: BLE001  (not on disk / torchvision missing)
        return None


def _synthetic(n: int, train: bool):
    g = torch.Generator().manual_seed(1234)
    templates = torch.randn(10, 28, 28, generator=g)
    g2 = torch.Generator().manual_seed(1 if train else 2)
    targets = torch.randint(0, 10, (n,), generator=g2)
    images = templates[targets] + 0.7 * torch.randn(n, 28, 28, generator=g2)
    return images, targets


class MNISTDataset(BaseDataset[MNISTDatasetItem, MNISTDatasetBatch, MNISTDatasetBatch]):
    def __init__(self, root: Path = Path("./.data"), train: bool = True, synthetic_size: int = 4096, seed: int = 42):
        loaded = _load_mnist(Path(root), train)
        self.synthetic = loaded is None
        self.images, self.targets = loaded if loaded is not None else _synthetic(synthetic_size if train else max(256, synthetic_size // 8), train)
        self.order = torch.arange(len(self.targets))
        super().__init__(seed=seed)

    def __len__(self) -> int:
        return len(self.targets)

    def __getitem__(self, index: int) -> MNISTDatasetItem:
        i = int(self.order[index])
        return MNISTDatasetItem(self.images[i], self.targets[i])

    def ident(self) -> str:
        return "MNIST" if not self.synthetic else "synthetic-digits"

    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        self.seed = seed
        self.order = (torch.randperm(len(self.targets), generator=torch.Generator().manual_seed(seed)) if shuffle
                      else torch.arange(len(self.targets)))

    def collate(self, batch: List[MNISTDatasetItem]) -> MNISTDatasetBatch:
        return MNISTDatasetBatch(inputs=torch.stack([b.input for b in batch]), targets=torch.stack([b.target for b in batch]))

    @staticmethod
    def sync_batch_to_model_parallel(topology: Topology, batch: Optional[MNISTDatasetBatch]) -> MNISTDatasetBatch:
        inputs = broadcast_data(None if batch is None else [batch.inputs], torch.float32, topology, n=1)[0]
        targets = broadcast_data(None if batch is None else [batch.targets], torch.long, topology, n=1)[0]
        return MNISTDatasetBatch(inputs=inputs, targets=targets)
