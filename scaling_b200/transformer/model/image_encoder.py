"""Image prefix encoder: CLIP RN50x16 "modified ResNet" trunk → Linear → dropout → LayerNorm
(reference: transformer/model/image_encoder/image_encoder.py:1-81, clip.py:1-192).

A 384x384 image becomes 12x12 = 144 tokens of 3072 channels, projected to the model width and spliced over the
placeholder tokens by ``EmbeddingInput``.  Module / parameter names follow the CLIP checkpoint layout
(``input_encoder.layer3.4.conv2.weight`` …, ``proj``, ``layernorm``) because they are the checkpoint contract.
The trunk is cold-path convolution work (a frozen or lightly tuned prefix encoder run once per image): it stays on
cuDNN; only the projection output joins the transformer's sm_100a path.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from ...core.nn.norms import LayerNorm, LayerNormConfig

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)
IMAGE_ENCODER_TOKEN_COUNTS = 144
IMAGE_SIZE = (384, 384)


class _ClipTransform:
    """PIL image → normalised float tensor [3, H, W]: bicubic resize to ``n_px``, RGB, /255, CLIP mean/std."""

    def __init__(self, n_px: Tuple[int, int]):
        self.n_px = (int(n_px[0]), int(n_px[1]))
        self.mean = torch.tensor(CLIP_MEAN).view(3, 1, 1)
        self.std = torch.tensor(CLIP_STD).view(3, 1, 1)

    def __call__(self, image) -> torch.Tensor:
        from PIL import Image

        h, w = self.n_px
        image = image.resize((w, h), resample=Image.BICUBIC).convert("RGB")
        x = torch.from_numpy(np.asarray(image, dtype=np.uint8).copy()).permute(2, 0, 1).float().div_(255.0)
        return (x - self.mean) / self.std


def clip_transform(n_px: Tuple[int, int]) -> _ClipTransform:
    return _ClipTransform(n_px)


def _conv_bn(cin: int, cout: int, k: int, **kw) -> Tuple[torch.nn.Conv2d, torch.nn.BatchNorm2d]:
    return torch.nn.Conv2d(cin, cout, k, bias=False, **kw), torch.nn.BatchNorm2d(cout)


class Bottleneck(torch.nn.Module):
    """1x1 → 3x3 → (avg-pool when strided) → 1x1 with an anti-aliased (avg-pool + 1x1) shortcut."""

    expansion = 4

    def __init__(self, inplanes: int, planes: int, stride: int = 1):
        super().__init__()
        out = planes * self.expansion
        self.conv1, self.bn1 = _conv_bn(inplanes, planes, 1)
        self.conv2, self.bn2 = _conv_bn(planes, planes, 3, padding=1)
        self.avgpool = torch.nn.AvgPool2d(stride) if stride > 1 else torch.nn.Identity()
        self.conv3, self.bn3 = _conv_bn(planes, out, 1)
        self.relu = torch.nn.ReLU(inplace=True)
        self.stride = stride
        self.downsample: Optional[torch.nn.Sequential] = None
        if stride > 1 or inplanes != out:
            conv, bn = _conv_bn(inplanes, out, 1, stride=1)
            self.downsample = torch.nn.Sequential(OrderedDict([("-1", torch.nn.AvgPool2d(stride)), ("0", conv), ("1", bn)]))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.avgpool(self.relu(self.bn2(self.conv2(y))))
        y = self.bn3(self.conv3(y))
        shortcut = x if self.downsample is None else self.downsample(x)
        return self.relu(y + shortcut)


class ClipModifiedResNet(torch.nn.Module):
    """CLIP's ResNet: 3-conv stem with average pooling, anti-aliased strided blocks, no classification pooling —
    the [b, C, h, w] feature map is returned as [b, h·w, C] tokens."""

    def __init__(self, layers: Sequence[int], num_init_channels: int = 64):
        super().__init__()
        c = num_init_channels
        self.conv1, self.bn1 = _conv_bn(3, c // 2, 3, stride=2, padding=1)
        self.conv2, self.bn2 = _conv_bn(c // 2, c // 2, 3, padding=1)
        self.conv3, self.bn3 = _conv_bn(c // 2, c, 3, padding=1)
        self.avgpool = torch.nn.AvgPool2d(2)
        self.relu = torch.nn.ReLU(inplace=True)
        inplanes = c
        for i, (blocks, mult, stride) in enumerate(zip(layers, (1, 2, 4, 8), (1, 2, 2, 2)), start=1):
            stage: List[torch.nn.Module] = []
            for b in range(blocks):
                stage.append(Bottleneck(inplanes, c * mult, stride if b == 0 else 1))
                inplanes = c * mult * Bottleneck.expansion
            setattr(self, f"layer{i}", torch.nn.Sequential(*stage))
        self.out_channels = inplanes

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.to(self.conv1.weight.dtype)
        for conv, bn in ((self.conv1, self.bn1), (self.conv2, self.bn2), (self.conv3, self.bn3)):
            x = self.relu(bn(conv(x)))
        x = self.avgpool(x)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return x.flatten(2).transpose(1, 2)


class ImageEncoder(torch.nn.Module):
    def __init__(self, out_features: int, device: torch.device, dropout_p: float = 0.0,
                 layernorm_config: Optional[LayerNormConfig] = None, image_encoder: str = "ClipRN50x16",
                 dtype: torch.dtype = torch.float32):
        super().__init__()
        assert image_encoder == "ClipRN50x16", "only clip implemented"
        self.input_encoder = ClipModifiedResNet(layers=[6, 8, 18, 8], num_init_channels=96).to(dtype).to(device)
        self.input_encoder_output_dim = self.input_encoder.out_channels  # 3072
        self.image_encoder_image_size = IMAGE_SIZE
        self.proj = torch.nn.Linear(self.input_encoder_output_dim, out_features, device=device, dtype=dtype)
        self.dropout = torch.nn.Dropout(dropout_p)
        self.has_layernorm = layernorm_config is not None
        if layernorm_config is not None:
            self.layernorm = LayerNorm(config=layernorm_config, normalized_shape=out_features, device=device, dtype=dtype)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """[n_images, 3, 384, 384] → [n_images, 144, out_features]."""
        x = self.dropout(self.proj(self.input_encoder(x)))
        return self.layernorm(x) if self.has_layernorm else x
