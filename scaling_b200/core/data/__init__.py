from .base import (  # noqa: F401
    BaseDataset,
    BaseDatasetBatch,
    BaseDatasetBatchBeforeSyncGeneric,
    BaseDatasetBatchGeneric,
    BaseDatasetItem,
    BaseDatasetItemGeneric,
    BaseLayerIO,
)
from .blended import (  # noqa: F401
    BaseBlendedDataset,
    BlendedDatasetConfig,
    blend_index,
    weights_by_num_docs,
    weights_examples_proportional,
)
from .broadcast import broadcast_data, sync_sizes  # noqa: F401
from .loader import DataLoader, RandomSampler  # noqa: F401
from .memmap import (  # noqa: F401
    FileDataset,
    FileHandle,
    MemoryMapDataset,
    MemoryMapDatasetBuilder,
    RetryableException,
)
