from .finetuning import (  # noqa: F401
    FinetuningChatBlendedDataset,
    FinetuningChatDataset,
    FinetuningTextBlendedDataset,
    FinetuningTextDataset,
)
from .inference_settings import Control, InferenceSettings, InferenceSuppressionParameters  # noqa: F401
from .text_dataset import TextBlendedDataset, TextDataset, TextDatasetItem, build_window_index  # noqa: F401
from .text_dataset_batch import TextDatasetBatch, TextDatasetBatchBeforeSync  # noqa: F401
from .utils import (  # noqa: F401
    add_cumulative_seq_lengths_padding,
    get_cumulative_seq_lengths,
    get_position_ids,
    remove_cumulative_seq_lengths_padding,
)
from .legacy import LegacyBlendedDataset  # noqa: F401,E402
