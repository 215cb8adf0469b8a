"""Every public name of the reference is importable from the same place under scaling_b200."""
import importlib

import pytest

CORE = """BaseConfig BaseContext BaseContextGeneric DeterminedBaseContext BaseBlendedDataset BaseDataset BaseDatasetBatch
BaseDatasetItem BaseLayerIO BlendedDatasetConfig DataLoader FileDataset MemoryMapDataset MemoryMapDatasetBuilder broadcast_data
ActivationFunction BaseLayer ColumnParallelLinear CoreParameterMeta LayerNorm LayerNormConfig LayerNormOptimizationType LayerSpec
LoRaConfig LoRAModuleType MaskedSoftmax MaskedSoftmaxConfig MaskedSoftmaxKernel NormType ParallelLoRa ParallelMLP ParallelModule
ParallelSelfAttention ParallelSwiGLUMLP PipelineScheduleInference PipelineScheduleTrain PipePartitionCoordinates
RelativePositionEmbeddingType RMSNorm RotaryConfig RotaryEmbedding RotaryEmbeddingComplex RowParallelLinear TiedLayerSpec
VocabParallelEmbedding get_activation_function get_norm pipe_partition_uniform BaseOptimizer LearningRateDecayStyle
LearningRateScheduler LearningRateSchedulerConfig LossScaler LossScalerConfig Optimizer OptimizerConfig OptimizerParamGroup
OptimizerParamGroupConfig Profiler ProfilerConfig SynchronizedTimer RunnerConfig RunnerDockerConfig RunnerType runner_main
PipePartitionMethod Topology TopologyConfig BaseTrainer DeterminedBaseTrainer TrainerConfig""".split()

TRANSFORMER = """TransformerConfig TransformerContext FinetuningChatBlendedDataset FinetuningChatDataset
FinetuningTextBlendedDataset FinetuningTextDataset LegacyBlendedDataset TextBlendedDataset TextDataset TextDatasetItem
TransformerLayerIO TransformerParallelModule init_model init_optimizer""".split()

SUB_PATHS = {
    "scaling_b200.core.logging": ["logger", "LoggerConfig"],
    "scaling_b200.core.runner": ["LaunchConfig"],
    "scaling_b200.core.utils.port": ["find_free_port"],
    "scaling_b200.core.nn": ["PipeCommunicator", "ModelParallelCommunicator", "TrainStepOutput", "EvaluationStepOutput",
                             "InferenceModule", "RecorderSetting", "HiddenStateRecorder", "SimulationEngine"],
    "scaling_b200.core.nn.inference_module": ["InferenceModule", "RecorderSetting"],
    "scaling_b200.transformer.train": ["main", "TransformerTrainer"],
    "scaling_b200.transformer.train_determined": ["main", "from_launcher_args_determined"],
    "scaling_b200.transformer.model.model": ["loss_function", "metrics_aggregation_fn", "get_transformer_layer_specs"],
    "scaling_b200.transformer.inference.inference_model": ["TransformerInferenceModule", "CompletionOutput"],
    "scaling_b200.transformer.inference": ["sample_argmax", "sample_temperature", "top_k_transform", "top_p_transform"],
    "scaling_b200.transformer.tokenizer": ["Tokenizer", "load_tokenizers"],
    "scaling_b200.transformer.utils.get_tflops": ["get_tflops_aleph_alpha", "get_tflops_bloom", "get_tflops_electra",
                                                  "get_tflops_megatron", "get_model_flop_utilization_palm", "HardwareType"],
    "scaling_b200.transformer.data.legacy": ["MMapIndexedDataset", "make_dataset", "make_builder"],
    "scaling_b200.transformer.model.image_encoder": ["ImageEncoder", "ClipModifiedResNet", "clip_transform"],
    "scaling_b200.core.data": ["RandomSampler", "sync_sizes", "FileHandle", "RetryableException",
                               "weights_by_num_docs", "weights_examples_proportional"],
    "scaling_b200.parallel": ["TPComm", "SymmetricArena", "get_tp_comm", "fused_zero_enabled"],
    "scaling_b200.models": ["architecture_preset", "PRESETS", "init_model"],
}


@pytest.mark.parametrize("name", CORE)
def test_core_name(name):
    import scaling_b200.core as core

    assert getattr(core, name) is not None
    assert name in dir(core)


@pytest.mark.parametrize("name", TRANSFORMER)
def test_transformer_name(name):
    import scaling_b200.transformer as tfm

    assert getattr(tfm, name) is not None


@pytest.mark.parametrize("module", sorted(SUB_PATHS))
def test_sub_path_names(module):
    mod = importlib.import_module(module)
    missing = [n for n in SUB_PATHS[module] if not hasattr(mod, n)]
    assert not missing, f"{module} lacks {missing}"


def test_train_step_output_fields_and_layer_contract():
    from scaling_b200.core import BaseLayer
    from scaling_b200.core.nn import EvaluationStepOutput, TrainStepOutput

    assert TrainStepOutput._fields[:3] == ("loss", "metrics", "global_grad_norm")
    assert {"learning_rates", "overflow", "no_overflow_steps", "current_loss_scale", "step_duration", "debug_dict",
            "global_grad_norm_clipped"} <= set(TrainStepOutput._fields)
    assert EvaluationStepOutput._fields == ("loss", "metrics", "step_duration")
    for hook in ("input_to_tuple", "tuple_to_input", "output_to_tuple", "tuple_to_last_stage_activation"):
        assert hasattr(BaseLayer, hook)
