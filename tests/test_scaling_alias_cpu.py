"""The ``scaling`` import namespace (reference module paths) resolves to scaling_b200 objects."""
import importlib
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parents[1]

# (reference module path, names the reference's examples / tests / users import from it)
PATHS = [
    ("scaling.core", ["Topology", "TopologyConfig", "ParallelModule", "Optimizer", "BaseTrainer", "runner_main", "BaseConfig"]),
    ("scaling.core.logging", ["logger", "LoggerConfig"]),
    ("scaling.core.logging.logging", ["logger"]),
    ("scaling.core.runner", ["LaunchConfig", "RunnerConfig"]),
    ("scaling.core.runner.launch_config", ["LaunchConfig"]),
    ("scaling.core.utils.port", ["find_free_port"]),
    ("scaling.core.nn.parallel_module", ["PipeCommunicator", "TrainStepOutput", "EvaluationStepOutput", "ParallelModule"]),
    ("scaling.core.nn.parallel_module.inference_module", ["InferenceModule", "RecorderSetting"]),
    ("scaling.core.nn.parallel_module.communicator", ["PipeCommunicator", "ModelParallelCommunicator"]),
    ("scaling.core.nn.linear.utils", ["all_concat", "all_reduce", "copy_to_tensor_model_parallel_region", "get_device"]),
    ("scaling.core.nn.pipeline_schedule.base", ["SimulationEngine"]),
    ("scaling.core.nn.pipeline_schedule.instructions", ["InstructionForwardPass", "InstructionOptimizerStep"]),
    ("scaling.core.topology.topology_config", ["TopologyConfig", "PipePartitionMethod", "ActivationCheckpointingType"]),
    ("scaling.core.topology.rng_tracker", ["CudaRNGStateTracker"]),
    ("scaling.core.data.memory_map", ["MemoryMapDataset", "MemoryMapDatasetBuilder"]),
    ("scaling.core.data.file_handles", ["FileHandle", "RetryableException"]),
    ("scaling.core.data.blended_dataset", ["BaseBlendedDataset", "weights_by_num_docs", "weights_examples_proportional"]),
    ("scaling.core.optimizer.parameter_group", ["OptimizerParamGroup"]),
    ("scaling.core.optimizer.learning_rate_scheduler", ["LearningRateScheduler", "LearningRateSchedulerConfig"]),
    ("scaling.core.utils.param_merge", ["merge_parameter", "split_parameter"]),
    ("scaling.transformer", ["TransformerConfig", "TransformerContext", "TextDataset", "init_model", "init_optimizer",
                             "TransformerLayerIO", "TransformerParallelModule", "LegacyBlendedDataset"]),
    ("scaling.transformer.context.config", ["TransformerConfig", "TransformerArchitectureConfig", "Precision", "MLPType"]),
    ("scaling.transformer.model.model", ["loss_function", "metrics_aggregation_fn", "get_transformer_layer_specs"]),
    ("scaling.transformer.model.layers.layer", ["TransformerLayer"]),
    ("scaling.transformer.model.layers.lm_head_tied", ["TransformerLMHeadTied"]),
    ("scaling.transformer.inference.inference_model", ["TransformerInferenceModule"]),
    ("scaling.transformer.tokenizer", ["Tokenizer", "load_tokenizers"]),
    ("scaling.transformer.utils.get_tflops", ["get_model_flop_utilization_palm", "HardwareType"]),
    ("scaling.transformer.train", ["main"]),
    ("scaling.transformer.data.legacy_dataset.indexed_dataset", ["MMapIndexedDataset"]),
]


@pytest.mark.parametrize("path,names", PATHS, ids=[p for p, _ in PATHS])
def test_reference_paths_resolve_to_scaling_b200_objects(path, names):
    mod = importlib.import_module(path)
    for n in names:
        obj = getattr(mod, n)
        owner = getattr(obj, "__module__", None) or type(obj).__module__
        assert owner.startswith("scaling_b200") or owner in ("typing", "logging"), (path, n, owner)


def test_alias_and_package_share_class_identity():
    from scaling.core import Topology as A
    from scaling.core.topology.topology import Topology as B
    from scaling_b200.core import Topology as C

    assert A is B is C
    with pytest.raises(ImportError):
        importlib.import_module("scaling.core.does_not_exist")


def test_alias_modules_run_as_scripts():
    """`python -m scaling.core.runner.launch` (the command line the runner builds, reference runner.py:41-115)."""
    r = subprocess.run([sys.executable, "-m", "scaling.core.runner.launch", "--help"], cwd=REPO, capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0 and "resource_pool" in r.stdout, r.stdout + r.stderr


def test_blended_dataset_loop_shim_writes_the_wheel_files(tmp_path):
    import json

    import numpy as np

    import blended_dataset_loop

    stem = str(tmp_path / "cache" / "index_seed_1")
    blended_dataset_loop.sample(np.array([3, 5]), stem)
    meta = json.load(open(stem + ".meta.json"))
    idx = np.fromfile(stem + ".bin", dtype=np.dtype(meta["dtype"])).reshape(meta["shape"])
    assert idx.shape == (8, 2) and sorted(idx[idx[:, 0] == 1][:, 1].tolist()) == [0, 1, 2, 3, 4]
    assert Path(stem + ".input.json").is_file()


def test_reference_unit_files_run_against_the_alias(tmp_path):
    """The reference's own test files are the oracle for API parity: a few fast ones (batch-broadcast shape table,
    finetuning parameter selection, retried file reads) run unchanged through the alias.  Skipped where the reference
    checkout is not available."""
    import json
    import subprocess
    import sys
    from pathlib import Path

    import pytest

    ref = Path("/root/reference")
    if not (ref / "tests").is_dir():
        pytest.skip("reference checkout not present")
    repo = Path(__file__).resolve().parents[1]
    out = tmp_path / "ref.json"
    subprocess.run([sys.executable, str(repo / "tools" / "run_reference_tests.py"), "--reference", str(ref), "--out", str(out),
                    "--targets", "tests/core/test_data/test_broadcast_data.py", "tests/core/test_data/test_file_dataset.py",
                    "tests/transformer/test_finetuning_parameter.py"], check=True, capture_output=True, timeout=600)
    res = json.loads(out.read_text())
    assert res["not_passing"] == [] and res["counts"].get("passed", 0) >= 30, res
