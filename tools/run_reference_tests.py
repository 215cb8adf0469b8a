"""Runs the REFERENCE's own CPU-capable test files against this framework through the ``scaling`` alias package.

The reference's tests are staged in a scratch directory (nothing of them is copied into this repository), the repo
root goes on PYTHONPATH so that ``import scaling`` resolves to the alias (scaling/__init__.py) and ``import
blended_dataset_loop`` to the C++-backed shim, and pytest's summary is written to ``profiles/reference_cpu_tests_r2.json``.

    python tools/run_reference_tests.py [--reference /root/reference]            # CPU unit files
    python tools/run_reference_tests.py --suite gpu                              # acceptance suites, on a GPU box
    python tools/run_reference_tests.py --suite gpu --fake-devices 1             # same on CPU / gloo (dry run)
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
# test files of the reference that need no GPU (SURVEY §4.1 "core unit (CPU)" plus the CPU-runnable transformer units)
TARGETS = [
    "tests/core/test_config", "tests/core/test_topology", "tests/core/test_logging", "tests/core/test_data",
    "tests/core/test_nn/test_pipeline_schedule.py", "tests/core/test_nn/test_pipeline_partitioning.py",
    "tests/core/test_nn/test_attention_helpers.py", "tests/core/test_optimizer/test_learning_rate_scheduler.py",
    "tests/core/unit", "tests/transformer/test_finetuning_parameter.py", "tests/transformer/test_data_utils.py",
    "tests/transformer/test_utils.py",
]

# the reference's accelerator test files (SURVEY §4.1 training / transformer suites): they skip a case when it needs more
# devices than the box has, so one GPU runs every world-size-1 case and N GPUs the rest
GPU_TARGETS = [
    "tests/transformer/test_training.py", "tests/transformer/test_training_flash_attention.py",
    "tests/transformer/test_training_local_attention.py", "tests/transformer/test_training_sequence_parallel.py",
    "tests/transformer/test_training_legacy.py", "tests/transformer/test_training_finetuning_chat.py",
    "tests/transformer/test_finetuning.py", "tests/transformer/test_inference.py",
    "tests/transformer/test_load_checkpoint_non_strict.py", "tests/transformer/test_backwards_compatibility.py",
    "tests/transformer/test_data.py", "tests/transformer/test_blended_dataset.py",
    "tests/core/test_training", "tests/core/test_nn", "tests/core/test_optimizer",
]
DETERMINED_STUB = '''"""Scratch stand-in so that test modules which import the Determined SDK at module level can be collected without
it; cases that really use Determined are deselected."""
import types


class _Obj:
    def __init__(self, *a, **k):
        self.__dict__.update(k)


ClusterInfo = TrialInfo = RendezvousInfo = _Obj


def get_cluster_info():
    return None


core = types.SimpleNamespace(Context=_Obj, DistributedContext=_Obj, DummyDistributedContext=_Obj, TrainContext=_Obj,
                             CheckpointContext=_Obj, init=lambda *a, **k: None)
'''
SKIP_EXPR = "world_size > torch.cuda.device_count()"

p = argparse.ArgumentParser()
p.add_argument("--reference", default="/root/reference")
p.add_argument("--suite", choices=["cpu", "gpu"], default="cpu",
               help="cpu: the CPU-capable unit files; gpu: the training / transformer acceptance suites")
p.add_argument("--fake-devices", type=int, default=0,
               help="gpu suite on a box without GPUs (gloo dry run): pretend this many devices in the tests' skip check")
p.add_argument("--cpu-shim", action="store_true",
               help="dry run only: a scratch sitecustomize makes Tensor.cuda()/Module.cuda() no-ops and torch.cuda.manual_seed "
                    "the CPU seed, so unit files that hard-code .cuda() exercise the gloo / CPU path")
p.add_argument("--surrogate-tokenizer", action="store_true",
               help="the checkout lacks tests/transformer/files/alpha-001-128k.json (a large blob); stage the llama2 "
                    "tokenizer file it does ship under that name so that the tests which only need SOME tokenizer can run")
p.add_argument("--select", default=None, help="pytest -k expression")
p.add_argument("--targets", nargs="*", default=None, help="override the file list of the suite")
p.add_argument("--timeout", type=int, default=900, help="per-test timeout (pytest-timeout)")
p.add_argument("--tb", default="auto", help="pytest --tb style")
p.add_argument("--out", default=None)
a = p.parse_args()
targets = a.targets or (TARGETS if a.suite == "cpu" else GPU_TARGETS)
out = a.out or str(REPO / "profiles" / f"reference_{a.suite}_tests_r2.json")

with tempfile.TemporaryDirectory(prefix="reftests_") as tmp:
    shutil.copytree(Path(a.reference) / "tests", Path(tmp) / "tests")
    try:
        import determined  # noqa: F401
        have_determined = True
    except ImportError:
        have_determined = False
        (Path(tmp) / "determined.py").write_text(DETERMINED_STUB)
    if a.fake_devices:
        for f in (Path(tmp) / "tests").rglob("*.py"):
            src = f.read_text()
            if SKIP_EXPR in src:
                f.write_text(src.replace(SKIP_EXPR, f"world_size > max(torch.cuda.device_count(), {a.fake_devices})"))
    if a.surrogate_tokenizer:
        files = Path(tmp) / "tests" / "transformer" / "files"
        if not (files / "alpha-001-128k.json").exists() and (files / "llama2-tokenizer.json").exists():
            shutil.copy(files / "llama2-tokenizer.json", files / "alpha-001-128k.json")
    if a.cpu_shim:
        (Path(tmp) / "sitecustomize.py").write_text(
            "import torch\n"
            "torch.Tensor.cuda = lambda self, *a, **k: self\n"
            "torch.nn.Module.cuda = lambda self, *a, **k: self\n"
            "torch.cuda.manual_seed = lambda s: torch.manual_seed(s)\n")
    select = a.select
    if a.suite == "gpu" and not have_determined:
        # cases that drive the Determined SDK itself (by name; test_train_with_deterministic_behaviour passes
        # use_determined=True)
        no_det = "not determined and not deterministic_behaviour"
        select = f"({select}) and {no_det}" if select else no_det
    env = dict(os.environ, PYTHONPATH=f"{tmp}{os.pathsep}{REPO}" if a.cpu_shim else str(REPO))
    cmd = [sys.executable, "-m", "pytest", *targets, "-q", "-p", "no:cacheprovider", "--continue-on-collection-errors",
           "-rfE", "--timeout", str(a.timeout), f"--tb={a.tb}"]
    if select:
        cmd += ["-k", select]
    r = subprocess.run(cmd, cwd=tmp, env=env, capture_output=True, text=True)
tail = r.stdout.strip().splitlines()
summary = tail[-1] if tail else ""
counts = {k: int(v) for v, k in re.findall(r"(\d+) (passed|failed|skipped|deselected|errors?|warnings?)", summary)}
not_passing = sorted({ln.split(" - ")[0] for ln in tail if ln.startswith(("FAILED ", "ERROR "))})
json.dump({"suite": a.suite, "targets": targets, "select": select, "fake_devices": a.fake_devices, "cpu_shim": a.cpu_shim, "surrogate_tokenizer": a.surrogate_tokenizer,
           "cuda_devices": __import__("torch").cuda.device_count(), "summary": summary, "counts": counts,
           "not_passing": not_passing,
           "note": "not_passing lists every failing or uncollectable case; see profiles/README.md for what each one needs"},
          open(out, "w"), indent=1)
print(summary)
for ln in not_passing:
    print(" ", ln)
if os.environ.get("REFTESTS_VERBOSE"):
    print("\n".join(tail[-200:]))
