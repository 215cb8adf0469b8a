"""Runs the REFERENCE's own CPU-capable test files against this framework through the ``scaling`` alias package.

The reference's tests are staged in a scratch directory (nothing of them is copied into this repository), the repo
root goes on PYTHONPATH so that ``import scaling`` resolves to the alias (scaling/__init__.py) and ``import
blended_dataset_loop`` to the C++-backed shim, and pytest's summary is written to ``profiles/reference_cpu_tests_r2.json``.

    python tools/run_reference_tests.py [--reference /root/reference]
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

p = argparse.ArgumentParser()
p.add_argument("--reference", default="/root/reference")
p.add_argument("--out", default=str(REPO / "profiles" / "reference_cpu_tests_r2.json"))
a = p.parse_args()

with tempfile.TemporaryDirectory(prefix="reftests_") as tmp:
    shutil.copytree(Path(a.reference) / "tests", Path(tmp) / "tests")
    env = dict(os.environ, PYTHONPATH=str(REPO))
    r = subprocess.run([sys.executable, "-m", "pytest", *TARGETS, "-q", "-p", "no:cacheprovider",
                        "--continue-on-collection-errors", "-rfE"], cwd=tmp, env=env, capture_output=True, text=True)
tail = r.stdout.strip().splitlines()
summary = tail[-1] if tail else ""
counts = {k: int(v) for v, k in re.findall(r"(\d+) (passed|failed|skipped|errors?|warnings?)", summary)}
not_passing = sorted({ln.split(" - ")[0] for ln in tail if ln.startswith(("FAILED ", "ERROR "))})
json.dump({"targets": TARGETS, "summary": summary, "counts": counts, "not_passing": not_passing,
           "note": "collection errors are test files that import private helpers of the reference's implementation "
                   "(names starting with an underscore) which are not part of its API"}, open(a.out, "w"), indent=1)
print(summary)
for ln in not_passing:
    print(" ", ln)
