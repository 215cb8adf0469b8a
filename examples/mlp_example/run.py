"""`python -m examples.mlp_example.run examples/mlp_example/config.yml`"""
from examples._launch import launch
from examples.mlp_example.config import MLPConfig

if __name__ == "__main__":
    raise SystemExit(launch(MLPConfig))
