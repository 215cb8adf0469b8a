"""`python -m examples.transformer_example.run examples/transformer_example/config.yml`"""
from examples._launch import launch
from scaling_b200.transformer import TransformerConfig

if __name__ == "__main__":
    raise SystemExit(launch(TransformerConfig))
