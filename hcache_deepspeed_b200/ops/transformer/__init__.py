from .transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer  # noqa: F401
