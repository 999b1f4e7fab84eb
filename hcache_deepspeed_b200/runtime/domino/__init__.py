from .transformer import DominoTransformerLayer, DominoAsyncLinear  # noqa: F401
