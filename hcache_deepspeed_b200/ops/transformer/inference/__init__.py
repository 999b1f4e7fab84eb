from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ops import InferenceOps  # noqa: F401
from .moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig, DeepSpeedMoEMLP  # noqa: F401,E402
from .diffusers_2d_transformer import Diffusers2DTransformerConfig  # noqa: F401,E402
from .diffusers_attention import DeepSpeedDiffusersAttention  # noqa: F401,E402
from .diffusers_transformer_block import DeepSpeedDiffusersTransformerBlock  # noqa: F401,E402
