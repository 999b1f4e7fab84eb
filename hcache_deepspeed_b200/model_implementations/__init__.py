"""v1 fused layers (``deepspeed/model_implementations``)."""
from .transformers.ds_transformer import (DeepSpeedBERTInference, DeepSpeedBloomInference, DeepSpeedGPTInference,  # noqa: F401
                                          DeepSpeedLlama2Inference, DeepSpeedMegatronGPTInference, DeepSpeedOPTInference,
                                          DeepSpeedTransformerInference)
from .transformers.clip_encoder import DSClipEncoder  # noqa: F401
