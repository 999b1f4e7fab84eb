"""``DeepSpeedGPTInference`` (``deepspeed/model_implementations/transformers/ds_gpt.py``): the reference has one thin subclass of
``DeepSpeedTransformerInference`` per model family; here one class covers them, configured by the layer wiring
(pre/post layer norm, parallel attention+MLP, rotary / ALiBi / learned positions) that the policy extracted."""
from .ds_transformer import DeepSpeedTransformerInference


class DeepSpeedGPTInference(DeepSpeedTransformerInference):
    pass
