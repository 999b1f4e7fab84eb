"""``DeepSpeedOPTInference`` (``deepspeed/model_implementations/transformers/ds_opt.py``): the reference has one thin subclass of
``DeepSpeedTransformerInference`` per model family; here one class covers them, configured by the layer wiring
(pre/post layer norm, parallel attention+MLP, rotary / ALiBi / learned positions) that the policy extracted."""
from .ds_transformer import DeepSpeedTransformerInference


class DeepSpeedOPTInference(DeepSpeedTransformerInference):
    pass
