"""Config of the fused diffusers transformer block (``deepspeed/ops/transformer/inference/diffusers_2d_transformer.py``)."""


class Diffusers2DTransformerConfig:

    def __init__(self, int8_quantization: bool = False):
        self.int8_quantization = int8_quantization
