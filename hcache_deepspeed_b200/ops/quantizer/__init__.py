from .quantizer import (Quantizer, dequantize_int, ds_quantizer, quantize_int, quantized_reduction,  # noqa: F401
                        swizzled_quantize)
