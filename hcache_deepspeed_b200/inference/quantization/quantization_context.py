"""``QuantizationContext`` (``deepspeed/inference/quantization/quantization_context.py:10-13``): a ZeRO-3 ``Init``
context for ZeRO-Inference whose parameters can also be backed by an NVMe parameter swapper."""
from __future__ import annotations

from ...runtime.zero.partition_parameters import Init


class QuantizationContext(Init):

    def __init__(self, config_dict_or_path, param_swapper=None) -> None:
        super().__init__(config_dict_or_path=config_dict_or_path, param_swapper=param_swapper)
        self.param_swapper = param_swapper
