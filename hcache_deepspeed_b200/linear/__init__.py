from .config import LoRAConfig, QuantizationConfig  # noqa: F401
from .optimized_linear import LoRAOptimizedLinear, OptimizedLinear  # noqa: F401
from .quantization import QuantizedLinear, QuantizedParameter  # noqa: F401
from .context_manager import Init, init_lora  # noqa: F401
