"""``deepspeed.linear.Init`` / ``init_lora`` (``deepspeed/linear/context_manager.py:17-90``): build a model with every
``nn.Linear`` replaced by :class:`OptimizedLinear` (sharded / quantised frozen base weight + LoRA adapters) at
construction time, so the full-precision base weights are never all resident.

The reference swaps ``torch.nn.Linear`` for a wrapper class and patches ``AutoModelForCausalLM`` constructors to run
``init_lora``. Here the context keeps a list of what it built and finishes them on exit (freeze everything except the
adapters), so it needs no hook into ``transformers``; bias-carrying or tiny layers fall through to the stock Linear."""
from __future__ import annotations

import torch

from .optimized_linear import LoRAOptimizedLinear, OptimizedLinear


def init_lora(model: torch.nn.Module) -> torch.nn.Module:
    """Freeze the model; only the LoRA adapter matrices train."""
    model.requires_grad_(False)
    for m in model.modules():
        if isinstance(m, LoRAOptimizedLinear):
            m.lora_weight_1.weight.requires_grad_(True)
            m.lora_weight_2.weight.requires_grad_(True)
    return model


class Init:

    def __init__(self, lora_config=None, quant_config=None, dtype=None, min_features: int = 0):
        self.lora_config, self.quant_config, self.dtype, self.min_features = lora_config, quant_config, dtype, min_features
        self.created = []
        self._orig = None
        self._building = False

    def __enter__(self):
        ctx, orig = self, torch.nn.Linear
        self._orig = orig

        class _LinearFactory(orig):        # isinstance(x, nn.Linear) checks inside model code keep working for stock ones

            def __new__(cls, in_features, out_features, bias=True, device=None, dtype=None):
                if ctx._building or bias or min(in_features, out_features) < ctx.min_features or \
                        (ctx.lora_config is None and ctx.quant_config is None):
                    return orig(in_features, out_features, bias=bias, device=device, dtype=dtype)
                ctx._building = True       # the adapters inside an OptimizedLinear are stock Linears
                try:
                    mod = OptimizedLinear(in_features, out_features, bias=False, lora_config=ctx.lora_config,
                                          quantization_config=ctx.quant_config, dtype=ctx.dtype or dtype or torch.bfloat16)
                finally:
                    ctx._building = False
                ctx.created.append(mod)
                return mod

        torch.nn.Linear = _LinearFactory
        return self

    def __exit__(self, *exc):
        torch.nn.Linear = self._orig
        return False

    def finalize(self, model: torch.nn.Module) -> torch.nn.Module:
        """Call on the built model (the reference does this from patched ``from_pretrained`` / ``from_config``)."""
        return init_lora(model) if self.lora_config is not None else model
