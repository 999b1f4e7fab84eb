"""Model policies: what to build and where its weights come from.

The reference needs a policy + container + model class per family (``inference_policy_base.py:28-220`` and the
ten ``model_implementations/<family>/`` packages). Here one policy class serves every family because the
architecture is data (:class:`ArchSpec`); the per-family names are kept as aliases so code written against
the reference (``Llama2Policy(config, checkpoint_engine=...)``) keeps working.
"""
from __future__ import annotations

import json
import os
import pickle
from typing import Any, Optional

import torch

from ....models import families
from ....models.arch import ArchSpec, arch_from_hf_config, supported_model_types
from ....models.flat_model import (make_metadata_filename, make_model_config_filename, make_param_filename,
                                   restore_model_weights)
from ....models.weights import load_weights, random_weights

POLICIES = {}


class InferenceV2Policy:
    """Builds a :class:`RaggedTransformer` from (a) a checkpoint engine, (b) a serialized DS inference checkpoint,
    or (c) random initialisation of a given architecture (benchmarks)."""

    def __init__(self, model_config: Any = None, checkpoint_engine=None, inf_checkpoint_path: Optional[str] = None,
                 arch: Optional[ArchSpec] = None, random_seed: Optional[int] = None,
                 dtype: Optional[torch.dtype] = None, device: Optional[torch.device] = None) -> None:
        if checkpoint_engine is None and inf_checkpoint_path is None and random_seed is None:
            raise ValueError("Either checkpoint_engine, inf_checkpoint_path or random_seed must be provided.")
        self._model_config = model_config
        self._checkpoint_engine = checkpoint_engine
        self._inf_checkpoint_path = inf_checkpoint_path
        self._arch = arch
        self._random_seed = random_seed
        self._dtype = dtype
        self._device = device

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        POLICIES[cls.__name__] = cls

    @property
    def arch(self) -> ArchSpec:
        if self._arch is None:
            self._arch = arch_from_hf_config(self._model_config)
        return self._arch

    def _resolve_dtype(self) -> torch.dtype:
        if self._dtype is not None:
            return self._dtype
        dt = getattr(self._model_config, "torch_dtype", None) or getattr(self._model_config, "dtype", None)
        if isinstance(dt, str):
            dt = getattr(torch, dt.replace("torch.", ""), None)
        if dt in (torch.float16, torch.bfloat16):
            return dt
        return torch.bfloat16

    def build_model(self, engine_config, mp_group) -> "RaggedTransformer":
        from ....models.transformer import RaggedTransformer, TPComm  # deferred: the executor imports this package's parent
        tp_size = engine_config.tensor_parallel.tp_size
        tp_rank = 0
        if tp_size > 1:
            import torch.distributed as dist
            tp_rank = dist.get_rank(group=mp_group)
        device = self._device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available()
                                  else torch.device("cpu"))
        dtype = self._resolve_dtype()
        if self._inf_checkpoint_path is not None:
            with open(make_model_config_filename(self._inf_checkpoint_path), "rb") as f:
                saved = pickle.load(f)
            self._arch = saved["arch"]
            dtype = getattr(torch, saved["dtype"])
            flat = torch.load(make_param_filename(self._inf_checkpoint_path, tp_rank, tp_size), map_location="cpu",
                              weights_only=True)
            with open(make_metadata_filename(self._inf_checkpoint_path, tp_rank, tp_size)) as f:
                meta = json.load(f)
            weights = restore_model_weights(flat, meta, device)
        elif self._checkpoint_engine is not None:
            weights = load_weights(self.arch, self._checkpoint_engine.parameters(), families.get_mapper(self.arch.model_type),
                                   tp_rank=tp_rank, tp_size=tp_size, dtype=dtype, device=device)
        else:
            weights = random_weights(self.arch, tp_rank=tp_rank, tp_size=tp_size, dtype=dtype, device=device,
                                     seed=self._random_seed)
        comm = TPComm(group=mp_group, size=tp_size, rank=tp_rank)
        if tp_size > 1 and device.type == "cuda" and engine_config.tensor_parallel.fused_collectives:
            from ....parallel.tp_comm import make_tp_comm
            comm = make_tp_comm(mp_group, tp_size, tp_rank, device)
        model = RaggedTransformer(self.arch, weights, engine_config.state_manager, dtype=dtype, device=device,
                                  tp_rank=tp_rank, tp_size=tp_size, comm=comm,
                                  capture_latents=engine_config.hcache.enabled and engine_config.hcache.capture != "off")
        mode = getattr(getattr(engine_config, "quantization", None), "quantization_mode", None)
        if mode == "wf8af8":
            model.quantize_weights_fp8()
        elif mode == "wf6af16":
            # the reference's FP6 mode (modules/implementations/linear/quantized_linear.py:96-205): here E3M2 weights are fed
            # to the tensor cores directly (kind::f8f6f4), with activations quantised to E4M3 per token instead of kept fp16
            model.quantize_weights_fp8("e3m2")
        elif mode is not None:
            raise ValueError(f"quantization_mode {mode!r} is not available: use 'wf8af8' or 'wf6af16'")
        return model


# names of the reference's per-family policies (all share the one implementation above)
class Llama2Policy(InferenceV2Policy):
    pass


class OPTPolicy(InferenceV2Policy):
    pass


class MistralPolicy(InferenceV2Policy):
    pass


class MixtralPolicy(InferenceV2Policy):
    pass


class FalconPolicy(InferenceV2Policy):
    pass


class PhiPolicy(InferenceV2Policy):
    pass


class Phi3Policy(InferenceV2Policy):
    pass


class QwenPolicy(InferenceV2Policy):
    pass


class Qwen2Policy(InferenceV2Policy):
    pass


class Qwen2MoePolicy(InferenceV2Policy):
    pass


class GPT2Policy(InferenceV2Policy):
    """GPT-2, GPT-J and GPT-NeoX have no v2 policy in the reference (v1 kernel injection only:
    ``module_inject/containers/{gpt2,gptj,gptneox}.py``); here they are ordinary families of the one executor."""


class GPTJPolicy(InferenceV2Policy):
    pass


class GPTNeoXPolicy(InferenceV2Policy):
    pass


class GPTNeoPolicy(InferenceV2Policy):
    """Alternating global / sliding-window layers (``hds_attention_next_window``), unscaled scores."""


class BloomPolicy(InferenceV2Policy):
    """ALiBi attention (``hds_attention_next_alibi``) + LayerNorm on the embeddings."""


MODEL_TYPE_TO_POLICY = {
    "gpt2": GPT2Policy, "gptj": GPTJPolicy, "gpt_neox": GPTNeoXPolicy, "bloom": BloomPolicy, "gpt_neo": GPTNeoPolicy,
    "llama": Llama2Policy, "opt": OPTPolicy, "mistral": MistralPolicy, "mixtral": MixtralPolicy, "falcon": FalconPolicy,
    "phi": PhiPolicy, "phi3": Phi3Policy, "qwen": QwenPolicy, "qwen2": Qwen2Policy, "qwen2_moe": Qwen2MoePolicy,
    "internlm": Llama2Policy,
}

__all__ = ["InferenceV2Policy", "POLICIES", "MODEL_TYPE_TO_POLICY", "supported_model_types"] + [c.__name__ for c in set(MODEL_TYPE_TO_POLICY.values())]

_LAZY = {"DSInferenceModelBase": ".inference_model_base", "DSTransformerModelBase": ".inference_model_base",
         "DSMoETransformerModelBase": ".inference_model_base", "ContainerMap": ".inference_policy_base",
         "LayerContainer": ".layer_container_base"}


def __getattr__(name):
    if name in _LAZY:
        import importlib
        value = getattr(importlib.import_module(_LAZY[name], __name__), name)
        globals()[name] = value
        return value
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


# the reference's per-family import paths (llama_v2.policy, opt.model, ...) resolve through one table: family_paths.py
from . import family_paths as _family_paths  # noqa: E402

_family_paths.register(__name__)
