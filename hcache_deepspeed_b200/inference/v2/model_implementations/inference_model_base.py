"""Model-side contracts for user-defined models built from :class:`LayerContainer`s.

* :class:`DSInferenceModelBase` — what the engine needs from a model: KV accounting and a forward over a ragged batch
  (``deepspeed/inference/v2/model_implementations/inference_model_base.py:45-272``).
* :class:`DSTransformerModelBase` — architecture accessors plus the ``transform_*_param`` hooks the common parameters
  call in ``finalize``: shard for this rank (``parallel/sharding.py``), cast to the activation dtype, move to the device,
  wrap as :class:`InferenceParameter` (``inference_transformer_base.py:48-538``, ``:541-623`` for MoE).

The built-in families do not go through this path — :class:`~hcache_deepspeed_b200.models.transformer.RaggedTransformer`
assembles its weights with ``models/weights.py::WeightAssembler`` from an :class:`ArchSpec` — but a model written
against the reference's container API (AddingAModel.md) ports over by subclassing these.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Iterable, Optional, Tuple

import torch

from ....models.arch import ArchSpec
from ....parallel import sharding
from ..inference_parameter import InferenceParameter


class DSInferenceModelBase(torch.nn.Module, ABC):

    def __init__(self, config, engine_config, base_mp_group=None) -> None:
        super().__init__()
        self._config, self._engine_config, self._base_mp_group = config, engine_config, base_mp_group
        self._transformer_params, self._non_transformer_params = None, None
        self._flattened_params, self._flattened_param_metadata = None, None
        self.state_manager = None

    @property
    def config(self):
        return self._config

    model_config = config

    @property
    def engine_config(self):
        return self._engine_config

    @property
    def tp_rank(self) -> int:
        import torch.distributed as dist
        return dist.get_rank(self._base_mp_group) if dist.is_available() and dist.is_initialized() else 0

    @property
    def tp_size(self) -> int:
        import torch.distributed as dist
        return dist.get_world_size(self._base_mp_group) if dist.is_available() and dist.is_initialized() else 1

    @property
    def flattened_params(self) -> Optional[torch.Tensor]:
        return self._flattened_params

    @property
    def flattened_param_metadata(self):
        return self._flattened_param_metadata

    def set_parameters(self, transformer: Iterable, non_transformer, flattened_param_buffer: Optional[torch.Tensor] = None,
                       flattened_param_metadata=None) -> None:
        self._transformer_params, self._non_transformer_params = list(transformer), non_transformer
        self._flattened_params, self._flattened_param_metadata = flattened_param_buffer, flattened_param_metadata

    def set_state_manager(self, state_manager) -> None:
        self.state_manager = state_manager

    @abstractmethod
    def get_kv_requirements(self, sequence, max_new_tokens: int, max_new_blocks) -> Tuple[int, torch.Tensor]:
        """(tokens that can be scheduled, KV blocks needed per cache group) given ``max_new_blocks`` free blocks."""

    @abstractmethod
    def get_remaining_block_capacity(self, sequence) -> int:
        ...

    @abstractmethod
    def maybe_allocate_kv(self, sequence, n_new_tokens: int) -> None:
        ...

    @abstractmethod
    def kv_cache_config(self):
        ...

    @property
    @abstractmethod
    def max_sequence_length(self) -> int:
        ...

    def maybe_free_kv(self, sequence) -> None:
        """Hook for sliding-window models to return blocks that fell out of the window; dense models keep everything."""

    @abstractmethod
    def prepare_batch(self, wrapped_batch) -> None:
        ...

    @abstractmethod
    def forward(self, wrapped_batch) -> torch.Tensor:
        ...


class DSTransformerModelBase(DSInferenceModelBase):
    """Accessors over an :class:`ArchSpec` (``self.arch``) and the parameter transforms."""

    def __init__(self, config, engine_config, base_mp_group=None, arch: Optional[ArchSpec] = None,
                 dtype: torch.dtype = torch.bfloat16, device: Optional[torch.device] = None) -> None:
        super().__init__(config, engine_config, base_mp_group)
        self.arch = arch if arch is not None else config
        self._dtype = dtype
        self._device = device if device is not None else \
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")

    # ---- architecture ------------------------------------------------------------------------------------------
    num_layers = property(lambda self: self.arch.n_layers)
    model_dim = property(lambda self: self.arch.hidden)
    vocab_size = property(lambda self: self.arch.vocab)
    head_size = property(lambda self: self.arch.head_dim)
    n_heads = property(lambda self: self.arch.n_heads)
    n_heads_q = property(lambda self: self.arch.n_heads)
    n_heads_kv = property(lambda self: self.arch.n_kv_heads)
    intermediate_dim = property(lambda self: self.arch.intermediate)
    gated_mlp = property(lambda self: self.arch.gated_mlp)
    activation_dtype = property(lambda self: self._dtype)
    n_experts = property(lambda self: self.arch.n_experts)
    n_top_k = property(lambda self: self.arch.top_k)
    normalize_expert_scores = property(lambda self: self.arch.norm_topk_prob)
    max_sequence_length = property(lambda self: self.arch.max_positions)

    @property
    def n_heads_q_local(self) -> int:
        return sharding.get_local_heads(self.tp_rank, self.tp_size, self.n_heads_q, self.n_heads_kv)[0]

    @property
    def n_heads_kv_local(self) -> int:
        return sharding.get_local_heads(self.tp_rank, self.tp_size, self.n_heads_q, self.n_heads_kv)[1]

    # ---- parameter transforms -------------------------------------------------------------------------------------
    def _wrap(self, t: Optional[torch.Tensor]) -> Optional[InferenceParameter]:
        if t is None:
            return None
        return InferenceParameter.initialize(t.to(device=self._device, dtype=self._dtype).contiguous())

    def transform_embedding_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(sharding.shard_embedding_param(param, self.tp_rank, self.tp_size))

    def transform_unembed_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(sharding.shard_unembed_param(param, self.tp_rank, self.tp_size))

    def transform_qkv_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(sharding.shard_qkv_param(param, self.tp_rank, self.tp_size, self.head_size, self.n_heads_q,
                                                   self.n_heads_kv))

    def transform_attn_out_param(self, param: torch.Tensor) -> Optional[InferenceParameter]:
        return self._wrap(sharding.shard_attn_out_param(param, self.tp_rank, self.tp_size, self.head_size, self.n_heads_q,
                                                        self.n_heads_kv))

    def _shard_mlp_1(self, param: torch.Tensor, is_moe: bool) -> torch.Tensor:
        """Column split. Gated weights arrive with (gate, up) rows INTERLEAVED, so a shard is a contiguous run of pairs."""
        if not self.gated_mlp or self.tp_size == 1:
            return sharding.shard_mlp_1_param(param, self.tp_rank, self.tp_size, gated=False, is_moe=is_moe)
        dim = 1 if is_moe else 0
        s, e = sharding.get_shard_endpoints(param.shape[dim] // 2, self.tp_rank, self.tp_size)
        return param.narrow(dim, 2 * s, 2 * (e - s)).contiguous()

    def transform_mlp_1_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(self._shard_mlp_1(param, is_moe=False))

    def transform_mlp_2_param(self, param: torch.Tensor) -> Optional[InferenceParameter]:
        return self._wrap(sharding.shard_mlp_2_param(param, self.tp_rank, self.tp_size))

    def transform_norm_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(param)

    # ---- mixture of experts ------------------------------------------------------------------------------------------
    def transform_moe_gate_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(param)

    def transform_moe_mlp_1_param(self, param: torch.Tensor) -> InferenceParameter:
        return self._wrap(self._shard_mlp_1(param, is_moe=True))

    def transform_moe_mlp_2_param(self, param: torch.Tensor) -> Optional[InferenceParameter]:
        return self._wrap(sharding.shard_mlp_2_param(param, self.tp_rank, self.tp_size, is_moe=True))


DSMoETransformerModelBase = DSTransformerModelBase
