"""``MoE``: the user-facing mixture-of-experts layer (gate + expert-parallel experts, optional residual MoE).
Reference: ``deepspeed/moe/layer.py:19-140``."""
from __future__ import annotations

from typing import Optional

import torch

from ..utils import groups
from ..utils.logging import log_dist
from .experts import Experts
from .sharded_moe import MOELayer, TopKGate


class MoE(torch.nn.Module):

    def __init__(self, hidden_size: int, expert: torch.nn.Module, num_experts: int = 1, ep_size: int = 1, k: int = 1,
                 capacity_factor: float = 1.0, eval_capacity_factor: float = 1.0, min_capacity: int = 4,
                 use_residual: bool = False, noisy_gate_policy: Optional[str] = None, drop_tokens: bool = True,
                 use_rts: bool = True, use_tutel: bool = False, enable_expert_tensor_parallelism: bool = False,
                 top2_2nd_expert_sampling: bool = True):
        super().__init__()
        self.use_residual = use_residual
        self.enable_expert_tensor_parallelism = enable_expert_tensor_parallelism
        assert num_experts % ep_size == 0, f"Number of experts ({num_experts}) should be divisible by expert parallel size ({ep_size})"
        self.ep_size, self.num_experts = ep_size, num_experts
        self.expert_group_name = f"ep_size_{ep_size}"
        self.num_local_experts = num_experts // ep_size
        assert noisy_gate_policy is None or noisy_gate_policy in ("None", "Jitter", "RSample"), \
            "Unsupported noisy_gate_policy: " + str(noisy_gate_policy)
        log_dist(f"Creating MoE layer with num_experts: {num_experts} | num_local_experts: {self.num_local_experts} | "
                 f"expert_parallel_size: {ep_size}", [0])
        experts = Experts(expert, self.num_local_experts, self.expert_group_name)
        self.deepspeed_moe = MOELayer(TopKGate(hidden_size, num_experts, k, capacity_factor, eval_capacity_factor, min_capacity,
                                               noisy_gate_policy, drop_tokens, use_rts, None, top2_2nd_expert_sampling),
                                      experts, self.expert_group_name, ep_size, self.num_local_experts, use_tutel=use_tutel)
        if use_residual:
            self.mlp = expert
            self.coefficient = torch.nn.Linear(hidden_size, 2)

    def set_deepspeed_parallelism(self, use_data_before_expert_parallel_: bool = False):
        self._create_process_groups(use_data_before_expert_parallel_)

    def _create_process_groups(self, use_data_before_expert_parallel_: bool = False):
        from .. import comm as dist
        if not dist.is_initialized():
            return
        if self.expert_group_name not in groups._get_expert_parallel_group_dict():
            if groups.mpu is None or not self.enable_expert_tensor_parallelism:
                groups._create_expert_and_data_parallel(self.ep_size, use_data_before_expert_parallel_)
            else:
                groups._create_expert_data_and_model_parallel(self.ep_size, groups.mpu, use_data_before_expert_parallel_)
        self.deepspeed_moe._set_ep_group(groups._get_expert_parallel_group(self.expert_group_name))

    def forward(self, hidden_states: torch.Tensor, used_token: Optional[torch.Tensor] = None):
        """Returns (output, auxiliary load-balancing loss, per-expert token counts)."""
        out = self.deepspeed_moe(hidden_states, used_token)
        if self.use_residual:
            mlp_out = self.mlp(hidden_states)
            if isinstance(mlp_out, tuple):
                mlp_out = mlp_out[0]
            coef = torch.nn.functional.softmax(self.coefficient(hidden_states), dim=-1)
            out = out * coef[..., 0:1] + mlp_out * coef[..., 1:]
        return out, self.deepspeed_moe.l_aux, self.deepspeed_moe.exp_counts
