"""Local experts of one expert-parallel rank. Reference: ``deepspeed/moe/experts.py:13-38``."""
from __future__ import annotations

import copy

import torch


class Experts(torch.nn.Module):

    def __init__(self, expert: torch.nn.Module, num_local_experts: int = 1, expert_group_name=None):
        super().__init__()
        self.deepspeed_experts = torch.nn.ModuleList([copy.deepcopy(expert) for _ in range(num_local_experts)])
        self.num_local_experts = num_local_experts
        for e in self.deepspeed_experts:
            for p in e.parameters():
                p.allreduce = False          # not replicated over the data-parallel group
                p.group_name = expert_group_name

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        """inputs: [ep_world, local_experts, capacity, model] — chunk over experts, run each, concatenate back."""
        chunks = inputs.chunk(self.num_local_experts, dim=1)
        outs = []
        for chunk, expert in zip(chunks, self.deepspeed_experts):
            out = expert(chunk)
            if isinstance(out, tuple):
                out = out[0]  # bias is ignored
            outs.append(out)
        return torch.cat(outs, dim=1)
