"""BF16_Optimizer: bf16 parameters with fp32 master weights and fp32 gradient accumulation, master weights and
optimizer state partitioned over the data-parallel group (ZeRO-1 layout, no loss scaling).
Built on the ZeRO stage-1 machinery with an fp32 accumulator that is fed after every micro-batch
(``immediate_grad_update``) or at the boundary. Reference: ``deepspeed/runtime/bf16_optimizer.py:35-564``."""
from __future__ import annotations

import torch

from .zero.stage_1_and_2 import DeepSpeedZeroOptimizer


class BF16_Optimizer(DeepSpeedZeroOptimizer):

    def __init__(self, init_optimizer, param_names=None, mpu=None, clip_grad: float = 0.0, norm_type: int = 2,
                 allgather_bucket_size: int = 5000000000, dp_process_group=None, timers=None,
                 grad_acc_dtype=torch.float32, graph_harvesting: bool = False, immediate_grad_update: bool = False,
                 has_moe_layers: bool = False):
        super().__init__(init_optimizer, param_names=param_names, timers=timers, static_loss_scale=1.0,
                         dynamic_loss_scale=False, mpu=mpu, clip_grad=clip_grad, dp_process_group=dp_process_group,
                         gradient_accumulation_dtype=grad_acc_dtype or torch.float32, partition_grads=False,
                         reduce_scatter=False, check_grad_overflow=False, verbose=False)
        self.norm_type = norm_type
        self.immediate_grad_update = immediate_grad_update
        self._hp_accum = [None] * len(self.bit16_groups)  # full-size fp32 gradient accumulators

    def backward(self, loss, retain_graph: bool = False, **kwargs):
        self.micro_step_id += 1
        loss.backward(retain_graph=retain_graph)
        self.update_hp_grads()

    @torch.no_grad()
    def update_hp_grads(self, clear_lp_grads: bool = True):
        """Fold the bf16 gradients of this micro-batch into the fp32 accumulators."""
        for gi, params in enumerate(self.bit16_groups):
            if not params:
                continue
            flat = self.bit16_groups_flat[gi]
            if self._hp_accum[gi] is None:
                self._hp_accum[gi] = torch.zeros(flat.numel(), dtype=self.gradient_accumulation_dtype, device=flat.device)
            off = 0
            for p in params:
                if p.grad is not None:
                    self._hp_accum[gi].narrow(0, off, p.numel()).add_(p.grad.detach().reshape(-1).to(self._hp_accum[gi].dtype))
                    if clear_lp_grads:
                        p.grad = None
                off += p.numel()

    def _flat_grads(self, gi: int) -> torch.Tensor:
        acc = self._hp_accum[gi]
        if acc is None:
            return super()._flat_grads(gi).float()
        self._hp_accum[gi] = None
        return acc

    def clear_hp_grads(self):
        self._hp_accum = [None] * len(self.bit16_groups)

    def clear_lp_grads(self):
        self.zero_grad()
