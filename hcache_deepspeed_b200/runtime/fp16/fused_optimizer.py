"""FP16_Optimizer: fp32 master weights (one flat buffer per param group) + loss scaling for fp16 training without
ZeRO. ``FP16_UnfusedOptimizer`` keeps per-parameter masters (needed by layer-wise optimizers such as LAMB).
Reference: ``deepspeed/runtime/fp16/fused_optimizer.py:32-500``, ``unfused_optimizer.py:26-420``."""
from __future__ import annotations

import math

import torch

from ... import comm as dist
from ...utils.logging import logger
from .loss_scaler import CreateLossScaler


class _FP16Base:
    def _init_common(self, init_optimizer, static_loss_scale, dynamic_loss_scale, dynamic_loss_args, clip_grad, mpu,
                     fused_adam_legacy=False, has_moe_layers=False, timers=None, deepspeed=None, initial_dynamic_scale=2 ** 32):
        self.optimizer = init_optimizer
        self.clip_grad = clip_grad
        self.mpu = mpu
        self.overflow = False
        self.skipped_steps = 0
        self._global_grad_norm = 0.0
        self.custom_loss_scaler = False
        self.external_loss_scale = None
        args = dict(dynamic_loss_args or {})
        if dynamic_loss_scale and "init_scale" not in args:
            args["init_scale"] = initial_dynamic_scale
        self.loss_scaler = CreateLossScaler(torch.half, static_loss_scale, dynamic_loss_scale, args)
        self.dynamic_loss_scale = self.loss_scaler.dynamic

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        return self.optimizer.state

    @property
    def cur_scale(self):
        return self.external_loss_scale if self.custom_loss_scaler else self.loss_scaler.cur_scale

    loss_scale = cur_scale

    def backward(self, loss, create_graph=False, retain_graph=False):
        (loss.float() * self.cur_scale).backward(create_graph=create_graph, retain_graph=retain_graph)

    def get_global_grad_norm(self):
        return self._global_grad_norm

    # ---- small public surface shared with the ZeRO wrappers (ref: fused_optimizer.py set_lr / get_lr /
    #      override_loss_scale / unscale_and_clip_grads / initialize_optimizer_states) ----------------------------------
    def set_lr(self, lr: float) -> None:
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    def get_lr(self) -> float:
        return self.optimizer.param_groups[0]["lr"]

    def override_loss_scale(self, loss_scale: float) -> None:
        self.custom_loss_scaler = True
        self.external_loss_scale = loss_scale

    def unscale_and_clip_grads(self, grad_groups_flat, total_norm: float, apply_scale: bool = True) -> float:
        """The divisor that removes the loss scale and clips to ``clip_grad`` in one pass; applied in place unless
        ``apply_scale`` is false."""
        combined = float(self.cur_scale)
        if self.clip_grad > 0.0:
            clip = (total_norm / self.cur_scale + 1e-6) / self.clip_grad
            if clip > 1:
                combined = clip * self.cur_scale
        if apply_scale:
            for g in grad_groups_flat:
                g.data.mul_(1.0 / combined)
        return combined

    def initialize_optimizer_states(self) -> None:
        """Create the inner optimizer's state tensors now (a step on zero gradients, masters restored afterwards)."""
        masters = [m for m in getattr(self, "fp32_groups_flat", []) if m.numel()]
        saved = [m.detach().clone() for m in masters]
        for m in masters:
            m.grad = torch.zeros_like(m)
        self.optimizer.step()
        for m, sv in zip(masters, saved):
            m.grad = None
            m.data.copy_(sv)
        for st in self.optimizer.state.values():
            if "step" in st:
                st["step"] = st["step"] * 0 if torch.is_tensor(st["step"]) else 0

    def step_fused_adam(self, closure=None):
        """Legacy entry of the reference for its apex-style fused Adam; the regular ``step`` already is one fused pass."""
        return self.step(closure)

    def _mp_reduce(self, v: float, op) -> float:
        if self.mpu is None or not dist.is_initialized():
            return v
        dev = self._device()
        t = torch.tensor([float(v)], device=dev)
        dist.all_reduce(t, op=op, group=self.mpu.get_model_parallel_group())
        return float(t[0])

    def _finish_overflow(self, overflow: bool) -> float:
        prev = self.cur_scale
        if not self.custom_loss_scaler:
            self.loss_scaler.update_scale(overflow)
        if overflow:
            self.skipped_steps += 1
            logger.info(f"[deepspeed] fp16 dynamic loss scale overflow! Skipping step. Attempted loss scale: {prev}, "
                        f"reducing to {self.cur_scale}")
        return prev


class FP16_Optimizer(_FP16Base):

    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 initial_dynamic_scale=2 ** 32, dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0,
                 fused_adam_legacy=False, has_moe_layers=False, timers=None):
        self._init_common(init_optimizer, static_loss_scale, dynamic_loss_scale, dynamic_loss_args, clip_grad, mpu,
                          initial_dynamic_scale=initial_dynamic_scale)
        self.fp16_groups, self.fp16_groups_flat, self.fp32_groups_flat = [], [], []
        for group in self.optimizer.param_groups:
            params = list(group["params"])
            self.fp16_groups.append(params)
            flat = torch.cat([p.data.contiguous().view(-1) for p in params]) if params else torch.empty(0)
            off = 0
            for p in params:
                p.data = flat.narrow(0, off, p.numel()).view_as(p.data)
                off += p.numel()
            self.fp16_groups_flat.append(flat)
            master = flat.detach().clone().float()
            master.requires_grad = True
            self.fp32_groups_flat.append(master)
            group["params"] = [master]

    def _device(self):
        return next((f.device for f in self.fp16_groups_flat if f.numel()), torch.device("cpu"))

    def zero_grad(self, set_to_none=True):
        for g in self.fp16_groups:
            for p in g:
                if set_to_none:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.detach_().zero_()

    def step(self, closure=None):
        flats = []
        for params in self.fp16_groups:
            flats.append(torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float()
                                    for p in params]) if params else torch.empty(0))
        sq = sum(float(f.norm(2)) ** 2 for f in flats if f.numel())
        bad = any((float(f.sum()) in (float("inf"), -float("inf")) or float(f.sum()) != float(f.sum())) for f in flats if f.numel())
        self.overflow = self._mp_reduce(1.0 if bad else 0.0, dist.ReduceOp.MAX) > 0
        prev = self._finish_overflow(self.overflow)
        for g in self.fp16_groups:
            for p in g:
                p.grad = None
        if self.overflow:
            return self.overflow
        scaled_norm = math.sqrt(self._mp_reduce(sq, dist.ReduceOp.SUM))
        self._global_grad_norm = scaled_norm / prev
        combined = prev
        if self.clip_grad > 0.0:
            clip = (scaled_norm / prev + 1e-6) / self.clip_grad
            if clip > 1:
                combined = clip * prev
        for master, f in zip(self.fp32_groups_flat, flats):
            master.grad = f.div_(combined) if f.numel() else None
        self.optimizer.step()
        for master, flat in zip(self.fp32_groups_flat, self.fp16_groups_flat):
            master.grad = None
            flat.copy_(master.detach())
        return self.overflow

    def state_dict(self):
        return {"loss_scaler": self.loss_scaler, "dynamic_loss_scale": self.dynamic_loss_scale, "overflow": self.overflow,
                "clip_grad": self.clip_grad, "optimizer_state_dict": self.optimizer.state_dict(),
                "fp32_groups_flat": [m.detach().clone() for m in self.fp32_groups_flat]}

    def load_state_dict(self, sd, load_optimizer_states=True):
        self.loss_scaler = sd.get("loss_scaler", self.loss_scaler)
        self.overflow = sd.get("overflow", False)
        self.clip_grad = sd.get("clip_grad", self.clip_grad)
        if load_optimizer_states:
            self.optimizer.load_state_dict(sd["optimizer_state_dict"])
        for cur, saved in zip(self.fp32_groups_flat, sd["fp32_groups_flat"]):
            cur.data.copy_(saved.data)
        self.refresh_fp16_params()

    def refresh_fp32_params(self):
        for master, flat in zip(self.fp32_groups_flat, self.fp16_groups_flat):
            master.data.copy_(flat.float())

    def refresh_fp16_params(self):
        for master, flat in zip(self.fp32_groups_flat, self.fp16_groups_flat):
            flat.copy_(master.detach())


class FP16_UnfusedOptimizer(_FP16Base):

    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0, fused_lamb_legacy=False):
        self._init_common(init_optimizer, static_loss_scale, dynamic_loss_scale, dynamic_loss_args, clip_grad, mpu,
                          initial_dynamic_scale=2 ** 16)
        self.fp16_groups, self.fp32_groups = [], []
        for group in self.optimizer.param_groups:
            params = list(group["params"])
            masters = [p.detach().clone().float().requires_grad_(True) for p in params]
            self.fp16_groups.append(params)
            self.fp32_groups.append(masters)
            group["params"] = masters

    def _device(self):
        return next((p.device for g in self.fp16_groups for p in g), torch.device("cpu"))

    def zero_grad(self, set_to_none=True):
        for g in self.fp16_groups:
            for p in g:
                p.grad = None

    def step(self, closure=None):
        grads = [[(p.grad.float() if p.grad is not None else torch.zeros_like(p, dtype=torch.float32)) for p in g]
                 for g in self.fp16_groups]
        bad = any(not bool(torch.isfinite(x).all()) for g in grads for x in g)
        self.overflow = self._mp_reduce(1.0 if bad else 0.0, dist.ReduceOp.MAX) > 0
        prev = self._finish_overflow(self.overflow)
        self.zero_grad()
        if self.overflow:
            return self.overflow
        sq = sum(float(x.norm(2)) ** 2 for g in grads for x in g)
        scaled_norm = math.sqrt(self._mp_reduce(sq, dist.ReduceOp.SUM))
        self._global_grad_norm = scaled_norm / prev
        combined = prev
        if self.clip_grad > 0.0:
            clip = (scaled_norm / prev + 1e-6) / self.clip_grad
            if clip > 1:
                combined = clip * prev
        for masters, gs in zip(self.fp32_groups, grads):
            for m, g in zip(masters, gs):
                m.grad = g.div_(combined)
        self.optimizer.step()
        for masters, params in zip(self.fp32_groups, self.fp16_groups):
            for m, p in zip(masters, params):
                m.grad = None
                p.data.copy_(m.detach())
        return self.overflow

    def state_dict(self):
        return {"loss_scaler": self.loss_scaler, "overflow": self.overflow, "optimizer_state_dict": self.optimizer.state_dict(),
                "fp32_groups": [[m.detach().clone() for m in g] for g in self.fp32_groups]}

    def load_state_dict(self, sd, load_optimizer_states=True):
        self.loss_scaler = sd.get("loss_scaler", self.loss_scaler)
        if load_optimizer_states:
            self.optimizer.load_state_dict(sd["optimizer_state_dict"])
        for cur_g, saved_g in zip(self.fp32_groups, sd["fp32_groups"]):
            for c, s in zip(cur_g, saved_g):
                c.data.copy_(s.data)
        for masters, params in zip(self.fp32_groups, self.fp16_groups):
            for m, p in zip(masters, params):
                p.data.copy_(m.detach())
