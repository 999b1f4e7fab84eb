"""Dense-op wrappers (``deepspeed/inference/v2/kernels/core_ops/``: ``blas_kernels``, ``cuda_layer_norm``,
``cuda_rms_norm``, ``bias_activations``, ``gated_activations``). In this framework most of these are EPILOGUES of the
tcgen05 GEMM or fused into the norm kernel; the stand-alone objects exist for code written against the reference's
kernel classes and route to the same native entry points."""
from __future__ import annotations

from typing import Optional

import torch

from ....ops import reference as ref
from .ds_kernel import DSKernelBase

ACT_IDS = {"identity": 0, "gelu": 1, "relu": 2, "silu": 3, "gelu_tanh": 4, "geglu": 1, "reglu": 2, "siglu": 3, "swiglu": 3}


def _act_id(act) -> int:
    name = str(getattr(act, "name", getattr(act, "value", act))).lower()
    if name not in ACT_IDS:
        raise ValueError(f"Unsupported activation: {act}")
    return ACT_IDS[name]


class BlasLibLinear(DSKernelBase):
    """``out = x @ w.T`` (weights ``[out, in]``)."""

    def __init__(self, fp_dtype: torch.dtype):
        self._check_dtype(fp_dtype)

    def __call__(self, output: torch.Tensor, hidden_states: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
        x2, o2 = hidden_states.reshape(-1, hidden_states.shape[-1]), output.view(-1, output.shape[-1])
        self._ops(x2).gemm(x2, weights, o2)
        return output


class _NormBase(DSKernelBase):
    layer_norm = False

    def __init__(self, channels: int, fp_dtype: torch.dtype, epsilon: float = 1e-5, gamma_dtype: Optional[torch.dtype] = None):
        self._check_dtype(fp_dtype)
        elem = torch.tensor([], dtype=fp_dtype).element_size()
        if channels * elem % 16 != 0:
            raise ValueError("channels must be divisible by 16 bytes")
        self.epsilon = epsilon


class CUDARMSNorm(_NormBase):

    def __call__(self, output: torch.Tensor, x: torch.Tensor, gamma: torch.Tensor) -> torch.Tensor:
        self._ops(x).norm(x.view(-1, x.shape[-1]), gamma, output.view(-1, x.shape[-1]), eps=self.epsilon)
        return output


class CUDARMSPreNorm(_NormBase):
    """``residual += hidden; out = rms_norm(residual)`` in one pass."""

    def __call__(self, z_res: torch.Tensor, z_hid: torch.Tensor, x_res: torch.Tensor, x_hid: torch.Tensor, gamma: torch.Tensor):
        H = x_res.shape[-1]
        self._ops(x_res).norm(x_res.view(-1, H), gamma, z_hid.view(-1, H), delta=x_hid.view(-1, H), res_out=z_res.view(-1, H),
                              eps=self.epsilon)
        return z_res, z_hid


class CUDAFPLN(_NormBase):
    layer_norm = True

    def __call__(self, output: torch.Tensor, x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor) -> torch.Tensor:
        self._ops(x).norm(x.view(-1, x.shape[-1]), gamma, output.view(-1, x.shape[-1]), beta=beta, eps=self.epsilon, layer_norm=True)
        return output


class CUDAFPPreLN(_NormBase):
    layer_norm = True

    def __call__(self, z_res, z_hid, x_res, x_hid, gamma, beta):
        H = x_res.shape[-1]
        self._ops(x_res).norm(x_res.view(-1, H), gamma, z_hid.view(-1, H), beta=beta, delta=x_hid.view(-1, H),
                              res_out=z_res.view(-1, H), eps=self.epsilon, layer_norm=True)
        return z_res, z_hid


class CUDAFPPostLN(_NormBase):
    """``out = layer_norm(residual + hidden)`` (post-LN blocks)."""
    layer_norm = True

    def __call__(self, z, x_res, x_hid, gamma, beta):
        H = x_res.shape[-1]
        tmp = torch.empty_like(x_res)
        self._ops(x_res).norm(x_res.view(-1, H), gamma, z.view(-1, H), beta=beta, delta=x_hid.view(-1, H),
                              res_out=tmp.view(-1, H), eps=self.epsilon, layer_norm=True)
        return z


class CUDABiasActivation(DSKernelBase):
    """In-place ``act(x + bias)``. (On the hot path this is the GEMM epilogue; the stand-alone form is elementwise.)"""

    def __init__(self, channels: int, dtype: torch.dtype, act_fn):
        self._check_dtype(dtype)
        self.act = _act_id(act_fn)

    def __call__(self, activation: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        y = activation.float() + (bias.float() if bias is not None else 0.0)
        activation.copy_(ref.apply_act(y, self.act))
        return activation


class CUDAGatedActivation(DSKernelBase):
    """``out = act(x[..., 0::2] + b) * (x[..., 1::2] + b)`` — gate and up interleaved, as the fused MLP-1 weight is laid out."""

    def __init__(self, channels: int, fp_dtype: torch.dtype, act_fn):
        self._check_dtype(fp_dtype)
        if channels % 2:
            raise ValueError("channels must be even for a gated activation")
        self.act = _act_id(act_fn)

    def __call__(self, output: torch.Tensor, input: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        y = input.float() + (bias.float() if bias is not None else 0.0)
        output.copy_(ref.apply_act(y[..., 0::2], self.act) * y[..., 1::2])
        return output
