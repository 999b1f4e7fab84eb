"""Mixture-of-experts layers for training and v1 inference with expert parallelism (``deepspeed.moe``)."""
from .layer import MoE  # noqa: F401
