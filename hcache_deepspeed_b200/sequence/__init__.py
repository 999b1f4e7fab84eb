"""Sequence parallelism: DeepSpeed-Ulysses (all-to-all between sequence and head sharding around attention) and the
FPDT chunked / offloaded variant."""
from .layer import DistributedAttention  # noqa: F401
