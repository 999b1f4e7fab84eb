"""Flat (single-buffer) model serialisation under its reference names
(``deepspeed/inference/v2/model_implementations/flat_model_helpers.py:59-282``): all weights of a rank are packed into
one aligned byte buffer + a JSON metadata tree, so a saved engine loads with one read and no per-tensor allocation.
Implementation: :mod:`hcache_deepspeed_b200.models.flat_model`."""
from ....models.flat_model import (flatten_model_weights, make_metadata_filename, make_model_config_filename,  # noqa: F401
                                   make_param_filename, pad_to_aligned_offset, restore_model_weights)


def flatten_inference_model(model_or_weights):
    """``(flat_buffer, metadata)`` of a model's weights (accepts the model or its weight tree)."""
    return flatten_model_weights(getattr(model_or_weights, "w", model_or_weights))


def restore_inference_model(buffer, metadata, device):
    return restore_model_weights(buffer, metadata, device)
