"""Gather / scatter of kept tokens for batch-first ``[batch, seq, hidden]`` activations
(``deepspeed/runtime/data_pipeline/data_routing/utils.py``)."""
import torch

from ..random_ltd import gather_tokens, scatter_tokens


def bsh_decoder_gather(reserved_length: int, hidden_states: torch.Tensor, mask):
    """Pick ``reserved_length`` random positions (sorted, so causality is preserved); returns
    ``(kept hidden states, indices, mask restricted to the kept positions)``."""
    batch, seq = hidden_states.shape[:2]
    idx = torch.stack([torch.randperm(seq, device=hidden_states.device)[:reserved_length].sort().values for _ in range(batch)])
    part = gather_tokens(hidden_states, idx, batch_first=True)
    part_mask = None
    if mask is not None:
        part_mask = mask[..., :reserved_length, :reserved_length]
    return part, idx, part_mask


def bsh_decoder_scatter(hidden_states: torch.Tensor, part_hidden_states: torch.Tensor, rand_list: torch.Tensor) -> torch.Tensor:
    return scatter_tokens(hidden_states, part_hidden_states, rand_list, batch_first=True)
