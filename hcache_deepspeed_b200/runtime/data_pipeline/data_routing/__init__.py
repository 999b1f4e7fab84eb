"""Package-name parity with ``deepspeed/runtime/data_pipeline/data_routing`` (random layer-wise token dropping)."""
from ..random_ltd import (RandomLayerTokenDrop, RandomLTDScheduler, convert_to_random_ltd, gather_tokens,  # noqa: F401
                          remove_random_ltd_state_dict, save_without_random_ltd, scatter_tokens)
