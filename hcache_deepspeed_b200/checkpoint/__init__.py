"""Checkpoint tooling: consolidate ZeRO shards to fp32, convert to / load from the "universal" (parallelism-agnostic)
format, inspect checkpoints (``deepspeed.checkpoint``)."""
from .constants import *  # noqa: F401,F403
from .deepspeed_checkpoint import (DeepSpeedCheckpoint, get_layer_ckpt_name_for_rank, get_model_ckpt_name_for_rank,  # noqa: F401
                                   get_zero_ckpt_name_for_rank)
from .ds_to_universal import convert_to_universal  # noqa: F401
from .universal_checkpoint import load_hp_checkpoint_state, load_universal_into_engine  # noqa: F401
from .zero_to_fp32 import (convert_zero_checkpoint_to_fp32_state_dict, get_fp32_state_dict_from_zero_checkpoint,  # noqa: F401
                           load_state_dict_from_zero_checkpoint)
from .reshape import (ZeROCheckpoint, get_model_3d_descriptor, meg_2d_parallel_map, model_3d_desc,  # noqa: F401
                      reshape_meg_2d_parallel)
from .reshape import clone_tensors_for_torch_save  # noqa: F401,E402
