"""``deepspeed/runtime/data_pipeline/data_routing/helper.py`` names."""
from ..random_ltd import convert_to_random_ltd, remove_random_ltd_state_dict, save_without_random_ltd  # noqa: F401
