"""``flops_profiler`` config section (``deepspeed/profiling/config.py``)."""
from ..runtime.config import FlopsProfilerConfig as DeepSpeedFlopsProfilerConfig  # noqa: F401


def get_flops_profiler_config(param_dict: dict):
    return DeepSpeedFlopsProfilerConfig(**(param_dict.get("flops_profiler") or {}))
