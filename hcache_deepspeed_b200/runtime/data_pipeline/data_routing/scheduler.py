"""``deepspeed/runtime/data_pipeline/data_routing/scheduler.py`` names."""
from ..random_ltd import RandomLTDScheduler  # noqa: F401


class BaseScheduler:
    """Fixed-linear schedule shared by the token-drop scheduler: value(step) grows from ``min`` to ``max`` in
    ``require_steps`` steps, in increments of ``seq_per_step``."""

    def __init__(self):
        self.state = {}

    def _fixed_root_get_value(self, global_steps: int, root_degree: int = None) -> int:
        s = self.state["schedule_config"]
        root_degree = root_degree or s.get("root_degree", 1)
        frac = min(1.0, (float(global_steps) / s["require_steps"]) ** (1.0 / root_degree))
        val = int(frac * (self.state["max_value"] - self.state["min_value"]) + self.state["min_value"])
        val -= val % s["seq_per_step"]
        return min(val, self.state["max_value"])

    def get_value(self, global_steps: int) -> int:
        if self.state.get("schedule_type") == "fixed_linear":
            return self._fixed_root_get_value(global_steps, 1)
        raise RuntimeError("Unsupported random LTD schedule type")
