"""``deepspeed/nvme/parse_nvme_stats.py``: read the sweep logs back, sorted by the chosen metric."""
from .tune import READ_SPEED, WRITE_SPEED, get_sorted_results  # noqa: F401
