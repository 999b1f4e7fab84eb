"""``deepspeed/nvme/perf_run_sweep.py`` entry point: run the block-size x queue-depth x threads sweep (see :mod:`.tune`)."""
from .tune import get_sweep_config_dict, run_sweep, sweep_main, sweep_points  # noqa: F401

if __name__ == "__main__":
    sweep_main()
