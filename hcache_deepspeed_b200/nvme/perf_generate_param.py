"""``deepspeed/nvme/perf_generate_param.py`` entry point: best ``aio`` config section from the sweep logs."""
from .tune import convert_to_param, generate_aio_param, generate_main  # noqa: F401

if __name__ == "__main__":
    import json
    print(json.dumps(generate_main(), indent=2))
