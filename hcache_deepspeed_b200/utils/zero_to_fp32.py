"""``deepspeed.utils.zero_to_fp32`` entry point (the implementation lives in :mod:`hcache_deepspeed_b200.checkpoint`)."""
from ..checkpoint.zero_to_fp32 import *  # noqa: F401,F403
from ..checkpoint.zero_to_fp32 import main  # noqa: F401

if __name__ == "__main__":
    main()
