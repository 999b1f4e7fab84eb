"""hcache_deepspeed_b200 — a B200-native (sm_100a) LLM inference / serving framework with the
capabilities of the HCache fork of DeepSpeed. See DESIGN.md for the architecture."""
from .version import __version__  # noqa: F401
