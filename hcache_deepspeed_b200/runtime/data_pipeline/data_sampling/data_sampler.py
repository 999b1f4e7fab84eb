"""Module-name parity with ``data_sampling/data_sampler.py``."""
from ..data_sampler import DeepSpeedDataSampler  # noqa: F401
