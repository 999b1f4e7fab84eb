"""``deepspeed/elasticity/config.py`` names."""
from .elasticity import ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize  # noqa: F401
