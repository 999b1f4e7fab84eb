from .elasticity import (ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize,  # noqa: F401
                         compute_elastic_config, elasticity_enabled, ensure_immutable_elastic_config, get_compatible_gpus,
                         is_torch_elastic_compatible)
from .constants import ELASTICITY, ENABLED, ENABLED_DEFAULT  # noqa: F401
from .elastic_agent import DSElasticAgent, WorkerState  # noqa: F401,E402
