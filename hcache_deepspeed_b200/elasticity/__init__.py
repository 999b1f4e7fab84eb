from .elasticity import (ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize,  # noqa: F401
                         compute_elastic_config, elasticity_enabled, get_compatible_gpus)
from .elastic_agent import DSElasticAgent, WorkerState  # noqa: F401,E402
