from . import offload_activation, offload_adam_states, offload_parameters, prefetch, selective_gather, zero1_compile, zero3_compile  # noqa: F401
