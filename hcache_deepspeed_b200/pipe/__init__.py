"""``deepspeed.pipe``: public names of pipeline parallelism."""
from ..runtime.pipe import LayerSpec, PipelineModule, TiedLayerSpec  # noqa: F401
