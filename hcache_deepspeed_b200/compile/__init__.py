"""Schedule compiler — this framework's counterpart of DeepCompile (``deepspeed/compile`` + ``csrc/compile``).

DeepCompile is a ``torch.compile`` back end with ZeRO-aware FX-graph passes. This runtime keeps tracing compilers off
its path (streams + CUDA graphs instead) and gets the same optimisations — prefetching of parameter
gathers, selective (persistent) gathering, optimizer-state / activation / parameter offload — by scheduling on the
MODULE trace that the ZeRO-3 hooks record: see :mod:`.backend`."""
from .backend import launch_compile_passes, make_backend, opt_passes, register_compile_pass  # noqa: F401
from .config import CompileConfig  # noqa: F401
from .init_z1 import init_z1  # noqa: F401
from .init_z3 import init_z3  # noqa: F401
from .trace import ModuleTrace, TraceEvent  # noqa: F401
from .util import is_deepcompile_supported  # noqa: F401
