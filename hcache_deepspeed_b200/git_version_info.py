"""Build information (``deepspeed/git_version_info.py``)."""
from .version import __version__ as version

git_hash = "unknown"
git_branch = "unknown"
installed_ops = {"hds_b200": True, "hds_host": True}
accelerator_name = "cuda"
torch_info = {"version": "0.0", "cuda_version": "0.0", "hip_version": "0.0", "bf16_support": True}
try:
    import torch as _torch
    torch_info["version"] = ".".join(_torch.__version__.split(".")[:2])
    torch_info["cuda_version"] = _torch.version.cuda or "0.0"
except Exception:       # pragma: no cover
    pass
