from .real_accelerator import get_accelerator, is_current_accelerator_supported, set_accelerator  # noqa: F401
