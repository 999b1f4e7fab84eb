"""DeepCompile (``deepspeed/compile`` + ``csrc/compile``) is a ``torch.compile`` back end with ZeRO-aware graph passes.
This framework deliberately has no tracing-compiler path (streams + CUDA graphs instead), so the entry
points exist and explain that rather than silently doing nothing."""


def is_deepcompile_supported() -> bool:
    return False


def init_z1(*args, **kwargs):
    raise NotImplementedError("DeepCompile is not part of the B200-native runtime: use ZeRO stage 1 without compile")


def init_z3(*args, **kwargs):
    raise NotImplementedError("DeepCompile is not part of the B200-native runtime: use ZeRO stage 3 without compile")
