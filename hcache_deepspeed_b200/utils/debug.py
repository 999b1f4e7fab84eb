"""Debug helpers: module / parameter naming maps and rank-prefixed printing (``deepspeed/utils/debug.py``)."""
import fcntl

module_names, param_names = {}, {}


def debug_clear_module_and_param_names():
    global module_names, param_names
    module_names, param_names = {}, {}


def debug_extract_module_and_param_names(model):
    global module_names, param_names
    module_names = {m: n for n, m in model.named_modules()}
    param_names = {p: n for n, p in model.named_parameters()}


def debug_module2name(module):
    return module_names.get(module, "unknown")


def debug_module2name_id(module):
    return f"name={debug_module2name(module)} id={getattr(module, 'id', id(module))}"


def debug_module2name_class(module):
    return f"name={debug_module2name(module)} {module.__class__.__name__}"


def debug_param2name(param):
    return param_names.get(param, "unknown")


def debug_param2name_id(param):
    return f"name={debug_param2name(param)} id={getattr(param, 'ds_id', id(param))}"


def debug_param2name_id_shape(param):
    return f"{debug_param2name_id(param)} shape={tuple(param.data.shape)}"


def debug_param2name_id_shape_device(param):
    return f"{debug_param2name_id_shape(param)} device={param.device}"


def debug_param2name_id_numel(param):
    return f"{debug_param2name_id(param)} numel={param.numel()}"


def debug_param2name_id_shape_status(param):
    return f"{debug_param2name_id_shape(param)} status={getattr(param, 'ds_status', None)}"


def printflock(*msgs):
    """Print with an exclusive file lock so lines of concurrent ranks do not interleave."""
    with open(__file__, "r") as fh:
        fcntl.flock(fh, fcntl.LOCK_EX)
        try:
            print(*msgs)
        finally:
            fcntl.flock(fh, fcntl.LOCK_UN)


def log_rank_file(rank, *msgs):
    with open(f"log_rank_{rank}.txt", "a") as fh:
        for m in msgs:
            fh.write(f"{m}\n")


def print_backward_tensors(tensor) -> None:
    """Walk the autograd graph below ``tensor`` and print every leaf variable it reaches (id, shape, value, grad)."""
    seen = set()

    def walk(fn):
        if fn is None or id(fn) in seen:
            return
        seen.add(id(fn))
        print(f"Backward tensors in {fn}")
        for nxt, _ in fn.next_functions:
            if nxt is None:
                continue
            var = getattr(nxt, "variable", None)
            if var is not None:
                print(nxt)
                print(f"Tensor - id: {id(var)}, shape: {var.shape}, data: {var}, grad: {var.grad}")
            else:
                walk(nxt)

    walk(getattr(tensor, "grad_fn", None))
