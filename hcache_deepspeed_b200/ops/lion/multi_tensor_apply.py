"""``multi_tensor_apply`` (``deepspeed/ops/{adam,lion}/multi_tensor_apply.py``): call a fused multi-tensor op over lists of
tensors in chunks."""


class MultiTensorApply:

    def __init__(self, chunk_size: int):
        self.chunk_size = chunk_size

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
