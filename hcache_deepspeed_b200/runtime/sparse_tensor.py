"""Row-sparse gradient representation for embedding layers (``sparse_gradients``): indices of non-zero rows + their
values; all-gathered instead of all-reduced. Reference: ``deepspeed/runtime/sparse_tensor.py:13-70``."""
import torch


class SparseTensor:

    def __init__(self, dense_tensor=None):
        self.orig_dense_tensor = dense_tensor
        self.dtype = dense_tensor.dtype if dense_tensor is not None else None
        if dense_tensor is not None:
            if dense_tensor.is_sparse:
                dense_tensor = dense_tensor.coalesce()
                self.indices = dense_tensor.indices().flatten()
                self.values = dense_tensor.values()
            else:
                rows = torch.sum(dense_tensor, dim=1)
                self.indices = rows.nonzero().flatten()
                self.values = dense_tensor[self.indices]
            self.dense_size = list(dense_tensor.size())
        else:
            self.indices = self.values = self.dense_size = None

    def to_coo_tensor(self):
        return torch.sparse_coo_tensor(self.indices.unsqueeze(0), self.values, self.dense_size)

    @staticmethod
    def type():
        return "deepspeed.SparseTensor"

    def to_dense(self):
        out = torch.zeros(self.dense_size, dtype=self.values.dtype, device=self.values.device)
        return out.index_add_(0, self.indices, self.values)

    def sparse_size(self):
        return self.indices.numel() + self.values.numel(), self.dense_size[0] * self.dense_size[1]

    def add(self, b):
        assert self.dense_size == b.dense_size
        self.indices = torch.cat([self.indices, b.indices])
        self.values = torch.cat([self.values, b.values])

    def __str__(self):
        s, d = self.sparse_size()
        return f"DeepSpeed.SparseTensor(indices_size={self.indices.size()}, values_size={self.values.size()}, dense_size={self.dense_size}, reduction_factor={d / max(1, s):.2f})"
