"""Memory-mapped indexed datasets in the Megatron ``.bin`` / ``.idx`` format (token data and the per-sample metric
files of the data analyzer). File layout is kept bit-compatible with the reference so existing datasets load:

``.idx``: ``b"MMIDIDX\\0\\0"`` | u64 version (1) | u8 dtype code | u64 n_items | u64 n_docs | i32 sizes[n_items]
| i64 byte_offsets[n_items] | i64 doc_idx[n_docs];  ``.bin``: the items back to back.

Reference: ``deepspeed/runtime/data_pipeline/data_sampling/indexed_dataset.py:100-135`` (dtype codes, paths),
``:369-573`` (reader), ``:575-627`` (builder). The legacy non-mmap ("lazy" / "cached") readers are not reproduced.
"""
from __future__ import annotations

import os
import shutil
import struct
from typing import List, Optional, Sequence

import numpy as np
import torch

_MAGIC = b"MMIDIDX\x00\x00"
dtypes = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.uint16, 7: np.uint32, 8: np.uint64}
_TORCH_TO_NP = {torch.uint8: np.uint8, torch.int8: np.int8, torch.int16: np.int16, torch.int32: np.int32, torch.int64: np.int64}


def code(dtype) -> int:
    dtype = _TORCH_TO_NP.get(dtype, dtype)
    for c, dt in dtypes.items():
        if np.dtype(dtype) == np.dtype(dt):
            return c
    raise ValueError(f"{dtype} not supported; supported: {sorted(str(np.dtype(d)) for d in dtypes.values())}")


def best_fitting_dtype(vocab_size: Optional[int] = None):
    return np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32


def index_file_path(prefix_path: str) -> str:
    return prefix_path + ".idx"


def data_file_path(prefix_path: str) -> str:
    return prefix_path + ".bin"


def create_doc_idx(sizes: Sequence[int]) -> List[int]:
    """Legacy convention: a zero-size item terminates a document."""
    return [0] + [i + 1 for i, s in enumerate(sizes) if s == 0]


class MMapIndexedDataset(torch.utils.data.Dataset):

    def __init__(self, path: str, skip_warmup: bool = True):
        self._path = path
        with open(index_file_path(path), "rb") as f:
            assert f.read(9) == _MAGIC, "index file does not match the mmap indexed dataset format"
            (version, ) = struct.unpack("<Q", f.read(8))
            assert version == 1
            (dcode, ) = struct.unpack("<B", f.read(1))
            n, n_docs = struct.unpack("<QQ", f.read(16))
            offset = f.tell()
        self._dtype = dtypes[dcode]
        self._index = np.memmap(index_file_path(path), mode="r", order="C")
        self._sizes = np.frombuffer(self._index, dtype=np.int32, count=n, offset=offset)
        self._pointers = np.frombuffer(self._index, dtype=np.int64, count=n, offset=offset + self._sizes.nbytes)
        self._doc_idx = np.frombuffer(self._index, dtype=np.int64, count=n_docs,
                                      offset=offset + self._sizes.nbytes + self._pointers.nbytes)
        self._data = np.memmap(data_file_path(path), mode="r", order="C") if os.path.getsize(data_file_path(path)) else \
            np.empty(0, dtype=np.uint8)

    def __getstate__(self):
        return self._path

    def __setstate__(self, state):
        self.__init__(state)

    def __len__(self) -> int:
        return len(self._sizes)

    def get(self, idx: int, offset: int = 0, length: Optional[int] = None) -> np.ndarray:
        """Items ``offset .. offset+length`` of sample ``idx`` (zero-copy view into the mapping)."""
        size = int(self._sizes[idx])
        length = size - offset if length is None else length
        start = int(self._pointers[idx]) + offset * np.dtype(self._dtype).itemsize
        return np.frombuffer(self._data, dtype=self._dtype, count=length, offset=start)

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            return self.get(int(idx))
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            sizes = self._sizes[start:stop]
            flat = np.frombuffer(self._data, dtype=self._dtype, count=int(sizes.sum()), offset=int(self._pointers[start])) \
                if stop > start else np.empty(0, dtype=self._dtype)
            return np.split(flat, np.cumsum(sizes)[:-1]) if stop > start else []
        raise TypeError(f"Unexpected index type {type(idx)}")

    @property
    def sizes(self) -> np.ndarray:
        return self._sizes

    def size(self, index: int) -> int:
        return int(self._sizes[index])

    @property
    def doc_idx(self) -> np.ndarray:
        return self._doc_idx

    def get_doc_idx(self) -> np.ndarray:
        return self._doc_idx

    def set_doc_idx(self, doc_idx_) -> None:
        self._doc_idx = doc_idx_

    @property
    def dtype(self):
        return self._dtype

    @property
    def supports_prefetch(self) -> bool:
        return False

    @staticmethod
    def exists(path: str) -> bool:
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))


class MMapIndexedDatasetBuilder:

    def __init__(self, out_file: str, dtype=np.int64):
        self._file = open(out_file, "wb")
        self._dtype = np.dtype(_TORCH_TO_NP.get(dtype, dtype)).type
        self._sizes: List[int] = []
        self._doc_idx: List[int] = [0]

    def add_item(self, tensor) -> None:
        self.add_item_numpy(tensor.numpy() if torch.is_tensor(tensor) else np.asarray(tensor))

    def add_item_numpy(self, np_array: np.ndarray) -> None:
        arr = np.ascontiguousarray(np_array, dtype=self._dtype)
        self._file.write(arr.tobytes(order="C"))
        self._sizes.append(arr.size)

    def add_items(self, arr_list) -> None:
        for a in arr_list:
            self.add_item_numpy(np.asarray(a))

    def end_document(self) -> None:
        self._doc_idx.append(len(self._sizes))

    def merge_file_(self, another_file: str) -> None:
        """Append another dataset (same dtype) built under prefix ``another_file``."""
        other = MMapIndexedDataset(another_file)
        assert np.dtype(other.dtype) == np.dtype(self._dtype)
        base = len(self._sizes)
        self._sizes.extend(int(s) for s in other.sizes)
        self._doc_idx.extend(int(d) + base for d in other.doc_idx[1:])
        with open(data_file_path(another_file), "rb") as f:
            shutil.copyfileobj(f, self._file)

    def finalize(self, index_file: str) -> None:
        self._file.close()
        sizes = np.asarray(self._sizes, dtype=np.int64)
        pointers = np.zeros(len(sizes), dtype=np.int64)
        if len(sizes) > 1:
            np.cumsum(sizes[:-1] * np.dtype(self._dtype).itemsize, out=pointers[1:])
        with open(index_file, "wb") as f:
            f.write(_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", code(self._dtype)))
            f.write(struct.pack("<QQ", len(sizes), len(self._doc_idx)))
            f.write(sizes.astype(np.int32).tobytes(order="C"))
            f.write(pointers.tobytes(order="C"))
            f.write(np.asarray(self._doc_idx, dtype=np.int64).tobytes(order="C"))


def get_available_dataset_impl() -> List[str]:
    return ["mmap"]


def infer_dataset_impl(path: str) -> Optional[str]:
    if not os.path.exists(index_file_path(path)):
        return None
    with open(index_file_path(path), "rb") as f:
        return "mmap" if f.read(9) == _MAGIC else None


def make_builder(out_file: str, impl: str = "mmap", vocab_size: Optional[int] = None) -> MMapIndexedDatasetBuilder:
    assert impl == "mmap", "only the mmap implementation is provided"
    return MMapIndexedDatasetBuilder(out_file, dtype=best_fitting_dtype(vocab_size))


def make_dataset(path: str, impl: str = "mmap", skip_warmup: bool = True) -> Optional[MMapIndexedDataset]:
    if not MMapIndexedDataset.exists(path):
        return None
    impl = infer_dataset_impl(path) if impl == "infer" else impl
    assert impl == "mmap", "only the mmap implementation is provided"
    return MMapIndexedDataset(path, skip_warmup)


def dataset_exists(path: str, impl: str = "mmap") -> bool:
    return MMapIndexedDataset.exists(path)
