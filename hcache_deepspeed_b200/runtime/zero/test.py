"""Manual check of the contiguous allocator (the reference keeps a script under this name): run as a module."""
import torch

from .contiguous_memory_allocator import ContiguousMemoryAllocator


def test1():
    mem = ContiguousMemoryAllocator(1024, torch.half, "cpu")
    mem.print_allocation(resolution=100)
    a1, a2, a3, a4 = mem.allocate_tensor(64), mem.allocate_tensor(64), mem.allocate_tensor(256), mem.allocate_tensor(128)
    for i, t in enumerate((a1, a2, a3, a4)):
        t.fill_(i + 1)
    mem.release_tensor(a2)
    mem.release_tensor(a4)
    a5 = mem.allocate_tensor(512 + 128)      # forces compaction
    mem.print_allocation(resolution=100)
    assert float(a1.sum()) == 64 * 1 and float(a3.sum()) == 256 * 3 and a5.numel() == 640


if __name__ == "__main__":
    test1()
