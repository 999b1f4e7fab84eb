"""Argument parsing of the aio benchmark (``deepspeed/nvme/ds_aio_args.py``)."""
import argparse

from .tune import parse_size


def parse_arguments(argv=None):
    p = argparse.ArgumentParser(description="asynchronous file I/O benchmark")
    p.add_argument("--folder", "--nvme_dir", dest="folder", type=str, required=True, help="directory on the device under test")
    p.add_argument("--io_size", type=str, default="400M")
    p.add_argument("--read", action="store_true", help="measure reads (default: writes)")
    p.add_argument("--block_size", type=str, default="1M")
    p.add_argument("--queue_depth", type=int, default=32)
    p.add_argument("--io_parallel", "--threads", dest="threads", type=int, default=1)
    p.add_argument("--single_submit", action="store_true")
    p.add_argument("--sequential_requests", dest="overlap_events", action="store_false")
    p.add_argument("--validate", action="store_true", help="read back and compare")
    p.add_argument("--handle", action="store_true", help="use the handle API (always true here)")
    p.add_argument("--loops", type=int, default=3)
    p.add_argument("--gpu", action="store_true", help="stage through device memory")
    p.add_argument("--use_gds", action="store_true")
    p.add_argument("--no_direct", dest="use_direct", action="store_false", help="go through the page cache instead of O_DIRECT")
    a = p.parse_args(argv)
    a.io_size, a.block_size = parse_size(a.io_size), parse_size(a.block_size)
    return a
