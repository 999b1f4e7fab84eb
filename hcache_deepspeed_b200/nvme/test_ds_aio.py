"""Command-line aio benchmark (``deepspeed/nvme/test_ds_aio.py``): ``python -m hcache_deepspeed_b200.nvme.test_ds_aio
--folder /mnt/nvme --read``; prints GB/s."""
from .ds_aio_args import parse_arguments
from .ds_aio_handle import aio_handle_multiprocessing


def main(argv=None) -> float:
    args = parse_arguments(argv)
    if args.read:                       # reads need a file: write it first
        aio_handle_multiprocessing(args, False)
    gbps = aio_handle_multiprocessing(args, args.read)
    print(f"{'read' if args.read else 'write'}: {gbps:.2f} GB/s (block {args.block_size}, depth {args.queue_depth}, threads {args.threads})")
    return gbps


if __name__ == "__main__":
    main()
