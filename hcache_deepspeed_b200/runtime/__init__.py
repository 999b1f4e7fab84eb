"""Training runtime (``deepspeed.initialize`` side of the reference: ``deepspeed/runtime``). Tier C in SURVEY.md §2.9:
kept functional and torch.distributed-native (NCCL on GPUs, gloo in CPU tests) rather than kernel-optimised."""
