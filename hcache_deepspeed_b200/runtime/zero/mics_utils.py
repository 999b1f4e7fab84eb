"""Module-name parity with ``deepspeed/runtime/zero/mics_utils.py``; the code lives in :mod:`.mics`."""
from .mics import MiCS_CommGroups, create_mics_comm_groups, mics_group_ranks  # noqa: F401
