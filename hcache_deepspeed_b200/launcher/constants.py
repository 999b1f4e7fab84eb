"""Launcher names (``deepspeed/launcher/constants.py``)."""
PDSH_LAUNCHER, PDSH_MAX_FAN_OUT = "pdsh", 1024
OPENMPI_LAUNCHER, MPICH_LAUNCHER, IMPI_LAUNCHER, SLURM_LAUNCHER, MVAPICH_LAUNCHER = "openmpi", "mpich", "impi", "slurm", "mvapich"
MVAPICH_TMP_HOSTFILE = "/tmp/deepspeed_mvapich_hostfile"
ELASTIC_TRAINING_ID_DEFAULT = "123456789"
