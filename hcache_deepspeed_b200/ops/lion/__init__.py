from .lion import DeepSpeedCPULion, FusedLion
