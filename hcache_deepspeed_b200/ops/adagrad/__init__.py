from .cpu_adagrad import DeepSpeedCPUAdagrad
