"""``comms_logger`` config (``deepspeed/comm/config.py``)."""
from __future__ import annotations

from typing import List

from ..config.config_utils import DeepSpeedConfigModel


class CommsConfig(DeepSpeedConfigModel):
    pass


class CommsLoggerConfig(CommsConfig):
    enabled: bool = False
    prof_all: bool = True
    prof_ops: List[str] = []
    verbose: bool = False
    debug: bool = False


class DeepSpeedCommsConfig:

    def __init__(self, ds_config: dict):
        self.comms_logger_enabled = "comms_logger" in ds_config
        if self.comms_logger_enabled:
            self.comms_logger = CommsLoggerConfig(**ds_config["comms_logger"])
