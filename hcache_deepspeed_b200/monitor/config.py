"""Monitor configuration (ref: deepspeed/monitor/config.py)."""
from typing import Optional

from pydantic import Field

from ..config import DeepSpeedConfigModel


class TensorBoardConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class WandbConfig(DeepSpeedConfigModel):
    enabled: bool = False
    group: Optional[str] = None
    team: Optional[str] = None
    project: str = "deepspeed"


class CSVConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class CometConfig(DeepSpeedConfigModel):
    enabled: bool = False
    samples_log_interval: int = 100
    project: Optional[str] = None
    workspace: Optional[str] = None
    api_key: Optional[str] = None
    experiment_name: Optional[str] = None
    experiment_key: Optional[str] = None
    online: Optional[bool] = None
    mode: Optional[str] = None


class DeepSpeedMonitorConfig(DeepSpeedConfigModel):
    tensorboard: TensorBoardConfig = Field(default_factory=TensorBoardConfig)
    comet: CometConfig = Field(default_factory=CometConfig)
    wandb: WandbConfig = Field(default_factory=WandbConfig)
    csv_monitor: CSVConfig = Field(default_factory=CSVConfig)

    @property
    def enabled(self) -> bool:
        return self.tensorboard.enabled or self.wandb.enabled or self.csv_monitor.enabled or self.comet.enabled


def get_monitor_config(param_dict) -> DeepSpeedMonitorConfig:
    keys = ("tensorboard", "wandb", "csv_monitor", "comet")
    return DeepSpeedMonitorConfig(**{k: param_dict.get(k, {}) for k in keys})
