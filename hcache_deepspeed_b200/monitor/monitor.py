"""Event monitors: ``write_events([(tag, value, step), ...])`` fanned out to TensorBoard / W&B / CSV on rank 0
(ref: deepspeed/monitor/monitor.py:30-59 and the per-backend modules). Used by the serving loop for TTFT / tokens-per-
second series and by the training engine for loss / lr."""
from __future__ import annotations

import csv
import os
from typing import Iterable, List, Tuple

from ..utils.logging import logger
from .config import DeepSpeedMonitorConfig

Event = Tuple[str, float, int]


def _rank() -> int:
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return int(os.environ.get("RANK", 0))


class Monitor:

    def __init__(self, config):
        self.enabled = bool(config.enabled)

    def write_events(self, event_list: Iterable[Event]) -> None:
        raise NotImplementedError


class CSVMonitor(Monitor):

    def __init__(self, config):
        super().__init__(config)
        self.filenames: List[str] = []
        self.log_dir = None
        if self.enabled and _rank() == 0:
            base = config.output_path or os.path.join(os.path.expanduser("~"), "csv_monitor")
            self.log_dir = os.path.join(base, config.job_name)
            os.makedirs(self.log_dir, exist_ok=True)

    def write_events(self, event_list: Iterable[Event]) -> None:
        if not (self.enabled and _rank() == 0):
            return
        for name, value, step in event_list:
            header = name.split("/")[-1]
            fname = os.path.join(self.log_dir, name.replace("/", "_") + ".csv")
            new = fname not in self.filenames and not os.path.exists(fname)
            with open(fname, "a+", newline="") as f:
                w = csv.writer(f)
                if new:
                    w.writerow(["step", header])
                w.writerow([step, value])
            if fname not in self.filenames:
                self.filenames.append(fname)


class TensorBoardMonitor(Monitor):

    def __init__(self, config):
        super().__init__(config)
        self.summary_writer = None
        if self.enabled and _rank() == 0:
            try:
                from torch.utils.tensorboard import SummaryWriter
                base = config.output_path or os.path.join(os.path.expanduser("~"), "tensorboard")
                self.summary_writer = SummaryWriter(log_dir=os.path.join(base, config.job_name))
            except Exception as e:  # tensorboard not installed
                logger.warning(f"TensorBoard monitor disabled: {e}")
                self.enabled = False

    def write_events(self, event_list: Iterable[Event], flush: bool = True) -> None:
        if self.summary_writer is None:
            return
        for name, value, step in event_list:
            self.summary_writer.add_scalar(name, value, step)
        if flush:
            self.summary_writer.flush()


class WandbMonitor(Monitor):

    def __init__(self, config):
        super().__init__(config)
        self._wandb = None
        if self.enabled and _rank() == 0:
            try:
                import wandb
                wandb.init(project=config.project, group=config.group, entity=config.team)
                self._wandb = wandb
            except Exception as e:
                logger.warning(f"W&B monitor disabled: {e}")
                self.enabled = False

    def write_events(self, event_list: Iterable[Event]) -> None:
        if self._wandb is None:
            return
        for name, value, step in event_list:
            self._wandb.log({name: value}, step=step)


class EventsLogScheduler:
    """Rate limiter: an event name is logged at most once per ``samples_log_interval`` samples (comet.py:70-92)."""

    def __init__(self, samples_log_interval: int):
        self._interval = samples_log_interval
        self._last = {}

    def needs_logging(self, name: str, current_sample: int) -> bool:
        last = self._last.get(name)
        if last is None or current_sample - last >= self._interval:
            self._last[name] = current_sample
            return True
        return False


class CometMonitor(Monitor):
    """Comet experiment sink (``deepspeed/monitor/comet.py:23-68``); disabled with a warning when ``comet_ml`` is absent."""

    def __init__(self, config, experiment=None):
        super().__init__(config)
        self._experiment = experiment
        self._scheduler = EventsLogScheduler(config.samples_log_interval)
        if self.enabled and _rank() == 0 and experiment is None:
            try:
                import comet_ml
                self._experiment = comet_ml.start(api_key=config.api_key, project=config.project, workspace=config.workspace,
                                                  experiment_key=config.experiment_key, mode=config.mode, online=config.online)
                if config.experiment_name is not None:
                    self._experiment.set_name(config.experiment_name)
            except Exception as e:
                logger.warning(f"Comet monitor disabled: {e}")
                self.enabled = False

    @property
    def experiment(self):
        return self._experiment

    @property
    def samples_log_interval(self) -> int:
        return self._scheduler._interval

    def write_events(self, event_list: Iterable[Event]) -> None:
        if not self.enabled or self._experiment is None or _rank() != 0:
            return
        for name, value, step in event_list:
            if self._scheduler.needs_logging(name, step):
                self._experiment.log_metric(name=name, value=value, step=step)


class MonitorMaster(Monitor):

    def __init__(self, monitor_config: DeepSpeedMonitorConfig):
        self.enabled = monitor_config.enabled
        self.tb_monitor = self.wandb_monitor = self.csv_monitor = self.comet_monitor = None
        if _rank() == 0:
            if monitor_config.tensorboard.enabled:
                self.tb_monitor = TensorBoardMonitor(monitor_config.tensorboard)
            if monitor_config.wandb.enabled:
                self.wandb_monitor = WandbMonitor(monitor_config.wandb)
            if monitor_config.csv_monitor.enabled:
                self.csv_monitor = CSVMonitor(monitor_config.csv_monitor)
            if monitor_config.comet.enabled:
                self.comet_monitor = CometMonitor(monitor_config.comet)

    def write_events(self, event_list: Iterable[Event]) -> None:
        if _rank() != 0:
            return
        event_list = list(event_list)
        for m in (self.tb_monitor, self.wandb_monitor, self.csv_monitor, self.comet_monitor):
            if m is not None:
                m.write_events(event_list)
