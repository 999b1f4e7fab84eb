"""Base classes of the module registry (``deepspeed/inference/v2/modules/ds_module.py``, ``module_registry.py``)."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict, Optional, Type

import torch

from ....config.config_utils import DeepSpeedConfigModel


class DSModuleConfig(DeepSpeedConfigModel):
    max_tokens: int = 768


class DSModuleBase(torch.nn.Module, ABC):
    """A module declares a ``name()``, the config class it accepts and whether it ``supports_config``."""

    @staticmethod
    @abstractmethod
    def name() -> str:
        ...

    @staticmethod
    @abstractmethod
    def config_class() -> Type[DSModuleConfig]:
        ...

    @staticmethod
    @abstractmethod
    def supports_config(config: DSModuleConfig) -> bool:
        ...

    def __init__(self, config: DSModuleConfig, implementation_config: Optional[Dict[str, Any]] = None) -> None:
        super().__init__()
        self._config = config
        self._implementation_config = implementation_config or {}


class ConfigBundle:
    """(implementation name, config, implementation options) triple handed to a registry."""

    def __init__(self, name: str, config: DSModuleConfig, implementation_config: Optional[Dict[str, Any]] = None):
        self.name, self.config, self.implementation_config = name, config, implementation_config or {}


class DSModuleRegistryBase(ABC):
    """One registry per interface; ``register_module`` is used as a class decorator."""
    registry: Dict[str, Type[DSModuleBase]]

    def __init_subclass__(cls, **kwargs):
        super().__init_subclass__(**kwargs)
        cls.registry = {}

    @classmethod
    def instantiate_config(cls, bundle: ConfigBundle) -> DSModuleBase:
        if bundle.name not in cls.registry:
            raise KeyError(f"Unknown DSModule: {bundle.name}, cls.registry={sorted(cls.registry)}")
        target = cls.registry[bundle.name]
        if not target.supports_config(bundle.config):
            raise ValueError(f"Config {bundle.config} is not supported by {target}")
        return target(bundle.config, bundle.implementation_config)

    @staticmethod
    @abstractmethod
    def associated_class() -> Type[DSModuleBase]:
        ...

    @classmethod
    def register_module(cls, child: Type[DSModuleBase]) -> Type[DSModuleBase]:
        if not issubclass(child, cls.associated_class()):
            raise TypeError(f"Can't register subclass {child} to registry of {cls.associated_class()}")
        cls.registry[child.name()] = child
        return child
