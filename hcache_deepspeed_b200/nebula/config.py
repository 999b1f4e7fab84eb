from .constants import *  # noqa: F401,F403
from . import constants as C


class DeepSpeedNebulaConfig:

    def __init__(self, param_dict):
        d = (param_dict or {}).get(C.NEBULA, {}) or {}
        self.enabled = d.get(C.NEBULA_ENABLED, C.NEBULA_ENABLED_DEFAULT)
        self.load_path = d.get(C.NEBULA_LOAD_PATH, C.NEBULA_LOAD_PATH_DEFAULT)
        self.enable_nebula_load = d.get(C.NEBULA_ENABLE_NEBULA_LOAD, C.NEBULA_ENABLE_NEBULA_LOAD_DEFAULT)
        self.persistent_storage_path = d.get(C.NEBULA_PERSISTENT_STORAGE_PATH, C.NEBULA_PERSISTENT_STORAGE_PATH_DEFAULT)
        self.persistent_time_interval = d.get(C.NEBULA_PERSISTENT_TIME_INTERVAL, C.NEBULA_PERSISTENT_TIME_INTERVAL_DEFAULT)
        self.num_of_version_in_retention = d.get(C.NEBULA_NUM_OF_VERSION_IN_RETENTION,
                                                 C.NEBULA_NUM_OF_VERSION_IN_RETENTION_DEFAULT)

    def repr(self):
        return self.__dict__

    def __repr__(self):
        return f"DeepSpeedNebulaConfig({self.__dict__})"
