from .config_utils import DeepSpeedConfigModel, DtypeEnum, get_config_default, pp_int  # noqa: F401
