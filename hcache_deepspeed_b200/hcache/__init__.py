from .format import StateHeader, plan_layout  # noqa: F401
from .planner import LinkModel, RestorePlan, measure_link, plan_restore, storage_bytes  # noqa: F401
from .session import HCacheSessionManager  # noqa: F401
from .store import HiddenStateStore, SessionState  # noqa: F401
