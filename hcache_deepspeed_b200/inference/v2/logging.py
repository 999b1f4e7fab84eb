"""Logger of the v2 inference stack (``deepspeed/inference/v2/logging.py``)."""
import logging

from ...utils.logging import logger as _root

inf_logger = None


def inference_logger(level: int = logging.INFO) -> logging.Logger:
    global inf_logger
    if inf_logger is None:
        inf_logger = _root.getChild("inference") if hasattr(_root, "getChild") else logging.getLogger("hds-inference")
        inf_logger.setLevel(level)
        inf_logger.debug("Inference logger created.")
    return inf_logger
