"""``deepspeed/utils/exceptions.py``."""


class DeprecatedException(Exception):
    pass
