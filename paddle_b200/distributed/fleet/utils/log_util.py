"""Fleet logger (reference: python/paddle/distributed/fleet/utils/log_util.py)."""
import logging

logger = logging.getLogger("paddle_b200.fleet")
if not logger.handlers:
    _h = logging.StreamHandler()
    _h.setFormatter(logging.Formatter("[%(asctime)s] [%(levelname)8s] %(filename)s:%(lineno)s - %(message)s"))
    logger.addHandler(_h)
    logger.propagate = False
logger.setLevel(logging.WARNING)


def set_log_level(level):
    if isinstance(level, int):
        logger.setLevel(level)
    else:
        logger.setLevel(getattr(logging, str(level).upper(), logging.INFO))


def get_log_level_code():
    return logger.getEffectiveLevel()


def get_log_level_name():
    return logging.getLevelName(logger.getEffectiveLevel())


def layer_to_str(base, *args, **kwargs):
    parts = [str(a) for a in args] + [f"{k}={v}" for k, v in kwargs.items()]
    return f"{base}({', '.join(parts)})"
