import logging


def get_logger(log_level, name="root"):
    logger = logging.getLogger(name)
    logger.propagate = False
    if not logger.handlers:
        logger.setLevel(log_level)
        h = logging.StreamHandler()
        h.setFormatter(logging.Formatter("%(levelname)s %(asctime)s %(filename)s:%(lineno)d] %(message)s"))
        logger.addHandler(h)
    return logger
