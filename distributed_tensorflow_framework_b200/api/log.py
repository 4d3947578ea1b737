"""Logging wrappers. Parity: /root/reference/distribute_log.py:8-27 (same format string, INFO level)."""
import logging

logging.basicConfig(level=logging.INFO, format="%(asctime)s  [ %(levelname)s ]: %(message)s")
_logger = logging.getLogger("dtf_b200")


def info(msg, *args):
    _logger.info(msg, *args)


def warn(msg, *args):
    _logger.warning(msg, *args)


def error(msg, *args):
    _logger.error(msg, *args)


def debug(msg, *args):
    _logger.debug(msg, *args)
