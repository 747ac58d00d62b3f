"""Same logger name/format as /root/reference/gance/logger_common.py:8-17."""
import logging

LOGGER_FORMAT = "[%(asctime)s - %(process)s - %(name)20s - %(levelname)s] %(message)s"
LOGGER_DATE_FORMAT = "%y-%m-%d %H:%M:%S"

logging.basicConfig(level=logging.INFO, format=LOGGER_FORMAT, datefmt=LOGGER_DATE_FORMAT)
LOGGER = logging.getLogger("gance_b200")
