"""Framework constants. Parity: /root/reference/distribute_constants.py:14-20."""
from . import flags as _flags

# Prefix of per-device replica scopes; stripped from summary names (distribute_tower.py:137-143).
TOWER_NAME = "tower"

# The reference freezes the flag value at import time (distribute_constants.py:17); that makes the flag useless when
# it is parsed later, so it is exposed as a function as well.
INITIAL_LEARNING_RATE = _flags.FLAGS.train_learning_rate


def initial_learning_rate():
    return _flags.FLAGS.train_learning_rate


# Fraction of one epoch kept in the shuffle queue (distribute_constants.py:20, distribute_input.py:101).
MIN_FRACTION_OF_EXAMPLE_IN_QUEUE = 0.05
