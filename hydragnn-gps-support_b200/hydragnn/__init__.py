"""hydragnn -- B200-native message-passing hot path behind the reference's Python API."""
from . import models  # noqa: F401
from .run_training import run_training  # noqa: F401
from .run_prediction import run_prediction  # noqa: F401
