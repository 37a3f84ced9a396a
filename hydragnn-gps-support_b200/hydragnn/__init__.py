"""hydragnn -- B200-native message-passing hot path behind the reference's Python API."""
from . import models  # noqa: F401
