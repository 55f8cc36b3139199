"""Name kept for parity with the reference module ``utils/tf_utils.py``."""
from .framework_utils import get_version  # noqa: F401
