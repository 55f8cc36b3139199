"""cloud_b200 - placeholder facade, filled in below."""
from .version import __version__  # noqa: F401
