"""Version of cloud_b200 (the reference keeps its own in src/python/tensorflow_cloud/version.py:16)."""
__version__ = "0.1.0"
