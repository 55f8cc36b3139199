"""``cloud_fit``: run ``model.fit`` of an in-memory model + dataset as a separately launched job
(reference: src/python/tensorflow_cloud/experimental/cloud_fit/{client,remote,utils}.py)."""
from .client import cloud_fit  # noqa: F401
