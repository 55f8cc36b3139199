"""Experimental APIs (reference: src/python/tensorflow_cloud/experimental)."""
