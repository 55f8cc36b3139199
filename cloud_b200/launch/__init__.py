"""Local job launcher: topology discovery, rank planning, a per-job supervisor process that spawns
one rank per GPU, tees logs, watches for failures, and a file-based job registry.

This is the local replacement of "submit to AI Platform Training"
(reference: src/python/tensorflow_cloud/core/deploy.py:28-218) - see ``cloud_b200.core.deploy``
for the ``deploy_job`` entry point that keeps the reference's shape.
"""
from .topology import discover, Topology  # noqa: F401
from .plan import plan_ranks, RankPlanError  # noqa: F401
from . import registry  # noqa: F401
