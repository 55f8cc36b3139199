"""machine_config -> ranks.  The AI Platform job model (one chief replica + ``workerCount``
identical workers, each with its own accelerator count; reference deploy.py:98-167 and appendix B5
of SURVEY.md) is mapped onto the GPUs of one box: ranks ``[0, chief.accelerator_count)`` form the
chief task, then ``worker_count`` groups of ``worker.accelerator_count`` ranks; a replica with no
accelerator still gets one (CPU) rank.  ``TF_CONFIG`` is synthesised per rank with the reference's
chief naming (``use_chief_in_tf_config``, deploy.py:159-161)."""
from __future__ import annotations

import json

from ..core import machine_config


class RankPlanError(RuntimeError):
    pass


def plan_ranks(chief_config, worker_count=0, worker_config=None, visible_gpus=None,
               gpu_ids=None, base_port=None):
    """Returns ``{"world_size", "ranks": [...], "cluster", "uses_gpu", "gpu_ids"}``."""
    if machine_config.is_tpu_config(worker_config) and worker_count > 0:
        raise RankPlanError("There is no TPU on this box: cloud_b200 runs on NVIDIA B200 GPUs. "
                            "Use e.g. COMMON_MACHINE_CONFIGS['B200_8X'] instead of a TPU worker.")
    tasks = [("chief", 0, chief_config)]
    for i in range(int(worker_count)):
        tasks.append(("worker", i, worker_config))
    need_gpus = sum(max(int(cfg.accelerator_count or 0), 0) for _t, _i, cfg in tasks
                    if cfg.accelerator_type != machine_config.AcceleratorType.NO_ACCELERATOR)
    if gpu_ids is None:
        gpu_ids = list(range(visible_gpus if visible_gpus is not None else need_gpus))
    if need_gpus > len(gpu_ids):
        raise RankPlanError("The job needs {} GPUs (chief {} + {} workers x {}) but only {} are "
                            "visible on this box.".format(
                                need_gpus, chief_config.accelerator_count, worker_count,
                                getattr(worker_config, "accelerator_count", 0), len(gpu_ids)))
    port0 = base_port or 20000
    cluster = {"chief": ["127.0.0.1:{}".format(port0)]}
    if worker_count:
        cluster["worker"] = ["127.0.0.1:{}".format(port0 + 1 + i) for i in range(int(worker_count))]
    ranks, next_gpu = [], 0
    for task_type, task_index, cfg in tasks:
        is_cpu = cfg.accelerator_type == machine_config.AcceleratorType.NO_ACCELERATOR
        n = 1 if is_cpu else max(int(cfg.accelerator_count), 1)
        for r in range(n):
            gpu = None
            if not is_cpu:
                gpu = gpu_ids[next_gpu]
                next_gpu += 1
            ranks.append({
                "rank": len(ranks), "task_type": task_type, "task_index": task_index,
                "rank_in_task": r, "gpu": gpu, "local_rank": (next_gpu - 1) if gpu is not None else 0,
                "tf_config": json.dumps({"cluster": cluster,
                                         "task": {"type": task_type, "index": task_index}}),
            })
    used = [r["gpu"] for r in ranks if r["gpu"] is not None]
    # LOCAL_RANK indexes into CUDA_VISIBLE_DEVICES (= the job's GPU subset, in plan order)
    for r in ranks:
        if r["gpu"] is not None:
            r["local_rank"] = used.index(r["gpu"])
    return {"world_size": len(ranks), "ranks": ranks, "cluster": cluster,
            "uses_gpu": bool(used), "gpu_ids": used}
