"""``cloud_b200.distribute`` - the strategies the run wrapper injects.

The reference only *names* ``tf.distribute`` strategies in the wrapper script it generates
(core/preprocess.py:124-149); everything below the name is TensorFlow.  Here the strategies are
real objects on top of one process per GPU:

* ``OneDeviceStrategy(device)``      - one replica, no communication (preprocess.py:144-146)
* ``MirroredStrategy()``             - one replica per local GPU (preprocess.py:142-143)
* ``MultiWorkerMirroredStrategy()``  - chief + workers; on the single B200 box the workers are
  rank groups of the same job, so it shares the engine with MirroredStrategy (preprocess.py:137-141)

All of them expose what user code touches: ``scope()``,
``num_replicas_in_sync``, ``experimental_distribute_dataset``, ``run`` /
``experimental_run_v2``, ``reduce``; ``experimental_set_strategy`` installs a process-global
default so that scripts run unmodified under the injected strategy.
"""
from .strategy import (  # noqa: F401
    ReduceOp, Strategy, OneDeviceStrategy, MirroredStrategy, MultiWorkerMirroredStrategy,
    TPUStrategy, experimental_set_strategy, get_strategy, has_strategy, in_cross_replica_context,
    get_replica_context, DistributedDataset, PerReplica,
)


class experimental:  # namespace parity: tf.distribute.experimental.*
    MultiWorkerMirroredStrategy = MultiWorkerMirroredStrategy
    TPUStrategy = TPUStrategy
