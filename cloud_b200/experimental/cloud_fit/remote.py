"""cloud_fit server side (reference: experimental/cloud_fit/remote.py:40-169).

    python -m cloud_b200.experimental.cloud_fit.remote --remote_dir D --distribution_strategy S

Instantiates the strategy by name, and inside its scope loads the training assets and the model,
runs ``model.fit(**fit_kwargs)`` and saves the result: every replica takes part in the save (state
may be sharded across replicas) but only the chief's ``<remote_dir>/output`` is kept
(remote.py:130-156; non-chief workers write nothing instead of writing to a temp dir and deleting).
"""
from __future__ import annotations

import argparse
import json
import logging
import os
import sys

import cloudpickle

from . import utils

logger = logging.getLogger(__name__)


def main(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument("--remote_dir", required=True,
                        help="Temporary directory holding the training assets and the model.")
    parser.add_argument("--distribution_strategy", required=True,
                        help="String representing distribution strategy.")
    flags = parser.parse_args(argv)
    logging.basicConfig(level=logging.INFO)
    run(flags.remote_dir, flags.distribution_strategy)


def run(remote_dir: str, distribution_strategy_text: str):
    """``model.fit`` with the serialised assets under the named strategy."""
    if distribution_strategy_text not in utils.SUPPORTED_DISTRIBUTION_STRATEGIES:
        raise ValueError("{} is not supported. Must be one of {}".format(
            distribution_strategy_text, list(utils.SUPPORTED_DISTRIBUTION_STRATEGIES.keys())))
    logger.info("Setting distribution strategy to %s", distribution_strategy_text)
    strategy = utils.SUPPORTED_DISTRIBUTION_STRATEGIES[distribution_strategy_text]()
    from ... import distribute
    from ...keras import models
    with strategy.scope():
        distribute.experimental_set_strategy(strategy)
        assets_dir = os.path.join(remote_dir, "training_assets")
        with open(os.path.join(assets_dir, "fit_kwargs.pkl"), "rb") as f:
            fit_kwargs = cloudpickle.loads(f.read())
        cb_path = os.path.join(assets_dir, "callbacks.pkl")
        if os.path.exists(cb_path):
            with open(cb_path, "rb") as f:
                fit_kwargs["callbacks"] = cloudpickle.loads(f.read())
        model = models.load_model(os.path.join(remote_dir, "model"))
        model._strategy = strategy
        model.fit(**fit_kwargs)
        # every replica participates (sharded optimizer state is gathered); only the chief writes
        model.save(os.path.join(remote_dir, "output"))
    strategy.barrier()
    return model


def _is_current_worker_chief():
    """TF_CONFIG task 'chief' or index 0 (remote.py:148-156)."""
    cfg = os.environ.get("TF_CONFIG")
    if cfg:
        parsed = json.loads(cfg)
        task = parsed.get("task", {})
        if "chief" in parsed.get("cluster", {}):
            return task.get("type") == "chief"
        return task.get("type") == "chief" or int(task.get("index", 0)) == 0
    return True


if __name__ == "__main__":
    main(sys.argv[1:])
