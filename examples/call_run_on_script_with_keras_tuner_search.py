"""A whole hyper-parameter search as one job on a multi-GPU machine (reference:
core/tests/examples/call_run_on_script_with_keras_tuner_search.py):

    python call_run_on_script_with_keras_tuner_search.py --path /tmp/keras-cifar
"""
import argparse
import os

import cloud_b200 as tfc

parser = argparse.ArgumentParser(description="Model save path arguments.")
parser.add_argument("--path", required=True, type=str, help="Keras model save path")
parser.add_argument("--max-trials", type=int, default=None)
parser.add_argument("--samples", type=int, default=None)
args = parser.parse_args()

entry_point_args = ["--path", args.path]
if args.max_trials is not None:
    entry_point_args += ["--max-trials", str(args.max_trials)]
if args.samples is not None:
    entry_point_args += ["--samples", str(args.samples)]

# Automated MirroredStrategy: chief config with multiple GPUs
tfc.run(entry_point="keras_tuner_cifar_example.py",
        distribution_strategy="auto",
        requirements_txt="requirements.txt",
        chief_config=tfc.COMMON_MACHINE_CONFIGS[os.environ.get("TFC_EXAMPLE_MACHINE", "B200_4X")],
        worker_count=0,
        entry_point_args=entry_point_args,
        stream_logs=True)
