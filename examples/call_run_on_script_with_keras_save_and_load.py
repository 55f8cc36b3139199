"""Checkpoint / save / load across strategies as a job, with arguments forwarded to the script
(reference: core/tests/examples/call_run_on_script_with_keras_save_and_load.py):

    python call_run_on_script_with_keras_save_and_load.py --path /tmp/saved_model
"""
import argparse
import os

import cloud_b200 as tfc

parser = argparse.ArgumentParser(description="Model save path arguments.")
parser.add_argument("--path", required=True, type=str, help="Keras model save path")
args = parser.parse_args()

tfc.run(entry_point="save_and_load.py",
        distribution_strategy=None,
        chief_config=tfc.COMMON_MACHINE_CONFIGS[os.environ.get("TFC_EXAMPLE_MACHINE", "B200_2X")],
        requirements_txt="requirements.txt",
        entry_point_args=["--path", args.path],
        stream_logs=True)
