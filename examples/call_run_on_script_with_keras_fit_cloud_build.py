"""The staged-image path: with ``docker_image_bucket_name`` the job image is built by the
``CloudContainerBuilder`` stand-in (staging area named after the bucket) instead of the local one
(reference: core/tests/examples/call_run_on_script_with_keras_fit_cloud_build.py):

    python call_run_on_script_with_keras_fit_cloud_build.py --bucket_name my-staging-area
"""
import argparse
import os

import cloud_b200 as tfc

parser = argparse.ArgumentParser(description="Model cloud bucket name argument.")
parser.add_argument("--bucket_name", required=True, type=str, help="Cloud bucket name")
args = parser.parse_args()

if "TFC_EXAMPLE_MACHINE" in os.environ:
    chief = tfc.COMMON_MACHINE_CONFIGS[os.environ["TFC_EXAMPLE_MACHINE"]]
else:       # automated MirroredStrategy: chief config with multiple GPUs
    chief = tfc.MachineConfig(cpu_cores=8, memory=30,
                              accelerator_type=tfc.AcceleratorType.NVIDIA_TESLA_T4, accelerator_count=2)
tfc.run(entry_point="mnist_example_using_fit.py",
        distribution_strategy="auto",
        chief_config=chief,
        worker_count=0,
        stream_logs=True,
        docker_image_bucket_name=args.bucket_name)
