"""A notebook as the entry point (reference:
core/tests/examples/call_run_on_notebook_with_keras_fit.py): its code cells are converted to a
script, the strategy for the chosen machine is injected in front, and the result runs as a job."""
import os

import cloud_b200 as tfc

# Automated MirroredStrategy: chief config with multiple GPUs
machine = os.environ.get("TFC_EXAMPLE_MACHINE", "B200_2X")
tfc.run(entry_point="mnist_example_using_fit.ipynb",
        distribution_strategy="auto",
        chief_config=tfc.COMMON_MACHINE_CONFIGS[machine],
        worker_count=0,
        stream_logs=True)
