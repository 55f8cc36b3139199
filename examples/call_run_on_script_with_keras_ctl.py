"""A custom-training-loop script launched with a user-managed strategy (reference:
core/tests/examples/call_run_on_script_with_keras_ctl.py).  ``distribution_strategy=None`` skips the
strategy injection: ``mnist_example_using_ctl.py`` builds its own MultiWorkerMirroredStrategy from
the ``TF_CONFIG`` the launcher writes for the chief and the one worker."""
import os

import cloud_b200 as tfc

machine = tfc.COMMON_MACHINE_CONFIGS[os.environ.get("TFC_EXAMPLE_MACHINE", "B200_1X")]
tfc.run(entry_point="mnist_example_using_ctl.py",
        distribution_strategy=None,
        chief_config=machine,
        worker_config=machine,
        worker_count=1,
        requirements_txt="requirements.txt",
        stream_logs=True)
