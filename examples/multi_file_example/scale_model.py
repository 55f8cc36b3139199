"""Submit train_model.py - and with it the whole directory, create_model.py included - as a job
(reference: core/tests/examples/multi_file_example/scale_model.py).  Run from this directory."""
import os

import cloud_b200 as tfc

tfc.run(entry_point="train_model.py",
        requirements_txt="requirements.txt",
        chief_config=tfc.COMMON_MACHINE_CONFIGS[os.environ.get("TFC_EXAMPLE_MACHINE", "B200_1X")],
        stream_logs=True)
