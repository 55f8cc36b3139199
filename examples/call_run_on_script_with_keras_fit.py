"""Launch a training script on the local B200s with one call (reference: README.md:88-91 and
core/tests/examples/call_run_on_script_with_keras_fit.py)."""
import cloud_b200 as tfc

tfc.run(entry_point="mnist_example_using_fit.py",
        chief_config=tfc.COMMON_MACHINE_CONFIGS["B200_1X"],
        stream_logs=True,
        job_labels={"job": "mnist-example", "team": "keras"})
