"""Model search inside a job (reference: core/tests/examples/call_run_within_script_with_autokeras.py):
`tfc.run()` at the top submits this very script, the rest runs on the box's GPU.  The AutoKeras
`ImageClassifier` of the reference is `cloud_b200.automl.ImageClassifier` here.

    python call_run_within_script_with_autokeras.py --path /tmp/autokeras_out
"""
import argparse
import os

import cloud_b200 as tfc
from cloud_b200 import automl as ak
from cloud_b200.keras.datasets import mnist

parser = argparse.ArgumentParser(description="Model save path arguments.")
parser.add_argument("--path", required=True, type=str, help="Keras model save path")
parser.add_argument("--samples", type=int, default=None, help="use only this many training images")
parser.add_argument("--epochs", type=int, default=10)
args = parser.parse_args()

tfc.run(chief_config=tfc.COMMON_MACHINE_CONFIGS[os.environ.get("TFC_EXAMPLE_MACHINE", "B200_1X")],
        entry_point_args=["--path", args.path, "--epochs", str(args.epochs)]
        + (["--samples", str(args.samples)] if args.samples else []),
        stream_logs=True)

# Prepare the dataset.
(x_train, y_train), (x_test, y_test) = mnist.load_data()
if args.samples:
    x_train, y_train = x_train[:args.samples], y_train[:args.samples]
    x_test, y_test = x_test[:max(64, args.samples // 4)], y_test[:max(64, args.samples // 4)]
print(x_train.shape)  # (60000, 28, 28)
print(y_train.shape)  # (60000,)
print(y_train[:3])

# Initialize the ImageClassifier.
clf = ak.ImageClassifier(max_trials=2, directory=os.path.join(args.path, "search"))
# Search for the best model.
clf.fit(x_train, y_train, epochs=args.epochs)
# Evaluate on the testing data.
print("Accuracy: {accuracy}".format(accuracy=clf.evaluate(x_test, y_test)[1]))

clf.export_model().save(os.path.join(args.path, "model.h5"))
