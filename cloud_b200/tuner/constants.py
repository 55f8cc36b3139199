"""Constants of the tuner package (reference: src/python/tensorflow_cloud/tuner/constants.py:17-31)."""
import os

# Service description of the local Vizier-compatible optimizer (the reference ships Google's
# discovery document here: tuner/api/ml_public_google_rest_v1.json)
OPTIMIZER_API_DOCUMENT_FILE = os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                           "api", "local_optimizer_v1.json")

# By default the optimizer returns 1 suggestion per request (constants.py:27).
SUGGESTION_COUNT_PER_REQUEST = 1

# Number of tries to load an existing study after a create() collision (constants.py:31).
NUM_TRIES_FOR_STUDIES = 3

# Hard cap of trials in one study; beyond it `suggest` answers 429 (optimizer_client.py:110-118).
MAX_TRIALS_PER_STUDY = 1000
