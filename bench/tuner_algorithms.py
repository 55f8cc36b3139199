"""Search quality of the local optimizer service's algorithms on a synthetic objective with a known
optimum (0): best value found after N trials, averaged over seeds, for RANDOM_SEARCH and the
Gaussian-process bandit (sequential and with 8 workers asking in parallel).  CPU only.

    python bench/tuner_algorithms.py --out profiles/r1_tuner_algorithms.json
"""
import argparse
import json
import math
import os
import statistics
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cloud_b200.tuner import gp_bandit, local_vizier  # noqa: E402

PARENT = "projects/bench/locations/local"
SPACE = [
    {"parameter": "lr", "type": "DOUBLE", "double_value_spec": {"min_value": 1e-5, "max_value": 1e-1},
     "scale_type": "UNIT_LOG_SCALE"},
    {"parameter": "units", "type": "INTEGER", "integer_value_spec": {"min_value": 8, "max_value": 128}},
    {"parameter": "dropout", "type": "DOUBLE", "double_value_spec": {"min_value": 0.0, "max_value": 0.9}},
    {"parameter": "act", "type": "CATEGORICAL", "categorical_value_spec": {"values": ["relu", "tanh", "elu"]}},
    {"parameter": "depth", "type": "DISCRETE", "discrete_value_spec": {"values": [1, 2, 4, 8]}},
]


def objective(parameters):
    v = {p["parameter"]: gp_bandit.param_value(p) for p in parameters}
    return -(math.log10(v["lr"]) + 3) ** 2 - ((v["units"] - 96) / 40) ** 2 - 4 * (v["dropout"] - 0.3) ** 2 \
        - {"relu": 0.0, "tanh": 0.5, "elu": 0.2}[v["act"]] - 0.05 * abs(math.log2(v["depth"]) - 1)


def search(svc, algo, seed, trials, parallel):
    sid = "{}_{}_{}".format(algo, seed, parallel)
    svc.create_study(PARENT, sid, {"study_config": {"algorithm": algo, "parameters": SPACE,
                                                    "metrics": [{"metric": "score", "goal": "MAXIMIZE"}]}})
    name = "{}/studies/{}".format(PARENT, sid)
    path = os.path.join(svc._study_dir(name), "study.json")
    study = json.load(open(path))
    study["seed"] = seed
    json.dump(study, open(path, "w"))
    best, curve, done = -float("inf"), [], 0
    while done < trials:
        batch = []
        for w in range(min(parallel, trials - done)):
            op = svc.suggest_trials(name, {"client_id": "w%d" % w, "suggestion_count": 1})
            batch.append(svc.get_operation(op["name"])["response"]["trials"][0])
        for t in batch:
            value = objective(t["parameters"])
            svc.add_measurement(t["name"], {"measurement": {"stepCount": 1, "metrics": [{"metric": "score", "value": value}]}})
            svc.complete_trial(t["name"], {})
            best = max(best, value)
            curve.append(best)
            done += 1
    return curve


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trials", type=int, default=40)
    ap.add_argument("--seeds", type=int, default=8)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    svc = local_vizier.LocalVizier(tempfile.mkdtemp(prefix="tuner_bench_"))
    rows = []
    for algo, parallel in (("RANDOM_SEARCH", 1), ("GAUSSIAN_PROCESS_BANDIT", 1), ("GAUSSIAN_PROCESS_BANDIT", 8)):
        t0 = time.time()
        curves = [search(svc, algo, seed, args.trials, parallel) for seed in range(args.seeds)]
        row = {"algorithm": algo, "parallel_workers": parallel, "trials": args.trials, "seeds": args.seeds,
               "seconds_per_suggestion": (time.time() - t0) / (args.trials * args.seeds)}
        for n in (10, 20, args.trials):
            row["best_after_%d" % n] = statistics.mean(c[n - 1] for c in curves)
        row["worst_seed_final"] = min(c[-1] for c in curves)
        rows.append(row)
        print(json.dumps(row), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"objective_optimum": 0.0, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
