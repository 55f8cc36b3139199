"""Job management from the shell - the local stand-in for ``gcloud ai-platform jobs ...`` (the
reference shells out to ``gcloud ai-platform jobs stream-logs <job_id>``, core/deploy.py:187-211,
and points users at the cloud console for everything else).

    python -m cloud_b200.launch list
    python -m cloud_b200.launch describe <job_id>
    python -m cloud_b200.launch stream-logs <job_id>
    python -m cloud_b200.launch cancel <job_id>
    python -m cloud_b200.launch topology
    python -m cloud_b200.launch gpus             # which job holds which GPU, who is waiting

Jobs live under ``$TFC_JOB_ROOT/jobs/<job_id>`` (see ``launch/registry.py``).
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import sys
import time

from . import leases, registry, topology


def _fmt_time(t):
    return time.strftime("%Y-%m-%d %H:%M:%S", time.localtime(t)) if t else "-"


def cmd_list(_args) -> int:
    rows = []
    for job_id in registry.list_jobs():
        st = registry.status(job_id)
        spec = registry.get(job_id)
        rows.append((job_id, st.get("state", "?"), spec.get("plan", {}).get("world_size", "?"),
                     "{:.1f}".format(st["seconds"]) if "seconds" in st else "-",
                     _fmt_time(spec.get("created"))))
    print("{:<58} {:<10} {:>5} {:>9}  {}".format("JOB_ID", "STATE", "RANKS", "SECONDS", "CREATED"))
    for r in rows:
        print("{:<58} {:<10} {:>5} {:>9}  {}".format(*r))
    return 0


def cmd_describe(args) -> int:
    spec = registry.get(args.job_id)
    st = registry.status(args.job_id)
    out = {"job_id": args.job_id, "status": st, "request": spec.get("request"), "plan": spec.get("plan"),
           "logs": registry.log_paths(args.job_id)}
    print(json.dumps(out, indent=1, default=str))
    return 0 if st.get("state") != "FAILED" else 1


def cmd_stream_logs(args) -> int:
    from ..core import deploy
    deploy._stream_logs(args.job_id)
    return 0 if registry.status(args.job_id).get("state") == "SUCCEEDED" else 1


def cmd_cancel(args) -> int:
    st = registry.status(args.job_id)
    if st.get("state") in registry.TERMINAL:
        print("Job {} is already {}.".format(args.job_id, st.get("state")))
        return 0
    pid = st.get("supervisor_pid")
    if not pid:
        print("Job {} has no running supervisor.".format(args.job_id), file=sys.stderr)
        return 1
    try:
        os.kill(int(pid), signal.SIGTERM)      # the supervisor kills its ranks and records CANCELLED
    except ProcessLookupError:
        pass
    deadline = time.time() + args.timeout
    while True:
        final = registry.status(args.job_id)
        if final.get("state") in registry.TERMINAL or time.time() > deadline:
            break
        if not leases._alive(pid):
            # the signal caught the supervisor before it could answer it (still starting up, or
            # gone without a trace): nothing is running, so the record is written here
            leases.release(args.job_id)
            registry.set_status(args.job_id, "CANCELLED", exit_codes=[], seconds=0.0)
            continue
        time.sleep(0.1)
    print("Job {}: {}".format(args.job_id, final.get("state")))
    return 0 if final.get("state") == "CANCELLED" else 1


def cmd_topology(_args) -> int:
    topo = topology.discover()
    print(json.dumps(topo.to_dict() if hasattr(topo, "to_dict") else topo.__dict__, indent=1, default=str))
    return 0


def cmd_gpus(_args) -> int:
    total = topology.gpu_count()
    table = leases.snapshot(total)
    print("{:<5} {:<58} {:>8}  {}".format("GPU", "HELD BY", "PID", "SINCE"))
    for g in range(total):
        lease = table["gpus"].get(str(g))
        if lease:
            print("{:<5} {:<58} {:>8}  {}".format(g, lease.get("job", "?"), lease.get("pid", "?"),
                                                  _fmt_time(lease.get("time"))))
        else:
            print("{:<5} {:<58}".format(g, "-"))
    for q in table["queue"]:
        print("waiting: {} wants {} GPU(s) since {}".format(q["owner"], q["n"], _fmt_time(q.get("time"))))
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m cloud_b200.launch", description=__doc__.split("\n\n")[0])
    sub = ap.add_subparsers(dest="cmd", required=True)
    sub.add_parser("list").set_defaults(fn=cmd_list)
    for name, fn in (("describe", cmd_describe), ("stream-logs", cmd_stream_logs)):
        p = sub.add_parser(name)
        p.add_argument("job_id")
        p.set_defaults(fn=fn)
    p = sub.add_parser("cancel")
    p.add_argument("job_id")
    p.add_argument("--timeout", type=float, default=30.0)
    p.set_defaults(fn=cmd_cancel)
    sub.add_parser("topology").set_defaults(fn=cmd_topology)
    sub.add_parser("gpus").set_defaults(fn=cmd_gpus)
    args = ap.parse_args(argv)
    try:
        return args.fn(args)
    except FileNotFoundError:
        print("No such job: {}".format(getattr(args, "job_id", "")), file=sys.stderr)
        return 2


if __name__ == "__main__":
    sys.exit(main())
