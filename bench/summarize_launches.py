"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time per kernel name for
the LAST of `--steps` equal slices of the capture (i.e. one steady-state training step)."""
import argparse
import collections
import csv
import re


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--top", type=int, default=40)
    args = ap.parse_args()
    rows = []
    with open(args.csv, newline="") as f:
        lines = [l for l in f if not l.startswith("==")]
    reader = csv.DictReader(lines)
    for r in reader:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
        rows.append((r["Kernel Name"], val * scale))
    # one step = from the last `optim_tick` launch (the first kernel of a step) to the end; without
    # that marker, the last of `--steps` equal slices
    ticks = [i for i, (name, _us) in enumerate(rows) if "optim_tick" in name]
    if len(ticks) >= 2:
        last = rows[ticks[-1]:]
    else:
        n = len(rows) // args.steps
        last = rows[-n:] if n else rows
    agg = collections.defaultdict(lambda: [0.0, 0])
    for name, us in last:
        short = re.sub(r"<.*", "", name)[:70]
        agg[short][0] += us
        agg[short][1] += 1
    total = sum(v[0] for v in agg.values())
    print(f"kernels in one step: {len(last)}   sum of kernel time: {total / 1e3:.3f} ms")
    for name, (us, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:args.top]:
        print(f"{us / 1e3:9.3f} ms {100 * us / total:5.1f}%  x{cnt:<4d} {name}")


if __name__ == "__main__":
    main()
