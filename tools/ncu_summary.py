"""Turn an `ncu --set full` report (read with `ncu -i X.ncu-rep --page raw --csv`) into the short
per-kernel table kept under profiles/: duration, achieved DRAM bandwidth against the measured copy
bandwidth (MEASURED_PEAKS.json), tensor-pipe utilisation, occupancy and the top warp-stall reasons."""
import argparse
import csv
import json
import os


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("raw_csv")
    ap.add_argument("--peaks", default=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                    "MEASURED_PEAKS.json"))
    args = ap.parse_args()
    peaks = json.load(open(args.peaks)) if os.path.exists(args.peaks) else {}
    hbm = float(peaks.get("hbm_gbs", 6555.0))
    rows = list(csv.reader(open(args.raw_csv)))
    hdr, units, data = rows[0], rows[1], rows[2:]

    def col(name):
        return hdr.index(name) if name in hdr else None

    def num(r, name):
        i = col(name)
        if i is None or r[i] in ("", "n/a"):
            return float("nan")
        v = float(r[i].replace(",", ""))
        u = units[i]
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-6, "ns": 1e-9, "ms": 1e-3,
                 "s": 1.0}.get(u.split("/")[0], 1.0)
        return v * scale

    stall_cols = [(i, h.split("issue_stalled_")[1].split("_per")[0]) for i, h in enumerate(hdr)
                  if "smsp__average_warps_issue_stalled" in h and h.endswith("_per_issue_active.ratio")
                  and "not_issued" not in h]
    print(f"roofline denominators: HBM copy {hbm:.0f} GB/s, bf16 {peaks.get('bf16_tflops', 'n/a')} TFLOP/s "
          f"(MEASURED_PEAKS.json)\n")
    print("| kernel | grid | regs | time us | DRAM GB | DRAM TB/s | % of copy BW | tensor pipe % | "
          "warps active % | top stalls (per issue) |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for r in data:
        name = r[col("Kernel Name")].replace("void ", "").replace("cb2::", "")
        name = name.split("(")[0][:44]
        t = num(r, "gpu__time_duration.sum")
        rd, wr = num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
        tot = rd + wr
        bw = tot / t / 1e12 if t > 0 else float("nan")
        tens = num(r, "sm__inst_executed_pipe_tensor_op_hmma.avg.pct_of_peak_sustained_active")
        if tens != tens:
            tens = num(r, "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active")
        if tens != tens:
            tens = num(r, "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active")
        occ = num(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
        st = sorted(((float(r[i].replace(",", "")), n) for i, n in stall_cols if r[i] not in ("", "n/a")),
                    reverse=True)[:3]
        print(f"| `{name}` | {r[col('launch__grid_size')]} | {r[col('launch__registers_per_thread')]} | "
              f"{t * 1e6:.1f} | {tot / 1e9:.3f} | {bw:.2f} | {100 * bw * 1e3 / hbm:.0f} | "
              f"{'' if tens != tens else f'{tens:.0f}'} | {occ:.0f} | "
              + ", ".join(f"{n} {v:.1f}" for v, n in st) + " |")


if __name__ == "__main__":
    main()
