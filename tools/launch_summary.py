"""Summarise an ncu launch list (ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X.csv CMD):
per kernel: launches, total us, share of all profiled GPU time.  The per-launch times are cold-cache and serialised, so
only the SHARE is comparable with bench.py's CUDA-event stage table.

    python tools/launch_summary.py gpurun_out/launches.csv profiles/rNN_launches_summary.csv [steps_profiled | view]

"view": normalise by the number of blend_backward launches = views rendered forward + backward by the process (warm-up,
graph capture and timed steps alike), which is the unit bench.py's stage table uses.
"""
import csv
import sys


def short(name):
    base = name.replace("void ", "").split("(")[0]
    return base if len(base) <= 90 else base[:87] + "..."


def main(src, dst, steps=1):
    rows = [r for r in csv.reader(l for l in open(src, errors="replace") if l.startswith('"'))]
    h = rows[0]
    kn, mn, mu, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value")
    unit = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}
    tot, cnt = {}, {}
    for r in rows[1:]:
        if r[mn] != "gpu__time_duration.sum":
            continue
        k = short(r[kn])
        tot[k] = tot.get(k, 0.0) + float(r[mv].replace(",", "")) * unit[r[mu]]
        cnt[k] = cnt.get(k, 0) + 1
    all_us = sum(tot.values())
    if steps == "view":
        steps = max([c for k, c in cnt.items() if "blend_backward" in k] or [1])
    steps = float(steps)
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "launches_per_unit", "us_per_unit", "share_of_profiled_time", "ours"])
        for k in sorted(tot, key=tot.get, reverse=True):
            w.writerow([k, f"{cnt[k] / steps:.2f}", f"{tot[k] / steps:.2f}", f"{tot[k] / all_us:.3f}", int(k.startswith("ghr::"))])
    ours = sum(v for k, v in tot.items() if k.startswith("ghr::"))
    print(f"{len(rows) - 1} launches, {steps:.0f} units, {all_us / steps:.1f} us per unit profiled, libghr share {ours / all_us:.3f}")


if __name__ == "__main__":
    main(*sys.argv[1:4])
