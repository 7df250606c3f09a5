"""Summarise an `ncu --set full` capture of one bench step (ncu -i X.ncu-rep --page raw --csv > raw.csv):
per kernel: duration, DRAM bytes read/written, issue activity, registers; per bench stage: DRAM bytes per launch
(profiles/traffic.json, what bench.py reports as roofline.traffic).

    python tools/ncu_traffic.py raw.csv profiles/rNN_ncu_full_summary.csv profiles/traffic.json [profiles/issue.json workload]

With the last two arguments the mean smsp__inst_executed.sum per launch of every stage kernel is written too (what
bench.py divides by the measured stage time for the issue-bound roofline).
"""
import csv, json, sys

STAGE_OF = {"blend_forward_wide_kernel": "blend_fwd_wide", "blend_backward_wide_kernel": "blend_bwd_wide",
            "lbs_mlp_forward_kernel": "lbs_mlp", "preprocess_kernel": "preprocess", "emit_keys_kernel": "emit_keys", "sort_histogram_kernel": "sort",
            "onesweep_pass_kernel": "sort", "tile_ranges_kernel": "tile_ranges", "tile_order_kernel": "tile_ranges",
            "blend_forward_kernel": "blend_fwd", "blend_backward_kernel": "blend_bwd",
            "preprocess_backward_kernel": "preprocess_bwd", "lbs_deform_forward_kernel": "lbs_fwd",
            "lbs_deform_backward_kernel": "lbs_bwd", "nearest_vertex_smem_kernel": "knn", "lbs_pose_kernel": "lbs_joints",
            "lbs_vertex_offsets_kernel": "lbs_joints", "lbs_pose_views_kernel": "lbs_joints",
            "deform_project_kernel": "deform_project", "project_deform_backward_kernel": "project_deform_bwd"}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
COLS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]


def main(raw, out_csv, out_json, issue_json=None, workload="zju_512"):
    rows = list(csv.reader(open(raw)))
    h, units = rows[0], rows[1]
    kn = h.index("Kernel Name")
    cols = [c for c in COLS if c in h]
    per_stage, seen, inst = {}, {}, {}
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        stall_cols = [c for c in h if c.startswith("smsp__average_warps_issue_stalled_") and "selected" not in c]
        w.writerow(["kernel", "stage"] + [f"{c} [{units[h.index(c)]}]" for c in cols] +
                   ["top stalls (warps stalled per issued instruction)"])
        for r in rows[2:]:
            name = r[kn]
            base = name.split("(")[0].replace("void ", "").split("<")[0].split("::")[-1].strip()
            stage = STAGE_OF.get(base, "")
            stalls = sorted(((float(r[h.index(c)].replace(",", "") or 0), c[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")])
                             for c in stall_cols), reverse=True)[:3]
            w.writerow([name[:70], stage] + [r[h.index(c)] for c in cols] +
                       ["; ".join(f"{n} {v:.2f}" for v, n in stalls)])
            if stage:
                rd = float(r[h.index("dram__bytes_read.sum")].replace(",", "")) * UNIT[units[h.index("dram__bytes_read.sum")]]
                wr = float(r[h.index("dram__bytes_write.sum")].replace(",", "")) * UNIT[units[h.index("dram__bytes_write.sum")]]
                key = (stage, base)
                seen.setdefault(key, []).append(rd + wr)
                if "smsp__inst_executed.sum" in h and base not in ("onesweep_pass_kernel", "sort_histogram_kernel",
                                                                   "tile_order_kernel", "lbs_pose_kernel"):
                    inst.setdefault(stage, []).append(float(r[h.index("smsp__inst_executed.sum")].replace(",", "")))
    # a stage's traffic = sum over its kernels of (mean per launch x launches per step); the sort has 1 histogram + its passes
    launches_per_step = {"onesweep_pass_kernel": None}
    for (stage, base), v in seen.items():
        n = len(v) if base == "onesweep_pass_kernel" else 1
        if base == "onesweep_pass_kernel":
            n = 5  # passes of a 43-45 bit key
        per_stage[stage] = per_stage.get(stage, 0.0) + n * sum(v) / len(v)
    json.dump({k: int(v) for k, v in sorted(per_stage.items())}, open(out_json, "w"), indent=1)
    print(json.dumps(per_stage, indent=1))
    if issue_json:
        try:
            cur = json.load(open(issue_json))
        except Exception:
            cur = {}
        cur[workload] = {st: {"inst_executed": int(sum(v) / len(v)), "launches_averaged": len(v),
                              "source": f"{out_csv} (ncu --set full, smsp__inst_executed.sum, mean over the captured launches)"}
                         for st, v in sorted(inst.items())}
        json.dump(cur, open(issue_json, "w"), indent=1)


if __name__ == "__main__":
    main(*sys.argv[1:6])
