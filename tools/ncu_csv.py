#!/usr/bin/env python
"""Per-launch CSV of the whole-render ncu captures (tools/r2_run26.sh): profiles/r2/ncu_NAME.csv from DIR/NAME.ncu-rep + NAME.json.
usage: python tools/ncu_csv.py gpurun_out/r2/ncu "comment appended to the header line" """
import csv
import os
import subprocess
import sys

COLS = ["Kernel Name", "gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct"]


def main():
    d, note = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for f in sorted(os.listdir(d)):
        if not f.endswith(".ncu-rep"):
            continue
        name = f[:-8]
        meta = open(os.path.join(d, name + ".json")).read().strip().splitlines()[-1]
        raw = subprocess.run(["ncu", "-i", os.path.join(d, f), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        hdr, units, data = rows[0], rows[1], rows[2:]
        idx = [hdr.index(c) for c in COLS]
        dst = os.path.join(root, "profiles", "r2", f"ncu_{name}.csv")
        with open(dst, "w", newline="") as out:
            w = csv.writer(out)
            w.writerow([f"# ncu --metrics ... --clock-control none --profile-from-start off python tools/profile_one.py {name} (every launch of one render{', ' + note if note else ''}; {meta})"])
            w.writerow(COLS)
            w.writerow([units[i] for i in idx])
            for r in data:
                w.writerow([r[i].split("(")[0].replace("void rtr::", "void ") + ("(WfArgs)" if "wf_" in r[i] and i == idx[0] else "") if i == idx[0] else r[i] for i in idx])
        print(dst, len(data), "launches")


if __name__ == "__main__":
    main()
