#!/usr/bin/env python
"""ncu reports of ONE render per workload (tools/profile_one.py under ncu --profile-from-start off) -> profiles/counters.json.
usage: python profiles/make_counters.py gpurun_out/r2/ncu   (directory holding NAME.ncu-rep + NAME.json)
Per workload: time-weighted smsp__issue_active / l1tex__data_pipe_lsu_wavefronts / gpu__dram_throughput (pct of peak),
instruction-weighted lanes per instruction, sum(dram bytes read + written) / rays, and the per-kernel time shares.
bench.py reads the file: roofline.bound = the busiest of the three units, roofline.traffic = dram_bytes_per_ray x rays."""
import csv
import json
import os
import subprocess
import sys

M = {"t": "gpu__time_duration.sum", "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "l1": "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
     "dram": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio", "inst": "smsp__inst_executed.sum",
     "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0,
        "nsecond": 1e-6, "second": 1e3, "s": 1e3}


def rows_of(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {k: hdr.index(v) for k, v in M.items()}
    ki = hdr.index("Kernel Name")
    out = []
    for r in data:
        num = lambda k: float(r[col[k]].replace(",", "")) * UNIT.get(units[col[k]], 1.0)
        out.append(dict(kernel=r[ki].split("(")[0].replace("void rtr::", "").replace("void ", ""), **{k: num(k) for k in M}))
    return out


def main():
    d = sys.argv[1]
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "counters.json")
    out = {"_how": __doc__.split("usage")[0].strip() + "  Captured with ncu --clock-control none; times are serialised and cold-cache, shares are what matters."}
    for f in sorted(os.listdir(d)):
        if not f.endswith(".ncu-rep"):
            continue
        name = f[:-8]
        meta = json.loads(open(os.path.join(d, name + ".json")).read().strip().splitlines()[-1])
        rows = rows_of(os.path.join(d, f))
        T = sum(r["t"] for r in rows)
        I = sum(r["inst"] for r in rows)
        shares = {}
        for r in rows:
            shares[r["kernel"]] = shares.get(r["kernel"], 0.0) + r["t"] / T
        out[name] = {
            "issue_active_pct": sum(r["issue"] * r["t"] for r in rows) / T,
            "l1_lsu_wavefront_pct": sum(r["l1"] * r["t"] for r in rows) / T,
            "dram_pct": sum(r["dram"] * r["t"] for r in rows) / T,
            "lanes_per_inst": sum(r["lanes"] * r["inst"] for r in rows) / I,
            "dram_bytes_per_ray": sum(r["rd"] + r["wr"] for r in rows) / meta["rays"],
            "launches": len(rows), "rays": meta["rays"], "spp": meta["spp"],
            "time_share": {k: round(v, 4) for k, v in sorted(shares.items(), key=lambda kv: -kv[1])},
            "source": f"profiles/r2/ncu_{name}.csv (ncu of tools/profile_one.py {name} {meta['spp']}: every launch of one render)",
        }
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
