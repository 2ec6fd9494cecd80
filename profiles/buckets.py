"""Warp-instruction share and lane occupancy of the render kernel's phases (node step / leaf step / scheduling /
shade+regenerate), from an .ncu-rep captured with --import-source on.  usage: buckets.py rep.ncu-rep [kernel-substring]"""
import csv
import re
import subprocess
import sys


def line_buckets():
    """(file, line) -> bucket, derived from marker comments / function starts in the sources."""
    import os
    root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "raytracingrust_b200", "csrc")
    out = {}
    for fn in ("device_core.cuh", "kernels.cu", "scan.cu", "wavefront.cu"):  # marker comments [bucket:NAME]
        path = os.path.join(root, fn)
        if not os.path.exists(path):
            continue
        cur = "other"
        for i, l in enumerate(open(path), 1):
            m = re.search(r"\[bucket:(\w+)\]", l)
            if m:
                cur = m.group(1)
            out[(fn, i)] = cur
    return out


def main():
    rep = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    lb = line_buckets()
    kernel, fname, hdr, cur_line = None, None, None, None
    acc = {}
    for r in csv.reader(raw.splitlines()):
        if not r:
            continue
        if r[0] == "Function Name":
            kernel = r[1]; acc.setdefault(kernel, {}); continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]; continue
        if r[0] == "Line No":
            hdr = r; continue
        if hdr is None or len(r) != len(hdr):
            continue
        if r[0] != "":
            cur_line = (fname, int(r[0])); continue
        if r[2] in ("...", "") or cur_line is None:
            continue
        ie, te, sm = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
        try:
            b = lb.get(cur_line, "other:" + cur_line[0])
            d = acc[kernel].setdefault(b, [0.0, 0.0, 0.0])
            d[0] += float(r[ie]); d[1] += float(r[te]); d[2] += float(r[sm])
        except ValueError:
            pass
    for k, bs in acc.items():
        if want not in k or not bs:
            continue
        tot = sum(v[0] for v in bs.values()); tots = sum(v[2] for v in bs.values())
        print(f"== {k[:100]}  warp-inst {tot:.3e}")
        for b, v in sorted(bs.items(), key=lambda kv: -kv[1][0]):
            print(f"{b:28s} {100 * v[0] / tot:6.2f}% warp-inst {100 * v[2] / max(tots, 1):6.2f}% stall-samples  lanes/inst {v[1] / max(v[0], 1):5.1f}")


if __name__ == "__main__":
    main()
