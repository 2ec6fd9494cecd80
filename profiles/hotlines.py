"""Top source lines of a kernel by executed warp-instructions, from an .ncu-rep captured with
--import-source on (kernels built with -lineinfo).  usage: hotlines.py rep.ncu-rep [kernel-substring] [N]"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    kernel, fname, hdr = None, None, None
    per_kernel = {}
    cur_line = None
    for r in csv.reader(raw.splitlines()):
        if not r:
            continue
        if r[0] == "Function Name":
            kernel = r[1]; per_kernel.setdefault(kernel, {}); continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]; continue
        if r[0] == "Line No":
            hdr = r; continue
        if hdr is None or len(r) != len(hdr):
            continue
        if r[0] != "":
            cur_line = (fname, int(r[0]), r[1].strip())
            continue
        if r[2] in ("...", "") or cur_line is None:
            continue
        ie = hdr.index("Instructions Executed")
        te = hdr.index("Thread Instructions Executed")
        sm = hdr.index("# Samples")
        try:
            d = per_kernel[kernel].setdefault(cur_line, [0.0, 0.0, 0.0])
            d[0] += float(r[ie]); d[1] += float(r[te]); d[2] += float(r[sm])
        except ValueError:
            pass
    for k, lines in per_kernel.items():
        if want not in k or not lines:
            continue
        tot = sum(v[0] for v in lines.values()); tots = sum(v[2] for v in lines.values())
        print(f"== {k[:100]}  warp-inst {tot:.3e}  samples {tots:.0f}")
        for (f, ln, src), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
            thr = v[1] / v[0] if v[0] else 0
            print(f"{100 * v[0] / tot:6.2f}% inst {100 * v[2] / max(tots, 1):6.2f}% stall-samples  thr/inst {thr:5.1f}  {f}:{ln}  {src[:90]}")


if __name__ == "__main__":
    main()
