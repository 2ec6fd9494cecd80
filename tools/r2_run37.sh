#!/bin/bash
# wf_extend occupancy re-swept under two wave sequences
mkdir -p gpurun_out/r2
WL="simple:256 mesh_hdri:64 volumes:64 bvh_stress:16"
{
timeout 600 python tools/ab.py "default_n8_v8" "thr12:RTR_WF_REFILL_THRESHOLD=12" "thr20:RTR_WF_REFILL_THRESHOLD=20" -- $WL 2>&1 | grep Mrays
for v in n9 n7 n6 v7 v6; do
  RTR_LIBRARY=$PWD/variants/librtr_$v.so timeout 600 python tools/ab.py "$v" -- $WL 2>&1 | grep Mrays
done
} > gpurun_out/r2/ab_extend_occupancy_two_sequences.txt 2>&1
python - <<'PY'
rows={}; order=[]
for l in open('gpurun_out/r2/ab_extend_occupancy_two_sequences.txt'):
    p=l.split()
    if 'Mrays/s' not in p: continue
    v,w=p[0],p[1]; mr=float(p[p.index('Mrays/s')-1])
    if v not in rows: rows[v]={}; order.append(v)
    rows[v][w]=mr
wl=list(rows[order[0]])
print(f"{'':16s}"+"".join(f"{w:>12s}" for w in wl))
for v in order: print(f"{v:16s}"+"".join(f"{rows[v].get(w,0):12.0f}" for w in wl))
PY
