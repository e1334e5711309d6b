#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread --clock-control none -k regex:'k_sort|k_gather|k_morton|k_centroid|k_emit_hier|k_refit|k_collapse_level' --launch-count 60 --csv --log-file gpurun_out/af_build_launches.csv python tools/build_sweep.py --sizes 10 --repeat 1 > gpurun_out/af.log 2>&1
python - <<'PY'
import csv,collections
rows=list(csv.reader(open("gpurun_out/af_build_launches.csv")))
i0=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
h=rows[i0]; K=h.index("Kernel Name"); M=h.index("Metric Name"); V=h.index("Metric Value"); I=h.index("ID")
d=collections.OrderedDict()
for r in rows[i0+1:]:
    if len(r)<=V: continue
    d.setdefault((int(r[I]),r[K][:40]),{})[r[M]]=r[V]
for (i,k),m in d.items():
    print(i,k,m.get("gpu__time_duration.sum"),m.get("dram__bytes_read.sum"),m.get("dram__bytes_write.sum"),m.get("launch__registers_per_thread"))
PY
