#!/bin/bash
# per-kernel split of the GroupNorm pair (ncu time + DRAM bytes) at a few shapes: tools/gn_split.sh [shapes]
S=${1:-4096x320,1024x640,1024x1280,256x1280,64x1280}
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:gn_ --csv --log-file gpurun_out/gn_ncu.csv python tools/gn_bench.py 38 $S 1 > /dev/null 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/gn_ncu.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); mi=hdr.index("Metric Name"); vi=hdr.index("Metric Value"); ii=hdr.index("ID"); gi=hdr.index("Grid Size")
d=collections.OrderedDict()
for r in rows[1:]:
    d.setdefault(r[ii],{"k":r[ki][:15],"g":r[gi]})[r[mi]]=float(r[vi].replace(",",""))
seg=[]
for i,v in d.items():
    key=(v["k"],v["g"])
    if seg and key in seg[-1] and len(seg[-1])<=2: seg[-1][key].append(v)
    elif seg and len(seg[-1])<2: seg[-1][key]=[v]
    else: seg.append({key:[v]})
for s in seg:
    for (k,g),vs in s.items():
        vs=vs[len(vs)//2:]; n=len(vs)
        print(k,g,"n=%d"%n,"us=%.1f"%(sum(v["gpu__time_duration.sum"] for v in vs)/n/1e3),"rdMB=%.1f"%(sum(v["dram__bytes_read.sum"] for v in vs)/n/1e6),"wrMB=%.1f"%(sum(v["dram__bytes_write.sum"] for v in vs)/n/1e6))
    print()
PY
