#!/bin/bash
# DRAM traffic + time of one SGEMM n^3 call per (K-chunk depth, tile-group rows); usage: traffic_sweep.sh N "kc:rows kc:rows ..."
n=$1; shift
for spec in $*; do
  kc=${spec%%:*}; rows=${spec##*:}
  GUDA_PIPE_KC=$kc GUDA_PIPE_GROUP_ROWS=$rows ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/traffic_${n}_${kc}_${rows}.csv python tools/prof_headline.py $n > /dev/null 2>&1
  python - <<PY
import csv,collections
rows=list(csv.reader(open("gpurun_out/traffic_${n}_${kc}_${rows}.csv")))
hdr=None; agg=collections.Counter()
for r in rows:
    if "Kernel Name" in r: hdr=r; continue
    if hdr and len(r)==len(hdr):
        rec=dict(zip(hdr,r))
        if "sgemm_p" in rec["Kernel Name"]: agg[rec["Metric Name"]]+=float(rec["Metric Value"].replace(",",""))
print("n=$n kc=$kc group_rows=$rows: time %.1f ms  dram read %.1f GB  write %.1f GB" % (agg["gpu__time_duration.sum"]/1e6, agg["dram__bytes_read.sum"]/1e9, agg["dram__bytes_write.sum"]/1e9))
PY
done
