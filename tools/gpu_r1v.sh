#!/bin/bash
# GPU call: alternating tile-walk direction (L2 reuse across kernel boundaries) A/B, parity tests with it on
mkdir -p gpurun_out
for rep in 1 2; do
F2F_ALT_DIR=0 timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed "s/^/alt off: /"
F2F_ALT_DIR=1 timeout 300 python tools/perf_dncnn.py 1080 1920 5 2>&1 | grep -v Warn | sed "s/^/alt on : /"
done | tee gpurun_out/altdir_ab.log
timeout 900 python -m pytest tests/test_dncnn_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_conv64t|k_wgrad64' --launch-skip 45 -c 45 --csv --log-file gpurun_out/alt_traffic.csv python tools/perf_dncnn.py 1080 1920 1 > gpurun_out/ncu_a.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/alt_traffic.csv')))
hdr=next(r for r in rows if r and r[0]=='ID'); idx={h:i for i,h in enumerate(hdr)}
agg=collections.OrderedDict()
for r in rows:
    if len(r)==len(hdr) and r[0]!='ID':
        k=r[idx['Kernel Name']].split('(')[0][-30:]; m=r[idx['Metric Name']]
        try: v=float(r[idx['Metric Value']].replace(',',''))
        except ValueError: continue
        a=agg.setdefault((k,m,r[idx['Metric Unit']]),[0,0.0]); a[0]+=1; a[1]+=v
for (k,m,u),(n,v) in agg.items(): print("%-32s %-28s %10.1f %s avg over %d" % (k,m,v/n,u,n))
PY
