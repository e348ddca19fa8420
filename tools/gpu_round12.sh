#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
ncu --set full --clock-control none --import-source on -k regex:franka_osc -s 4 -c 1 -o gpurun_out/r1_osc_262k_v2 -f python tools/bench_osc.py --rows 262144 > gpurun_out/ncu_f_osc2.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:franka_osc -c 40 --csv --log-file gpurun_out/r1_launches_osc.csv python tools/bench_osc.py --rows 16384 > gpurun_out/ncu_l_osc.log 2>&1
python tools/ncu_summary.py gpurun_out/r1_osc_262k_v2.ncu-rep > gpurun_out/r1_ncu_full_v11_osc.txt 2>&1
cat gpurun_out/r1_ncu_full_v11_osc.txt
ncu -i gpurun_out/r1_osc_262k_v2.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; r=rows[2]
for i,n in enumerate(h):
    if 'issue_stalled' in n and n.endswith('.pct') or 'issue_stalled' in n and 'ratio' in n:
        try:
            v=float(r[i])
        except: continue
        if v>0.3: print(n, v)
" | sort -k2 -n -r | head -12
