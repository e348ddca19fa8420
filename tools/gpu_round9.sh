#!/bin/bash
mkdir -p gpurun_out
for spec in "ShadowHand 16384 hand16k" "FrankaPPMA 8192 pp8k" "Cartpole 16384 cart16k"; do set -- $spec
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 64 --csv --log-file gpurun_out/launches_b_$3.csv python bench.py --task $1 --num-envs $2 --steps 32 --warmup 16 --no-sweep --no-cpu --no-e2e > gpurun_out/ncu_b_$3.log 2>&1
python - "$3" <<'PY'
import csv,sys,collections
rows=[r for r in csv.reader(open(f"gpurun_out/launches_b_{sys.argv[1]}.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
d=collections.defaultdict(list)
for r in rows[1:]: d[r[ki][:60]].append(float(r[vi].replace(",","")))
for k,v in d.items(): print(sys.argv[1], k, len(v), "mean_us", sum(v)/len(v)/1e3, "min", min(v)/1e3)
PY
done
