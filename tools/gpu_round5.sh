#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "franka or pipeline or mapped or shadow" > gpurun_out/pytest_gpu5.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu5.log
timeout 300 python bench.py --no-sweep --no-cpu --steps 512 --warmup 32 > gpurun_out/bench_e2e.json 2> gpurun_out/bench_e2e.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_e2e.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_e2e.json").read().strip().splitlines()[-1])
print("step_us", d["ms_per_step"]*1e3, {k:(round(v["ms_per_step"]*1e3,1), round(v["value"]/1e6,1)) for k,v in d["e2e"]["modes"].items()})
PY
for spec in "ShadowHand 16384" "FrankaReachMA 8192" "FrankaCollectMA 8192" "FrankaCombineMA 8192" "FrankaPPMA 8192" "FrankaPPMA 262144"; do set -- $spec
  timeout 300 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu --no-e2e | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['config']['task'], d['config']['num_envs_per_gpu'], 'step_us', round(d['ms_per_step']*1e3,2), 'frac', round(r['frac'],3))"
done
for spec in "ShadowHand 16384 hand16k" "FrankaPPMA 8192 pp8k" "FrankaReachMA 8192 reach8k"; do set -- $spec
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$3.csv python tools/prof.py --task $1 --num-envs $2 --steps 8 --replicas 4 > gpurun_out/ncu_l_$3.log 2>&1
  python - "$3" <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(f"gpurun_out/launches_{sys.argv[1]}.csv")) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
out=[(r[ki][:70], r[vi]) for r in rows[1:] if "ige" in r[ki] or "shadow" in r[ki] or "franka" in r[ki] or "locomotion" in r[ki]]
print(sys.argv[1], out[-6:])
PY
done
