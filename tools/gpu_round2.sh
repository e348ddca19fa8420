#!/bin/bash
# e2e mapped-mode probe, cartpole check, floors, properly filtered ncu captures
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "host_mapped or cartpole" > gpurun_out/pytest_mapped.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_mapped.log
timeout 300 python bench.py --no-sweep --no-cpu --steps 512 --warmup 32 > gpurun_out/bench_e2e_tma.json 2> gpurun_out/bench_e2e_tma.err; echo "bench tma rc=$?"
IGE_NO_TMA=1 timeout 300 python bench.py --no-sweep --no-cpu --steps 512 --warmup 32 > gpurun_out/bench_e2e_notma.json 2> gpurun_out/bench_e2e_notma.err; echo "bench notma rc=$?"
python - <<'PY'
import json
for f in ("tma","notma"):
    try:
        d=json.loads(open(f"gpurun_out/bench_e2e_{f}.json").read().strip().splitlines()[-1])
        print(f, "step_us", d["ms_per_step"]*1e3, {k:(round(v["ms_per_step"]*1e3,1), round(v["value"]/1e6,1)) for k,v in d["e2e"]["modes"].items()})
    except Exception as e: print(f, "failed", e)
PY
for spec in "Cartpole 16384" "Cartpole 4194304" "Humanoid 32768"; do set -- $spec
  timeout 300 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['config']['task'], d['config']['num_envs_per_gpu'], 'step_us', round(d['ms_per_step']*1e3,2), 'frac', round(r['frac'],3), 'e2e', {k:round(v['ms_per_step']*1e3,1) for k,v in d['e2e']['modes'].items()})"
done
timeout 200 python tools/floors.py > gpurun_out/floors.txt 2>&1; cat gpurun_out/floors.txt
for spec in "Ant 16384 ant16k" "Ant 1048576 ant1m" "Humanoid 32768 hum32k" "ShadowHand 16384 hand16k" "FrankaPPMA 8192 pp8k"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"locomotion|shadow_hand_post|franka_ma" -s 2 -c 1 -f -o gpurun_out/prof_$3 \
    python tools/prof.py --task $1 --num-envs $2 --steps 4 --replicas 4 > gpurun_out/ncu_$3.log 2>&1
  echo "ncu $3 rc=$?"
done
