#!/bin/bash
# quick kernel-only lines for the locomotion tasks: tools/bench_loco.sh [extra bench flags]
EXTRA=("$@")
for s in "Humanoid 32768" "Humanoid 1048576" "Ant 16384" "Ant 1048576"; do set -- $s
  timeout 300 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu --no-e2e "${EXTRA[@]}" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['config']['task'], d['config']['num_envs_per_gpu'], 'step_us', round(d['ms_per_step']*1e3,2), 'kernel_us', round(r['kernel_us'],2), 'frac', round(r['frac'],3))"
done
