#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "shadow_hand or generic" > gpurun_out/pytest_hand.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_hand.log
for spec in "ShadowHand 16384" "ShadowHand 262144"; do set -- $spec
  timeout 300 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu --no-e2e | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print(d['config']['task'], d['config']['num_envs_per_gpu'], 'step_us', round(d['ms_per_step']*1e3,2), 'frac', round(r['frac'],3))"
done
