#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q -k anymal 2>&1 | tail -2
for n in 16384 1048576; do
timeout 200 python bench.py --task Anymal --num-envs $n --steps 256 --warmup 16 --no-e2e --no-cpu --no-sweep 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['num_envs_per_gpu'], d['ms_per_step']*1e3, 'us', d['roofline']['frac'])"
done
