#!/bin/bash
for c in 3c9c1b4 5c145bc 4baecbe 51ec893 d45c123 HEAD; do
  if [ $c = HEAD ]; then d=.; else d=_bisect/$c; fi
  for s in "Ant 16384" "Ant 65536" "Ant 262144" "Ant 1048576" "Humanoid 32768"; do set -- $s
    (cd $d && timeout 120 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$c', '$1', $2, 'step_us', round(d['ms_per_step']*1e3,2), 'kernel_us', round(r['kernel_us'],2), 'frac', round(r['frac'],3))")
  done
done | tee gpurun_out/r2_bisect.txt
