#!/bin/bash
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:anymal_post -s 8 -c 1 -o gpurun_out/r1_anymal_1m -f python bench.py --task Anymal --num-envs 1048576 --steps 32 --warmup 4 --no-e2e --no-cpu --no-sweep > gpurun_out/ncu_f_anymal.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:anymal_post -c 48 --csv --log-file gpurun_out/r1_launches_anymal16k.csv python bench.py --task Anymal --num-envs 16384 --steps 32 --warmup 4 --no-e2e --no-cpu --no-sweep > gpurun_out/ncu_l_anymal.log 2>&1
python tools/ncu_summary.py gpurun_out/r1_anymal_1m.ncu-rep > gpurun_out/r1_ncu_full_v12_anymal.txt 2>&1
cat gpurun_out/r1_ncu_full_v12_anymal.txt | head -30
tail -3 gpurun_out/r1_launches_anymal16k.csv
