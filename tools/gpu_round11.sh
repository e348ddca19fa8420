#!/bin/bash
# ncu evidence for the 8f rows 3/4 kernels: launch list + one full capture each
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ma_rollout|ma_discount" -c 60 --csv --log-file gpurun_out/r1_launches_rollout.csv python tools/bench_rollout.py --rows 16384 > gpurun_out/ncu_l_rollout.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:franka_osc -c 40 --csv --log-file gpurun_out/r1_launches_osc.csv python tools/bench_osc.py --rows 16384 > gpurun_out/ncu_l_osc.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:franka_osc -s 4 -c 1 -o gpurun_out/r1_osc_262k -f python tools/bench_osc.py --rows 262144 > gpurun_out/ncu_f_osc.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ma_rollout_step -s 4 -c 1 -o gpurun_out/r1_rollout_4m -f python tools/bench_rollout.py --rows 4194304 > gpurun_out/ncu_f_rollout.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ma_rollout_step -s 20 -c 1 -o gpurun_out/r1_rollout_16k -f python tools/bench_rollout.py --rows 16384 > gpurun_out/ncu_f_rollout16k.log 2>&1
python tools/ncu_summary.py gpurun_out/r1_osc_262k.ncu-rep gpurun_out/r1_rollout_4m.ncu-rep gpurun_out/r1_rollout_16k.ncu-rep > gpurun_out/r1_ncu_full_v10_rows34.txt 2>&1
tail -60 gpurun_out/r1_ncu_full_v10_rows34.txt
