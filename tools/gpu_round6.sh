#!/bin/bash
mkdir -p gpurun_out
for spec in "FrankaReachMA 8192 reach8k franka_ma_pair" "ShadowHand 16384 hand16k shadow_hand_post" "Cartpole 16384 cart16k cartpole"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:"$4" -s 4 -c 1 -f -o gpurun_out/prof_$3 \
    python tools/prof.py --task $1 --num-envs $2 --steps 8 --replicas 4 > gpurun_out/ncu_$3.log 2>&1
  echo "ncu $3 rc=$?"
done
