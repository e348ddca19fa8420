#!/bin/bash
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"shadow_hand_post" -s 4 -c 1 -f -o gpurun_out/prof_hand16k_post python tools/prof.py --task ShadowHand --num-envs 16384 --steps 8 --replicas 4 > gpurun_out/ncu_a.log 2>&1; echo rc=$?
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"shadow_hand_reset" -s 4 -c 1 -f -o gpurun_out/prof_hand16k_reset python tools/prof.py --task ShadowHand --num-envs 16384 --steps 8 --replicas 4 --hot > gpurun_out/ncu_b.log 2>&1; echo rc=$?
