#!/bin/bash
# Per-task bench lines (BASELINE.json configs[1..3] + the other tasks) -> gpurun_out/bench_tasks.jsonl
out=gpurun_out/${ROUND:-r2}_bench_tasks.jsonl; : > $out
for spec in "Ant 16384" "Humanoid 32768" "ShadowHand 16384" "FrankaReachMA 8192" "FrankaCollectMA 8192" "FrankaCombineMA 8192" "FrankaPPMA 8192" "Cartpole 16384" "Humanoid 1048576" "ShadowHand 262144" "FrankaPPMA 262144" "Cartpole 4194304"; do
  set -- $spec
  timeout 300 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu ${E2E:---no-e2e} >> $out 2>gpurun_out/bench_$1_$2.err || echo "{\"task\": \"$1\", \"failed\": true}" >> $out
done
# SURVEY 8f row 2 tasks
for spec in "Anymal 16384" "AllegroHand 16384" "BallBalance 16384" "Quadcopter 16384" "Ingenuity 16384" "FrankaReach 16384" "FrankaCubeStack2 16384" "FrankaCubeStack 16384" "FrankaCabinet 16384" "Anymal 1048576" "FrankaCubeStack2 1048576"; do
  set -- $spec
  timeout 300 python bench.py --task $1 --num-envs $2 --steps 512 --warmup 32 --no-sweep --no-cpu --no-e2e >> $out 2>gpurun_out/bench_$1_$2.err || echo "{\"task\": \"$1\", \"failed\": true}" >> $out
done
python - <<'PY'
import json
for l in open("gpurun_out/" + __import__("os").environ.get("ROUND", "r2") + "_bench_tasks.jsonl"):
    d = json.loads(l)
    if d.get('failed'): print(d); continue
    c = d['config']; r = d['roofline']
    print(f"{c['task']:16s} N={c['num_envs_per_gpu']:8d} step {d['ms_per_step']*1e3:8.2f} us  {d['value']/1e9:7.3f} G env-steps/s  kernel {r['kernel_us']:8.2f} us  {r['achieved']:7.0f} GB/s  frac {r['frac']:.3f}  e2e {d.get('e2e', {}).get('value', 0)/1e6:7.1f} M/s")
PY
