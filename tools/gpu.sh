#!/bin/bash
# One parametrised GPU-box script (replaces the per-lease gpu_round*.sh of round 1).
#   bash tools/gpu.sh <action> [<action> ...]        actions run in order; logs go to gpurun_out/
# actions:
#   tests            pytest -m gpu (whole suite)            tests:<expr>   pytest -m gpu -k <expr>
#   smoke            __graft_entry__.smoke()
#   bench            default bench.py (driver's N=1 line)   ref            bench.py --impl reference
#   driver           bench.py --steps 20 --warmup 5 (the driver's own command, N=1)
#   scale:<N>        torchrun bench.py --gpus N --steps 20 --warmup 5 (+ a --steps 2048 line)
#   tasks            per-task bench lines (tools/bench_all.sh)
#   launches         ncu launch list of the default bench's timed region
#   ncu:<Task>:<N>:<tag>   ncu --set full of one launch (tools/prof.py), summary into gpurun_out/
#   sanitize         compute-sanitizer memcheck + racecheck over a reduced -m gpu selection
#   floors           tools/floors.py (empty-launch / copy floors)
mkdir -p gpurun_out
R=${ROUND:-r2}
for act in "$@"; do
  IFS=: read -r a b c d <<< "$act"
  case "$a" in
    tests)
      if [ -n "$b" ]; then timeout 600 python -m pytest tests -m gpu -x -q -k "$b" > gpurun_out/pytest_gpu_k.log 2>&1; echo "pytest -k '$b' rc=$?"; tail -5 gpurun_out/pytest_gpu_k.log
      else timeout 1500 python -m pytest tests -m gpu --maxfail=25 -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log; fi ;;
    smoke) timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log ;;
    bench) timeout 600 python bench.py > gpurun_out/${R}_bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_default.err; cut -c1-1500 gpurun_out/${R}_bench_default.json ;;
    driver) timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${R}_bench_driver_n1.json 2> gpurun_out/bench_driver.err; echo "driver-bench rc=$?"; tail -2 gpurun_out/bench_driver.err; cut -c1-900 gpurun_out/${R}_bench_driver_n1.json ;;
    ref) timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${R}_bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"; tail -2 gpurun_out/bench_reference.err; cut -c1-1200 gpurun_out/${R}_bench_reference.json ;;
    scale)
      for steps in 20 2048; do
        timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $b --master-addr 127.0.0.1 --master-port 29533 \
          bench.py --gpus $b --steps $steps --warmup 5 --no-sweep > gpurun_out/${R}_bench_n${b}_s${steps}.json 2> gpurun_out/bench_n${b}_s${steps}.err
        echo "scale N=$b steps=$steps rc=$?"; tail -3 gpurun_out/bench_n${b}_s${steps}.err | cut -c1-300; cut -c1-700 gpurun_out/${R}_bench_n${b}_s${steps}.json
      done ;;
    scalemodes)   # scalemodes:<N>  -- the driver's 20-step window with each statistics-reduction mode (+ none = no branch)
      for mode in peer nccl none; do
        for rep in 1 2; do
        timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $b --master-addr 127.0.0.1 --master-port 29544 \
          bench.py --gpus $b --steps 20 --warmup 5 --no-sweep --no-e2e --no-cpu --stats-reduce $mode > gpurun_out/${R}_bench_n${b}_s20_${mode}_$rep.json 2> gpurun_out/bench_n${b}_${mode}.err
        echo "scale N=$b mode=$mode rep=$rep rc=$?"; tail -2 gpurun_out/bench_n${b}_${mode}.err | cut -c1-300
        python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${R}_bench_n${b}_s20_${mode}_$rep.json").read().strip().splitlines()[-1])
    print(f"   ms_per_step {d['ms_per_step']*1e3:.3f} us  value {d['value']/1e9:.3f} G  kernel_us {d['roofline']['kernel_us']:.2f}  reduce: {d['how']['stats_reduce'][:60]}")
except Exception as e:
    print("   failed:", e)
PY
        done
      done ;;
    tasks) bash tools/bench_all.sh > gpurun_out/${R}_bench_tasks.txt 2>&1; cat gpurun_out/${R}_bench_tasks.txt ;;
    launches)
      ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
        --log-file gpurun_out/${R}_launches_bench.csv python bench.py --steps 64 --warmup 16 --no-sweep --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1
      echo "ncu launches rc=$?" ;;
    ncu)
      timeout 400 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-ige|locomotion|franka|shadow_hand|anymal|cartpole|rollout|task}" -s 2 -c 1 -f -o gpurun_out/${R}_prof_$d \
        python tools/prof.py --task $b --num-envs $c --steps 4 --replicas 4 --hot > gpurun_out/ncu_$d.log 2>&1
      echo "ncu $d rc=$?"; python tools/ncu_summary.py gpurun_out/${R}_prof_$d.ncu-rep > gpurun_out/${R}_ncu_full_$d.txt 2>&1; cat gpurun_out/${R}_ncu_full_$d.txt ;;
    sanitize)   # sanitize:<memcheck|racecheck>  -- ONE tool per gpurun call (B200_PROFILING.md)
      SEL="reference_golden or test_reset_extremes or test_stats_snapshot or test_reducer_as_graph"
      timeout 1500 compute-sanitizer --tool $b --error-exitcode 9 python -m pytest tests -m gpu -x -q -k "$SEL" > gpurun_out/${R}_sanitizer_$b.log 2>&1
      echo "compute-sanitizer $b rc=$?"; tail -8 gpurun_out/${R}_sanitizer_$b.log ;;
    roleclk)    # roleclk:<Task>:<N>  -- debug build with per-role cycle counters, then the normal build again
      python isaacgymenvs-ma_b200/build.py -DIGE_ROLE_CLOCKS > gpurun_out/build_clk.log 2>&1
      timeout 300 python tools/role_clocks.py --task $b --num-envs $c > gpurun_out/${R}_role_clocks_${b}_$c.txt 2>&1; cat gpurun_out/${R}_role_clocks_${b}_$c.txt
      python isaacgymenvs-ma_b200/build.py > gpurun_out/build.log 2>&1 ;;
    scandbg)    # scandbg:<Task>:<N>  -- debug build with the scanner timeline, then the normal build again
      python isaacgymenvs-ma_b200/build.py -DIGE_SCAN_DEBUG > gpurun_out/build_scan.log 2>&1
      timeout 300 python tools/scan_debug.py --task $b --num-envs $c > gpurun_out/${R}_scan_debug_${b}_$c.txt 2>&1; cat gpurun_out/${R}_scan_debug_${b}_$c.txt
      python isaacgymenvs-ma_b200/build.py > gpurun_out/build.log 2>&1 ;;
    kern)       # kern:<Task>:<N>[:ENV=VAL,...]  -- kernel-only timing of one task (no e2e / sweep / cpu legs)
      env ${d//,/ } timeout 300 python bench.py --task $b --num-envs $c --steps 2048 --warmup 64 --no-e2e --no-sweep --no-cpu > gpurun_out/kern_${b}_$c.json 2> gpurun_out/kern_${b}_$c.err
      echo "kern $b $c rc=$?"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/kern_${b}_$c.json")); r = d["roofline"]
    print(f"  ${b} N=$c: step {d['ms_per_step']*1e3:.2f} us  kernel {r['kernel_us']:.2f} us  frac {r['frac']:.3f}")
except Exception as e:
    print("  failed:", e); print(open("gpurun_out/kern_${b}_$c.err").read()[-800:])
PY
      ;;
    build)      # build:<nvcc flags separated by commas>  -- rebuild the library on the box (experiments)
      python isaacgymenvs-ma_b200/build.py ${b//,/ } > gpurun_out/build_$c.log 2>&1; echo "build [$b] rc=$?" ;;
    floors) python tools/floors.py > gpurun_out/${R}_floors.txt 2>&1; cat gpurun_out/${R}_floors.txt ;;
    *) echo "unknown action $act" ;;
  esac
done
ls -la gpurun_out | tail -30
