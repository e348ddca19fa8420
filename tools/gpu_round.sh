#!/bin/bash
# One GPU-box pass: parity tests, smoke, default bench, per-task lines, ncu launch list + full captures.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
bash tools/bench_all.sh > gpurun_out/bench_all.txt 2>&1; cat gpurun_out/bench_all.txt
# ncu: launch list of the timed region of the same command
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 64 --warmup 16 --no-sweep --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
for spec in "Ant 16384 ant16k" "Ant 1048576 ant1m" "Humanoid 32768 hum32k" "Humanoid 1048576 hum1m" "ShadowHand 16384 hand16k" "FrankaPPMA 8192 pp8k"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -s 2 -c 1 -f -o gpurun_out/prof_$3 \
    python tools/prof.py --task $1 --num-envs $2 --steps 4 --replicas 4 > gpurun_out/ncu_$3.log 2>&1
  echo "ncu $3 rc=$?"
done
ls -la gpurun_out
