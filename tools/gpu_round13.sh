#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "anymal_mirror" 2>&1 | tail -3
timeout 200 python bench.py --task Anymal --num-envs 16384 --no-e2e --no-cpu --no-sweep > gpurun_out/bench_anymal16k.json 2> gpurun_out/bench_anymal.err
timeout 200 python bench.py --task Anymal --num-envs 1048576 --steps 256 --warmup 16 --no-e2e --no-cpu --no-sweep > gpurun_out/bench_anymal1m.json 2>> gpurun_out/bench_anymal.err
tail -3 gpurun_out/bench_anymal.err
python - <<PY
import json
for f in ("gpurun_out/bench_anymal16k.json","gpurun_out/bench_anymal1m.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"]*1e3, "us", d["value"]/1e9, "G env-steps/s", d["roofline"]["frac"])
    except Exception as e: print(f, "ERR", e)
PY
