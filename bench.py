#!/usr/bin/env python
"""bench.py -- env-steps/sec of the fused post-physics step (obs + reward + done + reset).

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`, under
torchrun for N > 1 (one rank per GPU); rank 0 prints ONE JSON line.

Workload at every N: BASELINE.json configs[1] / configs[4] -- Ant post_physics_step, 16384 envs per
GPU, synthetic post-simulate state (isaacgymenvs-ma_b200/synthetic.py), natural reset rate.
A "step" is one pass of the hot path over all envs of a GPU: device-side `reset_buf.nonzero()`
inside ONE fused kernel (ige_ant_post_physics), replayed from a CUDA graph.

L2 hygiene: the timed loop rotates through `replicas` independent copies of the whole env state
whose total footprint exceeds 2x the 126 MB L2, so every step streams its tile from HBM.

Keys beyond the base contract: `roofline` (dominant kernel, algorithmic bytes / measured launch
duration vs MEASURED_PEAKS.json), `cpu_baseline` (CPU port on the host cores), `e2e` (host buffers
in pinned memory -> H2D -> step -> D2H every step), `clocks`, `gpu_launches`, `sweep`.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from ige_loader import load_package  # noqa: E402

# frozen algorithmic bytes per env-step (SURVEY.md 8d / BASELINE.md section 2)
ALGO_BYTES = {"Cartpole": 68, "Ant": 556, "Humanoid": 940, "ShadowHand": 1732, "FrankaReachMA": 376,
              "FrankaCollectMA": 488, "FrankaCombineMA": 692, "FrankaPPMA": 720, "Anymal": 544,
              # SURVEY 8f row 2 (compulsory bytes of the reset + obs + reward + done launches)
              "AllegroHand": 748, "BallBalance": 360, "Quadcopter": 396, "Ingenuity": 248, "FrankaReach": 160,
              "FrankaCubeStack2": 200, "FrankaCubeStack": 216, "FrankaCabinet": 380}
ALGO_BYTES_EPILOGUE = {"Cartpole": 85, "Ant": 797, "Humanoid": 1373, "ShadowHand": 2577, "FrankaReachMA": 530,
                       "FrankaCollectMA": 714, "FrankaCombineMA": 1078, "FrankaPPMA": 1122, "Anymal": 737,
                       "AllegroHand": 1101, "BallBalance": 457, "Quadcopter": 481, "Ingenuity": 301, "FrankaReach": 217,
                       "FrankaCubeStack2": 285, "FrankaCubeStack": 293, "FrankaCabinet": 473}
KERNEL_NAME = {"Cartpole": "cartpole_post_physics_kernel", "Ant": "locomotion_pipelined_kernel<Ant>",
               "Humanoid": "locomotion_pipelined_kernel<Humanoid>",
               "ShadowHand": "shadow_hand_post_physics_kernel (+ shadow_hand_reset_kernel)",
               "FrankaReachMA": "franka_ma_kernel", "FrankaCollectMA": "franka_ma_kernel",
               "FrankaCombineMA": "franka_ma_kernel (reset + compute)", "FrankaPPMA": "franka_ma_kernel (reset + compute)",
               "Anymal": "anymal_post_physics_kernel",
               "AllegroHand": "shadow_hand_post_physics_fast_kernel<6> (+ shadow_hand_reset_kernel)",
               "BallBalance": "ball_balance_pre_kernel + ball_balance_post_kernel",
               "Quadcopter": "quadcopter_pre_kernel + quadcopter_post_kernel",
               "Ingenuity": "ingenuity_pre_kernel + ingenuity_post_kernel", "FrankaReach": "franka_sa_kernel<0>",
               "FrankaCubeStack2": "franka_sa_kernel<1>", "FrankaCubeStack": "franka_sa_kernel<2>",
               "FrankaCabinet": "franka_cabinet_post_kernel"}
L2_BYTES = 126 * 1024 * 1024
HORIZON = 16  # cfg/train/AntPPO.yaml:65 horizon_length -> one stats all-reduce per horizon


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, enabled=True):
        self.idx, self.samples, self._stop, self._t = gpu_index, [], threading.Event(), None
        self.enabled = enabled

    def _once(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                  str(self.idx)], capture_output=True, text=True, timeout=5).stdout.strip()
            if out:
                self.samples.append([x.strip() for x in out.split(",")])
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self._once()
            self._stop.wait(0.1)

    def __enter__(self):
        self.t0 = time.time()
        if self.enabled:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        if not self.enabled:
            return
        self._stop.set()
        self._t.join(timeout=10)
        if not self.samples:
            self._once()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.samples[0][1]),
                "reasons": reasons, "samples": len(self.samples)}


def ncu_traffic(task, n):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json),
    or None when this (task, num_envs) was not captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(task, {}).get(str(n))
    except (OSError, ValueError):
        return None


def build_replicas(pkg, task, num_envs, replicas, device, seed, fall_frac, fuse_epilogue, rank):
    envs = []
    for r in range(replicas):
        st = pkg.synthetic.make_state(pkg.factory._SYN_NAME[task], num_envs, seed=seed + 1000 * rank + r,
                                      reset_prob=0.02, fall_frac=fall_frac)
        env = pkg.factory.from_state(task, st, device, rand_mode="philox", seed=seed + rank,
                                     fuse_epilogue=fuse_epilogue)
        env.sim.record = False
        envs.append(env)
    return envs


def capture(envs, fn, pre=None, post=None, error_mode="global"):
    """Capture fn(env) for every env of `envs`, in order, into one CUDA graph.  `pre` / `post` run inside the
    capture before / after the steps (the statistics branch forks and joins there)."""
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for e in envs:      # warm (module load etc.) outside capture
            fn(e)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, capture_error_mode=error_mode):
        if pre is not None:
            pre()
        for e in envs:
            fn(e)
        if post is not None:
            post()
    return g


def check_stats_reduce(pkg, reducer, chunks, dist, world, device):
    """Does the statistics all-reduce the timed graphs carry (peer-mailbox kernel or NCCL) really SUM over ranks?
    One horizon is accumulated (graph 0), its per-rank column sums are read back, the next graph folds and reduces
    it, and the result is compared with the float64 sum of all ranks' vectors gathered over plain NCCL -- a path
    that shares nothing with ige_stats_exchange.  Bar: 1e-5 relative (fp32 sums of <= 32 rows x N ranks)."""
    abi = pkg._abi
    torch.cuda.synchronize()
    for r in reducer.rows:
        r.zero_()
    chunks[0].replay()                       # accumulates horizon A into rows[0] (and folds the zeroed rows[1])
    torch.cuda.synchronize()
    local = reducer.rows[0].view(abi.IGE_STATS_ROWS, abi.IGE_STATS_STRIDE)[:, :abi.IGE_NUM_STATS].double().sum(0)
    chunks[1].replay()                       # folds + all-reduces horizon A
    torch.cuda.synchronize()
    got = reducer.reduced.double().clone()
    err = None
    try:
        reducer.check_error()
    except RuntimeError as e:
        err = str(e)
    want = local.clone()
    if dist is not None:
        parts = [torch.zeros_like(local) for _ in range(world)]
        dist.all_gather(parts, local)
        want = torch.stack(parts).sum(0)
    rel = float(((got - want).abs() / want.abs().clamp_min(1.0)).max())
    ok = err is None and rel <= 1e-5 and float(want[0]) > 0
    if dist is not None:                     # the line says ok only if EVERY rank's vector checked out
        flag = torch.tensor([1.0 if ok else 0.0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item() > 0.5)
    return {"ok": ok, "ranks": world, "mode": reducer.mode, "max_rel_err": rel, "error": err,
            "steps_summed": float(want[0]), "reward_sum": float(want[1])}   # IGE_STAT_STEPS, IGE_STAT_REWARD


def bench_config(task, n, world):
    """`config` of the JSON line -- the SAME dict for both arms (ours / --impl reference)."""
    return {"workload": f"{task} post_physics_step num_envs={n} per GPU" +
                        (" (BASELINE.json configs[1])" if task == "Ant" and n == 16384 else ""),
            "task": task, "num_envs_per_gpu": n, "global_envs": n * world,
            "parallelism": f"env-shard x{world}", "reset": "natural (2% fallen envs/step)",
            "stats": f"episode-statistics all-reduce every {HORIZON} steps"}


def run_ours(args):
    pkg = load_package()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the post-physics kernels have no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    numa_node = pkg.distributed.bind_to_gpu_numa_node(local_rank)   # before any pinned arena exists
    dist = None
    if world > 1:
        import torch.distributed as dist
        pkg.distributed.init_process_group(device)      # NCCL, one rank per GPU

    task, n = args.task, args.num_envs
    algo = ALGO_BYTES[task]
    state_bytes = ALGO_BYTES_EPILOGUE[task] * n
    replicas = args.replicas or max(2 * HORIZON, -(-2 * L2_BYTES // state_bytes))
    replicas = -(-replicas // (2 * HORIZON)) * (2 * HORIZON)      # an even number of horizons (double-buffered stats)
    envs = build_replicas(pkg, task, n, replicas, device, args.seed, args.fall_frac, False, rank)
    lib = pkg._lib

    # ---- graphs: one per horizon = HORIZON consecutive replicas + (forked branch) the statistics snapshot and
    #      NCCL all-reduce of the PREVIOUS horizon's accumulator; the remainder steps are one more graph ----
    reducer = pkg.distributed.StatsReducer(device, mode=None if args.stats_reduce in ("auto", "none") else args.stats_reduce)
    reducer.warm()
    # tasks whose kernels have no statistics accumulator (ACCUMULATES_STATS False) get no reduction branch
    use_branch = args.stats_reduce != "none" and envs[0].ACCUMULATES_STATS
    n_chunks = replicas // HORIZON
    for c in range(n_chunks):
        for e in envs[c * HORIZON:(c + 1) * HORIZON]:
            if use_branch:
                reducer.attach(e, c)
    mode = "thread_local" if world > 1 else "global"   # NCCL's watchdog thread may touch the device during capture
    chunks = [capture(envs[c * HORIZON:(c + 1) * HORIZON], lambda e: e.hot_step(),
                      pre=(lambda c=c: reducer.fork(reducer.rows[(c & 1) ^ 1])) if use_branch else None,
                      post=reducer.join if use_branch else None, error_mode=mode)
              for c in range(n_chunks)]
    full, rem = divmod(args.steps, HORIZON)
    # the remainder continues with the replicas (and the accumulator) of the horizon that would come next
    rem_envs = [envs[((full % n_chunks) * HORIZON + i) % replicas] for i in range(rem)]
    rem_graph = [capture(rem_envs, lambda e: e.hot_step(), error_mode=mode)] if rem else []

    # Short runs (the driver's 20-step window is ~120 us): ALL timed steps in ONE graph launch, so a rank's host
    # jitter is paid once, not once per horizon graph -- the MAX over ranks of several back-to-back launches was what
    # kept the 8-GPU curve under 0.95 in round 2.  Long runs keep one graph per horizon (launch cost amortised).
    whole = None
    if args.steps <= 8 * HORIZON:
        whole = torch.cuda.CUDAGraph()
        with torch.cuda.graph(whole, capture_error_mode=mode):
            for i in range(full):
                c = i % n_chunks
                if use_branch:
                    reducer.fork(reducer.rows[(c & 1) ^ 1])
                for e in envs[c * HORIZON:(c + 1) * HORIZON]:
                    e.hot_step()
                if use_branch:
                    reducer.join()
            for e in rem_envs:
                e.hot_step()

    def timed(steps, warmup):
        def body():
            if whole is not None:
                whole.replay()
                return
            for i in range(full):
                chunks[i % n_chunks].replay()
            for g in rem_graph:
                g.replay()
        w = max(-(-warmup // HORIZON), 3)
        for i in range(w):
            chunks[(i - w) % n_chunks].replay()      # ends on the last chunk: body() starts at chunk 0
        if whole is not None:                        # the first launch of an instantiated graph uploads it: not timed
            for _ in range(max(3, -(-warmup // max(steps, 1)))):
                whole.replay()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.cudart().cudaProfilerStart()     # `ncu --profile-from-start off` sees exactly the timed region
        t0.record()
        body()
        t1.record()
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStop()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        return t0.elapsed_time(t1) * 1e-3

    # Clocks: ONE sampler per box (local rank 0 forks nvidia-smi at 10 Hz for its own GPU); in round 2 every rank did,
    # 8 subprocess trains contending for the driver lock and the cores that issue the launches inside a 120-us window.
    clk = ClockSampler(local_rank, enabled=(local_rank == 0))
    clk.__enter__()
    time.sleep(0.5)          # the first nvidia-smi fork is over before the timed window opens
    elapsed = timed(args.steps, args.warmup)
    if dist is not None:
        t = torch.tensor([elapsed], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed = float(t.item())
    value = n * world * args.steps / elapsed
    gpu_launches = args.steps + full          # one fused step kernel per step + one statistics snapshot per horizon
    stats_check = check_stats_reduce(pkg, reducer, chunks, dist, world, device) if use_branch else None

    # ---- dominant kernel alone (roofline): fused kernel only, same rotation, CUDA events ----
    kgraph = capture(envs, lambda e: e.hot_step())
    for _ in range(3):
        kgraph.replay()
    torch.cuda.synchronize()
    reps = max(1, 4096 // replicas)
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record()
    for _ in range(reps):
        kgraph.replay()
    k1.record()
    torch.cuda.synchronize()
    kernel_s = k0.elapsed_time(k1) * 1e-3 / (reps * replicas)
    peak, peak_kind = measured_peaks()
    achieved = algo * n / kernel_s / 1e9

    out = {
        "metric": "env-steps/sec (obs+reward+reset)", "value": value, "unit": "env-steps/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(task, n, world),
        "how": {"l2": f"inputs larger than L2: rotating {replicas} state replicas = "
                      f"{replicas * state_bytes / 2**20:.0f} MiB > 126 MiB L2",
                "launch": f"one CUDA graph per horizon of {HORIZON} steps; the fold + all-reduce of the previous horizon's "
                          f"statistics is a forked branch of the same graph ({reducer.collectives} collectives captured, "
                          f"world={world}); remainder steps = one more graph",
                "stats_reduce": ("none (diagnostic, or a task without an accumulator)" if not use_branch else
                                 {"peer": "ige_stats_exchange: one kernel, fold + NVLink peer-mailbox all-reduce",
                                  "nccl": "ige_stats_snapshot + ncclAllReduce captured in the branch"
                                          + (f" (peer mapping unavailable: {reducer.peer_error})"
                                             if getattr(reducer, "peer_error", None) else ""),
                                  "local": "ige_stats_snapshot (single rank)"}[reducer.mode]),
                "reset_noise": "in-kernel Philox",
                "pdl": "programmatic dependent launch: on" if lib.lib().ige_use_pdl() else "programmatic dependent launch: off",
                "host_numa_node": numa_node},
        "gpu_launches": gpu_launches,
        "stats_check": stats_check,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(task, n), "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})",
                     "kernel": KERNEL_NAME[task], "kernel_us": kernel_s * 1e6,
                     "algorithmic_bytes_per_env": algo},
    }
    if not args.no_e2e:
        out["e2e"] = run_e2e(pkg, args, device, dist=dist, world=world)   # collective: every rank runs it
    if rank == 0 and world == 1 and not args.no_sweep:
        out["sweep"] = run_sweep(pkg, args, device, peak)
        out["variants"] = run_variants(pkg, args, device, peak)
    # keep sampling for >= 1 s of GPU load so that short runs still see clocks under load
    t_end = time.time() + max(0.0, 1.0 - (time.time() - clk.t0))
    while time.time() < t_end:
        kgraph.replay()          # step kernels only: no collective, so ranks need not agree on the count
        torch.cuda.synchronize()
    clk.__exit__()
    out["clocks"] = clk.summary()
    if rank == 0 and world == 1 and not args.no_cpu:
        out["cpu_baseline"] = host_baselines(pkg, args, seconds=args.cpu_seconds)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if dist is not None:
        # NCCL keeps a reference per captured graph and ncclCommDestroy waits for them: drop the graphs that hold
        # the captured all-reduce BEFORE tearing the process group down (a rank that skipped this hung forever in
        # destroy_process_group after printing its line).  The timer is a last resort against a wedged teardown.
        torch.cuda.synchronize()
        chunks.clear()
        rem_graph.clear()
        import gc
        gc.collect()
        torch.cuda.synchronize()
        threading.Timer(30.0, lambda: os._exit(0)).start()
        reducer.close()
        dist.barrier()
        dist.destroy_process_group()
        os._exit(0)


def _time_steps(step, steps, warmup, device, dist):
    for _ in range(max(3, warmup)):
        step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    el = max(e0.elapsed_time(e1) * 1e-3, 0.0)
    if dist is not None:
        t = torch.tensor([el, wall], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        el, wall = float(t[0].item()), float(t[1].item())
    return el, wall


def _nbytes(ts):
    return sum(t.numel() * t.element_size() for t in ts)


def run_e2e(pkg, args, device, dist=None, world=1):
    """Same metric through the public API with HOST buffers (the reference's sim_device=cpu pipeline=cpu
    situation): every step the simulator tensors + actions start in pinned host memory and obs / rew /
    reset end in pinned host memory.  Two ways to get them across PCIe, both timed, the faster reported:
      staged : cudaMemcpyAsync H2D of every input, the fused kernel, cudaMemcpyAsync D2H of the results;
      mapped : env.map_host_sim() -- the fused kernel itself streams the pinned host tensors in and the
               results out (reads and writes overlap on the duplex link, one launch, no staging pass)."""
    task, n = args.task, args.num_envs
    syn = pkg.factory._SYN_NAME[task]
    res = {}
    # ---- staged copies ----
    st = pkg.synthetic.make_state(syn, n, seed=args.seed + 77, reset_prob=0.02, fall_frac=args.fall_frac)
    env = pkg.factory.from_state(task, st, device, rand_mode="philox", seed=args.seed, fuse_epilogue=False)
    env.sim.record = False
    names = pkg.factory.sim_tensor_names(task) + (["actions"] if hasattr(env, "actions") else [])
    dev_in = [env.sim.tensors[k] if k in env.sim.tensors else getattr(env, k) for k in names]
    host_in = [torch.empty(t.shape, dtype=t.dtype).pin_memory().copy_(t.cpu()) for t in dev_in]
    dev_out = [env.obs_buf, env.rew_buf, env.reset_buf]
    host_out = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in dev_out]

    def step_staged():
        for d, h in zip(dev_in, host_in):
            d.copy_(h, non_blocking=True)
        env.hot_step()
        for d, h in zip(dev_out, host_out):
            h.copy_(d, non_blocking=True)

    el, wall = _time_steps(step_staged, args.steps, args.warmup, device, dist)
    res["staged"] = {"value": n * world * args.steps / el, "ms_per_step": el / args.steps * 1e3,
                     "wall_ms_per_step": wall / args.steps * 1e3, "h2d_bytes_per_step": _nbytes(host_in),
                     "d2h_bytes_per_step": _nbytes(host_out)}
    del env
    # ---- mapped: rotating host-resident replicas so no step re-reads lines a previous step left in L2 ----
    has_host = bool(getattr(pkg.tasks.task_map[task], "HOST_SIM_ATTRS", ()))
    if has_host:
        reps = 8
        envs, h2d, d2h = [], 0, 0
        for r in range(reps):
            st = pkg.synthetic.make_state(syn, n, seed=args.seed + 177 + r, reset_prob=0.02, fall_frac=args.fall_frac)
            e = pkg.factory.from_state(task, st, device, rand_mode="philox", seed=args.seed, fuse_epilogue=False)
            e.sim.record = False
            moved = e.map_host_sim()
            h2d = _nbytes([moved[k] for k in moved if k not in ("obs_buf", "rew_buf")])  # reset_buf is read too
            d2h = _nbytes([moved[k] for k in ("obs_buf", "rew_buf", "reset_buf")])
            envs.append(e)
        it = [0]

        def step_mapped():
            envs[it[0] % reps].hot_step()
            it[0] += 1

        el, wall = _time_steps(step_mapped, args.steps, args.warmup, device, dist)
        ok = all(bool(torch.isfinite(e.obs_buf).all()) and float(e.obs_buf.abs().sum()) > 0 for e in envs)
        res["mapped"] = {"value": n * world * args.steps / el, "ms_per_step": el / args.steps * 1e3,
                         "wall_ms_per_step": wall / args.steps * 1e3, "h2d_bytes_per_step": h2d,
                         "d2h_bytes_per_step": d2h, "results_on_host_ok": ok}
        del envs
    # ---- pipelined: G env groups, one arena copy per group and direction, one CUDA graph per step ----
    if has_host:
        for groups in (4, 8):
            if n % (groups * 32):
                continue
            pipe = pkg.host_pipeline.HostStagedPipeline(task, n, device, groups=groups, seed=args.seed + 277,
                                                        fuse_epilogue=False).capture()
            el, wall = _time_steps(pipe.step, args.steps, args.warmup, device, dist)
            ok = all(bool(torch.isfinite(h["obs_buf"]).all()) and float(h["obs_buf"].abs().sum()) > 0
                     for h in pipe.host_out)
            res[f"pipelined_g{groups}"] = {
                "value": n * world * args.steps / el, "ms_per_step": el / args.steps * 1e3,
                "wall_ms_per_step": wall / args.steps * 1e3, "h2d_bytes_per_step": pipe.h2d_bytes_per_step,
                "d2h_bytes_per_step": pipe.d2h_bytes_per_step, "results_on_host_ok": ok}
            del pipe
    # ---- the floor of this box: nothing but the copies.  One pinned buffer of the staged mode's H2D bytes and one of
    #      its D2H bytes per step, the two directions streaming concurrently on their own streams, all ranks at once.
    #      This is what an ideal pipeline (H2D of step t+1 || kernel of t || D2H of t-1) would cost on this host. ----
    nb_in, nb_out = res["staged"]["h2d_bytes_per_step"], res["staged"]["d2h_bytes_per_step"]
    h_in = torch.empty(nb_in, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nb_out, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(nb_in, dtype=torch.uint8, device=device)
    d_out = torch.zeros(nb_out, dtype=torch.uint8, device=device)
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)

    def copies_only(k):
        cur = torch.cuda.current_stream(device)
        s_in.wait_stream(cur)
        s_out.wait_stream(cur)
        with torch.cuda.stream(s_in):
            for _ in range(k):
                d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s_out):
            for _ in range(k):
                h_out.copy_(d_out, non_blocking=True)
        cur.wait_stream(s_in)
        cur.wait_stream(s_out)

    copies_only(max(3, args.warmup))
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    copies_only(args.steps)
    f1.record()
    torch.cuda.synchronize()
    floor_s = f0.elapsed_time(f1) * 1e-3
    if dist is not None:
        t = torch.tensor([floor_s], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        floor_s = float(t.item())
    floor_ms = floor_s / args.steps * 1e3
    best = max(res, key=lambda k: res[k]["value"])
    out = dict(res[best])
    out.update({"pcie_floor_ms": floor_ms,
                "pcie_floor": f"{nb_in} B H2D || {nb_out} B D2H per step, cudaMemcpyAsync only, {world} rank(s) at once: "
                              f"{nb_in / floor_ms / 1e6:.1f} + {nb_out / floor_ms / 1e6:.1f} GB/s per GPU",
                "vs_pcie_floor": res[best]["ms_per_step"] / floor_ms,
                "staged_ms_per_step": res["staged"]["ms_per_step"]})     # the SAME mode at every N, next to best-of
    out.update({"unit": "env-steps/s", "mode": best, "modes": res,
                "api": f"{task}.post_physics_step() (ctypes -> C ABI), pinned host buffers; "
                       + {"mapped": "map_host_sim(): the kernel reads/writes pinned host memory over PCIe",
                          "staged": "cudaMemcpyAsync H2D + kernel + D2H"}.get(
                              best, "HostStagedPipeline: env groups, arena H2D -> kernel -> arena D2H as parallel "
                                    "branches of one CUDA graph")})
    return out


def run_sweep(pkg, args, device, peak):
    """Roofline of the fused kernel vs num_envs: same kernel, larger batches."""
    res = []
    task = args.task
    for n in (16384, 65536, 262144, 1048576):
        per = ALGO_BYTES_EPILOGUE[task] * n
        reps = max(2, min(32, -(-2 * L2_BYTES // per)))
        envs = build_replicas(pkg, task, n, reps, device, args.seed + 5, args.fall_frac, False, 0)

        g = capture(envs, lambda e: e.hot_step())
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        loops = max(1, 512 // reps)
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(loops):
            g.replay()
        b_.record()
        torch.cuda.synchronize()
        t = a.elapsed_time(b_) * 1e-3 / (loops * reps)
        gbs = ALGO_BYTES[task] * n / t / 1e9
        res.append({"num_envs": n, "kernel_us": t * 1e6, "achieved_gbs": gbs, "frac": gbs / peak,
                    "env_steps_per_s": n / t})
        del envs, g
        torch.cuda.empty_cache()
    return res


def run_variants(pkg, args, device, peak):
    """SURVEY.md 8d timing-method items beside the headline (cold L2, natural reset rate): the same kernel with its
    state resident in L2 (one replica replayed: 13 MB at Ant 16384 << 126 MB) and at the worst-case reset rate
    p = 1.0 (every env below the termination height: each step flags every env, the next one resets it)."""
    task, n = args.task, args.num_envs
    out = {}

    def timed(envs, loops):
        g = capture(envs, lambda e: e.hot_step())
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(loops):
            g.replay()
        b_.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b_) * 1e-3 / (loops * len(envs))

    def row(t, note):
        gbs = ALGO_BYTES[task] * n / t / 1e9
        return {"kernel_us": t * 1e6, "achieved_gbs": gbs, "frac": gbs / peak, "env_steps_per_s": n / t, "how": note}

    warm = build_replicas(pkg, task, n, 1, device, args.seed + 9, args.fall_frac, False, 0)
    out["warm_l2"] = row(timed(warm * 16, 64), "one state replica replayed back to back: inputs and outputs stay in L2")
    del warm
    reps = max(2, min(32, -(-2 * L2_BYTES // (ALGO_BYTES_EPILOGUE[task] * n))))
    try:
        envs = build_replicas(pkg, task, n, reps, device, args.seed + 11, 1.0, False, 0)
        for e in envs:
            e.reset_buf.fill_(1)
        out["reset_p1.0"] = row(timed(envs, max(1, 512 // reps)),
                                "cold L2, every env flagged every step (fall_frac = 1.0): reset_idx runs for all envs")
        del envs
    except Exception as exc:          # tasks whose synthetic maker has no fall knob
        out["reset_p1.0"] = {"unavailable": str(exc)[:120]}
    torch.cuda.empty_cache()
    return out


_FULL_AFFINITY = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None   # before any NUMA binding


class host_affinity:
    """The CPU baselines get every core this process started with: run_ours() narrows the process to the GPU's NUMA
    node (distributed.bind_to_gpu_numa_node) for the pinned arenas, which must not shrink the reference's thread
    pool and understate it.  Restores the narrowed mask on exit."""

    def __enter__(self):
        self.saved = os.sched_getaffinity(0) if _FULL_AFFINITY is not None else None
        if self.saved is not None and self.saved != _FULL_AFFINITY:
            os.sched_setaffinity(0, _FULL_AFFINITY)
        return self

    def __exit__(self, *exc):
        if self.saved is not None and self.saved != _FULL_AFFINITY:
            os.sched_setaffinity(0, self.saved)


def host_cores():
    """Threads a CPU baseline may really use: the affinity mask, not the machine's core count."""
    return len(os.sched_getaffinity(0)) if _FULL_AFFINITY is not None else (os.cpu_count() or 1)


def cpu_baseline(pkg, args, seconds=6.0, steps=None, warmup=3):
    """The CPU port (oracle/ige_oracle.c, plain C + pthreads) of the same step on the host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import harness as H
    from oracle import oracle
    task, n = args.task, args.num_envs
    cores = host_cores()
    syn = pkg.factory._SYN_NAME[task]
    if syn not in H.TASK_NAME:
        return None
    oracle.lib().ige_oracle_set_threads(cores)
    st = pkg.synthetic.make_state(syn, n, seed=args.seed, reset_prob=0.02, fall_frac=args.fall_frac)
    obj = H.build(syn, st, "cpu", rand_mode="philox", fuse_epilogue=False)
    for _ in range(warmup):
        H.oracle_step(syn, obj)
    done, t0 = 0, time.perf_counter()
    while True:
        H.oracle_step(syn, obj)
        done += 1
        if steps is not None:
            if done >= steps:
                break
        elif time.perf_counter() - t0 >= seconds:
            break
    el = time.perf_counter() - t0
    return {"value": n * done / el, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": f"{done} steps of {task} num_envs={n} (oracle/ige_oracle.c, {cores} pthreads), "
                      f"{el:.1f} s", "ms_per_step": el / done * 1e3}


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def reference_torch_baseline(pkg, args, seconds=6.0, steps=None, warmup=5, threads=None, windows=1):
    """The reference's OWN implementation of the step on the host cores (north_star: "the reference torch path
    ... sim_device=cpu pipeline=cpu"): the unmodified task class from the reference tree (/root/reference in the
    build container, else the files tools/vendor_reference.py staged under oracle/_ref/), built around the same
    synthetic state with a recording-free FakeGym (tests/ref_harness.py), its
    ``post_physics_step()`` timed with time.perf_counter.  Returns None when the reference is not available."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_shim
    if not ref_shim.reference_available():
        return None
    import ref_harness
    task, n = args.task, args.num_envs
    syn = pkg.factory._SYN_NAME[task]
    if syn not in ref_harness.BUILDERS:
        return None
    cores = threads or host_cores()
    torch.set_num_threads(cores)
    st = pkg.synthetic.make_state(syn, n, seed=args.seed, reset_prob=0.02, fall_frac=args.fall_frac)
    cfg = pkg.factory.task_cfg(task)
    if syn.startswith("franka_"):
        cfg = pkg.synthetic.franka_ma_cfg(syn[len("franka_"):], int(st["cfg_num_agents"]), int(st["cfg_num_targets"]))
    obj = ref_harness.BUILDERS[syn](st, cfg)
    obj.gym.record = False
    if syn == "shadow_hand":       # its resets run in pre_physics_step (shadow_hand.py:670-682)
        act = obj.actions.clone()

        def step():
            obj.pre_physics_step(act)
            obj.post_physics_step()
    else:
        step = obj.post_physics_step
    with torch.no_grad():
        for _ in range(max(warmup, 3)):       # TorchScript's profiling executor specialises on the first calls
            step()
        def window():
            done, t0 = 0, time.perf_counter()
            while True:
                step()
                done += 1
                if steps is not None:
                    if done >= steps:
                        break
                elif time.perf_counter() - t0 >= seconds:
                    break
            return done, time.perf_counter() - t0
        done, el = window()
        # A 20-step window is ~90 ms of CPU time and swings by +-25 % run to run.  With `windows` > 1 the same K-step
        # window is repeated (while a window stays under 10 s) and the MEDIAN window is reported.
        extra = {}
        if windows > 1 and steps is not None and el < 10.0:
            runs = [(done, el)] + [window() for _ in range(windows - 1)]
            per_window = sorted(n * d / t for d, t in runs)
            extra = {"windows": len(runs), "first_window_value": n * done / el, "min_window_value": per_window[0],
                     "max_window_value": per_window[-1], "value_is": f"median of {len(runs)} windows of {steps} steps"}
            done, el = sorted(runs, key=lambda r: r[1] / r[0])[len(runs) // 2]
    return {**extra, "value": n * done / el, "unit": "env-steps/s", "cores": cores, "kind": "reference",
            "cpu_model": cpu_model(), "torch_threads": torch.get_num_threads(),
            "sample": f"{done} steps of the reference's own {task}.post_physics_step (unmodified "
                      f"isaacgymenvs/tasks/{syn}.py on CPU torch {torch.__version__}, FakeGym, num_envs={n}), "
                      f"{cores} threads, {el:.1f} s", "ms_per_step": el / done * 1e3,
            "reference_root": os.path.relpath(ref_shim.REFERENCE_ROOT, ROOT) if ref_shim.REFERENCE_ROOT.startswith(ROOT)
            else ref_shim.REFERENCE_ROOT}


def host_baselines(pkg, args, seconds, steps=None, warmup=3):
    """cpu_baseline object: the reference's torch path when available (kind "reference"), with the C port of
    the oracle beside it as `cpu_port`; the port alone (kind "port") otherwise."""
    with host_affinity():
        port = cpu_baseline(pkg, args, seconds=seconds, steps=steps, warmup=warmup)
        ref = reference_torch_baseline(pkg, args, seconds=seconds, steps=steps, warmup=max(warmup, 3))
        if ref is None:
            return port
        if port is None:
            return ref
        one = reference_torch_baseline(pkg, args, seconds=min(seconds, 3.0), steps=None if steps is None else max(2, steps // 4),
                                       warmup=3, threads=1)
    ref["single_thread"] = {"value": one["value"], "ms_per_step": one["ms_per_step"]}
    ref["cpu_port"] = port
    return ref


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pkg = load_package()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # The reference arm runs the reference and nothing of this repository: no C port beside it (building the port's
    # CPU mirror asks libige_b200.so for scratch sizes, which put the library into this process's maps), only rank 0's
    # 16 384-env shard whatever N is (stated in config.reference_arm_envs).  The port is the fallback when the
    # reference tree is absent.
    with host_affinity():
        cb = reference_torch_baseline(pkg, args, seconds=args.cpu_seconds, steps=args.steps, warmup=max(args.warmup, 3),
                                      windows=5)
    if cb is None:
        cb = host_baselines(pkg, args, seconds=args.cpu_seconds, steps=args.steps, warmup=max(args.warmup, 3))
    out = {"impl": "reference", "metric": "env-steps/sec (obs+reward+reset)", "value": cb["value"],
           "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic",
           "config": {**bench_config(args.task, args.num_envs, world), "reference_arm_envs": args.num_envs},
           "cpu_baseline": cb,
           "e2e": {"value": cb["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2048)
    ap.add_argument("--warmup", type=int, default=64)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--task", default="Ant", choices=list(ALGO_BYTES))
    ap.add_argument("--num-envs", type=int, default=16384)
    ap.add_argument("--replicas", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--fall-frac", type=float, default=0.02)
    ap.add_argument("--cpu-seconds", type=float, default=8.0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--stats-reduce", default="auto", choices=["auto", "peer", "nccl", "none"],
                    help="multi-GPU statistics all-reduce: peer-mailbox kernel (default), NCCL, or none (diagnostic)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
