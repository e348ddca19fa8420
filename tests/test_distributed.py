"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: env sharding, rank seeds and the
episode-statistics all-reduce.  The per-rank step here is the CPU oracle (the CUDA path needs a GPU);
what is tested is that shards + all-reduce reproduce the single-process statistics."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import harness as H


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_env_shard_partitions():
    d = H.pkg.distributed
    for n, w in [(16384 * 8, 8), (10, 3), (7, 8), (1, 2)]:
        spans = [d.env_shard(n, r, w) for r in range(w)]
        assert sum(c for _, c in spans) == n
        pos = 0
        for start, count in spans:
            assert start == pos
            pos += count
    assert d.rank_seed(42, 3) == 45


def _worker(rank, world, port, n_global, q):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    d = H.pkg.distributed
    d.init_process_group("cpu")
    start, count = d.env_shard(n_global, rank, world)
    full = H.pkg.synthetic.make_state("ant", n_global, seed=5, reset_prob=0.1)
    per_env = {k: v for k, v in full.items() if isinstance(v, np.ndarray) and v.ndim >= 1 and v.shape[0] in (n_global, n_global * 8)}
    st = dict(full)
    for k, v in per_env.items():
        rows = v.shape[0] // n_global
        st[k] = v[start * rows:(start + count) * rows].copy()
    env = H.build("ant", st, "cpu", rand_mode="philox", seed=d.rank_seed(0, 0))
    H.oracle_step("ant", env)
    red = d.StatsReducer("cpu")
    red.start(env.stats)
    out = red.result()
    if rank == 0:
        q.put(out)
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_sharded_stats_allreduce_matches_single_process():
    n_global, world = 4096, 2
    ref = H.build("ant", H.pkg.synthetic.make_state("ant", n_global, seed=5, reset_prob=0.1), "cpu", rand_mode="philox")
    H.oracle_step("ant", ref)
    want = ref.stats.numpy()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_global, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert out["steps"] == n_global
    # statistics that do not depend on the (per-env-id) reset noise
    np.testing.assert_allclose(out["done"], want[H.pkg._abi.STAT_DONE], rtol=0, atol=0)
    np.testing.assert_allclose(out["timeout"], want[H.pkg._abi.STAT_TIMEOUT], rtol=0, atol=0)
    np.testing.assert_allclose(out["task0"], want[H.pkg._abi.STAT_TASK0], rtol=1e-4)


def test_reducer_refuses_tasks_without_accumulator():
    """Anymal / BallBalance / ... take no `stats` pointer in their C ABI: StatsReducer must say so instead of
    all-reducing a vector of zeros (and instead of crashing on a None accumulator)."""
    import pytest
    pkg = H.pkg
    red = pkg.distributed.StatsReducer("cpu")

    class NoStats:
        ACCUMULATES_STATS = False
        stats_rows = None

    with pytest.raises(ValueError, match="no episode-statistics accumulator"):
        red.attach(NoStats(), 0)
    tm = pkg.tasks.task_map
    assert all(tm[k].ACCUMULATES_STATS for k in ("Cartpole", "Ant", "Humanoid", "ShadowHand", "AllegroHand", "FrankaPPMA"))
    assert not any(tm[k].ACCUMULATES_STATS for k in ("Anymal", "BallBalance", "Quadcopter", "Ingenuity", "FrankaCabinet"))
