"""GPU tests of the episode-statistics path: the snapshot kernel, the double-buffered
StatsReducer as a forked branch of a CUDA graph, and -- with >= 2 GPUs -- the NCCL all-reduce captured in that
graph: the reduced vector must equal the sum of the per-rank vectors."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import harness as H

pytestmark = pytest.mark.gpu
abi = H.pkg._abi


def _col_sums(rows):
    return rows.view(abi.IGE_STATS_ROWS, abi.IGE_STATS_STRIDE)[:, :abi.IGE_NUM_STATS].double().sum(0).cpu().numpy()


def test_stats_snapshot_kernel():
    dev = "cuda:0"
    g = torch.Generator().manual_seed(3)
    rows = torch.rand(abi.IGE_STATS_ROWS * abi.IGE_STATS_STRIDE, generator=g).to(dev)
    want = _col_sums(rows)
    out = torch.full((abi.IGE_NUM_STATS,), 7.0, device=dev)
    a = abi.IgeStatsSnapshotArgs()
    a.stats, a.out, a.zero_stats, a.accumulate = rows.data_ptr(), out.data_ptr(), 0, 0
    H.pkg._lib.call("ige_stats_snapshot", a, stream=torch.cuda.current_stream().cuda_stream)
    np.testing.assert_allclose(out.cpu().numpy(), want, rtol=1e-6)
    assert float(rows.abs().sum()) > 0                      # untouched
    a.zero_stats, a.accumulate = 1, 1
    H.pkg._lib.call("ige_stats_snapshot", a, stream=torch.cuda.current_stream().cuda_stream)
    np.testing.assert_allclose(out.cpu().numpy(), 2 * want, rtol=1e-6)
    # only the statistic columns are cleared; the padding columns are never written by the step kernels
    assert float(rows.view(abi.IGE_STATS_ROWS, abi.IGE_STATS_STRIDE)[:, :abi.IGE_NUM_STATS].abs().sum()) == 0.0


def _horizon_graph(envs, reducer, parity, mode="global"):
    """One horizon: every env of `envs` steps once, adding into rows[parity]; the forked branch folds + all-reduces
    rows[parity ^ 1]."""
    for e in envs:
        reducer.attach(e, parity)
        e.sim.record = False
        e._constants()        # the per-task constants are validated with torch ops once: not inside a capture
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s, capture_error_mode=mode):
            reducer.fork(reducer.rows[parity ^ 1])
            for e in envs:
                e.hot_step()
            reducer.join()
    return g


def _make_envs(n, count, seed, device):
    envs = []
    for r in range(count):
        st = H.pkg.synthetic.make_state("ant", n, seed=seed + r, reset_prob=0.05)
        envs.append(H.pkg.factory.from_state("Ant", st, device, rand_mode="philox", seed=seed, fuse_epilogue=False))
    return envs


def test_reducer_as_graph_branch_single_gpu():
    """Two horizons (A, B) alternate; after replaying A, B, A the reduced vector is what horizon B accumulated
    (the branch of the second A folds B's buffer), and nothing of horizon A leaks into it."""
    dev = torch.device("cuda:0")
    n, per = 2048, 4
    red = H.pkg.distributed.StatsReducer(dev)
    envs_a, envs_b = _make_envs(n, per, 10, dev), _make_envs(n, per, 20, dev)
    # expected per-horizon sums from un-graphed twins of the same states
    want = {}
    for name, seed in (("a", 10), ("b", 20)):
        tw = _make_envs(n, per, seed, dev)
        for e in tw:
            e.hot_step()
        torch.cuda.synchronize()
        want[name] = sum(_col_sums(e.stats_rows) for e in tw)
    ga, gb = _horizon_graph(envs_a, red, 0), _horizon_graph(envs_b, red, 1)
    torch.cuda.synchronize()
    for r in red.rows:
        r.zero_()
    # (capture does not execute; the eager warm-up above did not happen for these envs)
    ga.replay()                      # A1 -> rows[0]; folds rows[1] (empty)
    gb.replay()                      # B1 -> rows[1]; folds rows[0] = A1
    torch.cuda.synchronize()
    got_a = red.reduced.double().cpu().numpy()
    ga.replay()                      # A2 -> rows[0]; folds rows[1] = B1
    torch.cuda.synchronize()
    got_b = red.reduced.double().cpu().numpy()
    assert got_a[abi.STAT_STEPS] == n * per and got_b[abi.STAT_STEPS] == n * per
    np.testing.assert_allclose(got_a, want["a"], rtol=2e-3, atol=2e-2)
    np.testing.assert_allclose(got_b, want["b"], rtol=2e-3, atol=2e-2)
    out = red.result()
    assert out["steps"] == n * per


def test_reducer_eager_swap():
    dev = torch.device("cuda:0")
    env = _make_envs(1024, 1, 5, dev)[0]
    red = H.pkg.distributed.StatsReducer(dev)
    own = env.stats_rows
    env.hot_step()
    red.start(env)                   # env continues into the spare buffer; `own` is folded on the side stream
    assert env.stats_rows is not own
    env.hot_step()
    first = red.result()
    assert first["steps"] == 1024
    red.start(env)
    second = red.result()
    assert second["steps"] == 1024
    assert env.stats_rows is own or env.stats_rows is red.rows[0] or env.stats_rows is red.rows[1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _nccl_worker(rank, world, port, q, mode="nccl"):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import harness as H2
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    d = H2.pkg.distributed
    d.init_process_group(dev)
    red = d.StatsReducer(dev, mode=mode)
    assert red.mode == mode, (red.mode, getattr(red, "peer_error", None))
    red.warm()
    _make_envs(64, 1, 0, dev)[0].hot_step()     # load the step kernel outside any capture
    torch.cuda.synchronize()
    n, per = 1024 * (rank + 1), 2          # ranks own different env counts: the sum is not rank-symmetric
    envs_a, envs_b = _make_envs(n, per, 100 + rank, dev), _make_envs(n, per, 200 + rank, dev)
    ga = _horizon_graph(envs_a, red, 0, mode="thread_local")
    gb = _horizon_graph(envs_b, red, 1, mode="thread_local")
    torch.cuda.synchronize()
    for r in red.rows:
        r.zero_()
    ga.replay()
    torch.cuda.synchronize()
    local_a = _col_sums(red.rows[0])       # what this rank's horizon A accumulated
    gb.replay()                            # folds + all-reduces A
    torch.cuda.synchronize()
    got = red.reduced.double().cpu().numpy()
    for _ in range(50):                    # back-to-back replays: the collective never wedges the step stream
        ga.replay()
        gb.replay()
    torch.cuda.synchronize()
    q.put((rank, local_a, got, red.collectives))
    if mode == "peer":
        assert int(red.error.item()) == 0
    del ga, gb                             # ncclCommDestroy waits for the graphs that captured its collectives
    import gc
    gc.collect()
    torch.cuda.synchronize()
    red.close()
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode", ["peer", "nccl"])
def test_allreduce_inside_horizon_graph(mode):
    """2 GPUs: the statistics all-reduce captured as a branch of the horizon graph -- over NVLink peer mailboxes
    (ige_stats_exchange, the default) or NCCL -- equals the sum of the per-rank vectors."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, q, mode)) for r in range(world)]
    for p in procs:
        p.start()
    try:
        res = sorted((q.get(timeout=240) for _ in range(world)), key=lambda t: t[0])
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    total = sum(r[1] for r in res)
    assert total[abi.STAT_STEPS] == sum(1024 * (r + 1) * 2 for r in range(world))
    for rank, _, got, ncoll in res:
        np.testing.assert_allclose(got, total, rtol=1e-5, atol=1e-3, err_msg=f"rank {rank}")
        assert ncoll == 2                  # one collective captured per horizon graph


def _peer_worker_one_gpu(rank, world, port, q):
    """Both ranks on cuda:0 (gloo bootstrap): exercises the IPC mapping, the {epoch : float} protocol, the parity
    double-buffering and graph replays of ige_stats_exchange without needing a second GPU."""
    os.environ.update(RANK=str(rank), LOCAL_RANK="0", WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    import harness as H2
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    red = H2.pkg.distributed.StatsReducer(dev, mode="peer")
    assert red.mode == "peer", getattr(red, "peer_error", None)
    rows = red.rows[0]
    view = rows.view(abi.IGE_STATS_ROWS, abi.IGE_STATS_STRIDE)
    results = []
    for it in range(4):                                   # eager exchanges: both parities, twice
        view[:, :abi.IGE_NUM_STATS] = float(rank + 1) * (it + 1)
        red.issue(rows)
        torch.cuda.synchronize()
        results.append(red.reduced.cpu().numpy().copy())
        assert float(view[:, :abi.IGE_NUM_STATS].abs().sum()) == 0.0          # re-zeroed
    # captured: fill + exchange in one graph, replayed (the kernel counts its own epochs)
    fill = torch.full((abi.IGE_STATS_ROWS, abi.IGE_NUM_STATS), float(10 * (rank + 1)), device=dev)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
            view[:, :abi.IGE_NUM_STATS].copy_(fill)
            red.fork(rows)
            red.join()
    for _ in range(7):
        g.replay()
    torch.cuda.synchronize()
    results.append(red.reduced.cpu().numpy().copy())
    epoch, err = int(red.epoch.item()), int(red.error.item())
    # a rank that is late beyond timeout_ms: the others give up on it (zeros in its place) and check_error() says so
    import time
    red.timeout_ms = 300
    dist.barrier()
    if rank == 1:
        time.sleep(2.0)
    view[:, :abi.IGE_NUM_STATS] = 1.0
    red.issue(rows)
    torch.cuda.synchronize()
    late = None
    try:
        red.check_error()
    except RuntimeError as e:
        late = str(e)
    red.check_error()                                     # the flag was cleared by the raise
    q.put((rank, results, epoch, err, late, red.reduced.cpu().numpy().copy()))
    del g
    red.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_peer_exchange_processes_on_one_gpu(world):
    """world ranks on ONE GPU (gloo bootstrap): IPC mapping, {epoch : float} protocol, parity double-buffering, graph
    replays of ige_stats_exchange; the reduced vector is the sum over ranks and bit-identical on every rank."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_peer_worker_one_gpu, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    try:
        res = sorted((q.get(timeout=240) for _ in range(world)), key=lambda t: t[0])
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    R = abi.IGE_STATS_ROWS
    tri = world * (world + 1) / 2.0                     # sum over ranks of (rank + 1)
    for rank, results, epoch, err, late, late_vec in res:
        assert err == 0 and epoch == 1 + 4 + 7          # set-up probe + eager + graph replays
        if rank == 1:                                   # the late rank finds everybody's slots already filled
            assert late is None
            np.testing.assert_array_equal(late_vec, np.full(abi.IGE_NUM_STATS, R * world, np.float32))
        else:                                           # the others report it instead of continuing silently
            assert late is not None and "rank 1 did not push" in late, late
            np.testing.assert_array_equal(late_vec, np.full(abi.IGE_NUM_STATS, R * (world - 1), np.float32))
        for it in range(4):
            np.testing.assert_array_equal(results[it], np.full(abi.IGE_NUM_STATS, R * tri * (it + 1), np.float32))
        np.testing.assert_array_equal(results[4], np.full(abi.IGE_NUM_STATS, R * 10.0 * tri, np.float32))
        np.testing.assert_array_equal(results[4], res[0][1][4])     # identical bits on every rank
