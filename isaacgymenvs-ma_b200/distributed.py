"""Multi-GPU plumbing for the hot path: one process per GPU, environments sharded by env index,
NO collective in the step itself; the only exchange is a sum all-reduce of the <= 16-float
episode-statistics vector once per horizon.

Reference behaviour kept: per-rank device from LOCAL_RANK (utils/rlgames_utils.py:89-107), per-rank
seed = seed + rank (utils/utils.py:87-94); every rank owns an independent env shard.
"""
import os

import torch

from . import _abi

STAT_NAMES = ["steps", "reward_sum", "done", "timeout", "done_length_sum"] + \
             [f"task{i}" for i in range(_abi.IGE_NUM_STATS - 5)]


def env_info():
    """(rank, local_rank, world_size) from the torchrun environment (defaults: single process)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_process_group(device=None):
    """NCCL over NVLink for CUDA ranks, gloo for CPU ranks (tests).  No-op for world_size 1."""
    import torch.distributed as dist
    rank, local_rank, world = env_info()
    if world == 1 or dist.is_initialized():
        return rank, local_rank, world
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    use_cuda = device is not None and torch.device(device).type == "cuda"
    if use_cuda:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(device))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, local_rank, world


def env_shard(global_num_envs, rank, world):
    """Contiguous env-index range [start, start + count) owned by `rank` (agent rows env*A+a stay together)."""
    base, rem = divmod(int(global_num_envs), int(world))
    count = base + (1 if rank < rem else 0)
    start = rank * base + min(rank, rem)
    return start, count


def bind_to_gpu_numa_node(device_index):
    """Restrict this process to the CPUs of the NUMA node its GPU hangs off, BEFORE any pinned host memory is
    allocated: with pipeline=cpu every step moves megabytes between pinned arenas and the GPU, and with 8 ranks on a
    two-socket box arenas that land on the far socket share the inter-socket link (measured in round 1: per-GPU
    end-to-end throughput halved at 8 ranks).  Pages are placed on the node of the allocating thread (first touch
    happens inside cudaHostAlloc), so binding the process is enough.  Returns the node, or None when the platform
    does not say (virtualised PCI topology) -- then nothing is changed."""
    try:
        import ctypes
        from . import _lib
        buf = ctypes.create_string_buffer(32)
        if _lib.lib().ige_device_pci_bus_id(int(device_index), buf, 32) != 0:
            return None
        bdf = buf.value.decode()
        bdf = str(bdf).lower()
        if len(bdf.split(":")) == 2:
            bdf = "0000:" + bdf
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return node
    except (OSError, ValueError, AttributeError, RuntimeError):
        return None


def rank_seed(seed, rank):
    return int(seed) + int(rank)          # utils/utils.py:87-94


class StatsReducer:
    """Sum all-reduce of the episode-statistics vector, once per horizon, off the step's critical path.

    CUDA (the product path).  The accumulator the step kernels add into is DOUBLE-BUFFERED: horizon h adds
    into ``rows[h & 1]`` while the previous horizon's ``rows[h & 1 ^ 1]`` is folded into the 16-float vector
    (``ige_stats_snapshot``: column sums + re-zero, one tiny launch) and all-reduced over NCCL / NVLink.  Both
    are issued by :meth:`fork` on a side stream and re-joined by :meth:`join`; called while a CUDA graph is being
    captured they become a forked BRANCH of the horizon's graph, so a horizon costs the host exactly one graph
    launch -- no Python, no ProcessGroup enqueue and no stream join between two step kernels (the eager
    ``dist.all_reduce`` per horizon of round 1 held the host thread that feeds the step graphs for ~180 us:
    49 % weak-scaling efficiency at the driver's 20-step window).  The collective overlaps the first steps
    of the horizon; the reduced numbers lag by one horizon, which is what an asynchronous reduction means.

    The collective itself (``mode``):
      * ``"peer"`` (default for world > 1 on CUDA): ``ige_stats_exchange`` -- ONE single-CTA kernel that folds the
        rows, stores the vector into every rank's mailbox over NVLink peer memory (cudaIpc-mapped, 64-bit
        {epoch : float} words) and sums the local mailbox in rank order.  An NCCL kernel in the branch held SM
        resources next to the resident-capacity step grid for its whole wait (measured: 0.90 weak-scaling efficiency
        at the driver's 20-step window); this kernel is 512 threads polling local memory.  The process group (NCCL)
        is only the bootstrap: one all-gather of the 64-byte IPC handles and a barrier.
      * ``"nccl"`` (``IGE_STATS_REDUCE=nccl``, or when peer mapping is unavailable): snapshot kernel +
        ``dist.all_reduce`` captured in the branch.

    A rank that does not push within ``timeout_ms`` is NOT waited for any longer: the kernel substitutes zeros and
    records ``rank + 1`` in ``self.error`` -- unlike NCCL, which would block.  The default is therefore long
    (``IGE_STATS_TIMEOUT_MS``, 10 minutes: a checkpoint save or an evaluation pause on one rank must not drop its
    horizon), and :meth:`check_error` / :meth:`result` raise when it happened, so ranks never continue silently with
    different vectors.  Tests pass a short ``timeout_ms``.

    CPU (gloo, tests of the host logic): :meth:`start` with the 16-float vector, eager.
    """

    def __init__(self, device, mode=None, timeout_ms=None):
        self.timeout_ms = int(timeout_ms if timeout_ms is not None else os.environ.get("IGE_STATS_TIMEOUT_MS", 600000))
        if not 0 < self.timeout_ms < 2 ** 32:
            raise ValueError(f"timeout_ms={self.timeout_ms} outside (0, 2^32)")
        self.error = None
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.mode = "local"
        self._mailboxes, self._own_mailbox = [], None
        self.reduced = torch.zeros(_abi.IGE_NUM_STATS, dtype=torch.float32, device=device)
        self.snap = self.reduced          # round-1 name
        n = _abi.IGE_STATS_ROWS * _abi.IGE_STATS_STRIDE
        self.rows = [torch.zeros(n, dtype=torch.float32, device=device) for _ in range(2)] if self.cuda else None
        self.side = torch.cuda.Stream(self.device) if self.cuda else None
        self._free = torch.cuda.Event() if self.cuda else None     # the folded buffer is zeroed and reusable
        self._work = None
        self._cur = 0
        self.collectives = 0
        if self.cuda:      # load the snapshot kernel now: its first launch must not happen inside a graph capture
            with torch.cuda.device(self.device):
                self._snapshot(self.rows[0])
                torch.cuda.synchronize(self.device)
            if self._multi():
                want = mode or os.environ.get("IGE_STATS_REDUCE", "peer")
                self.mode = "nccl"
                if want == "peer":
                    self._setup_peer()

    # -- peer-memory mailboxes (ige_stats_exchange) ----------------------------------------------------------
    def _setup_peer(self):
        """Allocate this rank's mailbox, all-gather the IPC handles, map every peer's mailbox.  Falls back to the
        NCCL collective (with a note in ``self.peer_error``) if any rank cannot map any peer."""
        import ctypes as C
        import torch.distributed as dist
        from . import _lib
        lib = _lib.lib()
        rank, world = dist.get_rank(), dist.get_world_size()
        self.peer_error = None
        own, handle = C.c_void_p(), C.create_string_buffer(_abi.IGE_IPC_HANDLE_BYTES)
        ok = world <= _abi.IGE_MAX_RANKS
        with torch.cuda.device(self.device):
            if ok and lib.ige_mailbox_create(world, C.byref(own), handle) != 0:
                ok, self.peer_error = False, lib.ige_last_error_string().decode()
            handles = [None] * world
            dist.all_gather_object(handles, (bool(ok), handle.raw))
            ptrs = [None] * world
            if all(h[0] for h in handles):
                for r, (_, raw) in enumerate(handles):
                    if r == rank:
                        ptrs[r] = own.value
                        continue
                    p = C.c_void_p()
                    if lib.ige_mailbox_open(raw, C.byref(p)) != 0:
                        ok, self.peer_error = False, lib.ige_last_error_string().decode()
                        break
                    ptrs[r] = p.value
            else:
                ok = False
            flags = [None] * world
            dist.all_gather_object(flags, bool(ok))
            if all(flags):
                self._own_mailbox, self._mailboxes = own.value, ptrs
                self.epoch = torch.zeros(1, dtype=torch.int64, device=self.device)
                self.error = torch.zeros(1, dtype=torch.int32, device=self.device)
                self.local = torch.zeros(_abi.IGE_NUM_STATS, dtype=torch.float32, device=self.device)
                self.rank, self.world = rank, world
                self.mode = "peer"
                self._exchange(self.rows[0], timeout_ms=min(self.timeout_ms, 30000))   # first launch outside any capture; also checks the mapping
                torch.cuda.synchronize(self.device)
                missing = int(self.error.item())
                self.error.zero_()
                answers = [None] * world
                dist.all_gather_object(answers, missing)     # every rank learns of a failure: all raise, nobody is left in the barrier
                if any(answers):
                    bad = sorted({m - 1 for m in answers if m})
                    raise RuntimeError(f"ige_stats_exchange: rank(s) {bad} never answered the first exchange "
                                       f"(seen by rank(s) {[r for r, m in enumerate(answers) if m]})")
                self.reduced.zero_()
            else:
                for r, p in enumerate(ptrs):
                    if p and r != rank:
                        lib.ige_mailbox_close(p)
                if own.value:
                    lib.ige_mailbox_destroy(own)
            dist.barrier()

    def _exchange(self, rows, timeout_ms=None):
        from . import _lib
        a = _abi.IgeStatsExchangeArgs()
        a.stats, a.out, a.local = rows.data_ptr(), self.reduced.data_ptr(), self.local.data_ptr()
        for r, p in enumerate(self._mailboxes):
            a.mailbox[r] = p
        a.epoch, a.error = self.epoch.data_ptr(), self.error.data_ptr()
        a.rank, a.world, a.zero_stats, a.timeout_ms = self.rank, self.world, 1, int(timeout_ms or self.timeout_ms)
        _lib.call("ige_stats_exchange", a, stream=torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        """Unmap the peers' mailboxes and free our own (after a barrier: nobody may still be storing into it)."""
        if self.mode != "peer" or self._own_mailbox is None:
            return
        import torch.distributed as dist
        from . import _lib
        lib = _lib.lib()
        torch.cuda.synchronize(self.device)
        dist.barrier()
        for r, p in enumerate(self._mailboxes):
            if r != self.rank:
                lib.ige_mailbox_close(p)
        lib.ige_mailbox_destroy(self._own_mailbox)
        self._own_mailbox, self._mailboxes, self.mode = None, [], "nccl"

    @staticmethod
    def _multi():
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1

    # -- CUDA: graph-capturable pieces ---------------------------------------------------------------------
    def attach(self, env, parity):
        """Make `env`'s step kernels accumulate into rows[parity]."""
        self._require_accumulator(env)
        env.stats_rows = self.rows[parity & 1]

    @staticmethod
    def _require_accumulator(env):
        """Tasks whose C ABI has no `stats` pointer (Anymal, BallBalance, Quadcopter, Ingenuity, the single-agent Franka
        tasks) never write the accumulator: reducing it would ship zeros that look like statistics."""
        if not getattr(env, "ACCUMULATES_STATS", True) or getattr(env, "stats_rows", 0) is None:
            raise ValueError(f"{type(env).__name__} has no episode-statistics accumulator (its step kernels take no "
                             "`stats` pointer, or it was built with accumulate_stats=False): nothing to reduce")

    def warm(self):
        """One eager collective so the NCCL communicator exists before any graph capture."""
        if self._multi() and self.mode != "peer":
            import torch.distributed as dist
            dist.all_reduce(self.reduced, op=dist.ReduceOp.SUM)
            torch.cuda.synchronize(self.device)
            self.reduced.zero_()

    def _snapshot(self, rows):
        from . import _lib
        a = _abi.IgeStatsSnapshotArgs()
        a.stats, a.out, a.zero_stats, a.accumulate = rows.data_ptr(), self.reduced.data_ptr(), 1, 0
        _lib.call("ige_stats_snapshot", a, stream=torch.cuda.current_stream(self.device).cuda_stream)

    def issue(self, rows):
        """Fold `rows` into the vector (re-zeroing them) and all-reduce it, on the CURRENT stream."""
        if self.mode == "peer":
            self._exchange(rows)
            if not torch.cuda.is_current_stream_capturing():
                self._free.record(torch.cuda.current_stream(self.device))
            self.collectives += 1
            return
        self._snapshot(rows)
        if not torch.cuda.is_current_stream_capturing():
            self._free.record(torch.cuda.current_stream(self.device))
        if self._multi():
            import torch.distributed as dist
            dist.all_reduce(self.reduced, op=dist.ReduceOp.SUM)     # NCCL, capturable; in place
            self.collectives += 1

    def fork(self, rows):
        """Side branch: snapshot + all-reduce of `rows` (the PREVIOUS horizon's buffer).  Inside
        ``torch.cuda.graph`` capture this forks the graph; pair with :meth:`join` before the capture ends."""
        cur = torch.cuda.current_stream(self.device)
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side):
            self.issue(rows)

    def join(self):
        torch.cuda.current_stream(self.device).wait_stream(self.side)

    # -- eager use -----------------------------------------------------------------------------------------
    def start(self, stats):
        """End of a horizon, eager.  `stats` is an env (CUDA: its accumulator is swapped for the spare buffer
        and the full one is reduced on the side stream; the step stream only waits for the spare to be zeroed,
        never for the collective) or the 16-float vector / a callable returning it (CPU, gloo)."""
        import torch.distributed as dist
        if self.cuda and hasattr(stats, "stats_rows"):
            env = stats
            self._require_accumulator(env)
            full = env.stats_rows
            if full is self.rows[0] or full is self.rows[1]:
                self._cur = 0 if full is self.rows[0] else 1
            else:                                  # first horizon of an env that owns its accumulator
                self.rows[self._cur] = full
            torch.cuda.current_stream(self.device).wait_event(self._free)   # spare buffer: zeroed long ago
            env.stats_rows = self.rows[self._cur ^ 1]
            self.fork(full)
            return
        vec = stats() if callable(stats) else stats
        if self.cuda:
            self.side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.side):
                self.reduced.copy_(vec)
                if self._multi():
                    dist.all_reduce(self.reduced, op=dist.ReduceOp.SUM)
            return
        self.reduced.copy_(vec)
        if self._multi():
            self._work = dist.all_reduce(self.reduced, op=dist.ReduceOp.SUM, async_op=True)

    def wait(self):
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self.side is not None:
            self.join()
        return self.reduced

    def check_error(self):
        """Peer mode: raise if an exchange since the last check gave up on a rank (its statistics were replaced by
        zeros and the ranks' vectors may differ).  Synchronises with the device; clears the flag."""
        if self.mode != "peer" or self.error is None:
            return
        missing = int(self.error.item())
        if missing:
            self.error.zero_()
            raise RuntimeError(f"ige_stats_exchange: rank {missing - 1} did not push its statistics within "
                               f"{self.timeout_ms} ms; this horizon's reduced vector is incomplete on rank {self.rank}")

    def result(self):
        v = self.wait().detach().cpu().tolist()
        self.check_error()
        out = dict(zip(STAT_NAMES, v))
        steps, done = max(out["steps"], 1.0), max(out["done"], 1.0)
        out["mean_reward"] = out["reward_sum"] / steps
        out["mean_episode_length"] = out["done_length_sum"] / done
        return out
