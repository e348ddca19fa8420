"""ige_compact_resets_static: reset compaction for hosts that cannot give the library scratch memory.

Some users of this library run inside a process whose GPU memory is already spoken for: the simulator
reserved its heap when it started, the trainer's caching allocator holds the rest under a hard cap, or the
binding is written in a language that has no CUDA runtime to call.  Such a host owns the tensors it was
handed (reset flags in, id list out) and nothing else, so it cannot cudaMalloc the look-back words that
ige_compact_resets wants from its caller.  The `_static` entry point keeps those words in a small
__device__ array inside libige_b200.so instead (limit: IGE_COMPACT_STATIC_MAX_ENVS envs; all calls on a
device share the array, so overlapping calls race).

What is checked here, through the C ABI, against the ordinary entry point on the same flags:
  * two calls in a row (and a large call followed by a small one) give identical ids / ranks / counts --
    integer work, so the bar is bit-exact, as in test_compact_resets;
  * a size beyond the built-in array is refused with IGE_ESIZE and a message, nothing is launched.

This file sorts after every other test file on purpose: `pytest -x` stops at the first failure, and a
failure here must not hide the parity tests.  It makes no assumption about what ran before it (it passes
alone in a fresh process and at the end of the whole suite).
"""
import numpy as np
import pytest
import torch

import harness as H

pytestmark = pytest.mark.gpu

MAX_ENVS = 4194304   # IGE_COMPACT_STATIC_MAX_ENVS in include/ige_b200.h


def _args(reset, group, ids, rank, count, scratch=None, num_rows=None):
    a = H.pkg._abi.IgeCompactArgs()
    a.num_rows, a.group = (reset.numel() if num_rows is None else num_rows), group
    a.reset_buf, a.reset_ids, a.reset_rank = reset.data_ptr(), ids.data_ptr(), rank.data_ptr()
    a.count = count.data_ptr()
    a.scratch = scratch.data_ptr() if scratch is not None else 0
    return a


def _run(entry, flags, group, with_scratch):
    """One call of `entry` on `flags`; returns (ids[:count], rank, count) as numpy."""
    dev = "cuda:0"
    # an empty env set still needs non-NULL pointers: lend it a one-element buffer
    reset = torch.from_numpy(flags).to(dev) if flags.size else torch.zeros(1, dtype=torch.int64, device=dev)
    ne = flags.size // group
    ids = torch.full((max(ne, 1),), -7, dtype=torch.int32, device=dev)
    rank = torch.full((max(ne, 1),), -7, dtype=torch.int32, device=dev)
    count = torch.full((1,), -7, dtype=torch.int32, device=dev)
    scratch = None
    if with_scratch:   # the ordinary entry point: the CALLER allocates and zeroes the scratch words
        scratch = torch.zeros(H.pkg._lib.lib().ige_compact_scratch_words(ne), dtype=torch.int32, device=dev)
    a = _args(reset, group, ids, rank, count, scratch, num_rows=flags.size)
    H.pkg._lib.call(entry, a, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    k = int(count.item())
    return ids[:k].cpu().numpy(), rank[:ne].cpu().numpy(), k


def test_header_limit_matches():
    hdr = open(H.ROOT + "/include/ige_b200.h").read()
    assert f"#define IGE_COMPACT_STATIC_MAX_ENVS {MAX_ENVS}\n" in hdr


@pytest.mark.parametrize("n,group,prob", [(16384, 1, 0.02), (8192 * 2, 2, 0.7), (1 << 20, 1, 0.3), (MAX_ENVS, 1, 0.05),
                                          (2049, 1, 1.0), (0, 1, 0.5)])
def test_static_scratch_matches_caller_scratch_twice(n, group, prob):
    """Same flags through both entry points, two rounds back to back with fresh flags each round, then a
    small call right after the large one (the shared array has just been used by more chunks)."""
    rng = np.random.Generator(np.random.PCG64(1000 + n + group))
    with torch.cuda.device(0):
        for call in range(2):
            flags = (rng.random(n) < prob).astype(np.int64) * rng.integers(1, 5, n)
            want_ids, want_rank, want_k = _run("ige_compact_resets", flags, group, with_scratch=True)
            got_ids, got_rank, got_k = _run("ige_compact_resets_static", flags, group, with_scratch=False)
            assert got_k == want_k, f"call {call}: count {got_k} != {want_k}"
            np.testing.assert_array_equal(got_ids, want_ids, err_msg=f"call {call}: ids")
            np.testing.assert_array_equal(got_rank, want_rank, err_msg=f"call {call}: ranks")
            # and both agree with numpy's nonzero() of the AND over the agent axis
            env_flag = (flags.reshape(n // group, group) != 0).all(axis=1)
            np.testing.assert_array_equal(got_ids, np.flatnonzero(env_flag).astype(np.int32))
        small = (rng.random(300) < 0.5).astype(np.int64)
        a = _run("ige_compact_resets", small, 1, with_scratch=True)
        b = _run("ige_compact_resets_static", small, 1, with_scratch=False)
        assert a[2] == b[2]
        np.testing.assert_array_equal(a[0], b[0])
        np.testing.assert_array_equal(a[1], b[1])


def test_static_scratch_back_to_back_without_sync():
    """Two calls queued on one stream with no host synchronisation between them: stream order is enough."""
    rng = np.random.Generator(np.random.PCG64(77))
    dev = "cuda:0"
    n = 100000
    with torch.cuda.device(0):
        stream = torch.cuda.current_stream().cuda_stream
        keep, outs, wants = [], [], []
        for call in range(2):
            flags = (rng.random(n) < 0.25).astype(np.int64)
            reset = torch.from_numpy(flags).to(dev)
            ids = torch.full((n,), -7, dtype=torch.int32, device=dev)
            rank = torch.full((n,), -7, dtype=torch.int32, device=dev)
            count = torch.zeros(1, dtype=torch.int32, device=dev)
            keep.append(reset)
            outs.append((ids, rank, count))
            wants.append(flags)
        for (ids, rank, count), reset in zip(outs, keep):
            H.pkg._lib.call("ige_compact_resets_static", _args(reset, 1, ids, rank, count), stream=stream)
        torch.cuda.synchronize()
        for (ids, rank, count), flags in zip(outs, wants):
            want = np.flatnonzero(flags).astype(np.int32)
            assert int(count.item()) == want.size
            np.testing.assert_array_equal(ids[:want.size].cpu().numpy(), want)
            np.testing.assert_array_equal(rank.cpu().numpy(), np.cumsum(flags) - flags)


def test_static_scratch_rejects_oversize():
    """One env more than the built-in array covers: IGE_ESIZE, a message naming the limit, no launch."""
    with torch.cuda.device(0):
        t = torch.zeros(8, dtype=torch.int64, device="cuda:0")
        o = torch.zeros(8, dtype=torch.int32, device="cuda:0")
        a = H.pkg._abi.IgeCompactArgs()
        a.num_rows, a.group = MAX_ENVS + 1, 1
        a.reset_buf, a.reset_ids, a.reset_rank, a.count = t.data_ptr(), o.data_ptr(), o.data_ptr(), o.data_ptr()
        before = H.pkg._lib.launch_count()
        with pytest.raises(H.pkg._lib.IgeError) as ei:
            H.pkg._lib.call("ige_compact_resets_static", a, stream=torch.cuda.current_stream().cuda_stream)
        assert "rc=-3" in str(ei.value) and "IGE_COMPACT_STATIC_MAX_ENVS" in str(ei.value)
        assert H.pkg._lib.launch_count() == before
        torch.cuda.synchronize()
