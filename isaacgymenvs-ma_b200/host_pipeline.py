"""Host-resident simulator (``sim_device=cpu pipeline=cpu``, vec_task.py:82-88) in front of the fused kernels:
a staged, chunk-pipelined step.

The reference's CPU pipeline keeps the simulator tensors in host memory; the fused post-physics kernels run on
the GPU, so every step moves the post-simulate state + actions host -> device and obs / rew / reset device -> host.
Done naively (copy everything in, one launch, copy everything out) the two PCIe directions never overlap.
``HostStagedPipeline`` splits the environments into G independent *env groups* (contiguous env-index ranges --
the same partitioning as the multi-GPU sharding), gives every group

  * ONE pinned host arena for its inputs and ONE for its results (the host simulator / policy write and read
    views of them: ``host_in[g][name]``, ``host_out[g][name]``); tensors that reset_idx rewrites in place
    (``HOST_INOUT_ATTRS``: dof_state) are also copied back into their ``host_in`` view, which is what
    ``gym.set_dof_state_tensor_indexed`` does for the reference's simulator,
  * the mirror arenas in HBM (the task mirror's tensors are re-homed as views of them), and
  * its own stream,

and captures  H2D(arena) -> fused kernel -> D2H(arena)  of all groups as the parallel branches of ONE CUDA graph:
group g's results travel device -> host while group g+1's inputs travel host -> device (full-duplex link, both
copy engines busy), and a step costs one graph launch on the host, not 2 x tensors x groups API calls.

There is no CPU compute path here either: the host only owns the buffers.
"""
from __future__ import annotations

import torch

from . import factory, synthetic

_ALIGN = 256


def _layout(tensors):
    """Byte offsets (256-B aligned) of `tensors` packed into one arena.  -> (offsets, total_bytes)"""
    offs, cur = [], 0
    for t in tensors:
        offs.append(cur)
        cur += -(-t.numel() * t.element_size() // _ALIGN) * _ALIGN
    return offs, max(cur, _ALIGN)


def _views(arena, tensors, offs):
    out = []
    for t, o in zip(tensors, offs):
        nb = t.numel() * t.element_size()
        out.append(arena[o:o + nb].view(t.dtype).view(t.shape))
    return out


class HostStagedPipeline:
    def __init__(self, task, num_envs, device, groups=4, seed=0, reset_prob=0.02, states=None, **env_kw):
        if num_envs % groups:
            raise ValueError(f"num_envs={num_envs} is not a multiple of groups={groups}")
        self.task, self.num_envs, self.groups = task, num_envs, groups
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("HostStagedPipeline needs a CUDA device: the post-physics kernels have no CPU path")
        per = num_envs // groups
        syn = factory._SYN_NAME[task]
        env_kw.setdefault("rand_mode", "philox")
        self.envs, self.streams = [], []
        self.host_in, self.host_out = [], []          # per group: {name: pinned host view}
        self._h_in, self._d_in, self._h_out, self._d_out = [], [], [], []
        for g in range(groups):
            st = states[g] if states is not None else synthetic.make_state(syn, per, seed=seed + g, reset_prob=reset_prob)
            env = factory.from_state(task, st, self.device, seed=seed + g, **env_kw)
            env.sim.record = False
            in_names = [n for n in env.HOST_SIM_ATTRS if getattr(env, n, None) is not None]
            out_names = [n for n in env.HOST_OUT_ATTRS if getattr(env, n, None) is not None]
            if not in_names:
                raise NotImplementedError(f"{task}: the mirror does not declare HOST_SIM_ATTRS")
            for names, h_list, d_list, views in ((in_names, self._h_in, self._d_in, self.host_in),
                                                 (out_names, self._h_out, self._d_out, self.host_out)):
                olds = [getattr(env, n) for n in names]
                offs, total = _layout(olds)
                d_arena = torch.zeros(total, dtype=torch.uint8, device=self.device)
                h_arena = torch.zeros(total, dtype=torch.uint8, pin_memory=True)
                by_id = {id(o): v for o, v in zip(olds, _views(d_arena, olds, offs))}
                env._rehome(names, lambda t, _m=by_id: _m[id(t)])     # the mirror's tensors become arena views
                hv = dict(zip(names, _views(h_arena, olds, offs)))
                for n in names:
                    hv[n].copy_(getattr(env, n))
                h_list.append(h_arena)
                d_list.append(d_arena)
                views.append(hv)
            self.envs.append(env)
            self.streams.append(torch.cuda.Stream(self.device))
        torch.cuda.synchronize(self.device)
        self.h2d_bytes_per_step = sum(a.numel() for a in self._h_in)
        self.d2h_bytes_per_step = sum(a.numel() for a in self._h_out) + sum(
            h[n].numel() * h[n].element_size() for e, h in zip(self.envs, self.host_in) for n in e.HOST_INOUT_ATTRS)
        self._graph = None

    # -- one step, eager (also the body that gets captured) ------------------------------------------
    def _issue(self, main):
        for g, env in enumerate(self.envs):
            s = self.streams[g]
            s.wait_stream(main)
            with torch.cuda.stream(s):
                self._d_in[g].copy_(self._h_in[g], non_blocking=True)
                env.hot_step()
                self._h_out[g].copy_(self._d_out[g], non_blocking=True)
                for n in env.HOST_INOUT_ATTRS:     # rows reset_idx rewrote go back to the simulator's tensor
                    self.host_in[g][n].copy_(getattr(env, n), non_blocking=True)
        for s in self.streams:
            main.wait_stream(s)

    def capture(self):
        """Capture the G-branch step into one CUDA graph (after one eager warm-up step)."""
        main = torch.cuda.Stream(self.device)
        main.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(main):
            self._issue(main)
        torch.cuda.current_stream(self.device).wait_stream(main)
        torch.cuda.synchronize(self.device)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph, stream=main):
            self._issue(main)
        return self

    def step(self):
        """H2D of every group's inputs, the fused kernels, D2H of obs / rew / reset: asynchronous; call
        ``synchronize()`` (or record an event) before the host reads ``host_out``."""
        if self._graph is not None:
            self._graph.replay()
        else:
            self._issue(torch.cuda.current_stream(self.device))

    def synchronize(self):
        torch.cuda.current_stream(self.device).synchronize()
