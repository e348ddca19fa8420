"""Drive the rollout bookkeeping over a scripted horizon through the CPU oracle
(host pointers in the C struct) or the CUDA path (``RolloutBookkeeperMA`` -> C ABI); both return the
same dict of outputs as tests/golden/gen_golden_rollout.py records from the reference's play_steps."""
import glob
import os

import numpy as np
import torch

import harness as H

OUT_KEYS = ["out_mb_rewards", "out_mb_dones", "out_current_rewards", "out_current_shaped_rewards",
            "out_current_lengths", "out_done_rows", "out_done_count", "out_meter_mean", "out_meter_size",
            "out_mb_returns"]


def golden_files():
    return sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "rollout_*.npz")))


def _cfg(blob):
    return dict(A=int(blob["cfg_num_agents"]), Hz=int(blob["cfg_horizon"]), boot=int(blob["cfg_value_bootstrap"]),
                track=int(blob["cfg_games_to_track"]), gamma=float(blob["cfg_gamma"]), tau=float(blob["cfg_tau"]),
                scale=float(blob["cfg_scale"]), shift=float(blob["cfg_shift"]), lo=float(blob["cfg_min"]),
                hi=float(blob["cfg_max"]))


def oracle_rollout(blob):
    c = _cfg(blob)
    Hz, rows = blob["rewards"].shape
    abi = H.pkg._abi
    cur = [np.ascontiguousarray(blob[k]).astype(np.float32).copy()
           for k in ("init_current_rewards", "init_current_shaped_rewards", "init_current_lengths")]
    mb_rewards = np.zeros((Hz, rows), np.float32)
    mb_dones = np.zeros((Hz + 1, rows), np.uint8)
    mb_dones[0] = blob["init_dones"]
    done_rows = np.full((Hz, rows), -1, np.int32)
    done_count = np.zeros((Hz,), np.int32)
    meter_mean, meter_size = np.zeros((4,), np.float32), np.zeros((4,), np.int32)
    tr_mean, tr_size = np.zeros((Hz, 3), np.float32), np.zeros((Hz, 3), np.int32)
    for n in range(Hz):
        a = abi.IgeRolloutStepArgs()
        a.num_rows, a.num_agents, a.value_bootstrap, a.meter_max_size = rows, c["A"], c["boot"], c["track"]
        a.gamma, a.shaper_shift, a.shaper_scale, a.shaper_min, a.shaper_max = c["gamma"], c["shift"], c["scale"], c["lo"], c["hi"]
        rew, dn = np.ascontiguousarray(blob["rewards"][n]), np.ascontiguousarray(blob["dones"][n])
        to, val = np.ascontiguousarray(blob["time_outs"][n]), np.ascontiguousarray(blob["values"][n])
        ids, cnt = np.full((rows,), -1, np.int32), np.zeros((1,), np.int32)
        a.rewards, a.dones, a.time_outs, a.values = rew.ctypes.data, dn.ctypes.data, to.ctypes.data, val.ctypes.data
        a.mb_rewards, a.dones_out = mb_rewards[n].ctypes.data, mb_dones[n + 1].ctypes.data
        a.current_rewards, a.current_shaped_rewards, a.current_lengths = (x.ctypes.data for x in cur)
        a.done_rows, a.done_count = ids.ctypes.data, cnt.ctypes.data
        a.meter_mean, a.meter_size = meter_mean.ctypes.data, meter_size.ctypes.data
        H.oracle.call("ma_rollout_step", a)
        done_count[n] = cnt[0]
        done_rows[n, :cnt[0]] = ids[:cnt[0]]
        tr_mean[n], tr_size[n] = meter_mean[:3], meter_size[:3]
    d = abi.IgeDiscountArgs()
    d.num_rows, d.horizon, d.gamma, d.gamma_tau = rows, Hz, c["gamma"], c["gamma"] * c["tau"]
    last = np.ascontiguousarray(blob["last_values"])
    vals = np.ascontiguousarray(blob["values"])
    rets = np.zeros((Hz, rows), np.float32)
    d.dones, d.last_values, d.mb_dones = mb_dones[Hz].ctypes.data, last.ctypes.data, mb_dones.ctypes.data
    d.mb_values, d.mb_rewards, d.mb_returns = vals.ctypes.data, mb_rewards.ctypes.data, rets.ctypes.data
    H.oracle.call("ma_discount_values", d)
    return {"out_mb_rewards": mb_rewards, "out_mb_dones": mb_dones[:Hz], "out_current_rewards": cur[0],
            "out_current_shaped_rewards": cur[1], "out_current_lengths": cur[2], "out_done_rows": done_rows,
            "out_done_count": done_count, "out_meter_mean": tr_mean, "out_meter_size": tr_size,
            "out_mb_returns": rets}


def gpu_rollout(blob, device="cuda:0"):
    c = _cfg(blob)
    Hz, rows = blob["rewards"].shape
    bk = H.pkg.rollout.RolloutBookkeeperMA(rows, c["A"], Hz, device, gamma=c["gamma"], tau=c["tau"],
                                           value_bootstrap=bool(c["boot"]), games_to_track=c["track"],
                                           scale_value=c["scale"], shift_value=c["shift"], min_val=c["lo"], max_val=c["hi"])
    t = {k: torch.from_numpy(np.ascontiguousarray(blob[k])).to(device) for k in
         ("rewards", "dones", "time_outs", "values", "last_values")}
    bk.current_rewards.copy_(torch.from_numpy(blob["init_current_rewards"]).unsqueeze(1))
    bk.current_shaped_rewards.copy_(torch.from_numpy(blob["init_current_shaped_rewards"]).unsqueeze(1))
    bk.current_lengths.copy_(torch.from_numpy(blob["init_current_lengths"]))
    bk._mb_dones[0].copy_(torch.from_numpy(blob["init_dones"]))
    done_rows = np.full((Hz, rows), -1, np.int32)
    done_count = np.zeros((Hz,), np.int32)
    tr_mean, tr_size = np.zeros((Hz, 3), np.float32), np.zeros((Hz, 3), np.int32)
    bk.begin_rollout()
    for n in range(Hz):
        bk.after_env_step(t["rewards"][n], t["dones"][n],
                          time_outs=t["time_outs"][n] if c["boot"] else None, values=t["values"][n])
        ids = bk.env_done_indices().flatten().cpu().numpy()
        done_count[n] = ids.shape[0]
        done_rows[n, :ids.shape[0]] = ids
        tr_mean[n], tr_size[n] = bk.meter_mean[:3].cpu().numpy(), bk.meter_size[:3].cpu().numpy()
    rets = bk.discount_values(t["last_values"], t["values"])
    torch.cuda.synchronize()
    assert int(bk._scratch.abs().sum()) == 0 and float(bk._reduce.abs().sum()) == 0.0, "scratch not restored"
    return {"out_mb_rewards": bk.mb_rewards.squeeze(-1).cpu().numpy(), "out_mb_dones": bk.mb_dones.cpu().numpy(),
            "out_current_rewards": bk.current_rewards.squeeze(1).cpu().numpy(),
            "out_current_shaped_rewards": bk.current_shaped_rewards.squeeze(1).cpu().numpy(),
            "out_current_lengths": bk.current_lengths.cpu().numpy(), "out_done_rows": done_rows,
            "out_done_count": done_count, "out_meter_mean": tr_mean, "out_meter_size": tr_size,
            "out_mb_returns": rets.squeeze(-1).cpu().numpy()}


def compare(got, ref, what):
    """ids / counts / byte flags / meter sizes bit-exact; the element-wise floats are op-for-op restatements
    and must be bit-exact too; meter means (a mean over done rows, summation order free) within 1e-5."""
    for k in ("out_done_count", "out_done_rows", "out_mb_dones", "out_meter_size"):
        H.assert_equal(f"{what} {k}", got[k], ref[k])
    for k in ("out_mb_rewards", "out_current_rewards", "out_current_shaped_rewards", "out_current_lengths",
              "out_mb_returns"):
        H.assert_equal(f"{what} {k}", got[k].view(np.uint32), np.ascontiguousarray(ref[k]).view(np.uint32))
    scale = np.maximum(1.0, np.abs(ref["out_meter_mean"]))
    err = np.abs(got["out_meter_mean"].astype(np.float64) - ref["out_meter_mean"]) / scale
    assert err.max() <= H.RTOL, f"{what} meter means differ: {err.max()}"
