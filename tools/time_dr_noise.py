"""Times ige_dr_noise on the device: CUDA events around a graph of launches that rotate over more
buffer sets than fit in L2.  Prints one JSON line per size: us per launch, algorithmic GB/s (20 B/element: x, corr,
draw read; y, y_clamped written) and the fraction of MEASURED_PEAKS.json's HBM figure."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import harness as H  # noqa: E402

abi = H.pkg._abi
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
dev = "cuda:0"
for rows, cols, mode in ((16384, 60, "injected"), (16384, 60, "philox"), (32768, 108, "injected"), (1 << 20, 60, "injected"),
                         (1 << 20, 60, "philox")):
    n = rows * cols
    sets = max(4, int(2 * 126e6 // (n * 20)) + 1)
    bufs = [[torch.randn(n, device=dev) for _ in range(5)] for _ in range(sets)]
    args = []
    for x, corr, draw, y, yc in bufs:
        a = abi.IgeNoiseArgs()
        a.count, a.operation, a.distribution = n, abi.IGE_NOISE_ADDITIVE, abi.IGE_NOISE_GAUSSIAN
        a.rand_mode = abi.IGE_RAND_PHILOX if mode == "philox" else abi.IGE_RAND_INJECTED
        a.corr_scale, a.corr_shift, a.scale, a.shift, a.clip, a.seed = 0.05, 0.01, 0.3, 0.02, 5.0, 3
        a.x, a.y, a.y_clamped, a.corr, a.draw = x.data_ptr(), y.data_ptr(), yc.data_ptr(), corr.data_ptr(), draw.data_ptr()
        args.append(a)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for a in args:
            H.pkg._lib.call("ige_dr_noise", a, stream=s.cuda_stream)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for a in args:
                H.pkg._lib.call("ige_dr_noise", a, stream=torch.cuda.current_stream().cuda_stream)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        reps = max(1, 2048 // sets)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * sets)
    bytes_per = 20 if mode == "injected" else 16           # philox: no draw tensor is read
    gbs = n * bytes_per / us / 1e3
    print(json.dumps({"kernel": "dr_noise_kernel", "rows": rows, "cols": cols, "mode": mode, "us": round(us, 3),
                      "algorithmic_bytes_per_element": bytes_per, "GBps": round(gbs, 1), "frac_of_peak": round(gbs / peak, 4),
                      "peak_GBps": peak, "buffer_sets": sets}))
