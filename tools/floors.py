"""Launch / copy floors on this GPU for calibrating small-N kernel times (graph replay, cold L2)."""
import torch
dev = "cuda:0"
def timeit(fn, n_inner, loops=50):
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s): fn()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    with torch.cuda.graph(g): fn()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(loops): g.replay()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (loops * n_inner)
R = 32
tiny = [torch.zeros(32, device=dev) for _ in range(R)]
print("tiny kernel (add_ on 32 floats) us/launch:", timeit(lambda: [t.add_(1.0) for t in tiny], R))
for mb in (4.5, 18, 72, 291):
    n = int(mb * 1e6 / 4)
    reps = max(2, min(R, int(300e6 / (mb * 2e6)) + 1))
    src = [torch.randn(n, device=dev) for _ in range(reps)]; dst = [torch.empty(n, device=dev) for _ in range(reps)]
    t = timeit(lambda: [d.copy_(s_) for d, s_ in zip(dst, src)], reps, loops=20)
    print(f"copy {mb} MB in + {mb} MB out: {t:.2f} us/launch -> {2*mb*1e6/t/1e3:.0f} GB/s")
    del src, dst
