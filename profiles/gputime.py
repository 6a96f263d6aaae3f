"""GPU-side timing helper: the calls are captured into one CUDA graph (rotating over distinct input buffers whose
total size exceeds L2) and the graph replay is timed with CUDA events, so Python / ctypes launch overhead (~20 us per
call, more than several of the kernels) is not part of the number."""
import torch


def graph_time_us(make_call, n_variants: int, reps: int = 5, inner: int = 1) -> float:
    """make_call(i) -> callable running the kernel on input variant i.  Returns mean microseconds per call."""
    calls = [make_call(i) for i in range(n_variants)]
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for c in calls:
            c()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(inner):
            for c in calls:
                c()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * inner * n_variants)
