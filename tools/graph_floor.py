import torch, time
x = torch.zeros(205*32, device='cuda', dtype=torch.int32)
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    for _ in range(10): x.add_(1)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(1152): x.add_(1)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(17): g.replay()
    e1.record(); torch.cuda.synchronize()
    print("dependent tiny kernel nodes in a graph: %.3f us per node" % (e0.elapsed_time(e1) * 1e3 / (17 * 1152)))
