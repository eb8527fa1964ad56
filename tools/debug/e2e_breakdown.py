"""Where the end-to-end PageRank step (host buffers in, result out) spends its time: per-kernel device times + wall clock of the
import alone.  Run on the GPU box: python tools/debug/e2e_breakdown.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "graph-python_b200")]
import numpy as np, torch
import bench
import graphblas_b200 as gb
from graphblas_b200 import Matrix, Vector

gb.init("b200"); gb.use_torch_stream()
job = bench.PageRankJob(gb, 22, 0, 1)
ap, ai, ax, d = bench.host_arrays(gb, job)
host = [bench.pinned_copy(a) for a in (ap, ai, ax, d, np.full(d.size, 1.0 / d.size, dtype=np.float32))]
keep = [t for _, t in host]; host = [h for h, _ in host]
out_p, t = bench.pinned_copy(np.zeros(d.size, dtype=np.float32)); keep.append(t)
bench.e2e_pagerank(gb, host, out_p); torch.cuda.synchronize()
# pure copy floor
dst = torch.empty(ai.nbytes, dtype=torch.uint8, device="cuda")
src = torch.from_numpy(host[1].view(np.uint8))
torch.cuda.synchronize(); t0 = time.perf_counter(); dst.copy_(src, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"pinned H2D {ai.nbytes/1e6:.0f} MB: {dt*1e3:.2f} ms = {ai.nbytes/dt/1e9:.1f} GB/s")
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    A = Matrix.from_csc(host[0], host[1], host[2], nrows=d.size, ncols=d.size)
    gb.sync(); t1 = time.perf_counter()
    dv, r = Vector.from_dense(host[3]), Vector.from_dense(host[4]); gb.sync(); t2 = time.perf_counter()
    print(f"import matrix {1e3*(t1-t0):.2f} ms, vectors {1e3*(t2-t1):.2f} ms")
    del A
gb.lib.GxB_b200_kernel_time_reset(); gb.set_option("timing", 1)
torch.cuda.synchronize(); t0 = time.perf_counter()
bench.e2e_pagerank(gb, host, out_p); torch.cuda.synchronize()
print(f"e2e step {1e3*(time.perf_counter()-t0):.2f} ms")
gb.set_option("timing", 0); gb.lib.GxB_b200_kernel_time_dump()
