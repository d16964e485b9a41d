"""One warm LBVH build + one refit over 10^7 triangles (C5 soup) inside a cudaProfilerStart/Stop range: the radix sort
(rs_hist, rs_scan, rs_scatter) and the bottom-up fit (k_fit) at the size of C4 / C5's largest point."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import cudaraytracer_b200 as crt
from cudaraytracer_b200 import scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
sc = scenes.c5_triangles(n)
ctx = crt.Context(0)
ctx.set_scene(sc)
ctx.build()
ctx.refit()
ctx.sync()
ctx.reset_counters()
torch.cuda.profiler.start()
ctx.build()
ctx.refit()
ctx.sync()
torch.cuda.profiler.stop()
c = ctx.counters()
print("n = %d: build %.3f ms, refit %.3f ms" % (n, c["build_ms"], c["refit_ms"]))
ctx.close()
