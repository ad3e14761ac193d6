#!/usr/bin/env python3
"""Where the HOST time of a multi-view step goes (cProfile over step_resident), per workload.
Usage: python tools/host_profile.py <workload> <streams> [steps]"""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
wl_name, k = sys.argv[1], int(sys.argv[2])
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = torch.device("cuda", 0)
wl = bench.Workload(wl_name, dev, 0, 1, 4)
if k > 1:
    from soar_b200.views import view_streams
    wl.streams = view_streams(min(k, wl.vpr), dev)
arm = bench.SoarArm(wl)
for s in range(3):
    bench.step_resident(arm, wl, s)
torch.cuda.synchronize()
t0 = time.perf_counter()
for s in range(steps):
    bench.step_resident(arm, wl, 3 + s)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"{wl_name} streams={k}: host enqueue {1e3*(t1-t0)/steps:.3f} ms/step, +drain {1e3*(t2-t1)/steps:.3f} ms/step, {wl.vpr} views/step")
pr = cProfile.Profile()
pr.enable()
for s in range(steps):
    bench.step_resident(arm, wl, 100 + s)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
