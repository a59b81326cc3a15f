"""Host-side cost of one decode step of bench.py's workload (python + ctypes), measured without device syncs
inside the loop, next to the device time of the same steps; plus a cProfile of the step."""
import argparse, cProfile, pstats, sys, time, io
sys.path.insert(0, '/root/repo')
import torch
import bench
sys.argv = [sys.argv[0]]
args = bench.parse_args()
dev = torch.device('cuda', 0)
arm = bench.Arm(args, dev, 0, 1, host_io=False)
arm.prefill()
arm.make_runner()
for _ in range(8):
    arm.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
N = 24
# host cost of a step with an idle GPU queue (sync before every step): what the e2e leg exposes
hs = []
for _ in range(N):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); arm.step(); hs.append(time.perf_counter() - t0)
hs.sort()
print("host issue time per step, idle queue: p50 %.3f ms  min %.3f  max %.3f" % (hs[N // 2] * 1e3, hs[0] * 1e3, hs[-1] * 1e3))
torch.cuda.synchronize()
e0.record(); t0 = time.perf_counter()
for _ in range(N):
    arm.step()
t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize()
print("host ms/step %.3f  device-span ms/step %.3f" % ((t1 - t0) / N * 1e3, e0.elapsed_time(e1) / N))
pr = cProfile.Profile(); pr.enable()
for _ in range(8):
    arm.step()
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(28); print(s.getvalue()[:6000])
