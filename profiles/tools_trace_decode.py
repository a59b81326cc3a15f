import sys, json, ctypes, torch, numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_paged_compression_b200 import _lib, ops
from benchtools import microbench
lib = _lib.load()
lib.pc_debug_set_decode_trace.argtypes = [ctypes.c_void_p]
lib.pc_debug_set_decode_trace.restype = None
dev = torch.device('cuda', 0)
B, ctx = int(sys.argv[1]), int(sys.argv[2])
HQ, HKV = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (32, 8)
buf = torch.zeros(148 * 8 * 8, dtype=torch.int64, device=dev)
# warm without trace
print(microbench.time_shape(dev, B, ctx, HQ, HKV, reps=10)['us'])
lib.pc_debug_set_decode_trace(buf.data_ptr())
r = microbench.time_shape(dev, B, ctx, HQ, HKV, reps=1)
torch.cuda.synchronize()
lib.pc_debug_set_decode_trace(None)
raw = buf.cpu().numpy().reshape(-1, 8)
t = raw[:, :4].astype(np.float64)
act = t[:, 0] > 0
t = t[act]
raw = raw[act]
t0 = t[:, 0].min()
t = (t - t0) / 1e3
print("workers", len(t))
for i, n in enumerate(["entry", "sched_ready", "stream_done", "exit"]):
    print(f"{n:12s} min {t[:, i].min():7.2f} p50 {np.median(t[:, i]):7.2f} p90 {np.percentile(t[:, i], 90):7.2f} max {t[:, i].max():7.2f} us")
d = t[:, 2] - t[:, 1]
print("stream dur  min %.2f p50 %.2f max %.2f" % (d.min(), np.median(d), d.max()))
d = t[:, 3] - t[:, 2]
print("fixup dur   min %.2f p50 %.2f p90 %.2f max %.2f" % (d.min(), np.median(d), np.percentile(d, 90), d.max()))
# per-CTA (= per-SM) view: w = warp * grid + block
full = buf.cpu().numpy().reshape(-1, 8)[:, :4].astype(np.float64)
grid = 148
nw = full.shape[0] // grid
blk_done = np.zeros(grid)
blk_first = np.zeros(grid)
for bidx in range(grid):
    rows = [full[wp * grid + bidx] for wp in range(nw) if full[wp * grid + bidx][0] > 0]
    if rows:
        r = (np.array(rows) - t0) / 1e3
        blk_done[bidx] = r[:, 2].max()
        blk_first[bidx] = r[:, 2].min()
print("per-SM last-warp stream_done: min %.2f p50 %.2f max %.2f" % (blk_done.min(), np.median(blk_done), blk_done.max()))
print("per-SM first-warp stream_done: min %.2f p50 %.2f max %.2f" % (blk_first.min(), np.median(blk_first), blk_first.max()))
print("within-SM spread (last-first): p50 %.2f max %.2f" % (np.median(blk_done - blk_first), (blk_done - blk_first).max()))
print("sorted per-SM done:", np.round(np.sort(blk_done), 1).tolist()[::8])

ex = t[:, 3]
order = np.argsort(ex)
print("chunks/warp: min %d p50 %d max %d; total %d" % (raw[:, 4].min(), np.median(raw[:, 4]), raw[:, 4].max(), raw[:, 4].sum()))
print("slowest 12 warps: exit_us, stream_us, chunks, combines, smid, ticket_atomic_us")
for i in order[-12:]:
    print("  %.1f %.1f %d %d %d %.1f" % (t[i, 3], t[i, 2], raw[i, 4], raw[i, 5], raw[i, 6], raw[i, 7] / 1e3))
print("fastest 6 warps:")
for i in order[:6]:
    print("  %.1f %.1f %d %d %d %.1f" % (t[i, 3], t[i, 2], raw[i, 4], raw[i, 5], raw[i, 6], raw[i, 7] / 1e3))
print("ticket atomic total per warp us: p50 %.2f p90 %.2f max %.2f" % (np.median(raw[:, 7]) / 1e3, np.percentile(raw[:, 7], 90) / 1e3, raw[:, 7].max() / 1e3))
