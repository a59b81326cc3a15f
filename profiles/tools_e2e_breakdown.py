"""Where does an e2e step (host-resident activations, bench.py `e2e`) spend its time?

Per step of the timed window: wall clock of runner.step(), host time before the graph launch,
device time of the replayed graph (events around the replay), host time after the launch call
returns (bookkeeping + the wait for the device).  Run for several PinnedLayerIO copy schedules.
    python profiles/tools_e2e_breakdown.py [--io-group 0,1,2,4] [--steps 24]
"""
import argparse
import json
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402


def one(io_group, K, W, extra, after_value=False):
    sys.argv = ["bench.py", "--steps", str(K), "--warmup", str(W), "--io-group", str(io_group)] + extra
    args = bench.parse_args()
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    spec = bench.main_spec(args, 1)
    P = bench.preroll_steps(spec, K, W, True)
    if after_value:      # what bench.py does before its e2e leg: the resident-input arm, then free it
        arm = bench.Arm(spec, args, device, 0, 1, host_io=False, total_steps=P + W + K)
        arm.prefill()
        arm.make_runner(P + W + K)
        for _ in range(P + W + K):
            arm.step()
        torch.cuda.synchronize()
        del arm
        torch.cuda.empty_cache()
    arm = bench.Arm(spec, args, device, 0, 1, host_io=True, total_steps=P + W + K)
    arm.prefill()
    arm.make_runner(P + W + K)
    runner = arm.runner
    marks = {}
    inner = runner._graph

    class Wrapped:
        def __init__(self, g):
            self.g = g

        def replay(self):
            cur = torch.cuda.current_stream(device)
            marks["t_launch"] = time.perf_counter()
            marks["e0"] = torch.cuda.Event(enable_timing=True)
            marks["e1"] = torch.cuda.Event(enable_timing=True)
            marks["e0"].record(cur)
            self.g.replay()
            marks["e1"].record(cur)
            marks["t_launched"] = time.perf_counter()

    runner._graph = lambda trig, par=0: Wrapped(inner(trig, par))
    for _ in range(P + W):
        arm.step()
    torch.cuda.synchronize()
    rows = []
    for _ in range(K):
        t0 = time.perf_counter()
        n = arm.step()
        t1 = time.perf_counter()
        rows.append({"wall_us": (t1 - t0) * 1e6, "pre_us": (marks["t_launch"] - t0) * 1e6,
                     "launch_call_us": (marks["t_launched"] - marks["t_launch"]) * 1e6,
                     "graph_us": marks["e0"].elapsed_time(marks["e1"]) * 1e3,
                     "post_us": (t1 - marks["t_launched"]) * 1e6, "trigger": n > 0})
    plain = [r for r in rows if not r["trigger"]]
    out = {"io_group": io_group, "after_value_arm": after_value, "io": arm.io.describe(), "steps": K,
           "plain": {k: statistics.mean(r[k] for r in plain) for k in rows[0] if k != "trigger"},
           "trigger": {k: statistics.mean(r[k] for r in rows if r["trigger"]) for k in rows[0] if k != "trigger"}
           if any(r["trigger"] for r in rows) else None}
    del arm, runner
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--io-group", default="0,1,2,4")
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--after-value", action="store_true", help="run the resident-input arm first, like bench.py")
    a, extra = ap.parse_known_args()
    for g in a.io_group.split(","):
        print(json.dumps(one(int(g), a.steps, a.warmup, extra, a.after_value)), flush=True)


if __name__ == "__main__":
    main()
