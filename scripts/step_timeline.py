"""In-graph timeline of one decode step: kernel start / end stamps from CUPTI (torch.profiler), so that gaps and
programmatic-dependent-launch overlaps between the kernels of a replayed CUDA graph can be read directly (ncu serialises
kernels and CUDA events per launch need the eager path).  Numbers from a profiled run are never bench values.

usage: python scripts/step_timeline.py [workload] [out_prefix]
"""
import json
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
from torch.profiler import ProfilerActivity, profile

import bench


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b-4bit-b64-t256"
    out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/timeline_" + wl
    cfg_name, B, T = bench.WORKLOADS[wl]
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    cfg, model = bench.build_model(cfg_name, dev)
    cache, sess = bench.make_session(model, B, T, 64)
    for _ in range(4):
        sess.run()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            sess.run()
        torch.cuda.synchronize()
    prof.export_chrome_trace(out + "_trace.json")
    ev = json.load(open(out + "_trace.json"))["traceEvents"]
    ks = sorted((e for e in ev if e.get("cat") == "kernel"), key=lambda e: e["ts"])
    os.remove(out + "_trace.json")
    # last replay only: split on the embed kernel
    starts = [i for i, e in enumerate(ks) if "embed" in e["name"] or "dequant_kernel" in e["name"]]
    ks = ks[starts[-1]:] if starts else ks
    t0 = ks[0]["ts"]
    lines = []
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for i, e in enumerate(ks):
        nxt = ks[i + 1] if i + 1 < len(ks) else None
        end = e["ts"] + e["dur"]
        gap = (nxt["ts"] - end) if nxt else 0.0
        s2s = (nxt["ts"] - e["ts"]) if nxt else e["dur"]
        name = e["name"].split("(")[0].replace("b200p::", "")[:60]
        g = e.get("args", {}).get("grid")
        key = name + " grid=" + str(g)
        a = agg[key]
        a[0] += 1
        a[1] += e["dur"]
        a[2] += gap
        a[3] += s2s
        if i < 40:
            lines.append("%9.2f  dur %7.2f  gap_to_next %7.2f  %s grid=%s" % (e["ts"] - t0, e["dur"], gap, name, g))
    total = ks[-1]["ts"] + ks[-1]["dur"] - t0
    with open(out + ".txt", "w") as f:
        f.write("workload %s: %d kernels in the last replay, first start -> last end %.1f us (under CUPTI)\n" % (wl, len(ks), total))
        f.write("per kernel (count, mean duration us, mean gap end->next start us [negative = PDL overlap], mean start->next start us, sum start->start us)\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][3]):
            f.write("%4d  dur %7.2f  gap %7.2f  s2s %7.2f  sum_s2s %8.1f  %s\n" % (a[0], a[1] / a[0], a[2] / a[0], a[3] / a[0], a[3], k))
        f.write("\nfirst 40 kernels of the replay (start us, duration, gap to the next start)\n")
        f.write("\n".join(lines) + "\n")
    print(open(out + ".txt").read())


if __name__ == "__main__":
    main()
