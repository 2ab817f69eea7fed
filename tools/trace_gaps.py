"""Summarise a torch.profiler chrome trace: GPU busy time per stream, the largest idle gaps on the busiest
stream (with the kernels around them) and time per kernel name.  Diagnosis aid for pipeline bubbles.

    python tools/trace_gaps.py gpurun_out/trace/trace_rank0.json [--gaps 15]
"""
import collections
import json
import re
import sys


def main():
    path = sys.argv[1]
    ngaps = int(sys.argv[sys.argv.index("--gaps") + 1]) if "--gaps" in sys.argv else 15
    ev = json.load(open(path))["traceEvents"]
    k = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
    by_stream = collections.defaultdict(list)
    for e in k:
        by_stream[e["args"].get("stream", e.get("tid"))].append(e)
    t0 = min(e["ts"] for e in k)
    t1 = max(e["ts"] + e["dur"] for e in k)
    print(f"{path}: {len(k)} GPU events over {(t1 - t0) / 1e3:.1f} ms")
    main_stream, best = None, 0
    for s, es in by_stream.items():
        busy = sum(e["dur"] for e in es)
        print(f"  stream {s}: {len(es)} events, busy {busy / 1e3:.1f} ms ({100 * busy / (t1 - t0):.0f} %)")
        if busy > best:
            main_stream, best = s, busy
    es = sorted(by_stream[main_stream], key=lambda e: e["ts"])
    gaps = []
    for a, b in zip(es, es[1:]):
        g = b["ts"] - (a["ts"] + a["dur"])
        if g > 0:
            gaps.append((g, a, b))
    tot_gap = sum(g for g, _, _ in gaps)
    print(f"  main stream {main_stream}: idle {tot_gap / 1e3:.1f} ms in {len(gaps)} gaps; "
          f">100us: {sum(g for g, _, _ in gaps if g > 100) / 1e3:.1f} ms, >1ms: {sum(g for g, _, _ in gaps if g > 1000) / 1e3:.1f} ms")
    for g, a, b in sorted(gaps, key=lambda x: -x[0])[:ngaps]:
        print(f"    gap {g / 1e3:8.2f} ms at +{(a['ts'] - t0) / 1e3:8.1f} ms  after {a['name'][:50]!r}  before {b['name'][:50]!r}")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for e in k:
        n = re.sub(r"<.*", "", e["name"])[:60]
        agg[n][0] += 1
        agg[n][1] += e["dur"]
    print("  top kernels:")
    for n, (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:18]:
        print(f"    {d / 1e3:9.2f} ms  n={c:5d}  {n}")


if __name__ == "__main__":
    main()
