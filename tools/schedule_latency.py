"""Latency of ONE Schedule() (kgwe_schedule_queue with one workload: score, pick, commit) on the 10k-node image, for both
device forms of the queue kernel; and of short queues.  KGWE_QUEUE_FORM is read when an engine is created."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package()
nodes, pods, led = kg.synth.config(3)
for form in ("cluster", "tile"):
    os.environ["KGWE_QUEUE_FORM"] = form
    for n in (1, 8, 64):
        eng = kg.Engine(max_nodes=len(nodes)); eng.upsert_nodes(nodes); eng.seed_ledger(led)
        warm = 20
        calls = min(300, len(pods) // n - warm)        # every workload UID is used once (a repeated UID ends a kernel segment)
        for i in range(warm):
            eng.schedule_queue(pods[i * n:(i + 1) * n])
        ts = []
        for i in range(warm, warm + calls):
            q = pods[i * n:(i + 1) * n]
            a = time.perf_counter(); eng.schedule_queue(q); ts.append(time.perf_counter() - a)
        ts = np.array(ts) * 1e6
        print(f"{form}: {n:3d} workload(s) per call: p50 {np.percentile(ts, 50):7.1f} us  p99 {np.percentile(ts, 99):7.1f} us  ({np.percentile(ts, 50) / n:6.1f} us per workload)", flush=True)
        eng.close()
