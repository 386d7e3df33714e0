"""The stateful queue on the 10k-node image in both forms of the kernel (cluster form / grid-wide tile form): device time per
workload and a byte comparison of the decisions.  KGWE_QUEUE_FORM is read when an engine is created."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
nodes, pods, led = kg.synth.config(3)
res = {}
forms = sys.argv[2].split(",") if len(sys.argv) > 2 else ["cluster", "tile"]
for form in forms:
    os.environ["KGWE_QUEUE_FORM"] = form
    eng = kg.Engine(max_nodes=len(nodes)); eng.upsert_nodes(nodes); eng.seed_ledger(led)
    eng.schedule_queue(pods[:64])
    d = eng.schedule_queue(pods[64:64 + n])
    st = eng.stats()
    print("%s: %.3f us/pod, %d of %d placed" % (form, st["last_call_gpu_ms"] * 1e3 / n, int((d["node_idx"] != kg.abi.NO_NODE).sum()), n), flush=True)
    res[form] = d.tobytes()
    eng.close()
if len(res) == 2:
    print("same decisions:", len(set(res.values())) == 1)
