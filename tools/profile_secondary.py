"""Small driver for ncu captures of the secondary kernels: the stateful queue (K3, tile form) and the
MIG bin-packing kernel (E2) on the BASELINE-sized inputs.  Run plain first, then under ncu:
  ncu --set full --clock-control none -k regex:'schedule_tile_kernel|mig_pack_tree_kernel' -c 2 -o out python tools/profile_secondary.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package()
nodes, pods, led = kg.synth.config(3)
eng = kg.Engine(max_nodes=len(nodes)); eng.upsert_nodes(nodes); eng.seed_ledger(led)
d = eng.schedule_queue(pods[:2000])
print("queue: %.2f us/pod, %d placed" % (eng.stats()["last_call_gpu_ms"] * 1e3 / 2000, int((d["node_idx"] != kg.abi.NO_NODE).sum())))
g, p = kg.synth.make_mig(10000, 20000)
a, _ = eng.mig_pack(g, p)
print("mig: %.2f ms for %d requests" % (eng.stats()["last_call_gpu_ms"], len(p)))
