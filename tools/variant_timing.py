"""K1 on the headline workload in one of its other forms (ncu target): exhaustive | int32 | fp64 | memo | prune."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
import torch
kg = entry.load_package()
abi = kg.abi
what = sys.argv[1] if len(sys.argv) > 1 else "exhaustive"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nodes, pods, ledger = kg.synth.config(3)
flags = {"exhaustive": abi.CFG_SUBSET_TABLES, "memo": abi.CFG_MEMOISE_GROUPS, "prune": abi.CFG_PRUNE_TIED_STARTS}.get(what, 0)
rng = kg.synth.SplitMix64(77)
if what == "exhaustive":
    pods = pods.copy(); pods["flags"] = abi.POD_EXHAUSTIVE
if what == "int32":
    nodes = nodes.copy(); bw = nodes["nvl_bw"]; nodes["nvl_bw"] = np.where(bw > 0, np.floor(bw * (0.5 + rng.uniform(bw.size).reshape(bw.shape))), 0.0)
if what == "fp64":
    nodes = nodes.copy(); bw = nodes["nvl_bw"]; nodes["nvl_bw"] = bw * (0.5 + rng.uniform(bw.size).reshape(bw.shape))
eng = kg.Engine(max_nodes=len(nodes), flags=flags)
eng.upsert_nodes(nodes); eng.seed_ledger(ledger)
P, N = len(pods), len(nodes)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
d_pods = torch.from_numpy(pods.view(np.uint8).reshape(P, abi.POD_REQ.itemsize).copy()).cuda()
d_scores = torch.empty((P, N, 16), dtype=torch.uint8, device="cuda")
d_top = torch.empty((P, 32), dtype=torch.uint8, device="cuda")
for i in range(reps + 1):
    torch.cuda.synchronize(); a = time.perf_counter()
    eng.score_batch_device(d_pods.data_ptr(), P, None, d_scores.data_ptr(), 1, d_top.data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    if i: print(f"{what}: {(time.perf_counter() - a) * 1e3:.3f} ms per step")
eng.close()
