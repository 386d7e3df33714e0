"""Timing of ranked top-k (K2) on BASELINE configs[2] for a few k; device-resident."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package(); abi = kg.abi
nodes, pods, led = kg.synth.config(3)
N, P = len(nodes), len(pods)
eng = kg.Engine(max_nodes=N); eng.upsert_nodes(nodes); eng.seed_ledger(led)
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
d_pods = torch.from_numpy(pods.view(np.uint8).reshape(P, 32)).to(dev)
d_sc = torch.empty((P, N, 16), dtype=torch.uint8, device=dev)
for k in (1, 2, 4, 8, 16):
    d_top = torch.empty((P, k, 32), dtype=torch.uint8, device=dev)
    for _ in range(2):
        eng.score_batch_device(d_pods.data_ptr(), P, None, d_sc.data_ptr(), k, d_top.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(5):
        eng.score_batch_device(d_pods.data_ptr(), P, None, d_sc.data_ptr(), k, d_top.data_ptr(), st.cuda_stream)
    b.record(st); torch.cuda.synchronize()
    print(f"topk={k}: {a.elapsed_time(b)/5:.2f} ms per 10k x 10k batch", flush=True)
