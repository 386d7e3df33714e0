"""K1 + finalisation on one rank's share of the headline workload at N ranks (10,000 / N nodes x 10,000 pods, score matrix + top-1,
device-resident), on ONE GPU: what the scoring kernel costs per rank in the strong-scaling runs, without the exchange."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package(); abi = kg.abi
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
nodes, pods, led = kg.synth.config(3)
P = len(pods)
d_pods = torch.from_numpy(pods.view(np.uint8).reshape(P, 32).copy()).to(dev)
d_top = torch.empty((P, 32), dtype=torch.uint8, device=dev)
flush = torch.empty(128 << 20, dtype=torch.uint8, device=dev)
for R in (1, 2, 4, 8):
    n = len(nodes) // R
    eng = kg.Engine(max_nodes=n); eng.upsert_nodes(nodes[:n])
    d_scores = torch.empty((P, n, 16), dtype=torch.uint8, device=dev)
    for _ in range(3):
        eng.score_batch_device(d_pods.data_ptr(), P, None, d_scores.data_ptr(), 1, d_top.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    tot = 0.0; n_it = 20
    for _ in range(n_it):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        eng.score_batch_device(d_pods.data_ptr(), P, None, d_scores.data_ptr(), 1, d_top.data_ptr(), st.cuda_stream)
        b.record(st); torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    ms = tot / n_it
    print(f"share 1/{R}: {n} nodes: {ms:.4f} ms  (x{R} = {ms * R:.3f} ms)", flush=True)
    eng.close(); del d_scores
