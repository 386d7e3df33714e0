"""K1 rate vs cluster size (device-resident, top-1 only, no score matrix): 1k .. 1M nodes.
Beyond ~160k nodes the image (768 B/node) no longer fits the 126 MB L2; every node tile is still read
once per pod chunk only, so the kernel stays compute-bound."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package(); abi = kg.abi
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
base_nodes, pods, _ = kg.synth.config(3)
for N, P in ((1000, 10000), (10000, 10000), (100000, 10000), (1000000, 2000)):
    reps = (N + len(base_nodes) - 1) // len(base_nodes)
    nodes = np.tile(base_nodes, reps)[:N]
    eng = kg.Engine(max_nodes=N); eng.upsert_nodes(nodes)
    q = pods[:P]
    d_pods = torch.from_numpy(q.view(np.uint8).reshape(P, 32)).to(dev)
    d_top = torch.empty((P, 32), dtype=torch.uint8, device=dev)
    for _ in range(2):
        eng.score_batch_device(d_pods.data_ptr(), P, None, None, 1, d_top.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_it = 3
    a.record(st)
    for _ in range(n_it):
        eng.score_batch_device(d_pods.data_ptr(), P, None, None, 1, d_top.data_ptr(), st.cuda_stream)
    b.record(st); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n_it
    print(f"N={N:>8} P={P:>6}: {ms:9.3f} ms  {N*P/ms/1e6:8.2f} G evaluations/s  image {N*768/1e6:8.1f} MB", flush=True)
    eng.close()
