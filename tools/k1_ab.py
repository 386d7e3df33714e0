"""A/B harness for score_kernel changes: BASELINE configs[2] (10k x 10k, ledger), device-resident, with and
without the score matrix; prints ms per launch pair and a checksum of the outputs (must not change)."""
import os, sys, zlib, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package(); abi = kg.abi
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
nodes, pods, led = kg.synth.config(3)
N, P = len(nodes), len(pods)
flags = int(sys.argv[1]) if len(sys.argv) > 1 else 0
if len(sys.argv) > 2 and sys.argv[2] == "fp64":       # irregular fractional bandwidths: the FP64 tier
    r = kg.synth.SplitMix64(77)
    nodes["nvl_bw"] = nodes["nvl_bw"] * (0.5 + r.uniform(nodes["nvl_bw"].size).reshape(nodes["nvl_bw"].shape))
eng = kg.Engine(max_nodes=N, flags=flags); eng.upsert_nodes(nodes); eng.seed_ledger(led)
d_pods = torch.from_numpy(pods.view(np.uint8).reshape(P, 32)).to(dev)
d_top = torch.empty((P, 32), dtype=torch.uint8, device=dev)
d_sc = torch.empty((P, N, 16), dtype=torch.uint8, device=dev)
for name, sc in (("top-1 only", None), ("matrix + top-1", d_sc.data_ptr())):
    for _ in range(3):
        eng.score_batch_device(d_pods.data_ptr(), P, None, sc, 1, d_top.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_it = 20
    a.record(st)
    for _ in range(n_it):
        eng.score_batch_device(d_pods.data_ptr(), P, None, sc, 1, d_top.data_ptr(), st.cuda_stream)
    b.record(st); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n_it
    crc = zlib.crc32(d_top.cpu().numpy().tobytes())
    if sc: crc = zlib.crc32(d_sc[:64].cpu().numpy().tobytes(), crc)
    print(f"{name:>15}: {ms:7.3f} ms  {N*P/ms/1e6:7.2f} G evaluations/s  crc {crc:08x}", flush=True)
eng.close()
