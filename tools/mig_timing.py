"""E2 packer timing on BASELINE configs[3] (10,000 GPUs, 50,000 requests); also the ncu target for the MIG kernels."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as entry
kg = entry.load_package()
eng = kg.Engine(max_nodes=1)
mg, mp = kg.synth.make_mig(10000, 50000)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    a = time.perf_counter(); asg, after = eng.mig_pack(mg, mp); w = time.perf_counter() - a
    print("mig_pack 10k x 50k: wall %.2f ms kernel %.2f ms placed %d" % (w * 1e3, eng.stats()["last_call_gpu_ms"], (asg["gpu_idx"] != kg.abi.NO_NODE).sum()))
eng.close()
