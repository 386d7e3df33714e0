"""kgwe_score_top1_noalloc (the entry point that allocates nothing) against the existing
kgwe_score_batch_device(topk=1) on the same device buffers.

This file sorts last on purpose: a run with -x has then already run every other test.
The no-alloc variant keeps its arg-max key table in one fixed device array shared by all
calls in the process, so the interesting cases are two calls in a row (the second must not
see the first one's keys) and the size limit.  Tolerance is the one the other GPU tests
use: integers and bytes bit-exact, EstimatedBandwidth within 1e-6 relative (and in fact
identical)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
REL_TOL = 1e-6


def _same(abi, got, want):
    got, want = got.view(abi.PLACEMENT).reshape(-1), want.view(abi.PLACEMENT).reshape(-1)
    for f in abi.PLACEMENT.names:
        if f == "est_bandwidth_gbps":
            np.testing.assert_allclose(got[f], want[f], rtol=REL_TOL, atol=0)
        else:
            assert np.array_equal(got[f], want[f]), f


def test_noalloc_matches_device_entry_point_twice_in_a_row(kg):
    import torch
    abi, synth = kg.abi, kg.synth
    dev = torch.device("cuda", 0)
    nodes, pods, led = synth.config(2, 0.3)
    N = len(nodes)
    eng = kg.Engine(max_nodes=N)
    eng.upsert_nodes(nodes)
    eng.seed_ledger(led)
    hint = synth.make_hint(len(pods), N, 0.02)
    # second batch: fewer pods, bigger requests (lower scores, more "no node" answers), so a
    # key left over from the first call would win over the right answer if it were not cleared
    pods_b = pods[: len(pods) // 2].copy()
    pods_b["gpu_count"] = 8
    pods_b["min_memory_gb"] = 100
    stream = torch.cuda.Stream(device=dev)
    for batch, h in ((pods, hint), (pods_b, None)):
        P = len(batch)
        d_pods = torch.from_numpy(batch.view(np.uint8).reshape(P, abi.POD_REQ.itemsize)).to(dev)
        d_hint = torch.from_numpy(h).to(dev) if h is not None else None
        hp = d_hint.data_ptr() if d_hint is not None else None
        want_top = torch.zeros((P, abi.PLACEMENT.itemsize), dtype=torch.uint8, device=dev)
        want_sc = torch.zeros((P, N, abi.NODE_SCORE.itemsize), dtype=torch.uint8, device=dev)
        got_top = torch.full_like(want_top, 0xAB)
        got_sc = torch.full_like(want_sc, 0xCD)
        torch.cuda.synchronize()
        eng.score_batch_device(d_pods.data_ptr(), P, hp, want_sc.data_ptr(), 1, want_top.data_ptr(), stream.cuda_stream)
        eng.score_top1_noalloc(d_pods.data_ptr(), P, hp, got_sc.data_ptr(), got_top.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        _same(abi, got_top.cpu().numpy(), want_top.cpu().numpy())
        assert torch.equal(got_sc, want_sc)
        # and without the optional score matrix
        got_top2 = torch.full_like(want_top, 0x11)
        eng.score_top1_noalloc(d_pods.data_ptr(), P, hp, None, got_top2.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        _same(abi, got_top2.cpu().numpy(), want_top.cpu().numpy())
    assert (want_top.cpu().numpy().view(abi.PLACEMENT)["node_idx"] == abi.NO_NODE).any()
    eng.close()


def test_noalloc_rejects_batches_beyond_the_fixed_table(kg):
    import torch
    abi = kg.abi
    nodes, _, _ = kg.synth.config(1)
    eng = kg.Engine(max_nodes=len(nodes))
    eng.upsert_nodes(nodes)
    P = abi.NOALLOC_MAX_PODS + 1
    d_pods = torch.zeros((P, abi.POD_REQ.itemsize), dtype=torch.uint8, device="cuda:0")
    d_top = torch.zeros((P, abi.PLACEMENT.itemsize), dtype=torch.uint8, device="cuda:0")
    with pytest.raises(kg.EngineError) as ei:
        eng.score_top1_noalloc(d_pods.data_ptr(), P, None, None, d_top.data_ptr())
    assert ei.value.code == -8 and "KGWE_NOALLOC_MAX_PODS" in str(ei.value)
    # the limit itself is accepted
    eng.score_top1_noalloc(d_pods.data_ptr(), P - 1, None, None, d_top.data_ptr())
    torch.cuda.synchronize()
    eng.close()
