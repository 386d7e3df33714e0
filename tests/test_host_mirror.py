"""C++ mirror of the reference's TopologyAwareScheduler interface (host/): packer on CPU, the
full Schedule / ScoreNodes / ReleaseAllocation surface on the GPU against the oracle."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_TESTS = os.path.join(ROOT, "tests", "host")


def _build(target):
    import __graft_entry__ as entry
    entry.build_engine()
    entry.build_host()
    entry.build_oracle()
    subprocess.check_call(["make", "-s", "-C", HOST_TESTS, target])


def test_packer_and_tables_cpu():
    _build("host_packer_test")
    r = subprocess.run([os.path.join(HOST_TESTS, "host_packer_test")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "host_packer_test OK" in r.stdout


@pytest.mark.gpu
def test_scheduler_mirror_against_oracle_gpu():
    _build("host_scheduler_test")
    r = subprocess.run([os.path.join(HOST_TESTS, "host_scheduler_test")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "host_scheduler_test OK" in r.stdout


def test_extender_json_and_pod_mapping_cpu():
    _build("json_test")
    r = subprocess.run([os.path.join(HOST_TESTS, "json_test")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "json_test OK" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_extender_http_verbs_against_oracle(kg, orc, tmp_path):
    """The extender front over a real socket: filter / prioritize / bind with ExtenderArgs as kube-scheduler sends them
    (both the NodeList form of nodeCacheCapable: false and the nodenames form), answers checked against the oracle on the
    same packed node images; bind really takes the GPUs (a second pod no longer fits, a third bind is refused)."""
    import http.client
    import json
    import numpy as np
    _build("extender_server")
    abi = kg.abi
    img_path = str(tmp_path / "images.bin")
    N = 24
    proc = subprocess.Popen([os.path.join(HOST_TESTS, "extender_server"), str(N), img_path], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    try:
        line = proc.stdout.readline()
        assert line.startswith("PORT "), line + proc.stderr.read()
        port = int(line.split()[1])
        nodes = np.fromfile(img_path, dtype=abi.NODE_IMAGE)
        assert len(nodes) == N
        names = [f"node-{i:03d}" for i in range(N)]

        def call(method, path, body=None):
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=30)
            c.request(method, path, body=None if body is None else json.dumps(body), headers={"Content-Type": "application/json"})
            r = c.getresponse(); data = r.read().decode(); c.close()
            return r.status, data

        def pod(uid, gpus, pref="NVLinkOptimal", mem=10, arch=None):
            ann = {"kgwe.io/topology-preference": pref, "kgwe.io/min-memory-gb": str(mem)}
            if arch:
                ann["kgwe.io/gpu-architecture"] = arch
            return {"metadata": {"name": uid, "namespace": "ml", "uid": uid, "annotations": ann},
                    "spec": {"containers": [{"name": "main", "resources": {"limits": {"nvidia.com/gpu": str(gpus)}}}]}}

        def oracle_row(k, pref, mem, arch, ledger):
            p = np.zeros(1, dtype=abi.POD_REQ)
            p["wl_id"], p["gpu_count"], p["min_memory_gb"] = 999, k, mem
            p["topology_pref"] = {"None": 0, "NVLinkOptimal": 1, "SameNUMA": 3}[pref]
            p["arch"] = {None: 0, "Hopper": abi.ARCH_HOPPER, "Blackwell": abi.ARCH_BLACKWELL}[arch]
            s, _, _ = orc.score_batch(abi, nodes, ledger, p, want_scores=True)
            return s[0]

        assert call("GET", "/healthz") == (200, "ok\n")
        empty = np.zeros(0, dtype=abi.ALLOC)
        # filter, NodeList form (nodeCacheCapable: false), with a node the snapshot does not know
        st, body = call("POST", "/filter", {"pod": pod("p1", 4, arch="Hopper"), "nodes": {"items": [{"metadata": {"name": n}} for n in names + ["ghost"]]}})
        fr = json.loads(body)
        want = oracle_row(4, "NVLinkOptimal", 10, "Hopper", empty)
        kept = [names[i] for i in range(N) if want[i]["flags"] & abi.F_KEPT]
        assert st == 200 and fr["error"] == "" and [x["metadata"]["name"] for x in fr["nodes"]["items"]] == kept
        assert set(fr["failedNodes"]) == set(names + ["ghost"]) - set(kept) and 0 < len(kept) < N
        # prioritize, nodenames form
        st, body = call("POST", "/prioritize", {"Pod": pod("p1", 4, arch="Hopper"), "NodeNames": names})
        pr = json.loads(body)
        assert [x["host"] for x in pr] == names
        for i, x in enumerate(pr):
            exp = min(100, max(0, int(want[i]["total"]))) * 10 // 100 if want[i]["flags"] & abi.F_KEPT else 0
            assert x["score"] == exp
        # bind p1 to its best node: the GPUs are taken ...
        best = max(range(N), key=lambda i: (int(want[i]["total"]) if want[i]["flags"] & abi.F_KEPT else -1, -i))
        st, body = call("POST", "/bind", {"podName": "p1", "podNamespace": "ml", "podUID": "p1", "node": names[best]})
        br = json.loads(body)
        assert br["error"] == "" and len(br["selectedGPUs"]) == 4 and br["score"] == int(want[best]["total"])
        led = np.zeros(1, dtype=abi.ALLOC)
        led["wl_id"], led["node_idx"], led["gpu_mask"] = 1, best, int(want[best]["gpu_mask"])
        # ... so an 8-GPU pod no longer fits that node, and the oracle with that ledger agrees on every node
        st, body = call("POST", "/filter", {"pod": pod("p2", 8, pref="None"), "nodenames": names})
        fr2 = json.loads(body)
        want2 = oracle_row(8, "None", 10, None, led)
        assert fr2["nodenames"] == [names[i] for i in range(N) if want2[i]["flags"] & abi.F_KEPT] and names[best] in fr2["failedNodes"]
        st, body = call("POST", "/bind", {"podName": "p2", "podNamespace": "ml", "podUID": "p2", "node": names[best]})
        assert "failed to schedule workload p2" in json.loads(body)["error"]
        # a pod that was never seen, a malformed body, an unknown path
        assert "never filtered" in json.loads(call("POST", "/bind", {"podUID": "nobody", "node": names[0]})[1])["error"]
        assert "bad ExtenderArgs" in json.loads(call("POST", "/filter", None)[1])["error"]
        assert call("GET", "/nope")[0] == 404
        # metrics: the histogram has seen the calls
        st, text = call("GET", "/metrics")
        assert st == 200 and 'kgwe_scheduling_latency_us_bucket{le="+Inf"}' in text and "kgwe_scheduling_successes_total 1" in text
        count = int([l for l in text.splitlines() if l.startswith("kgwe_scheduling_latency_us_count")][0].split()[1])
        assert count >= 5
        call("POST", "/quitquitquit", {})
        assert proc.wait(timeout=30) == 0
    finally:
        if proc.poll() is None:
            proc.kill()
