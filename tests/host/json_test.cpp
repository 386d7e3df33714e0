// CPU: the extender front's JSON reader and its pod -> GPUWorkload mapping (host/kgwe_extender.hpp); no GPU, no engine call.
#include <cstdio>
#include <cstdlib>

#include "../../k8s-gpu-workload-enhancer_b200/host/kgwe_extender.hpp"

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); exit(1); } } while (0)
using namespace kgwe_host;

int main() {
    Json j; std::string err;
    CHECK(Json::Parse(R"({"a":[1,2.5,-3e2,true,false,null,"x\n\"y\" \u00e9\ud83d\ude00"],"B":{"c":{}}, "d":[]})", &j, &err));
    CHECK(j.type == Json::Object && j.get("a")->items.size() == 7 && j.get("a")->items[2].num == -300.0);
    CHECK(j.get("a")->items[6].str == "x\n\"y\" \xc3\xa9\xf0\x9f\x98\x80");
    CHECK(j.get("b") == j.get("B") && j.get("B")->get("c")->type == Json::Object && j.get("zzz") == nullptr);      // look-ups ignore case
    CHECK(Json::Quote("a\"b\\c\n\x01") == "\"a\\\"b\\\\c\\n\\u0001\"");
    for (const char* bad : {"", "{", "{\"a\":}", "[1,]", "{\"a\" 1}", "tru", "\"abc", "{} x", "[\"\\q\"]"}) CHECK(!Json::Parse(bad, &j, &err));
    // a pod as kube-scheduler sends it (abridged): two containers asking for 3 + 1 GPUs, one MIG resource, annotations
    const char* pod = R"({"metadata":{"name":"train-7","namespace":"ml","uid":"6f1c-77","annotations":{"kgwe.io/topology-preference":"NVLinkOptimal",
        "kgwe.io/min-memory-gb":"40","kgwe.io/gpu-architecture":"Hopper","kgwe.io/min-bandwidth-gbps":"450.5"}},
        "spec":{"containers":[{"name":"a","resources":{"limits":{"cpu":"4","nvidia.com/gpu":"3"}}},{"name":"b","resources":{"limits":{"nvidia.com/gpu":1,"nvidia.com/mig-1g.10gb":"2"}}},{"name":"c"}]}})";
    CHECK(Json::Parse(pod, &j, &err));
    GPUWorkload w;
    CHECK(WorkloadFromPod(j, &w).empty());
    CHECK(w.Name == "train-7" && w.Namespace == "ml" && w.UID == "6f1c-77" && w.Requirements.GPUCount == 4 && w.Requirements.MinMemoryGB == 40);
    CHECK(w.Requirements.TopologyPreference == "NVLinkOptimal" && w.Requirements.GPUArchitecture == "Hopper" && w.Requirements.MinBandwidthGBps == 450.5);
    CHECK(w.Requirements.MIGReq && w.Requirements.MIGReq->Profile == "1g.10gb" && w.Requirements.MIGReq->Count == 2);
    CHECK(Json::Parse(R"({"metadata":{"name":"x"}})", &j, &err) && !WorkloadFromPod(j, &w).empty());                  // no uid
    CHECK(Json::Parse(R"({"metadata":{"uid":"u"},"spec":{"containers":[{"resources":{"limits":{"nvidia.com/gpu":"two"}}}]}})", &j, &err) && !WorkloadFromPod(j, &w).empty());
    CHECK(Json::Parse(R"({"metadata":{"uid":"u"}})", &j, &err) && WorkloadFromPod(j, &w).empty() && w.Requirements.GPUCount == 0 && w.Requirements.TopologyPreference == "None");
    printf("json_test OK\n");
    return 0;
}
