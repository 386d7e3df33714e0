// Test driver of the extender front (host/kgwe_extender.hpp): serves the fixture cluster over HTTP until told to quit
// (POST /quitquitquit).
//   extender_server <n_nodes> <images_out>    prints "PORT <p>" once it listens; writes the packed node images (what the
//                                             oracle needs to check the answers) to <images_out>
#include <cstdio>
#include <cstdlib>

#include "../../k8s-gpu-workload-enhancer_b200/host/kgwe_extender.hpp"
#include "cluster_fixture.hpp"

using namespace kgwe_host;

static void announce(int port) { printf("PORT %d\n", port); fflush(stdout); }

int main(int argc, char** argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 24;
    auto made = TopologyAwareScheduler::New(SchedulerConfig{}, (uint32_t)n + 8);
    if (!made.ok()) { fprintf(stderr, "%s\n", made.error.c_str()); return 2; }
    auto cluster = fixture::make_cluster(n);
    FILE* f = argc > 2 ? fopen(argv[2], "wb") : nullptr;
    for (auto& node : cluster) {
        if (!made.value->UpsertNode(node).empty()) return 3;
        PackedNode pk;
        if (!PackNode(node, &pk).empty()) return 3;
        if (f) fwrite(&pk.image, sizeof pk.image, 1, f);
    }
    if (f) fclose(f);
    ExtenderServer srv(made.value.get());
    const std::string err = srv.Serve(0, announce);
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 4; }
    return 0;
}
