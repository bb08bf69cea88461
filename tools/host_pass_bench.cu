// Host-only timing of the node-sharded dependency pass of tppr_api.cu (no CUDA call is made, runs without a GPU):
//   nvcc -O3 -std=c++17 -fmad=false -gencode arch=compute_100a,code=sm_100a -o /tmp/host_pass_bench tools/host_pass_bench.cu && /tmp/host_pass_bench <world> [edges per rank]
// Synthetic stream: cubic-skewed endpoints over world x 1,140,149 nodes, uniform negatives, rank 0's share.
#include "../zebra_b200/csrc/tppr_api.cu"
#include <chrono>
#include <random>
#include <cstdio>
int main(int argc, char **argv)
{
    const int world = argc > 1 ? atoi(argv[1]) : 8;
    const int64_t per = argc > 2 ? atoll(argv[2]) : 7833140;
    const int64_t m = per * world, N = 1140149LL * world + 1;
    std::vector<int32_t> src(m), dst(m), neg(m);
    std::vector<double> ts(m);
    std::vector<int64_t> eidx(m);
    std::mt19937_64 rng(1);
    for (int64_t i = 0; i < m; ++i) {
        auto pick = [&]() { double u = (rng() >> 11) * (1.0 / 9007199254740992.0); int64_t v = (int64_t)(N * u * u * u); return (int32_t)(1 + v % (N - 1)); };
        src[i] = pick(); dst[i] = pick(); neg[i] = 1 + rng() % (N - 1); ts[i] = (double)i; eidx[i] = i + 1;
    }
    zebra_tppr h;
    h.num_nodes = N; h.world = world; h.rank = 0; h.world_log2 = 0; while ((1 << h.world_log2) < world) ++h.world_log2;
    h.dep.assign(N, NodeDep{-1, -1, 0, 0});
    std::vector<char> stage;
    for (int rep = 0; rep < 3; ++rep) {
        auto t0 = std::chrono::steady_clock::now();
        std::vector<int32_t> list; std::vector<int64_t> ofs;
        shard_work_list(&h, src.data(), dst.data(), neg.data(), eidx.data(), m, 200, list, ofs);
        auto t1 = std::chrono::steady_clock::now();
        const size_t lcap = (list.size() + 8) & ~(size_t)7;
        if (stage.size() < lcap * kEdgeBytes) stage.resize(lcap * kEdgeBytes);
        auto t2 = std::chrono::steady_clock::now();
        fill_soa_sharded(&h, stage.data(), lcap, src.data(), dst.data(), neg.data(), ts.data(), eidx.data(), list);
        auto t3 = std::chrono::steady_clock::now();
        printf("world %d edges %lld list %zu: work list %.3f s, stage alloc %.3f s, records %.3f s, total %.3f s (%.1f ns/edge)\n", world, (long long)m, list.size(),
               std::chrono::duration<double>(t1 - t0).count(), std::chrono::duration<double>(t2 - t1).count(), std::chrono::duration<double>(t3 - t2).count(),
               std::chrono::duration<double>(t3 - t0).count(), std::chrono::duration<double>(t3 - t0).count() / m * 1e9);
    }
    return 0;
}
