// gather_device.cuh — fused PPR-weighted top-k neighbour gather (sm_100a).
//
// Replaces, for all (alpha,beta) instances of one batch and in ONE launch, the per-instance chain of
// torch launches in LuckyLYM/Zebra modules/embedding_module.py:236-271:
//     memory[nodes_0]                                  (:243-245)  -> src_rows [R, Dm]
//     memory[selected_nodes]                           (:253)      \
//     edge_features[selected_edge_idxs]                (:257)       > X [n_tppr, R*k, Dm+De+Dt]
//     time_encoder(selected_delta_time) = cos(dt*w+b)  (:261, model/time_encoding.py:23-28)
//     torch.cat(...)                                   (:264)      /
//     weights / weights.sum(1), zero where the sum is 0 (:268-271) -> wn [n_tppr, R, k]
// The ids / edge ids / dt / weights are read straight from the span kernel's output buffer
// ([n_tppr][4][R*k] 4-byte planes), so nothing leaves the GPU between the T-PPR update and the
// torch projections.  Pure HBM streaming: one warp per output row, 128-bit loads and stores.
// cosf is the accurate libdevice routine (arguments reach 1e6 and more: no fast-math).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>

namespace zebra {

struct GatherParams {
    const char *topk;          // [n_tppr][4][R*k]: int32 node | int32 eidx | f32 dt | f32 w
    int32_t n_tppr, k;
    int64_t R;                 // rows per instance (3n or 2n)
    const float *memory;       // [num_nodes, Dm]
    const float *edge_feat;    // [n_edge_rows, De]
    const float *time_w, *time_b;   // [Dt]
    int64_t num_nodes, n_edge_rows;
    int32_t Dm, De, Dt;
    const int32_t *src_nodes;  // [R] or nullptr
    float *X;                  // [n_tppr, R*k, Dm+De+Dt]
    float *wn;                 // [n_tppr, R, k]
    float *src_rows;           // [R, Dm] or nullptr
    int32_t *status;           // [0] set to 1 on an out-of-range id (row skipped)
    // lazy memory update (all three NULL = plain gather): nodes of the pending set read their row from
    // upd_rows [U, Dm] instead of memory (see PendingSet below)
    const float *upd_rows;
    const uint32_t *bitmap;
    const int32_t *wordprefix;
};

// Pending set of the lazy memory update (modules/memory_updater.py:61-90): a bitmap over node ids plus,
// per 32-bit word, the number of set bits before it.  Node v is in the set iff bit v is set, and its row
// in the compacted (ascending) arrays is wordprefix[v >> 5] + popc(bits below v).
__device__ __forceinline__ int pending_row(const uint32_t *bitmap, const int32_t *wordprefix, int node)
{
    const uint32_t w = __ldg(bitmap + (node >> 5)), bit = 1u << (node & 31);
    return (w & bit) ? __ldg(wordprefix + (node >> 5)) + __popc(w & (bit - 1u)) : -1;
}

__device__ __forceinline__ const float *memory_row(const GatherParams &p, int node)
{
    if (p.bitmap) {
        const int r = pending_row(p.bitmap, p.wordprefix, node);
        if (r >= 0) return p.upd_rows + static_cast<size_t>(r) * p.Dm;
    }
    return p.memory + static_cast<size_t>(node) * p.Dm;
}

// Time encoding (model/time_encoding.py:23-28).  The reference runs torch.cos on the GPU, i.e. libdevice's
// accurate cosf; calling the same routine makes this part of X bit-identical to the reference's (an fp64
// range reduction for the large arguments was tried: it is as fast and marginally more accurate, but it
// differs from cosf in the last bit for 20 % of the arguments beyond 1e5 — parity wins).
__device__ __forceinline__ float cos_arg(float x) { return cosf(x); }

template <bool VEC>
__device__ __forceinline__ void copy_row(float *dst, const float *src, int n, int lane)
{
    if (VEC) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src);
        float4 *d4 = reinterpret_cast<float4 *>(dst);
        for (int c = lane; c < (n >> 2); c += 32) d4[c] = __ldg(s4 + c);
    } else {
        for (int c = lane; c < n; c += 32) dst[c] = __ldg(src + c);
    }
}

template <bool VEC>
__global__ void __launch_bounds__(256)
gather_fwd_kernel(const GatherParams p)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t slots = static_cast<int64_t>(p.n_tppr) * p.R * p.k;
    const int64_t plane = p.R * p.k;
    const int D = p.Dm + p.De + p.Dt;
    if (warp < slots) {
        // slot counts stay below 2^31 (checked on the host): 32-bit divisions
        const int64_t j = static_cast<uint32_t>(warp) / static_cast<uint32_t>(plane), o = warp - j * plane;
        const char *base = p.topk + static_cast<size_t>(j) * plane * 16;
        const int node = reinterpret_cast<const int32_t *>(base)[o];
        const int eidx = reinterpret_cast<const int32_t *>(base + plane * 4)[o];
        const float dt = reinterpret_cast<const float *>(base + plane * 8)[o];
        float *x = p.X + static_cast<size_t>(warp) * D;
        if (node < 0 || node >= p.num_nodes || eidx < 0 || eidx >= p.n_edge_rows) {
            if (lane == 0) atomicExch(p.status, 1);
        } else {
            copy_row<VEC>(x, memory_row(p, node), p.Dm, lane);
            copy_row<VEC>(x + p.Dm, p.edge_feat + static_cast<size_t>(eidx) * p.De, p.De, lane);
            float *xt = x + p.Dm + p.De;
            if (VEC) {
                for (int c = lane; c < (p.Dt >> 2); c += 32) {
                    const float4 w = __ldg(reinterpret_cast<const float4 *>(p.time_w) + c);
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(p.time_b) + c);
                    float4 r;
                    r.x = cos_arg(__fadd_rn(__fmul_rn(dt, w.x), b.x)); r.y = cos_arg(__fadd_rn(__fmul_rn(dt, w.y), b.y));
                    r.z = cos_arg(__fadd_rn(__fmul_rn(dt, w.z), b.z)); r.w = cos_arg(__fadd_rn(__fmul_rn(dt, w.w), b.w));
                    reinterpret_cast<float4 *>(xt)[c] = r;
                }
            } else {
                for (int c = lane; c < p.Dt; c += 32)
                    xt[c] = cos_arg(__fadd_rn(__fmul_rn(dt, __ldg(p.time_w + c)), __ldg(p.time_b + c)));
            }
        }
        // normalised weights: the warp of slot 0 of each row handles the row's k weights
        const int64_t row = static_cast<uint32_t>(o) / static_cast<uint32_t>(p.k);
        if (o - row * p.k == 0) {
            const float *w = reinterpret_cast<const float *>(base + plane * 12) + row * p.k;
            float *out = p.wn + static_cast<size_t>(j) * plane + row * p.k;
            float s = 0.f;
            for (int c = 0; c < p.k; ++c) s = __fadd_rn(s, w[c]);       // left-to-right fp32 sum
            for (int c = lane; c < p.k; c += 32) out[c] = (s == 0.f) ? 0.f : __fdiv_rn(w[c], s);
        }
    } else if (p.src_rows && warp < slots + p.R) {
        const int64_t r = warp - slots;
        const int node = p.src_nodes[r];
        if (node < 0 || node >= p.num_nodes) { if (lane == 0) atomicExch(p.status, 1); }
        else copy_row<VEC>(p.src_rows + static_cast<size_t>(r) * p.Dm, memory_row(p, node), p.Dm, lane);
    }
}

// The same gather with EIGHT consecutive slots per warp (VEC: rows are 16-byte multiples, float4 items; otherwise
// float items — configs without edge features have De = 1).  Why: with one slot per warp the row copies keep 25 / 43 of 32 lanes busy per
// pass and, worse, every warp runs cosf twice — the lanes of the slowest-decaying time dimensions (dt*w beyond
// 105,615: Payne-Hanek reduction) and all the others diverge.  Here
//   * the 8 x (Dm + De)/4 float4 items of the group are walked as ONE flat list (all 32 lanes busy, four
//     independent loads in flight per lane before the first store);
//   * the time encoding gives lane l the slot l & 7 and the dimension quads (l >> 3) + 4t: in pass t every lane of
//     the warp works on the same four quads, so the slow reduction is paid in the first pass (or two) by all
//     lanes together and the remaining passes are uniformly fast.
// Arithmetic per element is unchanged (fmul, fadd, cosf): bit-identical to the one-slot kernel.
constexpr int kGatherGroup = 8;
constexpr int kGatherLoads = 4;      // independent loads a lane has in flight before its first store (8: -10 % on large launches)

template <bool VEC>
__global__ void __launch_bounds__(256)
gather_fwd8_kernel(const GatherParams p)
{
    typedef typename std::conditional<VEC, float4, float>::type T;      // unit of the row copies
    constexpr int kSh = VEC ? 2 : 0;
    __shared__ const T *s_mem[8][kGatherGroup];
    __shared__ const T *s_edge[8][kGatherGroup];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t slots = static_cast<int64_t>(p.n_tppr) * p.R * p.k;
    const int64_t groups = (slots + kGatherGroup - 1) / kGatherGroup;
    const int64_t plane = p.R * p.k;
    const int Du = (p.Dm + p.De + p.Dt) >> kSh, Dmu = p.Dm >> kSh, W = (p.Dm + p.De) >> kSh;
    if (warp < groups) {
        // lane l (and l + 8, l + 16, l + 24) looks after slot first + (l & 7)
        const int64_t slot = warp * kGatherGroup + (lane & 7);
        bool ok = false;
        float dt = 0.f;
        if (slot < slots) {
            const int64_t j = static_cast<uint32_t>(slot) / static_cast<uint32_t>(plane), o = slot - j * plane;
            const char *base = p.topk + static_cast<size_t>(j) * plane * 16;
            const int node = reinterpret_cast<const int32_t *>(base)[o];
            const int eidx = reinterpret_cast<const int32_t *>(base + plane * 4)[o];
            dt = reinterpret_cast<const float *>(base + plane * 8)[o];
            ok = node >= 0 && node < p.num_nodes && eidx >= 0 && eidx < p.n_edge_rows;
            if (!ok && lane < kGatherGroup) atomicExch(p.status, 1);
            if (ok && lane < kGatherGroup) {
                s_mem[wib][lane] = reinterpret_cast<const T *>(memory_row(p, node));
                s_edge[wib][lane] = reinterpret_cast<const T *>(p.edge_feat + static_cast<size_t>(eidx) * p.De);
            }
            // normalised weights of the row this slot starts (left-to-right fp32 sum, as the one-slot kernel)
            const int64_t row = static_cast<uint32_t>(o) / static_cast<uint32_t>(p.k);
            if (lane < kGatherGroup && o - row * p.k == 0) {
                const float *w = reinterpret_cast<const float *>(base + plane * 12) + row * p.k;
                float *out = p.wn + static_cast<size_t>(j) * plane + row * p.k;
                float sum = 0.f;
                for (int c = 0; c < p.k; ++c) sum = __fadd_rn(sum, w[c]);
                for (int c = 0; c < p.k; ++c) out[c] = (sum == 0.f) ? 0.f : __fdiv_rn(w[c], sum);
            }
        }
        const unsigned vmask = __ballot_sync(0xffffffffu, ok) & 0xffu;
        __syncwarp();
        T *xg = reinterpret_cast<T *>(p.X) + static_cast<size_t>(warp) * kGatherGroup * Du;
        // ---- memory row | edge row of the 8 slots: one flat list of 8 * W items ----
        int sl = 0, c = lane;
        while (c >= W) { c -= W; ++sl; }
        while (sl < kGatherGroup) {
            T v[kGatherLoads];
            T *d[kGatherLoads];
            #pragma unroll
            for (int u = 0; u < kGatherLoads; ++u) {
                d[u] = nullptr;
                if (sl < kGatherGroup) {
                    if ((vmask >> sl) & 1u) {
                        const T *src = c < Dmu ? s_mem[wib][sl] + c : s_edge[wib][sl] + (c - Dmu);
                        v[u] = __ldg(src);
                        d[u] = xg + static_cast<size_t>(sl) * Du + c;
                    }
                    c += 32;
                    while (c >= W) { c -= W; ++sl; }
                }
            }
            #pragma unroll
            for (int u = 0; u < kGatherLoads; ++u)
                if (d[u]) *d[u] = v[u];
        }
        // ---- time encoding: lane -> slot (lane & 7); pass t covers dimensions [16t, 16t + 16) of all 8 slots ----
        if (ok) {
            if (VEC) {
                float4 *xt = reinterpret_cast<float4 *>(xg) + static_cast<size_t>(lane & 7) * Du + W;
                for (int c4 = lane >> 3; c4 < (p.Dt >> 2); c4 += 4) {
                    const float4 w = __ldg(reinterpret_cast<const float4 *>(p.time_w) + c4);
                    const float4 b = __ldg(reinterpret_cast<const float4 *>(p.time_b) + c4);
                    float4 r;
                    r.x = cos_arg(__fadd_rn(__fmul_rn(dt, w.x), b.x)); r.y = cos_arg(__fadd_rn(__fmul_rn(dt, w.y), b.y));
                    r.z = cos_arg(__fadd_rn(__fmul_rn(dt, w.z), b.z)); r.w = cos_arg(__fadd_rn(__fmul_rn(dt, w.w), b.w));
                    xt[c4] = r;
                }
            } else {
                // scalar rows: store u of a pass writes four adjacent floats per slot (lanes l, l+8, l+16, l+24)
                float *xt = reinterpret_cast<float *>(xg) + static_cast<size_t>(lane & 7) * Du + W;
                for (int c0 = 0; c0 < p.Dt; c0 += 16) {
                    #pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int cc = c0 + 4 * u + (lane >> 3);
                        if (cc < p.Dt)
                            xt[cc] = cos_arg(__fadd_rn(__fmul_rn(dt, __ldg(p.time_w + cc)), __ldg(p.time_b + cc)));
                    }
                }
            }
        }
    } else if (p.src_rows && warp < groups + p.R) {
        const int64_t r = warp - groups;
        const int node = p.src_nodes[r];
        if (node < 0 || node >= p.num_nodes) { if (lane == 0) atomicExch(p.status, 1); }
        else copy_row<VEC>(p.src_rows + static_cast<size_t>(r) * p.Dm, memory_row(p, node), p.Dm, lane);
    }
}

struct GatherBwdParams {
    const char *topk;
    int32_t n_tppr, k;
    int64_t R;
    const float *gX;           // [n_tppr, R*k, D] gradient of X (only the first Dm columns are used)
    const float *wn;           // normalised weights: slots with weight 0 receive an exactly-zero gradient
    const float *g_src;        // [R, Dm] or nullptr
    const int32_t *src_nodes;
    float *g_memory;           // [num_nodes, Dm], accumulated into
    int64_t num_nodes;
    int32_t Dm, D;
    // lazy memory update (sorted backward only): g_memory is [U, Dm], one row per node of the pending set;
    // gradients of nodes outside the set are not needed (memory itself takes no gradient) and are skipped
    const uint32_t *bitmap;
    const int32_t *wordprefix;
};

// Backward of the memory-row gathers (edge features and the time encoder are constants,
// model/tgn_model.py:41, model/time_encoding.py:18-23): scatter-add rows of gX into g_memory.
template <bool VEC>
__global__ void __launch_bounds__(256)
gather_bwd_kernel(const GatherBwdParams p)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t plane = p.R * p.k;
    const int64_t slots = static_cast<int64_t>(p.n_tppr) * plane;
    const float *g;
    int node;
    if (warp < slots) {
        // slot counts stay below 2^31 (checked on the host): 32-bit divisions
        const int64_t j = static_cast<uint32_t>(warp) / static_cast<uint32_t>(plane), o = warp - j * plane;
        if (p.wn[warp] == 0.f) return;
        node = reinterpret_cast<const int32_t *>(p.topk + static_cast<size_t>(j) * plane * 16)[o];
        g = p.gX + static_cast<size_t>(warp) * p.D;
    } else if (p.g_src && warp < slots + p.R) {
        node = p.src_nodes[warp - slots];
        g = p.g_src + static_cast<size_t>(warp - slots) * p.Dm;
    } else {
        return;
    }
    if (node < 0 || node >= p.num_nodes) return;
    float *dst = p.g_memory + static_cast<size_t>(node) * p.Dm;
    if (VEC) {
        for (int c = lane; c < (p.Dm >> 2); c += 32) {
            const float4 v = __ldg(reinterpret_cast<const float4 *>(g) + c);
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 4 * c), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
        }
    } else {
        for (int c = lane; c < p.Dm; c += 32) atomicAdd(dst + c, __ldg(g + c));
    }
}

// Reproducible backward, bit-identical to the reference's autograd.  The reference gathers memory
// rows with three (n_tppr + 1) separate index operations (embedding_module.py:243-245, 253); torch's
// backward of each is index_put(accumulate) with a stable sort — duplicates of a node are added one
// after the other in source order — and autograd then adds the pieces as
// ((g[n_tppr-1] + g[n_tppr-2]) + ... + g[0]) + g[src]  (reverse order of creation; checked on B200).
// Here the sources (slots [0, n_tppr*R*k) in buffer order, then the R src rows) arrive stably sorted
// by destination node (`keys` = node ids in sorted order, `order` = source indices).  The warp that
// sits on the first source of a node replays exactly that arithmetic: one accumulator per piece,
// filled in source order, pieces combined in autograd's order, result STORED (g_memory zero-filled
// beforehand).  No atomics; every run gives the same bits.
__global__ void __launch_bounds__(256)
gather_bwd_sorted_kernel(const GatherBwdParams p, const int32_t *keys, const int64_t *order, int64_t total, int chunks)
{
    // one warp per (sorted position, chunk of 32 columns): a hub node's long run of sources is walked by
    // `chunks` warps side by side, one column per lane, several row loads in flight
    const int lane = threadIdx.x & 31;
    const int64_t w = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t pos = w / chunks;
    const int col = static_cast<int>(w - pos * chunks) * 32 + lane;
    if (pos >= total) return;
    const int node = keys[pos];
    if (pos > 0 && keys[pos - 1] == node) return;            // not the first source of its node
    if (node < 0 || node >= p.num_nodes) return;
    int64_t out_row = node;
    if (p.bitmap) {
        out_row = pending_row(p.bitmap, p.wordprefix, node);
        if (out_row < 0) return;
    }
    const int64_t plane = p.R * p.k, slots = static_cast<int64_t>(p.n_tppr) * plane;
    int64_t end = pos + 1;
    while (end < total && keys[end] == node) ++end;
    const bool live = col < p.Dm;
    float tot = 0.f;
    bool first = true;
    // sources of a node are in ascending source index, i.e. piece 0, 1, ..., n_tppr-1, src: walk the
    // pieces in autograd's order; each piece is a contiguous part [a, b) of the run
    int64_t b = end;
    if (p.g_src) { while (b > pos && order[b - 1] >= slots) --b; }
    const int64_t src_begin = b;                               // [src_begin, end) are src rows
    for (int pc = p.n_tppr - 1; pc >= -1; --pc) {
        int64_t lo_q, hi_q;
        if (pc < 0) {
            if (!p.g_src) break;
            lo_q = src_begin; hi_q = end;
        } else {
            hi_q = b;
            const int64_t lo = pc * plane;
            while (b > pos && order[b - 1] >= lo) --b;
            lo_q = b;
        }
        float acc = 0.f;
        if (live) {
#pragma unroll 4
            for (int64_t q = lo_q; q < hi_q; ++q) {
                const int64_t src = order[q];
                const float *g = pc >= 0 ? p.gX + static_cast<size_t>(src) * p.D : p.g_src + static_cast<size_t>(src - slots) * p.Dm;
                acc = __fadd_rn(acc, __ldg(g + col));
            }
        }
        tot = first ? acc : __fadd_rn(tot, acc);
        first = false;
    }
    if (live) p.g_memory[static_cast<size_t>(out_row) * p.Dm + col] = tot;
}

// ---- pending-set selection: np.unique(neighbour ids) restricted to the nodes with a stored message, on the device ----
// Replaces, for the train branch of compute_embedding_tppr_ensemble (embedding_module.py:227-230), the D2H
// of the id planes, numba_unique on the host and the flag lookup of get_updated_memory
// (memory_updater.py:66-70).  Two launches: mark, then an ordered compaction by one block.
__global__ void __launch_bounds__(256)
pending_mark_kernel(const char *topk, int32_t n_tppr, int64_t plane, const uint8_t *flags, int64_t num_nodes,
                    uint32_t *bitmap, int32_t *status)
{
    const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= n_tppr * plane) return;
    const int64_t j = static_cast<uint32_t>(t) / static_cast<uint32_t>(plane), o = t - j * plane;
    const int node = reinterpret_cast<const int32_t *>(topk + static_cast<size_t>(j) * plane * 16)[o];
    if (node < 0 || node >= num_nodes) { atomicExch(status, 1); return; }
    if (flags[node]) atomicOr(bitmap + (node >> 5), 1u << (node & 31));
}

// One block of 1024 threads; thread t owns the contiguous words [t*per, (t+1)*per).  Writes the per-word
// prefix counts, the set's node ids in ascending order (what np.unique + np.where give) and the count.
__global__ void __launch_bounds__(1024)
pending_compact_kernel(const uint32_t *bitmap, int64_t n_words, int32_t *wordprefix, int64_t *ids, int32_t *count)
{
    __shared__ int32_t warp_sums[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t per = (n_words + 1023) / 1024;
    int64_t lo = static_cast<int64_t>(tid) * per, hi = lo + per;
    if (lo > n_words) lo = n_words;
    if (hi > n_words) hi = n_words;
    int32_t s = 0;
    for (int64_t w = lo; w < hi; ++w) s += __popc(bitmap[w]);
    int32_t inc = s;                                            // inclusive scan inside the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int32_t v = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += v;
    }
    if (lane == 31) warp_sums[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int32_t v = warp_sums[lane];
        int32_t i2 = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t u = __shfl_up_sync(0xffffffffu, i2, d);
            if (lane >= d) i2 += u;
        }
        warp_sums[lane] = i2 - v;                               // exclusive over warps
    }
    __syncthreads();
    int32_t base = warp_sums[wid] + inc - s;                    // set bits before this thread's words
    if (tid == 1023) *count = base + s;
    for (int64_t w = lo; w < hi; ++w) {
        uint32_t bits = bitmap[w];
        wordprefix[w] = base;
        while (bits) {
            const int b = __ffs(bits) - 1;
            ids[base++] = w * 32 + b;
            bits &= bits - 1u;
        }
    }
}

// ---- raw messages (model/tgn_model.py:204-226) --------------------------------
// After a batch the reference builds, for every distinct node u of [src ; dst] (its LAST interaction of the
// batch, partner v, edge e, time t):  message[u] = [ memory[u] | memory[v] | edge_features[e] |
// cos((float(t) - last_update[u]) * w + b) ], with four separate index operations, a time encoding and a
// torch.cat, each after its own small host-to-device copy.  One warp per node does all of it here.
struct RawMsgParams {
    const int32_t *u, *v, *e;      // [U] distinct nodes (ascending), their partners, edge ids
    const float *t;                // [U] float32 edge times
    int64_t U, num_nodes, n_edge_rows;
    const float *memory, *last_update, *edge_feat, *time_w, *time_b;
    int32_t Dm, De, Dt;
    float *msg;                    // [U, 2*Dm + De + Dt]
    int32_t *status;
};

template <bool VEC>
__global__ void __launch_bounds__(256)
raw_message_kernel(const RawMsgParams p)
{
    const int lane = threadIdx.x & 31;
    const int64_t r = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (r >= p.U) return;
    const int u = p.u[r], v = p.v[r], e = p.e[r];
    if (u < 0 || u >= p.num_nodes || v < 0 || v >= p.num_nodes || e < 0 || e >= p.n_edge_rows) {
        if (lane == 0) atomicExch(p.status, 1);
        return;
    }
    const int D = 2 * p.Dm + p.De + p.Dt;
    float *x = p.msg + static_cast<size_t>(r) * D;
    copy_row<VEC>(x, p.memory + static_cast<size_t>(u) * p.Dm, p.Dm, lane);
    copy_row<VEC>(x + p.Dm, p.memory + static_cast<size_t>(v) * p.Dm, p.Dm, lane);
    copy_row<VEC>(x + 2 * p.Dm, p.edge_feat + static_cast<size_t>(e) * p.De, p.De, lane);
    const float dt = __fsub_rn(p.t[r], __ldg(p.last_update + u));
    float *xt = x + 2 * p.Dm + p.De;
    for (int c = lane; c < p.Dt; c += 32)
        xt[c] = cos_arg(__fadd_rn(__fmul_rn(dt, __ldg(p.time_w + c)), __ldg(p.time_b + c)));
}

}  // namespace zebra
