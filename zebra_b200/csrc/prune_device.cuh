// prune_device.cuh — pruned power iteration ("PPI") top-k T-PPR query on a static time-sorted
// adjacency (sm_100a).  Restates LuckyLYM/Zebra utils/util.py:185-276 (NeighborFinder.get_pruned_topk)
// and :152-154 (find_before).
//
// The query is stateless and embarrassingly parallel over (node, time) pairs: one warp per query.
// Per query the reference walks `depth` levels; every frontier item (node, t, w) looks up the
// neighbours of `node` strictly earlier than t, spreads weight over the `width` most recent ones
// (w0, w0*beta, w0*beta^2, ... by repeated multiplication) and accumulates them in an insertion-
// ordered dictionary keyed by (eidx, node, ts); all generated states form the next frontier.
// Here a level is expanded lane-parallel (binary search and weights per frontier item, an exclusive
// scan for the output offsets), duplicates are resolved by "first occurrence owns the sum, added in
// generation order" (the dictionary's exact fp64 summation order), and the top-k is a rank count —
// unless a group of exactly equal weights reaches the surviving ranks, in which case one lane runs
// numba's argsort (median-of-three quicksort + insertion sort) verbatim so the surviving entries and
// their order are the reference's.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace zebra {

struct PruneParams {
    const int64_t *indptr;      // [num_nodes + 1]
    const int32_t *adj_node, *adj_eidx;
    const double *adj_ts;
    const int32_t *q_nodes;     // [nq]
    const double *q_ts;         // [nq]
    int64_t nq, num_nodes;
    int32_t width, depth, k, t_max;   // t_max = width + width^2 + ... + width^depth
    double alpha, beta;
    int32_t *o_node, *o_eidx;   // [nq, k]
    float *o_dt, *o_w;
    int32_t *o_filled;          // [nq] 1 = the reference would have written this row
    char *scratch;              // per-warp work areas in global memory, or NULL: the work areas are the block's
    size_t scratch_per_warp;    // dynamic shared memory (blockDim.x / 32 areas of scratch_per_warp bytes)
};

// bytes of per-warp scratch for t_max states
__host__ __device__ inline size_t prune_scratch_bytes(int64_t t_max)
{
    // state: eidx i32, node i32, ts f64, w f64, first i32 | dict: state index i32, value f64, perm i32
    return static_cast<size_t>((t_max + 2) * (4 + 4 + 8 + 8 + 4 + 4 + 8 + 4) + 64 + 127) / 128 * 128;
}

__device__ __forceinline__ double pow_int(double a, int64_t n)
{   // numba lowers float ** int to exponentiation by squaring; beyond 0x10000 it calls libm's pow(a, float(n)).
    // CUDA's pow is within 1 ulp of libm's, not guaranteed bit-identical: a node with more than 65,536 earlier
    // neighbours (and beta so close to 1 that beta^65537 has not underflowed) may differ in the last bit there.
    if (n > 0x10000) return pow(a, static_cast<double>(n));
    double r = 1.0;
    while (n != 0) {
        if (n & 1) r = __dmul_rn(r, a);
        n >>= 1;
        a = __dmul_rn(a, a);
    }
    return r;
}

// numba.misc.quicksort argsort on A[0..n), run by one thread (only on exact score ties)
__device__ inline void numba_argsort_seq(const double *A, int n, int32_t *R)
{
    for (int i = 0; i < n; ++i) R[i] = i;
    if (n < 2) return;
    int lo_stack[64], hi_stack[64];
    int sp = 1;
    lo_stack[0] = 0; hi_stack[0] = n - 1;
    while (sp > 0) {
        --sp;
        int low = lo_stack[sp], high = hi_stack[sp];
        while (high - low >= 15) {
            const int mid = (low + high) >> 1;
            int32_t t;
            if (A[R[mid]] < A[R[low]]) { t = R[low]; R[low] = R[mid]; R[mid] = t; }
            if (A[R[high]] < A[R[mid]]) { t = R[high]; R[high] = R[mid]; R[mid] = t; }
            if (A[R[mid]] < A[R[low]]) { t = R[low]; R[low] = R[mid]; R[mid] = t; }
            const double pivot = A[R[mid]];
            t = R[high]; R[high] = R[mid]; R[mid] = t;
            int i = low, j = high - 1;
            for (;;) {
                while (i < high && A[R[i]] < pivot) ++i;
                while (j >= low && pivot < A[R[j]]) --j;
                if (i >= j) break;
                t = R[i]; R[i] = R[j]; R[j] = t;
                ++i; --j;
            }
            t = R[i]; R[i] = R[high]; R[high] = t;
            if (high - i > i - low) {
                if (high > i) { lo_stack[sp] = i + 1; hi_stack[sp] = high; ++sp; }
                high = i - 1;
            } else {
                if (i > low) { lo_stack[sp] = low; hi_stack[sp] = i - 1; ++sp; }
                low = i + 1;
            }
        }
        for (int i = low + 1; i <= high; ++i) {
            const int32_t kk = R[i];
            const double v = A[kk];
            int j = i;
            while (j > low && v < A[R[j - 1]]) { R[j] = R[j - 1]; --j; }
            R[j] = kk;
        }
    }
}

__global__ void __launch_bounds__(256)
prune_topk_kernel(const PruneParams p)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const unsigned full = 0xffffffffu;
    const int T = p.t_max + 2;
    // the work area of a query is read and written dozens of times per state by dependent steps: in shared memory
    // when the host found room for it (the usual width 10-20, depth 2 needs 5-19 KB per warp), else in global memory
    extern __shared__ __align__(16) char prune_smem[];
    char *base = p.scratch ? p.scratch + static_cast<size_t>(warp) * p.scratch_per_warp
                           : prune_smem + static_cast<size_t>(threadIdx.x >> 5) * p.scratch_per_warp;
    double *s_ts = reinterpret_cast<double *>(base);
    double *s_w = s_ts + T;
    double *d_val = s_w + T;
    int32_t *s_e = reinterpret_cast<int32_t *>(d_val + T);
    int32_t *s_n = s_e + T;
    int32_t *s_first = s_n + T;
    int32_t *d_idx = s_first + T;
    int32_t *perm = d_idx + T;
    const int k = p.k;

    for (int64_t qi = warp; qi < p.nq; qi += n_warps) {
        const int q_node = p.q_nodes[qi];
        const double t_now = p.q_ts[qi];
        // ---- expand `depth` levels; level l's states are the frontier of level l+1 ----
        int total = 0, fs = 0, fcount = 1;
        for (int dep = 0; dep < p.depth; ++dep) {
            const int level_start = total;
            for (int c0 = 0; c0 < fcount; c0 += 32) {
                const int f = c0 + lane;
                int cnt = 0;
                int64_t last = 0;           // index of the most recent earlier neighbour
                double w = 0.0;
                if (f < fcount) {
                    int node; double ts, qw;
                    if (dep == 0) { node = q_node; ts = t_now; qw = 1.0; }
                    else { node = s_n[fs + f]; ts = s_ts[fs + f]; qw = s_w[fs + f]; }
                    const int64_t b = p.indptr[node];
                    int64_t lo = 0, hi = p.indptr[node + 1] - b;
                    while (lo < hi) {       // searchsorted side='left': strictly earlier entries
                        const int64_t mid = (lo + hi) >> 1;
                        if (p.adj_ts[b + mid] < ts) lo = mid + 1; else hi = mid;
                    }
                    if (lo > 0) {
                        const double norm = __dmul_rn(__ddiv_rn(p.beta, __dsub_rn(1.0, p.beta)),
                                                      __dsub_rn(1.0, pow_int(p.beta, lo)));
                        w = __ddiv_rn(__dmul_rn(__dmul_rn(qw, __dsub_rn(1.0, p.alpha)), p.beta), norm);
                        if (p.alpha != 0.0 && dep == 0) w = __dmul_rn(w, p.alpha);
                        cnt = lo < p.width ? static_cast<int>(lo) : p.width;
                        last = b + lo - 1;
                    }
                }
                // exclusive scan of cnt over the 32 items of this chunk
                int incl = cnt;
                for (int d = 1; d < 32; d <<= 1) {
                    const int y = __shfl_up_sync(full, incl, d);
                    if (lane >= d) incl += y;
                }
                int off = total + incl - cnt;
                total += __shfl_sync(full, incl, 31);
                for (int z = 0; z < cnt; ++z) {
                    const int64_t a = last - z;
                    s_e[off] = p.adj_eidx[a]; s_n[off] = p.adj_node[a]; s_ts[off] = p.adj_ts[a]; s_w[off] = w;
                    ++off;
                    w = __dmul_rn(w, p.beta);
                }
            }
            __syncwarp();
            fs = level_start; fcount = total - level_start;
            if (fcount == 0) break;
        }
        int32_t *o_node = p.o_node + qi * k, *o_eidx = p.o_eidx + qi * k;
        float *o_dt = p.o_dt + qi * k, *o_w = p.o_w + qi * k;
        if (total == 0) {
            // the reference leaves the caller's row untouched (util.py:241-242)
            if (lane < k) { o_node[lane] = 0; o_eidx[lane] = 0; o_dt[lane] = 0.f; o_w[lane] = 0.f; }
            if (lane == 0) p.o_filled[qi] = 0;
            continue;
        }
        // ---- dictionary: first occurrence of each key, in generation order ----
        int dn = 0;
        for (int g0 = 0; g0 < total; g0 += 32) {
            const int g = g0 + lane;
            int first = g;
            if (g < total) {
                const int e = s_e[g], nd = s_n[g];
                const double ts = s_ts[g];
                for (int h = 0; h < g; ++h)
                    if (s_e[h] == e && s_n[h] == nd && s_ts[h] == ts) { first = h; break; }
                s_first[g] = first;
            }
            const unsigned own = __ballot_sync(full, g < total && first == g);
            if (g < total && first == g) d_idx[dn + __popc(own & ((1u << lane) - 1))] = g;
            dn += __popc(own);
        }
        __syncwarp();
        // ---- value of an entry = its occurrences added in generation order ----
        for (int d0 = 0; d0 < dn; d0 += 32) {
            const int d = d0 + lane;
            if (d < dn) {
                const int f = d_idx[d];
                double v = s_w[f];
                for (int g = f + 1; g < total; ++g)
                    if (s_first[g] == f) v = __dadd_rn(v, s_w[g]);
                d_val[d] = v;
            }
        }
        __syncwarp();
        // ---- top-k (util.py:255-258): all if dn <= k, else the last k of argsort(values) ----
        const int m = dn <= k ? dn : k;
        if (dn <= k) {
            if (lane < m) perm[lane] = lane;
        } else {
            const int cut = dn - k;
            bool need_exact = false;
            for (int d0 = 0; d0 < dn; d0 += 32) {
                const int d = d0 + lane;
                if (d < dn) {
                    const double v = d_val[d];
                    int lt = 0, eq = 0;
                    for (int h = 0; h < dn; ++h) { const double u = d_val[h]; lt += u < v; eq += u == v; }
                    if (eq > 1 && lt + eq - 1 >= cut) need_exact = true;     // a tied group reaches the survivors
                    else if (lt >= cut) perm[lt - cut] = d;
                }
            }
            if (__any_sync(full, need_exact)) {
                __syncwarp();
                if (lane == 0) {
                    numba_argsort_seq(d_val, dn, s_first);       // s_first is free now
                    for (int j = 0; j < k; ++j) perm[j] = s_first[cut + j];
                }
            }
        }
        __syncwarp();
        if (lane < k) {
            int node = 0, eidx = 0;
            float w = 0.f, tsf = 0.f;
            if (lane < m) {
                const int d = perm[lane];
                const int f = d_idx[d];
                node = s_n[f]; eidx = s_e[f]; w = static_cast<float>(d_val[d]); tsf = static_cast<float>(s_ts[f]);
            }
            o_node[lane] = node; o_eidx[lane] = eidx; o_w[lane] = w;
            o_dt[lane] = static_cast<float>(__dsub_rn(t_now, static_cast<double>(tsf)));
        }
        if (lane == 0) p.o_filled[qi] = 1;
        __syncwarp();
    }
}

}  // namespace zebra
