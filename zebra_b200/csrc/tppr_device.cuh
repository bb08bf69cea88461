// tppr_device.cuh — device side of the streaming T-PPR span kernel (sm_100a).
//
// One warp owns one work item = (edge i, instance j) of a chronological span.  Items are claimed
// in stream order from a global ticket, so every dependency of an item is held by a warp that is
// already running: per-node completion counters (`done`) replace the barrier between the
// reference's conflict-free "waves" with point-to-point waits (dataflow over the same DAG).
//
// Semantics restated from the reference (LuckyLYM/Zebra utils/util.py:447-469, 495-574); the
// arithmetic is IEEE fp64 with separate multiply and add roundings (build with -fmad=false; the
// intrinsics below make it explicit) so scores are bit-identical to the reference's numba code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace zebra {

constexpr int kMaxInst = 8;
constexpr int kWarpsPerBlock = 8;
constexpr unsigned kFull = 0xffffffffu;

struct SpanParams {
    int64_t n_chunk;        // edges in this launch
    int64_t n_total;        // edges of the whole span (row stride of the output)
    int64_t i_offset;       // index of this launch's first edge inside the span
    int64_t rows_total;     // rows_per_edge * n_total
    int32_t k, n_tppr, rows_per_edge, inst_mask;
    int32_t stride;         // bytes per node record
    uint32_t ticket_base;
    const int32_t *src, *dst, *neg, *eidx;
    const uint32_t *thr_s, *thr_d, *thr_n;
    const double *ts;
    char *tables[kMaxInst];
    uint32_t *done[kMaxInst];
    double alpha[kMaxInst], beta[kMaxInst];
    char *out;              // [n_tppr][4][rows_total*k] 4-byte elements, or nullptr
    // batch-structured output (plan mode): batch_size > 0 => edge g = i_offset + i belongs to batch
    // b = g / batch_size and its rows go to slot (b - first_batch) % out_slots of `out`, laid out
    // as the reference lays out that batch: [n_tppr][4][rows_per_edge*nb*k], nb = edges in batch b
    int64_t batch_size, first_batch, out_slots, slot_bytes;
    uint32_t *ticket;
    int32_t *status;        // [0] spin time-out flag, [1] exact-tie prunes (statistics)
};

struct WarpScratch {
    unsigned long long V[64];   // candidate scores (fp64 bit patterns; non-negative => integer order)
    unsigned char R[64];        // position -> candidate (numba quicksort working permutation)
    unsigned char P[64];        // candidate -> position
    unsigned char X[32], Y[32]; // swap exchange
    unsigned char stk[16];      // (low, high) stack
};

struct VecR {
    int e, n;        // key parts 0,1 of this lane's entry
    double ts, sc;   // key part 2, value
    int cnt;         // entries in use
    double norm;
};

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ VecR load_vec(const char *rec, int k, int lane)
{
    VecR v;
    int l = lane < k ? lane : k - 1;
    int2 kk = __ldcg(reinterpret_cast<const int2 *>(rec) + l);
    v.e = kk.x; v.n = kk.y;
    v.ts = __ldcg(reinterpret_cast<const double *>(rec + 8 * k) + l);
    v.sc = __ldcg(reinterpret_cast<const double *>(rec + 16 * k) + l);
    v.norm = __ldcg(reinterpret_cast<const double *>(rec + 24 * k));
    v.cnt = __ldcg(reinterpret_cast<const int *>(rec + 24 * k + 8));
    return v;
}

__device__ __forceinline__ unsigned long long pack_key(int e, int n)
{
    return (static_cast<unsigned long long>(static_cast<unsigned>(e)) << 32) | static_cast<unsigned>(n);
}

// fp64 -> unsigned integer with the same ordering (scores are non-negative on this path; the sign
// fold keeps the order right even for a caller that passes alpha > 1)
__device__ __forceinline__ unsigned long long order_bits(double v)
{
    const long long b = __double_as_longlong(v);
    return static_cast<unsigned long long>(b ^ ((b >> 63) | (1ll << 63)));
}

// utils/util.py:447-469: one output row in dict order; dt = f32(t_now - f64(f32(ts))).
__device__ __forceinline__ void emit_row(const SpanParams &p, char *base, size_t plane, int64_t row, const VecR &v,
                                         double t_now, int lane)
{
    if (lane >= p.k) return;
    size_t o = static_cast<size_t>(row) * p.k + lane;
    int node = 0, eidx = 0;
    float dt = 0.f, w = 0.f;
    if (v.cnt != 0) {
        float tsf = 0.f;
        if (lane < v.cnt) { node = v.n; eidx = v.e; w = static_cast<float>(v.sc); tsf = static_cast<float>(v.ts); }
        dt = static_cast<float>(__dsub_rn(t_now, static_cast<double>(tsf)));
    }
    reinterpret_cast<int *>(base)[o] = node;
    reinterpret_cast<int *>(base + plane * 4)[o] = eidx;
    reinterpret_cast<float *>(base + plane * 8)[o] = dt;
    reinterpret_cast<float *>(base + plane * 12)[o] = w;
}

// One partition step of numba.misc.quicksort (median of three, pivot stashed at `high`, Hoare
// scan) on the permutation R over values V, computed warp-wide from two stopper bitmaps instead
// of a sequential scan.  Returns the pivot's final position.
__device__ __forceinline__ int qs_partition(WarpScratch &s, int low, int high, int lane)
{
    int mid = (low + high) >> 1;
    int rl = s.R[low], rm = s.R[mid], rh = s.R[high];
    unsigned long long vl = s.V[rl], vm = s.V[rm], vh = s.V[rh];
    if (vm < vl) { int t = rl; rl = rm; rm = t; unsigned long long u = vl; vl = vm; vm = u; }
    if (vh < vm) { int t = rh; rh = rm; rm = t; unsigned long long u = vh; vh = vm; vm = u; }
    if (vm < vl) { int t = rl; rl = rm; rm = t; unsigned long long u = vl; vl = vm; vm = u; }
    const unsigned long long pivot = vm;
    __syncwarp();
    if (lane == 0) { s.R[low] = rl; s.R[mid] = rh; s.R[high] = rm; }
    __syncwarp();
    const int p0 = lane, p1 = lane + 32;
    const bool in0 = p0 >= low && p0 < high, in1 = p1 >= low && p1 < high;
    const int r0 = s.R[p0], r1 = s.R[p1];
    const unsigned long long v0 = s.V[r0 & 63], v1 = s.V[r1 & 63];
    const bool L0 = in0 && !(v0 < pivot), L1 = in1 && !(v1 < pivot);   // i stops here
    const bool Q0 = in0 && !(pivot < v0), Q1 = in1 && !(pivot < v1);   // j stops here
    const unsigned long long Lm = __ballot_sync(kFull, L0) | (static_cast<unsigned long long>(__ballot_sync(kFull, L1)) << 32);
    const unsigned long long Qm = __ballot_sync(kFull, Q0) | (static_cast<unsigned long long>(__ballot_sync(kFull, Q1)) << 32);
    const int lb0 = __popcll(Lm & ((1ull << p0) - 1)), lb1 = __popcll(Lm & ((1ull << p1) - 1));
    const int qa0 = __popcll(Qm >> (p0 + 1)), qa1 = (p1 == 63) ? 0 : __popcll(Qm >> (p1 + 1));
    const bool swL0 = L0 && qa0 > lb0, swL1 = L1 && qa1 > lb1;
    const bool swQ0 = Q0 && lb0 > qa0, swQ1 = Q1 && lb1 > qa1;
    if (swL0) s.X[lb0] = r0;
    if (swL1) s.X[lb1] = r1;
    if (swQ0) s.Y[qa0] = r0;
    if (swQ1) s.Y[qa1] = r1;
    const int M = __popc(__ballot_sync(kFull, swL0)) + __popc(__ballot_sync(kFull, swL1));
    __syncwarp();
    if (swL0) s.R[p0] = s.Y[lb0];
    if (swL1) s.R[p1] = s.Y[lb1];
    if (swQ0) s.R[p0] = s.X[qa0];
    if (swQ1) s.R[p1] = s.X[qa1];
    // first unswapped left stopper (rank M) and the last swapped right stopper (rank M-1)
    const unsigned long long lM = __ballot_sync(kFull, L0 && lb0 == M) |
                                  (static_cast<unsigned long long>(__ballot_sync(kFull, L1 && lb1 == M)) << 32);
    const unsigned long long qM = __ballot_sync(kFull, Q0 && qa0 == M - 1) |
                                  (static_cast<unsigned long long>(__ballot_sync(kFull, Q1 && qa1 == M - 1)) << 32);
    int bound = (M > 0 && qM) ? (__ffsll(static_cast<long long>(qM)) - 1) : high;
    int i = lM ? (__ffsll(static_cast<long long>(lM)) - 1) : 64;
    if (bound < i) i = bound;
    __syncwarp();
    if (lane == 0) { unsigned char t = s.R[i]; s.R[i] = s.R[high]; s.R[high] = t; }
    __syncwarp();
    return i;
}

// Runs the partition phase of numba's argsort on n candidates; afterwards a stable sort of the
// permuted sequence equals numba's final order (its insertion sort on segments < 16 is stable).
__device__ __forceinline__ void qs_partition_phase(WarpScratch &s, int n, int lane)
{
    if (lane < n) s.R[lane] = lane;
    if (lane + 32 < n) s.R[lane + 32] = lane + 32;
    else s.R[lane + 32] = 0;
    __syncwarp();
    int sp = 0;
    int low = 0, high = n - 1;
    for (;;) {
        while (high - low >= 15) {
            int i = qs_partition(s, low, high, lane);
            if (high - i > i - low) {
                if (high > i) { if (lane == 0) { s.stk[2 * sp] = i + 1; s.stk[2 * sp + 1] = high; } ++sp; }
                high = i - 1;
            } else {
                if (i > low) { if (lane == 0) { s.stk[2 * sp] = low; s.stk[2 * sp + 1] = i - 1; } ++sp; }
                low = i + 1;
            }
        }
        if (sp == 0) break;
        --sp;
        __syncwarp();
        low = s.stk[2 * sp]; high = s.stk[2 * sp + 1];
    }
    __syncwarp();
    if (lane < n) s.P[s.R[lane]] = lane;
    if (lane + 32 < n) s.P[s.R[lane + 32]] = lane + 32;
    __syncwarp();
}

// New vector of one endpoint from the OLD vectors of itself (A) and its peer (B)
// (utils/util.py:509-559), written straight into the endpoint's record.
__device__ __forceinline__ void update_endpoint(const SpanParams &p, WarpScratch &s, const VecR &A, const VecR &B,
                                                int peer_id, int e_new, double t_new, double alpha, double beta,
                                                char *rec, int lane)
{
    const int k = p.k;
    const int c1 = A.cnt;
    const int c2 = (B.norm != 0.0) ? B.cnt : 0;
    const double new_norm = __dadd_rn(__dmul_rn(A.norm, beta), beta);
    const double s1c = __dmul_rn(__ddiv_rn(A.norm, new_norm), beta);
    const double s2c = __dmul_rn(__ddiv_rn(beta, new_norm), __dsub_rn(1.0, alpha));
    double vA = __dmul_rn(A.sc, s1c);
    double vB = __dmul_rn(B.sc, s2c);
    const unsigned long long keyA = pack_key(A.e, A.n), keyB = pack_key(B.e, B.n);

    // ---- duplicate keys between A and B: all-pairs equality through match.any ----
    bool dupB = false;
    const int chunk = 32 - c1;
    const unsigned ownmask = c1 ? (kFull >> (32 - c1)) : 0u;
    for (int base = 0; c1 > 0 && base < c2; base += chunk) {
        const int srcl = base + lane - c1;
        const bool present = lane >= c1 && srcl < c2;
        const int sl = present ? srcl : lane;
        const unsigned long long pk = __shfl_sync(kFull, keyB, sl);
        const double pts = __shfl_sync(kFull, B.ts, sl);
        const double padd = __shfl_sync(kFull, vB, sl);
        const unsigned long long mykey = lane < c1 ? keyA : (present ? pk : pack_key(-1 - lane, -1));
        const double myts = lane < c1 ? A.ts : pts;
        const unsigned m = __match_any_sync(kFull, mykey);
        unsigned cand = lane < c1 ? (m & ~ownmask) : (present ? (m & ownmask) : 0u);
        int partner = -1;
        while (__any_sync(kFull, cand != 0)) {
            const int t = cand ? (__ffs(cand) - 1) : lane;
            const double ots = __shfl_sync(kFull, myts, t);
            if (cand) {
                if (ots == myts) { partner = t; cand = 0; }
                else cand &= cand - 1;
            }
        }
        const double got = __shfl_sync(kFull, padd, partner >= 0 ? partner : lane);
        if (lane < c1 && partner >= 0) vA = __dadd_rn(vA, got);
        const bool inround = lane >= base && lane < base + chunk && lane < c2;
        const int pf = __shfl_sync(kFull, partner, inround ? (c1 + lane - base) : lane);
        if (inround && pf >= 0) dupB = true;
    }

    // ---- insertion-order positions: A's entries, then B's new keys, then the new key ----
    const bool ndB = lane < c2 && !dupB;
    const unsigned ndmask = __ballot_sync(kFull, ndB);
    int n = c1 + __popc(ndmask);
    int pos1 = c1 + __popc(ndmask & ((1u << lane) - 1));
    const unsigned long long keyN = pack_key(e_new, peer_id);
    const double w_new = (alpha != 0.0) ? __dmul_rn(s2c, alpha) : s2c;
    const bool hitA = lane < c1 && keyA == keyN && A.ts == t_new;
    const bool hitB = ndB && keyB == keyN && B.ts == t_new;
    const bool replaced = __any_sync(kFull, hitA || hitB);
    if (hitA) vA = w_new;                      // assignment, not accumulation (util.py:541)
    if (hitB) vB = w_new;
    // slot 1 of lane 31 is always free (c2 <= k <= 31): it carries the appended new key
    bool valid1 = ndB;
    int e1 = B.e, n1 = B.n;
    double ts1 = B.ts, v1 = vB;
    if (!replaced) {
        if (lane == 31) { valid1 = true; pos1 = n; e1 = e_new; n1 = peer_id; ts1 = t_new; v1 = w_new; }
        ++n;
    }
    const bool valid0 = lane < c1;

    int2 *keys = reinterpret_cast<int2 *>(rec);
    double *tsp = reinterpret_cast<double *>(rec + 8 * k);
    double *scp = reinterpret_cast<double *>(rec + 16 * k);
    if (n <= k) {
        // no prune: A's keys keep their slots, only their scores change
        if (valid0) scp[lane] = vA;
        if (valid1) { keys[pos1] = make_int2(e1, n1); tsp[pos1] = ts1; scp[pos1] = v1; }
    } else {
        // ---- keep the k largest, re-inserted in ascending order (util.py:553-559) ----
        const unsigned long long b0 = order_bits(vA);
        const unsigned long long b1 = order_bits(v1);
        __syncwarp();
        if (valid0) s.V[lane] = b0;
        if (valid1) s.V[pos1] = b1;
        __syncwarp();
        int r0 = 0, r1 = 0;
        bool tie = false;
        for (int q = 0; q < n; ++q) {
            const unsigned long long vq = s.V[q];
            r0 += (vq < b0) || (vq == b0 && q < lane);
            r1 += (vq < b1) || (vq == b1 && q < pos1);
            tie |= (valid0 && vq == b0 && q != lane) || (valid1 && vq == b1 && q != pos1);
        }
        if (__any_sync(kFull, tie)) {
            // exact numba argsort order among tied scores
            qs_partition_phase(s, n, lane);
            const int P0 = valid0 ? s.P[lane] : 0, P1 = valid1 ? s.P[pos1] : 0;
            r0 = 0; r1 = 0;
            for (int q = 0; q < n; ++q) {
                const unsigned long long vq = s.V[q];
                const int Pq = s.P[q];
                r0 += (vq < b0) || (vq == b0 && Pq < P0);
                r1 += (vq < b1) || (vq == b1 && Pq < P1);
            }
            if (lane == 0) atomicAdd(p.status + 1, 1);
        }
        const int cut = n - k;
        if (valid0 && r0 >= cut) { keys[r0 - cut] = make_int2(A.e, A.n); tsp[r0 - cut] = A.ts; scp[r0 - cut] = vA; }
        if (valid1 && r1 >= cut) { keys[r1 - cut] = make_int2(e1, n1); tsp[r1 - cut] = ts1; scp[r1 - cut] = v1; }
        n = k;
    }
    if (lane == 0) {
        *reinterpret_cast<double *>(rec + 24 * k) = new_norm;
        *reinterpret_cast<int *>(rec + 24 * k + 8) = n;
    }
}

__device__ __forceinline__ bool reached(uint32_t done, uint32_t thr)
{
    return static_cast<int32_t>(done - thr) >= 0;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
tppr_span_kernel(const __grid_constant__ SpanParams p)
{
    __shared__ WarpScratch scratch[kWarpsPerBlock];
    const int lane = threadIdx.x & 31;
    WarpScratch &s = scratch[threadIdx.x >> 5];
    const unsigned long long total = static_cast<unsigned long long>(p.n_chunk) * p.n_tppr;
    const int k = p.k;

    for (;;) {
        uint32_t tk = 0;
        if (lane == 0) tk = atomicAdd(p.ticket, 1u) - p.ticket_base;
        tk = __shfl_sync(kFull, tk, 0);
        if (tk >= total) break;
        const int64_t i = tk / p.n_tppr;
        const int j = tk - i * p.n_tppr;
        const int sN = p.src[i], dN = p.dst[i];
        const bool has_neg = p.rows_per_edge == 3;
        const int gN = has_neg ? p.neg[i] : sN;
        const bool loop = sN == dN;
        const bool neg_own = has_neg && gN != sN && gN != dN;
        uint32_t *done = p.done[j];

        if (!((p.inst_mask >> j) & 1)) {
            // instance not selected: keep its counters in step with the shared thresholds
            if (lane == 0) {
                atomicAdd(done + sN, 1u);
                if (!loop) atomicAdd(done + dN, 1u);
                if (neg_own) atomicAdd(done + gN, 1u);
            }
            continue;
        }

        // ---- wait for the producers of the three vectors (and earlier readers of src/dst) ----
        const uint32_t ths = p.thr_s[i], thd = p.thr_d[i], thn = neg_own ? p.thr_n[i] : 0u;
        bool okS = false, okD = loop, okG = !neg_own;
        unsigned spins = 0;
        for (;;) {
            uint32_t a = 0, b = 0, c = 0;
            if (!okS) a = ld_acquire(done + sN);
            if (!okD) b = ld_acquire(done + dN);
            if (!okG) c = ld_acquire(done + gN);
            okS = okS || reached(a, ths);
            okD = okD || reached(b, thd);
            okG = okG || reached(c, thn);
            if (okS && okD && okG) break;
            if ((++spins & 0x3ff) == 0) {
                if (*reinterpret_cast<volatile int32_t *>(p.status) != 0) break;
                if (spins > (1u << 22)) { atomicExch(p.status, 1); break; }
            }
            __nanosleep(20);
        }

        char *table = p.tables[j];
        char *recS = table + static_cast<size_t>(sN) * p.stride;
        char *recD = table + static_cast<size_t>(dN) * p.stride;
        const VecR S = load_vec(recS, k, lane);
        const VecR D = loop ? S : load_vec(recD, k, lane);
        const double t = p.ts[i];
        const int e = p.eidx[i];

        if (p.rows_per_edge >= 2) {
            int64_t row = p.i_offset + i, nrow = p.n_total;
            char *obase = p.out;
            if (p.batch_size > 0) {
                const int64_t b = row / p.batch_size;
                row -= b * p.batch_size;
                nrow = p.n_total - b * p.batch_size;
                if (nrow > p.batch_size) nrow = p.batch_size;
                obase += static_cast<size_t>((b - p.first_batch) % p.out_slots) * p.slot_bytes;
            }
            const size_t plane = static_cast<size_t>(nrow) * p.rows_per_edge * k;
            obase += static_cast<size_t>(j) * plane * 16;
            emit_row(p, obase, plane, row, S, t, lane);
            emit_row(p, obase, plane, row + nrow, D, t, lane);
            if (has_neg) {
                if (neg_own) {
                    const VecR G = load_vec(table + static_cast<size_t>(gN) * p.stride, k, lane);
                    emit_row(p, obase, plane, row + 2 * nrow, G, t, lane);
                    __threadfence();               // the read of G is complete before it is released
                    __syncwarp();
                    if (lane == 0) atomicAdd(done + gN, 1u);
                } else {
                    emit_row(p, obase, plane, row + 2 * nrow, gN == sN ? S : D, t, lane);
                }
            }
        }

        const double alpha = p.alpha[j], beta = p.beta[j];
        update_endpoint(p, s, S, D, dN, e, t, alpha, beta, recS, lane);
        if (!loop) update_endpoint(p, s, D, S, sN, e, t, alpha, beta, recD, lane);

        __threadfence();
        __syncwarp();
        if (lane == 0) {
            atomicAdd(done + sN, 1u);
            if (!loop) atomicAdd(done + dN, 1u);
        }
    }
}

__global__ void tppr_read_state_kernel(const char *table, int32_t stride, int32_t k, const int64_t *ids, int64_t n_ids,
                                       int32_t *counts, int32_t *eidx, int32_t *node, double *ts, double *score,
                                       double *norm)
{
    int64_t a = blockIdx.x;
    if (a >= n_ids) return;
    const char *rec = table + static_cast<size_t>(ids[a]) * stride;
    int l = threadIdx.x;
    if (l < k) {
        int2 kk = reinterpret_cast<const int2 *>(rec)[l];
        eidx[a * k + l] = kk.x; node[a * k + l] = kk.y;
        ts[a * k + l] = reinterpret_cast<const double *>(rec + 8 * k)[l];
        score[a * k + l] = reinterpret_cast<const double *>(rec + 16 * k)[l];
    }
    if (l == 0) {
        norm[a] = *reinterpret_cast<const double *>(rec + 24 * k);
        counts[a] = *reinterpret_cast<const int *>(rec + 24 * k + 8);
    }
}

}  // namespace zebra
