// tppr_device.cuh — device side of the streaming T-PPR span kernel (sm_100a).
//
// One warp owns one work item = (edge i, instance j) of a chronological span.  Items are claimed
// in stream order from a global ticket, so every dependency of an item is held by a warp that is
// already running; the barrier between the reference's conflict-free "waves" becomes point-to-point
// dataflow over the same DAG.
//
// Synchronisation is carried by the data itself.  A node record is a header and 2k 16-byte chunks,
// every one of which ends in a 32-bit tag = sequence number of the edge that wrote it.  A reader
// knows which edge wrote the version it needs (the host passes the index of the previous writer of
// each endpoint) and simply re-reads the record until every chunk carries that tag: one L2 round
// trip from the writer's stores to the reader's registers, no fence on the writer, no flag.
// Writers wait only for the (rare) negative-sample readers of the version they overwrite.
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
constexpr uint32_t kNear = 192;          // edges: closer than this to our turn -> poll the whole record

// record layout: hdr {f64 norm, i32 count, u32 tag} | A[k] {f64 score, i32 eidx, u32 tag} |
//                B[k] {f64 ts, i32 node, u32 tag} | pad to 128 B
__host__ __device__ inline size_t record_stride(int k) { return (static_cast<size_t>(32) * k + 16 + 127) / 128 * 128; }

enum : uint32_t { kFlagSignal = 1u, kFlagDstFirst = 2u };

struct SpanParams {
    int64_t n_chunk;        // edges in this launch
    int64_t n_total;        // edges of the whole span (row stride of the output)
    int64_t i_offset;       // index of this launch's first edge inside the span
    int32_t k, n_tppr, rows_per_edge, inst_mask;
    int32_t stride;         // bytes per node record
    uint32_t ticket_base;
    uint32_t seq0;          // tag of launch-local edge i = seq0 + tag_offset + i; a reader expects seq0 + prev
    int64_t tag_offset;     // index base of the prev_* arrays relative to this launch's first edge
    const int32_t *src, *dst, *neg, *eidx;          // indexed by launch-local edge
    const int32_t *prev_s, *prev_d, *prev_n;        // span index of the previous writer, -1 = none
    const uint32_t *rthr_s, *rthr_d, *flags;        // negative reads to wait for before overwriting
    const double *ts;
    char *tables[kMaxInst];
    int32_t *rcnt[kMaxInst];                        // [num_nodes] completed-but-unconsumed negative reads
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

__device__ __forceinline__ uint4 ld_chunk(const char *p)
{
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_chunk(char *p, double d, int i, uint32_t tag)
{
    const unsigned long long b = static_cast<unsigned long long>(__double_as_longlong(d));
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(static_cast<uint32_t>(b)),
                 "r"(static_cast<uint32_t>(b >> 32)), "r"(i), "r"(tag) : "memory");
}

__device__ __forceinline__ double chunk_f64(const uint4 &c)
{
    return __longlong_as_double(static_cast<long long>((static_cast<unsigned long long>(c.y) << 32) | c.x));
}

__device__ __forceinline__ int32_t ld_relaxed_s32(const int32_t *p)
{
    int32_t v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One attempt to read a whole record.  `any`: the record has no writer earlier in this span, so it
// is stable and any self-consistent content is the wanted version; otherwise every chunk must carry
// `expect`.  `cur` returns the header tag seen (progress estimate for the back-off).
__device__ __forceinline__ bool try_load(const char *rec, int k, int lane, bool any, uint32_t expect, VecR &v, uint32_t &cur)
{
    const int l = lane < k ? lane : k - 1;
    const uint4 h = ld_chunk(rec);
    const uint4 a = ld_chunk(rec + 16 + 16 * l);
    const uint4 b = ld_chunk(rec + 16 + 16 * (k + l));
    cur = h.w;
    const uint32_t want = any ? h.w : expect;
    const int cnt = static_cast<int>(h.z);
    const bool ok = h.w == want && (lane >= cnt || (a.w == want && b.w == want));
    v.sc = chunk_f64(a); v.e = static_cast<int>(a.z);
    v.ts = chunk_f64(b); v.n = static_cast<int>(b.z);
    v.norm = chunk_f64(h); v.cnt = cnt;
    return __all_sync(kFull, ok);
}

__device__ __forceinline__ uint32_t ld_tag(const char *rec) { return ld_chunk(rec).w; }

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


__device__ __forceinline__ void st_entry(char *rec, int k, int pos, int e, int n, double ts, double sc, uint32_t tag)
{
    st_chunk(rec + 16 + 16 * pos, sc, e, tag);
    st_chunk(rec + 16 + 16 * (k + pos), ts, n, tag);
}

// New vector of one endpoint from the OLD vectors of itself (A) and its peer (B)
// (utils/util.py:509-559), written straight into the endpoint's record with tag `tag`.
// `rwait` > 0: that many negative-sample reads of the old version must have finished first.
__device__ __forceinline__ bool update_endpoint(const SpanParams &p, WarpScratch &s, const VecR &A, const VecR &B,
                                                int peer_id, int e_new, double t_new, double alpha, double beta,
                                                char *rec, uint32_t tag, int32_t *rcnt, uint32_t rwait, int lane)
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
    int pos0 = lane;
    bool keep0 = valid0, keep1 = valid1;

    if (n > k) {
        // ---- keep the k largest, re-inserted in ascending order (util.py:553-559) ----
        const unsigned long long b0 = order_bits(vA);
        const unsigned long long b1 = order_bits(v1);
        __syncwarp();
        if (valid0) s.V[lane] = b0;
        if (valid1) s.V[pos1] = b1;
        __syncwarp();
        int r0 = 0, r1 = 0;
        bool eq0 = false, eq1 = false;
        for (int q = 0; q < n; ++q) {
            const unsigned long long vq = s.V[q];
            r0 += (vq < b0) || (vq == b0 && q < lane);
            r1 += (vq < b1) || (vq == b1 && q < pos1);
            eq0 |= vq == b0 && q != lane;
            eq1 |= vq == b1 && q != pos1;
        }
        const int cut = n - k;
        // a tie only matters if one of the tied candidates survives under the stable order: then
        // the survivor set or its slot order depends on numba's (unstable) argsort
        if (__any_sync(kFull, (valid0 && eq0 && r0 >= cut) || (valid1 && eq1 && r1 >= cut))) {
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
        keep0 = valid0 && r0 >= cut; pos0 = r0 - cut;
        keep1 = valid1 && r1 >= cut; pos1 = r1 - cut;
        n = k;
    }

    // ---- the version being overwritten may still have negative-sample readers ----
    bool ok = true;
    if (rwait != 0) {
        unsigned spins = 0;
        while (ld_relaxed_s32(rcnt) < static_cast<int32_t>(rwait)) {
            if ((++spins & 0xff) == 0) {
                if (*reinterpret_cast<volatile int32_t *>(p.status) != 0 || spins > (1u << 21)) { ok = false; break; }
            }
            __nanosleep(64);
        }
        if (lane == 0) atomicSub(rcnt, static_cast<int32_t>(rwait));
    }
    if (keep0) st_entry(rec, k, pos0, A.e, A.n, A.ts, vA, tag);
    if (keep1) st_entry(rec, k, pos1, e1, n1, ts1, v1, tag);
    __syncwarp();
    if (lane == 0) st_chunk(rec, new_norm, n, tag);
    return ok;
}

// how long to sleep before looking again at a record whose wanted version is `dist` edges away
__device__ __forceinline__ unsigned backoff_ns(uint32_t dist)
{
    const unsigned ns = dist * 8u;
    return ns > 4000u ? 4000u : ns;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
tppr_span_kernel(const __grid_constant__ SpanParams p)
{
    __shared__ WarpScratch scratch[kWarpsPerBlock];
    const int lane = threadIdx.x & 31;
    WarpScratch &s = scratch[threadIdx.x >> 5];
    const unsigned long long total = static_cast<unsigned long long>(p.n_chunk) * p.n_tppr;
    const int k = p.k;
    const uint32_t tag_base = p.seq0 + static_cast<uint32_t>(p.tag_offset);   // tag of launch-local edge 0

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
        const uint32_t flags = p.flags[i];
        int32_t *rcnt = p.rcnt[j];

        if (!((p.inst_mask >> j) & 1)) continue;   // instance not selected: its tables are not touched

        const int32_t pS = p.prev_s[i], pD = loop ? -1 : p.prev_d[i], pG = neg_own ? p.prev_n[i] : -1;
        char *table = p.tables[j];
        char *recS = table + static_cast<size_t>(sN) * p.stride;
        char *recD = table + static_cast<size_t>(dN) * p.stride;
        const char *recG = table + static_cast<size_t>(gN) * p.stride;
        const uint32_t my_tag = tag_base + static_cast<uint32_t>(i);

        // ---- read the three vectors; wait for the producers of src and dst (the chain) ----
        VecR S, D, G;
        uint32_t curS = 0, curD = 0, curG = 0;
        bool okS = try_load(recS, k, lane, pS < 0, p.seq0 + static_cast<uint32_t>(pS), S, curS);
        bool okD = loop ? true : try_load(recD, k, lane, pD < 0, p.seq0 + static_cast<uint32_t>(pD), D, curD);
        bool okG = neg_own ? try_load(recG, k, lane, pG < 0, p.seq0 + static_cast<uint32_t>(pG), G, curG) : true;
        bool alive = true;
        unsigned spins = 0;
        while (!(okS && okD)) {
            // distance (in edges) of the farther unsatisfied producer from the version visible now
            uint32_t dist = 0;
            if (!okS) {
                const uint32_t d = curS - tag_base, pe = static_cast<uint32_t>(pS - p.tag_offset);
                dist = (d < static_cast<uint32_t>(p.n_chunk) && d <= pe) ? pe - d : pe + 1;
            }
            if (!okD) {
                const uint32_t d = curD - tag_base, pe = static_cast<uint32_t>(pD - p.tag_offset);
                const uint32_t dd = (d < static_cast<uint32_t>(p.n_chunk) && d <= pe) ? pe - d : pe + 1;
                dist = dd > dist ? dd : dist;
            }
            if (dist > kNear) {
                // far from our turn: sleep, then look at the header tags only
                __nanosleep(backoff_ns(dist));
                if (!okS) curS = ld_tag(recS);
                if (!okD) curD = ld_tag(recD);
                const bool nearS = okS || curS == p.seq0 + static_cast<uint32_t>(pS);
                const bool nearD = okD || curD == p.seq0 + static_cast<uint32_t>(pD);
                if (!(nearS && nearD)) {
                    if ((++spins & 0xff) == 0) {
                        if (*reinterpret_cast<volatile int32_t *>(p.status) != 0) { alive = false; break; }
                        if (spins > (1u << 20)) { atomicExch(p.status, 1); alive = false; break; }
                    }
                    continue;
                }
            } else {
                __nanosleep(32);
            }
            if (!okS) okS = try_load(recS, k, lane, false, p.seq0 + static_cast<uint32_t>(pS), S, curS);
            if (!okD) okD = try_load(recD, k, lane, false, p.seq0 + static_cast<uint32_t>(pD), D, curD);
            if (!okG) okG = try_load(recG, k, lane, false, p.seq0 + static_cast<uint32_t>(pG), G, curG);
            if ((++spins & 0xff) == 0) {
                if (*reinterpret_cast<volatile int32_t *>(p.status) != 0) { alive = false; break; }
                if (spins > (1u << 22)) { atomicExch(p.status, 1); alive = false; break; }
            }
        }
        if (!alive) continue;
        if (loop) D = S;
        if (okG && neg_own && (flags & kFlagSignal) && lane == 0) atomicAdd(rcnt + gN, 1);
        bool sigG = okG;

        const double t = p.ts[i];
        const int e = p.eidx[i];
        const double alpha = p.alpha[j], beta = p.beta[j];
        const uint32_t rwS = p.rthr_s[i], rwD = loop ? 0u : p.rthr_d[i];

        // ---- the hotter endpoint first: its next writer is the one most likely to be waiting ----
        bool ok = true;
        if ((flags & kFlagDstFirst) && !loop) {
            ok &= update_endpoint(p, s, D, S, sN, e, t, alpha, beta, recD, my_tag, rcnt + dN, rwD, lane);
            ok &= update_endpoint(p, s, S, D, dN, e, t, alpha, beta, recS, my_tag, rcnt + sN, rwS, lane);
        } else {
            ok &= update_endpoint(p, s, S, D, dN, e, t, alpha, beta, recS, my_tag, rcnt + sN, rwS, lane);
            if (!loop) ok &= update_endpoint(p, s, D, S, sN, e, t, alpha, beta, recD, my_tag, rcnt + dN, rwD, lane);
        }
        if (!ok) { atomicExch(p.status, 1); continue; }

        // ---- off the chain: the negative's vector (if it was not ready yet) and the output rows ----
        while (!okG) {
            __nanosleep(64);
            okG = try_load(recG, k, lane, false, p.seq0 + static_cast<uint32_t>(pG), G, curG);
            if ((++spins & 0xff) == 0) {
                if (*reinterpret_cast<volatile int32_t *>(p.status) != 0) break;
                if (spins > (1u << 22)) { atomicExch(p.status, 1); break; }
            }
        }
        if (!okG) continue;
        if (!sigG && neg_own && (flags & kFlagSignal) && lane == 0) atomicAdd(rcnt + gN, 1);

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
            if (has_neg) emit_row(p, obase, plane, row + 2 * nrow, neg_own ? G : (gN == sN ? S : D), t, lane);
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
        const uint4 ca = *reinterpret_cast<const uint4 *>(rec + 16 + 16 * l);
        const uint4 cb = *reinterpret_cast<const uint4 *>(rec + 16 + 16 * (k + l));
        eidx[a * k + l] = static_cast<int32_t>(ca.z); node[a * k + l] = static_cast<int32_t>(cb.z);
        ts[a * k + l] = chunk_f64(cb);
        score[a * k + l] = chunk_f64(ca);
    }
    if (l == 0) {
        const uint4 h = *reinterpret_cast<const uint4 *>(rec);
        norm[a] = chunk_f64(h);
        counts[a] = static_cast<int32_t>(h.z);
    }
}

// Sets every tag of a table to zero (used when the 32-bit edge sequence wraps).
__global__ void tppr_retag_kernel(char *table, int32_t stride, int32_t k, int64_t num_nodes)
{
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int per = 2 * k + 1;
    const int64_t node = idx / per;
    if (node >= num_nodes) return;
    const int c = static_cast<int>(idx - node * per);
    *reinterpret_cast<uint32_t *>(table + static_cast<size_t>(node) * stride + 16 * c + 12) = 0u;
}

}  // namespace zebra
