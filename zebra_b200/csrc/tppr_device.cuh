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

#ifndef ZEBRA_ABLATE
#define ZEBRA_ABLATE 0   // developer timing builds only: bit 0 no duplicate scan, bit 1 no rank loop, bit 2 no tie path
#endif
#ifdef ZEBRA_PROFILE
#define ZP_DECL long long zp_t0 = clock64(), zp_t1
#define ZP_MARK(slot) do { zp_t1 = clock64(); if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long *>(p.prof) + (slot), static_cast<unsigned long long>(zp_t1 - zp_t0)); zp_t0 = zp_t1; } while (0)
#define ZP_MARK_IF(c, slot) do { zp_t1 = clock64(); if ((c) && lane == 0) atomicAdd(reinterpret_cast<unsigned long long *>(p.prof) + (slot), static_cast<unsigned long long>(zp_t1 - zp_t0)); zp_t0 = zp_t1; } while (0)
#else
#define ZP_DECL
#define ZP_MARK(slot)
#define ZP_MARK_IF(c, slot)
#endif

constexpr int kMaxInst = 8;
constexpr int kMaxRanks = 8;
constexpr int kWarpsPerBlock = 8;
constexpr unsigned kFull = 0xffffffffu;
constexpr uint32_t kNear = 192;          // edges: closer than this to our turn -> poll the whole record

// record layout: hdr {f64 norm, i32 count, u32 tag} | A[k] {f64 score, i32 eidx, u32 tag} |
//                B[k] {f64 ts, i32 node, u32 tag} | pad to 128 B
__host__ __device__ inline size_t record_stride(int k) { return (static_cast<size_t>(32) * k + 16 + 127) / 128 * 128; }

enum : uint32_t { kFlagSignal = 1u, kFlagDstFirst = 2u,
                  kFlagSrcChain = 4u, kFlagDstChain = 8u };   // that endpoint is updated by a resident chain warp
constexpr int kFlagSrcChainShift = 8, kFlagDstChainShift = 13;   // 5 bits each: which chain

// ---- hub chains -------------------------------------------------------------------------------
// A node with many edges in one launch serialises them through its vector: every update needs the
// previous one (the top-k prune is a non-linear recurrence).  With one warp per work item that chain
// pays, per step, an L2 round trip (writer's stores -> next reader's loads) plus the update of BOTH
// endpoints.  The host therefore marks the hottest nodes of a span as CHAIN nodes; each (chain node,
// instance) gets a resident warp that walks the node's edges in order with the vector in REGISTERS and
// computes only that node's side.  Everything else of such an edge (the partner's update, the three
// output rows) stays with the ordinary ticket warp, which receives the chain node's OLD vector through
// a launch-local mailbox slot (tagged 16-byte chunks, pushed from the chain warp's registers before it
// computes).  Every version is still published to the node's table record, so negative-sample readers
// and later launches see exactly what they saw before.
constexpr int kMaxChains = 24;
constexpr int kSmSlots = 512;            // SM ids are below this (148 on B200)
struct alignas(16) ChainStep {
    int32_t edge;        // index of the edge in the span the tags are drawn from (tag = seq0 + edge)
    int32_t peer;        // the other endpoint
    int32_t eidx;
    int32_t prev_peer;   // previous writer of the peer (span index), -1 = none
    uint32_t rthr;       // negative reads of the version this step overwrites
    uint32_t flags;      // kStepDst | kStepLoop | kStepPeerChain
    int32_t cidx;        // chain item index of the edge (mailbox slot = cidx - cidx_base)
    int32_t prev_self;   // previous writer of the chain node itself (used by the first step of a launch)
    double ts;
    int64_t rec;         // index of the edge's staging record in the image (node-sharded plans hold compact records)
};
enum : uint32_t { kStepDst = 1u, kStepLoop = 2u, kStepPeerChain = 4u,
                  kStepPeerRemote = 8u };     // node-sharded: the partner lives on another rank (its vector arrives by NVLink push)
struct ChainDesc { int32_t node, begin, end, owned; };   // steps [begin, end) of SpanParams::csteps; owned: this GPU runs the chain warp

// The norm of a node depends only on how many updates it has seen (norm <- norm*beta + beta,
// utils/util.py:522,574), so the factors of its next update are tabulated per update count instead
// of being divided out on the dependency chain.  Entry m serves the update that takes a node from m
// to m+1 updates; the recurrence reaches a floating-point fixed point, where the table ends.
struct DecayStep { double new_norm, s1, s2, w_new; };
constexpr int kCountBits = 6;            // header word: entries in use | update count << 6 (saturating)
constexpr int kMaxUpdates = (1 << 26) - 1;

__host__ __device__ inline DecayStep decay_step(double norm, double alpha, double beta)
{
    DecayStep d;
#ifdef __CUDA_ARCH__
    d.new_norm = __dadd_rn(__dmul_rn(norm, beta), beta);
    d.s1 = __dmul_rn(__ddiv_rn(norm, d.new_norm), beta);
    d.s2 = __dmul_rn(__ddiv_rn(beta, d.new_norm), __dsub_rn(1.0, alpha));
    d.w_new = (alpha != 0.0) ? __dmul_rn(d.s2, alpha) : d.s2;
#else
    d.new_norm = norm * beta + beta;          // host build uses -ffp-contract=off: same roundings
    d.s1 = norm / d.new_norm * beta;
    d.s2 = beta / d.new_norm * (1.0 - alpha);
    d.w_new = (alpha != 0.0) ? d.s2 * alpha : d.s2;
#endif
    return d;
}

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
    const DecayStep *decay[kMaxInst];               // per-instance table of update factors by update count
    int32_t decay_len[kMaxInst];                    // entries; the last one is the fixed point if decay_fixed
    int32_t decay_fixed[kMaxInst];
    char *out;              // [n_tppr][4][rows_total*k] 4-byte elements, or nullptr
    // batch-structured output (plan mode): batch_size > 0 => edge g = i_offset + i belongs to batch
    // b = g / batch_size and its rows go to slot (b - first_batch) % out_slots of `out`, laid out
    // as the reference lays out that batch: [n_tppr][4][rows_per_edge*nb*k], nb = edges in batch b
    int64_t batch_size, first_batch, out_slots, slot_bytes;
    uint32_t *ticket;
    int32_t *status;        // [0] spin time-out flag, [1] exact-tie prunes, [2] merge-ranked prunes (statistics)
    unsigned long long *prof;   // [8] cycle sums per phase (ZEBRA_PROFILE builds)
    // node-sharded mode (tppr_shard_device.cuh): node v lives on rank v & (world-1), local row v >> world_log2
    int32_t world_log2, rank, mbox_parity;
    int64_t rec_offset;                             // node-sharded: staging record of a chain step = ChainStep::rec - rec_offset
    int32_t shard_launch;                           // index of this sharded launch on the handle (same on every rank)
    const int32_t *shard_done;                      // own arena header: [r] = launches rank r has finished
    int32_t row_guard;                              // pipelined epoch: wait for the consumer rank before storing its rows
    int32_t row_mode;                               // 0: rows to every rank; 1: rows of batch b to rank b & (world-1) only
    const int32_t *list;                            // sharded: span indices of the edges this rank takes part in (this launch's range)
    int64_t n_list;
    int64_t mbox_half_bytes;                        // bytes of one parity half of a mailbox
    char *mbox_peer[kMaxRanks];                     // every rank's mailbox (peer-mapped), own included
    char *out_peer[kMaxRanks];                      // every rank's output region (peer-mapped), own included
    // hub chains (single-GPU kernel)
    const int32_t *cidx;                            // per launch-local edge: chain item index, -1 = no chain endpoint
    const ChainStep *csteps;
    char *cmbox;                                    // [(cidx - cidx_base) * n_tppr + j][2 sides] records
    int32_t cidx_base, n_chains;
    int32_t tie_mode;                               // 0: numba's argsort order inside groups of equal scores (exact); 1: insertion order
    int32_t *smflag;                                // [kSmSlots] running chain warps per SM, [kSmSlots] ticket warps yielding
    int32_t max_yield;                              // ticket warps that may yield at the same time
    uint32_t *dbg_owner;                            // debug (ZEBRA_DEBUG_ITEMS=1): [edge][instance] who processed the item
    uint32_t *cprog;                                // [chain][instance] tag of the step the chain warp is working on
    unsigned long long wait_ns;                     // dependency-wait time-out (globaltimer)
    ChainDesc chains[kMaxChains];
};

constexpr int kKeySlotBits = 11, kKeySlots = 1 << kKeySlotBits;

__device__ __forceinline__ unsigned key_slot(int e, int n)
{
    return (static_cast<unsigned>(e) * 0x9E3779B1u ^ static_cast<unsigned>(n) * 0x85EBCA6Bu) >> (32 - kKeySlotBits);
}

struct alignas(16) WarpScratch {
    unsigned long long V[64];   // candidate scores by candidate index (fp64 bit patterns in integer order)
    unsigned long long NK[32], NT[32], NS[32];   // chain warps: the new vector by slot (key, ts, score)
    unsigned long long K[32];   // peer keys of the duplicate scan
    // exact-tie path: numba's quicksort is replayed on (rank << 8 | candidate) items, the rank (number
    // of strictly smaller scores) being an order-isomorphic 6-bit stand-in for the fp64 score
    unsigned Mlo[64], Mhi[64];  // per rank: bitmap of the sort positions holding a candidate of that rank
    unsigned short Q[64];       // sort position -> item (numba quicksort working array)
    unsigned short X[32], Y[32];   // swap exchange
    unsigned char P[64];        // candidate -> sort position
    unsigned char T[kKeySlots]; // duplicate lookup: slot of a peer key -> lane holding it, 0xff = empty
};

struct VecR {
    int e, n;        // key parts 0,1 of this lane's entry
    double ts, sc;   // key part 2, value
    int cnt;         // entries in use
    int upd;         // updates seen (saturating)
    double norm;
};

// field-wise select: keeps the vectors in registers (a reference chosen at run time would not)
__device__ __forceinline__ VecR vsel(bool c, const VecR &a, const VecR &b)
{
    VecR r;
    r.e = c ? a.e : b.e; r.n = c ? a.n : b.n; r.ts = c ? a.ts : b.ts; r.sc = c ? a.sc : b.sc;
    r.cnt = c ? a.cnt : b.cnt; r.upd = c ? a.upd : b.upd; r.norm = c ? a.norm : b.norm;
    return r;
}

// Tagged chunks are read and written with the scalar 128-bit forms (ld/st.relaxed.*.b128, PTX ISA 8.3):
// the PTX memory model defines a .b128 access as ONE access, so a chunk's tag can never be seen with
// a payload of another version (a .v4.u32 access is formally four scalar accesses in unspecified order;
// both compile to LDG/STG.E.128.STRONG on sm_100a).
template <bool kSys = false>
__device__ __forceinline__ uint4 ld_chunk(const char *p)
{
    unsigned long long lo, hi;
    if (kSys)   // memory another GPU writes into (mailboxes of the node-sharded mode)
        asm volatile("{\n\t.reg .b128 r;\n\tld.relaxed.sys.global.b128 r, [%2];\n\tmov.b128 {%0, %1}, r;\n\t}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    else
        asm volatile("{\n\t.reg .b128 r;\n\tld.relaxed.gpu.global.b128 r, [%2];\n\tmov.b128 {%0, %1}, r;\n\t}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
    uint4 v;
    v.x = static_cast<uint32_t>(lo); v.y = static_cast<uint32_t>(lo >> 32);
    v.z = static_cast<uint32_t>(hi); v.w = static_cast<uint32_t>(hi >> 32);
    return v;
}

template <bool kSys = false>
__device__ __forceinline__ void st_chunk_t(char *p, double d, int i, uint32_t tag)
{
    const unsigned long long lo = static_cast<unsigned long long>(__double_as_longlong(d));
    const unsigned long long hi = (static_cast<unsigned long long>(tag) << 32) | static_cast<uint32_t>(i);
    if (kSys)
        asm volatile("{\n\t.reg .b128 r;\n\tmov.b128 r, {%1, %2};\n\tst.relaxed.sys.global.b128 [%0], r;\n\t}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
    else
        asm volatile("{\n\t.reg .b128 r;\n\tmov.b128 r, {%1, %2};\n\tst.relaxed.gpu.global.b128 [%0], r;\n\t}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}

__device__ __forceinline__ void st_chunk(char *p, double d, int i, uint32_t tag) { st_chunk_t<false>(p, d, i, tag); }

__device__ __forceinline__ double chunk_f64(const uint4 &c)
{
    return __longlong_as_double(static_cast<long long>((static_cast<unsigned long long>(c.y) << 32) | c.x));
}

__device__ __forceinline__ int32_t ld_relaxed_sys_s32(const int32_t *p)
{
    int32_t v;
    asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ int32_t ld_relaxed_s32(const int32_t *p)
{
    int32_t v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Bounded waiting: every 256 polls look at the abort flag and at the clock.  The limit is wall time
// (SpanParams::wait_ns, ZEBRA_WAIT_TIMEOUT_MS), not a poll count: a peer GPU that launches late or a
// long producer chain must not be mistaken for a dead-lock.
constexpr int kWaitLog = 96;   // abandoned waits listed in status[8 ...] (4 words each) after a time-out
struct Waiter {
    unsigned spins = 0;
    unsigned long long t0 = 0;
    // `code`, `a`, `b`: what was being waited for, recorded in status[4..6] by the first wait that expires
    // (1 table record of src/dst, 2 chain mailbox slot, 3 negative-sample readers, 4 negative's record,
    //  5 chain: own record, 6 chain: partner chain's push, 7 chain: partner's record, 8 peer GPU's push,
    //  9 consumer rank has not copied its rows out yet, 10 a rank has not finished the launch)
    __device__ __forceinline__ void log_wait(int32_t *status, int code, uint32_t a, uint32_t b, int32_t c)
    {
        if ((threadIdx.x & 31) != 0) return;
        const int idx = atomicAdd(status + 7, 1);          // every abandoned wait is listed (first kWaitLog of them)
        if (idx < kWaitLog) {
            int32_t *e = status + 8 + 4 * idx;
            e[0] = code; e[1] = static_cast<int32_t>(a); e[2] = static_cast<int32_t>(b);
            e[3] = c;
        }
    }
    __device__ __forceinline__ bool expired(int32_t *status, unsigned long long limit, int code = 0, uint32_t a = 0, uint32_t b = 0, int32_t c = -1)
    {
        if ((++spins & 0xff) != 0) return false;
        if (*reinterpret_cast<const volatile int32_t *>(status) != 0) { log_wait(status, code, a, b, c); return true; }
        const unsigned long long now = gtime();
        if (t0 == 0) t0 = now;
        else if (now - t0 > limit) {
            if ((threadIdx.x & 31) == 0 && atomicExch(status, 1) == 0) {
                status[4] = code; status[5] = static_cast<int32_t>(a); status[6] = static_cast<int32_t>(b);
            }
            log_wait(status, code, a, b, c);
            return true;
        }
        return false;
    }
};

// One attempt to read a whole record.  `any`: the record has no writer earlier in this span, so it
// is stable and any self-consistent content is the wanted version; otherwise every chunk must carry
// `expect`.  `cur` returns the header tag seen (progress estimate for the back-off).
struct RawRec { uint4 h, a, b; };   // header and this lane's two chunks, as loaded

template <bool kSys = false>
__device__ __forceinline__ RawRec load_raw(const char *rec, int k, int lane)
{
    const int l = lane < k ? lane : k - 1;
    RawRec r;
    r.h = ld_chunk<kSys>(rec);
    r.a = ld_chunk<kSys>(rec + 16 + 16 * l);
    r.b = ld_chunk<kSys>(rec + 16 + 16 * (k + l));
    return r;
}

__device__ __forceinline__ bool check_raw(const RawRec &r, int lane, bool any, uint32_t expect, VecR &v, uint32_t &cur)
{
    cur = r.h.w;
    const uint32_t want = any ? r.h.w : expect;
    const int cnt = static_cast<int>(r.h.z & ((1u << kCountBits) - 1));
    const bool ok = r.h.w == want && (lane >= cnt || (r.a.w == want && r.b.w == want));
    v.sc = chunk_f64(r.a); v.e = static_cast<int>(r.a.z);
    v.ts = chunk_f64(r.b); v.n = static_cast<int>(r.b.z);
    v.norm = chunk_f64(r.h); v.cnt = cnt; v.upd = static_cast<int>(r.h.z >> kCountBits);
    return __all_sync(kFull, ok);
}

template <bool kSys = false>
__device__ __forceinline__ bool try_load(const char *rec, int k, int lane, bool any, uint32_t expect, VecR &v, uint32_t &cur)
{
    return check_raw(load_raw<kSys>(rec, k, lane), lane, any, expect, v, cur);
}

// Copy of a vector held in registers into a mailbox slot (a peer GPU's: kSys; a chain item's: gpu scope).
template <bool kSys>
__device__ __forceinline__ void push_vector(char *slot, int k, const VecR &v, uint32_t tag, int lane)
{
    if (lane == 0) st_chunk_t<kSys>(slot, v.norm, v.cnt | (v.upd << kCountBits), tag);
    if (lane < v.cnt) {
        st_chunk_t<kSys>(slot + 16 + 16 * lane, v.sc, v.e, tag);
        st_chunk_t<kSys>(slot + 16 + 16 * (k + lane), v.ts, v.n, tag);
    }
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
// ---- exact ties: numba's argsort replayed -------------------------------------------------------
// np.argsort(values)[-k:] (utils/util.py:555) runs numba.misc.quicksort: median-of-three Hoare
// partitions while a range is longer than 15, then one insertion sort.  The insertion sort is stable,
// so the order inside a group of equal scores is their order after the partition phase.  The phase is
// replayed on 16-bit items (rank << 8 | candidate) — the rank (number of strictly smaller scores) is an
// order-isomorphic stand-in for the fp64 score.  (A register-resident variant that partitions two
// pending ranges per round was measured 20-35 % slower on the hub chain — more selects and 64-bit mask
// arithmetic per round than the two steps it replaces — and was dropped.)

// One partition step of numba.misc.quicksort (median of three, pivot stashed at `high`, Hoare
// scan) on positions [low, high] of the working array Q (items rank << 8 | candidate, compared by
// rank only).  Computed warp-wide from two stopper bitmaps instead of a sequential scan; two warp
// barriers and three shared-memory latencies per step.  Returns the pivot's position.
__device__ __forceinline__ int qs_partition(WarpScratch &s, int low, int high, int lane)
{
    const int mid = (low + high) >> 1;
    const int p0 = lane, p1 = lane + 32;
    unsigned il = s.Q[low], im = s.Q[mid], ih = s.Q[high];
    unsigned q0 = s.Q[p0], q1 = s.Q[p1];
    if ((im >> 8) < (il >> 8)) { const unsigned t = il; il = im; im = t; }
    if ((ih >> 8) < (im >> 8)) { const unsigned t = ih; ih = im; im = t; }
    if ((im >> 8) < (il >> 8)) { const unsigned t = il; il = im; im = t; }
    const unsigned pivot = im >> 8;
    // what the scan sees: low <- smallest, mid <- largest, high <- pivot (excluded from the scan)
    if (p0 == low) q0 = il; else if (p0 == mid) q0 = ih;
    if (p1 == low) q1 = il; else if (p1 == mid) q1 = ih;
    const bool in0 = p0 >= low && p0 < high, in1 = p1 >= low && p1 < high;
    const bool L0 = in0 && (q0 >> 8) >= pivot, L1 = in1 && (q1 >> 8) >= pivot;   // i stops here
    const bool Q0 = in0 && (q0 >> 8) <= pivot, Q1 = in1 && (q1 >> 8) <= pivot;   // j stops here
    const unsigned long long Lm = __ballot_sync(kFull, L0) | (static_cast<unsigned long long>(__ballot_sync(kFull, L1)) << 32);
    const unsigned long long Qm = __ballot_sync(kFull, Q0) | (static_cast<unsigned long long>(__ballot_sync(kFull, Q1)) << 32);
    const int lb0 = __popcll(Lm & ((1ull << p0) - 1)), lb1 = __popcll(Lm & ((1ull << p1) - 1));
    const int qa0 = __popcll(Qm >> (p0 + 1)), qa1 = (p1 == 63) ? 0 : __popcll(Qm >> (p1 + 1));
    const bool swL0 = L0 && qa0 > lb0, swL1 = L1 && qa1 > lb1;
    const bool swQ0 = Q0 && lb0 > qa0, swQ1 = Q1 && lb1 > qa1;
    if (swL0) s.X[lb0] = static_cast<unsigned short>(q0);
    if (swL1) s.X[lb1] = static_cast<unsigned short>(q1);
    if (swQ0) s.Y[qa0] = static_cast<unsigned short>(q0);
    if (swQ1) s.Y[qa1] = static_cast<unsigned short>(q1);
    // M = number of swaps; the swapped left stoppers are exactly those with lb < M
    const unsigned long long swLm = __ballot_sync(kFull, swL0) | (static_cast<unsigned long long>(__ballot_sync(kFull, swL1)) << 32);
    const int M = __popcll(swLm);
    // first unswapped left stopper (rank M) and the last swapped right stopper (rank M-1)
    const unsigned long long lM = __ballot_sync(kFull, L0 && lb0 == M) |
                                  (static_cast<unsigned long long>(__ballot_sync(kFull, L1 && lb1 == M)) << 32);
    const unsigned long long qM = __ballot_sync(kFull, Q0 && qa0 == M - 1) |
                                  (static_cast<unsigned long long>(__ballot_sync(kFull, Q1 && qa1 == M - 1)) << 32);
    const int bound = (M > 0 && qM) ? (__ffsll(static_cast<long long>(qM)) - 1) : high;
    int i = lM ? (__ffsll(static_cast<long long>(lM)) - 1) : 64;
    if (bound < i) i = bound;
    __syncwarp();
    if (swL0) q0 = s.Y[lb0];
    if (swL1) q1 = s.Y[lb1];
    if (swQ0) q0 = s.X[qa0];
    if (swQ1) q1 = s.X[qa1];
    // the pivot lands on i, the element that was there moves to `high`
    if (in0) {
        if (p0 == i) { s.Q[high] = static_cast<unsigned short>(q0); s.Q[p0] = static_cast<unsigned short>(im); }
        else s.Q[p0] = static_cast<unsigned short>(q0);
    } else if (p0 == high && i == high) s.Q[p0] = static_cast<unsigned short>(im);
    if (in1) {
        if (p1 == i) { s.Q[high] = static_cast<unsigned short>(q1); s.Q[p1] = static_cast<unsigned short>(im); }
        else s.Q[p1] = static_cast<unsigned short>(q1);
    } else if (p1 == high && i == high) s.Q[p1] = static_cast<unsigned short>(im);
    __syncwarp();
    return i;
}

// The same step for a range of at most 32 positions: one position per lane (position = low + lane),
// 32-bit stopper bitmaps, half the votes and population counts of the two-position version.
__device__ __forceinline__ int qs_partition32(WarpScratch &s, int low, int high, int lane)
{
    const int mid = (low + high) >> 1;
    const int pos = low + lane;
    unsigned il = s.Q[low], im = s.Q[mid], ih = s.Q[high];
    unsigned q = s.Q[pos <= high ? pos : high];
    if ((im >> 8) < (il >> 8)) { const unsigned t = il; il = im; im = t; }
    if ((ih >> 8) < (im >> 8)) { const unsigned t = ih; ih = im; im = t; }
    if ((im >> 8) < (il >> 8)) { const unsigned t = il; il = im; im = t; }
    const unsigned pivot = im >> 8;
    if (pos == low) q = il; else if (pos == mid) q = ih;
    const bool in = pos < high;
    const bool L = in && (q >> 8) >= pivot;          // i stops here
    const bool R = in && (q >> 8) <= pivot;          // j stops here
    const unsigned Lm = __ballot_sync(kFull, L), Rm = __ballot_sync(kFull, R);
    const int lb = __popc(Lm & ((1u << lane) - 1));
    const int ra = __popc((Rm >> lane) >> 1);
    const bool swL = L && ra > lb, swR = R && lb > ra;
    if (swL) s.X[lb] = static_cast<unsigned short>(q);
    if (swR) s.Y[ra] = static_cast<unsigned short>(q);
    const int M = __popc(__ballot_sync(kFull, swL));
    const unsigned lM = __ballot_sync(kFull, L && lb == M);
    const unsigned rM = __ballot_sync(kFull, R && ra == M - 1);
    const int bound = (M > 0 && rM) ? low + __ffs(rM) - 1 : high;
    int i = lM ? low + __ffs(lM) - 1 : 64;
    if (bound < i) i = bound;
    __syncwarp();
    if (swL) q = s.Y[lb];
    if (swR) q = s.X[ra];
    if (in) {
        if (pos == i) { s.Q[high] = static_cast<unsigned short>(q); s.Q[pos] = static_cast<unsigned short>(im); }
        else s.Q[pos] = static_cast<unsigned short>(q);
    } else if (pos == high && i == high) s.Q[pos] = static_cast<unsigned short>(im);
    __syncwarp();
    return i;
}

// Runs the partition phase of numba's argsort on the n items in Q; afterwards a stable sort of the
// permuted sequence equals numba's final order (its insertion sort on segments < 16 is stable).
// Leaves P[candidate] = position.
__device__ __forceinline__ void qs_partition_phase(WarpScratch &s, int n, int lane)
{
    // pending (low, high) ranges, packed low<<8|high; numba pushes the larger side, so with n <= 63
    // and ranges shorter than 16 never pushed-and-split again more than twice, four slots suffice
    int sp = 0, k0 = 0, k1 = 0, k2 = 0, k3 = 0;
    int low = 0, high = n - 1;
    for (;;) {
        while (high - low >= 15) {
            const int i = (high - low <= 31) ? qs_partition32(s, low, high, lane) : qs_partition(s, low, high, lane);
            int push = -1;
            if (high - i > i - low) {
                if (high > i) push = ((i + 1) << 8) | high;
                high = i - 1;
            } else {
                if (i > low) push = (low << 8) | (i - 1);
                low = i + 1;
            }
            if (push >= 0) { k3 = k2; k2 = k1; k1 = k0; k0 = push; ++sp; }
        }
        if (sp == 0) break;
        --sp;
        low = k0 >> 8; high = k0 & 0xff;
        k0 = k1; k1 = k2; k2 = k3;
    }
    if (lane < n) s.P[s.Q[lane] & 0xff] = lane;
    if (lane + 32 < n) s.P[s.Q[lane + 32] & 0xff] = lane + 32;
    __syncwarp();
}

// Final ranks (ascending argsort positions) of n candidates from their ranks (number of strictly
// smaller scores), with numba's order inside groups of equal ranks; groups whose final ranks all lie
// below `cut` keep their common rank (they are pruned anyway).  `occ` = bitmap of the occupied ranks.
// slot 0 = candidate `lane` (valid0), slot 1 = candidate `c1` (valid1).
// `stable` (tie_mode 1, "ties within tolerance" of the north star): equal scores keep their insertion
// order instead — no replay; the CPU checker's tie_mode 1 is the matching restatement.
__device__ __forceinline__ void tie_order(WarpScratch &s, int n, int cut, unsigned long long occ, bool valid0, int &r0,
                                          bool valid1, int c1, int &r1, int lane, bool stable = false)
{
    (void)cut; (void)occ;
    __syncwarp();
    s.Mlo[lane] = 0; s.Mlo[lane + 32] = 0; s.Mhi[lane] = 0; s.Mhi[lane + 32] = 0;
    int P0 = lane, P1 = c1;
    if (!stable) {
        if (valid0) s.Q[lane] = static_cast<unsigned short>((r0 << 8) | lane);
        if (valid1) s.Q[c1] = static_cast<unsigned short>((r1 << 8) | c1);
        __syncwarp();
        qs_partition_phase(s, n, lane);
        P0 = valid0 ? s.P[lane] : 0; P1 = valid1 ? s.P[c1] : 0;
    } else {
        __syncwarp();
    }
    // order inside a group of equal ranks = order of the positions after the partition phase: per rank a
    // bitmap of the positions holding it, a candidate's offset = the set bits below its own position
    if (valid0) atomicOr(P0 < 32 ? &s.Mlo[r0] : &s.Mhi[r0], 1u << (P0 & 31));
    if (valid1) atomicOr(P1 < 32 ? &s.Mlo[r1] : &s.Mhi[r1], 1u << (P1 & 31));
    __syncwarp();
    if (valid0) {
        const unsigned lo = s.Mlo[r0], hi = s.Mhi[r0];
        r0 += P0 < 32 ? __popc(lo & ((1u << P0) - 1)) : __popc(lo) + __popc(hi & ((1u << (P0 - 32)) - 1));
    }
    if (valid1) {
        const unsigned lo = s.Mlo[r1], hi = s.Mhi[r1];
        r1 += P1 < 32 ? __popc(lo & ((1u << P1) - 1)) : __popc(lo) + __popc(hi & ((1u << (P1 - 32)) - 1));
    }
}

// Test hook: final ranks of `n` candidates (values[] = small integers standing in for scores) exactly as
// the prune computes them; one warp per case.  candidates 0..31 in slot 0, 32..63 in slot 1.
__global__ void tie_order_test_kernel(const int32_t *values, const int32_t *ns, int32_t stride, int32_t *out)
{
    __shared__ WarpScratch s;
    const int lane = threadIdx.x;
    const int32_t *v = values + static_cast<size_t>(blockIdx.x) * stride;
    const int n = ns[blockIdx.x];
    const bool valid0 = lane < n, valid1 = lane + 32 < n;
    int r0 = 0, r1 = 0;
    for (int c = 0; c < n; ++c) {
        if (valid0 && v[c] < v[lane]) ++r0;
        if (valid1 && v[c] < v[lane + 32]) ++r1;
    }
    const unsigned long long mine = (valid0 ? 1ull << r0 : 0ull) | (valid1 ? 1ull << r1 : 0ull);
    const unsigned long long occ = __reduce_or_sync(kFull, static_cast<unsigned>(mine)) |
                                   (static_cast<unsigned long long>(__reduce_or_sync(kFull, static_cast<unsigned>(mine >> 32))) << 32);
    tie_order(s, n, 0, occ, valid0, r0, valid1, lane + 32, r1, lane);
    if (valid0) out[static_cast<size_t>(blockIdx.x) * stride + lane] = r0;
    if (valid1) out[static_cast<size_t>(blockIdx.x) * stride + lane + 32] = r1;
}

__device__ __forceinline__ void st_entry(char *rec, int k, int pos, int e, int n, double ts, double sc, uint32_t tag)
{
    st_chunk(rec + 16 + 16 * pos, sc, e, tag);
    st_chunk(rec + 16 + 16 * (k + pos), ts, n, tag);
}

// New vector of one endpoint from the OLD vectors of itself (A) and its peer (B)
// (utils/util.py:509-559), written straight into the endpoint's record with tag `tag`.
// `rwait` > 0: that many negative-sample reads of the old version must have finished first.  The
// counter of a node holds (reads finished) - (reads claimed by writers): the writer CLAIMS its `rwait`
// reads with a returning atomic at the start of the update and, before storing, waits until the counter
// is back to >= 0.  The claim has returned — i.e. has been performed — before the new version is stored,
// so the next writer of the node (which starts from that version) can never see the counter without it.
// (A fire-and-forget subtraction after the wait, as in round 1, is not ordered against the record stores:
// the next writer could read the counter before the subtraction landed and overwrite a version whose
// reader had not loaded it yet.)
// `stats`: prunes that took the exact-tie path (high half) / the merge ranking (low half) are counted per
// warp and added to the global counters once, when the warp leaves the kernel.
// This function is the step of every hub's dependency chain: it is written for the latency of a
// single warp (independent shared-memory loads, unrolled compare loops), not for throughput.
// kKeep (chain warps): the new vector is also returned in registers, in slot order, through `newv`.
template <bool kKeep = false>
__device__ __forceinline__ bool update_endpoint(const SpanParams &p, WarpScratch &s, const VecR &A, const VecR &B,
                                                int peer_id, int e_new, double t_new, int j,
                                                char *rec, uint32_t tag, int32_t *rcnt, uint32_t rwait, int lane,
                                                unsigned &stats, VecR *newv = nullptr)
{
    ZP_DECL;
    int claimed = 0;
    if (rwait != 0 && lane == 0) claimed = atomicSub(rcnt, static_cast<int32_t>(rwait));   // returns the value before
    const int k = p.k;
    const int c1 = A.cnt;
    const int c2 = (B.norm != 0.0) ? B.cnt : 0;
    DecayStep ds;
    {
        const int len = p.decay_len[j];
        int m = A.upd;
        if (m >= len && p.decay_fixed[j]) m = len - 1;
        if (m < len) {
            const double2 *t2 = reinterpret_cast<const double2 *>(p.decay[j] + m);
            const double2 x = __ldg(t2), y = __ldg(t2 + 1);
            ds.new_norm = x.x; ds.s1 = x.y; ds.s2 = y.x; ds.w_new = y.y;
        } else {
            ds = decay_step(A.norm, p.alpha[j], p.beta[j]);
        }
    }
    const double new_norm = ds.new_norm, s1c = ds.s1, s2c = ds.s2;
    double vA = __dmul_rn(A.sc, s1c);
    double vB = __dmul_rn(B.sc, s2c);
    const unsigned long long keyA = pack_key(A.e, A.n), keyB = pack_key(B.e, B.n);

    // ---- duplicate keys between A and B ----
    // The peer's keys are dropped into a 2048-slot byte table (slot = hash of (eidx, node), value =
    // lane); each own key looks its slot up and confirms the full key (eidx, node, ts) of the lane it
    // finds.  Exact as long as the peer's keys fall into distinct slots (checked by reading the slots
    // back); otherwise every own lane scans all peer keys in shared memory.  The table is kept clean
    // (0xff) between uses by erasing the slots just written.
    unsigned dupmask = 0;          // peer entries whose key A already holds
    if (c1 > 0 && c2 > 0 && !(ZEBRA_ABLATE & 1)) {
        const unsigned hA = key_slot(A.e, A.n), hB = key_slot(B.e, B.n);
        const bool isB = lane < c2;
        __syncwarp();
        if (isB) s.T[hB] = static_cast<unsigned char>(lane);
        __syncwarp();
        const int back = isB ? s.T[hB] : lane;
        int probe = lane < c1 ? s.T[hA] : 0xff;
        const bool clash = __any_sync(kFull, back != lane) || (ZEBRA_ABLATE & 16);
        const bool found = __any_sync(kFull, probe != 0xff);      // no own key maps to an occupied slot: no duplicate
        __syncwarp();                  // every lane has read its slots
        if (isB) s.T[hB] = 0xff;
        int hit = -1;
        if (!clash && !found) {
            // nothing to add, dupmask stays 0
        } else if (!clash) {
            const int c = probe < 32 ? probe : 0;
            const unsigned long long kb = __shfl_sync(kFull, keyB, c);
            const double tb = __shfl_sync(kFull, B.ts, c);
            if (probe < c2 && kb == keyA && tb == A.ts) hit = probe;
        } else {
            __syncwarp();
            s.K[lane] = isB ? keyB : 0xffffffffffffffffull;   // (eidx -1, node -1) never occurs
            __syncwarp();
            unsigned cand = 0;
            const ulonglong2 *K2 = reinterpret_cast<const ulonglong2 *>(s.K);
#pragma unroll 4
            for (int q = 0; q < (c2 + 1) >> 1; ++q) {
                const ulonglong2 kk = K2[q];
                cand |= (kk.x == keyA ? 1u : 0u) << (2 * q);
                cand |= (kk.y == keyA ? 1u : 0u) << (2 * q + 1);
            }
            if (lane >= c1) cand = 0;
            // keys are (eidx, node, ts): confirm the timestamp of the (rare) candidates
            while (__any_sync(kFull, cand != 0)) {
                const int t = cand ? (__ffs(cand) - 1) : 0;
                const double ots = __shfl_sync(kFull, B.ts, t);
                if (cand) {
                    if (ots == A.ts) { hit = t; cand = 0; }
                    else cand &= cand - 1;
                }
            }
        }
        if (clash || found) {
            const double got = __shfl_sync(kFull, vB, hit >= 0 ? hit : 0);
            if (hit >= 0) vA = __dadd_rn(vA, got);
            dupmask = __reduce_or_sync(kFull, hit >= 0 ? (1u << hit) : 0u);
        }
    }
    ZP_MARK(8);

    // ---- insertion-order positions: A's entries, then B's new keys, then the new key ----
    const bool ndB = lane < c2 && !((dupmask >> lane) & 1u);
    const unsigned ndmask = __ballot_sync(kFull, ndB);
    int n = c1 + __popc(ndmask);
    int pos1 = c1 + __popc(ndmask & ((1u << lane) - 1));
    const unsigned long long keyN = pack_key(e_new, peer_id);
    const double w_new = ds.w_new;
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
    ZP_MARK(9);

    if (n > k) {
        // ---- keep the k largest, re-inserted in ascending order (util.py:553-559) ----
        const unsigned long long b0 = order_bits(vA);
        const unsigned long long b1 = order_bits(v1);
        __syncwarp();
        if (valid0) s.V[lane] = b0;
        if (valid1) s.V[pos1] = b1;
        if (lane == 0) s.V[n] = 0xffffffffffffffffull;   // pair loads may touch one filler: above every score
        __syncwarp();
        // rank = number of candidates strictly below; equal scores share a rank, which shows up as
        // a hole in the set of occupied ranks
        int r0 = 0, r1 = 0;
        // Stored vectors are in ascending score order whenever their last update pruned (always, for
        // a full vector), and scaling keeps that order.  If neither list was disturbed (no duplicate
        // key added into A, new key not replacing an entry) and both are still sorted, ranks come from
        // a merge: the run start inside the own list plus a binary search in the other one.
        bool fast = dupmask == 0 && !replaced && !(ZEBRA_ABLATE & 8);
        unsigned long long pa = 0, pb = 0;
        if (fast) {
            pa = __shfl_up_sync(kFull, b0, 1); pb = __shfl_up_sync(kFull, b1, 1);
            const bool bad = lane > 0 && ((valid0 && b0 < pa) || (ndB && b1 < pb));
            fast = !__any_sync(kFull, bad);
        }
        if (fast) {
            const unsigned upto = 0xffffffffu >> (31 - lane);
            const unsigned SA = ~__ballot_sync(kFull, valid0 && lane > 0 && b0 == pa);
            const unsigned SB = ~__ballot_sync(kFull, ndB && lane > 0 && b1 == pb);
            const int firstA = 31 - __clz(SA & upto), firstB = 31 - __clz(SB & upto);
            const unsigned long long bw = order_bits(w_new);
            const unsigned long long x1 = lane == 31 ? bw : b0;     // looked up in the peer's part [c1, c1+c2)
            const unsigned long long x2 = b1;                       // looked up in the own part [0, c1)
            int lb = 0, la = 0;
#pragma unroll
            for (int st = 16; st > 0; st >>= 1) {
                const int ib = lb + st, ia = la + st;
                const unsigned long long vb = s.V[c1 + (ib <= c2 ? ib : 1) - 1];
                const unsigned long long va = s.V[(ia <= c1 ? ia : 1) - 1];
                if (ib <= c2 && vb < x1) lb = ib;
                if (ia <= c1 && va < x2) la = ia;
            }
            r0 = firstA + lb + (bw < b0 ? 1 : 0);
            r1 = lane == 31 ? la + lb : firstB + la + (bw < b1 ? 1 : 0);
            stats += 1u;
        } else {
            const ulonglong2 *V2 = reinterpret_cast<const ulonglong2 *>(s.V);
#pragma unroll 4
            for (int q = 0; q < (((ZEBRA_ABLATE & 2) ? 0 : n + 1) >> 1); ++q) {
                const ulonglong2 vv = V2[q];
                r0 += vv.x < b0; r1 += vv.x < b1;
                r0 += vv.y < b0; r1 += vv.y < b1;
            }
        }
        const int cut = n - k;
        if (ZEBRA_ABLATE & 2) { r0 = lane; r1 = pos1; }
        const unsigned long long mine = (valid0 ? 1ull << r0 : 0ull) | (valid1 ? 1ull << r1 : 0ull);
        const unsigned long long occ = __reduce_or_sync(kFull, static_cast<unsigned>(mine)) |
                                       (static_cast<unsigned long long>(__reduce_or_sync(kFull, static_cast<unsigned>(mine >> 32))) << 32);
        ZP_MARK(10);
        // a group of equal scores only matters if it can reach the surviving ranks [cut, n): then the
        // survivor set or its slot order depends on numba's (unstable) argsort
        const unsigned long long surv = ((1ull << k) - 1) << cut;
        if ((occ & surv) != surv && !(ZEBRA_ABLATE & 4)) {
            // numba's (unstable) argsort decides the order inside a group of equal scores
            tie_order(s, n, cut, occ, valid0, r0, valid1, pos1, r1, lane, p.tie_mode != 0);
            stats += 0x10000u;
        }
        ZP_MARK(11);
        keep0 = valid0 && r0 >= cut; pos0 = r0 - cut;
        keep1 = valid1 && r1 >= cut; pos1 = r1 - cut;
        n = k;
    }

    // ---- the version being overwritten may still have negative-sample readers ----
    bool ok = true;
    if (rwait != 0) {
        claimed = __shfl_sync(kFull, claimed, 0);
        if (claimed < static_cast<int32_t>(rwait)) {       // some of the claimed reads had not finished yet
            Waiter w;
            while (ld_relaxed_s32(rcnt) < 0) {
                if (w.expired(p.status, p.wait_ns, 3, tag, rwait, static_cast<int32_t>(rcnt - p.rcnt[j]))) { ok = false; break; }
                __nanosleep(64);
            }
        }
        if (!ok) return false;                             // (also makes the stores depend on the claim's result)
    }
    // every chunk validates itself by its tag, so the order of these stores does not matter
    const int upd = A.upd < kMaxUpdates ? A.upd + 1 : kMaxUpdates;
    if (lane == 0) st_chunk(rec, new_norm, n | (upd << kCountBits), tag);
    if (keep0) st_entry(rec, k, pos0, A.e, A.n, A.ts, vA, tag);
    if (keep1) st_entry(rec, k, pos1, e1, n1, ts1, v1, tag);
    if (kKeep) {
        // the chain warp carries the new vector in registers: scatter by slot, read back by lane
        __syncwarp();
        if (keep0) {
            s.NK[pos0] = keyA; s.NT[pos0] = static_cast<unsigned long long>(__double_as_longlong(A.ts));
            s.NS[pos0] = static_cast<unsigned long long>(__double_as_longlong(vA));
        }
        if (keep1) {
            s.NK[pos1] = pack_key(e1, n1); s.NT[pos1] = static_cast<unsigned long long>(__double_as_longlong(ts1));
            s.NS[pos1] = static_cast<unsigned long long>(__double_as_longlong(v1));
        }
        __syncwarp();
        const unsigned long long kk = s.NK[lane];
        newv->e = static_cast<int>(kk >> 32); newv->n = static_cast<int>(kk & 0xffffffffu);
        newv->ts = __longlong_as_double(static_cast<long long>(s.NT[lane]));
        newv->sc = __longlong_as_double(static_cast<long long>(s.NS[lane]));
        newv->cnt = n; newv->upd = upd; newv->norm = new_norm;
    }
    ZP_MARK(12);
    return ok;
}

__device__ __forceinline__ void scratch_init(WarpScratch &s, int lane)
{
    uint4 *t = reinterpret_cast<uint4 *>(s.T);
    for (int q = lane; q < kKeySlots / 16; q += 32) t[q] = make_uint4(~0u, ~0u, ~0u, ~0u);
    __syncwarp();
}

// how long to sleep before looking again at a record whose wanted version is `dist` edges away
__device__ __forceinline__ unsigned backoff_ns(uint32_t dist)
{
    const unsigned ns = dist * 8u;
    return ns > 4000u ? 4000u : ns;
}

__device__ __forceinline__ char *chain_slot(const SpanParams &p, int32_t cidx, int j, int side)
{
    return p.cmbox + ((static_cast<size_t>(cidx - p.cidx_base) * p.n_tppr + j) * 2 + side) * p.stride;
}

// word w (0..11) of a 48-byte chain step record sits in lane w; fields are broadcast on demand
__device__ __forceinline__ uint32_t load_step_word(const ChainStep *st, int lane)
{
    return lane < 12 ? __ldg(reinterpret_cast<const uint32_t *>(st) + lane) : 0u;
}

// The resident warp of chain `q`, instance `j`: walks the node's edges of this launch in order with the
// node's vector in registers (no L2 hand-off between consecutive edges of a hub) and computes only that
// node's side of each edge (utils/util.py:509-559).  Per step: take the partner's OLD vector (its table
// record, tagged by its previous writer; for a partner that is itself a chain node, the mailbox slot its
// chain warp pushes), push the own OLD vector into the item's mailbox slot for the ticket warp that does
// the rest of the edge, update, publish the new version to the table record.  The next partner's record
// is loaded into registers before the update, so that its latency overlaps the update.
// kShard (node-sharded kernel): records are addressed by local row (node >> log2 world); a partner owned by
// another rank sends its old vector into this rank's mailbox slot of the item and receives ours in its own
// (pushed straight from the chain warp's registers over NVLink, after the launch hand-shake with that rank).
template <bool kShard>
__device__ __noinline__ void run_chain(const SpanParams &p, WarpScratch &s, int q, int j, int lane, unsigned &stats)
{
    const ChainDesc c = p.chains[q];
    if (c.begin >= c.end) return;
    const int k = p.k;
    const int lg = kShard ? p.world_log2 : 0;
    char *table = p.tables[j];
    char *recH = table + static_cast<size_t>(c.node >> lg) * p.stride;
    unsigned peers_ok = 0;
    char *const my_mbox = kShard ? p.mbox_peer[p.rank] + static_cast<size_t>(p.mbox_parity) * p.mbox_half_bytes : nullptr;
    int32_t *rcntH = p.rcnt[j] + c.node;
    const ChainStep *steps = p.csteps;
    uint32_t w0 = load_step_word(steps + c.begin, lane);
    uint32_t w1 = c.begin + 1 < c.end ? load_step_word(steps + c.begin + 1, lane) : 0u;
    VecR H;
    {
        const int32_t prev_self = static_cast<int32_t>(__shfl_sync(kFull, w0, 7));
        uint32_t cur;
        Waiter wt;
        while (!try_load(recH, k, lane, prev_self < 0, p.seq0 + static_cast<uint32_t>(prev_self), H, cur)) {
            if (wt.expired(p.status, p.wait_ns, 5, static_cast<uint32_t>(c.node), static_cast<uint32_t>(prev_self))) return;
            __nanosleep(100);
        }
    }
    // the first partner record (later ones are loaded one step ahead)
    RawRec raw;
    {
        const uint32_t fl = __shfl_sync(kFull, w0, 5);
        const int peer = static_cast<int>(__shfl_sync(kFull, w0, 1));
        raw = load_raw(table + static_cast<size_t>(((fl & (kStepLoop | kStepPeerChain | kStepPeerRemote)) ? c.node : peer) >> lg) * p.stride, k, lane);
    }
    for (int x = c.begin; x < c.end; ++x) {
        ZP_DECL;
        const uint32_t w2 = x + 2 < c.end ? load_step_word(steps + x + 2, lane) : 0u;   // used two steps from now
        const int32_t edge = static_cast<int32_t>(__shfl_sync(kFull, w0, 0));
        const int peer = static_cast<int>(__shfl_sync(kFull, w0, 1));
        const int e_new = static_cast<int>(__shfl_sync(kFull, w0, 2));
        const int32_t prev_peer = static_cast<int32_t>(__shfl_sync(kFull, w0, 3));
        const uint32_t rthr = __shfl_sync(kFull, w0, 4);
        const uint32_t fl = __shfl_sync(kFull, w0, 5);
        const int32_t cidx = static_cast<int32_t>(__shfl_sync(kFull, w0, 6));
        const double t_new = __hiloint2double(static_cast<int>(__shfl_sync(kFull, w0, 9)), static_cast<int>(__shfl_sync(kFull, w0, 8)));
        const uint32_t tag = p.seq0 + static_cast<uint32_t>(edge);
        if (lane == 0) *reinterpret_cast<volatile uint32_t *>(p.cprog + q * p.n_tppr + j) = tag;   // how far the chain is (waiters pace their polling by it)
        const int side = static_cast<int>(fl & kStepDst);
        char *my_slot = chain_slot(p, cidx, j, side);
        VecR B;
        bool alive = true;
        if (fl & kStepLoop) {
            B = vsel(true, H, H);
            push_vector<false>(my_slot, k, H, tag, lane);
        } else if (kShard && (fl & kStepPeerRemote)) {
            // the partner lives on another rank: our old vector goes to the helper here (rows) and over NVLink to
            // the partner's owner; the partner's old vector arrives in this rank's mailbox slot of the item
            push_vector<false>(my_slot, k, H, tag, lane);
            const int partner = peer & ((1 << lg) - 1);
            const size_t slot_off = (static_cast<size_t>(edge - p.tag_offset) * p.n_tppr + j) * p.stride;   // launch-local edge
            if (!((peers_ok >> partner) & 1u)) {
                Waiter wh;
                while (ld_relaxed_sys_s32(p.shard_done + partner) < p.shard_launch - 1) {
                    if (wh.expired(p.status, p.wait_ns, 8, tag, static_cast<uint32_t>(partner))) { alive = false; break; }
                    __nanosleep(200);
                }
                peers_ok |= 1u << partner;
            }
            if (alive) {
                push_vector<true>(p.mbox_peer[partner] + static_cast<size_t>(p.mbox_parity) * p.mbox_half_bytes + slot_off, k, H, tag, lane);
                uint32_t cur;
                Waiter wt;
                while (!try_load<true>(my_mbox + slot_off, k, lane, false, tag, B, cur)) {
                    if (wt.expired(p.status, p.wait_ns, 8, tag, static_cast<uint32_t>(peer))) { alive = false; break; }
                    __nanosleep(40);
                }
            }
        } else if (fl & kStepPeerChain) {
            // both endpoints are chain nodes: each chain warp pushes first, then takes the other's push
            push_vector<false>(my_slot, k, H, tag, lane);
            const char *other = chain_slot(p, cidx, j, side ^ 1);
            uint32_t cur;
            Waiter wt;
            while (!try_load(other, k, lane, false, tag, B, cur)) {
                if (wt.expired(p.status, p.wait_ns, 6, tag, static_cast<uint32_t>(peer))) { alive = false; break; }
                __nanosleep(40);
            }
        } else {
            const char *recP = table + static_cast<size_t>(peer >> lg) * p.stride;
            const uint32_t expect = p.seq0 + static_cast<uint32_t>(prev_peer);
            uint32_t cur;
            bool ok = check_raw(raw, lane, prev_peer < 0, expect, B, cur);
#ifdef ZEBRA_PROFILE
            if (!ok && lane == 0 && q == 0) atomicAdd(p.prof + 14, 1ull);      // partner record not there when the step began
#endif
            Waiter wt;
            while (!ok) {
#ifdef ZEBRA_PROFILE
                if (lane == 0 && q == 0) atomicAdd(p.prof + 13, 1ull);
#endif
                if (wt.expired(p.status, p.wait_ns, 7, tag, expect, peer)) { alive = false; break; }
                __nanosleep(40);
                ok = try_load(recP, k, lane, prev_peer < 0, expect, B, cur);
            }
            // only now: the ticket warp overwrites the partner's record as soon as it has our old vector
            if (alive) push_vector<false>(my_slot, k, H, tag, lane);
        }
        if (!alive) return;
        ZP_MARK_IF(q == 0, 5);
        // next step's partner record on its way while this step computes
        if (x + 1 < c.end) {
            const uint32_t nfl = __shfl_sync(kFull, w1, 5);
            const int npeer = static_cast<int>(__shfl_sync(kFull, w1, 1));
            raw = load_raw(table + static_cast<size_t>(((nfl & (kStepLoop | kStepPeerChain | kStepPeerRemote)) ? c.node : npeer) >> lg) * p.stride, k, lane);
        }
        ZP_MARK_IF(q == 0, 6);
        VecR Hn;
        const bool ok = update_endpoint<true>(p, s, H, B, peer, e_new, t_new, j, recH, tag, rcntH, rthr, lane, stats, &Hn);
        if (!ok) { atomicExch(p.status, 1); return; }
        if ((stats & 0xffffu) > 0xff00u || (stats >> 16) > 0xff00u) {      // the halves are 16 bits wide
            if (lane == 0) { atomicAdd(p.status + 2, static_cast<int>(stats & 0xffffu)); atomicAdd(p.status + 1, static_cast<int>(stats >> 16)); }
            stats = 0;
        }
        H = Hn;
        w0 = w1; w1 = w2;
        ZP_MARK_IF(q == 0, 7);
    }
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32, 2)
tppr_span_kernel(const __grid_constant__ SpanParams p)
{
    __shared__ WarpScratch scratch[kWarpsPerBlock];
    const int lane = threadIdx.x & 31;
    WarpScratch &s = scratch[threadIdx.x >> 5];
    scratch_init(s, lane);
    const unsigned long long total = static_cast<unsigned long long>(p.n_chunk) * p.n_tppr;
    const int k = p.k;
    const uint32_t tag_base = p.seq0 + static_cast<uint32_t>(p.tag_offset);   // tag of launch-local edge 0

    // The first n_chains * n_tppr blocks each serve one (chain, instance): warp 0 walks the chain
    // (run_chain), the other warps are its HELPERS — they take the chain's edges in turn (helper h the
    // steps h-1, h-1+7, ...) and do everything of such an edge that is not the chain node's own update:
    // the partner's update and the rows.  Ticket warps skip edges with a chain endpoint, so the ticket
    // front never queues up behind a chain; when a block's chain is done all its warps draw tickets.
    // A chain warp is one warp's dependent instruction stream: every issue slot and instruction-cache line
    // it shares with other warps lengthens the chain.  While a chain warp runs on an SM, the TICKET warps
    // of the other block resident there stand aside (they hold no work while they do); at most max_yield
    // ticket warps stand aside at any time, so ticket work always has takers.
    unsigned stats = 0;
    int hx = 0, hx_end = 0, hj = 0;
    unsigned smid = 0;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    smid &= kSmSlots - 1;
    bool may_yield = p.n_chains > 0 && p.max_yield > 0;
    unsigned claims = 0;
    if (p.n_chains > 0 && blockIdx.x < static_cast<unsigned>(p.n_chains * p.n_tppr)) {
        const int q = blockIdx.x / p.n_tppr;
        hj = blockIdx.x - q * p.n_tppr;
        may_yield = false;
        if ((p.inst_mask >> hj) & 1) {
            if ((threadIdx.x >> 5) == 0) {
                if (lane == 0) atomicAdd(p.smflag + smid, 1);
                run_chain<false>(p, s, q, hj, lane, stats);
                if (lane == 0) atomicSub(p.smflag + smid, 1);
            } else { hx = p.chains[q].begin + (threadIdx.x >> 5) - 1; hx_end = p.chains[q].end; }
        }
    }

    for (;;) {
        int64_t i;
        int j;
        bool from_ticket = false;
        uint32_t my_tk = 0;
        if (hx < hx_end) {
            // helper: the next edge of the chain that is this warp's turn
            const ChainStep *st = p.csteps + hx;
            hx += kWarpsPerBlock - 1;
            const uint32_t fl = __ldg(&st->flags);
            // an edge between two chain nodes is in both chains: the helpers of its src side take it
            if ((fl & kStepPeerChain) && (fl & kStepDst)) continue;
            i = static_cast<int64_t>(__ldg(&st->edge)) - p.tag_offset;
            j = hj;
        } else {
            if (may_yield && (claims < 4 || (claims & 15) == 0)) {
                if (*reinterpret_cast<volatile int32_t *>(p.smflag + smid) > 0) {
                    int y = 0;
                    if (lane == 0) y = atomicAdd(p.smflag + kSmSlots, 1);
                    y = __shfl_sync(kFull, y, 0);
                    if (y < p.max_yield) {
                        while (*reinterpret_cast<volatile int32_t *>(p.smflag + smid) > 0) __nanosleep(2000);
                    }
                    if (lane == 0) atomicSub(p.smflag + kSmSlots, 1);
                    may_yield = false;          // the chains of this SM are done (or the quota was full)
                }
            }
            ++claims;
            uint32_t tk = 0;
            if (lane == 0) tk = atomicAdd(p.ticket, 1u) - p.ticket_base;
            tk = __shfl_sync(kFull, tk, 0);
            if (tk >= total) {
                if (lane == 0 && stats) {
                    if (stats & 0xffffu) atomicAdd(p.status + 2, static_cast<int>(stats & 0xffffu));
                    if (stats >> 16) atomicAdd(p.status + 1, static_cast<int>(stats >> 16));
                }
                break;
            }
            i = tk / static_cast<uint32_t>(p.n_tppr);     // 32-bit division: items < 2^32
            j = static_cast<int>(tk - i * p.n_tppr);
            from_ticket = true;
            my_tk = tk;
        }
        ZP_DECL;
        if (p.dbg_owner && lane == 0) {
            // debug: every (edge, instance) must be processed exactly once
            const uint32_t me = 1u + blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5) + (hx_end ? 0x40000000u : 0u);
            const uint32_t old = atomicExch(p.dbg_owner + i * p.n_tppr + j, me);
            if (old != 0) {
                atomicExch(p.status, 1);
                Waiter w;
                w.log_wait(p.status, 99, tag_base + static_cast<uint32_t>(i), old, static_cast<int32_t>(me));
            }
        }
        // the whole edge record in one batch of independent loads (one memory round trip), before any branch
        // that depends on it; a load issued after the wait below would sit on the dependency chain
        const int sN = p.src[i], dN = p.dst[i];
        const int gN_raw = p.neg[i];
        const uint32_t flags = p.flags[i];
        const int32_t pS = p.prev_s[i], pD_raw = p.prev_d[i], pG_raw = p.prev_n[i];
        const double t = p.ts[i];
        const int e = p.eidx[i];
        const uint32_t rwS = p.rthr_s[i], rwD_raw = p.rthr_d[i];
        const int32_t ci = p.cidx[i];
        // look-ahead (see below): the node whose record this lane will pull into L2, requested with the batch above
        const unsigned long long tk2 = static_cast<unsigned long long>(my_tk) + gridDim.x * kWarpsPerBlock;
        const int pf_lines = p.stride >> 7;
        const bool pf = from_ticket && tk2 < total && lane < 3 * pf_lines;
        const uint32_t pf_i = static_cast<uint32_t>(tk2) / static_cast<uint32_t>(p.n_tppr);
        const int pf_r = lane / pf_lines;
        int32_t pf_node = 0;
        if (pf) pf_node = (pf_r == 0 ? p.src : (pf_r == 1 ? p.dst : p.neg))[pf_i];
        const bool has_neg = p.rows_per_edge == 3;
        const int gN = has_neg ? gN_raw : sN;
        const bool loop = sN == dN;
        const bool neg_own = has_neg && gN != sN && gN != dN;
        int32_t *rcnt = p.rcnt[j];

        if (from_ticket && (flags & (kFlagSrcChain | kFlagDstChain))) continue;   // a chain block's helpers take it
        if (!((p.inst_mask >> j) & 1)) continue;   // instance not selected: its tables are not touched
        // update-only span and every side of the edge belongs to a chain warp: nothing to do here
        if (p.rows_per_edge == 0 && (flags & kFlagSrcChain) && (loop || (flags & kFlagDstChain))) continue;

        const int32_t pD = loop ? -1 : pD_raw, pG = neg_own ? pG_raw : -1;
        const uint32_t rwD = loop ? 0u : rwD_raw;
        char *table = p.tables[j];
        char *recS = table + static_cast<size_t>(sN) * p.stride;
        char *recD = table + static_cast<size_t>(dN) * p.stride;
        const char *recG = table + static_cast<size_t>(gN) * p.stride;
        const uint32_t my_tag = tag_base + static_cast<uint32_t>(i);
        // an endpoint that is a chain node: its OLD vector arrives in the item's mailbox slot, tagged with
        // this edge, and its update is the chain warp's business
        const bool cS = (flags & kFlagSrcChain) != 0, cD = !loop && (flags & kFlagDstChain) != 0;
        const char *srcS = recS, *srcD = recD;
        uint32_t expS = p.seq0 + static_cast<uint32_t>(pS), expD = p.seq0 + static_cast<uint32_t>(pD);
        bool anyS = pS < 0, anyD = pD < 0;
        if (cS | cD) {
            if (cS) { srcS = chain_slot(p, ci, j, 0); expS = my_tag; anyS = false; }
            if (cD) { srcD = chain_slot(p, ci, j, 1); expD = my_tag; anyD = false; }
        }

        // ---- read the three vectors; wait for the producers of src and dst (the chain) ----
        VecR S, D, G;
        uint32_t curS = 0, curD = 0, curG = 0;
        bool okS = try_load(srcS, k, lane, anyS, expS, S, curS);
        bool okD = loop ? true : try_load(srcD, k, lane, anyD, expD, D, curD);
        bool okG = neg_own ? try_load(recG, k, lane, pG < 0, p.seq0 + static_cast<uint32_t>(pG), G, curG) : true;
        bool alive = true;
        Waiter wt;
        if (pf) {
            // The item a whole front of ticket warps ahead will be claimed about one item-time from now: pull its
            // three records into L2 (no registers held beyond one id per lane, nothing waited for — the id was
            // requested together with this item's edge record and has arrived by now).  Without it a ticket warp
            // spends most of an item waiting for DRAM (ncu: long-scoreboard stalls on the three record loads dominate
            // the hub-free stream at 16 warps per SM).
            const int j2 = static_cast<int>(static_cast<uint32_t>(tk2) - pf_i * p.n_tppr);
            const char *a2 = p.tables[j2] + static_cast<size_t>(pf_node) * p.stride + 128 * (lane - pf_r * pf_lines);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(a2));
        }
        ZP_MARK(0);
        while (!(okS && okD)) {
            // distance (in edges) of the farther unsatisfied table producer from the version visible now
            uint32_t dist = 0;
            if (!okS && !cS) {
                const uint32_t d = curS - tag_base, pe = static_cast<uint32_t>(pS - p.tag_offset);
                dist = (d < static_cast<uint32_t>(p.n_chunk) && d <= pe) ? pe - d : pe + 1;
            }
            if (!okD && !cD) {
                const uint32_t d = curD - tag_base, pe = static_cast<uint32_t>(pD - p.tag_offset);
                const uint32_t dd = (d < static_cast<uint32_t>(p.n_chunk) && d <= pe) ? pe - d : pe + 1;
                dist = dd > dist ? dd : dist;
            }
            if (dist > kNear) {
                // far from our turn: sleep, then look at the header tags only
                __nanosleep(backoff_ns(dist));
                if (!okS && !cS) curS = ld_tag(recS);
                if (!okD && !cD) curD = ld_tag(recD);
                const bool nearS = okS || cS || curS == expS;
                const bool nearD = okD || cD || curD == expD;
                if (!(nearS && nearD)) {
                    if (wt.expired(p.status, p.wait_ns, 1, my_tag, okS ? expD : expS, okS ? dN : sN)) { alive = false; break; }
                    continue;
                }
            } else if ((!okS && cS) || (!okD && cD)) {
                // waiting for a chain warp to reach this edge: sleep in proportion to how far away it is
                // (it publishes the tag of the step it is at), then look at the slot header only.  The
                // partner's update of this edge may be the producer of the chain's NEXT step (reply
                // edges), so once the chain is here the polling is tight.
                uint32_t away = 0;
                if (!okS && cS) away = my_tag - *reinterpret_cast<volatile uint32_t *>(p.cprog + ((flags >> kFlagSrcChainShift) & 31u) * p.n_tppr + j);
                if (!okD && cD) {
                    const uint32_t a2 = my_tag - *reinterpret_cast<volatile uint32_t *>(p.cprog + ((flags >> kFlagDstChainShift) & 31u) * p.n_tppr + j);
                    away = a2 > away ? a2 : away;
                }
                if (away != 0) {
                    __nanosleep(away > 1024u ? 4000u : 50u + away * 4u);
                    const bool hS = okS || !cS || ld_tag(srcS) == my_tag;
                    const bool hD = okD || !cD || ld_tag(srcD) == my_tag;
                    if (!(hS && hD)) {
                        if (wt.expired(p.status, p.wait_ns, 2, my_tag, away)) { alive = false; break; }
                        continue;
                    }
                }
            }
            if (!okS) okS = try_load(srcS, k, lane, false, expS, S, curS);
            if (!okD) okD = try_load(srcD, k, lane, false, expD, D, curD);
            if (!okG) okG = try_load(recG, k, lane, false, p.seq0 + static_cast<uint32_t>(pG), G, curG);
            if (wt.expired(p.status, p.wait_ns, (!okS && cS) || (!okD && cD) ? 2 : 1, my_tag, okS ? expD : expS, okS ? dN : sN)) { alive = false; break; }
        }
        if (!alive) continue;
        ZP_MARK(1);
        if (loop) D = vsel(true, S, S);
        if (!neg_own) G = vsel(gN == sN, S, D);    // a negative equal to an endpoint: its OLD vector
        if (okG && neg_own && (flags & kFlagSignal) && lane == 0) atomicAdd(rcnt + gN, 1);
        bool sigG = okG;

        // ---- the hotter endpoint first: its next writer is the one most likely to be waiting ----
        // (one copy of the update code, run twice: the kernel is bound by instruction fetch otherwise)
        const bool df = (flags & kFlagDstFirst) && !loop;
        bool ok = true;
#pragma unroll 1
        for (int h = 0; h < (loop ? 1 : 2); ++h) {
            const bool isD = (h == 0) ? df : !df;
            if (isD ? cD : cS) continue;            // a chain node's side is computed by its chain warp
            const VecR A = vsel(isD, D, S), B = vsel(isD, S, D);
            ok &= update_endpoint(p, s, A, B, isD ? sN : dN, e, t, j, isD ? recD : recS, my_tag,
                                  rcnt + (isD ? dN : sN), isD ? rwD : rwS, lane, stats);
            if ((stats & 0xffffu) > 0xff00u || (stats >> 16) > 0xff00u) {      // the halves are 16 bits wide
                if (lane == 0) { atomicAdd(p.status + 2, static_cast<int>(stats & 0xffffu)); atomicAdd(p.status + 1, static_cast<int>(stats >> 16)); }
                stats = 0;
            }
            ZP_MARK(2 + h);
        }
        if (!ok) { atomicExch(p.status, 1); continue; }

        // ---- off the chain: the negative's vector (if it was not ready yet) and the output rows ----
        while (!okG) {
            __nanosleep(64);
            okG = try_load(recG, k, lane, false, p.seq0 + static_cast<uint32_t>(pG), G, curG);
            if (wt.expired(p.status, p.wait_ns, 4, my_tag, p.seq0 + static_cast<uint32_t>(pG), gN)) break;
        }
        if (!okG) continue;
        if (!sigG && neg_own && (flags & kFlagSignal) && lane == 0) atomicAdd(rcnt + gN, 1);

        if (p.rows_per_edge >= 2) {
            int64_t row = p.i_offset + i, nrow = p.n_total;
            char *obase = p.out;
            if (p.batch_size > 0) {
                // plans and epochs are shorter than 2^31 edges (checked on the host): 32-bit divisions
                const int64_t b = static_cast<uint32_t>(row) / static_cast<uint32_t>(p.batch_size);
                row -= b * p.batch_size;
                nrow = p.n_total - b * p.batch_size;
                if (nrow > p.batch_size) nrow = p.batch_size;
                obase += static_cast<size_t>(static_cast<uint32_t>(b - p.first_batch) % static_cast<uint32_t>(p.out_slots)) * p.slot_bytes;
            }
            const size_t plane = static_cast<size_t>(nrow) * p.rows_per_edge * k;
            obase += static_cast<size_t>(j) * plane * 16;
            emit_row(p, obase, plane, row, S, t, lane);
            emit_row(p, obase, plane, row + nrow, D, t, lane);
            if (has_neg) emit_row(p, obase, plane, row + 2 * nrow, G, t, lane);
        }
        ZP_MARK(4);
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
        counts[a] = static_cast<int32_t>(h.z & ((1u << kCountBits) - 1));
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
