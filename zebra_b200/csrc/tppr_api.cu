// tppr_api.cu — host side of libzebra_tppr.so: the C ABI declared in include/zebra_tppr.h.
//
// The host's only jobs are (a) turning a chronological span into per-access dependency
// thresholds (one pass, two uint32 per node), (b) staging the span through pinned memory and
// (c) launching the persistent span kernel.  All T-PPR arithmetic runs on the GPU; there is no
// CPU fallback: without a CUDA device every entry point fails with an error.
#include "common.cuh"
#include "tppr_device.cuh"
#include "tppr_shard_device.cuh"
#include "../../include/zebra_tppr.h"

#include <algorithm>
#include <array>
#include <cstdarg>
#include <cstdio>
#include <climits>
#include <cstring>
#include <memory>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include <stdlib.h>
#include <sys/mman.h>

namespace {
thread_local std::string g_err;
}

int zebra::fail(const char *fmt, ...)
{
    char buf[2048];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

namespace {

using zebra::fail;

constexpr int64_t kChunk = 1 << 18;  // edges per launch of a long span
constexpr int kStatusWords = 8 + 4 * zebra::kWaitLog;

struct Staging {
    char *host = nullptr;   // pinned
    char *dev = nullptr;
    size_t cap = 0;         // edges
    cudaEvent_t free_ev = nullptr;
    bool pending = false;
};

}  // namespace

// per-node scratch of the host dependency pass: last writer, head of the list of negative readers of
// the current version and their count (all -1 / 0 between passes)
struct NodeDep { int32_t lastw, rhead; uint32_t rcount, pad; };

// The per-node scratch of the host dependency pass is read at random (negative samples are uniform
// over the nodes): with 4 KB pages nearly every look-up is a TLB miss on top of the cache miss, and a
// software prefetch that misses the TLB is dropped.  2 MB-aligned and madvise(MADV_HUGEPAGE): a handful
// of huge pages when the kernel grants them, plain memory otherwise.
template <typename T>
struct HugeArray {
    T *p = nullptr;
    size_t n = 0;
    HugeArray() = default;
    HugeArray(const HugeArray &) = delete;
    HugeArray &operator=(const HugeArray &) = delete;
    ~HugeArray() { free(p); }
    void assign(size_t count, const T &v)
    {
        if (count != n) {
            free(p);
            p = nullptr; n = 0;
            constexpr size_t kHuge = size_t(2) << 20;
            const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + kHuge - 1) / kHuge * kHuge;
            void *q = nullptr;
            if (posix_memalign(&q, kHuge, bytes) != 0) throw std::bad_alloc();
            madvise(q, bytes, MADV_HUGEPAGE);
            p = static_cast<T *>(q); n = count;
        }
        std::fill(p, p + n, v);
    }
    T *data() { return p; }
    size_t size() const { return n; }
    const T *begin() const { return p; }
    const T *end() const { return p + n; }
};

// Scratch of one host dependency pass (fill_soa): per-node state and the chains of negative readers.
// The handle owns one; every worker thread of a pipelined epoch owns another.
struct HostScratch {
    HugeArray<NodeDep> dep;
    std::vector<int32_t> rnext;
};

struct zebra_tppr {
    int64_t num_nodes = 0;
    int32_t k = 0, n_tppr = 0, stride = 0, device = 0;
    double alpha[zebra::kMaxInst], beta[zebra::kMaxInst];
    char *tables[zebra::kMaxInst] = {};
    int32_t *rcnt = nullptr;                  // [n_tppr][num_nodes] negative reads finished, not yet consumed by a writer
    HugeArray<NodeDep> dep;                   // host scratch of the dependency pass, one cache line touch per node
    std::vector<int32_t> rnext;
    uint32_t epoch = 0;                       // tag epoch the bound tables were last written in
    zebra::DecayStep *decay[zebra::kMaxInst] = {};   // device: update factors by update count
    int32_t decay_len[zebra::kMaxInst] = {}, decay_fixed[zebra::kMaxInst] = {};
    uint32_t *ticket = nullptr;
    uint32_t ticket_base = 0;
    int32_t *status = nullptr;                // device [8]: [0] wait time-out, [1] tie prunes, [2] merge-ranked prunes, [4..6] what the expired wait was for
    int32_t *status_host = nullptr;           // pinned [8]
    unsigned long long *prof = nullptr;       // device [16] phase cycle sums (ZEBRA_PROFILE builds only)
    Staging stg[2];
    int stg_next = 0;
    char *out_dev = nullptr;
    size_t out_cap = 0;
    char *out_pinned = nullptr;
    size_t out_pinned_cap = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    int max_blocks = 0;
    int64_t launches = 0, last_items = 0, last_ties = 0, last_merge = 0, timeouts = 0;
    struct zebra_plan *plan_active = nullptr;   // a plan is part-way through its batches
    bool rcnt_dirty = false;                    // a span was abandoned part-way: unconsumed negative-read counts may remain
    int64_t plan_next = 0;
    struct zebra_epoch *epoch_active = nullptr; // a pipelined epoch is part-way through its windows
    // window ring of the pipelined epoch path, kept across epochs (pinned allocations are slow)
    struct EpochSlot {
        Staging stg;                 // edge records of the window (pinned + device)
        char *out_dev = nullptr;     // rows of the window on the device
        char *out_host = nullptr;    // the same in pinned host memory
        size_t out_cap = 0;
        cudaEvent_t ev_kernel = nullptr, ev_copy = nullptr;
        bool copy_pending = false;
    } eslot[3];
    cudaStream_t copy_stream = nullptr;
    // staging ring of the pipelined epoch: the dependency pass of window w runs (on a worker thread when the
    // epoch has any) into slot w % kPrepSlots while earlier windows are on the GPU
    static constexpr int kPrepSlots = 6;
    Staging prep[kPrepSlots];
    int host_threads = -1;                     // worker threads of an epoch (ZEBRA_HOST_THREADS; -1 = default)
    // node-sharded mode (zebra_tppr_shard_*): tables are library-owned local slices, the arena
    // (mailboxes + output region) is peer-mapped on every rank
    int32_t world = 1, rank = 0, world_log2 = 0;
    int64_t table_rows = 0;                    // records per table: num_nodes, or the local slice when sharded
    bool own_tables = false;
    char *arena = nullptr;
    size_t mbox_half = 0, out_region = 0;
    int64_t max_window_edges = 0;
    char *peer_arena[zebra::kMaxRanks] = {};
    bool peer_opened[zebra::kMaxRanks] = {};   // mapped with cudaIpcOpenMemHandle (to be closed)
    int64_t shard_launches = 0;               // sharded launches so far on this handle (the same on every rank)
    uint32_t shard_seq = 1;                    // handle-local tag sequence: identical on every rank
    int shard_max_blocks = 0;
    int32_t row_mode = 0;
    char *plan_stage = nullptr;                // pinned staging of sharded plans, kept across plans
    size_t plan_stage_cap = 0;
    // hub chains (tppr_device.cuh "hub chains"): the chain nodes of the span being streamed
    bool chains_enabled = true;
    int32_t max_chains = 8;                     // chain nodes per span (ZEBRA_CHAINS, at most zebra::kMaxChains): 1-8 measured 2.5 % faster than 24 on the Wiki-Talk-shaped stream, equal on the Reddit- and MOOC-shaped ones
    std::vector<int32_t> chain_cnt;            // [num_nodes] scratch of the detection pass (zero between passes)
    int32_t *smflag = nullptr;                 // device [2 * kSmSlots]: chain warps running per SM, ticket warps standing aside
    int32_t tie_mode = 0;                      // 0 exact (numba's argsort), 1 stable (insertion order among equal scores)
    bool shard_chains = false;                 // ZEBRA_SHARD_CHAINS=1: hub chains on node-sharded handles too
    bool sm_yield = true;                      // ZEBRA_SM_YIELD=0 switches the standing aside off
    uint32_t *dbg_owner = nullptr;             // debug (ZEBRA_DEBUG_ITEMS=1): who processed each item of the launch
    size_t dbg_cap = 0;
    bool dbg_items = false;
    uint32_t *cprog = nullptr;                 // device [kMaxChains][kMaxInst]: tag of the step each chain warp is at
    char *cmbox = nullptr;                     // chain mailbox (device), grown on demand
    size_t cmbox_cap = 0;
    unsigned long long wait_ns = 20000000000ull;
    int64_t chain_items = 0, chain_steps = 0;  // statistics of the last span
};

// The chain nodes of one span (chosen once per span, used by every dependency pass over its edges).
struct ChainSet {
    int32_t n = 0;
    int32_t node[zebra::kMaxChains] = {};
    bool owned[zebra::kMaxChains] = {};        // node-sharded: this rank owns the node (runs its chain warp); else it only helps
    std::vector<uint8_t> map;                  // [num_nodes] 0 = ordinary node, q + 1 = chain q; empty when n == 0
};

// What one host dependency pass found out about the chain nodes of its edges.
struct ChainBuild {
    std::vector<zebra::ChainStep> steps[zebra::kMaxChains];
    int32_t n_items = 0;
    void clear() { for (auto &v : steps) v.clear(); n_items = 0; }
    size_t total() const { size_t t = 0; for (auto &v : steps) t += v.size(); return t; }
};

// Chain geometry of one kernel launch.
struct ChainLaunch {
    int32_t n = 0;
    int32_t node[zebra::kMaxChains], begin[zebra::kMaxChains], end[zebra::kMaxChains], owned[zebra::kMaxChains];
    const zebra::ChainStep *dev_steps = nullptr;
    int32_t cidx_base = 0, n_items = 0;
};

// A span prepared once and kept resident in HBM (edge records + dependency thresholds).
struct zebra_plan {
    zebra_tppr *h = nullptr;
    int64_t n = 0, batch_size = 0, n_batches = 0;
    int32_t rows_per_edge = 0;
    char *dev = nullptr;
    size_t cap = 0;
    uint32_t seq0 = 0;                          // tag base of the run in progress
    // node-sharded handles: span indices of the edges this rank takes part in, and for every batch the
    // position in that list of its first edge
    int32_t *dev_list = nullptr;
    std::vector<int64_t> batch_ofs;
    // hub chains of the span: per chain the span indices of its steps (ascending) and where its step records
    // start inside the device image; per chain item the span index of its edge
    int32_t chain_n = 0;
    int32_t chain_node[zebra::kMaxChains] = {};
    bool chain_owned[zebra::kMaxChains] = {};
    std::vector<int32_t> chain_edges[zebra::kMaxChains];
    int32_t chain_off[zebra::kMaxChains + 1] = {};
    std::vector<int32_t> item_edge;
    size_t steps_bytes = 0;
};

// A pipelined pass over a host-resident span: windows of whole batches are staged, computed and
// copied back through a ring of three slots; the caller consumes window after window from pinned
// host memory while later windows are in flight.
struct zebra_epoch {
    zebra_tppr *h = nullptr;
    const int32_t *src = nullptr, *dst = nullptr, *neg = nullptr;   // caller memory, must outlive the epoch
    const double *ts = nullptr;
    const int64_t *eidx = nullptr;
    int64_t n = 0, batch_size = 0, n_batches = 0, window = 0, n_windows = 0;
    int32_t rows_per_edge = 0;
    size_t slot_bytes = 0;          // one batch's rows
    int64_t issued = 0, consumed = 0;
    cudaStream_t st = nullptr;
    ChainSet cs;                    // chain nodes of the span, chosen at epoch_begin
    // per staging slot: what the dependency pass of the window staged there produced
    struct Prep {
        ChainBuild cb; ChainLaunch cl; size_t img_bytes = 0, cap = 0; int64_t window = -1; bool ready = false, failed = false; std::string err;
        std::vector<int32_t> list; std::vector<int64_t> batch_ofs;      // node-sharded handles: this rank's work list of the window
    };
    Prep prep[zebra_tppr::kPrepSlots];
    // worker threads: window w is prepared by worker w % n_workers, as soon as its slot is free
    std::vector<std::thread> workers;
    std::vector<std::unique_ptr<HostScratch>> scratch;
    std::mutex mu;
    std::condition_variable cv;
    bool stop = false;
};

namespace {

size_t record_bytes(int k) { return zebra::record_stride(k); }

// bytes per edge in a staging slot: src,dst,neg,eidx,prev_s,prev_d,prev_n,rthr_s,rthr_d,flags (4 B each) + ts (8 B)
// + cidx (4 B, chain item index); the chain step records of the slot follow the arrays
constexpr size_t kEdgeBytes = 10 * 4 + 8 + 4;
constexpr size_t kCidxOffset = 10 * 4 + 8;          // byte offset of the cidx array, per unit of `cap`
constexpr size_t kStepBytesPerEdge = 2 * sizeof(zebra::ChainStep);   // worst case: both endpoints are chain nodes
constexpr size_t kArenaHeader = 256;                // sharded arena: [header: launches finished per rank][mailbox halves][rows]
constexpr int64_t kChainMinEdges = 4096;            // shorter spans are streamed without chain warps

// Tags are edge sequence numbers drawn from one process-wide counter, so a table written through one
// handle can be bound to another.  Tag 0 is what a zero-filled table carries.
std::mutex g_seq_mu;
uint32_t g_seq = 1, g_epoch = 0;

int ensure_staging(zebra_tppr *h, Staging &s, size_t edges)
{
    if (s.cap >= edges) return 0;
    if (s.pending) { CU(cudaEventSynchronize(s.free_ev)); s.pending = false; }
    if (s.host) CU(cudaFreeHost(s.host));
    if (s.dev) CU(cudaFree(s.dev));
    size_t cap = std::max<size_t>(edges, 1024);
    // arrays + (sharded epochs) the work list + the chain step records, worst case
    CU(cudaMallocHost(&s.host, cap * (kEdgeBytes + kStepBytesPerEdge + 8) + 64));
    CU(cudaMalloc(&s.dev, cap * (kEdgeBytes + kStepBytesPerEdge + 8) + 64));
    if (!s.free_ev) CU(cudaEventCreateWithFlags(&s.free_ev, cudaEventDisableTiming));
    s.cap = cap;
    return 0;
}

// Picks the chain nodes of a span: the nodes with the most incident edges (sampled count), at most
// h->max_chains of them, none below 1/16 of the hottest node or below 64 edges.  The choice (`cs`) is handed
// to the dependency passes over the span's edges.
template <typename NodeT>
void detect_chains(zebra_tppr *h, const NodeT *src, const NodeT *dst, int64_t n, ChainSet &cs)
{
    cs.n = 0;
    if (!h->chains_enabled || n < kChainMinEdges || h->max_chains <= 0) return;
    // node-sharded handles: chain warps (and helpers for other ranks' chains) are built and tested, but measured
    // on 2 B200 they gain 0-7 % of `value` (a cross-shard reply edge costs two NVLink hops plus the partner's
    // update whoever walks the chain) and cost 15 % of `e2e` (heavier host pass): off unless ZEBRA_SHARD_CHAINS=1
    if (h->world > 1 && !h->shard_chains) return;
    const uint64_t N = static_cast<uint64_t>(h->num_nodes);
    if (h->chain_cnt.size() != N) h->chain_cnt.assign(N, 0);
    int32_t *cnt = h->chain_cnt.data();
    const int64_t stride = std::max<int64_t>(1, n >> 20);
    int32_t top = 0;
    for (int64_t i = 0; i < n; i += stride) {
        const uint64_t a = static_cast<uint64_t>(static_cast<int64_t>(src[i])), b = static_cast<uint64_t>(static_cast<int64_t>(dst[i]));
        if (a < N) top = std::max(top, ++cnt[a]);
        if (b < N && b != a) top = std::max(top, ++cnt[b]);
    }
    const int32_t thr = std::max<int32_t>({4, top / 16, static_cast<int32_t>((64 + stride - 1) / stride)});
    std::vector<std::pair<int32_t, int32_t>> found;          // (count, node)
    for (int64_t i = 0; i < n; i += stride) {
        const uint64_t ab[2] = {static_cast<uint64_t>(static_cast<int64_t>(src[i])), static_cast<uint64_t>(static_cast<int64_t>(dst[i]))};
        for (uint64_t v : ab) {
            if (v >= N) continue;
            if (cnt[v] >= thr) { found.emplace_back(cnt[v], static_cast<int32_t>(v)); cnt[v] = -cnt[v]; }
        }
    }
    for (int64_t i = 0; i < n; i += stride) {                 // the scratch goes back to zero
        const uint64_t a = static_cast<uint64_t>(static_cast<int64_t>(src[i])), b = static_cast<uint64_t>(static_cast<int64_t>(dst[i]));
        if (a < N) cnt[a] = 0;
        if (b < N) cnt[b] = 0;
    }
    std::sort(found.begin(), found.end(), [](const std::pair<int32_t, int32_t> &x, const std::pair<int32_t, int32_t> &y) {
        return x.first != y.first ? x.first > y.first : x.second < y.second;
    });
    // every (chain, instance) gets a block of its own, resident next to the ticket blocks
    const int room = std::max(0, (h->world > 1 ? h->shard_max_blocks : h->max_blocks) / 2 / std::max(1, h->n_tppr));
    const int keep = std::min<int>({static_cast<int>(found.size()), h->max_chains, zebra::kMaxChains, room});
    // node-sharded: every rank picks the same nodes from the same replicated span; the owner of a chain node runs
    // its chain warp, the other ranks run helpers for the edges that join it to one of THEIR nodes
    if (keep > 0) cs.map.assign(N, 0);
    for (int q = 0; q < keep; ++q) {
        const int32_t v = found[static_cast<size_t>(q)].second;
        cs.node[q] = v;
        cs.owned[q] = h->world == 1 || (v & (h->world - 1)) == h->rank;
        cs.map[static_cast<size_t>(v)] = static_cast<uint8_t>(q + 1);
    }
    cs.n = keep;
}

// Fills one staging slot for m edges and derives their dependencies (one forward pass and a clean-up
// pass, host scratch only — no state survives the call):
//   prev_x   index (within these m edges) of the previous edge that wrote endpoint x, -1 = none;
//   rthr_x   negative-sample reads of x's current version that must finish before it is overwritten;
//   flags    kFlagSignal  : this edge's negative read is waited for by a later writer;
//            kFlagDstFirst: dst is written again sooner than src -> update it first.
// A negative equal to src or dst is served from the writer's registers.
// With a non-empty chain set `cs` the pass also marks the endpoints that are chain nodes (kFlagSrcChain /
// kFlagDstChain), numbers the edges that have one (cidx) and writes, per chain, the step records its
// resident warp walks.
template <typename NodeT>
int fill_soa(zebra_tppr *h, char *host, size_t cap, const NodeT *src, const NodeT *dst, const NodeT *neg,
             const double *ts, const int64_t *eidx, int64_t m, const ChainSet *cs = nullptr, ChainBuild *cb = nullptr,
             HostScratch *scratch = nullptr)
{
    const int64_t N = h->num_nodes;
    int32_t *o_src = reinterpret_cast<int32_t *>(host);
    int32_t *o_dst = o_src + cap, *o_neg = o_dst + cap, *o_e = o_neg + cap;
    int32_t *p_s = o_e + cap, *p_d = p_s + cap, *p_n = p_d + cap;
    uint32_t *r_s = reinterpret_cast<uint32_t *>(p_n + cap), *r_d = r_s + cap, *flg = r_d + cap;
    double *o_ts = reinterpret_cast<double *>(flg + cap);
    int32_t *o_cidx = reinterpret_cast<int32_t *>(o_ts + cap);
    const uint8_t *cmap = (cb && cs && cs->n > 0) ? cs->map.data() : nullptr;
    if (cb) cb->clear();
    // validate before touching the scratch (one branch-free sweep; the culprit is looked for only on failure)
    const uint64_t uN = static_cast<uint64_t>(N);
    uint64_t bad = 0;
    for (int64_t i = 0; i < m; ++i) {
        const uint64_t a = static_cast<uint64_t>(static_cast<int64_t>(src[i])), b = static_cast<uint64_t>(static_cast<int64_t>(dst[i]));
        const uint64_t g = neg ? static_cast<uint64_t>(static_cast<int64_t>(neg[i])) : a;
        bad |= (a >= uN) | (b >= uN) | (g >= uN) | (static_cast<uint64_t>(eidx[i]) > 0x7fffffffULL);
    }
    for (int64_t i = 0; bad && i < m; ++i) {
        const int64_t a = src[i], b = dst[i], g = neg ? static_cast<int64_t>(neg[i]) : a;
        const int64_t e = eidx[i];
        if (a < 0 || a >= N || b < 0 || b >= N || g < 0 || g >= N)
            return fail("node id out of range at edge %lld (num_nodes=%lld)", (long long)i, (long long)N);
        if (e < 0 || e > 0x7fffffffLL) return fail("edge id %lld does not fit int32", (long long)e);
    }
    NodeDep *dep = scratch ? scratch->dep.data() : h->dep.data();
    std::vector<int32_t> &rn = scratch ? scratch->rnext : h->rnext;
    if (rn.size() < static_cast<size_t>(m)) rn.resize(static_cast<size_t>(m));
    int32_t *rnext = rn.data();
    // kFlagDstFirst without a backward pass: NodeDep::pad remembers, per node, the latest edge (index + 1)
    // that wrote it and does not know yet which of its endpoints is rewritten first; the first later
    // write of either endpoint settles it (one edge rewriting both: no preference)
    auto settle = [&](uint32_t qe, int32_t x) {
        const int32_t s = o_src[qe], d = o_dst[qe];
        if (x == d) flg[qe] |= zebra::kFlagDstFirst;
        NodeDep &other = dep[x == d ? s : d];
        if (other.pad == qe + 1) other.pad = 0;
        dep[x].pad = 0;
    };
    auto overwrite = [&](NodeDep &n, int64_t i, int32_t &prev, uint32_t &rthr) {
        prev = n.lastw; rthr = n.rcount;
        if (n.rcount) {                                               // its pending negative reads must signal
            for (int32_t r = n.rhead; r >= 0; r = rnext[r]) flg[r] |= zebra::kFlagSignal;
            n.rhead = -1; n.rcount = 0;
        }
        n.lastw = static_cast<int32_t>(i);
    };
    constexpr int64_t kAhead = 12;                                    // the scratch look-ups are cache misses: fetch ahead
    for (int64_t i = 0; i < m; ++i) {
        if (i + kAhead < m) {
            __builtin_prefetch(dep + src[i + kAhead], 1);
            __builtin_prefetch(dep + dst[i + kAhead], 1);
            if (neg) __builtin_prefetch(dep + neg[i + kAhead], 1);
        }
        const int32_t a = static_cast<int32_t>(src[i]), b = static_cast<int32_t>(dst[i]);
        const int32_t g = neg ? static_cast<int32_t>(neg[i]) : a;
        o_src[i] = a; o_dst[i] = b; o_neg[i] = g;
        o_e[i] = static_cast<int32_t>(eidx[i]);
        o_ts[i] = ts[i];
        flg[i] = 0;
        if (neg && g != a && g != b) {
            p_n[i] = dep[g].lastw;
            rnext[i] = dep[g].rhead; dep[g].rhead = static_cast<int32_t>(i); dep[g].rcount++;
        } else {
            p_n[i] = -1;
        }
        NodeDep *da = dep + a, *db = b != a ? dep + b : nullptr;
        const uint32_t qa = da->pad, qb = db ? db->pad : 0u;
        if (qa && qa == qb) {
            da->pad = 0; db->pad = 0;
        } else {
            if (qa) settle(qa - 1, a);
            if (qb) settle(qb - 1, b);
        }
        overwrite(*da, i, p_s[i], r_s[i]);
        if (db) {
            overwrite(*db, i, p_d[i], r_d[i]);
            da->pad = static_cast<uint32_t>(i) + 1; db->pad = static_cast<uint32_t>(i) + 1;
        } else {
            p_d[i] = -1; r_d[i] = 0;
        }
        o_cidx[i] = -1;
        if (cmap) {
            const uint8_t ca = cmap[a], cbq = db ? cmap[b] : 0;
            if (ca | cbq) {
                const int32_t ci = cb->n_items++;
                o_cidx[i] = ci;
                const int32_t ei = static_cast<int32_t>(i), ee = static_cast<int32_t>(eidx[i]);
                if (ca) {
                    flg[i] |= zebra::kFlagSrcChain | (static_cast<uint32_t>(ca - 1) << zebra::kFlagSrcChainShift);
                    cb->steps[ca - 1].push_back(zebra::ChainStep{ei, b, ee, db ? p_d[i] : -1, r_s[i],
                                                                 (db ? 0u : zebra::kStepLoop) | (cbq ? zebra::kStepPeerChain : 0u), ci, p_s[i], ts[i], 0});
                }
                if (cbq) {
                    flg[i] |= zebra::kFlagDstChain | (static_cast<uint32_t>(cbq - 1) << zebra::kFlagDstChainShift);
                    cb->steps[cbq - 1].push_back(zebra::ChainStep{ei, a, ee, p_s[i], r_d[i],
                                                                  zebra::kStepDst | (ca ? zebra::kStepPeerChain : 0u), ci, p_d[i], ts[i], 0});
                }
            }
        }
    }
    for (int64_t i = 0; i < m; ++i) {                                 // the scratch goes back to "no writer, no reader"
        dep[o_src[i]].lastw = -1; dep[o_src[i]].pad = 0;
        dep[o_dst[i]].lastw = -1; dep[o_dst[i]].pad = 0;
        dep[o_neg[i]].rhead = -1; dep[o_neg[i]].rcount = 0;
    }
    return 0;
}

// Node-sharded variant of fill_soa, in two passes.  A rank only ever reads and writes the records of
// the nodes it owns, so only the edges it takes part in (it owns src, dst or a distinct negative) get
// a staging record, and the records are COMPACT: indexed by position in the rank's work list, not by
// span index (1/3 of the span at world = 8: a third of the host writes, page faults and H2D bytes).
//
// Pass A builds the work list (span indices, ascending) and, per batch, the list position of its
// first edge; it validates every id before any scratch is touched.
int shard_work_list(zebra_tppr *h, const int32_t *src, const int32_t *dst, const int32_t *neg, const int64_t *eidx, int64_t m,
                    int64_t batch_size, std::vector<int32_t> &list, std::vector<int64_t> &batch_ofs)
{
    const uint32_t N = static_cast<uint32_t>(std::min<int64_t>(h->num_nodes, 0x7fffffffLL));
    const int32_t pm = h->world - 1, me = h->rank;
    list.clear();
    list.reserve(static_cast<size_t>(m / h->world) * 3 + 16);
    // chunks of 1024 edges, no data-dependent branch inside: ownership is a coin flip per edge
    constexpr int64_t kChunk = 1024;
    int32_t tmp[kChunk];
    for (int64_t lo = 0; lo < m; lo += kChunk) {
        const int64_t hi = std::min(m, lo + kChunk);
        uint32_t bad = 0;
        int cnt = 0;
        if (neg) {
            for (int64_t i = lo; i < hi; ++i) {
                const int32_t a = src[i], b = dst[i], g = neg[i];
                bad |= (static_cast<uint32_t>(a) >= N) | (static_cast<uint32_t>(b) >= N) | (static_cast<uint32_t>(g) >= N);
                const int own = ((a & pm) == me) | ((b & pm) == me) | ((g != a) & (g != b) & ((g & pm) == me));
                tmp[cnt] = static_cast<int32_t>(i);
                cnt += own;
            }
        } else {
            for (int64_t i = lo; i < hi; ++i) {
                const int32_t a = src[i], b = dst[i];
                bad |= (static_cast<uint32_t>(a) >= N) | (static_cast<uint32_t>(b) >= N);
                tmp[cnt] = static_cast<int32_t>(i);
                cnt += ((a & pm) == me) | ((b & pm) == me);
            }
        }
        if (bad) {
            for (int64_t i = lo; i < hi; ++i) {
                const int64_t a = src[i], b = dst[i], g = neg ? static_cast<int64_t>(neg[i]) : a;
                if (a < 0 || a >= h->num_nodes || b < 0 || b >= h->num_nodes || g < 0 || g >= h->num_nodes)
                    return fail("node id out of range at edge %lld (num_nodes=%lld)", (long long)i, (long long)h->num_nodes);
            }
        }
        for (int c = 0; c < cnt; ++c) {
            const int64_t e = eidx[tmp[c]];
            if (e < 0 || e > 0x7fffffffLL) return fail("edge id %lld does not fit int32", (long long)e);
        }
        list.insert(list.end(), tmp, tmp + cnt);
    }
    // per batch, the list position of its first edge (the list is ascending)
    const int64_t n_batches = (m + batch_size - 1) / batch_size;
    batch_ofs.assign(static_cast<size_t>(n_batches) + 1, 0);
    size_t pos = 0;
    for (int64_t b = 0; b < n_batches; ++b) {
        const int64_t first = b * batch_size;
        while (pos < list.size() && list[pos] < first) ++pos;
        batch_ofs[static_cast<size_t>(b)] = static_cast<int64_t>(pos);
    }
    batch_ofs[static_cast<size_t>(n_batches)] = static_cast<int64_t>(list.size());
    return 0;
}

// Pass B fills the compact staging image (arrays of `cap` >= list.size() entries, entry q describes
// edge list[q]) and derives the dependencies of OWNED endpoints (scratch indexed by local row: 1/world
// of the nodes, cache resident).  A non-owned endpoint arrives through the mailbox, tagged with the
// edge itself: no prev, no reader threshold.  prev_x are SPAN indices (the kernel turns them into
// tags); reader chains and the "both endpoints here" bookkeeping use list positions.
// kFlagDstFirst (only meaningful when both endpoints are owned): set iff dst is rewritten strictly
// before src.  NodeDep::pad remembers, per node, the latest such edge still waiting for that answer;
// the first later write of either endpoint settles it (a single edge rewriting both: no preference).
void fill_soa_sharded(zebra_tppr *h, char *host, size_t cap, const int32_t *src, const int32_t *dst, const int32_t *neg,
                      const double *ts, const int64_t *eidx, const std::vector<int32_t> &list, HostScratch *scratch = nullptr,
                      const ChainSet *cs = nullptr, ChainBuild *cb = nullptr)
{
    const int32_t pm = h->world - 1, me = h->rank, lg = h->world_log2;
    int32_t *o_src = reinterpret_cast<int32_t *>(host);
    int32_t *o_dst = o_src + cap, *o_neg = o_dst + cap, *o_e = o_neg + cap;
    int32_t *p_s = o_e + cap, *p_d = p_s + cap, *p_n = p_d + cap;
    uint32_t *r_s = reinterpret_cast<uint32_t *>(p_n + cap), *r_d = r_s + cap, *flg = r_d + cap;
    double *o_ts = reinterpret_cast<double *>(flg + cap);
    int32_t *o_cidx = reinterpret_cast<int32_t *>(o_ts + cap);
    const uint8_t *cmap = (cb && cs && cs->n > 0) ? cs->map.data() : nullptr;
    if (cb) cb->clear();
    NodeDep *dep = scratch ? scratch->dep.data() : h->dep.data();     // sized num_nodes >= local rows
    const size_t L = list.size();
    std::vector<int32_t> &rn = scratch ? scratch->rnext : h->rnext;
    if (rn.size() < L) rn.resize(L);
    int32_t *rnext = rn.data();
    auto settle = [&](uint32_t qe, int32_t x) {                       // node x of edge qe is the first to be rewritten
        const int32_t s = o_src[qe], d = o_dst[qe];
        if (x == d) flg[qe] |= zebra::kFlagDstFirst;
        NodeDep &other = dep[(x == d ? s : d) >> lg];
        if (other.pad == qe + 1) other.pad = 0;
        dep[x >> lg].pad = 0;
    };
    auto overwrite = [&](NodeDep &n, int32_t i, int32_t &prev, uint32_t &rthr) {
        prev = n.lastw; rthr = n.rcount;
        if (n.rcount) {                                               // its pending negative reads must signal
            for (int32_t r = n.rhead; r >= 0; r = rnext[r]) flg[r] |= zebra::kFlagSignal;
            n.rhead = -1; n.rcount = 0;
        }
        n.lastw = i;
    };
    constexpr size_t kAhead = 12;                                     // the scratch look-ups are cache misses: fetch ahead
    for (size_t q = 0; q < L; ++q) {
        if (q + kAhead < L) {
            const int32_t i2 = list[q + kAhead];
            const int32_t a2 = src[i2], b2 = dst[i2];
            if ((a2 & pm) == me) __builtin_prefetch(dep + (a2 >> lg), 1);
            if ((b2 & pm) == me) __builtin_prefetch(dep + (b2 >> lg), 1);
            if (neg && (neg[i2] & pm) == me) __builtin_prefetch(dep + (neg[i2] >> lg), 1);
        }
        const int32_t i = list[q];
        const int32_t a = src[i], b = dst[i], g = neg ? neg[i] : a;
        const bool oa = (a & pm) == me, wb = b != a && (b & pm) == me;
        const bool og = neg && g != a && g != b && (g & pm) == me;
        o_src[q] = a; o_dst[q] = b; o_neg[q] = g; o_e[q] = static_cast<int32_t>(eidx[i]); o_ts[q] = ts[i];
        flg[q] = 0; p_s[q] = -1; p_d[q] = -1; p_n[q] = -1; r_s[q] = 0; r_d[q] = 0;
        if (og) {
            NodeDep &n = dep[g >> lg];
            p_n[q] = n.lastw;
            rnext[q] = n.rhead; n.rhead = static_cast<int32_t>(q); n.rcount++;
        }
        NodeDep *da = oa ? dep + (a >> lg) : nullptr, *db = wb ? dep + (b >> lg) : nullptr;
        const uint32_t qa = da ? da->pad : 0u, qb = db ? db->pad : 0u;
        if (qa && qa == qb) {
            da->pad = 0; db->pad = 0;
        } else {
            if (qa) settle(qa - 1, a);
            if (qb) settle(qb - 1, b);
        }
        if (oa) overwrite(*da, i, p_s[q], r_s[q]);
        if (wb) overwrite(*db, i, p_d[q], r_d[q]);
        if (oa && wb) { da->pad = static_cast<uint32_t>(q) + 1; db->pad = static_cast<uint32_t>(q) + 1; }
        o_cidx[q] = -1;
        if (cmap) {
            // chain nodes this rank owns: their side of the edge goes to the resident chain warp.  A chain node of
            // ANOTHER rank joined to one of this rank's nodes: the edge goes to this rank's helpers of that chain (a
            // ticket warp would sit on it until the remote chain warp gets there, and with it the ticket front)
            const uint8_t ga = cmap[a], gb = (b != a) ? cmap[b] : 0;
            const uint8_t ca = (ga && oa) ? ga : 0, cbq = (gb && wb) ? gb : 0;
            if (ca | cbq) {
                const int32_t ci = cb->n_items++;
                o_cidx[q] = ci;
                const int32_t ee = static_cast<int32_t>(eidx[i]);
                const bool loop = b == a, b_here = (b & pm) == me, a_here = oa;
                if (ca) {
                    flg[q] |= zebra::kFlagSrcChain | (static_cast<uint32_t>(ca - 1) << zebra::kFlagSrcChainShift);
                    cb->steps[ca - 1].push_back(zebra::ChainStep{i, b, ee, (wb ? p_d[q] : -1), r_s[q],
                                                                 (loop ? zebra::kStepLoop : 0u) | (cbq ? zebra::kStepPeerChain : 0u) |
                                                                     ((!loop && !b_here) ? zebra::kStepPeerRemote : 0u),
                                                                 ci, p_s[q], ts[i], static_cast<int64_t>(q)});
                }
                if (cbq) {
                    flg[q] |= zebra::kFlagDstChain | (static_cast<uint32_t>(cbq - 1) << zebra::kFlagDstChainShift);
                    cb->steps[cbq - 1].push_back(zebra::ChainStep{i, a, ee, (oa ? p_s[q] : -1), r_d[q],
                                                                  zebra::kStepDst | (ca ? zebra::kStepPeerChain : 0u) |
                                                                      (!a_here ? zebra::kStepPeerRemote : 0u),
                                                                  ci, p_d[q], ts[i], static_cast<int64_t>(q)});
                }
            } else if (ga && !oa && wb) {
                flg[q] |= zebra::kFlagSrcChain | (static_cast<uint32_t>(ga - 1) << zebra::kFlagSrcChainShift);
                cb->steps[ga - 1].push_back(zebra::ChainStep{i, b, static_cast<int32_t>(eidx[i]), -1, 0u, 0u, -1, -1, ts[i], static_cast<int64_t>(q)});
            } else if (gb && !wb && (b & pm) != me && oa) {
                flg[q] |= zebra::kFlagDstChain | (static_cast<uint32_t>(gb - 1) << zebra::kFlagDstChainShift);
                cb->steps[gb - 1].push_back(zebra::ChainStep{i, a, static_cast<int32_t>(eidx[i]), -1, 0u, 0u, -1, -1, ts[i], static_cast<int64_t>(q)});
            }
        }
    }
    // the scratch goes back to "no writer, no reader": the local rows are few enough to refill wholesale
    const size_t local_rows = static_cast<size_t>((h->num_nodes - 1) >> lg) + 1;
    std::fill(dep, dep + local_rows, NodeDep{-1, -1, 0u, 0u});
}

struct BatchOut {
    int64_t batch_size = 0, first_batch = 0, out_slots = 1, slot_bytes = 0;
};

// Reserves n consecutive tags.  When the 32-bit sequence is about to wrap, it restarts at 1 and every
// handle zeroes the tags of its bound tables before its next launch.
int alloc_seq(zebra_tppr *h, int64_t n, cudaStream_t st, uint32_t *base)
{
    if (h->world > 1) {
        // every rank must draw the same tags for the same span: a per-handle sequence, advanced by the
        // same calls on every rank (SPMD), instead of the process-wide one
        if (static_cast<uint64_t>(h->shard_seq) + static_cast<uint64_t>(n) >= 0xF0000000ull) {
            const int per = 2 * h->k + 1;
            const int64_t total = h->table_rows * per;
            for (int j = 0; j < h->n_tppr; ++j)
                zebra::tppr_retag_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(h->tables[j], h->stride, h->k,
                                                                                                    h->table_rows);
            CU(cudaGetLastError());
            h->shard_seq = 1;
        }
        *base = h->shard_seq;
        h->shard_seq += static_cast<uint32_t>(n);
        return 0;
    }
    uint32_t epoch;
    {
        std::lock_guard<std::mutex> lk(g_seq_mu);
        if (static_cast<uint64_t>(g_seq) + static_cast<uint64_t>(n) >= 0xF0000000ull) { g_seq = 1; ++g_epoch; }
        *base = g_seq;
        g_seq += static_cast<uint32_t>(n);
        epoch = g_epoch;
    }
    if (h->epoch != epoch) {
        const int per = 2 * h->k + 1;
        const int64_t total = h->table_rows * per;
        for (int j = 0; j < h->n_tppr; ++j)
            zebra::tppr_retag_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(h->tables[j], h->stride, h->k,
                                                                                                h->table_rows);
        CU(cudaGetLastError());
        if (h->cmbox) CU(cudaMemsetAsync(h->cmbox, 0, h->cmbox_cap, st));    // stale mailbox tags could repeat as well
        h->epoch = epoch;
    }
    return 0;
}

template <typename NodeT>
int prepare_chunk(zebra_tppr *h, Staging &s, const NodeT *src, const NodeT *dst, const NodeT *neg, const double *ts,
                  const int64_t *eidx, int64_t m, const ChainSet &cs, ChainBuild &cb)
{
    const size_t cap = (static_cast<size_t>(m) + 7) & ~static_cast<size_t>(7);   // SoA stride of this launch
    return fill_soa<NodeT>(h, s.host, cap, src, dst, neg, ts, eidx, m, &cs, &cb);
}

// Launches the span kernel over m edges whose SoA arrays (stride `cap`) start at `dev`+`lo`.
int launch_kernel(zebra_tppr *h, const char *dev, size_t cap, int64_t lo, int64_t m, int64_t n_total, int64_t i_offset,
                  int32_t rows_per_edge, int32_t inst_mask, char *out, const BatchOut &bo, uint32_t seq0, int64_t tag_offset,
                  cudaStream_t st, const int32_t *list = nullptr, int64_t n_list = 0, const ChainLaunch *cl = nullptr)
{
    zebra::SpanParams p;
    memset(&p, 0, sizeof p);
    p.n_chunk = m; p.n_total = n_total; p.i_offset = i_offset;
    p.k = h->k; p.n_tppr = h->n_tppr; p.rows_per_edge = rows_per_edge; p.inst_mask = inst_mask; p.stride = h->stride;
    p.ticket_base = h->ticket_base;
    p.seq0 = seq0; p.tag_offset = tag_offset;
    const int32_t *a0 = reinterpret_cast<const int32_t *>(dev);
    p.src = a0 + lo; p.dst = a0 + cap + lo; p.neg = a0 + 2 * cap + lo; p.eidx = a0 + 3 * cap + lo;
    p.prev_s = a0 + 4 * cap + lo; p.prev_d = a0 + 5 * cap + lo; p.prev_n = a0 + 6 * cap + lo;
    const uint32_t *t0 = reinterpret_cast<const uint32_t *>(a0 + 7 * cap);
    p.rthr_s = t0 + lo; p.rthr_d = t0 + cap + lo; p.flags = t0 + 2 * cap + lo;
    p.ts = reinterpret_cast<const double *>(t0 + 3 * cap) + lo;
    for (int j = 0; j < h->n_tppr; ++j) {
        p.tables[j] = h->tables[j];
        p.rcnt[j] = h->rcnt + static_cast<size_t>(j) * h->num_nodes;
        p.alpha[j] = h->alpha[j]; p.beta[j] = h->beta[j];
        p.decay[j] = h->decay[j]; p.decay_len[j] = h->decay_len[j]; p.decay_fixed[j] = h->decay_fixed[j];
    }
    p.out = out; p.ticket = h->ticket; p.status = h->status; p.prof = h->prof;
    p.wait_ns = h->wait_ns;
    p.tie_mode = h->tie_mode;
    if (h->dbg_items) {
        const size_t need = static_cast<size_t>(std::max<int64_t>(h->world > 1 ? n_list : m, 1)) * h->n_tppr * sizeof(uint32_t);
        if (h->dbg_cap < need) {
            if (h->dbg_owner) CU(cudaFreeAsync(h->dbg_owner, st));
            CU(cudaMallocAsync(reinterpret_cast<void **>(&h->dbg_owner), need, st));
            h->dbg_cap = need;
        }
        CU(cudaMemsetAsync(h->dbg_owner, 0, need, st));
        p.dbg_owner = h->dbg_owner;
    }
    p.cidx = reinterpret_cast<const int32_t *>(dev + cap * kCidxOffset) + lo;
    int chain_warps = 0;
    if (cl && cl->n > 0) {
        // the mailbox of this launch: one slot per (chain item, instance, side)
        const size_t need = static_cast<size_t>(std::max(cl->n_items, 1)) * h->n_tppr * 2 * h->stride;
        if (h->cmbox_cap < need) {
            // stream-ordered (cudaFree / cudaMalloc would wait for every kernel on the device — also for another
            // rank's span kernel in the same process, which may be waiting for THIS launch)
            if (h->cmbox) CU(cudaFreeAsync(h->cmbox, st));
            h->cmbox = nullptr; h->cmbox_cap = 0;
            const size_t grow = need + need / 4;
            CU(cudaMallocAsync(reinterpret_cast<void **>(&h->cmbox), grow, st));
            CU(cudaMemsetAsync(h->cmbox, 0, grow, st));
            h->cmbox_cap = grow;
        }
        p.cprog = h->cprog;
        p.csteps = cl->dev_steps; p.cmbox = h->cmbox; p.cidx_base = cl->cidx_base; p.n_chains = cl->n;
        for (int q = 0; q < cl->n; ++q) p.chains[q] = zebra::ChainDesc{cl->node[q], cl->begin[q], cl->end[q], cl->owned[q]};
        chain_warps = cl->n * h->n_tppr;
        h->chain_items += cl->n_items;
        for (int q = 0; q < cl->n; ++q) h->chain_steps += cl->end[q] - cl->begin[q];
    }
    p.batch_size = bo.batch_size; p.first_batch = bo.first_batch; p.out_slots = bo.out_slots; p.slot_bytes = bo.slot_bytes;
    const bool sharded = h->world > 1;
    if (sharded && !list) return fail("node-sharded launch without a work list");
    p.list = list; p.n_list = n_list;
    const int64_t items = (sharded ? n_list : m) * h->n_tppr;
    int blocks = static_cast<int>(std::min<int64_t>((items + zebra::kWarpsPerBlock - 1) / zebra::kWarpsPerBlock,
                                                    sharded ? h->shard_max_blocks : h->max_blocks));
    if (blocks < 1) blocks = 1;
    if (chain_warps) {
        // one block per (chain, instance): its warp 0 walks the chain, the others help; all resident, and
        // always some blocks of ticket warps next to them (the chains wait for what those produce)
        const int cap_blocks = sharded ? h->shard_max_blocks : h->max_blocks;
        blocks = std::max(blocks, std::min(cap_blocks, chain_warps + std::max(16, chain_warps)));
        if (chain_warps + 8 > blocks) return fail("%d chain blocks do not fit a resident grid of %d blocks", chain_warps, cap_blocks);
        CU(cudaMemsetAsync(h->smflag, 0, 2 * zebra::kSmSlots * sizeof(int32_t), st));
        p.smflag = h->smflag;
        p.max_yield = (h->sm_yield && !sharded) ? std::max(0, (blocks - chain_warps) * zebra::kWarpsPerBlock - 256) : 0;
    }
    if (sharded) {
        if (m > h->max_window_edges) return fail("sharded launch of %lld edges exceeds the mailbox window (%lld)", (long long)m, (long long)h->max_window_edges);
        const size_t out_off = reinterpret_cast<size_t>(out);   // sharded: dev_out carries a byte OFFSET into the region
        if (out_off % 16) return fail("output offset must be a multiple of 16");
        for (int r = 0; r < h->world; ++r)
            if (!h->peer_arena[r]) return fail("zebra_tppr_shard_connect has not been called");
        if (rows_per_edge) {
            const size_t need = bo.batch_size > 0 ? static_cast<size_t>(bo.out_slots) * bo.slot_bytes
                                                  : zebra_tppr_out_bytes(h, n_total, rows_per_edge);
            if (need + out_off > h->out_region) return fail("output of %zu bytes at offset %zu exceeds the shared output region (%zu)", need, out_off, h->out_region);
        }
        p.world_log2 = h->world_log2; p.rank = h->rank; p.mbox_parity = static_cast<int32_t>(h->shard_launches & 1);
        p.row_mode = h->row_mode;
        p.shard_launch = static_cast<int32_t>(h->shard_launches);
        p.rec_offset = lo;                              // chain steps name their staging record by its index in the image
        p.row_guard = (h->epoch_active && rows_per_edge) ? 1 : 0;
        p.shard_done = reinterpret_cast<const int32_t *>(h->arena);     // header of the own arena: launches every rank has finished
        p.mbox_half_bytes = static_cast<int64_t>(h->mbox_half);
        for (int r = 0; r < h->world; ++r) {
            p.mbox_peer[r] = h->peer_arena[r] + kArenaHeader;
            p.out_peer[r] = h->peer_arena[r] + kArenaHeader + 2 * h->mbox_half + out_off;
        }
        p.out = nullptr;
        // pipelined epoch: the copy stream runs two one-warp kernels per window (wait for every rank, publish the
        // copy) while the span kernel of a later window may be spinning for exactly that publication — the
        // persistent grid must not fill the GPU to the last block slot, or those kernels could never start
        if (p.row_guard) blocks = std::max(1, std::min(blocks, h->shard_max_blocks - 8));
        zebra::tppr_span_shard_kernel<<<blocks, zebra::kWarpsPerBlock * 32, 0, st>>>(p);
        // stream-ordered behind the kernel: tell every rank that this rank has finished (and stopped reading
        // the mailbox half of) this launch; a peer pushes into that half again only after seeing it
        zebra::ShardDoneParams dp;
        for (int r = 0; r < zebra::kMaxRanks; ++r) dp.hdr[r] = r < h->world ? reinterpret_cast<int32_t *>(h->peer_arena[r]) : nullptr;
        dp.world = h->world; dp.rank = h->rank; dp.value = static_cast<int32_t>(h->shard_launches + 1); dp.slot = 0;
        zebra::shard_done_kernel<<<1, 32, 0, st>>>(dp);
        h->shard_launches++;
    } else {
        zebra::tppr_span_kernel<<<blocks, zebra::kWarpsPerBlock * 32, 0, st>>>(p);
    }
    CU(cudaGetLastError());
    // every warp of the grid draws exactly one ticket past the end
    h->ticket_base += static_cast<uint32_t>(items) + static_cast<uint32_t>(blocks) * zebra::kWarpsPerBlock;
    h->launches++;
    h->last_items += items;
    return 0;
}

// Appends the chain step records of a single-launch staging image right behind its arrays (chain after
// chain) and describes them for the launch.  Returns the bytes of the image to upload.
size_t place_chain_steps(const ChainSet &cs, char *host, char *dev, size_t cap, const ChainBuild &cb, ChainLaunch &cl,
                         size_t extra = 0)
{
    size_t bytes = cap * kEdgeBytes + extra;          // `extra`: bytes between the arrays and the step records (sharded: the work list)
    cl.n = 0;
    if (cs.n == 0 || cb.total() == 0) return bytes;
    zebra::ChainStep *dst = reinterpret_cast<zebra::ChainStep *>(host + bytes);
    int32_t pos = 0;
    cl.n = cs.n;
    for (int q = 0; q < cs.n; ++q) {
        cl.node[q] = cs.node[q];
        cl.owned[q] = cs.owned[q] ? 1 : 0;
        cl.begin[q] = pos;
        if (!cb.steps[q].empty()) memcpy(dst + pos, cb.steps[q].data(), cb.steps[q].size() * sizeof(zebra::ChainStep));
        pos += static_cast<int32_t>(cb.steps[q].size());
        cl.end[q] = pos;
    }
    cl.dev_steps = reinterpret_cast<const zebra::ChainStep *>(dev + bytes);
    cl.cidx_base = 0; cl.n_items = cb.n_items;
    return bytes + static_cast<size_t>(pos) * sizeof(zebra::ChainStep);
}

int launch_chunk(zebra_tppr *h, Staging &s, int64_t m, int64_t n_total, int64_t i_offset, int32_t rows_per_edge,
                 int32_t inst_mask, char *out, cudaStream_t st, const ChainSet &cs, const ChainBuild &cb)
{
    const size_t cap = (static_cast<size_t>(m) + 7) & ~static_cast<size_t>(7);
    uint32_t seq0 = 0;
    if (alloc_seq(h, m, st, &seq0)) return 1;
    ChainLaunch cl;
    const size_t bytes = place_chain_steps(cs, s.host, s.dev, cap, cb, cl);
    CU(cudaMemcpyAsync(s.dev, s.host, bytes, cudaMemcpyHostToDevice, st));
    if (i_offset == 0) CU(cudaEventRecord(h->ev0, st));
    if (launch_kernel(h, s.dev, cap, 0, m, n_total, i_offset, rows_per_edge, inst_mask, out, BatchOut(), seq0, 0, st, nullptr, 0, &cl)) return 1;
    CU(cudaEventRecord(s.free_ev, st));
    s.pending = true;
    return 0;
}

// Error text of a wait time-out: the first expired wait and the abandoned waits with the smallest edge tags
// (the root cause is normally the oldest edge that could not proceed).
int fail_timeout(zebra_tppr *h)
{
    const int32_t *sh = h->status_host;
    std::vector<std::array<int32_t, 4>> log;
    const int cnt = std::min<int>(sh[7], zebra::kWaitLog);
    for (int q = 0; q < cnt; ++q) log.push_back({sh[8 + 4 * q], sh[9 + 4 * q], sh[10 + 4 * q], sh[11 + 4 * q]});
    std::sort(log.begin(), log.end(), [](const std::array<int32_t, 4> &x, const std::array<int32_t, 4> &y) {
        return static_cast<uint32_t>(x[1]) < static_cast<uint32_t>(y[1]);
    });
    std::string txt;
    char buf[96];
    for (size_t q = 0; q < log.size() && q < 10; ++q) {
        snprintf(buf, sizeof buf, " [kind %d tag %u detail %u node %d]", log[q][0], static_cast<unsigned>(log[q][1]),
                 static_cast<unsigned>(log[q][2]), log[q][3]);
        txt += buf;
    }
    return fail("T-PPR span kernel: dependency wait timed out (first: kind %d, edge tag %u, detail %u; %d waits abandoned, oldest:%s; kinds: see Waiter in tppr_device.cuh)",
                sh[4], static_cast<unsigned>(sh[5]), static_cast<unsigned>(sh[6]), sh[7], txt.c_str());
}

int check_status(zebra_tppr *h, cudaStream_t st)
{
    CU(cudaMemcpyAsync(h->status_host, h->status, kStatusWords * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    h->last_ties = h->status_host[1];
    h->last_merge = h->status_host[2];
    if (h->status_host[0] != 0) {
        h->timeouts++;
        return fail_timeout(h);
    }
    return 0;
}

// Every way of streaming a span starts here.  The per-node counters of finished negative reads return
// to zero by themselves over a COMPLETE span (the host marks exactly the reads a later writer consumes);
// a span abandoned part-way (plan destroyed or re-started, failed launch, epoch closed early) leaves
// counts behind that would let a later writer overwrite a version its real readers have not loaded yet,
// so the counters are cleared whenever the previous span did not run to its end.
int begin_span(zebra_tppr *h, cudaStream_t st)
{
    if (h->rcnt_dirty) CU(cudaMemsetAsync(h->rcnt, 0, sizeof(int32_t) * h->num_nodes * h->n_tppr, st));
    h->rcnt_dirty = true;
    return 0;
}

template <typename NodeT>
int run_span(zebra_tppr *h, const NodeT *src, const NodeT *dst, const NodeT *neg, const double *ts, const int64_t *eidx,
             int64_t n, int32_t rows_per_edge, int32_t inst_mask, char *out, cudaStream_t st)
{
    for (int j = 0; j < h->n_tppr; ++j)
        if (!h->tables[j]) return fail("zebra_tppr_bind_tables has not been called");
    if (h->world > 1) return fail("node-sharded handles stream through plans (zebra_tppr_plan_run): dependencies must be span-global across ranks");
    if (h->plan_active || h->epoch_active) return fail("a plan or epoch is part-way through its batches; finish it before streaming other spans");
    CU(cudaSetDevice(h->device));
    CU(cudaMemsetAsync(h->status + 1, 0, 3 * sizeof(int32_t), st));   // [0] (time-out) is sticky until a reset
    h->last_items = 0;
    if (begin_span(h, st)) return 1;
    ChainSet cs;
    detect_chains<NodeT>(h, src, dst, n, cs);
    h->chain_items = 0; h->chain_steps = 0;
    ChainBuild cb;
    for (int64_t lo = 0; lo < n; lo += kChunk) {
        const int64_t m = std::min(kChunk, n - lo);
        Staging &s = h->stg[h->stg_next];
        h->stg_next ^= 1;
        if (ensure_staging(h, s, static_cast<size_t>(m) + 8)) return 1;
        if (s.pending) { CU(cudaEventSynchronize(s.free_ev)); s.pending = false; }
        if (prepare_chunk(h, s, src + lo, dst + lo, neg ? neg + lo : nullptr, ts + lo, eidx + lo, m, cs, cb)) return 1;
        if (launch_chunk(h, s, m, n, lo, rows_per_edge, inst_mask, out, st, cs, cb)) return 1;
    }
    CU(cudaEventRecord(h->ev1, st));
    h->timed = true;
    h->rcnt_dirty = false;
    return 0;
}

}  // namespace

extern "C" {

const char *zebra_last_error(void) { return g_err.c_str(); }

const char *zebra_build_info(void) { return "libzebra_tppr sm_100a fp64 no-fma " __DATE__ " " __TIME__; }

int zebra_tppr_create(zebra_tppr_t **out, int64_t num_nodes, int32_t k, int32_t n_tppr, const double *alpha_list,
                      const double *beta_list, int32_t device)
{
    if (!out) return fail("null out pointer");
    *out = nullptr;
    if (num_nodes <= 0) return fail("num_nodes must be positive");
    if (k < 1 || k > 31) return fail("k=%d unsupported: this build keeps one candidate pair per warp lane (1 <= k <= 31)", k);
    if (n_tppr < 1 || n_tppr > zebra::kMaxInst) return fail("n_tppr=%d unsupported (1..%d)", n_tppr, zebra::kMaxInst);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail("no CUDA device: libzebra_tppr has no CPU path (%s)", e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail("device %d out of range", device);
    CU(cudaSetDevice(device));
    zebra_tppr *h = new zebra_tppr();
    h->num_nodes = num_nodes; h->k = k; h->n_tppr = n_tppr; h->device = device;
    h->stride = static_cast<int32_t>(record_bytes(k));
    h->table_rows = num_nodes;
    for (int j = 0; j < n_tppr; ++j) { h->alpha[j] = alpha_list[j]; h->beta[j] = beta_list[j]; }
    h->dep.assign(static_cast<size_t>(num_nodes), NodeDep{-1, -1, 0u, 0u});
    CU(cudaMalloc(&h->rcnt, sizeof(int32_t) * num_nodes * n_tppr));
    CU(cudaMemset(h->rcnt, 0, sizeof(int32_t) * num_nodes * n_tppr));
    { std::lock_guard<std::mutex> lk(g_seq_mu); h->epoch = g_epoch; }
    for (int j = 0; j < n_tppr; ++j) {
        // tabulate the norm recurrence up to its floating-point fixed point (or a cap; beyond a capped
        // table the kernel divides as the reference does)
        constexpr int kCap = 1 << 16;
        std::vector<zebra::DecayStep> tab;
        double norm = 0.0;
        int fixed = 0;
        while (static_cast<int>(tab.size()) < kCap) {
            const zebra::DecayStep d = zebra::decay_step(norm, h->alpha[j], h->beta[j]);
            tab.push_back(d);
            if (d.new_norm == norm) { fixed = 1; break; }
            norm = d.new_norm;
        }
        h->decay_len[j] = static_cast<int32_t>(tab.size());
        h->decay_fixed[j] = fixed;
        CU(cudaMalloc(&h->decay[j], tab.size() * sizeof(zebra::DecayStep)));
        CU(cudaMemcpy(h->decay[j], tab.data(), tab.size() * sizeof(zebra::DecayStep), cudaMemcpyHostToDevice));
    }
    CU(cudaMalloc(&h->cprog, sizeof(uint32_t) * zebra::kMaxChains * zebra::kMaxInst));
    CU(cudaMemset(h->cprog, 0, sizeof(uint32_t) * zebra::kMaxChains * zebra::kMaxInst));
    CU(cudaMalloc(&h->smflag, 2 * zebra::kSmSlots * sizeof(int32_t)));
    CU(cudaMemset(h->smflag, 0, 2 * zebra::kSmSlots * sizeof(int32_t)));
    CU(cudaMalloc(&h->ticket, sizeof(uint32_t)));
    CU(cudaMemset(h->ticket, 0, sizeof(uint32_t)));
    CU(cudaMalloc(&h->status, kStatusWords * sizeof(int32_t)));
    CU(cudaMemset(h->status, 0, kStatusWords * sizeof(int32_t)));
    CU(cudaMallocHost(&h->status_host, kStatusWords * sizeof(int32_t)));
    CU(cudaMalloc(&h->prof, 16 * sizeof(unsigned long long)));
    CU(cudaMemset(h->prof, 0, 16 * sizeof(unsigned long long)));
    CU(cudaEventCreate(&h->ev0));
    CU(cudaEventCreate(&h->ev1));
    int per_sm = 0, sms = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, zebra::tppr_span_kernel, zebra::kWarpsPerBlock * 32, 0));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    h->max_blocks = std::max(1, per_sm * sms);
    if (const char *e = getenv("ZEBRA_CHAINS")) {
        h->max_chains = std::max(0, std::min(atoi(e), zebra::kMaxChains));
        h->chains_enabled = h->max_chains > 0;
    }
    if (const char *e = getenv("ZEBRA_HOST_THREADS")) h->host_threads = std::max(0, atoi(e));
    if (const char *e = getenv("ZEBRA_SHARD_CHAINS")) h->shard_chains = atoi(e) != 0;
    if (const char *e = getenv("ZEBRA_SM_YIELD")) h->sm_yield = atoi(e) != 0;
    if (const char *e = getenv("ZEBRA_DEBUG_ITEMS")) h->dbg_items = atoi(e) != 0;
    if (const char *e = getenv("ZEBRA_WAIT_TIMEOUT_MS")) h->wait_ns = static_cast<unsigned long long>(std::max(1, atoi(e))) * 1000000ull;
    *out = h;
    return 0;
}

int zebra_tppr_destroy(zebra_tppr_t *h)
{
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (auto &s : h->stg) {
        if (s.host) cudaFreeHost(s.host);
        if (s.dev) cudaFree(s.dev);
        if (s.free_ev) cudaEventDestroy(s.free_ev);
    }
    for (auto &sl : h->eslot) {
        if (sl.stg.host) cudaFreeHost(sl.stg.host);
        if (sl.stg.dev) cudaFree(sl.stg.dev);
        if (sl.stg.free_ev) cudaEventDestroy(sl.stg.free_ev);
        if (sl.out_dev) cudaFree(sl.out_dev);
        if (sl.out_host) cudaFreeHost(sl.out_host);
        if (sl.ev_kernel) cudaEventDestroy(sl.ev_kernel);
        if (sl.ev_copy) cudaEventDestroy(sl.ev_copy);
    }
    for (auto &ps : h->prep) {
        if (ps.host) cudaFreeHost(ps.host);
        if (ps.dev) cudaFree(ps.dev);
        if (ps.free_ev) cudaEventDestroy(ps.free_ev);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (int r = 0; r < zebra::kMaxRanks; ++r)
        if (h->peer_opened[r] && h->peer_arena[r]) cudaIpcCloseMemHandle(h->peer_arena[r]);
    if (h->arena) cudaFree(h->arena);
    if (h->own_tables)
        for (auto &t : h->tables) if (t) cudaFree(t);
    if (h->rcnt) cudaFree(h->rcnt);
    if (h->cmbox) cudaFree(h->cmbox);
    if (h->cprog) cudaFree(h->cprog);
    if (h->smflag) cudaFree(h->smflag);
    if (h->dbg_owner) cudaFree(h->dbg_owner);
    for (auto d : h->decay) if (d) cudaFree(d);
    if (h->ticket) cudaFree(h->ticket);
    if (h->status) cudaFree(h->status);
    if (h->status_host) cudaFreeHost(h->status_host);
    if (h->prof) cudaFree(h->prof);
    if (h->out_dev) cudaFree(h->out_dev);
    if (h->out_pinned) cudaFreeHost(h->out_pinned);
    if (h->plan_stage) cudaFreeHost(h->plan_stage);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    delete h;
    return 0;
}

size_t zebra_tppr_record_bytes(const zebra_tppr_t *h) { return h ? static_cast<size_t>(h->stride) : 0; }
size_t zebra_tppr_table_bytes(const zebra_tppr_t *h) { return h ? static_cast<size_t>(h->stride) * h->table_rows : 0; }

int zebra_tppr_bind_tables(zebra_tppr_t *h, void *const *dev_tables)
{
    if (!h || !dev_tables) return fail("null argument");
    if (h->own_tables) return fail("node-sharded handles own their tables (zebra_tppr_shard_reset zero-fills them)");
    for (int j = 0; j < h->n_tppr; ++j) {
        if (!dev_tables[j]) return fail("table %d is null", j);
        if (reinterpret_cast<uintptr_t>(dev_tables[j]) % 16) return fail("table %d is not 16-byte aligned", j);
        h->tables[j] = static_cast<char *>(dev_tables[j]);
    }
    return 0;
}

int zebra_tppr_reset_tickets(zebra_tppr_t *h, void *stream)
{
    if (!h) return fail("null handle");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(st));
    CU(cudaMemsetAsync(h->rcnt, 0, sizeof(int32_t) * h->num_nodes * h->n_tppr, st));
    CU(cudaMemsetAsync(h->status, 0, kStatusWords * sizeof(int32_t), st));       // also clears a recorded wait time-out
    h->plan_active = nullptr;
    h->rcnt_dirty = false;
    return 0;
}

size_t zebra_tppr_out_bytes(const zebra_tppr_t *h, int64_t n, int32_t rows_per_edge)
{
    if (!h) return 0;
    return static_cast<size_t>(h->n_tppr) * 16 * static_cast<size_t>(rows_per_edge) * n * h->k;
}

int zebra_tppr_stream(zebra_tppr_t *h, const int32_t *host_nodes, const double *host_ts, const int64_t *host_eidx,
                      int64_t n, int32_t rows_per_edge, int32_t inst, void *dev_out, void *host_out, void *stream)
{
    if (!h) return fail("null handle");
    if (n < 0) return fail("negative edge count");
    if (rows_per_edge != 0 && rows_per_edge != 2 && rows_per_edge != 3) return fail("rows_per_edge must be 0, 2 or 3");
    if (inst < -1 || inst >= h->n_tppr) return fail("instance %d out of range", inst);
    if (n == 0) return 0;
    if (!host_nodes || !host_ts || !host_eidx) return fail("null input");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CU(cudaSetDevice(h->device));
    const size_t ob = zebra_tppr_out_bytes(h, n, rows_per_edge);
    char *out = static_cast<char *>(dev_out);
    if (rows_per_edge && !out) {
        if (h->out_cap < ob) {
            CU(cudaStreamSynchronize(st));
            if (h->out_dev) CU(cudaFree(h->out_dev));
            CU(cudaMalloc(&h->out_dev, ob));
            h->out_cap = ob;
        }
        out = h->out_dev;
    }
    const int32_t mask = inst < 0 ? ((1 << h->n_tppr) - 1) : (1 << inst);
    if (rows_per_edge && inst >= 0) CU(cudaMemsetAsync(out, 0, ob, st));
    const int32_t *neg = rows_per_edge == 3 ? host_nodes + 2 * n : nullptr;
    if (run_span<int32_t>(h, host_nodes, host_nodes + n, neg, host_ts, host_eidx, n, rows_per_edge, mask, out, st)) return 1;
    if (host_out && rows_per_edge) {
        CU(cudaMemcpyAsync(host_out, out, ob, cudaMemcpyDeviceToHost, st));
        return check_status(h, st);
    }
    if (host_out || !rows_per_edge) return check_status(h, st);
    return 0;
}

int zebra_tppr_update_only(zebra_tppr_t *h, const int64_t *host_src, const int64_t *host_dst, const double *host_ts,
                           const int64_t *host_eidx, int64_t n_edges, void *stream)
{
    if (!h) return fail("null handle");
    if (n_edges < 0) return fail("negative edge count");
    if (n_edges == 0) return 0;
    if (!host_src || !host_dst || !host_ts || !host_eidx) return fail("null input");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (run_span<int64_t>(h, host_src, host_dst, nullptr, host_ts, host_eidx, n_edges, 0, (1 << h->n_tppr) - 1, nullptr, st))
        return 1;
    return check_status(h, st);
}

int zebra_tppr_read_state(zebra_tppr_t *h, int32_t inst, const int64_t *host_ids, int64_t n_ids, int32_t *counts,
                          int32_t *eidx, int32_t *node, double *ts, double *score, double *norm, void *stream)
{
    if (!h) return fail("null handle");
    if (inst < 0 || inst >= h->n_tppr || !h->tables[inst]) return fail("bad instance");
    if (n_ids <= 0) return 0;
    for (int64_t a = 0; a < n_ids; ++a)
        if (host_ids[a] < 0 || host_ids[a] >= h->num_nodes) return fail("node id out of range");
    std::vector<int64_t> local;
    if (h->world > 1) {          // node-sharded: only this rank's nodes, addressed by local row
        local.resize(static_cast<size_t>(n_ids));
        for (int64_t a = 0; a < n_ids; ++a) {
            if ((host_ids[a] & (h->world - 1)) != h->rank) return fail("node %lld lives on rank %lld", (long long)host_ids[a], (long long)(host_ids[a] & (h->world - 1)));
            local[static_cast<size_t>(a)] = host_ids[a] >> h->world_log2;
        }
        host_ids = local.data();
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CU(cudaSetDevice(h->device));
    const int k = h->k;
    const size_t nk = static_cast<size_t>(n_ids) * k;
    const size_t bytes = n_ids * 8 + n_ids * 4 + nk * 4 * 2 + nk * 8 * 2 + n_ids * 8;
    char *d = nullptr;
    CU(cudaMalloc(&d, bytes));
    int64_t *d_ids = reinterpret_cast<int64_t *>(d);
    double *d_ts = reinterpret_cast<double *>(d_ids + n_ids), *d_sc = d_ts + nk, *d_norm = d_sc + nk;
    int32_t *d_cnt = reinterpret_cast<int32_t *>(d_norm + n_ids), *d_e = d_cnt + n_ids, *d_n = d_e + nk;
    cudaError_t e = cudaMemcpyAsync(d_ids, host_ids, n_ids * 8, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        zebra::tppr_read_state_kernel<<<static_cast<unsigned>(n_ids), 32, 0, st>>>(h->tables[inst], h->stride, k, d_ids, n_ids,
                                                                                  d_cnt, d_e, d_n, d_ts, d_sc, d_norm);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(counts, d_cnt, n_ids * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(eidx, d_e, nk * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(node, d_n, nk * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ts, d_ts, nk * 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(score, d_sc, nk * 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(norm, d_norm, n_ids * 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d);
    if (e != cudaSuccess) return fail("read_state failed: %s", cudaGetErrorString(e));
    return 0;
}

int zebra_tppr_plan_create(zebra_tppr_t *h, const int32_t *host_src, const int32_t *host_dst, const int32_t *host_neg,
                           const double *host_ts, const int64_t *host_eidx, int64_t n, int32_t rows_per_edge,
                           int64_t batch_size, zebra_plan_t **out)
{
    if (!h || !out) return fail("null argument");
    *out = nullptr;
    if (n <= 0 || batch_size <= 0) return fail("plan needs n > 0 and batch_size > 0");
    if (rows_per_edge != 0 && rows_per_edge != 2 && rows_per_edge != 3) return fail("rows_per_edge must be 0, 2 or 3");
    if (!host_src || !host_dst || !host_ts || !host_eidx || (rows_per_edge == 3 && !host_neg)) return fail("null input");
    CU(cudaSetDevice(h->device));
    zebra_plan *pl = new zebra_plan();
    pl->h = h; pl->n = n; pl->batch_size = batch_size; pl->rows_per_edge = rows_per_edge;
    pl->n_batches = (n + batch_size - 1) / batch_size;
    pl->cap = (static_cast<size_t>(n) + 7) & ~static_cast<size_t>(7);
    if (n >= 0x70000000LL || batch_size >= 0x70000000LL) { delete pl; return fail("plan too long"); }
    std::vector<int32_t> list;
    cudaError_t e = cudaSuccess;
    if (h->world > 1) {
        // compact records, one per edge of this rank's work list, staged in a pinned buffer the handle keeps
        if (shard_work_list(h, host_src, host_dst, rows_per_edge == 3 ? host_neg : nullptr, host_eidx, n, batch_size, list,
                            pl->batch_ofs)) {
            delete pl;
            return 1;
        }
        pl->cap = (list.size() + 8) & ~static_cast<size_t>(7);
        const size_t bytes = pl->cap * kEdgeBytes;
        if (h->plan_stage_cap < bytes) {
            if (h->plan_stage) cudaFreeHost(h->plan_stage);
            h->plan_stage = nullptr; h->plan_stage_cap = 0;
            e = cudaMallocHost(&h->plan_stage, bytes + bytes / 8);
            if (e != cudaSuccess) { delete pl; return fail("pinned staging of %zu bytes: %s", bytes, cudaGetErrorString(e)); }
            h->plan_stage_cap = bytes + bytes / 8;
        }
        ChainSet cs;
        ChainBuild cb;
        detect_chains<int32_t>(h, host_src, host_dst, n, cs);
        fill_soa_sharded(h, h->plan_stage, pl->cap, host_src, host_dst, rows_per_edge == 3 ? host_neg : nullptr, host_ts, host_eidx, list,
                         nullptr, &cs, &cb);
        std::vector<zebra::ChainStep> all;
        if (cs.n > 0 && cb.total() > 0) {       // (helper steps of another rank's chain count too)
            pl->chain_n = cs.n;
            all.reserve(cb.total());
            pl->item_edge.assign(static_cast<size_t>(cb.n_items), 0);
            for (int q = 0; q < cs.n; ++q) {
                pl->chain_node[q] = cs.node[q];
                pl->chain_owned[q] = cs.owned[q];
                pl->chain_off[q] = static_cast<int32_t>(all.size());
                pl->chain_edges[q].reserve(cb.steps[q].size());
                for (const zebra::ChainStep &st : cb.steps[q]) {
                    pl->chain_edges[q].push_back(st.edge);
                    if (st.cidx >= 0) pl->item_edge[static_cast<size_t>(st.cidx)] = st.edge;
                    all.push_back(st);
                }
            }
            pl->chain_off[cs.n] = static_cast<int32_t>(all.size());
        }
        pl->steps_bytes = all.size() * sizeof(zebra::ChainStep);
        e = cudaMalloc(&pl->dev, bytes + std::max<size_t>(pl->steps_bytes, 16));
        if (e == cudaSuccess) e = cudaMemcpy(pl->dev, h->plan_stage, bytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && pl->steps_bytes) e = cudaMemcpy(pl->dev + bytes, all.data(), pl->steps_bytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMalloc(&pl->dev_list, std::max<size_t>(list.size(), 1) * sizeof(int32_t));
        if (e == cudaSuccess) e = cudaMemcpy(pl->dev_list, list.data(), list.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
    } else {
        // staging image of the span, followed by the step records of its chain nodes (chain after chain)
        std::unique_ptr<char[]> host(new char[pl->cap * kEdgeBytes]);
        ChainSet cs;
        ChainBuild cb;
        detect_chains<int32_t>(h, host_src, host_dst, n, cs);
        if (fill_soa<int32_t>(h, host.get(), pl->cap, host_src, host_dst, rows_per_edge == 3 ? host_neg : nullptr, host_ts,
                              host_eidx, n, &cs, &cb)) {
            delete pl;
            return 1;
        }
        std::vector<zebra::ChainStep> all;
        if (cs.n > 0 && cb.total() > 0) {       // (helper steps of another rank's chain count too)
            pl->chain_n = cs.n;
            all.reserve(cb.total());
            pl->item_edge.assign(static_cast<size_t>(cb.n_items), 0);
            for (int q = 0; q < cs.n; ++q) {
                pl->chain_node[q] = cs.node[q];
                pl->chain_owned[q] = cs.owned[q];
                pl->chain_off[q] = static_cast<int32_t>(all.size());
                pl->chain_edges[q].reserve(cb.steps[q].size());
                for (const zebra::ChainStep &st : cb.steps[q]) {
                    pl->chain_edges[q].push_back(st.edge);
                    if (st.cidx >= 0) pl->item_edge[static_cast<size_t>(st.cidx)] = st.edge;
                    all.push_back(st);
                }
            }
            pl->chain_off[cs.n] = static_cast<int32_t>(all.size());
        }
        pl->steps_bytes = all.size() * sizeof(zebra::ChainStep);
        e = cudaMalloc(&pl->dev, pl->cap * kEdgeBytes + std::max<size_t>(pl->steps_bytes, 16));
        if (e == cudaSuccess) e = cudaMemcpy(pl->dev, host.get(), pl->cap * kEdgeBytes, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && pl->steps_bytes)
            e = cudaMemcpy(pl->dev + pl->cap * kEdgeBytes, all.data(), pl->steps_bytes, cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        if (pl->dev) cudaFree(pl->dev);
        if (pl->dev_list) cudaFree(pl->dev_list);
        delete pl;
        return fail("plan upload failed: %s", cudaGetErrorString(e));
    }
    *out = pl;
    return 0;
}

int zebra_tppr_plan_destroy(zebra_plan_t *pl)
{
    if (!pl) return 0;
    if (pl->h && pl->h->plan_active == pl) pl->h->plan_active = nullptr;
    if (pl->dev) { cudaSetDevice(pl->h->device); cudaDeviceSynchronize(); cudaFree(pl->dev); }
    if (pl->dev_list) cudaFree(pl->dev_list);
    delete pl;
    return 0;
}

int64_t zebra_tppr_plan_batches(const zebra_plan_t *pl) { return pl ? pl->n_batches : 0; }

// Test hook (host only, no CUDA call): the staging records a plan would upload for this span.
int zebra_debug_host_pass(int64_t num_nodes, int32_t world, int32_t rank, const int32_t *host_src, const int32_t *host_dst,
                          const int32_t *host_neg, const double *host_ts, const int64_t *host_eidx, int64_t n,
                          int64_t batch_size, int32_t *out_fields, double *out_ts, int32_t *out_list, int64_t *out_batch_ofs,
                          int64_t *out_count)
{
    if (num_nodes <= 0 || n <= 0 || batch_size <= 0 || world < 1 || (world & (world - 1)) || world > zebra::kMaxRanks ||
        rank < 0 || rank >= world)
        return fail("bad host-pass geometry");
    if (!host_src || !host_dst || !host_ts || !host_eidx || !out_fields || !out_ts || !out_list || !out_batch_ofs || !out_count)
        return fail("null argument");
    zebra_tppr h;
    h.num_nodes = num_nodes; h.world = world; h.rank = rank;
    while ((1 << h.world_log2) < world) ++h.world_log2;
    h.dep.assign(static_cast<size_t>(num_nodes), NodeDep{-1, -1, 0u, 0u});
    std::vector<int32_t> list;
    std::vector<int64_t> ofs;
    size_t cap = 0, count = 0;
    std::unique_ptr<char[]> img;
    if (world > 1) {
        if (shard_work_list(&h, host_src, host_dst, host_neg, host_eidx, n, batch_size, list, ofs)) return 1;
        count = list.size();
        cap = (count + 8) & ~static_cast<size_t>(7);
        img.reset(new char[cap * kEdgeBytes]);
        ChainSet cs;
        ChainBuild cb;
        h.shard_max_blocks = 296;
        h.shard_chains = true;
        detect_chains<int32_t>(&h, host_src, host_dst, n, cs);
        fill_soa_sharded(&h, img.get(), cap, host_src, host_dst, host_neg, host_ts, host_eidx, list, nullptr, &cs, &cb);
        {
            // every list entry with a chain flag must be taken by exactly one helper step; no other entry by any
            const uint32_t *fl = reinterpret_cast<const uint32_t *>(img.get()) + 9 * cap;
            std::vector<int> taken(count, 0);
            for (int q = 0; q < cs.n; ++q)
                for (const zebra::ChainStep &st : cb.steps[q]) {
                    if ((st.flags & zebra::kStepPeerChain) && (st.flags & zebra::kStepDst)) continue;
                    if (st.rec < 0 || st.rec >= static_cast<int64_t>(count)) return fail("chain step names record %lld of %zu", (long long)st.rec, count);
                    taken[static_cast<size_t>(st.rec)]++;
                }
            for (size_t q = 0; q < count; ++q) {
                const bool flagged = (fl[q] & (zebra::kFlagSrcChain | zebra::kFlagDstChain)) != 0;
                if (taken[q] != (flagged ? 1 : 0))
                    return fail("list entry %zu (edge %d): chain flags %u but taken by %d helper steps", q, list[q], fl[q] & 12u, taken[q]);
            }
        }
    } else {
        count = static_cast<size_t>(n);
        cap = (count + 7) & ~static_cast<size_t>(7);
        img.reset(new char[cap * kEdgeBytes]);
        if (fill_soa<int32_t>(&h, img.get(), cap, host_src, host_dst, host_neg, host_ts, host_eidx, n)) return 1;
        list.resize(count);
        for (size_t q = 0; q < count; ++q) list[q] = static_cast<int32_t>(q);
        const int64_t nb = (n + batch_size - 1) / batch_size;
        ofs.resize(static_cast<size_t>(nb) + 1);
        for (int64_t b = 0; b <= nb; ++b) ofs[static_cast<size_t>(b)] = std::min(n, b * batch_size);
    }
    for (const NodeDep &d : h.dep)
        if (d.lastw != -1 || d.rhead != -1 || d.rcount != 0 || d.pad != 0) return fail("host pass left its scratch dirty");
    const int32_t *f = reinterpret_cast<const int32_t *>(img.get());
    for (int a = 0; a < 10; ++a) memcpy(out_fields + static_cast<size_t>(a) * count, f + static_cast<size_t>(a) * cap, count * sizeof(int32_t));
    memcpy(out_ts, f + 10 * cap, count * sizeof(double));
    memcpy(out_list, list.data(), count * sizeof(int32_t));
    memcpy(out_batch_ofs, ofs.data(), ofs.size() * sizeof(int64_t));
    *out_count = static_cast<int64_t>(count);
    return 0;
}

// Test hook: the exact-tie ordering of the prune in isolation (tppr_device.cuh tie_order): final argsort
// ranks of n_cases arrays of small integers, to be compared with numba's argsort on the host.
int zebra_debug_tie_order(const int32_t *host_values, const int32_t *host_n, int32_t n_cases, int32_t stride, int32_t *host_ranks)
{
    if (!host_values || !host_n || !host_ranks || n_cases <= 0 || stride < 64) return fail("bad argument");
    for (int c = 0; c < n_cases; ++c)
        if (host_n[c] < 2 || host_n[c] > 63) return fail("case %d: n must be 2..63", c);
    int32_t *dv = nullptr, *dn = nullptr, *dout = nullptr;
    const size_t bytes = static_cast<size_t>(n_cases) * stride * sizeof(int32_t);
    cudaError_t e = cudaMalloc(&dv, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&dn, n_cases * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMalloc(&dout, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(dv, host_values, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dn, host_n, n_cases * sizeof(int32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(dout, 0xff, bytes);
    if (e == cudaSuccess) {
        zebra::tie_order_test_kernel<<<n_cases, 32>>>(dv, dn, stride, dout);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(host_ranks, dout, bytes, cudaMemcpyDeviceToHost);
    cudaFree(dv); cudaFree(dn); cudaFree(dout);
    if (e != cudaSuccess) return fail("tie order test: %s", cudaGetErrorString(e));
    return 0;
}

// Test / debug hook: the raw bytes of one node record (header and chunks with their tags).
int zebra_debug_read_record(zebra_tppr_t *h, int32_t inst, int64_t node, void *host_out)
{
    if (!h || !host_out || inst < 0 || inst >= h->n_tppr || !h->tables[inst]) return fail("bad argument");
    if (node < 0 || node >= h->table_rows) return fail("node out of range");
    CU(cudaSetDevice(h->device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(host_out, h->tables[inst] + static_cast<size_t>(node) * h->stride, h->stride, cudaMemcpyDeviceToHost));
    return 0;
}

size_t zebra_tppr_plan_bytes(const zebra_plan_t *pl)
{
    if (!pl) return 0;
    return pl->cap * kEdgeBytes + pl->steps_bytes + (pl->dev_list ? static_cast<size_t>(pl->batch_ofs.back()) * sizeof(int32_t) : 0);
}

int zebra_tppr_plan_run(zebra_tppr_t *h, zebra_plan_t *pl, int64_t first_batch, int64_t n_batches,
                        int64_t batches_per_launch, void *dev_out, int64_t out_slots, void *stream)
{
    if (!h || !pl || pl->h != h) return fail("plan does not belong to this handle");
    if (n_batches <= 0) return 0;
    if (batches_per_launch <= 0) return fail("batches_per_launch must be positive");
    if (first_batch < 0 || first_batch + n_batches > pl->n_batches) return fail("batch range out of the plan");
    for (int j = 0; j < h->n_tppr; ++j)
        if (!h->tables[j]) return fail("zebra_tppr_bind_tables has not been called");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CU(cudaSetDevice(h->device));
    if (first_batch == 0) {
        if (begin_span(h, st)) return 1;   // clears what a run abandoned part-way left behind
        if (alloc_seq(h, pl->n, st, &pl->seq0)) return 1;
        h->plan_active = pl;
        h->plan_next = 0;
    }
    if (h->plan_active != pl || h->plan_next != first_batch)
        return fail("plan batches must run in order starting from 0 (next expected %lld)", (long long)h->plan_next);
    if (pl->rows_per_edge) {
        if (out_slots <= 0 || (h->world == 1 && !dev_out)) return fail("plan with output rows needs dev_out and out_slots");
        if (out_slots < std::min(batches_per_launch, n_batches))
            return fail("out_slots must cover one launch window");
    }
    BatchOut bo;
    bo.batch_size = pl->batch_size; bo.first_batch = first_batch; bo.out_slots = std::max<int64_t>(out_slots, 1);
    bo.slot_bytes = static_cast<int64_t>(zebra_tppr_out_bytes(h, pl->batch_size, pl->rows_per_edge));
    CU(cudaMemsetAsync(h->status + 1, 0, 3 * sizeof(int32_t), st));
    h->last_items = 0;
    h->chain_items = 0; h->chain_steps = 0;
    CU(cudaEventRecord(h->ev0, st));
    const int64_t last = first_batch + n_batches;
    for (int64_t b = first_batch; b < last; b += batches_per_launch) {
        const int64_t be = std::min(last, b + batches_per_launch);
        const int64_t lo = b * pl->batch_size, hi = std::min(pl->n, be * pl->batch_size);
        const int32_t *list = nullptr;
        int64_t n_list = 0, rec0 = lo;                 // rec0: first staging record of the launch
        if (h->world > 1) {
            // sharded plans hold compact records: entry q of the work list describes edge list[q]
            rec0 = pl->batch_ofs[static_cast<size_t>(b)];
            list = pl->dev_list + rec0;
            n_list = pl->batch_ofs[static_cast<size_t>(be)] - rec0;
        }
        ChainLaunch cl;
        if (pl->chain_n > 0) {
            // the chain steps and chain items of this launch: sub-ranges of the plan's (ascending) lists
            cl.n = pl->chain_n;
            const int32_t lo32 = static_cast<int32_t>(lo), hi32 = static_cast<int32_t>(hi);
            for (int q = 0; q < pl->chain_n; ++q) {
                const std::vector<int32_t> &ce = pl->chain_edges[q];
                cl.node[q] = pl->chain_node[q];
                cl.owned[q] = pl->chain_owned[q] ? 1 : 0;
                cl.begin[q] = pl->chain_off[q] + static_cast<int32_t>(std::lower_bound(ce.begin(), ce.end(), lo32) - ce.begin());
                cl.end[q] = pl->chain_off[q] + static_cast<int32_t>(std::lower_bound(ce.begin(), ce.end(), hi32) - ce.begin());
            }
            cl.dev_steps = reinterpret_cast<const zebra::ChainStep *>(pl->dev + pl->cap * kEdgeBytes);
            cl.cidx_base = static_cast<int32_t>(std::lower_bound(pl->item_edge.begin(), pl->item_edge.end(), lo32) - pl->item_edge.begin());
            cl.n_items = static_cast<int32_t>(std::lower_bound(pl->item_edge.begin(), pl->item_edge.end(), hi32) - pl->item_edge.begin()) - cl.cidx_base;
        }
        if (launch_kernel(h, pl->dev, pl->cap, rec0, hi - lo, pl->n, lo, pl->rows_per_edge, (1 << h->n_tppr) - 1,
                          static_cast<char *>(dev_out), bo, pl->seq0, lo, st, list, n_list, &cl)) {
            h->plan_active = nullptr;          // a failed launch abandons the run
            return 1;
        }
    }
    CU(cudaEventRecord(h->ev1, st));
    h->timed = true;
    h->plan_next = last;
    if (last == pl->n_batches) { h->plan_active = nullptr; h->rcnt_dirty = false; }
    return 0;
}

/* ---- pipelined epoch ---- */

// Host dependency pass of window w into its staging slot (pinned memory): the edge records, the chain step
// records behind them and the launch geometry.  Runs on the caller's thread (scratch = the handle's) or on a
// worker thread of the epoch (scratch = the worker's own); touches no CUDA state.
static void epoch_prepare(zebra_epoch *ep, int64_t w, HostScratch *scratch)
{
    zebra_tppr *h = ep->h;
    zebra_epoch::Prep &pr = ep->prep[w % zebra_tppr::kPrepSlots];
    Staging &stg = h->prep[w % zebra_tppr::kPrepSlots];
    const int64_t b0 = w * ep->window, b1 = std::min(ep->n_batches, b0 + ep->window);
    const int64_t lo = b0 * ep->batch_size, hi = std::min(ep->n, b1 * ep->batch_size);
    const int64_t m = hi - lo;
    if (h->world > 1) {
        // node-sharded: the work list of this rank within the window, compact records for it, then the list
        // itself (as span indices) behind the records
        const int32_t *ng = ep->neg ? ep->neg + lo : nullptr;
        pr.failed = shard_work_list(h, ep->src + lo, ep->dst + lo, ng, ep->eidx + lo, m, ep->batch_size, pr.list, pr.batch_ofs) != 0;
        if (pr.failed) { pr.err = zebra_last_error(); pr.window = w; return; }
        pr.cap = (pr.list.size() + 8) & ~static_cast<size_t>(7);
        fill_soa_sharded(h, stg.host, pr.cap, ep->src + lo, ep->dst + lo, ng, ep->ts + lo, ep->eidx + lo, pr.list, scratch,
                         &ep->cs, &pr.cb);
        int32_t *dl = reinterpret_cast<int32_t *>(stg.host + pr.cap * kEdgeBytes);
        for (size_t q = 0; q < pr.list.size(); ++q) dl[q] = pr.list[q] + static_cast<int32_t>(lo);
        // [compact records][work list][chain steps]
        const size_t list_bytes = ((pr.list.size() * sizeof(int32_t)) + 15) & ~static_cast<size_t>(15);
        pr.img_bytes = place_chain_steps(ep->cs, stg.host, stg.dev, pr.cap, pr.cb, pr.cl, list_bytes);
        pr.window = w;
        return;
    }
    pr.cap = (static_cast<size_t>(m) + 7) & ~static_cast<size_t>(7);
    pr.failed = fill_soa<int32_t>(h, stg.host, pr.cap, ep->src + lo, ep->dst + lo, ep->neg ? ep->neg + lo : nullptr, ep->ts + lo,
                                  ep->eidx + lo, m, &ep->cs, &pr.cb, scratch) != 0;
    if (pr.failed) pr.err = zebra_last_error();
    else pr.img_bytes = place_chain_steps(ep->cs, stg.host, stg.dev, pr.cap, pr.cb, pr.cl);
    pr.window = w;
}

// Worker `tid` of `n_workers`: prepares windows tid, tid + n_workers, ... each as soon as its staging slot is
// free, i.e. once the window that used the slot before has been handed to the caller (its kernel is done).
static void epoch_worker(zebra_epoch *ep, int tid, int n_workers)
{
    HostScratch *sc = ep->scratch[static_cast<size_t>(tid)].get();
    for (int64_t w = tid; w < ep->n_windows; w += n_workers) {
        {
            std::unique_lock<std::mutex> lk(ep->mu);
            ep->cv.wait(lk, [&] { return ep->stop || w < ep->consumed + zebra_tppr::kPrepSlots; });
            if (ep->stop) return;
        }
        epoch_prepare(ep, w, sc);
        {
            std::lock_guard<std::mutex> lk(ep->mu);
            ep->prep[w % zebra_tppr::kPrepSlots].ready = true;
        }
        ep->cv.notify_all();
    }
}

static int epoch_issue(zebra_epoch *ep)
{
    zebra_tppr *h = ep->h;
    const int64_t w = ep->issued;
    zebra_tppr::EpochSlot &sl = h->eslot[w % 3];
    zebra_epoch::Prep &pr = ep->prep[w % zebra_tppr::kPrepSlots];
    Staging &stg = h->prep[w % zebra_tppr::kPrepSlots];
    const int64_t b0 = w * ep->window, b1 = std::min(ep->n_batches, b0 + ep->window);
    const int64_t lo = b0 * ep->batch_size, hi = std::min(ep->n, b1 * ep->batch_size);
    const int64_t m = hi - lo;
    // the slot's previous window (w-3) has been handed to the caller and released, so its copy is done
    if (sl.copy_pending) { CU(cudaEventSynchronize(sl.ev_copy)); sl.copy_pending = false; }
    const bool sharded = h->world > 1;
    const size_t ob = static_cast<size_t>(ep->window) * ep->slot_bytes;
    if (sl.out_cap < ob) {
        if (sl.out_dev) CU(cudaFree(sl.out_dev));
        if (sl.out_host) CU(cudaFreeHost(sl.out_host));
        sl.out_dev = nullptr; sl.out_host = nullptr; sl.out_cap = 0;
        if (ob) {
            if (!sharded) CU(cudaMalloc(&sl.out_dev, ob));      // sharded: rows land in the shared output region
            CU(cudaMallocHost(&sl.out_host, ob));
        }
        sl.out_cap = ob;
    }
    if (!sl.ev_kernel) {
        CU(cudaEventCreateWithFlags(&sl.ev_kernel, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sl.ev_copy, cudaEventDisableTiming));
    }
    // the staging image of the window: prepared by a worker thread, or right here
    if (ep->workers.empty()) {
        epoch_prepare(ep, w, nullptr);
    } else {
        std::unique_lock<std::mutex> lk(ep->mu);
        ep->cv.wait(lk, [&] { return pr.ready && pr.window == w; });
        pr.ready = false;
    }
    if (pr.failed) return fail("%s", pr.err.c_str());
    const size_t cap = pr.cap;
    const ChainLaunch &cl = pr.cl;
    uint32_t seq0 = 0;
    if (alloc_seq(h, m, ep->st, &seq0)) return 1;
    CU(cudaMemcpyAsync(stg.dev, stg.host, pr.img_bytes, cudaMemcpyHostToDevice, ep->st));
    if (w == 0) CU(cudaEventRecord(h->ev0, ep->st));
    BatchOut bo;
    bo.batch_size = ep->batch_size; bo.first_batch = b0; bo.out_slots = ep->window;
    bo.slot_bytes = static_cast<int64_t>(ep->slot_bytes);
    if (sharded) {
        // rows of window w go to half w & 1 of every consumer's output region; the consumer copies its batches
        // out once EVERY rank has finished the launch, and tells the others when its half may be stored into again
        const int64_t launch = h->shard_launches;
        const size_t half_off = static_cast<size_t>(w & 1) * ob;
        const int32_t *dev_list = reinterpret_cast<const int32_t *>(stg.dev + cap * kEdgeBytes);
        if (launch_kernel(h, stg.dev, cap, 0, m, ep->n, lo, ep->rows_per_edge, (1 << h->n_tppr) - 1,
                          reinterpret_cast<char *>(half_off), bo, seq0, 0, ep->st, dev_list, static_cast<int64_t>(pr.list.size()), &cl)) return 1;
        CU(cudaEventRecord(sl.ev_kernel, ep->st));
        if (ep->rows_per_edge) {
            CU(cudaStreamWaitEvent(h->copy_stream, sl.ev_kernel, 0));
            zebra::shard_wait_kernel<<<1, 32, 0, h->copy_stream>>>(reinterpret_cast<const int32_t *>(h->arena), h->world,
                                                                  static_cast<int32_t>(launch + 1), h->status, h->wait_ns);
            const char *region = h->arena + kArenaHeader + 2 * h->mbox_half + half_off;
            const int64_t nb = b1 - b0;
            if (h->row_mode == 1) {
                // this rank's batches of the window: b = rank (mod world); windows start at multiples of world
                const int64_t mine = nb > h->rank ? (nb - h->rank + h->world - 1) / h->world : 0;
                if (mine > 0)
                    CU(cudaMemcpy2DAsync(sl.out_host, ep->slot_bytes, region + static_cast<size_t>(h->rank) * ep->slot_bytes,
                                         static_cast<size_t>(h->world) * ep->slot_bytes, ep->slot_bytes, static_cast<size_t>(mine),
                                         cudaMemcpyDeviceToHost, h->copy_stream));
            } else {
                CU(cudaMemcpyAsync(sl.out_host, region, static_cast<size_t>(nb) * ep->slot_bytes, cudaMemcpyDeviceToHost, h->copy_stream));
            }
            zebra::ShardDoneParams dp;
            for (int r = 0; r < zebra::kMaxRanks; ++r) dp.hdr[r] = r < h->world ? reinterpret_cast<int32_t *>(h->peer_arena[r]) : nullptr;
            dp.world = h->world; dp.rank = h->rank; dp.value = static_cast<int32_t>(launch + 1); dp.slot = 1;
            zebra::shard_done_kernel<<<1, 32, 0, h->copy_stream>>>(dp);
            CU(cudaGetLastError());
            CU(cudaEventRecord(sl.ev_copy, h->copy_stream));
        } else {
            CU(cudaEventRecord(sl.ev_copy, ep->st));
        }
        sl.copy_pending = true;
        if (w + 1 == ep->n_windows) { CU(cudaEventRecord(h->ev1, ep->st)); h->timed = true; }
        ep->issued = w + 1;
        return 0;
    }
    if (launch_kernel(h, stg.dev, cap, 0, m, ep->n, lo, ep->rows_per_edge, (1 << h->n_tppr) - 1, sl.out_dev, bo, seq0, 0,
                      ep->st, nullptr, 0, &cl)) return 1;
    CU(cudaEventRecord(sl.ev_kernel, ep->st));
    if (ep->rows_per_edge) {
        CU(cudaStreamWaitEvent(h->copy_stream, sl.ev_kernel, 0));
        CU(cudaMemcpyAsync(sl.out_host, sl.out_dev, static_cast<size_t>(b1 - b0) * ep->slot_bytes, cudaMemcpyDeviceToHost,
                           h->copy_stream));
        CU(cudaEventRecord(sl.ev_copy, h->copy_stream));
        // no stream-side wait: this slot is staged again only after the caller has taken the window,
        // i.e. after the host has synchronised on ev_copy
    } else {
        CU(cudaEventRecord(sl.ev_copy, ep->st));
    }
    sl.copy_pending = true;
    if (w + 1 == ep->n_windows) { CU(cudaEventRecord(h->ev1, ep->st)); h->timed = true; }
    ep->issued = w + 1;
    return 0;
}

int zebra_tppr_epoch_begin(zebra_tppr_t *h, const int32_t *host_src, const int32_t *host_dst, const int32_t *host_neg,
                           const double *host_ts, const int64_t *host_eidx, int64_t n, int32_t rows_per_edge,
                           int64_t batch_size, int64_t batches_per_window, void *stream, zebra_epoch_t **out)
{
    if (!h || !out) return fail("null argument");
    *out = nullptr;
    if (n <= 0 || batch_size <= 0 || batches_per_window <= 0) return fail("epoch needs n, batch_size, batches_per_window > 0");
    if (rows_per_edge != 0 && rows_per_edge != 2 && rows_per_edge != 3) return fail("rows_per_edge must be 0, 2 or 3");
    if (!host_src || !host_dst || !host_ts || !host_eidx || (rows_per_edge == 3 && !host_neg)) return fail("null input");
    if (batch_size * batches_per_window > (1 << 22)) return fail("window too long (max 2^22 edges)");
    if (n >= 0x70000000LL) return fail("epoch too long");
    for (int j = 0; j < h->n_tppr; ++j)
        if (!h->tables[j]) return fail("zebra_tppr_bind_tables has not been called");
    if (h->plan_active || h->epoch_active) return fail("another plan or epoch is in progress on this handle");
    if (h->world > 1) {
        if (rows_per_edge && (batches_per_window % h->world) != 0 && h->row_mode == 1)
            return fail("node-sharded epoch: batches_per_window must be a multiple of the world size (batch b is consumed by rank b mod world)");
        if (batch_size * batches_per_window > h->max_window_edges) return fail("window exceeds the mailbox window given to zebra_tppr_shard_init");
        if (rows_per_edge && 2 * static_cast<size_t>(batches_per_window) * zebra_tppr_out_bytes(h, batch_size, rows_per_edge) > h->out_region)
            return fail("node-sharded epoch needs an output region of two windows of rows (zebra_tppr_shard_init out_bytes)");
        for (int r = 0; r < h->world; ++r)
            if (!h->peer_arena[r]) return fail("zebra_tppr_shard_connect has not been called");
    }
    CU(cudaSetDevice(h->device));
    if (!h->copy_stream) CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    zebra_epoch *ep = new zebra_epoch();
    ep->h = h; ep->src = host_src; ep->dst = host_dst; ep->neg = rows_per_edge == 3 ? host_neg : nullptr;
    ep->ts = host_ts; ep->eidx = host_eidx; ep->n = n; ep->batch_size = batch_size; ep->rows_per_edge = rows_per_edge;
    ep->n_batches = (n + batch_size - 1) / batch_size;
    ep->window = std::min(batches_per_window, ep->n_batches);
    ep->n_windows = (ep->n_batches + ep->window - 1) / ep->window;
    ep->slot_bytes = zebra_tppr_out_bytes(h, batch_size, rows_per_edge);
    ep->st = static_cast<cudaStream_t>(stream);
    if (h->world > 1 && rows_per_edge) {
        // everything this rank received before this epoch has been consumed: both halves may be stored into
        zebra::ShardDoneParams dp;
        for (int r = 0; r < zebra::kMaxRanks; ++r) dp.hdr[r] = r < h->world ? reinterpret_cast<int32_t *>(h->peer_arena[r]) : nullptr;
        dp.world = h->world; dp.rank = h->rank; dp.value = static_cast<int32_t>(h->shard_launches); dp.slot = 1;
        zebra::shard_done_kernel<<<1, 32, 0, ep->st>>>(dp);
        CU(cudaGetLastError());
    }
    detect_chains<int32_t>(h, host_src, host_dst, n, ep->cs);
    h->chain_items = 0; h->chain_steps = 0;
    CU(cudaMemsetAsync(h->status + 1, 0, 3 * sizeof(int32_t), ep->st));
    h->last_items = 0;
    if (begin_span(h, ep->st)) { delete ep; return 1; }
    h->epoch_active = ep;
    for (auto &ps : h->prep)
        if (ensure_staging(h, ps, static_cast<size_t>(ep->window * ep->batch_size) + 8)) { h->epoch_active = nullptr; delete ep; return 1; }
    // worker threads for the host dependency pass (the GPU needs ~45 ns per edge of a hub-bound stream, one
    // host thread ~55 ns): windows are independent of each other, so they are prepared in parallel, ahead
    int n_workers = h->host_threads;
    if (n_workers < 0) {
        // share the host cores with the other ranks of the box (torchrun sets LOCAL_WORLD_SIZE)
        unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        if (const char *e = getenv("LOCAL_WORLD_SIZE")) hw = std::max(1u, hw / static_cast<unsigned>(std::max(1, atoi(e))));
        n_workers = static_cast<int>(std::min<unsigned>(3, hw - 1));
    }
    if (ep->n_windows < 4) n_workers = 0;
    n_workers = std::min<int>(n_workers, zebra_tppr::kPrepSlots - 2);
    for (int t = 0; t < n_workers; ++t) {
        ep->scratch.emplace_back(new HostScratch());
        ep->scratch.back()->dep.assign(static_cast<size_t>(h->num_nodes), NodeDep{-1, -1, 0u, 0u});
    }
    for (int t = 0; t < n_workers; ++t) ep->workers.emplace_back(epoch_worker, ep, t, n_workers);
    // two windows in flight before the caller asks for the first
    for (int q = 0; q < 2 && ep->issued < ep->n_windows; ++q)
        if (epoch_issue(ep)) { zebra_tppr_epoch_end(ep); return 1; }
    *out = ep;
    return 0;
}

int zebra_tppr_epoch_next(zebra_epoch_t *ep, void **host_rows, int64_t *first_batch, int64_t *n_batches)
{
    if (!ep || !host_rows || !first_batch || !n_batches) return fail("null argument");
    zebra_tppr *h = ep->h;
    *host_rows = nullptr; *first_batch = 0; *n_batches = 0;
    if (ep->consumed >= ep->n_windows) return 0;       // end of the epoch: n_batches = 0
    CU(cudaSetDevice(h->device));
    // keep the device busy: stage and launch the window after the next before blocking
    if (ep->issued < ep->n_windows && ep->issued < ep->consumed + 3) {
        // slot (issued % 3) was handed out at call consumed-1 and is released by this call
        if (epoch_issue(ep)) return 1;
    }
    const int64_t w = ep->consumed;
    zebra_tppr::EpochSlot &sl = h->eslot[w % 3];
    CU(cudaEventSynchronize(sl.ev_copy));
    sl.copy_pending = false;
    const int64_t b0 = w * ep->window;
    *first_batch = b0;
    *n_batches = std::min(ep->n_batches, b0 + ep->window) - b0;
    *host_rows = ep->rows_per_edge ? sl.out_host : nullptr;
    {
        std::lock_guard<std::mutex> lk(ep->mu);
        ep->consumed = w + 1;
    }
    ep->cv.notify_all();
    if (ep->consumed == ep->n_windows) {
        CU(cudaMemcpyAsync(h->status_host, h->status, kStatusWords * sizeof(int32_t), cudaMemcpyDeviceToHost, ep->st));
        CU(cudaStreamSynchronize(ep->st));
        h->last_ties = h->status_host[1];
        h->last_merge = h->status_host[2];
        if (h->status_host[0] != 0) {
            h->timeouts++;
            return fail_timeout(h);
        }
        h->rcnt_dirty = false;
    }
    return 0;
}

int zebra_tppr_epoch_end(zebra_epoch_t *ep)
{
    if (!ep) return 0;
    zebra_tppr *h = ep->h;
    {
        std::lock_guard<std::mutex> lk(ep->mu);
        ep->stop = true;
    }
    ep->cv.notify_all();
    for (auto &t : ep->workers) t.join();
    cudaSetDevice(h->device);
    cudaStreamSynchronize(ep->st);
    if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    for (auto &sl : h->eslot) sl.copy_pending = false;
    if (h->epoch_active == ep) h->epoch_active = nullptr;
    delete ep;
    return 0;
}

int64_t zebra_tppr_epoch_slot_bytes(const zebra_epoch_t *ep) { return ep ? static_cast<int64_t>(ep->slot_bytes) : 0; }

/* ---- node-sharded mode ---- */

int zebra_tppr_shard_init(zebra_tppr_t *h, int32_t rank, int32_t world, int64_t max_window_edges, size_t out_bytes,
                          void *ipc_handle_out)
{
    if (!h) return fail("null handle");
    if (h->own_tables || h->arena) return fail("handle is already sharded");
    if (world != 2 && world != 4 && world != 8) return fail("world must be 2, 4 or 8 (node v lives on rank v & (world-1))");
    if (rank < 0 || rank >= world) return fail("rank out of range");
    if (max_window_edges <= 0) return fail("max_window_edges must be positive");
    CU(cudaSetDevice(h->device));
    int lg = 0;
    while ((1 << lg) < world) ++lg;
    h->world = world; h->rank = rank; h->world_log2 = lg;
    h->table_rows = (h->num_nodes + world - 1) >> lg;
    for (int j = 0; j < h->n_tppr; ++j) {
        h->tables[j] = nullptr;
        CU(cudaMalloc(&h->tables[j], static_cast<size_t>(h->table_rows) * h->stride));
        CU(cudaMemset(h->tables[j], 0, static_cast<size_t>(h->table_rows) * h->stride));
    }
    h->own_tables = true;
    h->max_window_edges = max_window_edges;
    h->mbox_half = static_cast<size_t>(max_window_edges) * h->n_tppr * h->stride;
    h->out_region = (out_bytes + 255) & ~static_cast<size_t>(255);
    CU(cudaMalloc(&h->arena, kArenaHeader + 2 * h->mbox_half + h->out_region));
    CU(cudaMemset(h->arena, 0, kArenaHeader + 2 * h->mbox_half + h->out_region));
    CU(cudaDeviceSynchronize());
    h->peer_arena[rank] = h->arena;
    int per_sm = 0, sms = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, zebra::tppr_span_shard_kernel, zebra::kWarpsPerBlock * 32, 0));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
    h->shard_max_blocks = std::max(1, per_sm * sms);
    {
        // load the small kernel now: a first launch would load it lazily, and that load can wait for running
        // kernels — e.g. this rank's span kernel, which itself waits for a peer the host has not launched yet
        cudaFuncAttributes fa;
        CU(cudaFuncGetAttributes(&fa, zebra::shard_done_kernel));
        CU(cudaFuncGetAttributes(&fa, zebra::shard_wait_kernel));
    }
    if (ipc_handle_out) {
        cudaIpcMemHandle_t mh;
        CU(cudaIpcGetMemHandle(&mh, h->arena));
        static_assert(sizeof(mh) == 64, "cudaIpcMemHandle_t is 64 bytes");
        memcpy(ipc_handle_out, &mh, sizeof mh);
    }
    return 0;
}

int zebra_tppr_shard_connect(zebra_tppr_t *h, const void *ipc_handles)
{
    if (!h || !ipc_handles) return fail("null argument");
    if (!h->arena) return fail("zebra_tppr_shard_init has not been called");
    CU(cudaSetDevice(h->device));
    for (int r = 0; r < h->world; ++r) {
        if (r == h->rank || h->peer_arena[r]) continue;
        cudaIpcMemHandle_t mh;
        memcpy(&mh, static_cast<const char *>(ipc_handles) + 64 * r, sizeof mh);
        void *ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, mh, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return fail("cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
        h->peer_arena[r] = static_cast<char *>(ptr);
        h->peer_opened[r] = true;
    }
    return 0;
}

int zebra_tppr_shard_connect_local(zebra_tppr_t *h, zebra_tppr_t *const *ranks)
{
    if (!h || !ranks) return fail("null argument");
    if (!h->arena) return fail("zebra_tppr_shard_init has not been called");
    CU(cudaSetDevice(h->device));
    for (int r = 0; r < h->world; ++r) {
        const zebra_tppr *o = ranks[r];
        if (!o || !o->arena || o->world != h->world || o->rank != r) return fail("rank %d: not a sharded handle of this world", r);
        if (o->mbox_half != h->mbox_half || o->out_region != h->out_region) return fail("rank %d: arena geometry differs", r);
        if (o->device != h->device) {
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, h->device, o->device));
            if (!can) return fail("device %d cannot access device %d", h->device, o->device);
            cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail("enable peer access: %s", cudaGetErrorString(e));
            cudaGetLastError();
        }
        h->peer_arena[r] = o->arena;
    }
    return 0;
}

int zebra_tppr_shard_disconnect(zebra_tppr_t *h)
{
    if (!h) return 0;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < zebra::kMaxRanks; ++r) {
        if (r == h->rank) continue;
        if (h->peer_opened[r] && h->peer_arena[r]) cudaIpcCloseMemHandle(h->peer_arena[r]);
        h->peer_arena[r] = nullptr;
        h->peer_opened[r] = false;
    }
    return 0;
}

void *zebra_tppr_shard_out(zebra_tppr_t *h) { return (h && h->arena) ? h->arena + kArenaHeader + 2 * h->mbox_half : nullptr; }
size_t zebra_tppr_shard_out_bytes(const zebra_tppr_t *h) { return h ? h->out_region : 0; }

int zebra_tppr_shard_reset(zebra_tppr_t *h, void *stream)
{
    if (!h || !h->own_tables) return fail("not a sharded handle");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CU(cudaSetDevice(h->device));
    for (int j = 0; j < h->n_tppr; ++j)
        CU(cudaMemsetAsync(h->tables[j], 0, static_cast<size_t>(h->table_rows) * h->stride, st));
    CU(cudaMemsetAsync(h->rcnt, 0, sizeof(int32_t) * h->num_nodes * h->n_tppr, st));
    CU(cudaMemsetAsync(h->status, 0, kStatusWords * sizeof(int32_t), st));   // also clears a recorded wait time-out
    h->plan_active = nullptr;
    h->rcnt_dirty = false;
    return 0;
}

int zebra_tppr_shard_row_mode(zebra_tppr_t *h, int32_t mode)
{
    if (!h || !h->own_tables) return fail("not a sharded handle");
    if (mode != 0 && mode != 1) return fail("row mode must be 0 (replicate) or 1 (batch b to rank b & (world-1))");
    h->row_mode = mode;
    return 0;
}

int zebra_tppr_set_max_blocks(zebra_tppr_t *h, int32_t blocks)
{
    if (!h || blocks < 1) return fail("bad argument");
    h->max_blocks = std::min(h->max_blocks, static_cast<int>(blocks));
    if (h->shard_max_blocks) h->shard_max_blocks = std::min(h->shard_max_blocks, static_cast<int>(blocks));
    return 0;
}

int zebra_tppr_set_tie_mode(zebra_tppr_t *h, int32_t mode)
{
    if (!h || (mode != 0 && mode != 1)) return fail("tie mode must be 0 (exact: numba's argsort order) or 1 (stable: insertion order)");
    h->tie_mode = mode;
    return 0;
}

int zebra_tppr_sync_check(zebra_tppr_t *h, void *stream)
{
    if (!h) return fail("null handle");
    return check_status(h, static_cast<cudaStream_t>(stream));
}

int zebra_tppr_counters(zebra_tppr_t *h, int64_t *out8)
{
    if (!h || !out8) return fail("null argument");
    for (int q = 0; q < 8; ++q) out8[q] = 0;
    out8[0] = h->last_items; out8[1] = h->last_ties; out8[2] = h->launches; out8[3] = h->timeouts;
    out8[4] = h->last_merge; out8[5] = h->chain_items; out8[6] = h->chain_steps;
    return 0;
}

int zebra_tppr_profile(zebra_tppr_t *h, uint64_t *out16)
{
    if (!h || !out16) return fail("null argument");
    CU(cudaSetDevice(h->device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(out16, h->prof, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    CU(cudaMemset(h->prof, 0, 16 * sizeof(unsigned long long)));
    return 0;
}

double zebra_tppr_last_kernel_ms(zebra_tppr_t *h)
{
    if (!h || !h->timed) return -1.0;
    if (cudaEventSynchronize(h->ev1) != cudaSuccess) return -1.0;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) return -1.0;
    return ms;
}

}  // extern "C"
