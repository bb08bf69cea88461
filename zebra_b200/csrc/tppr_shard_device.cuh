// tppr_shard_device.cuh — node-sharded variant of the span kernel: one process per GPU, T-PPR state
// partitioned by node id, cross-shard endpoint vectors exchanged INSIDE the kernel through
// peer-mapped mailboxes over NVLink (no collective call between waves, no host round trip).
//
// Ownership.  Node v lives on rank v & (world-1) at local row v >> log2(world).  A record is read
// and written only by its owner's warps, so the single-GPU hand-off protocol (tagged records,
// tppr_device.cuh) is unchanged and never crosses NVLink.
//
// Exchange.  Every rank walks the same replicated edge stream.  For edge i = (s, d, neg) an
// endpoint's update needs the OLD vector of the other endpoint (utils/util.py:533-538).  If the
// two endpoints have different owners, each owner, as soon as it holds the right version of its
// endpoint, PUSHES that old vector from registers into the partner's mailbox slot of item (i, j)
// (16-byte chunks tagged with the edge's sequence number, the same self-validating format as a
// record) and then polls its OWN mailbox slot — local memory — for the partner's vector.  Nothing
// is ever pulled over NVLink, and no version can be overwritten under a remote reader.
//
// Work list.  The host gives every rank the (sorted) list of edges it takes part in — it owns src,
// dst or a distinct negative — and the tickets run over that list only.  The staging records are
// compact as well: entry q of the arrays (src, dst, ..., ts) describes edge list[q].
//
// Mailboxes are indexed by launch-local item and double-buffered by launch parity.  A rank may push into
// a peer's half of launch n only after that peer has FINISHED launch n-2 (the previous user of the half):
// every rank publishes, stream-ordered behind each of its launches, the number of launches it has finished
// into the arena header of every rank (shard_done_kernel), and a pushing warp checks the partner's count in
// its own header first (once per warp and partner).  Without it two ranks that share no edge for a launch
// or two could drift apart by more than one launch and overwrite a slot a third rank's push is still
// waited for in.
//
// Output rows (utils/util.py:447-469) are written by the owner of the node they describe, straight
// into the output region of the consuming rank(s) (peer stores): row_mode 0 replicates every row on
// every rank, row_mode 1 delivers the rows of batch b to rank b & (world-1) only — what data-parallel
// training consumes (each rank works on its own batches), with 1/world of the NVLink store traffic.
#pragma once
#include "tppr_device.cuh"

namespace zebra {

// Arena header: int32 [0..kMaxRanks) launches each rank has FINISHED, [kMaxRanks..2 kMaxRanks) launches whose
// rows each rank has COPIED out of its output region (pipelined epochs).
struct ShardDoneParams { int32_t *hdr[kMaxRanks]; int32_t world, rank, value, slot; };

// One warp, launched behind a sharded span kernel (slot 0) or behind the copy of a window's rows (slot 1) on
// the same stream: lane r stores this rank's count into rank r's arena header (release at system scope).
__global__ void shard_done_kernel(const ShardDoneParams d)
{
    const int r = threadIdx.x;
    if (r < d.world) {
        __threadfence_system();
        asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(d.hdr[r] + d.slot * kMaxRanks + d.rank), "r"(d.value) : "memory");
    }
}

// Holds a stream until every rank has finished `value` launches (their rows for this rank have landed).
__global__ void shard_wait_kernel(const int32_t *done, int32_t world, int32_t value, int32_t *status, unsigned long long wait_ns)
{
    const int r = threadIdx.x;
    if (r < world) {
        Waiter w;
        while (ld_relaxed_sys_s32(done + r) < value) {
            if (w.expired(status, wait_ns, 10, static_cast<uint32_t>(value), static_cast<uint32_t>(r))) break;
            __nanosleep(1000);
        }
        __threadfence_system();
    }
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
tppr_span_shard_kernel(const __grid_constant__ SpanParams p)
{
    __shared__ WarpScratch scratch[kWarpsPerBlock];
    const int lane = threadIdx.x & 31;
    WarpScratch &s = scratch[threadIdx.x >> 5];
    scratch_init(s, lane);
    // tickets run over this rank's own list of edges (built by the host from the ownership rule), so
    // edges that do not touch this shard cost nothing here
    const unsigned long long total = static_cast<unsigned long long>(p.n_list) * p.n_tppr;
    const int k = p.k;
    const uint32_t tag_base = p.seq0 + static_cast<uint32_t>(p.tag_offset);
    const int lg = p.world_log2, me = p.rank, pmask = (1 << lg) - 1, world = 1 << lg;
    unsigned stats = 0;
    unsigned rows_ok = 0;                   // ranks known to have drained the output half this launch stores into
    unsigned peers_ok = 0;                  // partners known to have left the mailbox half of this launch's parity
    char *const my_mbox = p.mbox_peer[me] + static_cast<size_t>(p.mbox_parity) * p.mbox_half_bytes;

    // hub chains of the nodes THIS rank owns (tppr_device.cuh "hub chains"): the first n_chains * n_tppr blocks
    // serve one (chain node, instance) each — warp 0 walks the chain, the others take the chain's edges in turn
    int hx = 0, hx_end = 0, hj = 0, hstride = kWarpsPerBlock - 1;
    if (p.n_chains > 0 && blockIdx.x < static_cast<unsigned>(p.n_chains * p.n_tppr)) {
        const int q = blockIdx.x / p.n_tppr;
        hj = blockIdx.x - q * p.n_tppr;
        if ((p.inst_mask >> hj) & 1) {
            // a chain node of another rank: no chain warp here, all eight warps help with the edges that join it to
            // this rank's nodes (their pushes are what the remote chain warp waits for)
            const bool owned = p.chains[q].owned != 0;
            const int w = threadIdx.x >> 5;
            if (owned && w == 0) run_chain<true>(p, s, q, hj, lane, stats);
            else { hx = p.chains[q].begin + (owned ? w - 1 : w); hx_end = p.chains[q].end; hstride = owned ? kWarpsPerBlock - 1 : kWarpsPerBlock; }
        }
    }

    for (;;) {
        int64_t li;
        int j;
        bool from_ticket = false;
        if (hx < hx_end) {
            const ChainStep *st = p.csteps + hx;
            hx += hstride;
            const uint32_t fl = __ldg(&st->flags);
            // an edge between two chain nodes of this rank is in both chains: the helpers of its src side take it
            if ((fl & kStepPeerChain) && (fl & kStepDst)) continue;
            li = __ldg(&st->rec) - p.rec_offset;
            j = hj;
        } else {
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
            li = tk / static_cast<uint32_t>(p.n_tppr);    // 32-bit division: items < 2^32
            j = static_cast<int>(tk - li * p.n_tppr);
            from_ticket = true;
        }
        if (!((p.inst_mask >> j) & 1)) continue;
        // the staging records of a sharded plan are compact: entry li describes edge list[li].  The whole
        // record in one batch of independent loads, before any branch that depends on it.
        const int64_t i = static_cast<int64_t>(p.list[li]) - p.i_offset;     // launch-local edge index
        const int sN = p.src[li], dN = p.dst[li];
        const int gN_raw = p.neg[li];
        const uint32_t flags = p.flags[li];
        const int32_t pS = p.prev_s[li], pD_raw = p.prev_d[li], pG_raw = p.prev_n[li];
        const double t = p.ts[li];
        const int e = p.eidx[li];
        const uint32_t rwS = p.rthr_s[li], rwD_raw = p.rthr_d[li];
        const int32_t ci = p.cidx[li];
        if (from_ticket && (flags & (kFlagSrcChain | kFlagDstChain))) continue;   // a chain block's helpers take it
        if (p.dbg_owner && lane == 0) {
            // debug: every (list entry, instance) must be processed exactly once
            const uint32_t me_id = 1u + blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5) + (from_ticket ? 0u : 0x40000000u);
            const uint32_t old = atomicExch(p.dbg_owner + li * p.n_tppr + j, me_id);
            if (old != 0) {
                atomicExch(p.status, 1);
                Waiter w;
                w.log_wait(p.status, 99, tag_base + static_cast<uint32_t>(p.list[li] - p.i_offset), old, static_cast<int32_t>(me_id));
            }
        }
        const bool has_neg = p.rows_per_edge == 3;
        const int gN = has_neg ? gN_raw : sN;
        const bool loop = sN == dN;
        const bool neg_own = has_neg && gN != sN && gN != dN;
        const bool oS = (sN & pmask) == me, oD = loop ? oS : (dN & pmask) == me;
        const bool oG = neg_own && (gN & pmask) == me;
        if (!(oS || oD || oG)) continue;          // nothing of this edge lives here
        int32_t *rcnt = p.rcnt[j];
        const int32_t pD = loop ? -1 : pD_raw, pG = neg_own ? pG_raw : -1;
        const uint32_t rwD = loop ? 0u : rwD_raw;
        char *table = p.tables[j];
        char *recS = table + static_cast<size_t>(sN >> lg) * p.stride;
        char *recD = table + static_cast<size_t>(dN >> lg) * p.stride;
        const char *recG = table + static_cast<size_t>(gN >> lg) * p.stride;
        const uint32_t my_tag = tag_base + static_cast<uint32_t>(i);

        // an OWNED endpoint that is a chain node: its old vector comes from the item's chain mailbox slot, its update
        // and the push to a remote partner are the chain warp's business
        const bool cS = oS && (flags & kFlagSrcChain) != 0, cD = !loop && oD && (flags & kFlagDstChain) != 0;
        const char *srcS = recS, *srcD = recD;
        uint32_t expS = p.seq0 + static_cast<uint32_t>(pS), expD = p.seq0 + static_cast<uint32_t>(pD);
        bool anyS = pS < 0, anyD = pD < 0;
        if (cS) { srcS = chain_slot(p, ci, j, 0); expS = my_tag; anyS = false; }
        if (cD) { srcD = chain_slot(p, ci, j, 1); expD = my_tag; anyD = false; }
        if (p.rows_per_edge == 0 && !oG && (!oS || cS) && (loop || !oD || cD)) continue;   // update-only: nothing left to do here

        const bool needSD = oS || oD;
        const bool cross = needSD && !(oS && oD);              // the other endpoint lives on another rank
        const size_t slot_off = (static_cast<size_t>(i) * p.n_tppr + j) * p.stride;
        const char *mslot = my_mbox + slot_off;                 // partner's push for this item lands here
        char *pslot = nullptr;
        if (cross) {
            const int partner = (oS ? dN : sN) & pmask;
            pslot = p.mbox_peer[partner] + static_cast<size_t>(p.mbox_parity) * p.mbox_half_bytes + slot_off;
        }

        VecR S, D, G;
        S.e = S.n = S.cnt = S.upd = 0; S.ts = S.sc = S.norm = 0.0;
        D = S; G = S;
        uint32_t curS = 0, curD = 0, curG = 0;
        bool okS = !needSD, okD = !needSD || loop, okG = !oG;
        // the only owned endpoint is a chain node: nothing here needs the remote partner's vector (its row and its
        // update belong to its owner, our update to the chain warp)
        if (cross && cS) okD = true;
        if (cross && cD) okS = true;
        bool pushed = !cross || (oS ? cS : cD), alive = true;      // a chain warp pushes its node's vector itself
        Waiter wt;
        for (;;) {
            if (!okS) okS = oS ? try_load<false>(srcS, k, lane, anyS, expS, S, curS)
                               : try_load<true>(mslot, k, lane, false, my_tag, S, curS);
            if (!okD) okD = oD ? try_load<false>(srcD, k, lane, anyD, expD, D, curD)
                               : try_load<true>(mslot, k, lane, false, my_tag, D, curD);
            if (!okG) okG = try_load<false>(recG, k, lane, pG < 0, p.seq0 + static_cast<uint32_t>(pG), G, curG);
            if (!pushed && (oS ? okS : okD)) {      // our endpoint's OLD vector is here: send it before computing
                const int partner = (oS ? dN : sN) & pmask;
                if (!((peers_ok >> partner) & 1u)) {
                    // the partner must have finished the launch that used this half before (launch - 2)
                    if (ld_relaxed_sys_s32(p.shard_done + partner) >= p.shard_launch - 1) peers_ok |= 1u << partner;
                }
                if ((peers_ok >> partner) & 1u) {
                    push_vector<true>(pslot, k, oS ? S : D, my_tag, lane);
                    pushed = true;
                }
            }
            if (okS && okD && pushed) break;
            // back off while the local producer is far away (same estimate as the single-GPU kernel)
            uint32_t dist = 0;
            if (!okS && oS && !cS) {
                const uint32_t d = curS - tag_base, pe = static_cast<uint32_t>(pS - p.tag_offset);
                dist = (d < static_cast<uint32_t>(p.n_chunk) && d <= pe) ? pe - d : pe + 1;
            }
            if (!okD && oD && !cD) {
                const uint32_t d = curD - tag_base, pe = static_cast<uint32_t>(pD - p.tag_offset);
                const uint32_t dd = (d < static_cast<uint32_t>(p.n_chunk) && d <= pe) ? pe - d : pe + 1;
                dist = dd > dist ? dd : dist;
            }
            if (dist > kNear) __nanosleep(backoff_ns(dist));
            else if ((!okS && (!oS || cS)) || (!okD && (!oD || cD)) || !pushed) __nanosleep(100);   // waiting for a peer's push, or for the peer to leave the mailbox half
            if (wt.expired(p.status, p.wait_ns, (!okS && !oS) || (!okD && !oD) ? 8 : 1, my_tag, 0)) { alive = false; break; }
        }
        if (!alive) continue;
        if (loop) D = vsel(true, S, S);
        if (!neg_own) G = vsel(gN == sN, S, D);
        bool sigG = okG;
        if (okG && oG && (flags & kFlagSignal) && lane == 0) atomicAdd(rcnt + gN, 1);

        bool ok = true;
        if (needSD) {
            const bool df = (flags & kFlagDstFirst) && !loop;
#pragma unroll 1
            for (int h = 0; h < (loop ? 1 : 2); ++h) {
                const bool isD = (h == 0) ? df : !df;
                if (!(isD ? oD : oS)) continue;
                if (isD ? cD : cS) continue;            // a chain node's side is computed by its chain warp
                const VecR A = vsel(isD, D, S), B = vsel(isD, S, D);
                ok &= update_endpoint(p, s, A, B, isD ? sN : dN, e, t, j, isD ? recD : recS, my_tag,
                                      rcnt + (isD ? dN : sN), isD ? rwD : rwS, lane, stats);
                if ((stats & 0xffffu) > 0xff00u || (stats >> 16) > 0xff00u) {
                    if (lane == 0) { atomicAdd(p.status + 2, static_cast<int>(stats & 0xffffu)); atomicAdd(p.status + 1, static_cast<int>(stats >> 16)); }
                    stats = 0;
                }
            }
        }
        if (!ok) { atomicExch(p.status, 1); continue; }

        while (!okG) {
            __nanosleep(64);
            okG = try_load<false>(recG, k, lane, pG < 0, p.seq0 + static_cast<uint32_t>(pG), G, curG);
            if (wt.expired(p.status, p.wait_ns, 4, my_tag, 0)) break;
        }
        if (!okG) continue;
        if (!sigG && oG && (flags & kFlagSignal) && lane == 0) atomicAdd(rcnt + gN, 1);

        if (p.rows_per_edge >= 2) {
            int64_t row = p.i_offset + i, nrow = p.n_total;
            size_t off = 0;
            int q_lo = 0, q_hi = world;              // destination ranks of this edge's rows
            if (p.batch_size > 0) {
                // plans and epochs are shorter than 2^31 edges (checked on the host): 32-bit divisions
                const int64_t b = static_cast<uint32_t>(row) / static_cast<uint32_t>(p.batch_size);
                if (p.row_mode == 1) { q_lo = static_cast<int>(b & pmask); q_hi = q_lo + 1; }
                row -= b * p.batch_size;
                nrow = p.n_total - b * p.batch_size;
                if (nrow > p.batch_size) nrow = p.batch_size;
                off = static_cast<size_t>(static_cast<uint32_t>(b - p.first_batch) % static_cast<uint32_t>(p.out_slots)) * p.slot_bytes;
            }
            const size_t plane = static_cast<size_t>(nrow) * p.rows_per_edge * k;
            off += static_cast<size_t>(j) * plane * 16;
            // a row is written by the owner of the node it describes, into every rank's output region
            const bool eS = oS, eD = oD;
            const bool eG = has_neg && (neg_own ? oG : (gN == sN ? oS : oD));
            for (int q = q_lo; q < q_hi; ++q) {
                if (p.row_guard && !((rows_ok >> q) & 1u)) {
                    // pipelined epochs: rank q must have copied out the rows of the launch that used this half
                    // of its output region before (launch - 2); it publishes how far it has copied
                    Waiter wr;
                    while (ld_relaxed_sys_s32(p.shard_done + kMaxRanks + q) < p.shard_launch - 1) {
                        if (wr.expired(p.status, p.wait_ns, 9, my_tag, static_cast<uint32_t>(q))) break;
                        __nanosleep(500);
                    }
                    rows_ok |= 1u << q;
                }
                char *obase = p.out_peer[q] + off;
                if (eS) emit_row(p, obase, plane, row, S, t, lane);
                if (eD) emit_row(p, obase, plane, row + nrow, D, t, lane);
                if (eG) emit_row(p, obase, plane, row + 2 * nrow, G, t, lane);
            }
        }
    }
}

}  // namespace zebra
