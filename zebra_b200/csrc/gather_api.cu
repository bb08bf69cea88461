// gather_api.cu — C ABI of the fused neighbour gather (include/zebra_tppr.h, "neighbour gather").
// Stateless: every pointer is caller-owned device memory; the call only enqueues kernels on `stream`.
#include "common.cuh"
#include <cstdlib>
#include "gather_device.cuh"
#include "../../include/zebra_tppr.h"

#include <cstdint>
#include <cstring>

using zebra::fail;

namespace {
bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
int32_t *g_status = nullptr;      // one device word per process: sticky out-of-range flag
int g_status_dev = -1;

int ensure_status()
{
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (g_status && g_status_dev == dev) return 0;
    CU(cudaMalloc(&g_status, sizeof(int32_t)));
    CU(cudaMemset(g_status, 0, sizeof(int32_t)));
    g_status_dev = dev;
    return 0;
}
}  // namespace

namespace {

int gather_fwd_impl(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_memory,
                    int64_t num_nodes, int32_t dim_mem, const float *dev_edge_feat, int64_t n_edge_rows,
                    int32_t dim_edge, const float *dev_time_w, const float *dev_time_b, int32_t dim_time,
                    const int32_t *dev_src_nodes, float *dev_x, float *dev_wn, float *dev_src_rows,
                    const float *dev_upd_rows, const uint32_t *dev_bitmap, const int32_t *dev_wordprefix, void *stream)
{
    if (!dev_topk || !dev_memory || !dev_edge_feat || !dev_time_w || !dev_time_b || !dev_x || !dev_wn)
        return fail("null device pointer");
    if (n_tppr < 1 || rows < 0 || k < 1 || dim_mem < 1 || dim_edge < 1 || dim_time < 1) return fail("bad gather shape");
    if ((dev_src_rows != nullptr) != (dev_src_nodes != nullptr)) return fail("src_nodes and src_rows go together");
    if (static_cast<double>(n_tppr) * static_cast<double>(rows) * k >= 2147483648.0) return fail("gather of 2^31 slots or more");
    if (rows == 0) return 0;
    if (ensure_status()) return 1;
    zebra::GatherParams p;
    memset(&p, 0, sizeof p);
    p.topk = static_cast<const char *>(dev_topk); p.n_tppr = n_tppr; p.k = k; p.R = rows;
    p.memory = dev_memory; p.edge_feat = dev_edge_feat; p.time_w = dev_time_w; p.time_b = dev_time_b;
    p.num_nodes = num_nodes; p.n_edge_rows = n_edge_rows; p.Dm = dim_mem; p.De = dim_edge; p.Dt = dim_time;
    p.src_nodes = dev_src_nodes; p.X = dev_x; p.wn = dev_wn; p.src_rows = dev_src_rows; p.status = g_status;
    p.upd_rows = dev_upd_rows; p.bitmap = dev_bitmap; p.wordprefix = dev_wordprefix;
    const bool vec = dim_mem % 4 == 0 && dim_edge % 4 == 0 && dim_time % 4 == 0 && aligned16(dev_memory) &&
                     aligned16(dev_edge_feat) && aligned16(dev_time_w) && aligned16(dev_time_b) && aligned16(dev_x) &&
                     (!dev_src_rows || aligned16(dev_src_rows)) && (!dev_upd_rows || aligned16(dev_upd_rows));
    const int64_t slots = static_cast<int64_t>(n_tppr) * rows * k;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    static const bool one_slot = getenv("ZEBRA_GATHER_ONE_SLOT") && atoi(getenv("ZEBRA_GATHER_ONE_SLOT")) != 0;   // developer knob
    if (!one_slot) {
        // eight slots per warp (gather_fwd8_kernel)
        const int64_t warps = (slots + zebra::kGatherGroup - 1) / zebra::kGatherGroup + (dev_src_rows ? rows : 0);
        const unsigned blocks = static_cast<unsigned>((warps + 7) / 8);
        if (vec) zebra::gather_fwd8_kernel<true><<<blocks, 256, 0, st>>>(p);
        else zebra::gather_fwd8_kernel<false><<<blocks, 256, 0, st>>>(p);
    } else {
        const int64_t warps = slots + (dev_src_rows ? rows : 0);
        const unsigned blocks = static_cast<unsigned>((warps + 7) / 8);
        if (vec) zebra::gather_fwd_kernel<true><<<blocks, 256, 0, st>>>(p);
        else zebra::gather_fwd_kernel<false><<<blocks, 256, 0, st>>>(p);
    }
    CU(cudaGetLastError());
    return 0;
}

int gather_bwd_sorted_impl(int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx, const float *dev_wn,
                           const float *dev_gsrc, const int32_t *dev_sorted_nodes, const int64_t *dev_order,
                           const uint32_t *dev_bitmap, const int32_t *dev_wordprefix, float *dev_g,
                           int64_t num_nodes, int32_t dim_mem, int32_t dim_total, void *stream)
{
    if (!dev_gx || !dev_wn || !dev_g || !dev_sorted_nodes || !dev_order) return fail("null device pointer");
    if (dim_mem < 1 || dim_mem > 4096) return fail("bad dim_mem");
    if (static_cast<double>(n_tppr) * static_cast<double>(rows) * k >= 2147483648.0) return fail("gather of 2^31 slots or more");
    if (rows == 0) return 0;
    zebra::GatherBwdParams p;
    memset(&p, 0, sizeof p);
    p.n_tppr = n_tppr; p.k = k; p.R = rows;
    p.gX = dev_gx; p.wn = dev_wn; p.g_src = dev_gsrc; p.g_memory = dev_g;
    p.num_nodes = num_nodes; p.Dm = dim_mem; p.D = dim_total;
    p.bitmap = dev_bitmap; p.wordprefix = dev_wordprefix;
    const int64_t total = static_cast<int64_t>(n_tppr) * rows * k + (dev_gsrc ? rows : 0);
    const int chunks = (dim_mem + 31) / 32;
    const unsigned blocks = static_cast<unsigned>((total * chunks + 7) / 8);
    zebra::gather_bwd_sorted_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(p, dev_sorted_nodes, dev_order, total, chunks);
    CU(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" {

int zebra_gather_fwd(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_memory,
                     int64_t num_nodes, int32_t dim_mem, const float *dev_edge_feat, int64_t n_edge_rows,
                     int32_t dim_edge, const float *dev_time_w, const float *dev_time_b, int32_t dim_time,
                     const int32_t *dev_src_nodes, float *dev_x, float *dev_wn, float *dev_src_rows, void *stream)
{
    return gather_fwd_impl(dev_topk, n_tppr, rows, k, dev_memory, num_nodes, dim_mem, dev_edge_feat, n_edge_rows, dim_edge,
                           dev_time_w, dev_time_b, dim_time, dev_src_nodes, dev_x, dev_wn, dev_src_rows, nullptr, nullptr,
                           nullptr, stream);
}

int zebra_raw_messages(const int32_t *dev_u, const int32_t *dev_v, const int32_t *dev_e, const float *dev_t, int64_t n_rows,
                       const float *dev_memory, int64_t num_nodes, int32_t dim_mem, const float *dev_last_update,
                       const float *dev_edge_feat, int64_t n_edge_rows, int32_t dim_edge, const float *dev_time_w,
                       const float *dev_time_b, int32_t dim_time, float *dev_messages, void *stream)
{
    if (n_rows <= 0) return 0;
    if (!dev_u || !dev_v || !dev_e || !dev_t || !dev_memory || !dev_last_update || !dev_edge_feat || !dev_time_w || !dev_time_b ||
        !dev_messages)
        return fail("null device pointer");
    if (ensure_status()) return 1;
    zebra::RawMsgParams p;
    memset(&p, 0, sizeof p);
    p.u = dev_u; p.v = dev_v; p.e = dev_e; p.t = dev_t; p.U = n_rows; p.num_nodes = num_nodes; p.n_edge_rows = n_edge_rows;
    p.memory = dev_memory; p.last_update = dev_last_update; p.edge_feat = dev_edge_feat; p.time_w = dev_time_w; p.time_b = dev_time_b;
    p.Dm = dim_mem; p.De = dim_edge; p.Dt = dim_time; p.msg = dev_messages; p.status = g_status;
    const int D = 2 * dim_mem + dim_edge + dim_time;
    const bool vec = dim_mem % 4 == 0 && dim_edge % 4 == 0 && D % 4 == 0 && aligned16(dev_memory) && aligned16(dev_edge_feat) &&
                     aligned16(dev_messages);
    const unsigned blocks = static_cast<unsigned>((n_rows + 7) / 8);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec) zebra::raw_message_kernel<true><<<blocks, 256, 0, st>>>(p);
    else zebra::raw_message_kernel<false><<<blocks, 256, 0, st>>>(p);
    CU(cudaGetLastError());
    return 0;
}

int zebra_gather_bwd(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx,
                     const float *dev_wn, const float *dev_gsrc, const int32_t *dev_src_nodes, float *dev_gmemory,
                     int64_t num_nodes, int32_t dim_mem, int32_t dim_total, void *stream)
{
    if (!dev_topk || !dev_gx || !dev_wn || !dev_gmemory) return fail("null device pointer");
    if ((dev_gsrc != nullptr) != (dev_src_nodes != nullptr)) return fail("src_nodes and g_src go together");
    if (static_cast<double>(n_tppr) * static_cast<double>(rows) * k >= 2147483648.0) return fail("gather of 2^31 slots or more");
    if (rows == 0) return 0;
    zebra::GatherBwdParams p;
    memset(&p, 0, sizeof p);
    p.topk = static_cast<const char *>(dev_topk); p.n_tppr = n_tppr; p.k = k; p.R = rows;
    p.gX = dev_gx; p.wn = dev_wn; p.g_src = dev_gsrc; p.src_nodes = dev_src_nodes; p.g_memory = dev_gmemory;
    p.num_nodes = num_nodes; p.Dm = dim_mem; p.D = dim_total;
    const bool vec = dim_mem % 4 == 0 && dim_total % 4 == 0 && aligned16(dev_gx) && aligned16(dev_gmemory) &&
                     (!dev_gsrc || aligned16(dev_gsrc));
    const int64_t warps = static_cast<int64_t>(n_tppr) * rows * k + (dev_gsrc ? rows : 0);
    const unsigned blocks = static_cast<unsigned>((warps + 7) / 8);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec) zebra::gather_bwd_kernel<true><<<blocks, 256, 0, st>>>(p);
    else zebra::gather_bwd_kernel<false><<<blocks, 256, 0, st>>>(p);
    CU(cudaGetLastError());
    return 0;
}

int zebra_gather_bwd_sorted(int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx, const float *dev_wn,
                            const float *dev_gsrc, const int32_t *dev_sorted_nodes, const int64_t *dev_order,
                            float *dev_gmemory, int64_t num_nodes, int32_t dim_mem, int32_t dim_total, void *stream)
{
    return gather_bwd_sorted_impl(n_tppr, rows, k, dev_gx, dev_wn, dev_gsrc, dev_sorted_nodes, dev_order, nullptr, nullptr,
                                  dev_gmemory, num_nodes, dim_mem, dim_total, stream);
}

int zebra_pending_select(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const uint8_t *dev_flags,
                         int64_t num_nodes, uint32_t *dev_bitmap, int32_t *dev_wordprefix, int64_t *dev_ids,
                         int32_t *dev_count, void *stream)
{
    if (!dev_topk || !dev_flags || !dev_bitmap || !dev_wordprefix || !dev_ids || !dev_count) return fail("null device pointer");
    if (n_tppr < 1 || rows < 0 || k < 1 || num_nodes < 1 || num_nodes >= 2147483648LL) return fail("bad pending_select shape");
    if (static_cast<double>(n_tppr) * static_cast<double>(rows) * k >= 2147483648.0) return fail("gather of 2^31 slots or more");
    if (ensure_status()) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t n_words = (num_nodes + 31) / 32, plane = rows * k, n = n_tppr * plane;
    CU(cudaMemsetAsync(dev_bitmap, 0, static_cast<size_t>(n_words) * sizeof(uint32_t), st));
    if (n > 0) {
        zebra::pending_mark_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(
            static_cast<const char *>(dev_topk), n_tppr, plane, dev_flags, num_nodes, dev_bitmap, g_status);
        CU(cudaGetLastError());
    }
    zebra::pending_compact_kernel<<<1, 1024, 0, st>>>(dev_bitmap, n_words, dev_wordprefix, dev_ids, dev_count);
    CU(cudaGetLastError());
    return 0;
}

int zebra_gather_fwd_lazy(const void *dev_topk, int32_t n_tppr, int64_t rows, int32_t k, const float *dev_memory,
                          int64_t num_nodes, int32_t dim_mem, const float *dev_edge_feat, int64_t n_edge_rows,
                          int32_t dim_edge, const float *dev_time_w, const float *dev_time_b, int32_t dim_time,
                          const int32_t *dev_src_nodes, float *dev_x, float *dev_wn, float *dev_src_rows,
                          const float *dev_upd_rows, const uint32_t *dev_bitmap, const int32_t *dev_wordprefix,
                          void *stream)
{
    if (!dev_upd_rows || !dev_bitmap || !dev_wordprefix) return fail("null device pointer (pending set)");
    return gather_fwd_impl(dev_topk, n_tppr, rows, k, dev_memory, num_nodes, dim_mem, dev_edge_feat, n_edge_rows, dim_edge,
                           dev_time_w, dev_time_b, dim_time, dev_src_nodes, dev_x, dev_wn, dev_src_rows, dev_upd_rows,
                           dev_bitmap, dev_wordprefix, stream);
}

int zebra_gather_bwd_lazy(int32_t n_tppr, int64_t rows, int32_t k, const float *dev_gx, const float *dev_wn,
                          const float *dev_gsrc, const int32_t *dev_sorted_nodes, const int64_t *dev_order,
                          const uint32_t *dev_bitmap, const int32_t *dev_wordprefix, float *dev_gupd,
                          int64_t num_nodes, int32_t dim_mem, int32_t dim_total, void *stream)
{
    if (!dev_bitmap || !dev_wordprefix) return fail("null device pointer (pending set)");
    return gather_bwd_sorted_impl(n_tppr, rows, k, dev_gx, dev_wn, dev_gsrc, dev_sorted_nodes, dev_order, dev_bitmap,
                                  dev_wordprefix, dev_gupd, num_nodes, dim_mem, dim_total, stream);
}

int zebra_gather_check(void *stream)
{
    if (ensure_status()) return 1;
    int32_t v = 0;
    CU(cudaMemcpyAsync(&v, g_status, sizeof v, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
    CU(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    if (v != 0) {
        CU(cudaMemset(g_status, 0, sizeof(int32_t)));
        return fail("neighbour gather: a node id or edge id was out of range (row skipped)");
    }
    return 0;
}

}  // extern "C"
