// prune_api.cu — C ABI of the pruning-strategy query (include/zebra_tppr.h, "pruned top-k").
// Host work: upload the time-sorted adjacency once, stage the queries, launch, copy back.
#include "common.cuh"
#include "prune_device.cuh"
#include "../../include/zebra_tppr.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>

using zebra::fail;

struct zebra_csr {
    int64_t num_nodes = 0, nnz = 0;
    int32_t device = 0;
    int64_t *indptr = nullptr;
    int32_t *adj_node = nullptr, *adj_eidx = nullptr;
    double *adj_ts = nullptr;
    char *q_dev = nullptr;          // staged queries
    char *q_host = nullptr;         // pinned
    size_t q_cap = 0;
    char *out_dev = nullptr;
    size_t out_cap = 0;
    char *scratch = nullptr;
    size_t scratch_cap = 0;
    int sms = 0;
    int smem_max = 0;               // opt-in shared memory per block of the device
    size_t smem_set = 0;            // dynamic shared memory the kernel has been allowed so far
    int64_t launches = 0;
};

extern "C" {

int zebra_csr_create(zebra_csr_t **out, const int64_t *host_indptr, const int32_t *host_adj_node,
                     const int32_t *host_adj_eidx, const double *host_adj_ts, int64_t num_nodes, int64_t nnz,
                     int32_t device)
{
    if (!out) return fail("null out pointer");
    *out = nullptr;
    if (num_nodes <= 0 || nnz < 0 || !host_indptr) return fail("bad adjacency arguments");
    if (nnz > 0 && (!host_adj_node || !host_adj_eidx || !host_adj_ts)) return fail("null adjacency arrays");
    if (host_indptr[0] != 0 || host_indptr[num_nodes] != nnz) return fail("indptr does not span the adjacency arrays");
    for (int64_t v = 0; v < num_nodes; ++v)
        if (host_indptr[v + 1] < host_indptr[v]) return fail("indptr is not monotone at node %lld", (long long)v);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail("no CUDA device: libzebra_tppr has no CPU path (%s)", e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail("device %d out of range", device);
    CU(cudaSetDevice(device));
    zebra_csr *c = new zebra_csr();
    c->num_nodes = num_nodes; c->nnz = nnz; c->device = device;
    const size_t n1 = static_cast<size_t>(std::max<int64_t>(nnz, 1));
    CU(cudaMalloc(&c->indptr, sizeof(int64_t) * (num_nodes + 1)));
    CU(cudaMalloc(&c->adj_node, sizeof(int32_t) * n1));
    CU(cudaMalloc(&c->adj_eidx, sizeof(int32_t) * n1));
    CU(cudaMalloc(&c->adj_ts, sizeof(double) * n1));
    CU(cudaMemcpy(c->indptr, host_indptr, sizeof(int64_t) * (num_nodes + 1), cudaMemcpyHostToDevice));
    if (nnz) {
        CU(cudaMemcpy(c->adj_node, host_adj_node, sizeof(int32_t) * nnz, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(c->adj_eidx, host_adj_eidx, sizeof(int32_t) * nnz, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(c->adj_ts, host_adj_ts, sizeof(double) * nnz, cudaMemcpyHostToDevice));
    }
    CU(cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, device));
    *out = c;
    return 0;
}

int zebra_csr_destroy(zebra_csr_t *c)
{
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    cudaFree(c->indptr); cudaFree(c->adj_node); cudaFree(c->adj_eidx); cudaFree(c->adj_ts);
    if (c->q_dev) cudaFree(c->q_dev);
    if (c->q_host) cudaFreeHost(c->q_host);
    if (c->out_dev) cudaFree(c->out_dev);
    if (c->scratch) cudaFree(c->scratch);
    delete c;
    return 0;
}

size_t zebra_prune_out_bytes(int64_t nq, int32_t k) { return static_cast<size_t>(nq) * k * 16 + static_cast<size_t>(nq) * 4; }

int zebra_prune_topk(zebra_csr_t *c, const int32_t *host_q_nodes, const double *host_q_ts, int64_t nq, int32_t width,
                     int32_t depth, double alpha, double beta, int32_t k, void *dev_out, void *host_out, void *stream)
{
    if (!c) return fail("null adjacency handle");
    if (nq < 0) return fail("negative query count");
    if (k < 1 || k > 32) return fail("k=%d unsupported by the pruning kernel (1 <= k <= 32: one output slot per lane)", k);
    if (width < 1 || depth < 1) return fail("width and depth must be positive");
    if (nq == 0) return 0;
    if (!host_q_nodes || !host_q_ts) return fail("null query arrays");
    int64_t t_max = 0, lvl = 1;
    for (int d = 0; d < depth; ++d) {
        lvl *= width;
        t_max += lvl;
        if (t_max > (1 << 22)) return fail("width=%d depth=%d visits up to %lld states per query: too many", width, depth, (long long)t_max);
    }
    for (int64_t i = 0; i < nq; ++i)
        if (host_q_nodes[i] < 0 || host_q_nodes[i] >= c->num_nodes)
            return fail("query node %d out of range at %lld (num_nodes=%lld)", host_q_nodes[i], (long long)i, (long long)c->num_nodes);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CU(cudaSetDevice(c->device));
    const size_t qbytes = static_cast<size_t>(nq) * 12;
    if (c->q_cap < qbytes) {
        CU(cudaStreamSynchronize(st));
        if (c->q_dev) CU(cudaFree(c->q_dev));
        if (c->q_host) CU(cudaFreeHost(c->q_host));
        c->q_cap = std::max<size_t>(qbytes * 2, 1 << 16);
        CU(cudaMalloc(&c->q_dev, c->q_cap));
        CU(cudaMallocHost(&c->q_host, c->q_cap));
    } else {
        CU(cudaStreamSynchronize(st));      // the previous call's upload must have drained the pinned buffer
    }
    double *hq_ts = reinterpret_cast<double *>(c->q_host);
    int32_t *hq_n = reinterpret_cast<int32_t *>(hq_ts + nq);
    memcpy(hq_ts, host_q_ts, sizeof(double) * nq);
    memcpy(hq_n, host_q_nodes, sizeof(int32_t) * nq);
    CU(cudaMemcpyAsync(c->q_dev, c->q_host, qbytes, cudaMemcpyHostToDevice, st));
    const size_t ob = zebra_prune_out_bytes(nq, k);
    char *outp = static_cast<char *>(dev_out);
    if (!outp) {
        if (c->out_cap < ob) {
            CU(cudaStreamSynchronize(st));
            if (c->out_dev) CU(cudaFree(c->out_dev));
            c->out_cap = ob * 2;
            CU(cudaMalloc(&c->out_dev, c->out_cap));
        }
        outp = c->out_dev;
    }
    const size_t per_warp = zebra::prune_scratch_bytes(t_max);
    // the per-warp work areas live in shared memory when a block of 8 (4, 2, 1) warps fits its share of the SM
    // (ZEBRA_PRUNE_GLOBAL_SCRATCH=1: the round-1 layout in global memory, a developer knob)
    static const bool force_global = getenv("ZEBRA_PRUNE_GLOBAL_SCRATCH") && atoi(getenv("ZEBRA_PRUNE_GLOBAL_SCRATCH")) != 0;
    int warps_per_block = 8;
    size_t smem = 0;
    if (!force_global) {
        if (!c->smem_max) CU(cudaDeviceGetAttribute(&c->smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device));
        const size_t room = std::min<size_t>(static_cast<size_t>(c->smem_max), 200u << 10);
        for (int w = 8; w >= 1 && !smem; w >>= 1)
            if (per_warp * w <= room) { warps_per_block = w; smem = per_warp * w; }
    }
    int64_t blocks;
    if (smem) {
        if (smem > c->smem_set) {
            CU(cudaFuncSetAttribute(zebra::prune_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
            c->smem_set = smem;
        }
        const int64_t per_sm = std::max<int64_t>(1, std::min<int64_t>(8, (220u << 10) / (smem + 1024)));
        blocks = std::min<int64_t>((nq + warps_per_block - 1) / warps_per_block, static_cast<int64_t>(c->sms) * per_sm);
    } else {
        // grid: enough warps to cover the queries, at most 8 blocks of 8 warps per SM
        blocks = std::min<int64_t>((nq + warps_per_block - 1) / warps_per_block, static_cast<int64_t>(c->sms) * 8);
        // very wide/deep queries: bound the work area to 2 GB by running fewer warps
        while (blocks > 1 && per_warp * warps_per_block * static_cast<size_t>(blocks) > (2ull << 30)) blocks = (blocks + 1) / 2;
        const size_t need = per_warp * warps_per_block * static_cast<size_t>(blocks);
        if (c->scratch_cap < need) {
            CU(cudaStreamSynchronize(st));
            if (c->scratch) CU(cudaFree(c->scratch));
            CU(cudaMalloc(&c->scratch, need));
            c->scratch_cap = need;
        }
    }
    zebra::PruneParams p;
    memset(&p, 0, sizeof p);
    p.indptr = c->indptr; p.adj_node = c->adj_node; p.adj_eidx = c->adj_eidx; p.adj_ts = c->adj_ts;
    p.q_ts = reinterpret_cast<const double *>(c->q_dev);
    p.q_nodes = reinterpret_cast<const int32_t *>(p.q_ts + nq);
    p.nq = nq; p.num_nodes = c->num_nodes;
    p.width = width; p.depth = depth; p.k = k; p.t_max = static_cast<int32_t>(t_max);
    p.alpha = alpha; p.beta = beta;
    const size_t plane = static_cast<size_t>(nq) * k;
    p.o_node = reinterpret_cast<int32_t *>(outp);
    p.o_eidx = p.o_node + plane;
    p.o_dt = reinterpret_cast<float *>(p.o_eidx + plane);
    p.o_w = p.o_dt + plane;
    p.o_filled = reinterpret_cast<int32_t *>(p.o_w + plane);
    p.scratch = smem ? nullptr : c->scratch; p.scratch_per_warp = per_warp;
    zebra::prune_topk_kernel<<<static_cast<unsigned>(blocks), warps_per_block * 32, smem, st>>>(p);
    CU(cudaGetLastError());
    c->launches++;
    if (host_out) {
        CU(cudaMemcpyAsync(host_out, outp, ob, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    return 0;
}

int64_t zebra_prune_launches(const zebra_csr_t *c) { return c ? c->launches : 0; }

}  // extern "C"

// ---- zebra_prune_topk_static: the same query without any allocation ------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  prune_topk_kernel's per-warp work area and the staged
// queries live in ONE file-scope device array of ZEBRA_PRUNE_STATIC_BYTES that every GPU loading the library
// holds until exit.  Every call on a device shares it: calls that overlap on the GPU race.  Nothing in it has to
// start at zero (the kernel writes every state, dictionary slot and rank before it reads it, and the staging
// area is rewritten for [0, nq) by every call), so nothing is cleared between calls.
namespace zebra {

constexpr size_t kStaticBytes = ZEBRA_PRUNE_STATIC_BYTES;
constexpr size_t kStaticStageBytes = 1u << 20;                       // [0, 1 MiB): q_ts f64[max] | q_nodes i32[max]
constexpr int64_t kStaticMaxQueries = ZEBRA_PRUNE_STATIC_MAX_QUERIES;
constexpr size_t kStaticWorkBytes = kStaticBytes - kStaticStageBytes; // per-warp work areas
static_assert(kStaticMaxQueries * 12 <= (int64_t)kStaticStageBytes, "staging area too small");

static __device__ __align__(256) unsigned char g_prune_static[kStaticBytes];

// Copies the caller's queries into the staging area; a node id outside [0, num_nodes) becomes a query that
// finds no neighbour (node 0 at time -inf) instead of an out-of-bounds read of indptr.
__global__ void prune_static_stage_kernel(const int32_t *q_nodes, const double *q_ts, int64_t nq, int64_t num_nodes,
                                          double *s_ts, int32_t *s_nodes)
{
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nq;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int32_t v = q_nodes[i];
        const bool ok = v >= 0 && v < num_nodes;
        s_nodes[i] = ok ? v : 0;
        s_ts[i] = ok ? q_ts[i] : -__longlong_as_double(0x7ff0000000000000ll);
    }
}

// filled[i] = -1 for the queries the staging kernel neutralised
__global__ void prune_static_mark_kernel(const int32_t *q_nodes, int64_t nq, int64_t num_nodes, int32_t *o_filled)
{
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < nq;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int32_t v = q_nodes[i];
        if (v < 0 || v >= num_nodes) o_filled[i] = -1;
    }
}

constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char *g_static_ptr[kStaticMaxDevices];      // device address of g_prune_static, per device
static int g_static_sms[kStaticMaxDevices];

}  // namespace zebra

extern "C" int zebra_prune_topk_static(const int64_t *dev_indptr, const int32_t *dev_adj_node, const int32_t *dev_adj_eidx,
                                       const double *dev_adj_ts, int64_t num_nodes, const int32_t *dev_q_nodes,
                                       const double *dev_q_ts, int64_t nq, int32_t width, int32_t depth, double alpha,
                                       double beta, int32_t k, void *dev_out, void *stream)
{
    using namespace zebra;
    if (nq < 0) return fail("negative query count");
    if (num_nodes <= 0) return fail("num_nodes must be positive");
    if (k < 1 || k > 32) return fail("k=%d unsupported by the pruning kernel (1 <= k <= 32: one output slot per lane)", k);
    if (width < 1 || depth < 1) return fail("width and depth must be positive");
    if (nq == 0) return 0;
    if (!dev_indptr || !dev_adj_node || !dev_adj_eidx || !dev_adj_ts || !dev_q_nodes || !dev_q_ts || !dev_out)
        return fail("null argument");
    if (nq > kStaticMaxQueries)
        return fail("nq=%lld exceeds the %lld queries per call of the allocation-free entry point (split the batch or use "
                    "zebra_prune_topk)", (long long)nq, (long long)kStaticMaxQueries);
    int64_t t_max = 0, lvl = 1;
    for (int d = 0; d < depth; ++d) {
        lvl *= width;
        t_max += lvl;
        if (t_max > (1 << 22)) break;
    }
    const size_t per_warp = t_max > (1 << 22) ? kStaticWorkBytes + 1 : prune_scratch_bytes(t_max);
    if (per_warp > kStaticWorkBytes)
        return fail("width=%d depth=%d needs %zu bytes of work area per warp, more than the %zu-byte static work area of the "
                    "allocation-free entry point (use zebra_prune_topk)", width, depth, per_warp, kStaticWorkBytes);
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kStaticMaxDevices) return fail("device %d out of range", dev);
    unsigned char *base = nullptr;
    int sms = 0;
    {   // once per device, with that device current: the array's NAME is a host-side stand-in, its device
        // address comes from the runtime
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ptr[dev]) {
            void *ptr = nullptr;
            CU(cudaGetSymbolAddress(&ptr, g_prune_static));
            if (!ptr) return fail("cudaGetSymbolAddress returned a null address for the static work area");
            int n = 0;
            CU(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
            g_static_sms[dev] = n;
            g_static_ptr[dev] = static_cast<unsigned char *>(ptr);
        }
        base = g_static_ptr[dev];
        sms = g_static_sms[dev];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    double *s_ts = reinterpret_cast<double *>(base);
    int32_t *s_nodes = reinterpret_cast<int32_t *>(s_ts + kStaticMaxQueries);
    const unsigned aux_blocks = static_cast<unsigned>(std::min<int64_t>((nq + 255) / 256, sms));
    prune_static_stage_kernel<<<aux_blocks, 256, 0, st>>>(dev_q_nodes, dev_q_ts, nq, num_nodes, s_ts, s_nodes);
    CU(cudaGetLastError());
    // as many warps as fit the work area, at most one per query and 64 per SM
    int64_t warps = std::min<int64_t>(nq, static_cast<int64_t>(kStaticWorkBytes / per_warp));
    warps = std::min<int64_t>(warps, static_cast<int64_t>(sms) * 64);
    const int wpb = warps >= 8 ? 8 : static_cast<int>(warps);
    const int64_t blocks = warps / wpb;
    PruneParams p;
    memset(&p, 0, sizeof p);
    p.indptr = dev_indptr; p.adj_node = dev_adj_node; p.adj_eidx = dev_adj_eidx; p.adj_ts = dev_adj_ts;
    p.q_ts = s_ts; p.q_nodes = s_nodes;
    p.nq = nq; p.num_nodes = num_nodes;
    p.width = width; p.depth = depth; p.k = k; p.t_max = static_cast<int32_t>(t_max);
    p.alpha = alpha; p.beta = beta;
    const size_t plane = static_cast<size_t>(nq) * k;
    p.o_node = reinterpret_cast<int32_t *>(dev_out);
    p.o_eidx = p.o_node + plane;
    p.o_dt = reinterpret_cast<float *>(p.o_eidx + plane);
    p.o_w = p.o_dt + plane;
    p.o_filled = reinterpret_cast<int32_t *>(p.o_w + plane);
    p.scratch = reinterpret_cast<char *>(base + kStaticStageBytes); p.scratch_per_warp = per_warp;
    prune_topk_kernel<<<static_cast<unsigned>(blocks), wpb * 32, 0, st>>>(p);
    CU(cudaGetLastError());
    prune_static_mark_kernel<<<aux_blocks, 256, 0, st>>>(dev_q_nodes, nq, num_nodes, p.o_filled);
    CU(cudaGetLastError());
    return 0;
}
