// metrics_api.cu — link-prediction metrics of a batch on the device.
//
// The reference scores every batch on the host: a D2H of the 2n probabilities, then scikit-learn's
// average_precision_score, roc_auc_score and accuracy_score (evaluation/evaluation.py:34-45; the same block
// in the training loop, train.py:218-227).  Here one block per batch computes the three numbers from the
// probabilities where they are, into a device buffer that is read back once per pass.
//
// Definitions restated from scikit-learn 1.x for binary labels (first n scores positive, last n negative):
//   average precision = sum over the distinct score values v, descending, of
//                       (recall(v) - recall(previous v)) * precision(v),
//                       precision(v) = tp(v) / (tp(v) + fp(v)), recall(v) = tp(v) / P,
//                       tp(v) = #positives with score >= v, fp(v) = #negatives with score >= v
//                     = (1 / P) * sum over the positives p of tp(s_p) / (tp(s_p) + fp(s_p));
//   ROC AUC           = (#{(p, q): s_p > s_q} + 0.5 * #{(p, q): s_p == s_q}) / (P * N)
//                       (the trapezoid rule over the ROC curve at the distinct thresholds);
//   accuracy          = mean(pos_prob >= neg_prob)   (argmax over [pos, neg] with the first maximum winning,
//                       against an all-zero label vector: evaluation.py:41-42).
// Scores are compared as the float32 values they are; sums run in fp64 in a fixed order (reproducible).
#include "common.cuh"
#include "../../include/zebra_tppr.h"

#include <stdint.h>

using zebra::fail;

namespace zebra {

constexpr int kMetricThreads = 256;
constexpr int kMetricMaxBatch = 4096;      // scores of one batch held in shared memory: 2 * 4096 floats

__global__ void __launch_bounds__(kMetricThreads)
edge_metrics_kernel(const float *pos, const float *neg, int64_t n_total, int64_t batch_size, double *out)
{
    extern __shared__ float sc[];           // [n] positives, [n] negatives
    __shared__ double red[3][kMetricThreads];
    const int64_t b = blockIdx.x;
    const int64_t lo = b * batch_size;
    const int n = static_cast<int>(min(batch_size, n_total - lo));
    for (int i = threadIdx.x; i < n; i += blockDim.x) { sc[i] = pos[lo + i]; sc[n + i] = neg[lo + i]; }
    __syncthreads();
    double ap = 0.0, u = 0.0, acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float s = sc[i];
        int tp = 0, fp = 0, gt = 0, eq = 0;
        for (int q = 0; q < n; ++q) {
            const float a = sc[q], c = sc[n + q];
            tp += a >= s;
            fp += c >= s;
            gt += s > c;
            eq += s == c;
        }
        ap += static_cast<double>(tp) / static_cast<double>(tp + fp);
        u += static_cast<double>(gt) + 0.5 * static_cast<double>(eq);
        acc += s >= sc[n + i] ? 1.0 : 0.0;
    }
    red[0][threadIdx.x] = ap; red[1][threadIdx.x] = u; red[2][threadIdx.x] = acc;
    __syncthreads();
    for (int st = kMetricThreads / 2; st > 0; st >>= 1) {
        if (threadIdx.x < st)
            for (int m = 0; m < 3; ++m) red[m][threadIdx.x] += red[m][threadIdx.x + st];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double P = static_cast<double>(n);
        out[3 * b + 0] = n ? red[0][0] / P : 0.0;
        out[3 * b + 1] = n ? red[1][0] / (P * P) : 0.0;
        out[3 * b + 2] = n ? red[2][0] / P : 0.0;
    }
}

}  // namespace zebra

extern "C" int zebra_edge_metrics(const float *dev_pos, const float *dev_neg, int64_t n_total, int64_t batch_size,
                                  double *dev_out, void *stream)
{
    if (n_total <= 0) return 0;
    if (!dev_pos || !dev_neg || !dev_out) return fail("null argument");
    if (batch_size <= 0 || batch_size > zebra::kMetricMaxBatch) return fail("batch_size must be 1..%d", zebra::kMetricMaxBatch);
    const int64_t nb = (n_total + batch_size - 1) / batch_size;
    const size_t smem = 2 * static_cast<size_t>(batch_size) * sizeof(float);
    zebra::edge_metrics_kernel<<<static_cast<unsigned>(nb), zebra::kMetricThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        dev_pos, dev_neg, n_total, batch_size, dev_out);
    CU(cudaGetLastError());
    return 0;
}
