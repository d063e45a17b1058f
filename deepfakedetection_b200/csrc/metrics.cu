// On-device metrics tail: the integer counting work behind the reference's
// post-evaluation block, operating directly on the gathered probs / labels:
//   * threshold sweep (/root/reference/orchestration/orchestrator.py:536-546): 501 thresholds x
//     balanced_accuracy_score -- O(501 N) sklearn calls on the CPU in the reference.  Here ONE pass builds,
//     per class, a histogram of "how many thresholds is this score >= to" (binary search over the exact
//     float64 thresholds, same `score >= thr` comparison in float64), from which TP/FP for all 501
//     thresholds follow by a suffix sum on the host (evaluate.py) -- bit-identical decisions;
//   * confusion matrix (orchestrator.py:631): counts[target][pred] with integer atomics;
//   * ROC-AUC (orchestrator.py:619-628): Mann-Whitney counts  sum_pos #{neg < s} and #{neg == s}
//     against the sorted negative scores (AUC = (less + equal/2) / (P N)).
// HBM-bound integer work, a few bytes per sample: coalesced loads, shared-memory privatised histograms,
// one global atomic per (block, bin).
#include "common.cuh"
#include "kernels.h"

namespace b200df {

constexpr int kMxThreads = 256;
constexpr int kMaxThr = 1024;

__global__ void __launch_bounds__(kMxThreads)
threshold_hist_kernel(const float* __restrict__ scores, const int64_t* __restrict__ labels, int n,
                      const double* __restrict__ thr, int nthr, unsigned long long* __restrict__ hist) {
  __shared__ double sthr[kMaxThr];
  __shared__ unsigned int sh[2 * (kMaxThr + 1)];
  for (int i = threadIdx.x; i < nthr; i += kMxThreads) sthr[i] = thr[i];
  for (int i = threadIdx.x; i < 2 * (nthr + 1); i += kMxThreads) sh[i] = 0u;
  __syncthreads();
  for (int i = blockIdx.x * kMxThreads + threadIdx.x; i < n; i += gridDim.x * kMxThreads) {
    const double s = (double)scores[i];  // numpy promotes the float32 score to float64 for `>=`
    int lo = 0, hi = nthr;               // number of thresholds t_k with s >= t_k (thresholds ascending)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (s >= sthr[mid]) lo = mid + 1;
      else hi = mid;
    }
    const int cls = labels[i] == 1 ? 1 : 0;
    if (s == s) atomicAdd(&sh[cls * (nthr + 1) + lo], 1u);  // NaN >= thr is false for every thr: bin 0
    else atomicAdd(&sh[cls * (nthr + 1)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * (nthr + 1); i += kMxThreads)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

__global__ void __launch_bounds__(kMxThreads)
confusion_kernel(const int64_t* __restrict__ targets, const int64_t* __restrict__ preds, int n, int nc,
                 unsigned long long* __restrict__ cm) {
  extern __shared__ unsigned int scm[];  // [nc*nc] when it fits, else straight to global
  const bool use_s = nc * nc <= 4096;
  if (use_s) {
    for (int i = threadIdx.x; i < nc * nc; i += kMxThreads) scm[i] = 0u;
    __syncthreads();
  }
  for (int i = blockIdx.x * kMxThreads + threadIdx.x; i < n; i += gridDim.x * kMxThreads) {
    const int64_t t = targets[i], p = preds[i];
    if (t < 0 || t >= nc || p < 0 || p >= nc) continue;
    if (use_s) atomicAdd(&scm[t * nc + p], 1u);
    else atomicAdd(&cm[t * nc + p], 1ull);
  }
  if (use_s) {
    __syncthreads();
    for (int i = threadIdx.x; i < nc * nc; i += kMxThreads)
      if (scm[i]) atomicAdd(&cm[i], (unsigned long long)scm[i]);
  }
}

// out[0] += #{(p, q): neg[q] < pos[p]},  out[1] += #{(p, q): neg[q] == pos[p]};  neg sorted ascending
__global__ void __launch_bounds__(kMxThreads)
auc_count_kernel(const float* __restrict__ pos, int np, const float* __restrict__ neg_sorted, int nn,
                 unsigned long long* __restrict__ out) {
  unsigned long long less = 0, equal = 0;
  for (int i = blockIdx.x * kMxThreads + threadIdx.x; i < np; i += gridDim.x * kMxThreads) {
    const float s = pos[i];
    int lo = 0, hi = nn;  // first index with neg >= s
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (neg_sorted[mid] < s) lo = mid + 1;
      else hi = mid;
    }
    const int first_ge = lo;
    hi = nn;              // first index with neg > s
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (neg_sorted[mid] <= s) lo = mid + 1;
      else hi = mid;
    }
    less += (unsigned long long)first_ge;
    equal += (unsigned long long)(lo - first_ge);
  }
  // warp reduce, then one atomic pair per warp
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    less += __shfl_xor_sync(0xffffffffu, less, o);
    equal += __shfl_xor_sync(0xffffffffu, equal, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (less) atomicAdd(&out[0], less);
    if (equal) atomicAdd(&out[1], equal);
  }
}

static int mx_grid(int n) {
  long long b = ((long long)n + kMxThreads - 1) / kMxThreads;
  const long long cap = (long long)kNumSMs * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

cudaError_t launch_threshold_hist(const float* scores, const int64_t* labels, int n, const double* thr, int nthr,
                                  unsigned long long* hist, cudaStream_t s) {
  if (nthr < 1 || nthr > kMaxThr) return cudaErrorInvalidValue;
  if (n <= 0) return cudaSuccess;
  threshold_hist_kernel<<<mx_grid(n), kMxThreads, 0, s>>>(scores, labels, n, thr, nthr, hist);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_confusion(const int64_t* targets, const int64_t* preds, int n, int nc, unsigned long long* cm,
                             cudaStream_t s) {
  if (nc < 1) return cudaErrorInvalidValue;
  if (n <= 0) return cudaSuccess;
  const size_t smem = nc * nc <= 4096 ? (size_t)nc * nc * sizeof(unsigned int) : 0;
  confusion_kernel<<<mx_grid(n), kMxThreads, smem, s>>>(targets, preds, n, nc, cm);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_auc_count(const float* pos, int np, const float* neg_sorted, int nn, unsigned long long* out,
                             cudaStream_t s) {
  if (np <= 0 || nn <= 0) return cudaSuccess;
  auc_count_kernel<<<mx_grid(np), kMxThreads, 0, s>>>(pos, np, neg_sorted, nn, out);
  count_launch();
  return cudaGetLastError();
}

}  // namespace b200df
