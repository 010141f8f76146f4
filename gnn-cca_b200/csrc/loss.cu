// Multi-step BCE loss + per-class loss / precision statistics of the reference's compute_loss_acc (train.py:51-133) in
// two launches (one pass over the logits + a one-block finalize) and its backward seed in one, replacing ~10 eager ops
// and >= 8 synchronising .cpu() copies per batch.  SURVEY.md section 8(f) rank 1: the consumer of the MPN's logits in
// the training step (train.py:458,493) and in validation (train.py:800).
//
// Per sample and classified step s (torch.nn.BCEWithLogitsLoss, optional scalar pos_weight, main_training.py:263-268):
//   l = (1 - y) x + (1 + (pw - 1) y) (log1p(exp(-|x|)) + max(-x, 0))
// optional per-sample focal transform (libs/utils.py:264-293, reduction 'none'): b (1 - exp(-l))^gamma l.
// Sums are accumulated per thread, reduced per block in a fixed tree and folded over blocks in a fixed order in double:
// results are bit-reproducible.  HBM-bound: 4 B/edge/step read (+4 B/edge/step for the optional probabilities).
#include "common.cuh"

namespace gnncca {
namespace loss {

constexpr int MAX_STEPS = 8;
constexpr int THREADS = 256;
constexpr int PER = 5;       // accumulators per classified step

struct Ptrs { const float* logits[MAX_STEPS]; float* out[MAX_STEPS]; };

__device__ __forceinline__ float bce(float x, float y, float pw) {
  const float sp = log1pf(expf(-fabsf(x))) + fmaxf(-x, 0.f);       // softplus(-x), stable
  return (1.f - y) * x + (1.f + (pw - 1.f) * y) * sp;
}
__device__ __forceinline__ float bce_grad(float x, float y, float pw) {   // d l / d x
  const float s = 1.f / (1.f + expf(x));                            // sigmoid(-x)
  return (1.f - y) - (1.f + (pw - 1.f) * y) * s;
}

// partials[block][5k + 4]: per step {sum l, sum_{y=1} pl, sum_{y=0} pl, sum_{y=1} sigmoid, sum_{y=0} sigmoid}, then n1, n0,
// correct1, correct0 (last step)
__global__ void __launch_bounds__(THREADS) stats_kernel(const Ptrs p, int k, const float* __restrict__ labels, int64_t E, float pw, int focal,
                                                        float gamma, float balance, float* __restrict__ probs, double* __restrict__ partials) {
  float acc[PER * MAX_STEPS + 4];
#pragma unroll
  for (int i = 0; i < PER * MAX_STEPS + 4; ++i) acc[i] = 0.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < E; i += stride) {
    const float y = labels[i];
    const bool pos = y == 1.f, neg = y == 0.f;
#pragma unroll
    for (int s = 0; s < MAX_STEPS; ++s) {
      if (s < k) {
        const float x = p.logits[s][i];
        const float l = bce(x, y, pw);
        const float pl = focal ? balance * powf(1.f - expf(-l), gamma) * l : l;
        acc[PER * s] += l;
        if (pos) acc[PER * s + 1] += pl;
        if (neg) acc[PER * s + 2] += pl;
        const float pr = 1.f / (1.f + expf(-x));
        if (pos) acc[PER * s + 3] += pr;
        if (neg) acc[PER * s + 4] += pr;
        if (probs) probs[(int64_t)s * E + i] = pr;
        if (s == k - 1) {
          const float pred = pr >= 0.5f ? 1.f : 0.f;     // (sigmoid(x) >= 0.5) * 1, train.py:103-105
          if (pos) { acc[PER * MAX_STEPS] += 1.f; if (pred == y) acc[PER * MAX_STEPS + 2] += 1.f; }
          if (neg) { acc[PER * MAX_STEPS + 1] += 1.f; if (pred == y) acc[PER * MAX_STEPS + 3] += 1.f; }
        }
      }
    }
  }
  __shared__ float red[THREADS / 32][PER * MAX_STEPS + 4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < PER * MAX_STEPS + 4; ++i) {
    float v = acc[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  const int width = PER * k + 4;
  for (int i = threadIdx.x; i < width; i += THREADS) {
    const int src = i < PER * k ? i : PER * MAX_STEPS + (i - PER * k);
    double v = 0.0;
    for (int w = 0; w < THREADS / 32; ++w) v += (double)red[w][src];
    partials[(size_t)blockIdx.x * width + i] = v;
  }
}

// out[5k + 4] (double): [k mean l][k mean_{y=1} pl][k mean_{y=0} pl][k mean_{y=1} sigmoid][k mean_{y=0} sigmoid],
// then n1, n0, correct1, correct0.  Empty classes give NaN means, as torch.mean of an empty tensor does.
__global__ void finalize_kernel(const double* __restrict__ partials, int blocks, int k, int64_t E, double* __restrict__ out,
                                float* __restrict__ loss_sum) {
  const int width = PER * k + 4;
  __shared__ double sums[PER * MAX_STEPS + 4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int i = warp; i < width; i += nwarps) {       // a warp per column: lanes stride the blocks, then a fixed xor tree
    double v = 0.0;
#pragma unroll 4
    for (int b = lane; b < blocks; b += 32) v += partials[(size_t)b * width + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) sums[i] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const double n1 = sums[PER * k], n0 = sums[PER * k + 1];
    for (int s = 0; s < k; ++s) {
      out[s] = sums[PER * s] / (double)E;
      out[k + s] = sums[PER * s + 1] / n1;
      out[2 * k + s] = sums[PER * s + 2] / n0;
      out[3 * k + s] = sums[PER * s + 3] / n1;
      out[4 * k + s] = sums[PER * s + 4] / n0;
    }
    for (int i = 0; i < 4; ++i) out[PER * k + i] = sums[PER * k + i];
    if (loss_sum) {                       // the summed loss of train.py:80-97 for the non-focal criteria: sum of the step means
      double tot = 0.0;
      for (int s = 0; s < k; ++s) tot += out[s];
      *loss_sum = (float)tot;
    }
  }
}

// g[s][i] = up[s] / E * d l / d x: the seed of the MPN's backward for loss = sum_s w_s mean_i l(x_s,i) (up[s] = d loss / d mean_s)
__global__ void __launch_bounds__(THREADS) grad_kernel(const Ptrs p, int k, const float* __restrict__ labels, int64_t E, float pw,
                                                       const double* __restrict__ up, const float* __restrict__ up_sum) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float scale[MAX_STEPS];
  const double common = up_sum ? (double)*up_sum : 0.0;     // d loss / d loss_sum reaches every step mean with weight 1
#pragma unroll
  for (int s = 0; s < MAX_STEPS; ++s) scale[s] = s < k ? (float)(((up ? up[s] : 0.0) + common) / (double)E) : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < E; i += stride) {
    const float y = labels[i];
#pragma unroll
    for (int s = 0; s < MAX_STEPS; ++s)
      if (s < k) p.out[s][i] = scale[s] * bce_grad(p.logits[s][i], y, pw);
  }
}

}  // namespace loss
}  // namespace gnncca

using namespace gnncca;

extern "C" int32_t gnncca_loss_blocks(int64_t n_edges) { return (int32_t)grid_for(n_edges, loss::THREADS, 4); }

extern "C" int gnncca_bce_stats_fwd(const float* const* host_logits, int32_t n_steps, const float* labels, int64_t n_edges, float pos_weight,
                                    int32_t focal, float focus, float balance, float* probs, double* partials, double* out, float* loss_sum,
                                    void* stream) {
  if (n_steps < 1 || n_steps > loss::MAX_STEPS || n_edges < 1 || !host_logits || !labels || !partials || !out) return GNNCCA_ERR_BAD_ARG;
  loss::Ptrs p{};
  for (int s = 0; s < n_steps; ++s) { if (!host_logits[s]) return GNNCCA_ERR_BAD_ARG; p.logits[s] = host_logits[s]; }
  const int blocks = gnncca_loss_blocks(n_edges);
  cudaStream_t st = (cudaStream_t)stream;
  loss::stats_kernel<<<blocks, loss::THREADS, 0, st>>>(p, n_steps, labels, n_edges, pos_weight, focal, focus, balance, probs, partials);
  loss::finalize_kernel<<<1, 256, 0, st>>>(partials, blocks, n_steps, n_edges, out, loss_sum);
  return check_launch(2);
}

extern "C" int gnncca_bce_stats_bwd(const float* const* host_logits, float* const* host_grads, int32_t n_steps, const float* labels,
                                    int64_t n_edges, float pos_weight, const double* upstream, const float* upstream_sum, void* stream) {
  if (n_steps < 1 || n_steps > loss::MAX_STEPS || n_edges < 1 || !host_logits || !host_grads || !labels || (!upstream && !upstream_sum))
    return GNNCCA_ERR_BAD_ARG;
  loss::Ptrs p{};
  for (int s = 0; s < n_steps; ++s) {
    if (!host_logits[s] || !host_grads[s]) return GNNCCA_ERR_BAD_ARG;
    p.logits[s] = host_logits[s]; p.out[s] = host_grads[s];
  }
  loss::grad_kernel<<<gnncca_loss_blocks(n_edges), loss::THREADS, 0, (cudaStream_t)stream>>>(p, n_steps, labels, n_edges, pos_weight, upstream, upstream_sum);
  return check_launch();
}
