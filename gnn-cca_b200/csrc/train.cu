// Backward of the message-passing network (autograd of reference models/mpn.py:250-299 as driven by
// train.py:80-97,493: loss = sum over classified steps of BCE-with-logits, loss.backward()).
//
// Same split algebra as the forward kernels, transposed.  For step t (inputs h = h^{t-1}, e_prev = e^{t-1};
// outputs e = e^t, h^t), with P/Q/R the per-node projections of h:
//   z_m = R[src] + W_ne e         m = relu(z_m)        h^t[v] = agg_{src=v} m
//   z_e = P[src] + Q[dst] + W_ee [e0,] e_prev          e = relu(z_e)
// the gradients are
//   gz_m = dAgg(gh^t[src]) * [z_m > 0]                  (sum: gh; mean: gh/deg; max: gh where m == h^t[src])
//   gR[v] = sum_{src=v} gz_m                            ge = W_ne^T gz_m + ge_in + ge_cls
//   gz_e = ge * [e > 0]      gP[v] = sum_{src=v} gz_e    gQ[v] = sum_{dst=v} gz_e     ge_prev = W_ee^T gz_e
//   gh^{t-1} = gP W_s + gQ W_d + gR W_nx
// z_m is RECOMPUTED from the saved e^t and h^{t-1} (nothing of width 32 per edge is ever stored).
// Every reduction has a fixed order (per-thread register accumulators -> per-block partials -> one in-order
// pass), so gradients are bit-identical run to run; there are no floating-point atomics.
#include "common.cuh"

namespace gnncca {
namespace train {

constexpr int FN = 32, FE = 6, HC = 4, EP = 8;
constexpr int TB = 256;                 // threads per block of the per-edge kernels
constexpr int NW = TB / 32;

// Sum K per-thread accumulators over the block in a fixed order and store them at out[0..K).
template <int K>
__device__ __forceinline__ void block_reduce_store(float (&acc)[K], float* smem /* [NW][K] */, float* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    float v = acc[k];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) smem[warp * K + k] = v;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float v = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += smem[w * K + k];
    out[k] = v;
  }
  __syncthreads();
}

// dAgg: gradient of the aggregation wrt one message, given the gradient of the aggregated node feature
__device__ __forceinline__ float agg_grad(int agg, float gh, float z, float hval, float inv_deg) {
  if (z <= 0.f) return 0.f;                                   // ReLU gate of the node MLP (mlp.py:18)
  if (agg == GNNCCA_AGG_SUM) return gh;
  if (agg == GNNCCA_AGG_MEAN) return gh * inv_deg;
  return z == hval ? gh : 0.f;                                // max: the attaining edge (ties only at 0, where the gate is closed)
}

// ------------------------------------------------------------------------------------------------
// A. node side of the node model: warp per node, lane = feature.
//    gR[v][c] = sum_{j in seg(v)} gz_m[j][c];   per-block partial of gW_node[c][K + k] += gz_m[j][c] * e[j][k]
// ------------------------------------------------------------------------------------------------
struct NodeAggBwd {
  const int32_t* rowptr; const float* R; const float* e; const float* gh; const float* h_out;
  const float* w_node; int node_row, K;       // w_node[c * node_row + K + k]
  int64_t n_nodes; int agg;
  float* gR; float* partial;                  // partial [gridDim][FN*FE], index c * FE + k
  float* gh_eff;                              // max only: gh / (number of edges attaining the max), what gnncca_edge_bwd must be given as gh
};

__global__ void __launch_bounds__(TB) node_agg_bwd_kernel(const NodeAggBwd p) {
  __shared__ float s_red[NW][FE * FN];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float wne[FE], gw[FE];
#pragma unroll
  for (int k = 0; k < FE; ++k) { wne[k] = p.w_node[lane * p.node_row + p.K + k]; gw[k] = 0.f; }
  for (int64_t v = (int64_t)blockIdx.x * NW + warp; v < p.n_nodes; v += (int64_t)gridDim.x * NW) {
    const int j0 = p.rowptr[v], j1 = p.rowptr[v + 1];
    const float r = p.R[v * FN + lane], gh_in = p.gh[v * FN + lane];
    const float inv_deg = j1 > j0 ? 1.f / (float)(j1 - j0) : 0.f;
    float gh = gh_in, hval = 0.f;
    if (p.agg == GNNCCA_AGG_MAX) {
      // The attaining edges are found against the max of the messages as THIS pass recomputes them (one fmaf chain per message, the
      // same chain in edge_bwd_kernel), not against the forward's h_out: the forward computes the messages on the tensor cores
      // (3xTF32), whose last bits differ.  Several edges can attain the max exactly (every edge whose updated feature row is all
      // zero has z == R[v]); the gradient is shared between them (torch amax semantics; the total is what torch_scatter's argmax
      // gives too).  The recomputed max goes to the second half of gh_eff for gnncca_edge_bwd.
      int ties = 0;
      for (int j = j0; j < j1; ++j) {
        float z = r;
#pragma unroll
        for (int k = 0; k < FE; ++k) z = fmaf(p.e[(int64_t)j * FE + k], wne[k], z);
        if (z > hval) { hval = z; ties = 1; }
        else if (z > 0.f && z == hval) ++ties;
      }
      gh = ties > 0 ? gh_in / (float)ties : 0.f;
      p.gh_eff[v * FN + lane] = gh;
      p.gh_eff[(p.n_nodes + v) * FN + lane] = hval;
    }
    float acc = 0.f;
    for (int j = j0; j < j1; ++j) {
      float ev[FE];
#pragma unroll
      for (int k = 0; k < FE; ++k) ev[k] = p.e[(int64_t)j * FE + k];
      float z = r;
#pragma unroll
      for (int k = 0; k < FE; ++k) z = fmaf(ev[k], wne[k], z);
      const float g = agg_grad(p.agg, gh, z, hval, inv_deg);
      acc += g;
#pragma unroll
      for (int k = 0; k < FE; ++k) gw[k] = fmaf(g, ev[k], gw[k]);
    }
    p.gR[v * FN + lane] = acc;
  }
#pragma unroll
  for (int k = 0; k < FE; ++k) s_red[warp][lane * FE + k] = gw[k];
  __syncthreads();
  for (int i = threadIdx.x; i < FE * FN; i += TB) {
    float v = 0.f;
    for (int w = 0; w < NW; ++w) v += s_red[w][i];
    p.partial[(int64_t)blockIdx.x * (FE * FN) + i] = v;
  }
}

// ------------------------------------------------------------------------------------------------
// B. per-edge backward of node model (edge side) + edge model: thread per edge.
// ------------------------------------------------------------------------------------------------
struct EdgeBwd {
  const int32_t* rowptr; const int32_t* src;
  const float* R; const float* e; const float* e_prev; const float* e0;
  const float* gh; const float* h_out;          // gh == nullptr: the node update of this step is dead (last step)
  const float* ge_in; const float* ge_cls;      // optional [E][FE] gradients wrt e arriving from step t+1 / the classifier
  const float* w_node; const float* w_edge; int node_row, edge_row, K;
  int64_t n_edges; int agg;
  float* gze;        // [E][EP]  gradient wrt the edge-update pre-activation (consumed by the P/Q reductions)
  float* ge_prev;    // [E][FE]  gradient wrt e^{t-1}
  float* ge0_acc;    // [E][FE]  += gradient wrt the encoder's edge output through the reattached block (REATT_E)
  float* partial;    // [gridDim][FE * EIN], index o * EIN + kk  -> gW_edge[o][2K + kk]
};

template <bool REATT_E>
__global__ void __launch_bounds__(TB) edge_bwd_kernel(const EdgeBwd p) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  constexpr int OFF = REATT_E ? FE : 0;
  __shared__ float s_wne[FE][FN];
  __shared__ float s_wee[FE][EIN];
  __shared__ float s_red[NW * FE * EIN];
  for (int i = threadIdx.x; i < FE * FN; i += TB) s_wne[i / FN][i % FN] = p.w_node[(i % FN) * p.node_row + p.K + i / FN];
  for (int i = threadIdx.x; i < FE * EIN; i += TB) s_wee[i / EIN][i % EIN] = p.w_edge[(i / EIN) * p.edge_row + 2 * p.K + i % EIN];
  __syncthreads();
  float gw[FE * EIN];
#pragma unroll
  for (int i = 0; i < FE * EIN; ++i) gw[i] = 0.f;
  const int64_t stride = (int64_t)gridDim.x * TB;
  for (int64_t j = (int64_t)blockIdx.x * TB + threadIdx.x; j < p.n_edges; j += stride) {
    float ev[FE], ge[FE];
#pragma unroll
    for (int k = 0; k < FE; ++k) {
      ev[k] = p.e[j * FE + k];
      ge[k] = (p.ge_in ? p.ge_in[j * FE + k] : 0.f) + (p.ge_cls ? p.ge_cls[j * FE + k] : 0.f);
    }
    if (p.gh != nullptr) {
      const int s = p.src[j];
      const float inv_deg = 1.f / (float)max(p.rowptr[s + 1] - p.rowptr[s], 1);
      const float4* R4 = reinterpret_cast<const float4*>(p.R + (int64_t)s * FN);
      const float4* G4 = reinterpret_cast<const float4*>(p.gh + (int64_t)s * FN);
      const float4* H4 = reinterpret_cast<const float4*>(p.h_out + (int64_t)s * FN);
#pragma unroll
      for (int c4 = 0; c4 < FN / 4; ++c4) {
        const float4 r = R4[c4], g = G4[c4];
        const float4 hv = p.agg == GNNCCA_AGG_MAX ? H4[c4] : make_float4(0.f, 0.f, 0.f, 0.f);
        const float rr[4] = {r.x, r.y, r.z, r.w}, gg[4] = {g.x, g.y, g.z, g.w}, hh[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int c = c4 * 4 + u;
          float z = rr[u];
#pragma unroll
          for (int k = 0; k < FE; ++k) z = fmaf(ev[k], s_wne[k][c], z);
          const float gz = agg_grad(p.agg, gg[u], z, hh[u], inv_deg);
#pragma unroll
          for (int k = 0; k < FE; ++k) ge[k] = fmaf(gz, s_wne[k][c], ge[k]);
        }
      }
    }
    float gze[FE];
#pragma unroll
    for (int o = 0; o < FE; ++o) gze[o] = ev[o] > 0.f ? ge[o] : 0.f;          // ReLU gate of the edge MLP
    *reinterpret_cast<float4*>(p.gze + j * EP) = make_float4(gze[0], gze[1], gze[2], gze[3]);
    *reinterpret_cast<float4*>(p.gze + j * EP + 4) = make_float4(gze[4], gze[5], 0.f, 0.f);
    float ein[EIN];
    if (REATT_E) {
#pragma unroll
      for (int k = 0; k < FE; ++k) ein[k] = p.e0[j * FE + k];
    }
#pragma unroll
    for (int k = 0; k < FE; ++k) ein[OFF + k] = p.e_prev[j * FE + k];
#pragma unroll
    for (int k = 0; k < FE; ++k) {
      float a = 0.f;
#pragma unroll
      for (int o = 0; o < FE; ++o) a = fmaf(gze[o], s_wee[o][OFF + k], a);
      p.ge_prev[j * FE + k] = a;
    }
    if (REATT_E) {
#pragma unroll
      for (int k = 0; k < FE; ++k) {
        float a = p.ge0_acc[j * FE + k];
#pragma unroll
        for (int o = 0; o < FE; ++o) a = fmaf(gze[o], s_wee[o][k], a);
        p.ge0_acc[j * FE + k] = a;
      }
    }
#pragma unroll
    for (int o = 0; o < FE; ++o)
#pragma unroll
      for (int kk = 0; kk < EIN; ++kk) gw[o * EIN + kk] = fmaf(gze[o], ein[kk], gw[o * EIN + kk]);
  }
  block_reduce_store<FE * EIN>(gw, s_red, p.partial + (int64_t)blockIdx.x * (FE * EIN));
}

// ------------------------------------------------------------------------------------------------
// C. classifier (mpn.py:290-293; mlp.py:12-23: Linear(FE,HC) [BatchNorm1d] ReLU Linear(HC,1)).
//    The BatchNorm (absent / eval / train) enters as a per-feature affine  y = c * scale + shift  with
//    c_hat = (c - mean) * invstd; train mode additionally needs the batch sums S1, S2 of its backward.
// ------------------------------------------------------------------------------------------------
struct ClsParams {
  const float* e; int64_t n_edges;
  const float* w1; const float* b1; const float* w2; const float* b2;
  const float* scale; const float* shift; const float* mean; const float* invstd;   // [HC] each (device)
};

__device__ __forceinline__ void cls_load(const ClsParams& p, float (*s_w1)[FE], float* s_b1, float* s_w2, float* s_aff) {
  for (int i = threadIdx.x; i < HC * FE; i += blockDim.x) s_w1[i / FE][i % FE] = p.w1[i];
  for (int i = threadIdx.x; i < HC; i += blockDim.x) {
    s_b1[i] = p.b1[i]; s_w2[i] = p.w2[i];
    s_aff[i] = p.scale ? p.scale[i] : 1.f; s_aff[HC + i] = p.shift ? p.shift[i] : 0.f;
    s_aff[2 * HC + i] = p.mean ? p.mean[i] : 0.f; s_aff[3 * HC + i] = p.invstd ? p.invstd[i] : 1.f;
  }
  __syncthreads();
}

// per-feature sums of (c - center) and (c - center)^2 -> partial [gridDim][2*HC]
__global__ void __launch_bounds__(TB) cls_moments_kernel(const ClsParams p, const float* center, float* partial) {
  __shared__ float s_w1[HC][FE], s_b1[HC], s_w2[HC], s_aff[4 * HC], s_red[NW * 2 * HC];
  cls_load(p, s_w1, s_b1, s_w2, s_aff);
  float acc[2 * HC];
#pragma unroll
  for (int i = 0; i < 2 * HC; ++i) acc[i] = 0.f;
  float ctr[HC];
#pragma unroll
  for (int f = 0; f < HC; ++f) ctr[f] = center ? center[f] : 0.f;
  const int64_t stride = (int64_t)gridDim.x * TB;
  for (int64_t j = (int64_t)blockIdx.x * TB + threadIdx.x; j < p.n_edges; j += stride) {
    float ev[FE];
#pragma unroll
    for (int k = 0; k < FE; ++k) ev[k] = p.e[j * FE + k];
#pragma unroll
    for (int f = 0; f < HC; ++f) {
      float c = s_b1[f];
#pragma unroll
      for (int k = 0; k < FE; ++k) c = fmaf(ev[k], s_w1[f][k], c);
      const float d = c - ctr[f];
      acc[f] += d; acc[HC + f] = fmaf(d, d, acc[HC + f]);
    }
  }
  block_reduce_store<2 * HC>(acc, s_red, partial + (int64_t)blockIdx.x * (2 * HC));
}

// BatchNorm1d training-mode statistics (torch semantics: biased variance normalises, unbiased one updates the
// running estimate with momentum).  mode 0: sums -> mean.  mode 1: centred sums -> invstd, scale, shift, running stats.
__global__ void cls_bn_prepare_kernel(int mode, const float* sums, int64_t n, const float* gamma, const float* beta, float eps,
                                      float momentum, float* running_mean, float* running_var, float* mean, float* invstd,
                                      float* scale, float* shift) {
  const int f = threadIdx.x;
  if (f >= HC) return;
  if (mode == 0) { mean[f] = sums[f] / (float)n; return; }
  const float var = sums[HC + f] / (float)n;
  const float is = 1.f / sqrtf(var + eps);
  invstd[f] = is;
  scale[f] = gamma[f] * is;
  shift[f] = beta[f] - mean[f] * gamma[f] * is;
  if (running_mean) {
    running_mean[f] = (1.f - momentum) * running_mean[f] + momentum * mean[f];
    const float unbiased = n > 1 ? var * ((float)n / (float)(n - 1)) : var;
    running_var[f] = (1.f - momentum) * running_var[f] + momentum * unbiased;
  }
}

__global__ void __launch_bounds__(TB) cls_fwd_kernel(const ClsParams p, float* logits) {
  __shared__ float s_w1[HC][FE], s_b1[HC], s_w2[HC], s_aff[4 * HC];
  cls_load(p, s_w1, s_b1, s_w2, s_aff);
  const float b2 = p.b2[0];
  const int64_t stride = (int64_t)gridDim.x * TB;
  for (int64_t j = (int64_t)blockIdx.x * TB + threadIdx.x; j < p.n_edges; j += stride) {
    float ev[FE];
#pragma unroll
    for (int k = 0; k < FE; ++k) ev[k] = p.e[j * FE + k];
    float logit = b2;
#pragma unroll
    for (int f = 0; f < HC; ++f) {
      float c = s_b1[f];
#pragma unroll
      for (int k = 0; k < FE; ++k) c = fmaf(ev[k], s_w1[f][k], c);
      logit = fmaf(fmaxf(fmaf(c, s_aff[f], s_aff[HC + f]), 0.f), s_w2[f], logit);
    }
    logits[j] = logit;
  }
}

// accumulator layout of the classifier backward: [0,24) gW1 | [24,28) gb1 | [28,32) gw2 | 32 gb2 | [33,37) S2 = g_gamma | [37,41) S1 = g_beta
constexpr int CLS_ACC = HC * FE + HC + HC + 1 + HC + HC;
// pass 0: no BatchNorm / eval-mode BatchNorm, everything in one sweep.  pass 1 / 2: training-mode BatchNorm
// (1: batch sums S1 = sum gy, S2 = sum gy * c_hat and the last layer; 2: gc = scale * (gy - S1/E - c_hat * S2/E), ge, gW1, gb1).
__global__ void __launch_bounds__(TB) cls_bwd_kernel(const ClsParams p, int pass, const float* g_logits, const float* S, float* ge_cls,
                                                     float* partial) {
  __shared__ float s_w1[HC][FE], s_b1[HC], s_w2[HC], s_aff[4 * HC], s_red[NW * CLS_ACC];
  cls_load(p, s_w1, s_b1, s_w2, s_aff);
  float acc[CLS_ACC];
#pragma unroll
  for (int i = 0; i < CLS_ACC; ++i) acc[i] = 0.f;
  float s1[HC], s2[HC];
  const float inv_n = 1.f / (float)p.n_edges;
#pragma unroll
  for (int f = 0; f < HC; ++f) {
    s1[f] = (pass == 2) ? S[HC * FE + 3 * HC + 1 + f] * inv_n : 0.f;     // mean of gy
    s2[f] = (pass == 2) ? S[HC * FE + 2 * HC + 1 + f] * inv_n : 0.f;     // mean of gy * c_hat
  }
  const int64_t stride = (int64_t)gridDim.x * TB;
  for (int64_t j = (int64_t)blockIdx.x * TB + threadIdx.x; j < p.n_edges; j += stride) {
    float ev[FE], ge[FE];
#pragma unroll
    for (int k = 0; k < FE; ++k) { ev[k] = p.e[j * FE + k]; ge[k] = 0.f; }
    const float g = g_logits[j];
    if (pass != 2) acc[HC * FE + 2 * HC] += g;
#pragma unroll
    for (int f = 0; f < HC; ++f) {
      float c = s_b1[f];
#pragma unroll
      for (int k = 0; k < FE; ++k) c = fmaf(ev[k], s_w1[f][k], c);
      const float y = fmaf(c, s_aff[f], s_aff[HC + f]);
      const float chat = (c - s_aff[2 * HC + f]) * s_aff[3 * HC + f];
      const float gy = y > 0.f ? g * s_w2[f] : 0.f;
      if (pass != 2) {
        acc[HC * FE + HC + f] = fmaf(g, fmaxf(y, 0.f), acc[HC * FE + HC + f]);        // gw2
        acc[HC * FE + 2 * HC + 1 + f] = fmaf(gy, chat, acc[HC * FE + 2 * HC + 1 + f]);  // S2 / g_gamma
        acc[HC * FE + 3 * HC + 1 + f] += gy;                                              // S1 / g_beta
      }
      if (pass != 1) {
        const float gc = s_aff[f] * (gy - s1[f] - chat * s2[f]);
        acc[HC * FE + f] += gc;
#pragma unroll
        for (int k = 0; k < FE; ++k) { acc[f * FE + k] = fmaf(gc, ev[k], acc[f * FE + k]); ge[k] = fmaf(gc, s_w1[f][k], ge[k]); }
      }
    }
    if (pass != 1) {
#pragma unroll
      for (int k = 0; k < FE; ++k) ge_cls[j * FE + k] = ge[k];
    }
  }
  block_reduce_store<CLS_ACC>(acc, s_red, partial + (int64_t)blockIdx.x * CLS_ACC);
}

// ------------------------------------------------------------------------------------------------
// D. gP[v] = sum over out-edges of gz_e, gQ[v] = sum over in-edges (transposed CSR: colptr + tperm).  Warp per node:
//    lane = (edge slot 0..3) x (component 0..7); fixed tree over the 4 slots.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TB) pq_reduce_kernel(const int32_t* rowptr, const int32_t* colptr, const int32_t* tperm,
                                                       const float* gze, int64_t n_nodes, float* gP, float* gQ) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, o = lane & 7, slot = lane >> 3;
  for (int64_t v = (int64_t)blockIdx.x * NW + warp; v < n_nodes; v += (int64_t)gridDim.x * NW) {
    float a = 0.f, b = 0.f;
    for (int j = rowptr[v] + slot; j < rowptr[v + 1]; j += 4) a += gze[(int64_t)j * EP + o];
    for (int t = colptr[v] + slot; t < colptr[v + 1]; t += 4) b += gze[(int64_t)tperm[t] * EP + o];
    a += __shfl_xor_sync(0xffffffffu, a, 8); a += __shfl_xor_sync(0xffffffffu, a, 16);
    b += __shfl_xor_sync(0xffffffffu, b, 8); b += __shfl_xor_sync(0xffffffffu, b, 16);
    if (slot == 0) { gP[v * EP + o] = a; gQ[v * EP + o] = b; }
  }
}

// ------------------------------------------------------------------------------------------------
// E. backward of the per-node projections: gh_prev = gP W_s + gQ W_d + gR W_nx, weight-gradient partials.
//    One block = 64 (16) nodes.  partial [gridDim][44 * (K + 1)], index o * (K+1) + k; column K is the bias sum.
// ------------------------------------------------------------------------------------------------
// nodes per block of the projection backward (= rows of its `partial`): 16 for small graphs (more blocks than SMs)
__host__ __device__ inline int proj_bwd_nodes(int64_t n_nodes) { return n_nodes <= 64 * 148 * 4 ? 16 : 64; }
template <int NODES>
__global__ void __launch_bounds__(TB) node_project_bwd_kernel(const float* w_edge, const float* w_node, int edge_row, int node_row, int K,
                                                              const float* gP, const float* gQ, const float* gR, const float* h0,
                                                              const float* h, int64_t n_nodes, float* gh_prev, float* gh0_acc,
                                                              float* partial) {
  constexpr int OUT = 2 * FE + FN;
  extern __shared__ float sm[];
  float* w = sm;                       // [OUT][K]
  float* g = w + OUT * K;              // [NODES][OUT + 1]
  float* hs = g + NODES * (OUT + 1);   // [NODES][K + 1], last column = 1 (bias)
  for (int i = threadIdx.x; i < OUT * K; i += TB) {
    const int o = i / K, k = i - o * K;
    w[i] = o < FE ? w_edge[o * edge_row + k] : (o < 2 * FE ? w_edge[(o - FE) * edge_row + K + k] : w_node[(o - 2 * FE) * node_row + k]);
  }
  const int64_t n0 = (int64_t)blockIdx.x * NODES;
  for (int i = threadIdx.x; i < NODES * OUT; i += TB) {
    const int v = i / OUT, o = i - v * OUT;
    const int64_t node = n0 + v;
    float val = 0.f;
    if (node < n_nodes) val = o < FE ? gP[node * EP + o] : (o < 2 * FE ? gQ[node * EP + o - FE] : gR[node * FN + o - 2 * FE]);
    g[v * (OUT + 1) + o] = val;
  }
  for (int i = threadIdx.x; i < NODES * (K + 1); i += TB) {
    const int v = i / (K + 1), k = i - v * (K + 1);
    const int64_t node = n0 + v;
    float val = 0.f;
    if (node < n_nodes) val = k == K ? 1.f : (K == 2 * FN ? (k < FN ? h0[node * FN + k] : h[node * FN + k - FN]) : h[node * FN + k]);
    hs[i] = val;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NODES * K; i += TB) {
    const int v = i / K, k = i - v * K;
    const int64_t node = n0 + v;
    if (node >= n_nodes) continue;
    float a = 0.f;
    for (int o = 0; o < OUT; ++o) a = fmaf(g[v * (OUT + 1) + o], w[o * K + k], a);
    if (K == 2 * FN && k < FN) gh0_acc[node * FN + k] += a;        // the reattached block is the encoder output (mpn.py:285)
    else gh_prev[node * FN + (K == 2 * FN ? k - FN : k)] = a;
  }
  for (int i = threadIdx.x; i < OUT * (K + 1); i += TB) {
    const int o = i / (K + 1), k = i - o * (K + 1);
    float a = 0.f;
    for (int v = 0; v < NODES; ++v) a = fmaf(g[v * (OUT + 1) + o], hs[v * (K + 1) + k], a);
    partial[(int64_t)blockIdx.x * (OUT * (K + 1)) + i] = a;
  }
}

// ------------------------------------------------------------------------------------------------
// F. edge encoder (single Linear(fa, FE) + ReLU, mpn.py:116-126 with edge_fc_dims = []): weight-gradient partials.
//    accumulators [FE][MAX_FA] then gb [FE]
// ------------------------------------------------------------------------------------------------
constexpr int MAX_FA = 8;
__global__ void __launch_bounds__(TB) edge_enc_bwd_kernel(const float* edge_attr, int fa, const float* e0, const float* ge0, int64_t n_edges,
                                                          float* partial) {
  __shared__ float s_red[NW * (FE * MAX_FA + FE)];
  float acc[FE * MAX_FA + FE];
#pragma unroll
  for (int i = 0; i < FE * MAX_FA + FE; ++i) acc[i] = 0.f;
  const int64_t stride = (int64_t)gridDim.x * TB;
  for (int64_t j = (int64_t)blockIdx.x * TB + threadIdx.x; j < n_edges; j += stride) {
    float ea[MAX_FA];
#pragma unroll
    for (int k = 0; k < MAX_FA; ++k) ea[k] = k < fa ? edge_attr[j * fa + k] : 0.f;
#pragma unroll
    for (int o = 0; o < FE; ++o) {
      const float gz = e0[j * FE + o] > 0.f ? ge0[j * FE + o] : 0.f;
      acc[FE * MAX_FA + o] += gz;
#pragma unroll
      for (int k = 0; k < MAX_FA; ++k) acc[o * MAX_FA + k] = fmaf(gz, ea[k], acc[o * MAX_FA + k]);
    }
  }
  block_reduce_store<FE * MAX_FA + FE>(acc, s_red, partial + (int64_t)blockIdx.x * (FE * MAX_FA + FE));
}

// ------------------------------------------------------------------------------------------------
// G. dense layer backward (node encoder, mlp.py:14,18): gz = gy * [y > 0]; gb partials; gW = gz^T x (split over rows);
//    gx = gz W.
// ------------------------------------------------------------------------------------------------
constexpr int RG_ROWS = 16;
// 256 threads = (256 / out_dim) row lanes x out_dim columns: every lane sums its rows of the chunk in order, the lanes are
// folded in lane order (fixed summation order).  out_dim > 256: one row lane, columns strided.
__global__ void __launch_bounds__(256) relu_grad_bias_kernel(const float* __restrict__ y, const float* __restrict__ gy, int relu, int64_t rows,
                                                             int out_dim, float* __restrict__ gz, float* __restrict__ partial /* [chunks][out] */) {
  __shared__ float red[256];
  const int64_t r0 = (int64_t)blockIdx.x * RG_ROWS, r1 = min(r0 + RG_ROWS, rows);
  const int lanes = out_dim <= 256 ? 256 / out_dim : 1;
  const int lr = threadIdx.x / out_dim, c0 = threadIdx.x - lr * out_dim;
  for (int cb = 0; cb < out_dim; cb += 256) {
    const int c = cb + c0;
    float acc = 0.f;
    if (lr < lanes && c < out_dim) {
#pragma unroll 4
      for (int64_t r = r0 + lr; r < r1; r += lanes) {
        const float g = (!relu || y[r * out_dim + c] > 0.f) ? gy[r * out_dim + c] : 0.f;
        gz[r * out_dim + c] = g;
        acc += g;
      }
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    if (lr == 0 && c < out_dim) {
      float tot = red[c0];
      for (int l = 1; l < lanes; ++l) tot += red[l * out_dim + c0];
      partial[(int64_t)blockIdx.x * out_dim + c] = tot;
    }
    __syncthreads();
  }
}

// C[m][n] = sum_r A[r][m] * B[r][n] over this split's rows (the weight gradient gz^T x of a dense layer).  64 x 64 tile per CTA,
// 32 rows per iteration, on the tensor cores with fp32-level accuracy: every operand is split into its tf32 prefix and the
// (tf32-rounded) rest, and hi.hi + hi.lo + lo.hi are three mma.sync m16n8k8 into one fp32 accumulator (the dropped lo.lo term is
// below 2^-22 of the product).  8 warps = 2 (m) x 4 (n), a warp owns 32 x 16 outputs.  Fixed summation order: bit-reproducible.
__device__ __forceinline__ uint32_t f2tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32_u(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
constexpr int ATB_ROWS = 32, ATB_LD = 72;      // row stride 72: the fragment loads (k = lane & 3, m = lane >> 2) hit 32 banks
// PASSES = 3: the operand split described above (math = fp32).  PASSES = 1: one tf32 mma per tile, the arithmetic of the forward's
// tcgen05 encoder (math = tf32, what torch does under allow_tf32): legacy mma.sync issues a tf32 m16n8k8 every ~16 cycles per SM
// sub-partition, so the three-pass form is bound by exactly that pipe.
template <int PASSES>
__global__ void __launch_bounds__(256) atb_kernel(const float* __restrict__ A, const float* __restrict__ B, int64_t rows, int M, int Nn,
                                                  int64_t rows_per_split, float* __restrict__ partial /* [splits][M*Nn] */) {
  __shared__ __align__(16) float as[ATB_ROWS][ATB_LD], bs[ATB_ROWS][ATB_LD];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int wm = (warp >> 2) * 32, wn = (warp & 3) * 16;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const int64_t r_begin = (int64_t)blockIdx.z * rows_per_split, r_end = min(r_begin + rows_per_split, rows);
  const bool vec = (M % 4 == 0) && (Nn % 4 == 0) && ((((uintptr_t)A | (uintptr_t)B) & 15) == 0);
  float acc[2][2][4] = {};
  // 32 rows x 64 columns of each operand per iteration: 512 float4 per operand, 2 per thread; the next iteration's tile
  // travels in registers while this one is in the tensor cores
  float4 va[2], vb[2];
  auto fetch = [&](const int64_t r0) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int idx = tid + u * 256, lr = idx >> 4, lc = (idx & 15) * 4;
      const int64_t r = r0 + lr;
      va[u] = make_float4(0.f, 0.f, 0.f, 0.f); vb[u] = va[u];
      if (r < r_end) {
        if (vec) {
          if (m0 + lc < M) va[u] = __ldg(reinterpret_cast<const float4*>(A + r * M + m0 + lc));
          if (n0 + lc < Nn) vb[u] = __ldg(reinterpret_cast<const float4*>(B + r * Nn + n0 + lc));
        } else {
          float ta[4], tb[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            ta[q] = (m0 + lc + q < M) ? A[r * M + m0 + lc + q] : 0.f;
            tb[q] = (n0 + lc + q < Nn) ? B[r * Nn + n0 + lc + q] : 0.f;
          }
          va[u] = make_float4(ta[0], ta[1], ta[2], ta[3]); vb[u] = make_float4(tb[0], tb[1], tb[2], tb[3]);
        }
      }
    }
  };
  if (r_begin < r_end) fetch(r_begin);
  for (int64_t r0 = r_begin; r0 < r_end; r0 += ATB_ROWS) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int idx = tid + u * 256, lr = idx >> 4, lc = (idx & 15) * 4;
      *reinterpret_cast<float4*>(&as[lr][lc]) = va[u];
      *reinterpret_cast<float4*>(&bs[lr][lc]) = vb[u];
    }
    __syncthreads();
    if (r0 + ATB_ROWS < r_end) fetch(r0 + ATB_ROWS);
#pragma unroll
    for (int k0 = 0; k0 < ATB_ROWS; k0 += 8) {
      uint32_t ah[2][4], al[2][4], bh[2][2], bl[2][2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {          // A fragment of C^T's row block: element (m, k) = A[k][m]
        const float v[4] = {as[k0 + t][wm + 16 * i + g], as[k0 + t][wm + 16 * i + g + 8], as[k0 + t + 4][wm + 16 * i + g], as[k0 + t + 4][wm + 16 * i + g + 8]};
#pragma unroll
        for (int q = 0; q < 4; ++q) { ah[i][q] = f2tf32(v[q]); al[i][q] = PASSES == 3 ? f2tf32(v[q] - __uint_as_float(ah[i][q])) : 0u; }
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {          // B fragment: element (k, n) = B[k][n]
        const float v[2] = {bs[k0 + t][wn + 8 * j + g], bs[k0 + t + 4][wn + 8 * j + g]};
#pragma unroll
        for (int q = 0; q < 2; ++q) { bh[j][q] = f2tf32(v[q]); bl[j][q] = PASSES == 3 ? f2tf32(v[q] - __uint_as_float(bh[j][q])) : 0u; }
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (PASSES == 3) {
            mma_tf32_u(acc[i][j], al[i], bh[j][0], bh[j][1]);     // small terms first
            mma_tf32_u(acc[i][j], ah[i], bl[j][0], bl[j][1]);
          }
          mma_tf32_u(acc[i][j], ah[i], bh[j][0], bh[j][1]);
        }
    }
    __syncthreads();
  }
  float* out = partial + (int64_t)blockIdx.z * M * Nn;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int m = m0 + wm + 16 * i + g + ((q & 2) ? 8 : 0), n = n0 + wn + 8 * j + 2 * t + (q & 1);
        if (m < M && n < Nn) out[(int64_t)m * Nn + n] = acc[i][j][q];
      }
}

// gx[r][i] = sum_o gz[r][o] * W[o][i]
__global__ void __launch_bounds__(256) gz_times_w_kernel(const float* gz, const float* W, int64_t rows, int out_dim, int in_dim, float* gx) {
  const int64_t total = rows * in_dim, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int64_t r = t / in_dim;
    const int i = (int)(t - r * in_dim);
    float a = 0.f;
    for (int o = 0; o < out_dim; ++o) a = fmaf(gz[r * out_dim + o], W[(int64_t)o * in_dim + i], a);
    gx[t] = a;
  }
}

// ------------------------------------------------------------------------------------------------
// H. dst[r * dst_stride + c] (+)= sum_b partial[b * width + src_off + r * src_stride + c], b ascending: the one place
//    where per-block partial sums meet, in a fixed order.  One block per table entry.
// ------------------------------------------------------------------------------------------------
struct FinTable { gnncca_fin_seg s[GNNCCA_MAX_FIN_SEGS]; int n; };
constexpr int FIN_WARP_MODE_BLOCKS = 64;   // from this many partial rows on, a WARP folds one element (lanes stride the rows)
__global__ void __launch_bounds__(256) finalize_kernel(const FinTable t) {
  const gnncca_fin_seg sg = t.s[blockIdx.x];
  const int total = sg.rows * sg.cols;
  if (sg.n_blocks >= FIN_WARP_MODE_BLOCKS) {
    // few elements, many partial rows (per-edge / per-node kernels): lane l sums rows l, l + 32, ... in order, then a fixed
    // xor tree over the lanes.  The order is a function of n_blocks only.
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = blockIdx.y * 8 + warp; i < total; i += 8 * gridDim.y) {
      const int r = i / sg.cols, c = i - r * sg.cols;
      const float* src = sg.partial + sg.src_off + r * sg.src_stride + c;
      float acc = 0.f;
#pragma unroll 4
      for (int b = lane; b < sg.n_blocks; b += 32) acc += __ldg(src + (int64_t)b * sg.width);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) {
        float* d = sg.dst + (int64_t)r * sg.dst_stride + c;
        *d = sg.accumulate ? *d + acc : acc;
      }
    }
    return;
  }
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < total; i += blockDim.x * gridDim.y) {
    const int r = i / sg.cols, c = i - r * sg.cols;
    float acc = 0.f;
    const float* src = sg.partial + sg.src_off + r * sg.src_stride + c;
#pragma unroll 8
    for (int b = 0; b < sg.n_blocks; ++b) acc += __ldg(src + (int64_t)b * sg.width);     // same order, 8 loads in flight
    float* d = sg.dst + (int64_t)r * sg.dst_stride + c;
    *d = sg.accumulate ? *d + acc : acc;
  }
}


// ------------------------------------------------------------------------------------------------
// I. The whole message-passing backward of a batch of SMALL frames in ONE launch: a CTA per frame walks the steps
//    t = L .. 1 through the phases A-E above (same arithmetic per element: the bodies below are those kernels with the
//    frame's node / edge range in place of the grid), separated by __syncthreads() instead of kernel boundaries.  A frame is
//    closed (no edge leaves it, Batch.from_data_list, train.py:450), so every buffer row a phase reads was written by this
//    CTA.  For EPFL-sized frames (tens of nodes, ~10^3 edges) a kernel boundary costs more than the phase's work: 18 launches
//    per backward become one.  Weight-gradient partials: one row per FRAME and step, folded by gnncca_finalize_grads.
// ------------------------------------------------------------------------------------------------
constexpr int PB_NODES = 16;
struct FramesBwd {
  const int64_t* graph_ptr; const int64_t* edge_ptr;
  const int32_t* rowptr; const int32_t* src; const int32_t* tperm;
  const float* w_edge; const float* w_node; int edge_row, node_row, K;
  const float* w1; const float* b1; const float* w2; const float* b2; const float* b_node;
  const float* e0; const float* h0;
  int n_steps, agg; int64_t n_nodes; int max_n, max_e;
  gnncca_frame_bwd_step st[GNNCCA_FRAMES_BWD_MAX_STEPS];
  float* ge_buf[2]; float* ge0_acc; float* gh0_acc;
};

constexpr int TBF = 512, NWF = TBF / 32;      // a frame's CTA: 16 warps (8 left each scheduler with two warps of long dependent chains)
template <bool REATT_E, int K>
__global__ void __launch_bounds__(TBF) frames_bwd_kernel(const FramesBwd p) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  constexpr int OFF = REATT_E ? FE : 0;
  constexpr int OUT = 2 * FE + FN;
  __shared__ float s_wne[FE][FN];
  __shared__ float s_wee[FE][EIN];
  __shared__ float s_w1[HC][FE], s_b1[HC], s_w2[HC], s_aff[4 * HC];
  __shared__ float s_red[NWF * (FE * EIN > CLS_ACC ? FE * EIN : CLS_ACC)];
  __shared__ float s_agg[NWF][FE * FN];
  __shared__ float s_pw[OUT * K];                    // projection weights [OUT][K]
  __shared__ float s_ph[PB_NODES * (K + 1)];
  // the frame's state of the current step (dynamic): every phase reads and writes it here -- through global memory each
  // dependent access would cost an L2 round trip (stores do not allocate in L1), and a CTA per frame has nothing to hide it with
  extern __shared__ __align__(16) float dyn[];
  float* s_e = dyn;                                     // [max_e][FE]   e^t
  float* s_gze = s_e + (size_t)p.max_e * FE;            // [max_e][EP]   classifier gradient, then gz_e
  float* s_R = s_gze + (size_t)p.max_e * EP;            // [max_n][FN]
  float* s_gh = s_R + p.max_n * FN;                     // [max_n][FN]   gradient wrt h^t (max: divided by the ties)
  float* s_hv = s_gh + p.max_n * FN;                    // [max_n][FN]   max only: the recomputed maxima
  float* s_gR = s_hv + p.max_n * FN;                    // [max_n][FN]
  float* s_gP = s_gR + p.max_n * FN;                    // [max_n][EP]
  float* s_gQ = s_gP + p.max_n * EP;                    // [max_n][EP]
  float* s_pg = s_gQ + p.max_n * EP;                    // [max_n][OUT + 1]  phase E's left operand
  int* s_rp = reinterpret_cast<int*>(s_pg + p.max_n * (OUT + 1));   // [max_n + 1] local row offsets
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int f = blockIdx.x;
  const int64_t n0 = p.graph_ptr[f], eb = p.edge_ptr[f];
  const int n = (int)(p.graph_ptr[f + 1] - n0), m = (int)(p.edge_ptr[f + 1] - eb);
  if (n > p.max_n || m > p.max_e) return;               // the caller sized the launch from the batch maxima: cannot happen
  for (int i = tid; i < FE * FN; i += TBF) s_wne[i / FN][i % FN] = p.w_node[(i % FN) * p.node_row + K + i / FN];
  for (int i = tid; i < FE * EIN; i += TBF) s_wee[i / EIN][i % EIN] = p.w_edge[(i / EIN) * p.edge_row + 2 * K + i % EIN];
  for (int i = tid; i < OUT * K; i += TBF) {
    const int o = i / K, k = i - o * K;
    s_pw[i] = o < FE ? p.w_edge[o * p.edge_row + k] : (o < 2 * FE ? p.w_edge[(o - FE) * p.edge_row + K + k] : p.w_node[(o - 2 * FE) * p.node_row + k]);
  }
  if (p.w1) {
    for (int i = tid; i < HC * FE; i += TBF) s_w1[i / FE][i % FE] = p.w1[i];
    for (int i = tid; i < HC; i += TBF) { s_b1[i] = p.b1[i]; s_w2[i] = p.w2[i]; }
  }
  for (int i = tid; i <= n; i += TBF) s_rp[i] = (int)(p.rowptr[n0 + i] - eb);
  float wne[FE];                                        // lane = feature (phase A)
#pragma unroll
  for (int k = 0; k < FE; ++k) wne[k] = p.w_node[lane * p.node_row + K + k];
  __syncthreads();

  for (int t = p.n_steps; t >= 1; --t) {
    const gnncca_frame_bwd_step& st = p.st[t - 1];
    const float* gh = t < p.n_steps ? p.st[t].gh_prev : nullptr;       // gradient wrt h^t: what step t + 1 wrote
    const float* ge_in = t < p.n_steps ? p.ge_buf[t & 1] : nullptr;
    float* ge_out = p.ge_buf[(t - 1) & 1];
    const bool has_cls = st.g_logits != nullptr;
    // ---- stage the step: e^t, R, gh (coalesced, independent loads) ----
    for (int i = tid; i < m * FE; i += TBF) s_e[i] = st.e[eb * FE + i];
    if (gh != nullptr) {
      if (st.R != nullptr) {
        for (int i = tid; i < n * FN; i += TBF) s_R[i] = st.R[n0 * FN + i];
      } else {
        // the forward kept no projections (the all-steps forward of small frames): R = [h0,] h^{t-1} W_nx^T + b_n here
        for (int i = tid; i < n * FN; i += TBF) {
          const int v = i / FN, c = i - v * FN;
          const float* wr = s_pw + (2 * FE + c) * K;
          float a = p.b_node[c];
          if (K == 2 * FN) {
#pragma unroll 8
            for (int k = 0; k < FN; ++k) a = fmaf(p.h0[(n0 + v) * FN + k], wr[k], a);
          }
#pragma unroll 8
          for (int k = 0; k < FN; ++k) a = fmaf(st.h_prev[(n0 + v) * FN + k], wr[(K == 2 * FN ? FN : 0) + k], a);
          s_R[i] = a;
        }
      }
      for (int i = tid; i < n * FN; i += TBF) s_gh[i] = gh[n0 * FN + i];
    }
    if (has_cls && tid < HC) {
      s_aff[tid] = st.aff ? st.aff[tid] : 1.f; s_aff[HC + tid] = st.aff ? st.aff[HC + tid] : 0.f;
      s_aff[2 * HC + tid] = st.aff ? st.aff[2 * HC + tid] : 0.f; s_aff[3 * HC + tid] = st.aff ? st.aff[3 * HC + tid] : 1.f;
    }
    __syncthreads();

    // ---- C. classifier backward (no BatchNorm / eval-mode BatchNorm): its gradient wrt e^t goes to s_gze ----
    if (has_cls) {
      float acc[CLS_ACC];
#pragma unroll
      for (int i = 0; i < CLS_ACC; ++i) acc[i] = 0.f;
      for (int j = tid; j < m; j += TBF) {
        float ev[FE], ge[FE];
#pragma unroll
        for (int k = 0; k < FE; ++k) { ev[k] = s_e[j * FE + k]; ge[k] = 0.f; }
        const float g = st.g_logits[eb + j];
        acc[HC * FE + 2 * HC] += g;
#pragma unroll
        for (int ff = 0; ff < HC; ++ff) {
          float c = s_b1[ff];
#pragma unroll
          for (int k = 0; k < FE; ++k) c = fmaf(ev[k], s_w1[ff][k], c);
          const float y = fmaf(c, s_aff[ff], s_aff[HC + ff]);
          const float chat = (c - s_aff[2 * HC + ff]) * s_aff[3 * HC + ff];
          const float gy = y > 0.f ? g * s_w2[ff] : 0.f;
          acc[HC * FE + HC + ff] = fmaf(g, fmaxf(y, 0.f), acc[HC * FE + HC + ff]);
          acc[HC * FE + 2 * HC + 1 + ff] = fmaf(gy, chat, acc[HC * FE + 2 * HC + 1 + ff]);
          acc[HC * FE + 3 * HC + 1 + ff] += gy;
          const float gc = s_aff[ff] * gy;
          acc[HC * FE + ff] += gc;
#pragma unroll
          for (int k = 0; k < FE; ++k) { acc[ff * FE + k] = fmaf(gc, ev[k], acc[ff * FE + k]); ge[k] = fmaf(gc, s_w1[ff][k], ge[k]); }
        }
#pragma unroll
        for (int k = 0; k < FE; ++k) s_gze[j * EP + k] = ge[k];
      }
      block_reduce_store<CLS_ACC>(acc, s_red, st.pa_cls + (int64_t)f * CLS_ACC);
    }

    // ---- A. node side of the node model: warp per node, lane = feature ----
    if (gh != nullptr) {
      float gw[FE];
#pragma unroll
      for (int k = 0; k < FE; ++k) gw[k] = 0.f;
      for (int v = warp; v < n; v += NWF) {
        const int j0 = s_rp[v], j1 = s_rp[v + 1];
        const float r = s_R[v * FN + lane], gh_in = s_gh[v * FN + lane];
        const float inv_deg = j1 > j0 ? 1.f / (float)(j1 - j0) : 0.f;
        float ghv = gh_in, hval = 0.f;
        if (p.agg == GNNCCA_AGG_MAX) {
          int ties = 0;
          for (int j = j0; j < j1; ++j) {
            float z = r;
#pragma unroll
            for (int k = 0; k < FE; ++k) z = fmaf(s_e[j * FE + k], wne[k], z);
            if (z > hval) { hval = z; ties = 1; }
            else if (z > 0.f && z == hval) ++ties;
          }
          ghv = ties > 0 ? gh_in / (float)ties : 0.f;
          s_gh[v * FN + lane] = ghv;                      // what the edge phase must see as gh
          s_hv[v * FN + lane] = hval;
        }
        float acc = 0.f;
        for (int j = j0; j < j1; ++j) {
          float ev[FE];
#pragma unroll
          for (int k = 0; k < FE; ++k) ev[k] = s_e[j * FE + k];
          float z = r;
#pragma unroll
          for (int k = 0; k < FE; ++k) z = fmaf(ev[k], wne[k], z);
          const float g = agg_grad(p.agg, ghv, z, hval, inv_deg);
          acc += g;
#pragma unroll
          for (int k = 0; k < FE; ++k) gw[k] = fmaf(g, ev[k], gw[k]);
        }
        s_gR[v * FN + lane] = acc;
      }
#pragma unroll
      for (int k = 0; k < FE; ++k) s_agg[warp][lane * FE + k] = gw[k];
      __syncthreads();
      for (int i = tid; i < FE * FN; i += TBF) {
        float v = 0.f;
        for (int w = 0; w < NWF; ++w) v += s_agg[w][i];
        st.pa_agg[(int64_t)f * (FE * FN) + i] = v;
      }
    } else {
      for (int i = tid; i < n * FN; i += TBF) s_gR[i] = 0.f;     // the node update of the last step is dead
    }
    __syncthreads();

    // ---- B. per-edge backward of the node model (edge side) and of the edge model ----
    {
      float gw[FE * EIN];
#pragma unroll
      for (int i = 0; i < FE * EIN; ++i) gw[i] = 0.f;
      for (int j = tid; j < m; j += TBF) {
        const int64_t jg = eb + j;
        float ev[FE], ge[FE];
#pragma unroll
        for (int k = 0; k < FE; ++k) {
          ev[k] = s_e[j * FE + k];
          ge[k] = (ge_in ? ge_in[jg * FE + k] : 0.f) + (has_cls ? s_gze[j * EP + k] : 0.f);
        }
        if (gh != nullptr) {
          const int s = (int)(p.src[jg] - n0);
          const float inv_deg = 1.f / (float)max(s_rp[s + 1] - s_rp[s], 1);
          const float4* R4 = reinterpret_cast<const float4*>(s_R + s * FN);
          const float4* G4 = reinterpret_cast<const float4*>(s_gh + s * FN);
          const float4* H4 = reinterpret_cast<const float4*>(s_hv + s * FN);
#pragma unroll
          for (int c4 = 0; c4 < FN / 4; ++c4) {
            const float4 r = R4[c4], g = G4[c4];
            const float4 hv = p.agg == GNNCCA_AGG_MAX ? H4[c4] : make_float4(0.f, 0.f, 0.f, 0.f);
            const float rr[4] = {r.x, r.y, r.z, r.w}, gg[4] = {g.x, g.y, g.z, g.w}, hh[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int c = c4 * 4 + u;
              float z = rr[u];
#pragma unroll
              for (int k = 0; k < FE; ++k) z = fmaf(ev[k], s_wne[k][c], z);
              const float gz = agg_grad(p.agg, gg[u], z, hh[u], inv_deg);
#pragma unroll
              for (int k = 0; k < FE; ++k) ge[k] = fmaf(gz, s_wne[k][c], ge[k]);
            }
          }
        }
        float gze[FE];
#pragma unroll
        for (int o = 0; o < FE; ++o) gze[o] = ev[o] > 0.f ? ge[o] : 0.f;          // ReLU gate of the edge MLP
        *reinterpret_cast<float4*>(s_gze + j * EP) = make_float4(gze[0], gze[1], gze[2], gze[3]);
        *reinterpret_cast<float4*>(s_gze + j * EP + 4) = make_float4(gze[4], gze[5], 0.f, 0.f);
        float ein[EIN];
        if (REATT_E) {
#pragma unroll
          for (int k = 0; k < FE; ++k) ein[k] = p.e0[jg * FE + k];
        }
#pragma unroll
        for (int k = 0; k < FE; ++k) ein[OFF + k] = st.e_prev[jg * FE + k];
#pragma unroll
        for (int k = 0; k < FE; ++k) {
          float a = 0.f;
#pragma unroll
          for (int o = 0; o < FE; ++o) a = fmaf(gze[o], s_wee[o][OFF + k], a);
          ge_out[jg * FE + k] = a;
        }
        if (REATT_E) {
#pragma unroll
          for (int k = 0; k < FE; ++k) {
            float a = p.ge0_acc[jg * FE + k];
#pragma unroll
            for (int o = 0; o < FE; ++o) a = fmaf(gze[o], s_wee[o][k], a);
            p.ge0_acc[jg * FE + k] = a;
          }
        }
#pragma unroll
        for (int o = 0; o < FE; ++o)
#pragma unroll
          for (int kk = 0; kk < EIN; ++kk) gw[o * EIN + kk] = fmaf(gze[o], ein[kk], gw[o * EIN + kk]);
      }
      block_reduce_store<FE * EIN>(gw, s_red, st.pa_edge + (int64_t)f * (FE * EIN));      // ends with __syncthreads()
    }

    // ---- D. gP / gQ: warp per node, lane = (edge slot 0..3) x (component 0..7); in-edges through the reverse-edge table ----
    {
      const int o = lane & 7, slot = lane >> 3;
      for (int v = warp; v < n; v += NWF) {
        float a = 0.f, b = 0.f;
        for (int j = s_rp[v] + slot; j < s_rp[v + 1]; j += 4) {
          a += s_gze[j * EP + o];
          b += s_gze[(int)(p.tperm[eb + j] - eb) * EP + o];          // symmetric frames: colptr == rowptr
        }
        a += __shfl_xor_sync(0xffffffffu, a, 8); a += __shfl_xor_sync(0xffffffffu, a, 16);
        b += __shfl_xor_sync(0xffffffffu, b, 8); b += __shfl_xor_sync(0xffffffffu, b, 16);
        if (slot == 0) { s_gP[v * EP + o] = a; s_gQ[v * EP + o] = b; }
      }
    }
    __syncthreads();

    // ---- E. backward of the per-node projections, 16 nodes at a time; the weight-gradient partial of the whole frame ----
    {
      constexpr int PER = (OUT * (K + 1) + TBF - 1) / TBF;
      float pacc[PER];
#pragma unroll
      for (int q = 0; q < PER; ++q) pacc[q] = 0.f;
      const float* h_prev = st.h_prev;
      for (int i = tid; i < n * OUT; i += TBF) {                // [gP | gQ | gR] of every node of the frame, row stride OUT + 1
        const int v = i / OUT, o = i - v * OUT;
        s_pg[v * (OUT + 1) + o] = o < FE ? s_gP[v * EP + o] : (o < 2 * FE ? s_gQ[v * EP + o - FE] : s_gR[v * FN + o - 2 * FE]);
      }
      for (int c0 = 0; c0 < n; c0 += PB_NODES) {
        const int nc = min(PB_NODES, n - c0);
        for (int i = tid; i < PB_NODES * (K + 1); i += TBF) {
          const int v = i / (K + 1), k = i - v * (K + 1);
          const int64_t node = n0 + c0 + v;
          float val = 0.f;
          if (v < nc) val = k == K ? 1.f : (K == 2 * FN ? (k < FN ? p.h0[node * FN + k] : h_prev[node * FN + k - FN]) : h_prev[node * FN + k]);
          s_ph[i] = val;
        }
        __syncthreads();
        for (int i = tid; i < nc * K; i += TBF) {
          const int v = i / K, k = i - v * K;
          const int64_t node = n0 + c0 + v;
          float a = 0.f;
          for (int o = 0; o < OUT; ++o) a = fmaf(s_pg[(c0 + v) * (OUT + 1) + o], s_pw[o * K + k], a);
          if (K == 2 * FN && k < FN) p.gh0_acc[node * FN + k] += a;        // the reattached block is the encoder output (mpn.py:285)
          else st.gh_prev[node * FN + (K == 2 * FN ? k - FN : k)] = a;
        }
#pragma unroll
        for (int q = 0; q < PER; ++q) {
          const int i = tid + q * TBF;
          if (i < OUT * (K + 1)) {
            const int o = i / (K + 1), k = i - o * (K + 1);
            float a = pacc[q];
            for (int v = 0; v < nc; ++v) a = fmaf(s_pg[(c0 + v) * (OUT + 1) + o], s_ph[v * (K + 1) + k], a);
            pacc[q] = a;
          }
        }
        __syncthreads();
      }
#pragma unroll
      for (int q = 0; q < PER; ++q) {
        const int i = tid + q * TBF;
        if (i < OUT * (K + 1)) st.pa_proj[(int64_t)f * (OUT * (K + 1)) + i] = pacc[q];
      }
    }
    __syncthreads();
  }
}

}  // namespace train
}  // namespace gnncca

using namespace gnncca;
namespace tr = gnncca::train;

static int blocks_for(int64_t n, int per_block) {
  int64_t b = ceil_div(n > 0 ? n : 1, per_block);
  const int64_t cap = (int64_t)kNumSMs * 4;
  return (int)(b < cap ? b : cap);
}

extern "C" int32_t gnncca_train_blocks(int64_t n_items, int32_t kind) {
  switch (kind) {
    case GNNCCA_BLOCKS_PER_EDGE: return blocks_for(n_items, tr::TB);
    case GNNCCA_BLOCKS_PER_NODE_WARP: return blocks_for(n_items, tr::NW);
    case GNNCCA_BLOCKS_NODE_PROJECT: return (int32_t)ceil_div(n_items > 0 ? n_items : 1, tr::proj_bwd_nodes(n_items));
    case GNNCCA_BLOCKS_ROW_CHUNKS: return (int32_t)ceil_div(n_items > 0 ? n_items : 1, tr::RG_ROWS);
    default: return -1;
  }
}

static bool dims_ok(const gnncca_mp_cfg* cfg) { return cfg && cfg->f_node == tr::FN && cfg->f_edge == tr::FE; }

extern "C" int gnncca_node_agg_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, int64_t n_nodes,
                                   const float* R, const float* e, const float* gh, const float* h_out, float* gR, float* gh_eff,
                                   float* partial, void* stream) {
  if (!dims_ok(cfg)) return cfg ? GNNCCA_ERR_UNSUPPORTED : GNNCCA_ERR_BAD_ARG;
  if (n_nodes == 0) return GNNCCA_OK;
  if (!w || !rowptr || !R || !e || !gh || !gR || !partial || (cfg->agg == GNNCCA_AGG_MAX && (!h_out || !gh_eff))) return GNNCCA_ERR_BAD_ARG;
  const int K = cfg->reattach_nodes ? 2 * tr::FN : tr::FN;
  tr::NodeAggBwd p{rowptr, R, e, gh, h_out, w->w_node, K + tr::FE, K, n_nodes, cfg->agg, gR, partial, gh_eff};
  tr::node_agg_bwd_kernel<<<blocks_for(n_nodes, tr::NW), tr::TB, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

extern "C" int gnncca_edge_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, const int32_t* src,
                               int64_t n_edges, const float* R, const float* e, const float* e_prev, const float* e0, const float* gh,
                               const float* h_out, const float* ge_in, const float* ge_cls, float* gze, float* ge_prev, float* ge0_acc,
                               float* partial, void* stream) {
  if (!dims_ok(cfg)) return cfg ? GNNCCA_ERR_UNSUPPORTED : GNNCCA_ERR_BAD_ARG;
  if (n_edges == 0) return GNNCCA_OK;
  if (!w || !rowptr || !src || !e || !e_prev || !gze || !ge_prev || !partial) return GNNCCA_ERR_BAD_ARG;
  if (gh && (!R || (cfg->agg == GNNCCA_AGG_MAX && !h_out))) return GNNCCA_ERR_BAD_ARG;
  if (cfg->reattach_edges && (!e0 || !ge0_acc)) return GNNCCA_ERR_BAD_ARG;
  const int K = cfg->reattach_nodes ? 2 * tr::FN : tr::FN;
  const int EIN = cfg->reattach_edges ? 2 * tr::FE : tr::FE;
  tr::EdgeBwd p{rowptr, src, R, e, e_prev, e0, gh, h_out, ge_in, ge_cls, w->w_node, w->w_edge, K + tr::FE, 2 * K + EIN, K,
                n_edges, cfg->agg, gze, ge_prev, ge0_acc, partial};
  const int grid = blocks_for(n_edges, tr::TB);
  if (cfg->reattach_edges) tr::edge_bwd_kernel<true><<<grid, tr::TB, 0, (cudaStream_t)stream>>>(p);
  else tr::edge_bwd_kernel<false><<<grid, tr::TB, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

static tr::ClsParams cls_params(const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* aff) {
  // aff: device [4][HC] = scale | shift | mean | invstd, or NULL for "no BatchNorm"
  return tr::ClsParams{e, n_edges, w->w_c1, w->b_c1, w->w_c2, w->b_c2, aff, aff ? aff + tr::HC : nullptr, aff ? aff + 2 * tr::HC : nullptr,
                       aff ? aff + 3 * tr::HC : nullptr};
}

extern "C" int gnncca_cls_moments(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* center,
                                  float* partial, void* stream) {
  if (!cfg || !w || cfg->cls_hidden != tr::HC || cfg->f_edge != tr::FE) return GNNCCA_ERR_UNSUPPORTED;
  if (n_edges <= 0 || !e || !partial || !w->w_c1 || !w->b_c1) return GNNCCA_ERR_BAD_ARG;
  tr::cls_moments_kernel<<<blocks_for(n_edges, tr::TB), tr::TB, 0, (cudaStream_t)stream>>>(cls_params(w, e, n_edges, nullptr), center, partial);
  return check_launch();
}

extern "C" int gnncca_cls_bn_prepare(int32_t mode, const float* sums, int64_t n_edges, const float* gamma, const float* beta, float eps,
                                     float momentum, float* running_mean, float* running_var, float* aff, void* stream) {
  if (!sums || !aff || n_edges <= 0 || (mode != 0 && mode != 1) || (mode == 1 && (!gamma || !beta))) return GNNCCA_ERR_BAD_ARG;
  tr::cls_bn_prepare_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(mode, sums, n_edges, gamma, beta, eps, momentum, running_mean, running_var,
                                                               aff + 2 * tr::HC, aff + 3 * tr::HC, aff, aff + tr::HC);
  return check_launch();
}

extern "C" int gnncca_cls_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* aff,
                              float* logits, void* stream) {
  if (!cfg || !w || cfg->cls_hidden != tr::HC || cfg->f_edge != tr::FE) return GNNCCA_ERR_UNSUPPORTED;
  if (n_edges == 0) return GNNCCA_OK;
  if (!e || !logits || !w->w_c1 || !w->b_c1 || !w->w_c2 || !w->b_c2) return GNNCCA_ERR_BAD_ARG;
  tr::cls_fwd_kernel<<<blocks_for(n_edges, tr::TB), tr::TB, 0, (cudaStream_t)stream>>>(cls_params(w, e, n_edges, aff), logits);
  return check_launch();
}

extern "C" int gnncca_cls_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* aff,
                              int32_t pass, const float* g_logits, const float* batch_sums, float* ge_cls, float* partial, void* stream) {
  if (!cfg || !w || cfg->cls_hidden != tr::HC || cfg->f_edge != tr::FE) return GNNCCA_ERR_UNSUPPORTED;
  if (n_edges == 0) return GNNCCA_OK;
  if (!e || !g_logits || !partial || pass < 0 || pass > 2 || (pass != 1 && !ge_cls) || (pass == 2 && !batch_sums)) return GNNCCA_ERR_BAD_ARG;
  tr::cls_bwd_kernel<<<blocks_for(n_edges, tr::TB), tr::TB, 0, (cudaStream_t)stream>>>(cls_params(w, e, n_edges, aff), pass, g_logits,
                                                                                       batch_sums, ge_cls, partial);
  return check_launch();
}

extern "C" int gnncca_pq_reduce(const int32_t* rowptr, const int32_t* colptr, const int32_t* tperm, const float* gze, int64_t n_nodes,
                                float* gP, float* gQ, void* stream) {
  if (n_nodes == 0) return GNNCCA_OK;
  if (!rowptr || !colptr || !tperm || !gze || !gP || !gQ) return GNNCCA_ERR_BAD_ARG;
  tr::pq_reduce_kernel<<<blocks_for(n_nodes, tr::NW), tr::TB, 0, (cudaStream_t)stream>>>(rowptr, colptr, tperm, gze, n_nodes, gP, gQ);
  return check_launch();
}

extern "C" int gnncca_node_project_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* gP, const float* gQ, const float* gR,
                                       const float* h0, const float* h, int64_t n_nodes, float* gh_prev, float* gh0_acc, float* partial,
                                       void* stream) {
  if (!dims_ok(cfg)) return cfg ? GNNCCA_ERR_UNSUPPORTED : GNNCCA_ERR_BAD_ARG;
  if (n_nodes == 0) return GNNCCA_OK;
  if (!w || !gP || !gQ || !gR || !h || !gh_prev || !partial || (cfg->reattach_nodes && (!h0 || !gh0_acc))) return GNNCCA_ERR_BAD_ARG;
  const int K = cfg->reattach_nodes ? 2 * tr::FN : tr::FN, EIN = cfg->reattach_edges ? 2 * tr::FE : tr::FE, OUT = 2 * tr::FE + tr::FN;
  const int nodes = tr::proj_bwd_nodes(n_nodes);
  const size_t smem = sizeof(float) * (OUT * K + nodes * (OUT + 1) + nodes * (K + 1));
  auto kern = nodes == 16 ? tr::node_project_bwd_kernel<16> : tr::node_project_bwd_kernel<64>;
  if (smem > 48 * 1024) {
    int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (rc) return rc;
  }
  kern<<<(unsigned)ceil_div(n_nodes, nodes), tr::TB, smem, (cudaStream_t)stream>>>(w->w_edge, w->w_node, 2 * K + EIN, K + tr::FE, K, gP, gQ, gR, h0,
                                                                               h, n_nodes, gh_prev, gh0_acc, partial);
  return check_launch();
}

extern "C" int64_t gnncca_mp_frames_bwd_smem_bytes(int32_t max_frame_nodes, int32_t max_frame_edges) {
  max_frame_edges = (max_frame_edges + 3) & ~3;          // keeps the arrays behind e^t 16-byte aligned
  return 4 * ((int64_t)max_frame_edges * (tr::FE + tr::EP) + (int64_t)max_frame_nodes * (4 * tr::FN + 2 * tr::EP + 2 * tr::FE + tr::FN + 1) +
              max_frame_nodes + 1 + 4);
}

extern "C" int gnncca_mp_frames_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int64_t* graph_ptr, const int64_t* edge_ptr,
                                    int64_t n_graphs, const int32_t* rowptr, const int32_t* src, const int32_t* tperm, int64_t n_nodes,
                                    int64_t n_edges, int32_t max_frame_nodes, int32_t max_frame_edges,
                                    const gnncca_frame_bwd_step* host_steps, int32_t n_steps, const float* e0, const float* h0,
                                    float* ge_a, float* ge_b, float* ge0_acc, float* gh0_acc, void* stream) {
  if (!dims_ok(cfg)) return cfg ? GNNCCA_ERR_UNSUPPORTED : GNNCCA_ERR_BAD_ARG;
  if (n_steps < 1 || n_steps > GNNCCA_FRAMES_BWD_MAX_STEPS || n_graphs < 0 || n_graphs > INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  if (n_graphs == 0 || n_edges == 0) return GNNCCA_OK;
  if (!w || !graph_ptr || !edge_ptr || !rowptr || !src || !tperm || !host_steps || !ge_a || !ge_b) return GNNCCA_ERR_BAD_ARG;
  if (max_frame_nodes < 1 || max_frame_edges < 1) return GNNCCA_ERR_BAD_ARG;
  const int64_t dyn_bytes = gnncca_mp_frames_bwd_smem_bytes(max_frame_nodes, max_frame_edges);
  if (dyn_bytes > GNNCCA_FRAMES_BWD_SMEM_LIMIT) return GNNCCA_ERR_UNSUPPORTED;         // frames too large for a CTA each: the per-step entries
  if (cfg->reattach_edges && (!e0 || !ge0_acc)) return GNNCCA_ERR_BAD_ARG;
  if (cfg->reattach_nodes && (!h0 || !gh0_acc)) return GNNCCA_ERR_BAD_ARG;
  const int K = cfg->reattach_nodes ? 2 * tr::FN : tr::FN;
  const int EIN = cfg->reattach_edges ? 2 * tr::FE : tr::FE;
  tr::FramesBwd p{};
  p.graph_ptr = graph_ptr; p.edge_ptr = edge_ptr; p.rowptr = rowptr; p.src = src; p.tperm = tperm;
  p.w_edge = w->w_edge; p.w_node = w->w_node; p.edge_row = 2 * K + EIN; p.node_row = K + tr::FE; p.K = K;
  p.w1 = w->w_c1; p.b1 = w->b_c1; p.w2 = w->w_c2; p.b2 = w->b_c2; p.b_node = w->b_node;
  p.e0 = e0; p.h0 = h0; p.n_steps = n_steps; p.agg = cfg->agg; p.n_nodes = n_nodes;
  bool any_cls = false;
  for (int t = 0; t < n_steps; ++t) {
    const gnncca_frame_bwd_step& st = host_steps[t];
    if (!st.e || !st.e_prev || !st.h_prev || !st.pa_edge || !st.pa_proj || !st.gh_prev) return GNNCCA_ERR_BAD_ARG;
    if (t + 1 < n_steps && (!st.pa_agg || (!st.R && !w->b_node))) return GNNCCA_ERR_BAD_ARG;
    if (st.g_logits) { if (!st.pa_cls) return GNNCCA_ERR_BAD_ARG; any_cls = true; }
    p.st[t] = st;
  }
  if (any_cls && (cfg->cls_hidden != tr::HC || !w->w_c1 || !w->b_c1 || !w->w_c2)) return GNNCCA_ERR_UNSUPPORTED;
  p.ge_buf[0] = ge_a; p.ge_buf[1] = ge_b; p.ge0_acc = ge0_acc; p.gh0_acc = gh0_acc;
  p.max_n = max_frame_nodes; p.max_e = (max_frame_edges + 3) & ~3;
  const int which = (cfg->reattach_edges ? 1 : 0) + (cfg->reattach_nodes ? 2 : 0);
  const void* fn = which == 0 ? (const void*)tr::frames_bwd_kernel<false, tr::FN> : which == 1 ? (const void*)tr::frames_bwd_kernel<true, tr::FN>
                 : which == 2 ? (const void*)tr::frames_bwd_kernel<false, 2 * tr::FN> : (const void*)tr::frames_bwd_kernel<true, 2 * tr::FN>;
  if (dyn_bytes > 48 * 1024) {       // opt in to the large dynamic shared memory (per device and function: set at every such launch)
    const cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, GNNCCA_FRAMES_BWD_SMEM_LIMIT);
    if (e != cudaSuccess) return check_cuda(e);
  }
  void* args[] = {(void*)&p};
  const cudaError_t le = cudaLaunchKernel(fn, dim3((unsigned)n_graphs), dim3(tr::TBF), args, (size_t)dyn_bytes, (cudaStream_t)stream);
  if (le != cudaSuccess) return check_cuda(le);
  return check_launch();
}

extern "C" int gnncca_edge_enc_bwd(const float* edge_attr, int32_t edge_in_dim, const float* e0, const float* ge0, int64_t n_edges,
                                   float* partial, void* stream) {
  if (edge_in_dim < 1 || edge_in_dim > tr::MAX_FA) return GNNCCA_ERR_UNSUPPORTED;
  if (n_edges == 0) return GNNCCA_OK;
  if (!edge_attr || !e0 || !ge0 || !partial) return GNNCCA_ERR_BAD_ARG;
  tr::edge_enc_bwd_kernel<<<blocks_for(n_edges, tr::TB), tr::TB, 0, (cudaStream_t)stream>>>(edge_attr, edge_in_dim, e0, ge0, n_edges, partial);
  return check_launch();
}

extern "C" int gnncca_linear_bwd(const float* x, const float* W, const float* y, const float* gy, int32_t relu, int64_t rows, int32_t in_dim,
                                 int32_t out_dim, float* gz, float* gx, float* gb_partial, float* gw_partial, int32_t n_splits,
                                 int32_t math, void* stream) {
  if (rows <= 0 || in_dim <= 0 || out_dim <= 0 || n_splits < 1) return GNNCCA_ERR_BAD_ARG;
  if (!x || !y || !gy || !gz || !gb_partial || !gw_partial || (gx && !W)) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  tr::relu_grad_bias_kernel<<<(unsigned)ceil_div(rows, tr::RG_ROWS), 256, 0, st>>>(y, gy, relu, rows, out_dim, gz, gb_partial);
  const int64_t rps = ceil_div(ceil_div(rows, n_splits), tr::ATB_ROWS) * tr::ATB_ROWS;
  dim3 grid((unsigned)ceil_div(in_dim, 64), (unsigned)ceil_div(out_dim, 64), (unsigned)n_splits);
  if (math == GNNCCA_MATH_TF32) tr::atb_kernel<1><<<grid, 256, 0, st>>>(gz, x, rows, out_dim, in_dim, rps, gw_partial);
  else tr::atb_kernel<3><<<grid, 256, 0, st>>>(gz, x, rows, out_dim, in_dim, rps, gw_partial);
  int n = 2;
  if (gx) { tr::gz_times_w_kernel<<<grid_for(rows * in_dim, 256), 256, 0, st>>>(gz, W, rows, out_dim, in_dim, gx); ++n; }
  return check_launch(n);
}

extern "C" int gnncca_finalize_grads(const gnncca_fin_seg* host_segs, int32_t n_segs, void* stream) {
  if (n_segs < 0 || n_segs > GNNCCA_MAX_FIN_SEGS || (n_segs > 0 && !host_segs)) return GNNCCA_ERR_BAD_ARG;
  if (n_segs == 0) return GNNCCA_OK;
  tr::FinTable t{};
  t.n = n_segs;
  int max_elems = 1;
  for (int i = 0; i < n_segs; ++i) {
    t.s[i] = host_segs[i];
    if (!t.s[i].partial || !t.s[i].dst || t.s[i].rows < 0 || t.s[i].cols < 0 || t.s[i].n_blocks < 0) return GNNCCA_ERR_BAD_ARG;
    // blocks along y so that the largest entry gets ~4 elements per thread (thread mode) or per warp (warp mode)
    const int elems = t.s[i].rows * t.s[i].cols * (t.s[i].n_blocks >= tr::FIN_WARP_MODE_BLOCKS ? 32 : 1);
    if (elems > max_elems) max_elems = elems;
  }
  int gy = (int)ceil_div(max_elems, 256 * 4);
  if (gy > 256) gy = 256;
  tr::finalize_kernel<<<dim3((unsigned)n_segs, (unsigned)gy), 256, 0, (cudaStream_t)stream>>>(t);
  return check_launch();
}
