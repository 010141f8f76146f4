// Dense pieces of the path: models/mlp.py:12-23 Linear[+BN eval][+ReLU] as an FP32 FFMA tile kernel
// (the exact-fp32 path; the tcgen05 TF32 path for the wide node-encoder layer lives in encoder_tc.cu),
// the per-node projections P/Q/R of the split edge/node update, and the small per-edge MLP chain.
#include "common.cuh"

namespace gnncca {

// ------------------------------------------------------------------------------------------------
// y[rows,out] = act(x[rows,in] W^T + b) ; 64x64 output tile, BK = 16, 256 threads x (4x4) accumulators
// ------------------------------------------------------------------------------------------------
constexpr int BM = 64, BN = 64, BK = 16;

struct LinearEpilogue {
  const float* b;
  const float* bn_w; const float* bn_b; const float* bn_m; const float* bn_v;
  float bn_eps;
  int relu;
};

__global__ void __launch_bounds__(256) linear_fp32_kernel(const float* __restrict__ x, const float* __restrict__ W,
                                                          float* __restrict__ y, int64_t rows, int in_dim, int out_dim,
                                                          LinearEpilogue ep) {
  __shared__ float xs[BK][BM + 4];
  __shared__ float ws[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;            // 16 x 16 thread grid, 4x4 outputs each
  const int64_t row0 = (int64_t)blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;
  float acc[4][4] = {};
  const int lr = tid >> 2;         // 0..63 : tile row loaded by this thread
  const int lk = (tid & 3) * 4;    // 0,4,8,12 : first k of its 4 consecutive values
  const bool vec_ok = ((in_dim & 3) == 0) && ((((uintptr_t)x) & 15) == 0) && ((((uintptr_t)W) & 15) == 0);
  for (int k0 = 0; k0 < in_dim; k0 += BK) {
    float xv[4], wv[4];
    const int64_t xr = row0 + lr;
    const int wr = col0 + lr;
    if (vec_ok && k0 + lk + 3 < in_dim) {
      float4 a = (xr < rows) ? *reinterpret_cast<const float4*>(x + xr * in_dim + k0 + lk) : make_float4(0, 0, 0, 0);
      float4 b = (wr < out_dim) ? *reinterpret_cast<const float4*>(W + (int64_t)wr * in_dim + k0 + lk) : make_float4(0, 0, 0, 0);
      xv[0] = a.x; xv[1] = a.y; xv[2] = a.z; xv[3] = a.w;
      wv[0] = b.x; wv[1] = b.y; wv[2] = b.z; wv[3] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k0 + lk + j;
        xv[j] = (xr < rows && k < in_dim) ? x[xr * in_dim + k] : 0.f;
        wv[j] = (wr < out_dim && k < in_dim) ? W[(int64_t)wr * in_dim + k] : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { xs[lk + j][lr] = xv[j]; ws[lk + j][lr] = wv[j]; }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&xs[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&ws[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = col0 + tx * 4 + j;
    if (c >= out_dim) continue;
    const float bias = ep.b ? ep.b[c] : 0.f;
    float scale = 1.f, shift = 0.f;
    if (ep.bn_w) {  // BatchNorm1d eval: (z - mean) / sqrt(var + eps) * gamma + beta
      scale = ep.bn_w[c] / sqrtf(ep.bn_v[c] + ep.bn_eps);
      shift = ep.bn_b[c] - ep.bn_m[c] * scale;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t r = row0 + ty * 4 + i;
      if (r >= rows) continue;
      float v = acc[i][j] + bias;
      if (ep.bn_w) v = v * scale + shift;
      if (ep.relu) v = fmaxf(v, 0.f);
      y[r * out_dim + c] = v;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// P/Q/R node projections.  One block = 64 nodes (16 for small graphs: more blocks than SMs); weights staged in shared
// memory as [k][out].
// ------------------------------------------------------------------------------------------------
template <int FN, int FE, int NODES>
__global__ void __launch_bounds__(256) node_project_kernel(const float* __restrict__ w_edge, const float* __restrict__ b_edge,
                                                           const float* __restrict__ w_node, const float* __restrict__ b_node,
                                                           const float* __restrict__ h0, const float* __restrict__ h, int64_t n_nodes,
                                                           int nf, int ef, float* __restrict__ P, float* __restrict__ Q,
                                                           float* __restrict__ R) {
  constexpr int OUT = 2 * FE + FN;
  const int K = nf * FN;                       // node input width (FN, or 2*FN when the encoder output is reattached)
  const int edge_in = 2 * K + ef * FE;         // row length of w_edge
  const int node_in = K + FE;                  // row length of w_node
  extern __shared__ float smem[];
  float* wt = smem;                            // [K][OUT]  (transposed: consecutive outputs -> consecutive banks)
  float* bias = wt + K * OUT;                  // [OUT]
  float* hs = bias + OUT;                      // [NODES][K + 1]
  for (int idx = threadIdx.x; idx < K * OUT; idx += blockDim.x) {
    const int k = idx / OUT, o = idx - k * OUT;
    float v;
    if (o < FE) v = w_edge[o * edge_in + k];                     // source block
    else if (o < 2 * FE) v = w_edge[(o - FE) * edge_in + K + k]; // target block
    else v = w_node[(o - 2 * FE) * node_in + k];                 // node block of the node model
    wt[idx] = v;
  }
  for (int o = threadIdx.x; o < OUT; o += blockDim.x)
    bias[o] = (o < FE) ? b_edge[o] : (o < 2 * FE ? 0.f : b_node[o - 2 * FE]);
  const int64_t n0 = (int64_t)blockIdx.x * NODES;
  for (int idx = threadIdx.x; idx < NODES * K; idx += blockDim.x) {
    const int v = idx / K, k = idx - v * K;
    const int64_t node = n0 + v;
    float val = 0.f;
    if (node < n_nodes) {
      if (nf == 2) val = (k < FN) ? h0[node * FN + k] : h[node * FN + (k - FN)];   // cat(initial, current), mpn.py:285
      else val = h[node * FN + k];
    }
    hs[v * (K + 1) + k] = val;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < NODES * OUT; idx += blockDim.x) {
    const int v = idx / OUT, o = idx - v * OUT;
    const int64_t node = n0 + v;
    if (node >= n_nodes) continue;
    float acc = bias[o];
    const float* hv = hs + v * (K + 1);
    for (int k = 0; k < K; ++k) acc = fmaf(hv[k], wt[k * OUT + o], acc);
    if (o < FE) P[node * 8 + o] = acc;
    else if (o < 2 * FE) Q[node * 8 + (o - FE)] = acc;
    else R[node * FN + (o - 2 * FE)] = acc;
  }
  // zero the padding lanes of P/Q so vector loads never see garbage
  if (FE < 8) {
    for (int idx = threadIdx.x; idx < NODES * (8 - FE); idx += blockDim.x) {
      const int v = idx / (8 - FE), o = FE + idx % (8 - FE);
      const int64_t node = n0 + v;
      if (node < n_nodes) { P[node * 8 + o] = 0.f; Q[node * 8 + o] = 0.f; }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Per-row small MLP chain (edge encoder 4 -> 6, stand-alone classifier): thread per row.
// ------------------------------------------------------------------------------------------------
constexpr int MLP_MAX_LAYERS = 4, MLP_MAX_WIDTH = 32;
struct SmallMlp {
  int n_layers;
  int dims[MLP_MAX_LAYERS + 1];
  const float* W[MLP_MAX_LAYERS];
  const float* b[MLP_MAX_LAYERS];
  int relu[MLP_MAX_LAYERS];
};

__global__ void __launch_bounds__(256) edge_mlp_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n_rows, SmallMlp m) {
  extern __shared__ float smem[];
  // stage all weights: layer l at offset off[l] : W (out*in) then b (out)
  int off[MLP_MAX_LAYERS + 1];
  off[0] = 0;
  for (int l = 0; l < m.n_layers; ++l) off[l + 1] = off[l] + m.dims[l + 1] * m.dims[l] + m.dims[l + 1];
  for (int l = 0; l < m.n_layers; ++l) {
    const int nw = m.dims[l + 1] * m.dims[l];
    for (int i = threadIdx.x; i < nw; i += blockDim.x) smem[off[l] + i] = m.W[l][i];
    for (int i = threadIdx.x; i < m.dims[l + 1]; i += blockDim.x) smem[off[l] + nw + i] = m.b[l] ? m.b[l][i] : 0.f;
  }
  __syncthreads();
  const int in0 = m.dims[0], outL = m.dims[m.n_layers];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    float a[MLP_MAX_WIDTH], c[MLP_MAX_WIDTH];
    for (int k = 0; k < in0; ++k) a[k] = x[r * in0 + k];
    for (int l = 0; l < m.n_layers; ++l) {
      const int din = m.dims[l], dout = m.dims[l + 1];
      const float* Wl = smem + off[l];
      const float* bl = Wl + dout * din;
      for (int o = 0; o < dout; ++o) {
        float acc = bl[o];
        for (int k = 0; k < din; ++k) acc = fmaf(a[k], Wl[o * din + k], acc);
        c[o] = m.relu[l] ? fmaxf(acc, 0.f) : acc;
      }
      for (int o = 0; o < dout; ++o) a[o] = c[o];
    }
    for (int o = 0; o < outL; ++o) y[r * outL + o] = a[o];
  }
}

// The shape every shipped configuration uses for the edge encoder (4 attributes -> 6 latents, one layer): one 16-byte
// streaming load and three 8-byte streaming stores per edge, weights in registers.  HBM-bound: 40 B/edge.
template <int IN, int OUT>
__global__ void __launch_bounds__(256) edge_linear_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n_rows,
                                                          const float* __restrict__ W, const float* __restrict__ b, int relu) {
  static_assert(IN == 4 && OUT % 2 == 0, "rows are read as one float4 and written as float2s");
  __shared__ float sw[OUT * IN + OUT];
  for (int i = threadIdx.x; i < OUT * IN; i += blockDim.x) sw[i] = W[i];
  for (int i = threadIdx.x; i < OUT; i += blockDim.x) sw[OUT * IN + i] = b ? b[i] : 0.f;
  __syncthreads();
  float w[OUT][IN], bias[OUT];
#pragma unroll
  for (int o = 0; o < OUT; ++o) {
    bias[o] = sw[OUT * IN + o];
#pragma unroll
    for (int k = 0; k < IN; ++k) w[o][k] = sw[o * IN + k];
  }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride) {
    const float4 v = ld_stream_f4(x + r * IN);
    float out[OUT];
#pragma unroll
    for (int o = 0; o < OUT; ++o) {
      float acc = bias[o];
      acc = fmaf(v.x, w[o][0], acc); acc = fmaf(v.y, w[o][1], acc); acc = fmaf(v.z, w[o][2], acc); acc = fmaf(v.w, w[o][3], acc);
      out[o] = relu ? fmaxf(acc, 0.f) : acc;
    }
#pragma unroll
    for (int o = 0; o < OUT; o += 2) st_stream_f2(y + r * OUT + o, make_float2(out[o], out[o + 1]));
  }
}

}  // namespace gnncca

using namespace gnncca;

int gnncca_node_encoder_tf32(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1, float* h,
                             int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* stream);   // encoder_tc.cu

extern "C" int gnncca_linear(const float* x, const float* W, const float* b, float* y, int64_t rows, int32_t in_dim,
                             int32_t out_dim, int32_t relu, const float* bn_weight, const float* bn_bias,
                             const float* bn_mean, const float* bn_var, float bn_eps, int32_t math, void* stream) {
  if (rows < 0 || in_dim <= 0 || out_dim <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!x || !W || !y) return GNNCCA_ERR_BAD_ARG;
  if (bn_weight && (!bn_bias || !bn_mean || !bn_var)) return GNNCCA_ERR_BAD_ARG;
  if (math == GNNCCA_MATH_TF32) return GNNCCA_ERR_UNSUPPORTED;   // the tensor-core path is gnncca_node_encoder (fused two layers)
  if (math != GNNCCA_MATH_FP32) return GNNCCA_ERR_BAD_ARG;
  LinearEpilogue ep{b, bn_weight, bn_bias, bn_mean, bn_var, bn_eps, relu};
  dim3 grid((unsigned)ceil_div(rows, BM), (unsigned)ceil_div(out_dim, BN));
  linear_fp32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, W, y, rows, in_dim, out_dim, ep);
  return check_launch();
}

extern "C" int gnncca_node_encoder(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h,
                                   int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, int32_t math, void* stream) {
  if (rows < 0 || in_dim <= 0 || hidden <= 0 || out_dim <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!x || !W1 || !b1 || !W2 || !b2 || !h) return GNNCCA_ERR_BAD_ARG;
  if (math != GNNCCA_MATH_TF32) return GNNCCA_ERR_UNSUPPORTED;   // FP32: two gnncca_linear calls
  return gnncca_node_encoder_tf32(x, W1, b1, W2, b2, nullptr, h, rows, in_dim, hidden, out_dim, stream);
}

extern "C" int gnncca_node_encoder_train(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1,
                                         float* h, int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* stream) {
  if (rows < 0 || in_dim <= 0 || hidden <= 0 || out_dim <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!x || !W1 || !b1 || !W2 || !b2 || !h || !h1) return GNNCCA_ERR_BAD_ARG;
  return gnncca_node_encoder_tf32(x, W1, b1, W2, b2, h1, h, rows, in_dim, hidden, out_dim, stream);
}

int gnncca_node_encoder_tf32_ws(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1, float* h,
                                int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* scratch, int64_t scratch_bytes, void* stream);

extern "C" int gnncca_node_encoder_ws(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1, float* h,
                                      int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* scratch, int64_t scratch_bytes,
                                      void* stream) {
  if (rows < 0 || in_dim <= 0 || hidden <= 0 || out_dim <= 0 || scratch_bytes < 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!x || !W1 || !b1 || !W2 || !b2 || !h) return GNNCCA_ERR_BAD_ARG;
  return gnncca_node_encoder_tf32_ws(x, W1, b1, W2, b2, h1, h, rows, in_dim, hidden, out_dim, scratch, scratch_bytes, stream);
}

extern "C" int gnncca_node_project(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* h0, const float* h,
                                   int64_t n_nodes, float* P, float* Q, float* R, void* stream) {
  if (!cfg || !w || n_nodes < 0) return GNNCCA_ERR_BAD_ARG;
  if (n_nodes == 0) return GNNCCA_OK;
  if (!h || !P || !Q || !R || !w->w_edge || !w->b_edge || !w->w_node || !w->b_node) return GNNCCA_ERR_BAD_ARG;
  if (cfg->reattach_nodes && !h0) return GNNCCA_ERR_BAD_ARG;
  if (cfg->f_node != 32 || cfg->f_edge != 6) return GNNCCA_ERR_UNSUPPORTED;
  const int nf = cfg->reattach_nodes ? 2 : 1, ef = cfg->reattach_edges ? 2 : 1;
  const int K = nf * 32, OUT = 2 * 6 + 32;
  const int nodes = n_nodes <= 64 * kNumSMs * 4 ? 16 : 64;      // small graphs: 4x the blocks, a quarter of the work each
  const size_t smem = sizeof(float) * (K * OUT + OUT + nodes * (K + 1));
  auto kern = nodes == 16 ? node_project_kernel<32, 6, 16> : node_project_kernel<32, 6, 64>;
  if (smem > 48 * 1024) {
    int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (rc) return rc;
  }
  kern<<<(unsigned)ceil_div(n_nodes, nodes), 256, smem, (cudaStream_t)stream>>>(w->w_edge, w->b_edge, w->w_node, w->b_node, h0, h,
                                                                                n_nodes, nf, ef, P, Q, R);
  return check_launch();
}

extern "C" int gnncca_edge_mlp(const float* x, float* y, int64_t n_rows, int32_t n_layers, const int32_t* host_dims,
                               const float* const* host_W, const float* const* host_b, const int32_t* host_relu, void* stream) {
  if (n_rows < 0 || n_layers < 1 || !host_dims || !host_W || !host_b || !host_relu) return GNNCCA_ERR_BAD_ARG;
  if (n_layers > MLP_MAX_LAYERS) return GNNCCA_ERR_UNSUPPORTED;
  if (n_rows == 0) return GNNCCA_OK;
  if (!x || !y) return GNNCCA_ERR_BAD_ARG;
  SmallMlp m{};
  m.n_layers = n_layers;
  size_t smem = 0;
  for (int l = 0; l <= n_layers; ++l) {
    if (host_dims[l] < 1 || host_dims[l] > MLP_MAX_WIDTH) return GNNCCA_ERR_UNSUPPORTED;
    m.dims[l] = host_dims[l];
  }
  for (int l = 0; l < n_layers; ++l) {
    if (!host_W[l]) return GNNCCA_ERR_BAD_ARG;
    m.W[l] = host_W[l]; m.b[l] = host_b[l]; m.relu[l] = host_relu[l];
    smem += sizeof(float) * (m.dims[l + 1] * m.dims[l] + m.dims[l + 1]);
  }
  if (n_layers == 1 && m.dims[0] == 4 && m.dims[1] == 6 && (((uintptr_t)x & 15) | ((uintptr_t)y & 7)) == 0) {
    edge_linear_kernel<4, 6><<<grid_for(n_rows, 256, 8), 256, 0, (cudaStream_t)stream>>>(x, y, n_rows, m.W[0], m.b[0], m.relu[0]);
    return check_launch();
  }
  edge_mlp_kernel<<<grid_for(n_rows, 256), 256, smem, (cudaStream_t)stream>>>(x, y, n_rows, m);
  return check_launch();
}
