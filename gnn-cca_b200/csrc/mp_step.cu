// One fused message-passing step (reference models/mpn.py:32-54 MetaLayer = EdgeModel :59-69 +
// NodeModel :71-101, plus the classifier call of MOTMPNet.forward :290-293), forward.
//
// Algebra.  Linear over a concatenation is a sum of per-block products, so with
//   P[v] = W_e[:, src block] h[v] + b_e,  Q[v] = W_e[:, dst block] h[v],  R[v] = W_n[:, node block] h[v] + b_n
// (gnncca_node_project, once per NODE) the per-EDGE work is
//   e'[i] = relu(P[src] + Q[dst] + W_e[:, edge block] e[i])          (FE x FE  FMAs)
//   m[i]  = relu(R[src] + W_n[:, edge block] e'[i])                  (FN x FE  FMAs)
//   h'[v] = agg_{i: src[i]==v} m[i]
//   logit[i] = w_c2 . relu(W_c1 e'[i] + b_c1) + b_c2
// and nothing of width 70 / 38 is ever materialised: 64 B/edge of HBM traffic instead of ~2.3 KB.
//
// Mapping.  Edges are source-sorted.  A CTA owns tiles of 256 consecutive edges (persistent, tiles
// strided over the grid).  Phase A is thread-per-edge (gather, edge update, classifier, write e' once,
// stage e' in shared memory).  Phase B is lane-per-feature: warp w walks edges [32w, 32w+32) of the tile
// and keeps the running aggregate of the current source segment in a register.  Segments are closed
//   - inside a warp chunk: written directly;
//   - across warps of the tile: chained in warp order through shared memory;
//   - across tiles: each tile publishes its partial, bumps an integer counter keyed by the segment's first
//     tile, and the LAST tile to arrive folds the partials in tile order.
// No floating-point atomics anywhere: the summation order is a pure function of (E, rowptr), so results are
// bit-identical run to run (the reference's torch_scatter atomicAdd path is not).
#include "common.cuh"

namespace gnncca {

constexpr int TILE = 256;     // edges per tile == threads per CTA
constexpr int NWARP = TILE / 32;
constexpr int FN = 32;        // latent node width handled by lane-per-feature phase B

struct StepParams {
  const int32_t* rowptr; const int32_t* src; const int32_t* dst;
  int64_t n_nodes, n_edges;
  const float* P; const float* Q; const float* R;
  const float* e0; const float* e_in; float* e_out; float* h_out; float* logits;
  gnncca_mp_weights w;
  int agg, nf;
  int num_tiles;
  float* tile_part;      // [num_tiles][2][FN]  slot 0: continuation piece (started in an earlier tile), slot 1: open tail
  int32_t* counters;     // [num_tiles] arrival counters, zeroed by the host entry before the launch
  int phase;             // position of local edge 0 inside a 256-edge tile of the GLOBAL edge numbering (edge-partitioned graphs)
  const int32_t* guard;  // optional CSR flags (gnncca_csr_from_edge_index): the kernel is a no-op when any is set
};

__device__ __forceinline__ float agg_init(int agg) { return agg == GNNCCA_AGG_MAX ? -INFINITY : 0.f; }
__device__ __forceinline__ float agg_comb(int agg, float a, float b) { return agg == GNNCCA_AGG_MAX ? fmaxf(a, b) : a + b; }
__device__ __forceinline__ void finalize_node(const StepParams& p, int node, float v, int deg, int lane) {
  if (p.agg == GNNCCA_AGG_MEAN) v = v / (float)max(deg, 1);   // scatter_mean: sum / clamp(count, 1)
  p.h_out[(int64_t)node * FN + lane] = v;
}

// Cross-tile combination.  kind 0 = this tile continues a segment begun earlier, kind 1 = this tile opens one.
__device__ __forceinline__ void publish_partial(const StepParams& p, int tile, int kind, int node, float total, int lane) {
  const int rs = p.rowptr[node], re = p.rowptr[node + 1];
  const int t0 = (rs + p.phase) / TILE, t1 = (re - 1 + p.phase) / TILE;
  p.tile_part[((int64_t)tile * 2 + kind) * FN + lane] = total;
  __threadfence();
  __syncwarp();
  int old = 0;
  if (lane == 0) old = atomicAdd(p.counters + t0, 1);
  old = __shfl_sync(0xffffffffu, old, 0);
  if (old == t1 - t0) {  // every tile of [t0, t1] has published: fold in tile order
    __threadfence();
    float tot = ld_cg_f(p.tile_part + ((int64_t)t0 * 2 + 1) * FN + lane);
    for (int t = t0 + 1; t <= t1; ++t) tot = agg_comb(p.agg, tot, ld_cg_f(p.tile_part + ((int64_t)t * 2) * FN + lane));
    finalize_node(p, node, tot, re - rs, lane);
  }
}

template <int FE, int HC, bool REATT_E>
__global__ void __launch_bounds__(TILE, 3) mp_step_fwd_kernel(const StepParams p) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  constexpr int HCS = HC > 0 ? HC : 1;
  __shared__ float s_wee[FE][EIN];
  __shared__ float s_wne[FE][FN];
  __shared__ float s_c1[HCS][FE], s_b1[HCS], s_c2[HCS], s_b2;
  __shared__ __align__(16) float s_e[TILE][FE];
  __shared__ int s_src[TILE];
  __shared__ float s_C[NWARP][FN], s_O[NWARP][FN];
  __shared__ int s_Cnode[NWARP], s_Cclosed[NWARP], s_Onode[NWARP];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int K = p.nf * FN;
  if (p.guard != nullptr && (p.guard[0] | p.guard[1]) != 0) return;   // CSR not valid (unsorted / out of range): uniform exit
  {  // stage weights (read from the live parameters at every launch)
    const int edge_row = 2 * K + EIN, node_row = K + FE;
    for (int idx = tid; idx < FE * EIN; idx += TILE) s_wee[idx / EIN][idx % EIN] = p.w.w_edge[(idx / EIN) * edge_row + 2 * K + idx % EIN];
    for (int idx = tid; idx < FE * FN; idx += TILE) s_wne[idx / FN][idx % FN] = p.w.w_node[(idx % FN) * node_row + K + idx / FN];
    if (HC > 0 && p.logits != nullptr) {
      for (int c = tid; c < HC; c += TILE) {
        float scale = 1.f, shift = 0.f;
        if (p.w.bn_weight) {  // BatchNorm1d (eval) folded into the hidden layer
          scale = p.w.bn_weight[c] / sqrtf(p.w.bn_var[c] + p.w.bn_eps);
          shift = p.w.bn_bias[c] - p.w.bn_mean[c] * scale;
        }
        for (int k = 0; k < FE; ++k) s_c1[c][k] = p.w.w_c1[c * FE + k] * scale;
        s_b1[c] = p.w.b_c1[c] * scale + shift;
        s_c2[c] = p.w.w_c2[c];
      }
      if (tid == 0) s_b2 = p.w.b_c2[0];
    }
  }
  __syncthreads();
  float wne[FE];
#pragma unroll
  for (int k = 0; k < FE; ++k) wne[k] = s_wne[k][lane];

  for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
    const int64_t base = (int64_t)tile * TILE - p.phase;   // tiles follow the global numbering: same chunking, same bits, on any partition
    const int64_t i = base + tid;
    // ------------------------------ phase A: thread per edge ------------------------------
    if (i >= 0 && i < p.n_edges) {
      const int s = ld_stream_i32(p.src + i), d = ld_stream_i32(p.dst + i);
      float ein[EIN];
      if (REATT_E) {  // cat(initial_edge_feats, latent_edge_feats), mpn.py:283
#pragma unroll
        for (int k = 0; k < FE; k += 2) { float2 v = ld_stream_f2(p.e0 + i * FE + k); ein[k] = v.x; ein[k + 1] = v.y; }
      }
      constexpr int OFF = REATT_E ? FE : 0;
#pragma unroll
      for (int k = 0; k < FE; k += 2) { float2 v = ld_stream_f2(p.e_in + i * FE + k); ein[OFF + k] = v.x; ein[OFF + k + 1] = v.y; }
      const float4 p0 = *reinterpret_cast<const float4*>(p.P + (int64_t)s * 8), p1 = *reinterpret_cast<const float4*>(p.P + (int64_t)s * 8 + 4);
      const float4 q0 = *reinterpret_cast<const float4*>(p.Q + (int64_t)d * 8), q1 = *reinterpret_cast<const float4*>(p.Q + (int64_t)d * 8 + 4);
      const float pq[8] = {p0.x + q0.x, p0.y + q0.y, p0.z + q0.z, p0.w + q0.w, p1.x + q1.x, p1.y + q1.y, p1.z + q1.z, p1.w + q1.w};
      float en[FE];
#pragma unroll
      for (int o = 0; o < FE; ++o) {
        float acc = pq[o];
#pragma unroll
        for (int k = 0; k < EIN; ++k) acc = fmaf(ein[k], s_wee[o][k], acc);
        en[o] = fmaxf(acc, 0.f);
      }
#pragma unroll
      for (int k = 0; k < FE; k += 2) {
        st_stream_f2(p.e_out + i * FE + k, make_float2(en[k], en[k + 1]));
        *reinterpret_cast<float2*>(&s_e[tid][k]) = make_float2(en[k], en[k + 1]);
      }
      s_src[tid] = s;
      if (HC > 0 && p.logits != nullptr) {
        float logit = s_b2;
#pragma unroll
        for (int c = 0; c < HC; ++c) {
          float hcv = s_b1[c];
#pragma unroll
          for (int k = 0; k < FE; ++k) hcv = fmaf(en[k], s_c1[c][k], hcv);
          logit = fmaf(fmaxf(hcv, 0.f), s_c2[c], logit);
        }
        st_stream_f1(p.logits + i, logit);
      }
    } else {
      s_src[tid] = -1;
    }
    if (p.h_out == nullptr) continue;  // last step: the node update is dead (mpn.py:288-299); no shared state reused
    __syncthreads();
    // ------------------------------ phase B: lane per feature ------------------------------
    {
      const int64_t tile_end = min(base + TILE, p.n_edges);
      const int64_t c0 = base + warp * 32;
      const int64_t c1 = min(c0 + 32, tile_end);
      int Cnode = -1, Cclosed = 0, Onode = -1;
      int64_t j0 = max(c0, (int64_t)0);
      while (j0 < c1) {
        const int s = s_src[j0 - base];
        const int rs = p.rowptr[s], re = p.rowptr[s + 1];
        const int64_t pe = min((int64_t)re, c1);
        const float r = p.R[(int64_t)s * FN + lane];
        float acc = agg_init(p.agg);
        for (int64_t j = j0; j < pe; ++j) {
          const float* ev = s_e[j - base];
          float m = r;
#pragma unroll
          for (int k = 0; k < FE; ++k) m = fmaf(ev[k], wne[k], m);
          acc = agg_comb(p.agg, acc, fmaxf(m, 0.f));
        }
        const bool starts = rs >= c0, ends = re <= c1;
        if (starts && ends) finalize_node(p, s, acc, re - rs, lane);
        else if (!starts) { s_C[warp][lane] = acc; Cnode = s; Cclosed = ends ? 1 : 0; }
        else { s_O[warp][lane] = acc; Onode = s; }
        j0 = pe;
      }
      if (lane == 0) { s_Cnode[warp] = Cnode; s_Cclosed[warp] = Cclosed; s_Onode[warp] = Onode; }
    }
    __syncthreads();
    // chain the pieces of segments that cross warp chunks, in warp order
    if (s_Onode[warp] >= 0) {
      const int s = s_Onode[warp];
      float total = s_O[warp][lane];
      bool closed = false;
      for (int w2 = warp + 1; w2 < NWARP; ++w2) {
        if (s_Cnode[w2] != s) break;
        total = agg_comb(p.agg, total, s_C[w2][lane]);
        if (s_Cclosed[w2]) { closed = true; break; }
      }
      if (closed) finalize_node(p, s, total, p.rowptr[s + 1] - p.rowptr[s], lane);
      else publish_partial(p, tile, 1, s, total, lane);
    }
    if (warp == 0 && s_Cnode[0] >= 0) {  // the tile begins inside a segment opened by an earlier tile
      const int s = s_Cnode[0];
      float total = s_C[0][lane];
      if (!s_Cclosed[0]) {
        for (int w2 = 1; w2 < NWARP; ++w2) {
          if (s_Cnode[w2] != s) break;
          total = agg_comb(p.agg, total, s_C[w2][lane]);
          if (s_Cclosed[w2]) break;
        }
      }
      publish_partial(p, tile, 0, s, total, lane);
    }
    __syncthreads();  // s_e / s_C / s_O are rewritten by the next tile
  }
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace gnncca

using namespace gnncca;

extern "C" int64_t gnncca_mp_step_workspace_bytes(const gnncca_mp_cfg* cfg, int64_t n_nodes, int64_t n_edges) {
  (void)cfg; (void)n_nodes;
  const int64_t tiles = ceil_div((n_edges > 0 ? n_edges : 1) + TILE, TILE);   // + one tile: the partition phase shifts the tiling
  return (int64_t)(align256(sizeof(int32_t) * tiles) + align256(sizeof(float) * tiles * 2 * FN) + 256);
}

template <int FE, int HC>
static int launch_step(const StepParams& sp, bool reatt_e, cudaStream_t st) {
  const int grid = (int)min((int64_t)sp.num_tiles, (int64_t)kNumSMs * 3);
  if (reatt_e) mp_step_fwd_kernel<FE, HC, true><<<grid, TILE, 0, st>>>(sp);
  else mp_step_fwd_kernel<FE, HC, false><<<grid, TILE, 0, st>>>(sp);
  return check_launch();
}

extern "C" int gnncca_mp_step_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr,
                                  const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P,
                                  const float* Q, const float* R, const float* e0, const float* e_in, float* e_out,
                                  float* h_out, float* logits, const int32_t* guard_flags, void* workspace, int64_t workspace_bytes, void* stream) {
  if (!cfg || !w || n_nodes < 0 || n_edges < 0 || n_edges >= INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (h_out && n_nodes > 0) {  // nodes without out-edges aggregate to 0 (torch_scatter semantics)
    int rc = check_cuda(cudaMemsetAsync(h_out, 0, sizeof(float) * n_nodes * FN, st));
    if (rc) return rc;
  }
  if (n_edges == 0) return GNNCCA_OK;
  if (!rowptr || !src || !dst || !P || !Q || !R || !e_in || !e_out || !w->w_edge || !w->w_node) return GNNCCA_ERR_BAD_ARG;
  if (cfg->reattach_edges && !e0) return GNNCCA_ERR_BAD_ARG;
  if (logits && cfg->cls_hidden > 0 && (!w->w_c1 || !w->b_c1 || !w->w_c2 || !w->b_c2)) return GNNCCA_ERR_BAD_ARG;
  if (cfg->agg < GNNCCA_AGG_SUM || cfg->agg > GNNCCA_AGG_MAX) return GNNCCA_ERR_BAD_ARG;
  if (cfg->f_node != FN) return GNNCCA_ERR_UNSUPPORTED;
  if (((uintptr_t)e_in | (uintptr_t)e_out | (uintptr_t)(e0 ? e0 : e_in)) & 7) return GNNCCA_ERR_BAD_ARG;
  if (((uintptr_t)P | (uintptr_t)Q) & 15) return GNNCCA_ERR_BAD_ARG;
  if (workspace_bytes < gnncca_mp_step_workspace_bytes(cfg, n_nodes, n_edges) || !workspace) return GNNCCA_ERR_WORKSPACE;

  StepParams sp{};
  sp.rowptr = rowptr; sp.src = src; sp.dst = dst; sp.n_nodes = n_nodes; sp.n_edges = n_edges;
  sp.P = P; sp.Q = Q; sp.R = R; sp.e0 = e0; sp.e_in = e_in; sp.e_out = e_out; sp.h_out = h_out; sp.logits = logits;
  sp.guard = guard_flags;
  sp.w = *w; sp.agg = cfg->agg; sp.nf = cfg->reattach_nodes ? 2 : 1;
  if (cfg->edge_phase < 0 || cfg->edge_phase >= TILE) return GNNCCA_ERR_BAD_ARG;
  sp.phase = cfg->edge_phase;
  sp.num_tiles = (int)ceil_div(n_edges + sp.phase, TILE);
  char* ws = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  sp.counters = (int32_t*)ws;
  sp.tile_part = (float*)(ws + align256(sizeof(int32_t) * sp.num_tiles));
  if (h_out) {
    int rc = check_cuda(cudaMemsetAsync(sp.counters, 0, sizeof(int32_t) * sp.num_tiles, st));
    if (rc) return rc;
  }
  const bool re = cfg->reattach_edges != 0;
  if (cfg->f_edge == 6 && cfg->cls_hidden == 4) return launch_step<6, 4>(sp, re, st);
  return GNNCCA_ERR_UNSUPPORTED;
}
