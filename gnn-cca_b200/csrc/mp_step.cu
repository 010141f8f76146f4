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
// Mapping.  Edges are source-sorted.  A CTA of 256 threads owns tiles of 512 consecutive edges (persistent, tiles
// strided over the grid); edges travel as PAIRS (2p, 2p+1 of the global numbering) so that the arithmetic is packed
// FFMA2 with the weights as broadcast scalars.  Phase A is thread-per-pair (gather P/Q rows, edge update, classifier,
// e' written once and staged in shared memory as [pair][k][2]); the raw inputs of the NEXT tile are requested before
// phase B starts, so their latency hides behind it.  Phase B is lane-per-feature: warp w walks edges [64w, 64w+64) of
// the tile pair by pair (3 LDS.128 + 6 FFMA2 + 2 FMNMX + 1 FADD2 per pair) with the running aggregate of the current
// source segment in registers (even / odd edges separately, folded at the end).  Segment boundaries come from one
// ballot over the staged source ids, not from rowptr.  A warp stages and reduces the SAME 64 edges (a "chunk" of the global
// edge numbering), so after the weights are staged the warps of a CTA never synchronise with each other:
//   - a segment that lies inside one chunk is finished and written by its warp;
//   - a segment that crosses chunks leaves one piece per chunk in the workspace (plain stores), and a second small
//     kernel (warp per node) folds the pieces of each such node in chunk order; it also zero-fills the nodes without
//     out-edges.  (A CTA barrier per tile plus fence / counter hand-over between tiles cost 0.17 of 0.57 ms at 8 x 64.)
// No floating-point atomics anywhere: the summation order is a pure function of (E, rowptr), so results are
// bit-identical run to run (the reference's torch_scatter atomicAdd path is not).
#include "common.cuh"

namespace gnncca {

constexpr int TILE = 512;     // edges per tile
constexpr int THREADS = 256;  // one thread per edge pair
constexpr int NWARP = THREADS / 32;
constexpr int CHUNK = TILE / NWARP;   // edges per warp in phase B
constexpr int FN = 32;        // latent node width handled by lane-per-feature phase B
#ifndef GNNCCA_STEP_MIN_CTAS
#define GNNCCA_STEP_MIN_CTAS 3
#endif
constexpr int STEP_MIN_CTAS = GNNCCA_STEP_MIN_CTAS;   // resident CTAs per SM the register allocation is held to

struct StepParams {
  const int32_t* rowptr; const int32_t* src; const int32_t* dst;
  int64_t n_nodes, n_edges;
  const float* P; const float* Q; const float* R;
  const float* e0; const float* e_in; float* e_out; float* h_out; float* logits;
  gnncca_mp_weights w;
  int agg, nf;
  int num_tiles;
  float* chunk_part;     // [num_tiles * NWARP][2][FN]  slot 0: piece of a segment begun in an earlier chunk, slot 1: open tail
  int phase;             // position of local edge 0 inside a tile of the GLOBAL edge numbering (edge-partitioned graphs)
  const int32_t* guard;  // optional CSR flags (gnncca_csr_from_edge_index): the kernel is a no-op when any is set
  // ATTR variants: the encoded edge features e0 = ReLU(W_enc attr + b_enc) (mpn.py:116-126, 4 -> FE) are computed here from
  // the raw attributes (16 B/edge) instead of being written by one kernel and read back by this one (2 x 24 B/edge)
  const float* attr; const float* w_enc; const float* b_enc;
};


__device__ __forceinline__ void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }
template <bool IS_MAX> __device__ __forceinline__ float comb(float a, float b) { return IS_MAX ? fmaxf(a, b) : a + b; }
template <bool IS_MAX> __device__ __forceinline__ float2 comb2(float2 a, float2 b) {
  return IS_MAX ? make_float2(fmaxf(a.x, b.x), fmaxf(a.y, b.y)) : pk::fadd2(a, b);
}
// lanes 0..15 hold features (2 fl, 2 fl + 1) of the node's aggregate
__device__ __forceinline__ void finalize_node(const StepParams& p, int node, float2 v, int fl) {
  if (p.agg == GNNCCA_AGG_MEAN) {   // scatter_mean: sum / clamp(count, 1)
    const float cnt = (float)max(p.rowptr[node + 1] - p.rowptr[node], 1);
    v.x = v.x / cnt; v.y = v.y / cnt;
  }
  *reinterpret_cast<float2*>(p.h_out + (int64_t)node * FN + 2 * fl) = v;
}

constexpr int FA = 4;   // raw attribute width of the ATTR variants (one 16-byte load per edge)

template <int FE, bool REATT_E, bool ATTR>
struct RawPair {       // what phase A reads for one edge pair, requested one tile ahead
  int sa, sb, da, db;
  float2 ea[FE / 2], eb[FE / 2];                                  // latent edge features (not read in step 1 of an ATTR variant)
  float2 xa[(REATT_E && !ATTR) ? FE / 2 : 1], xb[(REATT_E && !ATTR) ? FE / 2 : 1];   // stored initial features
  float4 ta[ATTR ? 1 : 1], tb[ATTR ? 1 : 1];                      // raw attributes
};

template <int FE, bool REATT_E, bool ATTR>
__device__ __forceinline__ void load_raw(const StepParams& p, int64_t i0, RawPair<FE, REATT_E, ATTR>& r) {
  const bool va = i0 >= 0 && i0 < p.n_edges, vb = i0 + 1 >= 0 && i0 + 1 < p.n_edges;
  const bool latent = !ATTR || p.e_in != nullptr;
  r.sa = r.sb = r.da = r.db = -1;
  if (va) {
    r.sa = ld_stream_i32(p.src + i0); r.da = ld_stream_i32(p.dst + i0);
    if (latent) {
#pragma unroll
      for (int k = 0; k < FE / 2; ++k) r.ea[k] = ld_stream_f2(p.e_in + i0 * FE + 2 * k);
    }
    if (REATT_E && !ATTR) {
#pragma unroll
      for (int k = 0; k < FE / 2; ++k) r.xa[k] = ld_stream_f2(p.e0 + i0 * FE + 2 * k);
    }
    if (ATTR) r.ta[0] = ld_stream_f4(p.attr + i0 * FA);
  }
  if (vb) {
    r.sb = ld_stream_i32(p.src + i0 + 1); r.db = ld_stream_i32(p.dst + i0 + 1);
    if (latent) {
#pragma unroll
      for (int k = 0; k < FE / 2; ++k) r.eb[k] = ld_stream_f2(p.e_in + (i0 + 1) * FE + 2 * k);
    }
    if (REATT_E && !ATTR) {
#pragma unroll
      for (int k = 0; k < FE / 2; ++k) r.xb[k] = ld_stream_f2(p.e0 + (i0 + 1) * FE + 2 * k);
    }
    if (ATTR) r.tb[0] = ld_stream_f4(p.attr + (i0 + 1) * FA);
  }
}

// The same for a pair that lies wholly inside the edge range at an even edge index with 16-byte aligned arrays (every pair of an
// interior tile when the partition phase is even): the pair's 2 x FE floats are contiguous, so they travel as float4s and the two
// ids as one 8-byte load; the pairing of the registers (k, k+1 of one edge) is the same as in load_raw.
template <int FE, bool REATT_E, bool ATTR>
__device__ __forceinline__ void load_raw_fast(const StepParams& p, int64_t i0, RawPair<FE, REATT_E, ATTR>& r) {
  static_assert((2 * FE) % 4 == 0, "a pair's features are read as float4");
  const int2 s2 = ld_stream_i2(p.src + i0), d2 = ld_stream_i2(p.dst + i0);
  r.sa = s2.x; r.sb = s2.y; r.da = d2.x; r.db = d2.y;
  if (!ATTR || p.e_in != nullptr) {
    float t[2 * FE];
#pragma unroll
    for (int q = 0; q < 2 * FE / 4; ++q) {
      const float4 v = ld_stream_f4(p.e_in + i0 * FE + 4 * q);
      t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < FE / 2; ++k) { r.ea[k] = make_float2(t[2 * k], t[2 * k + 1]); r.eb[k] = make_float2(t[FE + 2 * k], t[FE + 2 * k + 1]); }
  }
  if (REATT_E && !ATTR) {
    float t[2 * FE];
#pragma unroll
    for (int q = 0; q < 2 * FE / 4; ++q) {
      const float4 v = ld_stream_f4(p.e0 + i0 * FE + 4 * q);
      t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < FE / 2; ++k) { r.xa[k] = make_float2(t[2 * k], t[2 * k + 1]); r.xb[k] = make_float2(t[FE + 2 * k], t[FE + 2 * k + 1]); }
  }
  if (ATTR) { r.ta[0] = ld_stream_f4(p.attr + i0 * FA); r.tb[0] = ld_stream_f4(p.attr + (i0 + 1) * FA); }
}

template <int FE, int HC, bool REATT_E, bool IS_MAX, bool ATTR>
__global__ void __launch_bounds__(THREADS, STEP_MIN_CTAS) mp_step_fwd_kernel(const StepParams p) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  constexpr int HCS = HC > 0 ? HC : 1;
  static_assert(FE % 2 == 0 && (FE * EIN) % 4 == 0, "weights are read as float4");
  __shared__ __align__(16) float s_wee[FE * EIN];
  __shared__ __align__(8) float s_wne[FE][FN];
  __shared__ __align__(16) float s_wenc[ATTR ? FE * FA : 4];
  __shared__ float s_benc[ATTR ? FE : 1];
  __shared__ __align__(16) float s_c1[HCS * FE];
  __shared__ float s_b1[HCS], s_c2[HCS], s_b2;
  __shared__ __align__(16) float s_e[TILE / 2][2 * FE];      // [pair][k][2]
  __shared__ __align__(8) int s_src[TILE];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int K = p.nf * FN;
  if (p.guard != nullptr && (p.guard[0] | p.guard[1]) != 0) return;   // CSR not valid (unsorted / out of range): uniform exit
  const bool classify = HC > 0 && p.logits != nullptr;
  {  // stage weights (read from the live parameters at every launch)
    const int edge_row = 2 * K + EIN, node_row = K + FE;
    for (int idx = tid; idx < FE * EIN; idx += THREADS) s_wee[idx] = p.w.w_edge[(idx / EIN) * edge_row + 2 * K + idx % EIN];
    for (int idx = tid; idx < FE * FN; idx += THREADS) s_wne[idx / FN][idx % FN] = p.w.w_node[(idx % FN) * node_row + K + idx / FN];
    if (ATTR) {
      for (int idx = tid; idx < FE * FA; idx += THREADS) s_wenc[idx] = p.w_enc[idx];
      for (int idx = tid; idx < FE; idx += THREADS) s_benc[idx] = p.b_enc[idx];
    }
    if (classify) {
      for (int c = tid; c < HC; c += THREADS) {
        float scale = 1.f, shift = 0.f;
        if (p.w.bn_weight) {  // BatchNorm1d (eval) folded into the hidden layer
          scale = p.w.bn_weight[c] / sqrtf(p.w.bn_var[c] + p.w.bn_eps);
          shift = p.w.bn_bias[c] - p.w.bn_mean[c] * scale;
        }
        for (int k = 0; k < FE; ++k) s_c1[c * FE + k] = p.w.w_c1[c * FE + k] * scale;
        s_b1[c] = p.w.b_c1[c] * scale + shift;
        s_c2[c] = p.w.w_c2[c];
      }
      if (tid == 0) s_b2 = p.w.b_c2[0];
    }
  }
  // interior tiles of an even partition phase read their pairs as 16-byte vectors (uniform over the launch)
  const bool vec_ok = (p.phase & 1) == 0 && (((uintptr_t)p.e_in | (uintptr_t)p.e0 | (uintptr_t)p.attr) & 15) == 0 &&
                      (((uintptr_t)p.src | (uintptr_t)p.dst) & 7) == 0;
  RawPair<FE, REATT_E, ATTR> cur;
  load_raw<FE, REATT_E, ATTR>(p, (int64_t)blockIdx.x * TILE - p.phase + 2 * tid, cur);
  __syncthreads();
  // phase B: the two half-warps take alternate edge pairs, a lane owns features (2 fl, 2 fl + 1)
  const int half = lane >> 4, fl = lane & 15;
  float2 w2[FE];
#pragma unroll
  for (int k = 0; k < FE; ++k) w2[k] = *reinterpret_cast<const float2*>(&s_wne[k][2 * fl]);

  for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
    const int64_t base = (int64_t)tile * TILE - p.phase;   // tiles follow the global numbering: same chunking, same bits, on any partition
    const int64_t i0 = base + 2 * tid;
    const bool va = cur.sa >= 0, vb = cur.sb >= 0;
    // A warp stages and then reduces the SAME 64 edges, so the warps never wait for each other; the sources of the edges
    // just outside the chunk (is its first / last piece part of a longer segment?) come from global memory.
    const int lo = (int)max((int64_t)0, -base), hi = (int)min((int64_t)TILE, p.n_edges - base);   // valid tile-local edge range
    const int c0 = warp * CHUNK;
    const int a = max(c0, lo), b = min(c0 + CHUNK, hi);
    int nb = -1;     // lane 0: source of the edge before the chunk, lane 1: of the edge after it (-1: none)
    float2 r_next = make_float2(0.f, 0.f);   // R row of the chunk's first segment, requested before phase A so that phase B does not wait for it
    if (p.h_out != nullptr && a < b) {
      if (lane == 0 && base + a > 0) nb = p.src[base + a - 1];
      if (lane == 1 && base + b < p.n_edges) nb = p.src[base + b];
      const int s_first = __shfl_sync(0xffffffffu, (a & 1) ? cur.sb : cur.sa, (a - c0) >> 1);
      r_next = *reinterpret_cast<const float2*>(p.R + (int64_t)s_first * FN + 2 * fl);
    }
    // ------------------------------ phase A: thread per edge pair ------------------------------
    {
      float2 ein[EIN];
      constexpr int OFF = REATT_E ? FE : 0;
      float2 e0p[ATTR ? FE : 1];
      if (ATTR) {     // e0 = ReLU(W_enc attr + b_enc) for both edges of the pair (the FMA order of the stand-alone encoder kernel)
        const float2 t0 = make_float2(cur.ta[0].x, cur.tb[0].x), t1 = make_float2(cur.ta[0].y, cur.tb[0].y);
        const float2 t2 = make_float2(cur.ta[0].z, cur.tb[0].z), t3 = make_float2(cur.ta[0].w, cur.tb[0].w);
#pragma unroll
        for (int o = 0; o < FE; ++o) {
          const float4 w4 = *reinterpret_cast<const float4*>(&s_wenc[o * FA]);
          float2 acc = pk::bc(s_benc[o]);
          acc = pk::ffma2(t0, pk::bc(w4.x), acc); acc = pk::ffma2(t1, pk::bc(w4.y), acc);
          acc = pk::ffma2(t2, pk::bc(w4.z), acc); acc = pk::ffma2(t3, pk::bc(w4.w), acc);
          e0p[o] = pk::relu2(acc);
        }
      }
      if (REATT_E) {  // cat(initial_edge_feats, latent_edge_feats), mpn.py:283
        if (ATTR) {
#pragma unroll
          for (int k = 0; k < FE; ++k) ein[k] = e0p[k];
        } else {
#pragma unroll
          for (int k = 0; k < FE / 2; ++k) {
            ein[2 * k] = make_float2(cur.xa[k].x, cur.xb[k].x); ein[2 * k + 1] = make_float2(cur.xa[k].y, cur.xb[k].y);
          }
        }
      }
      if (ATTR && p.e_in == nullptr) {   // step 1: the latent features are the encoded ones
#pragma unroll
        for (int k = 0; k < FE; ++k) ein[OFF + k] = e0p[k];
      } else {
#pragma unroll
        for (int k = 0; k < FE / 2; ++k) {
          ein[OFF + 2 * k] = make_float2(cur.ea[k].x, cur.eb[k].x); ein[OFF + 2 * k + 1] = make_float2(cur.ea[k].y, cur.eb[k].y);
        }
      }
      float2 en[8];
      {
        float4 pa0 = make_float4(0.f, 0.f, 0.f, 0.f), pa1 = pa0, qa0 = pa0, qa1 = pa0, pb0 = pa0, pb1 = pa0, qb0 = pa0, qb1 = pa0;
        if (va) {
          pa0 = *reinterpret_cast<const float4*>(p.P + (int64_t)cur.sa * 8); pa1 = *reinterpret_cast<const float4*>(p.P + (int64_t)cur.sa * 8 + 4);
          qa0 = *reinterpret_cast<const float4*>(p.Q + (int64_t)cur.da * 8); qa1 = *reinterpret_cast<const float4*>(p.Q + (int64_t)cur.da * 8 + 4);
        }
        if (vb) {
          pb0 = *reinterpret_cast<const float4*>(p.P + (int64_t)cur.sb * 8); pb1 = *reinterpret_cast<const float4*>(p.P + (int64_t)cur.sb * 8 + 4);
          qb0 = *reinterpret_cast<const float4*>(p.Q + (int64_t)cur.db * 8); qb1 = *reinterpret_cast<const float4*>(p.Q + (int64_t)cur.db * 8 + 4);
        }
        en[0] = make_float2(pa0.x + qa0.x, pb0.x + qb0.x); en[1] = make_float2(pa0.y + qa0.y, pb0.y + qb0.y);
        en[2] = make_float2(pa0.z + qa0.z, pb0.z + qb0.z); en[3] = make_float2(pa0.w + qa0.w, pb0.w + qb0.w);
        en[4] = make_float2(pa1.x + qa1.x, pb1.x + qb1.x); en[5] = make_float2(pa1.y + qa1.y, pb1.y + qb1.y);
        en[6] = make_float2(pa1.z + qa1.z, pb1.z + qb1.z); en[7] = make_float2(pa1.w + qa1.w, pb1.w + qb1.w);
      }
#pragma unroll
      for (int o = 0; o < FE; ++o) {
        float2 acc = en[o];
#pragma unroll
        for (int k = 0; k < EIN; k += 2) {     // FE * EIN is a multiple of 4 and EIN is even: (o * EIN + k) is even
          const float2 w2 = *reinterpret_cast<const float2*>(&s_wee[o * EIN + k]);
          acc = pk::ffma2(ein[k], pk::bc(w2.x), acc);
          acc = pk::ffma2(ein[k + 1], pk::bc(w2.y), acc);
        }
        en[o] = pk::relu2(acc);
      }
      if (va && p.e_out != nullptr) {       // NULL in the last step of an inference forward: nothing reads e^L (mpn.py:288-299)
#pragma unroll
        for (int k = 0; k < FE; k += 2) st_stream_f2(p.e_out + i0 * FE + k, make_float2(en[k].x, en[k + 1].x));
      }
      if (vb && p.e_out != nullptr) {
#pragma unroll
        for (int k = 0; k < FE; k += 2) st_stream_f2(p.e_out + (i0 + 1) * FE + k, make_float2(en[k].y, en[k + 1].y));
      }
#pragma unroll
      for (int k = 0; k < FE; k += 2)
        *reinterpret_cast<float4*>(&s_e[tid][2 * k]) = make_float4(en[k].x, en[k].y, en[k + 1].x, en[k + 1].y);
      *reinterpret_cast<int2*>(&s_src[2 * tid]) = make_int2(cur.sa, cur.sb);
      if (classify) {
        float2 logit = pk::bc(s_b2);
#pragma unroll
        for (int c = 0; c < HC; ++c) {
          float2 hcv = pk::bc(s_b1[c]);
#pragma unroll
          for (int k = 0; k < FE; k += 2) {
            const float2 w2 = *reinterpret_cast<const float2*>(&s_c1[c * FE + k]);
            hcv = pk::ffma2(en[k], pk::bc(w2.x), hcv);
            hcv = pk::ffma2(en[k + 1], pk::bc(w2.y), hcv);
          }
          logit = pk::ffma2(pk::relu2(hcv), pk::bc(s_c2[c]), logit);
        }
        if (va) st_stream_f1(p.logits + i0, logit.x);
        if (vb) st_stream_f1(p.logits + i0 + 1, logit.y);
      }
    }
    const int next_tile = tile + gridDim.x;
    if (next_tile < p.num_tiles) {
      const int64_t nbase = (int64_t)next_tile * TILE - p.phase;
      if (vec_ok && nbase >= 0 && nbase + TILE <= p.n_edges) load_raw_fast<FE, REATT_E, ATTR>(p, nbase + 2 * tid, cur);
      else load_raw<FE, REATT_E, ATTR>(p, nbase + 2 * tid, cur);
    }
    {  // the tile after that: pull its lines into L2 now (one 128-byte line per thread), DRAM latency off the critical path
      const int64_t far = (int64_t)next_tile + gridDim.x;
      if (far < p.num_tiles) {
        const int64_t fb = far * TILE - p.phase;        // > 0: far >= 2
        constexpr int LE = TILE * FE / 32, LI = TILE / 32, LA = TILE * FA / 32;      // lines of e_in (e0), of src / dst, of attr
        const char* ptr = nullptr; const char* end = nullptr;
        if (tid < LE) { if (p.e_in != nullptr) { ptr = (const char*)(p.e_in + fb * FE); end = (const char*)(p.e_in + p.n_edges * FE); } }
        else if (tid < LE + LI) { ptr = (const char*)(p.src + fb) - LE * 128; end = (const char*)(p.src + p.n_edges); }
        else if (tid < LE + 2 * LI) { ptr = (const char*)(p.dst + fb) - (LE + LI) * 128; end = (const char*)(p.dst + p.n_edges); }
        else if (ATTR && tid < LE + 2 * LI + LA) { ptr = (const char*)(p.attr + fb * FA) - (LE + 2 * LI) * 128; end = (const char*)(p.attr + p.n_edges * FA); }
        else if (REATT_E && !ATTR && tid < 2 * LE + 2 * LI) { ptr = (const char*)(p.e0 + fb * FE) - (LE + 2 * LI) * 128; end = (const char*)(p.e0 + p.n_edges * FE); }
        if (ptr != nullptr) { ptr += (size_t)tid * 128; if (ptr < end) prefetch_l2(ptr); }
      }
    }
    if (p.h_out == nullptr) continue;  // last step: the node update is dead (mpn.py:288-299); no shared state reused
    __syncwarp();
    // ------------------------------ phase B: lane per feature ------------------------------
    {
      if (a < b) {
        const int nb_prev = __shfl_sync(0xffffffffu, nb, 0), nb_next = __shfl_sync(0xffffffffu, nb, 1);
        const int ie = c0 + 2 * lane;
        const int2 my = *reinterpret_cast<const int2*>(&s_src[ie]);
        const int prev = ie - 1 >= a ? s_src[max(ie - 1, 0)] : nb_prev;
        const bool st_e = ie >= a && ie < b && (ie == a || my.x != prev);
        const bool st_o = ie + 1 >= a && ie + 1 < b && (ie + 1 == a || my.y != my.x);
        const unsigned bits = (st_e ? 1u : 0u) | (st_o ? 2u : 0u);
        const unsigned m_lo = __reduce_or_sync(0xffffffffu, lane < 16 ? bits << (2 * lane) : 0u);
        const unsigned m_hi = __reduce_or_sync(0xffffffffu, lane >= 16 ? bits << (2 * lane - 32) : 0u);
        unsigned long long mask = (unsigned long long)m_lo | ((unsigned long long)m_hi << 32);   // bit i: edge c0 + i opens a piece
        const bool first_cont = nb_prev == s_src[a];
        const bool last_cont = nb_next == s_src[b - 1];
        while (mask) {
          const int j0 = c0 + __ffsll((long long)mask) - 1;
          mask &= mask - 1;
          const int j1 = mask ? c0 + __ffsll((long long)mask) - 1 : b;
          const int s = s_src[j0];
          const float2 r = r_next;
          if (mask) r_next = *reinterpret_cast<const float2*>(p.R + (int64_t)s_src[j1] * FN + 2 * fl);   // next segment's row: in flight during this one
          float2 acc2 = pk::bc(IS_MAX ? -INFINITY : 0.f);
          int j = j0;
          if (j & 1) {                      // a piece that begins on an odd edge: that edge goes to half 0, ahead of its pairs
            if (half == 0) {
              const float* ev = &s_e[j >> 1][1];
              float2 m = r;
#pragma unroll
              for (int k = 0; k < FE; ++k) m = pk::ffma2(w2[k], pk::bc(ev[2 * k]), m);
              acc2 = comb2<IS_MAX>(acc2, pk::relu2(m));
            }
            ++j;
          }
          {   // whole pairs, alternately to the two halves; one pointer walks the staged rows (3 x LDS.128 per pair, immediate offsets)
            const int npairs = (j1 - j) >> 1;
            const int nq = (npairs - half + 1) >> 1;
            const float4* ev = reinterpret_cast<const float4*>(&s_e[(j >> 1) + half][0]);
#pragma unroll 4
            for (int q = 0; q < nq; ++q, ev += FE) {
              float2 ma = r, mb = r;
#pragma unroll
              for (int k = 0; k < FE / 2; ++k) {
                const float4 v = ev[k];
                ma = pk::ffma2(w2[2 * k], pk::bc(v.x), ma);
                mb = pk::ffma2(w2[2 * k], pk::bc(v.y), mb);
                ma = pk::ffma2(w2[2 * k + 1], pk::bc(v.z), ma);
                mb = pk::ffma2(w2[2 * k + 1], pk::bc(v.w), mb);
              }
              acc2 = comb2<IS_MAX>(acc2, pk::relu2(ma));
              acc2 = comb2<IS_MAX>(acc2, pk::relu2(mb));
            }
            j += 2 * npairs;
          }
          if (j < j1 && half == 1) {         // a piece that ends on an even edge: that edge goes to half 1, after its pairs
            const float* ev = &s_e[j >> 1][0];
            float2 m = r;
#pragma unroll
            for (int k = 0; k < FE; ++k) m = pk::ffma2(w2[k], pk::bc(ev[2 * k]), m);
            acc2 = comb2<IS_MAX>(acc2, pk::relu2(m));
          }
          float2 other;
          other.x = __shfl_xor_sync(0xffffffffu, acc2.x, 16); other.y = __shfl_xor_sync(0xffffffffu, acc2.y, 16);
          const float2 acc = half == 0 ? comb2<IS_MAX>(acc2, other) : comb2<IS_MAX>(other, acc2);     // half 0's part first, on every lane
          const bool starts = !(j0 == a && first_cont), ends = !(j1 == b && last_cont);
          if (half == 0) {
            if (starts && ends) finalize_node(p, s, acc, fl);
            else *reinterpret_cast<float2*>(p.chunk_part + (((int64_t)tile * NWARP + warp) * 2 + (starts ? 1 : 0)) * FN + 2 * fl) = acc;   // folded by mp_step_fold_kernel
          }
        }
      }
    }
    __syncwarp();      // the warp's rows of s_e / s_src are rewritten by its next phase A
  }
}

// Second kernel of a step: one warp per node.  Nodes without out-edges aggregate to 0 (torch_scatter semantics); a node whose
// segment crosses chunk boundaries gets the fold of its pieces in chunk order (first chunk's open tail, then one piece per
// later chunk); everything else was written by the step kernel.
__global__ void __launch_bounds__(256) mp_step_fold_kernel(const StepParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t v = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (v >= p.n_nodes) return;
  if (p.guard != nullptr && (p.guard[0] | p.guard[1]) != 0) return;
  const int rs = p.rowptr[v], re = p.rowptr[v + 1];
  if (re <= rs) { p.h_out[v * FN + lane] = 0.f; return; }
  const int64_t c0 = ((int64_t)rs + p.phase) / CHUNK, c1 = ((int64_t)re - 1 + p.phase) / CHUNK;
  if (c0 == c1) return;
  const bool is_max = p.agg == GNNCCA_AGG_MAX;
  // sums are carried in double across the pieces (a handful of adds per node): the rounding of a long segment is then that
  // of its 64-edge chunks alone
  double tot = p.chunk_part[(c0 * 2 + 1) * FN + lane];
  int64_t c = c0 + 1;
  for (; c + 3 <= c1; c += 4) {        // four loads in flight, folded in order
    const float v0 = p.chunk_part[(c * 2) * FN + lane], v1 = p.chunk_part[((c + 1) * 2) * FN + lane];
    const float v2 = p.chunk_part[((c + 2) * 2) * FN + lane], v3 = p.chunk_part[((c + 3) * 2) * FN + lane];
    if (is_max) tot = fmax(fmax(fmax(fmax(tot, (double)v0), (double)v1), (double)v2), (double)v3);
    else tot = (((tot + v0) + v1) + v2) + v3;
  }
  for (; c <= c1; ++c) {
    const float v0 = p.chunk_part[(c * 2) * FN + lane];
    tot = is_max ? fmax(tot, (double)v0) : tot + v0;
  }
  float out = (float)tot;
  if (p.agg == GNNCCA_AGG_MEAN) out = out / (float)(re - rs);      // scatter_mean: sum / count, in fp32 like finalize_node
  p.h_out[v * FN + lane] = out;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace gnncca

using namespace gnncca;

extern "C" int64_t gnncca_mp_step_workspace_bytes(const gnncca_mp_cfg* cfg, int64_t n_nodes, int64_t n_edges) {
  (void)cfg; (void)n_nodes;
  const int64_t tiles = ceil_div((n_edges > 0 ? n_edges : 1) + TILE, TILE);   // + one tile: the partition phase shifts the tiling
  return (int64_t)(align256(sizeof(float) * tiles * NWARP * 2 * FN) + 256);
}

template <int FE, int HC>
static int launch_step(const StepParams& sp, bool reatt_e, cudaStream_t st) {
  const int grid = (int)min((int64_t)sp.num_tiles, (int64_t)kNumSMs * STEP_MIN_CTAS);
  const bool is_max = sp.agg == GNNCCA_AGG_MAX;
  const bool attr = sp.attr != nullptr;
#define GNNCCA_STEP(R, M, A) mp_step_fwd_kernel<FE, HC, R, M, A><<<grid, THREADS, 0, st>>>(sp)
  if (reatt_e) {
    if (is_max) { if (attr) GNNCCA_STEP(true, true, true); else GNNCCA_STEP(true, true, false); }
    else { if (attr) GNNCCA_STEP(true, false, true); else GNNCCA_STEP(true, false, false); }
  } else {
    if (is_max) { if (attr) GNNCCA_STEP(false, true, true); else GNNCCA_STEP(false, true, false); }
    else { if (attr) GNNCCA_STEP(false, false, true); else GNNCCA_STEP(false, false, false); }
  }
#undef GNNCCA_STEP
  if (sp.h_out == nullptr) return check_launch();
  mp_step_fold_kernel<<<(unsigned)ceil_div(sp.n_nodes, (int64_t)8), 256, 0, st>>>(sp);
  return check_launch(2);
}

extern "C" int32_t gnncca_mp_step_tile(void) { return TILE; }

static int mp_step_impl(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr,
                        const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P,
                        const float* Q, const float* R, const float* e0, const float* e_in, float* e_out,
                        float* h_out, float* logits, const int32_t* guard_flags, void* workspace, int64_t workspace_bytes, void* stream,
                        const float* attr, const float* w_enc, const float* b_enc) {
  if (!cfg || !w || n_nodes < 0 || n_edges < 0 || n_edges >= INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (n_edges == 0) {          // nodes without out-edges aggregate to 0 (torch_scatter semantics)
    if (h_out && n_nodes > 0) return check_cuda(cudaMemsetAsync(h_out, 0, sizeof(float) * n_nodes * FN, st));
    return GNNCCA_OK;
  }
  if (!rowptr || !src || !dst || !P || !Q || !R || !w->w_edge || !w->w_node) return GNNCCA_ERR_BAD_ARG;
  if (!e_out && (h_out || !logits)) return GNNCCA_ERR_BAD_ARG;      // e_out may only be dropped when the step just classifies
  if (attr ? (!w_enc || !b_enc || (((uintptr_t)attr) & 15) != 0) : (!e_in || (cfg->reattach_edges && !e0))) return GNNCCA_ERR_BAD_ARG;
  if (logits && cfg->cls_hidden > 0 && (!w->w_c1 || !w->b_c1 || !w->w_c2 || !w->b_c2)) return GNNCCA_ERR_BAD_ARG;
  if (cfg->agg < GNNCCA_AGG_SUM || cfg->agg > GNNCCA_AGG_MAX) return GNNCCA_ERR_BAD_ARG;
  if (cfg->f_node != FN) return GNNCCA_ERR_UNSUPPORTED;
  if (((uintptr_t)e_in | (uintptr_t)e_out | (uintptr_t)e0) & 7) return GNNCCA_ERR_BAD_ARG;
  if (((uintptr_t)P | (uintptr_t)Q) & 15) return GNNCCA_ERR_BAD_ARG;
  if (workspace_bytes < gnncca_mp_step_workspace_bytes(cfg, n_nodes, n_edges) || !workspace) return GNNCCA_ERR_WORKSPACE;

  StepParams sp{};
  sp.rowptr = rowptr; sp.src = src; sp.dst = dst; sp.n_nodes = n_nodes; sp.n_edges = n_edges;
  sp.P = P; sp.Q = Q; sp.R = R; sp.e0 = e0; sp.e_in = e_in; sp.e_out = e_out; sp.h_out = h_out; sp.logits = logits;
  sp.guard = guard_flags;
  sp.attr = attr; sp.w_enc = w_enc; sp.b_enc = b_enc;
  sp.w = *w; sp.agg = cfg->agg; sp.nf = cfg->reattach_nodes ? 2 : 1;
  if (cfg->edge_phase < 0 || cfg->edge_phase >= TILE) return GNNCCA_ERR_BAD_ARG;
  sp.phase = cfg->edge_phase;
  sp.num_tiles = (int)ceil_div(n_edges + sp.phase, TILE);
  char* ws = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  sp.chunk_part = (float*)ws;
  const bool re = cfg->reattach_edges != 0;
  if (cfg->f_edge == 6 && cfg->cls_hidden == 4) return launch_step<6, 4>(sp, re, st);
  return GNNCCA_ERR_UNSUPPORTED;
}

extern "C" int gnncca_mp_step_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr,
                                  const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P,
                                  const float* Q, const float* R, const float* e0, const float* e_in, float* e_out,
                                  float* h_out, float* logits, const int32_t* guard_flags, void* workspace, int64_t workspace_bytes, void* stream) {
  return mp_step_impl(cfg, w, rowptr, src, dst, n_nodes, n_edges, P, Q, R, e0, e_in, e_out, h_out, logits, guard_flags, workspace, workspace_bytes,
                      stream, nullptr, nullptr, nullptr);
}

extern "C" int gnncca_mp_step_fwd_attr(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr,
                                       const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P,
                                       const float* Q, const float* R, const float* edge_attr, int32_t edge_in_dim, const float* w_enc_edge,
                                       const float* b_enc_edge, const float* e_in, float* e_out, float* h_out, float* logits,
                                       const int32_t* guard_flags, void* workspace, int64_t workspace_bytes, void* stream) {
  if (edge_in_dim != FA) return GNNCCA_ERR_UNSUPPORTED;
  if (!edge_attr && n_edges > 0) return GNNCCA_ERR_BAD_ARG;
  return mp_step_impl(cfg, w, rowptr, src, dst, n_nodes, n_edges, P, Q, R, nullptr, e_in, e_out, h_out, logits, guard_flags, workspace,
                      workspace_bytes, stream, edge_attr, w_enc_edge, b_enc_edge);
}
