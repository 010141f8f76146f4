// All message-passing steps of MOTMPNet.forward (reference models/mpn.py:270-299) for batches of SMALL graphs,
// as ONE kernel: a CTA takes one frame graph at a time, keeps its node and edge latents in shared memory across
// the L steps (edge encoder mlp.py / mpn.py:128-142, EdgeModel :59-69, NodeModel :71-101 with the aggregation of
// :192-202, classifier :290-293) and touches HBM only for the inputs (encoded nodes, edge_attr, CSR) and the
// per-step logits.  EPFL frames are 2..40 detections / <= 1200 directed edges (~30 KB of state), so nothing
// of size E is ever written between steps: compulsory traffic is 44 B/edge instead of 268 B/edge.
//
// Arithmetic is the same split form as mp_step.cu (P/Q/R per node, 6x6 + 6x32 FMAs per edge), FP32 FFMA, with
// the same fixed summation order (a warp owns a node's whole out-edge segment), so results are bit-identical
// run to run and equal to the per-step kernels' up to nothing at all (same operation order).
//
// Validity is checked on the device (frame fits the capacity class, every target lies inside the frame); a
// violation raises flags[2] and the host re-runs the batch through the general per-step kernels.
#include "common.cuh"

namespace gnncca {
namespace fused {

constexpr int FN = 32, FE = 6, HC = 4, EP = 8;   // node width, edge width, classifier hidden, padded edge row
// Edge latents live in shared memory as PAIRS of edges, component-interleaved: pair p = edges (2p, 2p+1) occupies 12
// floats [k][2], so one 16-byte load yields two components of both edges in the operand order of the packed FFMA2
// (fma.rn.f32x2) that evaluates two messages per instruction.
__device__ __forceinline__ int eidx(int j, int k) { return (j >> 1) * 12 + k * 2 + (j & 1); }
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t ra, rb, rc, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
}
constexpr int HS = 36;                            // shared-memory row stride of the node latents: rows of different nodes start in different banks
constexpr int MAX_FA = 8;                         // widest raw edge attribute row supported

struct Params {
  const int64_t* graph_ptr; int64_t n_graphs; int64_t n_nodes; int64_t n_edges;
  const int32_t* rowptr; const int32_t* src; const int32_t* dst;
  const float* h_enc;                    // [N][FN] node encoder output
  const float* edge_attr; int fa;        // [E][fa], source-sorted order
  const float* w_enc; const float* b_enc;   // edge encoder Linear(fa, FE)
  gnncca_mp_weights w;
  int L, first_cls, agg, reatt_n, reatt_e, has_cls;
  int cap_n, cap_e;
  float* logits;                         // [(L - first_cls + 1)][E], classified steps in order
  float* h_dump; float* e_dump;          // optional per-step latents [L+1][N][FN], [L+1][E][FE] (tests / saved for backward)
  int32_t* flags;                        // [0],[1]: CSR invalid -> no-op; [2]: raised here when a frame does not fit
};

__device__ __forceinline__ float agg_init(int agg) { return agg == GNNCCA_AGG_MAX ? -INFINITY : 0.f; }
__device__ __forceinline__ float agg_comb(int agg, float a, float b) { return agg == GNNCCA_AGG_MAX ? fmaxf(a, b) : a + b; }

// Offsets (in floats) from the dynamic shared-memory base: integer offsets instead of pointers keep every access
// provably in the shared address space (LDS/STS, not generic LD/ST with runtime address-space checks).
struct Smem {
  int wproj;   // [K][44]   transposed projection weights: outputs 0..5 P, 6..11 Q, 12..43 R
  int bproj;   // [44]
  int wee;     // [FE][EIN]
  int wne;     // [FE][FN]
  int wenc;    // [FE][MAX_FA] , benc [FE]
  int benc;
  int c1;      // [HC][FE] (BatchNorm eval folded), b1 [HC], c2 [HC], b2 [1]
  int b1, c2, b2;
  int h;       // [cap_n][HS]
  int h0;      // [cap_n][HS] (reattach_nodes)
  int pqr;     // [cap_n][44]  per node: P (6, edge-update source part + bias) | Q (6, target part) | R (32, node-update part + bias)
  int e;       // [cap_e / 2][12]  pair-interleaved, see eidx()
  int e0;      // same layout (reattach_edges)
  int rp;      // int [cap_n + 1]
  int sl, dl;  // uint16 [cap_e] local source / target
};

__host__ __device__ inline size_t carve(Smem& s, int K, int EIN, int cap_n, int cap_e, bool rn, bool re) {
  int p = 0;
  s.wproj = p; p += K * 44;
  s.bproj = p; p += 44;
  s.wee = p; p += FE * EIN;
  s.wne = p; p += FE * FN;
  s.wenc = p; p += FE * MAX_FA;
  s.benc = p; p += 8;
  s.c1 = p; p += HC * FE;
  s.b1 = p; p += HC; s.c2 = p; p += HC; s.b2 = p; p += 4;
  p = (p + 3) & ~3;
  s.h = p; p += cap_n * HS;
  s.h0 = p; if (rn) p += cap_n * HS;
  s.pqr = p; p += cap_n * 44;
  s.e = p; p += ((cap_e + 1) / 2) * 12;
  s.e0 = p; if (re) p += ((cap_e + 1) / 2) * 12;
  s.rp = p; p += ((cap_n + 1 + 3) & ~3);
  s.sl = p; p += ((cap_e + 1) / 2 + 3) & ~3;
  s.dl = p; p += ((cap_e + 1) / 2 + 3) & ~3;
  return (size_t)p * sizeof(float) + 16;
}

template <bool REATT_E, int AGG>
__global__ void __launch_bounds__(256) mp_fused_fwd_kernel(const Params p) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  extern __shared__ __align__(16) float smem_f[];
  if ((p.flags[0] | p.flags[1]) != 0) return;      // CSR not valid (unsorted / out-of-range edge_index): uniform exit
  const int tid = threadIdx.x, NT = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarps = NT >> 5;
  const int K = p.reatt_n ? 2 * FN : FN;
  Smem so;
  carve(so, K, EIN, p.cap_n, p.cap_e, p.reatt_n != 0, REATT_E);
  int* const smem_i = reinterpret_cast<int*>(smem_f);
  uint16_t* const smem_h = reinterpret_cast<uint16_t*>(smem_f);

  // ---- weights, once per CTA (read from the live parameters at every launch)
  {
    const int edge_row = 2 * K + EIN, node_row = K + FE;
    for (int idx = tid; idx < K * 44; idx += NT) {
      const int k = idx / 44, o = idx - k * 44;
      float v;
      if (o < FE) v = p.w.w_edge[o * edge_row + k];                        // source block of the edge update
      else if (o < 2 * FE) v = p.w.w_edge[(o - FE) * edge_row + K + k];    // target block
      else v = p.w.w_node[(o - 2 * FE) * node_row + k];                    // node block of the node update
      (smem_f + so.wproj)[idx] = v;
    }
    for (int o = tid; o < 44; o += NT) (smem_f + so.bproj)[o] = o < FE ? p.w.b_edge[o] : (o < 2 * FE ? 0.f : p.w.b_node[o - 2 * FE]);
    for (int idx = tid; idx < FE * EIN; idx += NT) (smem_f + so.wee)[idx] = p.w.w_edge[(idx / EIN) * edge_row + 2 * K + idx % EIN];
    for (int idx = tid; idx < FE * FN; idx += NT) (smem_f + so.wne)[idx] = p.w.w_node[(idx % FN) * node_row + K + idx / FN];
    for (int idx = tid; idx < FE * MAX_FA; idx += NT) { const int o = idx / MAX_FA, k = idx % MAX_FA; (smem_f + so.wenc)[idx] = k < p.fa ? p.w_enc[o * p.fa + k] : 0.f; }
    for (int o = tid; o < FE; o += NT) (smem_f + so.benc)[o] = p.b_enc[o];
    if (p.has_cls) {
      for (int c = tid; c < HC; c += NT) {
        float scale = 1.f, shift = 0.f;
        if (p.w.bn_weight) {   // BatchNorm1d (eval) folded into the hidden layer
          scale = p.w.bn_weight[c] / sqrtf(p.w.bn_var[c] + p.w.bn_eps);
          shift = p.w.bn_bias[c] - p.w.bn_mean[c] * scale;
        }
        for (int k = 0; k < FE; ++k) (smem_f + so.c1)[c * FE + k] = p.w.w_c1[c * FE + k] * scale;
        (smem_f + so.b1)[c] = p.w.b_c1[c] * scale + shift;
        (smem_f + so.c2)[c] = p.w.w_c2[c];
      }
      if (tid == 0) (smem_f + so.b2)[0] = p.w.b_c2[0];
    }
  }
  __syncthreads();
  float wlo[FE], whi[FE];      // node-update weights of this lane's two features (c, c + 16)
#pragma unroll
  for (int k = 0; k < FE; ++k) { wlo[k] = (smem_f + so.wne)[k * FN + (lane & 15)]; whi[k] = (smem_f + so.wne)[k * FN + (lane & 15) + 16]; }

  for (int64_t g = blockIdx.x; g < p.n_graphs; g += gridDim.x) {
    const int64_t a = p.graph_ptr[g], b = p.graph_ptr[g + 1];
    if (a < 0 || b < a || b > p.n_nodes) { if (tid == 0) atomicOr(p.flags + 2, 1); continue; }
    const int n = (int)min(b - a, (int64_t)INT32_MAX);
    const int64_t eb = p.rowptr[a];
    const int ne = p.rowptr[b] - (int)eb;
    if (n > p.cap_n || ne > p.cap_e || ne < 0) { if (tid == 0) atomicOr(p.flags + 2, 1); continue; }   // CTA-uniform
    // ---- load the frame: encoded nodes, local CSR, edge encoder
    for (int idx = tid; idx < n * (FN / 4); idx += NT) {
      const float4 v = *reinterpret_cast<const float4*>(p.h_enc + a * FN + (int64_t)idx * 4);
      const int hoff = (idx >> 3) * HS + (idx & 7) * 4;
      *reinterpret_cast<float4*>((smem_f + so.h) + hoff) = v;
      if (p.reatt_n) *reinterpret_cast<float4*>((smem_f + so.h0) + hoff) = v;
    }
    for (int v = tid; v <= n; v += NT) (smem_i + so.rp)[v] = p.rowptr[a + v] - (int)eb;
    bool bad = false;
    for (int j = tid; j < ne; j += NT) {
      const int sg = ld_stream_i32(p.src + eb + j) - (int)a, dg = ld_stream_i32(p.dst + eb + j) - (int)a;
      const bool ok = (sg >= 0) & (sg < n) & (dg >= 0) & (dg < n);
      bad |= !ok;
      (smem_h + 2 * so.sl)[j] = (uint16_t)(ok ? sg : 0);
      (smem_h + 2 * so.dl)[j] = (uint16_t)(ok ? dg : 0);
      float ea[MAX_FA];
      if (p.fa == 4) {   // the shipped 4-attribute row: one 16-byte streaming load
        const float4 v4 = ld_stream_f4(p.edge_attr + (eb + j) * 4);
        ea[0] = v4.x; ea[1] = v4.y; ea[2] = v4.z; ea[3] = v4.w;
#pragma unroll
        for (int k = 4; k < MAX_FA; ++k) ea[k] = 0.f;
      } else {
#pragma unroll
        for (int k = 0; k < MAX_FA; ++k) ea[k] = k < p.fa ? p.edge_attr[(eb + j) * p.fa + k] : 0.f;
      }
      float en[EP];
#pragma unroll
      for (int o = 0; o < FE; ++o) {
        float acc = (smem_f + so.benc)[o];
#pragma unroll
        for (int k = 0; k < MAX_FA; ++k) acc = fmaf(ea[k], (smem_f + so.wenc)[o * MAX_FA + k], acc);
        en[o] = fmaxf(acc, 0.f);                     // Linear + ReLU (edge_out_dim != 1), mlp.py:14,18
      }
      en[6] = 0.f; en[7] = 0.f;
#pragma unroll
      for (int o = 0; o < FE; ++o) {
        (smem_f + so.e)[eidx(j, o)] = en[o];
        if (REATT_E) (smem_f + so.e0)[eidx(j, o)] = en[o];
      }
      if (p.e_dump) {
#pragma unroll
        for (int o = 0; o < FE; ++o) p.e_dump[(eb + j) * FE + o] = en[o];
      }
    }
    if (bad) atomicOr(p.flags + 2, 1);               // a target outside the frame: not a disjoint union of frames
    if (p.h_dump) for (int idx = tid; idx < n * FN; idx += NT) p.h_dump[a * FN + idx] = p.h_enc[a * FN + idx];
    __syncthreads();

    for (int step = 1; step <= p.L; ++step) {
      // ---- per-node projections P | Q | R  (Linear over a concatenation = sum of per-block products)
      // register tile: 2 nodes x 4 outputs per thread -- every weight quad fetched from shared memory feeds 8 FMAs
      // (the shared-memory pipe, not the FMA pipe, is what this kernel saturates first)
      const int npairs = (n + 1) >> 1;
      for (int item = tid; item < npairs * 11; item += NT) {
        const int vp = item / 11, q = item - vp * 11;
        const int v0 = 2 * vp, v1 = min(2 * vp + 1, n - 1);
        float4 a = *reinterpret_cast<const float4*>((smem_f + so.bproj) + 4 * q), b = a;
        const float* wq = (smem_f + so.wproj) + 4 * q;
        auto block = [&](const float* hbase) {
          const float4* ha = reinterpret_cast<const float4*>(hbase + v0 * HS);
          const float4* hb = reinterpret_cast<const float4*>(hbase + v1 * HS);
#pragma unroll 2
          for (int k4 = 0; k4 < FN / 4; ++k4) {
            const float4 xa = ha[k4], xb = hb[k4];
            const float xs[4] = {xa.x, xa.y, xa.z, xa.w}, ys[4] = {xb.x, xb.y, xb.z, xb.w};
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const float4 w = *reinterpret_cast<const float4*>(wq + (4 * k4 + kk) * 44);
              a.x = fmaf(xs[kk], w.x, a.x); a.y = fmaf(xs[kk], w.y, a.y); a.z = fmaf(xs[kk], w.z, a.z); a.w = fmaf(xs[kk], w.w, a.w);
              b.x = fmaf(ys[kk], w.x, b.x); b.y = fmaf(ys[kk], w.y, b.y); b.z = fmaf(ys[kk], w.z, b.z); b.w = fmaf(ys[kk], w.w, b.w);
            }
          }
          wq += FN * 44;
        };
        if (p.reatt_n) block(smem_f + so.h0);          // cat(initial, current), mpn.py:285
        block(smem_f + so.h);
        *reinterpret_cast<float4*>((smem_f + so.pqr) + v0 * 44 + 4 * q) = a;
        if (2 * vp + 1 < n) *reinterpret_cast<float4*>((smem_f + so.pqr) + v1 * 44 + 4 * q) = b;
      }
      __syncthreads();
      // ---- phase A: thread per edge -- edge update (+ classifier)
      const bool classify = p.has_cls && step >= p.first_cls;
      float* logit_out = classify ? p.logits + (int64_t)(step - p.first_cls) * p.n_edges + eb : nullptr;
      for (int j = tid; j < ne; j += NT) {
        const float* Ps = (smem_f + so.pqr) + (int)(smem_h + 2 * so.sl)[j] * 44;
        const float* Qd = (smem_f + so.pqr) + (int)(smem_h + 2 * so.dl)[j] * 44 + FE;
        const float4 p0 = *reinterpret_cast<const float4*>(Ps);
        const float2 p1 = *reinterpret_cast<const float2*>(Ps + 4);
        const float2 q0 = *reinterpret_cast<const float2*>(Qd), q1 = *reinterpret_cast<const float2*>(Qd + 2), q2 = *reinterpret_cast<const float2*>(Qd + 4);
        const float pq[FE] = {p0.x + q0.x, p0.y + q0.y, p0.z + q1.x, p0.w + q1.y, p1.x + q2.x, p1.y + q2.y};
        float ein[EIN];
        if (REATT_E) {   // cat(initial_edge_feats, latent_edge_feats), mpn.py:283
#pragma unroll
          for (int k = 0; k < FE; ++k) ein[k] = (smem_f + so.e0)[eidx(j, k)];
        }
        constexpr int OFF = REATT_E ? FE : 0;
        {
#pragma unroll
          for (int k = 0; k < FE; ++k) ein[OFF + k] = (smem_f + so.e)[eidx(j, k)];
        }
        float en[FE];
#pragma unroll
        for (int o = 0; o < FE; ++o) {
          float acc = pq[o];
#pragma unroll
          for (int k = 0; k < EIN; ++k) acc = fmaf(ein[k], (smem_f + so.wee)[o * EIN + k], acc);
          en[o] = fmaxf(acc, 0.f);
        }
#pragma unroll
        for (int o = 0; o < FE; ++o) (smem_f + so.e)[eidx(j, o)] = en[o];
        if (p.e_dump) {
#pragma unroll
          for (int o = 0; o < FE; ++o) p.e_dump[((int64_t)step * p.n_edges + eb + j) * FE + o] = en[o];
        }
        if (classify) {
          float logit = (smem_f + so.b2)[0];
#pragma unroll
          for (int c = 0; c < HC; ++c) {
            float hcv = (smem_f + so.b1)[c];
#pragma unroll
            for (int k = 0; k < FE; ++k) hcv = fmaf(en[k], (smem_f + so.c1)[c * FE + k], hcv);
            logit = fmaf(fmaxf(hcv, 0.f), (smem_f + so.c2)[c], logit);
          }
          st_stream_f1(logit_out + j, logit);
        }
      }
      if (step == p.L && !p.h_dump) break;           // the last node update is dead (mpn.py:288-299)
      __syncthreads();
      // ---- phase B: a HALF-warp owns a node's whole out-edge segment, lane = two features (c, c + 16); fixed
      //      summation order.  Two nodes per warp: one shared-memory wavefront serves an edge pair of each node.
      {
        const int half = lane >> 4, c0 = lane & 15;
        const float* eb_ = smem_f + so.e;
        for (int vb = 2 * warp; vb < n; vb += 2 * nwarps) {
          const int v = vb + half;
          const bool live = v < n;
          const int j0 = live ? (smem_i + so.rp)[v] : 0, j1 = live ? (smem_i + so.rp)[v + 1] : 0;
          const float r0 = live ? (smem_f + so.pqr)[v * 44 + 2 * FE + c0] : 0.f, r1 = live ? (smem_f + so.pqr)[v * 44 + 2 * FE + c0 + 16] : 0.f;
          float acc0 = agg_init(AGG), acc1 = agg_init(AGG);
          auto single = [&](int j) {          // one edge, scalar FFMA (odd segment ends)
            float t0 = r0, t1 = r1;
#pragma unroll
            for (int k = 0; k < FE; ++k) { const float x = eb_[eidx(j, k)]; t0 = fmaf(x, wlo[k], t0); t1 = fmaf(x, whi[k], t1); }
            acc0 = agg_comb(AGG, acc0, fmaxf(t0, 0.f)); acc1 = agg_comb(AGG, acc1, fmaxf(t1, 0.f));
          };
          int j = j0;
          if (j < j1 && (j & 1)) { single(j); ++j; }
          const float2 rr0 = make_float2(r0, r0), rr1 = make_float2(r1, r1);
          for (; j + 2 <= j1; j += 2) {        // an edge pair per iteration, messages folded in edge order
            const float4* pa = reinterpret_cast<const float4*>(eb_ + (j >> 1) * 12);
            const float4 a0 = pa[0], a1 = pa[1], a2 = pa[2];
            const float2 x0 = make_float2(a0.x, a0.y), x1 = make_float2(a0.z, a0.w), x2 = make_float2(a1.x, a1.y),
                         x3 = make_float2(a1.z, a1.w), x4 = make_float2(a2.x, a2.y), x5 = make_float2(a2.z, a2.w);
            float2 m0 = rr0, m1 = rr1;
            m0 = ffma2(x0, make_float2(wlo[0], wlo[0]), m0); m1 = ffma2(x0, make_float2(whi[0], whi[0]), m1);
            m0 = ffma2(x1, make_float2(wlo[1], wlo[1]), m0); m1 = ffma2(x1, make_float2(whi[1], whi[1]), m1);
            m0 = ffma2(x2, make_float2(wlo[2], wlo[2]), m0); m1 = ffma2(x2, make_float2(whi[2], whi[2]), m1);
            m0 = ffma2(x3, make_float2(wlo[3], wlo[3]), m0); m1 = ffma2(x3, make_float2(whi[3], whi[3]), m1);
            m0 = ffma2(x4, make_float2(wlo[4], wlo[4]), m0); m1 = ffma2(x4, make_float2(whi[4], whi[4]), m1);
            m0 = ffma2(x5, make_float2(wlo[5], wlo[5]), m0); m1 = ffma2(x5, make_float2(whi[5], whi[5]), m1);
            acc0 = agg_comb(AGG, acc0, fmaxf(m0.x, 0.f)); acc0 = agg_comb(AGG, acc0, fmaxf(m0.y, 0.f));
            acc1 = agg_comb(AGG, acc1, fmaxf(m1.x, 0.f)); acc1 = agg_comb(AGG, acc1, fmaxf(m1.y, 0.f));
          }
          if (j < j1) single(j);
          if (j1 == j0) { acc0 = 0.f; acc1 = 0.f; }                        // no out-edge: torch_scatter leaves 0
          else if (AGG == GNNCCA_AGG_MEAN) { const float inv = 1.f / (float)(j1 - j0); acc0 = acc0 * inv; acc1 = acc1 * inv; }
          if (live) {
            (smem_f + so.h)[v * HS + c0] = acc0;
            (smem_f + so.h)[v * HS + c0 + 16] = acc1;
            if (p.h_dump) {
              p.h_dump[((int64_t)step * p.n_nodes + a + v) * FN + c0] = acc0;
              p.h_dump[((int64_t)step * p.n_nodes + a + v) * FN + c0 + 16] = acc1;
            }
          }
        }
      }
      __syncthreads();
    }
    __syncthreads();   // the next frame overwrites every buffer
  }
}

}  // namespace fused
}  // namespace gnncca

using namespace gnncca;

extern "C" int gnncca_mp_fused_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* w_enc_edge,
                                   const float* b_enc_edge, int32_t edge_in_dim, const int64_t* graph_ptr, int64_t n_graphs,
                                   const int32_t* rowptr, const int32_t* src, const int32_t* dst, int64_t n_nodes,
                                   int64_t n_edges, const float* h_enc, const float* edge_attr, int32_t num_steps,
                                   int32_t first_class_step, int32_t cap_nodes, int32_t cap_edges, float* logits,
                                   float* h_dump, float* e_dump, int32_t* flags, void* stream) {
  if (!cfg || !w || n_graphs < 0 || n_nodes < 0 || n_edges < 0 || num_steps < 1 || !flags) return GNNCCA_ERR_BAD_ARG;
  if (n_graphs == 0 || n_nodes == 0) return GNNCCA_OK;
  if (!graph_ptr || !rowptr || !h_enc || !w->w_edge || !w->b_edge || !w->w_node || !w->b_node || !w_enc_edge || !b_enc_edge)
    return GNNCCA_ERR_BAD_ARG;
  if (n_edges > 0 && (!src || !dst || !edge_attr)) return GNNCCA_ERR_BAD_ARG;
  if (cfg->f_node != fused::FN || cfg->f_edge != fused::FE) return GNNCCA_ERR_UNSUPPORTED;
  if (edge_in_dim < 1 || edge_in_dim > fused::MAX_FA) return GNNCCA_ERR_UNSUPPORTED;
  if (cfg->agg < GNNCCA_AGG_SUM || cfg->agg > GNNCCA_AGG_MAX) return GNNCCA_ERR_BAD_ARG;
  if (cap_nodes < 1 || cap_nodes > 65535 || cap_edges < 1) return GNNCCA_ERR_BAD_ARG;
  const bool has_cls = first_class_step <= num_steps;
  if (has_cls && (cfg->cls_hidden != fused::HC || !w->w_c1 || !w->b_c1 || !w->w_c2 || !w->b_c2 || !logits || first_class_step < 1))
    return cfg->cls_hidden != fused::HC ? GNNCCA_ERR_UNSUPPORTED : GNNCCA_ERR_BAD_ARG;
  if ((((uintptr_t)h_enc | (uintptr_t)edge_attr) & 15) != 0) return GNNCCA_ERR_BAD_ARG;

  fused::Params p{};
  p.graph_ptr = graph_ptr; p.n_graphs = n_graphs; p.n_nodes = n_nodes; p.n_edges = n_edges;
  p.rowptr = rowptr; p.src = src; p.dst = dst; p.h_enc = h_enc; p.edge_attr = edge_attr; p.fa = edge_in_dim;
  p.w_enc = w_enc_edge; p.b_enc = b_enc_edge; p.w = *w;
  p.L = num_steps; p.first_cls = first_class_step; p.agg = cfg->agg;
  p.reatt_n = cfg->reattach_nodes ? 1 : 0; p.reatt_e = cfg->reattach_edges ? 1 : 0; p.has_cls = has_cls ? 1 : 0;
  p.cap_n = cap_nodes; p.cap_e = cap_edges; p.logits = logits; p.h_dump = h_dump; p.e_dump = e_dump; p.flags = flags;

  fused::Smem lay;
  const int K = p.reatt_n ? 64 : 32, EIN = p.reatt_e ? 12 : 6;
  const size_t smem = fused::carve(lay, K, EIN, cap_nodes, cap_edges, p.reatt_n != 0, p.reatt_e != 0);
  if (smem > 227 * 1024) return GNNCCA_ERR_UNSUPPORTED;   // frame class too large for one CTA: use the per-step kernels
  const int threads = cap_edges <= 512 ? 128 : 256;
  cudaStream_t st = (cudaStream_t)stream;
  void (*kern)(const fused::Params) = nullptr;
  if (p.reatt_e) kern = p.agg == GNNCCA_AGG_MAX ? fused::mp_fused_fwd_kernel<true, GNNCCA_AGG_MAX>
                      : p.agg == GNNCCA_AGG_MEAN ? fused::mp_fused_fwd_kernel<true, GNNCCA_AGG_MEAN> : fused::mp_fused_fwd_kernel<true, GNNCCA_AGG_SUM>;
  else kern = p.agg == GNNCCA_AGG_MAX ? fused::mp_fused_fwd_kernel<false, GNNCCA_AGG_MAX>
            : p.agg == GNNCCA_AGG_MEAN ? fused::mp_fused_fwd_kernel<false, GNNCCA_AGG_MEAN> : fused::mp_fused_fwd_kernel<false, GNNCCA_AGG_SUM>;
  if (smem > 48 * 1024) {
    int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    if (rc) return rc;
  }
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem) != cudaSuccess || occ < 1) { (void)cudaGetLastError(); occ = 1; }
  const int64_t want = (int64_t)kNumSMs * occ;        // every frame of the wave resident at once when the batch allows
  const int grid = (int)(n_graphs < want ? n_graphs : want);
  kern<<<grid, threads, smem, st>>>(p);
  return check_launch();
}
