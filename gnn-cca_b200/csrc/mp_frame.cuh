// All message-passing steps of MOTMPNet.forward (reference models/mpn.py:270-299) for ONE small frame graph, run
// by a GROUP of warps (up to 7 groups per CTA in mp_fused.cu) that
// keeps the frame's node and edge latents in shared memory across the L steps: edge encoder (mpn.py:128-142),
// EdgeModel (:59-69), NodeModel (:71-101) with the aggregation of :192-202, classifier (:290-293).  HBM sees the
// encoded nodes, edge_attr, the edge list and the per-step logits, nothing of size E between steps.
//
// Arithmetic (split form of DESIGN.md section 2):
//   per node   P|Q|R = [h0,] h . Wp^T + b      [n x K] x [K x 44]   mma.sync m16n8k8 with 3xTF32 operand splitting
//                                                                   (hi*hi + hi*lo + lo*hi, fp32 accumulate: fp32-level
//                                                                   accuracy, a tenth of the FFMA instruction count)
//   per edge   e' = relu(P[src] + Q[dst] + Wee e)                   packed FFMA2 over edge PAIRS (thread = 2 edges)
//              logit = w2 . relu(c1 e' + b1) + b2                   same pairs
//   per node   h' = agg_j relu(R[v] + Wne e'_j)                     half-warp per node, lane = 2 features, FFMA2 over
//                                                                   edge pairs, messages folded in edge order
// Fixed operation order everywhere: results are bit-identical run to run and between the two host kernels.
#pragma once
#include "common.cuh"

namespace gnncca {
namespace mpf {

constexpr int FN = 32, FE = 6, HC = 4;
constexpr int HS = 36;        // row stride of the node latents: A-fragment loads (4*row + col) hit 32 distinct banks
constexpr int PS = 44;        // P (6) | Q (6) | R (32) per node
constexpr int MAX_FA = 8;     // widest raw edge attribute row

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
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  uint64_t ra, rb, rd;
  asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
  float2 d;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
  return d;
}
__device__ __forceinline__ float2 bc(float w) { return make_float2(w, w); }   // scalar operand of FFMA2 (no extra register)
__device__ __forceinline__ float2 relu2(float2 v) { return make_float2(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f)); }
__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const float (&a)[4], float b0, float b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}

// A group of warps with its own named barrier (id 0 is left to __syncthreads)
struct Group {
  int tid, size, warp, nwarps, lane, bar;
  __device__ __forceinline__ void sync() const { asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(size) : "memory"); }
};

struct Net {                   // what the network is (uniform over the launch)
  gnncca_mp_weights w;
  const float* w_enc; const float* b_enc; int fa;     // edge encoder Linear(fa, FE)
  int L, first_cls, has_cls, agg, reatt_n, reatt_e;
  int cap_n, cap_e;
};

struct Batch {                 // where the batch lives
  const int64_t* graph_ptr; int64_t n_graphs, n_nodes, n_edges;
  const int64_t* edge_index;   // RAW mode: caller's [2][E] int64, must be source-major (checked here); else nullptr
  const int64_t* edge_ptr;     // RAW mode: [n_graphs + 1] edge offsets of the frames
  const int32_t* rowptr; const int32_t* src; const int32_t* dst;   // CSR mode (gnncca_csr_from_edge_index)
  const float* h_enc;          // [N][FN]
  const float* edge_attr;      // [E][fa]
  float* logits;               // [(L - first_cls + 1)][E]
  float* h_dump; float* e_dump;
  int32_t* flags;              // [2] raised when a frame does not fit / is not a closed, source-major frame
};

// ---- shared-memory layouts, as float offsets (integer offsets keep every access provably in the shared space)
struct WOff { int wB, bproj, wee, wne, wenc, benc, c1, b1, c2, b2, total; };
struct SOff { int h, h0, pqr, e, e0, rp, sl, dl, bad, total; };

__host__ __device__ inline WOff weight_layout(int K, int EIN) {
  WOff o; int p = 0;
  o.wB = p; p += K * 96;           // [K/8][6 n-tiles][32 lanes][b0 hi, b1 hi, b0 lo, b1 lo]
  o.bproj = p; p += 48;
  o.wee = p; p += FE * EIN;
  o.wne = p; p += FE * FN;         // [k][c]
  o.wenc = p; p += FE * MAX_FA;
  o.benc = p; p += 8;
  o.c1 = p; p += HC * FE;
  o.b1 = p; p += HC; o.c2 = p; p += HC; o.b2 = p; p += 4;
  o.total = (p + 3) & ~3;
  return o;
}
__host__ __device__ inline SOff state_layout(int cap_n, int cap_e, bool rn, bool re) {
  SOff o; int p = 0;
  const int pairs = (cap_e + 1) / 2;
  o.h = p; p += cap_n * HS;
  o.h0 = p; if (rn) p += cap_n * HS;
  o.pqr = p; p += cap_n * PS;
  o.e = p; p += pairs * 12;        // pair-interleaved: pair q = edges (2q, 2q+1) holds [k][2]
  o.e0 = p; if (re) p += pairs * 12;
  o.rp = p; p += (cap_n + 1 + 3) & ~3;
  o.sl = p; p += (pairs + 3) & ~3; // uint16 local source ids, two per word
  o.dl = p; p += (pairs + 3) & ~3;
  o.bad = p; p += 4;
  o.total = (p + 3) & ~3;
  return o;
}

// ---- weights: read from the live parameters, re-laid out for the kernels above (once per CTA)
__device__ inline void load_weights(float* sw, const WOff& wo, const Net& nt, int tid, int nthreads) {
  const int K = nt.reatt_n ? 2 * FN : FN, EIN = nt.reatt_e ? 2 * FE : FE;
  const int edge_row = 2 * K + EIN, node_row = K + FE;
  for (int idx = tid; idx < K * 24; idx += nthreads) {          // one (k-step, n-tile, lane) slot = 4 floats
    const int lane = idx & 31, t = lane & 3, g = lane >> 2;
    const int nti = (idx >> 5) % 6, ks = idx / 192;
    const int o = nti * 8 + g;
    float b[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int k = ks * 8 + t + 4 * i;
      float v = 0.f;
      if (o < FE) v = nt.w.w_edge[o * edge_row + k];                            // source block of the edge update
      else if (o < 2 * FE) v = nt.w.w_edge[(o - FE) * edge_row + K + k];        // target block
      else if (o < PS) v = nt.w.w_node[(o - 2 * FE) * node_row + k];            // node block of the node update
      b[i] = v;
    }
    const float h0 = tf32_rna(b[0]), h1 = tf32_rna(b[1]);
    *reinterpret_cast<float4*>(sw + wo.wB + idx * 4) = make_float4(h0, h1, tf32_rna(b[0] - h0), tf32_rna(b[1] - h1));
  }
  for (int o = tid; o < 48; o += nthreads) sw[wo.bproj + o] = o < FE ? nt.w.b_edge[o] : (o < 2 * FE || o >= PS ? 0.f : nt.w.b_node[o - 2 * FE]);
  for (int idx = tid; idx < FE * EIN; idx += nthreads) sw[wo.wee + idx] = nt.w.w_edge[(idx / EIN) * edge_row + 2 * K + idx % EIN];
  for (int idx = tid; idx < FE * FN; idx += nthreads) sw[wo.wne + idx] = nt.w.w_node[(idx % FN) * node_row + K + idx / FN];
  for (int idx = tid; idx < FE * MAX_FA; idx += nthreads) { const int o = idx / MAX_FA, k = idx % MAX_FA; sw[wo.wenc + idx] = k < nt.fa ? nt.w_enc[o * nt.fa + k] : 0.f; }
  for (int o = tid; o < 8; o += nthreads) sw[wo.benc + o] = o < FE ? nt.b_enc[o] : 0.f;
  if (nt.has_cls) {
    for (int c = tid; c < HC; c += nthreads) {
      float scale = 1.f, shift = 0.f;
      if (nt.w.bn_weight) {   // BatchNorm1d (eval) folded into the hidden layer
        scale = nt.w.bn_weight[c] / sqrtf(nt.w.bn_var[c] + nt.w.bn_eps);
        shift = nt.w.bn_bias[c] - nt.w.bn_mean[c] * scale;
      }
      for (int k = 0; k < FE; ++k) sw[wo.c1 + c * FE + k] = nt.w.w_c1[c * FE + k] * scale;
      sw[wo.b1 + c] = nt.w.b_c1[c] * scale + shift;
      sw[wo.c2 + c] = nt.w.w_c2[c];
    }
    if (tid == 0) sw[wo.b2] = nt.w.b_c2[0];
  }
}

__device__ __forceinline__ float agg_init(int agg) { return agg == GNNCCA_AGG_MAX ? -INFINITY : 0.f; }
__device__ __forceinline__ float agg_comb(int agg, float a, float b) { return agg == GNNCCA_AGG_MAX ? fmaxf(a, b) : a + b; }

// A frame that breaks the contract of its batch (larger than the capacity class, an edge leaving the frame, sources not
// ascending) is skipped with flags[2] raised; its logits are filled with NaN so that nobody mistakes uninitialised memory for a
// result (the host re-runs such a batch on the checked path, or raises: engine.FastPath.verify).
__device__ __forceinline__ void poison_frame_logits(const Group& gr, const Net& nt, const Batch& bt, int64_t e0, int64_t e1) {
  if (!nt.has_cls || bt.logits == nullptr || e0 < 0 || e1 > bt.n_edges || e1 <= e0) return;
  const float qnan = __int_as_float(0x7fc00000);
  for (int s = 0; s <= nt.L - nt.first_cls; ++s)
    for (int64_t j = e0 + gr.tid; j < e1; j += gr.size) bt.logits[(int64_t)s * bt.n_edges + j] = qnan;
}

// One frame, all steps.  sw: the CTA's weights; ss: this group's state.  Every thread of the group calls it with
// the same arguments; returns after the group's last barrier (the state may be overwritten at once).
template <bool REATT_E, int AGG>
__device__ __forceinline__ void mp_frame(const Group& gr, const Net& nt, const Batch& bt, const float* sw, const WOff& wo,
                                         float* ss, const SOff& so, int64_t g, bool dep_wait = false) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  int* const ssi = reinterpret_cast<int*>(ss);
  const int tid = gr.tid, NT = gr.size, lane = gr.lane;
  const int K = nt.reatt_n ? 2 * FN : FN;

  const int64_t a = bt.graph_ptr[g], b = bt.graph_ptr[g + 1];
  const bool raw = bt.edge_index != nullptr;
  bool fits = !(a < 0 || b < a || b > bt.n_nodes);
  const int n = fits ? (int)min(b - a, (int64_t)INT32_MAX) : 0;
  int64_t eb = 0; int ne = 0;
  int64_t pe0 = -1, pe1 = -1;      // the frame's edge range as far as it can be trusted (for the NaN fill of a skipped frame)
  if (fits) {
    if (raw) {
      const int64_t e0 = bt.edge_ptr[g], e1 = bt.edge_ptr[g + 1];
      if (e0 >= 0 && e1 >= e0 && e1 <= bt.n_edges) { pe0 = e0; pe1 = e1; }
      fits = e0 >= 0 && e1 >= e0 && e1 <= bt.n_edges && e1 - e0 <= nt.cap_e;
      eb = e0; ne = fits ? (int)(e1 - e0) : 0;
    } else {
      eb = bt.rowptr[a]; ne = bt.rowptr[b] - (int)eb;
      if (ne >= 0) { pe0 = eb; pe1 = eb + ne; }
      fits = ne >= 0 && ne <= nt.cap_e;
    }
  }
  if (!fits || n > nt.cap_n) {     // group-uniform
    if (tid == 0) atomicOr(bt.flags + 2, 1);
    poison_frame_logits(gr, nt, bt, pe0, pe1);
    return;
  }

  // ---- load the frame: local CSR + edge encoder first (inputs of the forward), the encoded nodes last: when the kernel is a
  //      programmatic dependent of the node encoder (dep_wait), everything before the wait overlaps the encoder's tail
  if (tid == 0) ssi[so.bad] = 0;
  if (!raw) for (int v = tid; v <= n; v += NT) ssi[so.rp + v] = bt.rowptr[a + v] - (int)eb;
  else if (ne == 0) for (int v = tid; v <= n; v += NT) ssi[so.rp + v] = 0;
  gr.sync();                                       // bad flag cleared before anyone raises it
  const int npairs = (ne + 1) >> 1;
  {
    bool bad = false;
    for (int q = tid; q < npairs; q += NT) {
      const int j0 = 2 * q;
      const bool has1 = j0 + 1 < ne;
      int s[2], d[2];
      if (raw) {
        const int64_t* es = bt.edge_index + eb + j0;
        const int64_t* ed = es + bt.n_edges;
        const int64_t s0 = es[0] - a, d0 = ed[0] - a, s1 = has1 ? es[1] - a : s0, d1 = has1 ? ed[1] - a : d0;
        const int64_t pv = j0 > 0 ? es[-1] - a : -1;
        const bool ok = (s0 >= 0) & (s0 < n) & (d0 >= 0) & (d0 < n) & (s1 >= 0) & (s1 < n) & (d1 >= 0) & (d1 < n) & (pv >= -1) & (pv <= s0) & (s0 <= s1);
        bad |= !ok;
        s[0] = ok ? (int)s0 : 0; s[1] = ok ? (int)s1 : 0; d[0] = ok ? (int)d0 : 0; d[1] = ok ? (int)d1 : 0;
        if (ok) {      // edge j opens every row in (source of j-1, source of j]; the last edge closes the rows after it
          for (int v = (int)pv + 1; v <= s[0]; ++v) ssi[so.rp + v] = j0;
          if (has1) for (int v = s[0] + 1; v <= s[1]; ++v) ssi[so.rp + v] = j0 + 1;
          if (j0 + 2 >= ne) for (int v = s[1] + 1; v <= n; ++v) ssi[so.rp + v] = ne;
        }
      } else {
        const int s0 = ld_stream_i32(bt.src + eb + j0) - (int)a, d0 = ld_stream_i32(bt.dst + eb + j0) - (int)a;
        const int s1 = has1 ? ld_stream_i32(bt.src + eb + j0 + 1) - (int)a : s0, d1 = has1 ? ld_stream_i32(bt.dst + eb + j0 + 1) - (int)a : d0;
        const bool ok = (s0 >= 0) & (s0 < n) & (d0 >= 0) & (d0 < n) & (s1 >= 0) & (s1 < n) & (d1 >= 0) & (d1 < n);
        bad |= !ok;
        s[0] = ok ? s0 : 0; s[1] = ok ? s1 : 0; d[0] = ok ? d0 : 0; d[1] = ok ? d1 : 0;
      }
      ssi[so.sl + q] = s[0] | (s[1] << 16);
      ssi[so.dl + q] = d[0] | (d[1] << 16);
      float2 en[FE];
      if (nt.fa == 4) {   // the shipped 4-attribute row: one 16-byte streaming load per edge
        const float4 x0 = ld_stream_f4(bt.edge_attr + (eb + j0) * 4);
        const float4 x1 = has1 ? ld_stream_f4(bt.edge_attr + (eb + j0 + 1) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float2 xa = make_float2(x0.x, x1.x), xb = make_float2(x0.y, x1.y), xc = make_float2(x0.z, x1.z), xd = make_float2(x0.w, x1.w);
#pragma unroll
        for (int o = 0; o < FE; ++o) {              // Linear + ReLU (edge_out_dim != 1), mlp.py:14,18
          const float4 w4 = *reinterpret_cast<const float4*>(sw + wo.wenc + o * MAX_FA);
          float2 acc = bc(sw[wo.benc + o]);
          acc = ffma2(xa, bc(w4.x), acc); acc = ffma2(xb, bc(w4.y), acc);
          acc = ffma2(xc, bc(w4.z), acc); acc = ffma2(xd, bc(w4.w), acc);
          en[o] = relu2(acc);
        }
      } else {
#pragma unroll
        for (int o = 0; o < FE; ++o) {
          float a0 = sw[wo.benc + o], a1 = a0;
          for (int k = 0; k < nt.fa; ++k) {
            const float w = sw[wo.wenc + o * MAX_FA + k];
            a0 = fmaf(bt.edge_attr[(eb + j0) * nt.fa + k], w, a0);
            if (has1) a1 = fmaf(bt.edge_attr[(eb + j0 + 1) * nt.fa + k], w, a1);
          }
          en[o] = make_float2(fmaxf(a0, 0.f), fmaxf(a1, 0.f));
        }
      }
      if (!has1) {
#pragma unroll
        for (int o = 0; o < FE; ++o) en[o].y = 0.f;
      }
      float4* ep = reinterpret_cast<float4*>(ss + so.e + q * 12);
      ep[0] = make_float4(en[0].x, en[0].y, en[1].x, en[1].y);
      ep[1] = make_float4(en[2].x, en[2].y, en[3].x, en[3].y);
      ep[2] = make_float4(en[4].x, en[4].y, en[5].x, en[5].y);
      if (REATT_E) {
        float4* ep0 = reinterpret_cast<float4*>(ss + so.e0 + q * 12);
        ep0[0] = ep[0]; ep0[1] = ep[1]; ep0[2] = ep[2];
      }
      if (bt.e_dump) {
#pragma unroll
        for (int o = 0; o < FE; ++o) {
          bt.e_dump[(eb + j0) * FE + o] = en[o].x;
          if (has1) bt.e_dump[(eb + j0 + 1) * FE + o] = en[o].y;
        }
      }
    }
    if (bad) ssi[so.bad] = 1;
  }
  if (dep_wait) asm volatile("griddepcontrol.wait;" ::: "memory");      // the encoder's output is visible from here on
  for (int idx = tid; idx < n * (FN / 4); idx += NT) {
    const float4 v = __ldcg(reinterpret_cast<const float4*>(bt.h_enc + a * FN) + idx);
    const int hoff = (idx >> 3) * HS + (idx & 7) * 4;
    *reinterpret_cast<float4*>(ss + so.h + hoff) = v;
    if (nt.reatt_n) *reinterpret_cast<float4*>(ss + so.h0 + hoff) = v;
    if (bt.h_dump) *(reinterpret_cast<float4*>(bt.h_dump + a * FN) + idx) = v;
  }
  gr.sync();
  if (ssi[so.bad]) {                                 // not a closed, source-major frame: group-uniform exit
    if (tid == 0) atomicOr(bt.flags + 2, 1);
    poison_frame_logits(gr, nt, bt, pe0, pe1);
    gr.sync();
    return;
  }

  float wlo[FE], whi[FE];      // node-update weights of this lane's two features (c, c + 16)
#pragma unroll
  for (int k = 0; k < FE; ++k) { wlo[k] = sw[wo.wne + k * FN + (lane & 15)]; whi[k] = sw[wo.wne + k * FN + (lane & 15) + 16]; }

  for (int step = 1; step <= nt.L; ++step) {
    // ---- per-node projections P | Q | R on the tensor cores (Linear over a concatenation = sum of per-block
    //      products): a unit = 16 nodes x 24 outputs, 3xTF32
    {
      const int gq = lane >> 2, t = lane & 3;
      const int units = ((n + 15) >> 4) * 2;
      for (int u = gr.warp; u < units; u += gr.nwarps) {
        const int mt = u >> 1, half = u & 1;
        float c[3][4];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const float2 bb = *reinterpret_cast<const float2*>(sw + wo.bproj + (half * 3 + i) * 8 + 2 * t);
          c[i][0] = bb.x; c[i][1] = bb.y; c[i][2] = bb.x; c[i][3] = bb.y;
        }
        const int r0 = mt * 16 + gq;
        for (int ks = 0; ks < K / 8; ++ks) {
          const float* hb = (nt.reatt_n && ks < FN / 8) ? ss + so.h0 : ss + so.h;     // cat(initial, current), mpn.py:285
          const int k0 = (ks * 8) & (FN - 1);
          float ah[4], al[4];
          const float av[4] = {hb[r0 * HS + k0 + t], hb[(r0 + 8) * HS + k0 + t], hb[r0 * HS + k0 + t + 4], hb[(r0 + 8) * HS + k0 + t + 4]};
#pragma unroll
          for (int i = 0; i < 4; ++i) {     // hi = the tf32 prefix of the fp32 bits (what the tensor core reads anyway), lo = the exact rest
            ah[i] = __uint_as_float(__float_as_uint(av[i]) & 0xffffe000u);
            al[i] = av[i] - ah[i];
          }
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const float4 w = *reinterpret_cast<const float4*>(sw + wo.wB + ((ks * 6 + half * 3 + i) * 32 + lane) * 4);
            mma_tf32(c[i], ah, w.z, w.w);     // hi * lo
            mma_tf32(c[i], al, w.x, w.y);     // lo * hi
            mma_tf32(c[i], ah, w.x, w.y);     // hi * hi
          }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const int col = (half * 3 + i) * 8 + 2 * t;
          if (col < PS) {
            if (r0 < n) *reinterpret_cast<float2*>(ss + so.pqr + r0 * PS + col) = make_float2(c[i][0], c[i][1]);
            if (r0 + 8 < n) *reinterpret_cast<float2*>(ss + so.pqr + (r0 + 8) * PS + col) = make_float2(c[i][2], c[i][3]);
          }
        }
      }
    }
    gr.sync();
    // ---- phase A: thread per edge PAIR -- edge update (+ classifier)
    {
      const bool classify = nt.has_cls && step >= nt.first_cls;
      float* logit_out = classify ? bt.logits + (int64_t)(step - nt.first_cls) * bt.n_edges + eb : nullptr;
      for (int q = tid; q < npairs; q += NT) {
        const uint32_t s2 = (uint32_t)ssi[so.sl + q], d2 = (uint32_t)ssi[so.dl + q];
        const float* Pa = ss + so.pqr + (int)(s2 & 0xffffu) * PS;
        const float* Pb = ss + so.pqr + (int)(s2 >> 16) * PS;
        const float* Qa = ss + so.pqr + (int)(d2 & 0xffffu) * PS + FE;
        const float* Qb = ss + so.pqr + (int)(d2 >> 16) * PS + FE;
        const float4 pa0 = *reinterpret_cast<const float4*>(Pa), pb0 = *reinterpret_cast<const float4*>(Pb);
        const float2 pa1 = *reinterpret_cast<const float2*>(Pa + 4), pb1 = *reinterpret_cast<const float2*>(Pb + 4);
        const float2 qa0 = *reinterpret_cast<const float2*>(Qa), qa1 = *reinterpret_cast<const float2*>(Qa + 2), qa2 = *reinterpret_cast<const float2*>(Qa + 4);
        const float2 qb0 = *reinterpret_cast<const float2*>(Qb), qb1 = *reinterpret_cast<const float2*>(Qb + 2), qb2 = *reinterpret_cast<const float2*>(Qb + 4);
        float2 acc[FE] = {make_float2(pa0.x + qa0.x, pb0.x + qb0.x), make_float2(pa0.y + qa0.y, pb0.y + qb0.y),
                          make_float2(pa0.z + qa1.x, pb0.z + qb1.x), make_float2(pa0.w + qa1.y, pb0.w + qb1.y),
                          make_float2(pa1.x + qa2.x, pb1.x + qb2.x), make_float2(pa1.y + qa2.y, pb1.y + qb2.y)};
        float2 ein[EIN];
        if (REATT_E) {   // cat(initial_edge_feats, latent_edge_feats), mpn.py:283
          const float4* e0p = reinterpret_cast<const float4*>(ss + so.e0 + q * 12);
          const float4 v0 = e0p[0], v1 = e0p[1], v2 = e0p[2];
          ein[0] = make_float2(v0.x, v0.y); ein[1] = make_float2(v0.z, v0.w); ein[2] = make_float2(v1.x, v1.y);
          ein[3] = make_float2(v1.z, v1.w); ein[4] = make_float2(v2.x, v2.y); ein[5] = make_float2(v2.z, v2.w);
        }
        constexpr int OFF = REATT_E ? FE : 0;
        float4* ep = reinterpret_cast<float4*>(ss + so.e + q * 12);
        {
          const float4 v0 = ep[0], v1 = ep[1], v2 = ep[2];
          ein[OFF + 0] = make_float2(v0.x, v0.y); ein[OFF + 1] = make_float2(v0.z, v0.w); ein[OFF + 2] = make_float2(v1.x, v1.y);
          ein[OFF + 3] = make_float2(v1.z, v1.w); ein[OFF + 4] = make_float2(v2.x, v2.y); ein[OFF + 5] = make_float2(v2.z, v2.w);
        }
        float2 en[FE];
#pragma unroll
        for (int o = 0; o < FE; ++o) {
          float2 v = acc[o];
#pragma unroll
          for (int k = 0; k < EIN; ++k) v = ffma2(ein[k], bc(sw[wo.wee + o * EIN + k]), v);
          en[o] = relu2(v);
        }
        ep[0] = make_float4(en[0].x, en[0].y, en[1].x, en[1].y);
        ep[1] = make_float4(en[2].x, en[2].y, en[3].x, en[3].y);
        ep[2] = make_float4(en[4].x, en[4].y, en[5].x, en[5].y);
        const int j0 = 2 * q;
        const bool has1 = j0 + 1 < ne;
        if (bt.e_dump) {
#pragma unroll
          for (int o = 0; o < FE; ++o) {
            bt.e_dump[((int64_t)step * bt.n_edges + eb + j0) * FE + o] = en[o].x;
            if (has1) bt.e_dump[((int64_t)step * bt.n_edges + eb + j0 + 1) * FE + o] = en[o].y;
          }
        }
        if (classify) {
          float2 logit = bc(sw[wo.b2]);
#pragma unroll
          for (int c = 0; c < HC; ++c) {
            float2 hcv = bc(sw[wo.b1 + c]);
#pragma unroll
            for (int k = 0; k < FE; ++k) hcv = ffma2(en[k], bc(sw[wo.c1 + c * FE + k]), hcv);
            logit = ffma2(relu2(hcv), bc(sw[wo.c2 + c]), logit);
          }
          st_stream_f1(logit_out + j0, logit.x);
          if (has1) st_stream_f1(logit_out + j0 + 1, logit.y);
        }
      }
    }
    if (step == nt.L && !bt.h_dump) break;           // the last node update is dead (mpn.py:288-299)
    gr.sync();
    // ---- phase B: a HALF-warp owns a node's whole out-edge segment, lane = two features (c, c + 16); fixed
    //      summation order.  Two nodes per warp: one shared-memory wavefront serves an edge pair of each node.
    {
      const int half = lane >> 4, c0 = lane & 15;
      const float* eb_ = ss + so.e;
      for (int vb = 2 * gr.warp; vb < n; vb += 2 * gr.nwarps) {
        const int v = vb + half;
        const bool live = v < n;
        const int j0 = live ? ssi[so.rp + v] : 0, j1 = live ? ssi[so.rp + v + 1] : 0;
        const float r0 = live ? ss[so.pqr + v * PS + 2 * FE + c0] : 0.f, r1 = live ? ss[so.pqr + v * PS + 2 * FE + c0 + 16] : 0.f;
        // even / odd edges of the segment are folded into the two halves of a packed accumulator (one FADD2 per
        // feature and edge pair); the halves are combined at the end -- a fixed order either way
        float2 acc0 = bc(agg_init(AGG)), acc1 = bc(agg_init(AGG));
        auto single = [&](int j) {          // one edge, scalar FFMA (odd segment ends)
          float t0 = r0, t1 = r1;
#pragma unroll
          for (int k = 0; k < FE; ++k) { const float x = eb_[(j >> 1) * 12 + k * 2 + (j & 1)]; t0 = fmaf(x, wlo[k], t0); t1 = fmaf(x, whi[k], t1); }
          acc0.x = agg_comb(AGG, acc0.x, fmaxf(t0, 0.f)); acc1.x = agg_comb(AGG, acc1.x, fmaxf(t1, 0.f));
        };
        int j = j0;
        if (j < j1 && (j & 1)) { single(j); ++j; }
        const float2 rr0 = make_float2(r0, r0), rr1 = make_float2(r1, r1);
        {                                    // an edge pair per iteration; pointer-stepped so the unrolled body has no index math
          const float4* pa = reinterpret_cast<const float4*>(eb_ + (j >> 1) * 12);
          const int npair = (j1 - j) >> 1;
#pragma unroll 4
          for (int i = 0; i < npair; ++i, pa += 3) {
            const float4 a0 = pa[0], a1 = pa[1], a2 = pa[2];
            const float2 x0 = make_float2(a0.x, a0.y), x1 = make_float2(a0.z, a0.w), x2 = make_float2(a1.x, a1.y),
                         x3 = make_float2(a1.z, a1.w), x4 = make_float2(a2.x, a2.y), x5 = make_float2(a2.z, a2.w);
            float2 m0 = rr0, m1 = rr1;
            m0 = ffma2(x0, bc(wlo[0]), m0); m1 = ffma2(x0, bc(whi[0]), m1);
            m0 = ffma2(x1, bc(wlo[1]), m0); m1 = ffma2(x1, bc(whi[1]), m1);
            m0 = ffma2(x2, bc(wlo[2]), m0); m1 = ffma2(x2, bc(whi[2]), m1);
            m0 = ffma2(x3, bc(wlo[3]), m0); m1 = ffma2(x3, bc(whi[3]), m1);
            m0 = ffma2(x4, bc(wlo[4]), m0); m1 = ffma2(x4, bc(whi[4]), m1);
            m0 = ffma2(x5, bc(wlo[5]), m0); m1 = ffma2(x5, bc(whi[5]), m1);
            if (AGG == GNNCCA_AGG_MAX) {
              acc0.x = fmaxf(acc0.x, fmaxf(m0.x, 0.f)); acc0.y = fmaxf(acc0.y, fmaxf(m0.y, 0.f));
              acc1.x = fmaxf(acc1.x, fmaxf(m1.x, 0.f)); acc1.y = fmaxf(acc1.y, fmaxf(m1.y, 0.f));
            } else {
              acc0 = fadd2(acc0, relu2(m0)); acc1 = fadd2(acc1, relu2(m1));
            }
          }
          j += 2 * npair;
        }
        if (j < j1) single(j);
        float s0 = agg_comb(AGG, acc0.x, acc0.y), s1 = agg_comb(AGG, acc1.x, acc1.y);
        if (j1 == j0) { s0 = 0.f; s1 = 0.f; }                            // no out-edge: torch_scatter leaves 0
        else if (AGG == GNNCCA_AGG_MEAN) { const float inv = 1.f / (float)(j1 - j0); s0 = s0 * inv; s1 = s1 * inv; }
        if (live) {
          ss[so.h + v * HS + c0] = s0;
          ss[so.h + v * HS + c0 + 16] = s1;
          if (bt.h_dump) {
            bt.h_dump[((int64_t)step * bt.n_nodes + a + v) * FN + c0] = s0;
            bt.h_dump[((int64_t)step * bt.n_nodes + a + v) * FN + c0 + 16] = s1;
          }
        }
      }
    }
    gr.sync();
  }
  gr.sync();   // the next frame overwrites every buffer
}

}  // namespace mpf
}  // namespace gnncca
