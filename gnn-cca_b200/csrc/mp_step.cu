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
// Mapping.  Edges are source-sorted.  Persistent CTAs of 256 threads take contiguous runs of 512-edge tiles; warp w owns the 64-edge
// "chunk" w of every tile (a chunk of the GLOBAL edge numbering) and never synchronises with the other warps after the set-up.
//   inputs   the chunk's src (+ a halo of 4 ids on either side), dst, e (and edge_attr / e0) arrive in a per-warp shared-memory ring as
//            1-D bulk copies (cp.async.bulk + mbarrier complete_tx, lane 0, two tiles ahead); chunks at the ends of the edge range, or on
//            a partition whose phase breaks the 16-byte alignment, are loaded through registers.
//   phase A  thread per edge pair: gathers of the P[src] / Q[dst] rows, edge update and classifier with packed FFMA2 over OUTPUT pairs
//            (weight pair = aligned LDS result, edge value = broadcast scalar), e' written once (STG.128) and staged row-major.
//   phase B  messages on the tensor cores: m^T [32 features x 8 edges] = W_ne^T [32 x 8] . e'^T [8 x 8 edges], two m-tiles.  Operands are
//            split into tf32 prefix + exact rest; hi.hi runs as mma.sync m16n8k8 tf32, the two cross terms hi.lo + lo.hi share ONE
//            m16n8k16 bf16 mma (their factors of size 2^-11 need 8 bits: fp32-level accuracy at 4 instead of 6 tensor instructions
//            per 8 edges -- the loop is bound by the mma.sync issue rate); the two k-slots the FE = 6 edge features leave free carry the R[src] row; ReLU and
//            the running sum / max are applied to the D fragment, lanes folded once per piece (a piece = the edges of one source inside
//            the chunk; boundaries from one ballot over the staged source ids, skipped when the chunk has a single source).
//   - a segment that lies inside one chunk is finished and written by its warp;
//   - a segment that crosses chunks leaves one piece per chunk in the workspace (plain stores), and a second small
//     kernel (warp per node) folds the pieces of each such node in chunk order; it also zero-fills the nodes without
//     out-edges.  (A CTA barrier per tile plus fence / counter hand-over between tiles cost 0.17 of 0.57 ms at 8 x 64.)
// A step without node update (h_out == NULL: the last step of an inference forward) runs a variant without phase B (NODE = false).
// No floating-point atomics anywhere: the summation order is a pure function of (E, rowptr, phase), so results are
// bit-identical run to run and between a partition and the whole graph (the reference's torch_scatter atomicAdd path is not).
#include "common.cuh"
#include "tc_helpers.cuh"

namespace gnncca {

constexpr int TILE = 512;     // edges per tile (the modulus of edge_phase: the reduction tiling of the GLOBAL edge numbering)
#ifndef GNNCCA_STEP_CHUNK
#define GNNCCA_STEP_CHUNK 64
#endif
constexpr int CHUNK = GNNCCA_STEP_CHUNK;   // edges a warp stages and reduces per tile.  64 (8 warps per CTA) measured faster than 128 (4 warps, half
                                           // the per-chunk bookkeeping per edge): 275 vs 312 us per step at cfg 4 (profiles/r02_step_experiments.md)
#ifndef GNNCCA_STEP_UNROLL_SUB
#define GNNCCA_STEP_UNROLL_SUB 1
#endif
constexpr int kUnrollSub = GNNCCA_STEP_UNROLL_SUB;
constexpr int SUB = CHUNK / 64;            // phase A walks a chunk in sub-chunks of 64 edges (a lane's pair per sub-chunk)
constexpr int NWARP = TILE / CHUNK;
constexpr int THREADS = NWARP * 32;
static_assert(CHUNK % 64 == 0 && TILE % CHUNK == 0, "a chunk is a whole number of 64-edge sub-chunks");
constexpr int FN = 32;        // latent node width handled by lane-per-feature phase B
#ifndef GNNCCA_STEP_WARPS_PER_SM
#define GNNCCA_STEP_WARPS_PER_SM 24
#endif
constexpr int STEP_MIN_CTAS = GNNCCA_STEP_WARPS_PER_SM / NWARP;   // resident CTAs per SM the register allocation is held to (24 warps: 85 registers)
constexpr int STEP_MIN_CTAS_EDGE = (GNNCCA_STEP_WARPS_PER_SM + 8) / NWARP;   // the variant without node update: 32 warps (64 registers)

struct StepParams {
  const int32_t* rowptr; const int32_t* src; const int32_t* dst;
  int64_t n_nodes, n_edges;
  const float* P; const float* Q; const float* R;
  const float* e0; const float* e_in; float* e_out; float* h_out; float* logits;
  gnncca_mp_weights w;
  int agg, nf;
  int edge_row;          // row length of w_edge: 2 K + EIN
  int num_tiles;
  float* chunk_part;     // [num_tiles * NWARP][2][FN]  slot 0: piece of a segment begun in an earlier chunk, slot 1: open tail
  int phase;             // position of local edge 0 inside a tile of the GLOBAL edge numbering (edge-partitioned graphs)
  const int32_t* guard;  // optional CSR flags (gnncca_csr_from_edge_index): the kernel is a no-op when any is set
  // ATTR variants: the encoded edge features e0 = ReLU(W_enc attr + b_enc) (mpn.py:116-126, 4 -> FE) are computed here from
  // the raw attributes (16 B/edge) instead of being written by one kernel and read back by this one (2 x 24 B/edge)
  const float* attr; const float* w_enc; const float* b_enc;
  // projections of the NEXT step (gnncca_mp_step_fwd_next): the fold kernel, which visits every node once the step kernel is done, also
  // computes P | Q | R of the updated latents -- the stand-alone node_project launch between two steps disappears.  May alias P, Q, R.
  float* Pn; float* Qn; float* Rn; const float* h0;
  int64_t node_lo, node_hi;      // the fold launch covers [node_lo, node_hi): all nodes, or the owned source range of an edge partition
};


__device__ __forceinline__ void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }
template <bool IS_MAX> __device__ __forceinline__ float comb(float a, float b) { return IS_MAX ? fmaxf(a, b) : a + b; }
template <bool IS_MAX> __device__ __forceinline__ float2 comb2(float2 a, float2 b) {
  return IS_MAX ? make_float2(fmaxf(a.x, b.x), fmaxf(a.y, b.y)) : pk::fadd2(a, b);
}
constexpr int FA = 4;   // raw attribute width of the ATTR variants (one 16-byte load per edge)

template <int FE, bool REATT_E, bool ATTR>
struct RawPair {       // what phase A reads for one edge pair, requested one tile ahead
  int sa, sb, da, db;
  float2 ea[FE / 2], eb[FE / 2];                                  // latent edge features (not read in step 1 of an ATTR variant)
  float2 xa[(REATT_E && !ATTR) ? FE / 2 : 1], xb[(REATT_E && !ATTR) ? FE / 2 : 1];   // stored initial features
  float4 ta[ATTR ? 1 : 1], tb[ATTR ? 1 : 1];                      // raw attributes
};


// ---- per-warp input ring: the raw inputs of a warp's 64-edge chunk travel global -> shared memory as 1-D bulk copies (TMA engine,
// completion on an mbarrier), two tiles ahead of their use.  No registers are held across phase B for them and nobody waits for DRAM.
#ifndef GNNCCA_STEP_STAGES
#define GNNCCA_STEP_STAGES 2
#endif
constexpr int NSTAGE = GNNCCA_STEP_STAGES;
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(enc::smem_u32(dst)), "l"(src), "r"(bytes), "r"(enc::smem_u32(bar)) : "memory");
}
template <int FE, bool REATT_E, bool ATTR> struct Slot {
  // src carries a halo of 4 ids on either side (16-byte granules): the sources of the edges just outside the chunk tell whether its
  // first / last piece belongs to a longer segment
  static constexpr int HALO = 4;
  static constexpr int SRC = 0, DST = (CHUNK + 2 * HALO) * 4, E = DST + CHUNK * 4, X = E + CHUNK * FE * 4;
  static constexpr int BYTES = (X + (ATTR ? CHUNK * FA * 4 : (REATT_E ? CHUNK * FE * 4 : 0)) + 127) & ~127;
};
template <int FE, bool REATT_E, bool ATTR, bool NODE> constexpr int step_dyn_smem() {      // ring | e' rows and sources of the tile (NODE) | mbarriers
  return NWARP * NSTAGE * Slot<FE, REATT_E, ATTR>::BYTES + (NODE ? TILE * FE * 4 + TILE * 4 : 0) + NWARP * NSTAGE * 8;
}

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

// Requests the raw inputs of the chunk [c0, c0 + CHUNK) of `tile` into `slot`.  Interior chunks of a 16-byte aligned layout go as bulk
// copies (returns true: the consumer waits on `bar`); chunks that touch the ends of the edge range, or any chunk of a partition whose
// phase breaks the alignment, are loaded through registers with the invalid edges marked src = -1 (returns false).
template <int FE, bool REATT_E, bool ATTR>
__device__ __forceinline__ bool fill_slot(const StepParams& p, int64_t tile, int c0, int lane, unsigned char* slot, uint64_t* bar, bool bulk_ok) {
  using SL = Slot<FE, REATT_E, ATTR>;
  const int cb = (int)tile * TILE - p.phase + c0;     // global index of the chunk's first edge (E < 2^31 - 2 TILE)
  const int n_edges = (int)p.n_edges;
  const bool latent = !ATTR || p.e_in != nullptr;
  if (bulk_ok && cb >= SL::HALO && cb + CHUNK + SL::HALO <= n_edges) {
    if (lane == 0) {
      // the lanes' reads (and, on the register path, writes) of the slot went through the generic proxy and are ordered before this
      // point by the __syncwarp that precedes the call: the proxy fence orders them before the bulk engine's writes.  (Without it the
      // refill overtook the reads when the data came from L2: run-to-run differences on a 1 M-edge graph.)
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      const uint32_t bytes = 2 * (CHUNK + SL::HALO) * 4 + (latent ? CHUNK * FE * 4 : 0) + (ATTR ? CHUNK * FA * 4 : (REATT_E ? CHUNK * FE * 4 : 0));
      enc::mbar_expect_tx(bar, bytes);
      bulk_load(slot + SL::SRC, p.src + cb - SL::HALO, (CHUNK + 2 * SL::HALO) * 4, bar);
      bulk_load(slot + SL::DST, p.dst + cb, CHUNK * 4, bar);
      if (latent) bulk_load(slot + SL::E, p.e_in + (int64_t)cb * FE, CHUNK * FE * 4, bar);
      if (ATTR) bulk_load(slot + SL::X, p.attr + (int64_t)cb * FA, CHUNK * FA * 4, bar);
      else if (REATT_E) bulk_load(slot + SL::X, p.e0 + (int64_t)cb * FE, CHUNK * FE * 4, bar);
    }
    return true;
  }
  if (lane < 2) {      // the two halo ids next to the chunk (-1: no such edge)
    const int j = lane == 0 ? cb - 1 : cb + CHUNK;
    *reinterpret_cast<int*>(slot + SL::SRC + (lane == 0 ? SL::HALO - 1 : SL::HALO + CHUNK) * 4) = (j >= 0 && j < n_edges) ? p.src[j] : -1;
  }
#pragma unroll 1
  for (int h = 0; h < SUB; ++h) {
    const int pl = 32 * h + lane;      // this lane's pair of the sub-chunk
    RawPair<FE, REATT_E, ATTR> r{};
    load_raw<FE, REATT_E, ATTR>(p, (int64_t)cb + 2 * pl, r);
    *reinterpret_cast<int2*>(slot + SL::SRC + SL::HALO * 4 + 8 * pl) = make_int2(r.sa, r.sb);
    *reinterpret_cast<int2*>(slot + SL::DST + 8 * pl) = make_int2(r.da, r.db);
    float* e = reinterpret_cast<float*>(slot + SL::E) + 2 * FE * pl;
#pragma unroll
    for (int k = 0; k < FE / 2; ++k) {
      *reinterpret_cast<float2*>(e + 2 * k) = r.ea[k];
      *reinterpret_cast<float2*>(e + FE + 2 * k) = r.eb[k];
    }
    if (ATTR) {
      float4* x = reinterpret_cast<float4*>(slot + SL::X) + 2 * pl;
      x[0] = r.ta[0]; x[1] = r.tb[0];
    } else if (REATT_E) {
      float* x = reinterpret_cast<float*>(slot + SL::X) + 2 * FE * pl;
#pragma unroll
      for (int k = 0; k < FE / 2; ++k) {
        *reinterpret_cast<float2*>(x + 2 * k) = r.xa[k];
        *reinterpret_cast<float2*>(x + FE + 2 * k) = r.xb[k];
      }
    }
  }
  __syncwarp();
  return false;
}

// A lane's edge pair out of the slot (the register naming of load_raw)
template <int FE, bool REATT_E, bool ATTR>
__device__ __forceinline__ void read_slot(const unsigned char* slot, int lane, bool latent, RawPair<FE, REATT_E, ATTR>& r) {
  using SL = Slot<FE, REATT_E, ATTR>;
  static_assert((2 * FE) % 4 == 0, "a pair's features are read as float4");
  const int2 s2 = *reinterpret_cast<const int2*>(slot + SL::SRC + SL::HALO * 4 + 8 * lane), d2 = *reinterpret_cast<const int2*>(slot + SL::DST + 8 * lane);
  r.sa = s2.x; r.sb = s2.y; r.da = d2.x; r.db = d2.y;
  if (latent) {
    float t[2 * FE];
#pragma unroll
    for (int q = 0; q < 2 * FE / 4; ++q) {
      const float4 v = *reinterpret_cast<const float4*>(slot + SL::E + (2 * FE * lane + 4 * q) * 4);
      t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < FE / 2; ++k) { r.ea[k] = make_float2(t[2 * k], t[2 * k + 1]); r.eb[k] = make_float2(t[FE + 2 * k], t[FE + 2 * k + 1]); }
  }
  if (ATTR) {
    const float4* x = reinterpret_cast<const float4*>(slot + SL::X) + 2 * lane;
    r.ta[0] = x[0]; r.tb[0] = x[1];
  } else if (REATT_E) {
    float t[2 * FE];
#pragma unroll
    for (int q = 0; q < 2 * FE / 4; ++q) {
      const float4 v = *reinterpret_cast<const float4*>(slot + SL::X + (2 * FE * lane + 4 * q) * 4);
      t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < FE / 2; ++k) { r.xa[k] = make_float2(t[2 * k], t[2 * k + 1]); r.xb[k] = make_float2(t[FE + 2 * k], t[FE + 2 * k + 1]); }
  }
}

// mma.sync m16n8k8 (tf32 operands, fp32 accumulate): d = a . b + c with separate d / c registers, and in place
__device__ __forceinline__ void mma_tf32_dc(float (&d)[4], const float (&a)[4], float b0, float b1, float c0, float c1, float c2, float c3) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
               : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
               : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)), "f"(c0), "f"(c1), "f"(c2), "f"(c3));
}
__device__ __forceinline__ void mma_tf32_acc(float (&d)[4], const float (&a)[4], float b0, float b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}
// mma.sync m16n8k16 (bf16 operands, fp32 accumulate), zero C: d = a . b
__device__ __forceinline__ void mma_bf16_d(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
               : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(0.f));
}
// two fp32 values rounded to bf16 and packed: `lo` in bits 0..15 (the lower k index of an mma fragment register)
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// the tf32 prefix of an fp32 value (what the tensor core reads of it); value - prefix is exact in fp32
__device__ __forceinline__ float tf32_prefix(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

// lane (g, t) = (lane >> 2, lane & 3) holds feature 8 t + g of a node's aggregate
__device__ __forceinline__ void finalize_node_feat(const StepParams& p, int node, float v, int feat) {
  if (p.agg == GNNCCA_AGG_MEAN) v = v / (float)max(p.rowptr[node + 1] - p.rowptr[node], 1);   // scatter_mean: sum / clamp(count, 1)
  p.h_out[(int64_t)node * FN + feat] = v;
}

// NODE = false: the step only updates and classifies the edges (last step of an inference forward, h_out == NULL): no phase B, so
// the register allocation is held to one more resident CTA per SM (this variant is bound by memory-instruction latency)
template <int FE, int HC, bool REATT_E, bool IS_MAX, bool ATTR, bool NODE>
__global__ void __launch_bounds__(THREADS, NODE ? STEP_MIN_CTAS : STEP_MIN_CTAS_EDGE) mp_step_fwd_kernel(const StepParams p) {
  constexpr int EIN = REATT_E ? 2 * FE : FE;
  constexpr int HCS = HC > 0 ? HC : 2;
  constexpr int FEP = (FE + 3) & ~3;      // row stride of the transposed weight tables (16-byte rows)
  static_assert(FE % 2 == 0 && FE <= 6 && HCS % 2 == 0, "outputs are handled as (o, o + 1) pairs; an edge row is one k-step of the mma");
  __shared__ __align__(16) float s_weeT[EIN * FE];           // [k][o] flat (read as float4 chunks): the pair (o, o + 1) is one FFMA2 operand
  __shared__ __align__(16) float4 s_afrag[4][32];             // phase B: W_ne^T as mma A fragments, [tile i][hi | lo] per lane
  __shared__ __align__(16) float s_wencT[ATTR ? FA : 1][FEP];
  __shared__ __align__(8) float s_benc[ATTR ? FE : 2];
  __shared__ __align__(16) float s_c1T[FE][HCS];             // [k][c]
  __shared__ __align__(16) float s_b1[HCS], s_c2[HCS];
  __shared__ float s_b2;
  __shared__ __align__(8) float s_one[2];                    // B operand of the k-slots that carry the R row
  using SL = Slot<FE, REATT_E, ATTR>;
  extern __shared__ __align__(128) unsigned char step_dyn[];
  float (*s_e)[FE] = reinterpret_cast<float (*)[FE]>(step_dyn + NWARP * NSTAGE * SL::BYTES);      // e' of the tile, one row per edge (B operand of phase B)
  int* s_src = reinterpret_cast<int*>(step_dyn + NWARP * NSTAGE * SL::BYTES + TILE * FE * 4);
  uint64_t* s_bar = reinterpret_cast<uint64_t*>(step_dyn + NWARP * NSTAGE * SL::BYTES + (NODE ? TILE * FE * 4 + TILE * 4 : 0));

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // warp-uniform by construction: per-warp bookkeeping runs on the uniform datapath
  const int K = p.nf * FN;
  if (p.guard != nullptr && (p.guard[0] | p.guard[1]) != 0) return;   // CSR not valid (unsorted / out of range): uniform exit
  const bool classify = HC > 0 && p.logits != nullptr;
  {  // stage weights (read from the live parameters at every launch)
    const int edge_row = 2 * K + EIN;
    for (int idx = tid; idx < FE * EIN; idx += THREADS) s_weeT[(idx % EIN) * FE + idx / EIN] = p.w.w_edge[(idx / EIN) * edge_row + 2 * K + idx % EIN];
    if (ATTR) {
      for (int idx = tid; idx < FE * FA; idx += THREADS) s_wencT[idx % FA][idx / FA] = p.w_enc[idx];
      for (int idx = tid; idx < FE; idx += THREADS) s_benc[idx] = p.b_enc[idx];
    }
    if (classify) {
      for (int c = tid; c < HC; c += THREADS) {
        float scale = 1.f, shift = 0.f;
        if (p.w.bn_weight) {  // BatchNorm1d (eval) folded into the hidden layer
          scale = p.w.bn_weight[c] / sqrtf(p.w.bn_var[c] + p.w.bn_eps);
          shift = p.w.bn_bias[c] - p.w.bn_mean[c] * scale;
        }
        for (int k = 0; k < FE; ++k) s_c1T[k][c] = p.w.w_c1[c * FE + k] * scale;
        s_b1[c] = p.w.b_c1[c] * scale + shift;
        s_c2[c] = p.w.w_c2[c];
      }
      if (tid == 0) s_b2 = p.w.b_c2[0];
    }
    if (tid < 2) s_one[tid] = 1.f;
  }
  // interior chunks of a 16-byte aligned layout arrive as bulk copies (uniform over the launch)
  const bool bulk_ok = (p.phase & 3) == 0 && (((uintptr_t)p.e_in | (uintptr_t)p.e0 | (uintptr_t)p.attr | (uintptr_t)p.src | (uintptr_t)p.dst) & 15) == 0;
  const bool vec_out = (p.phase & 1) == 0 && (((uintptr_t)p.e_out) & 15) == 0;
  const bool latent = !ATTR || p.e_in != nullptr;
  const int n_edges = (int)p.n_edges;
  if (tid < NWARP * NSTAGE) enc::mbar_init(&s_bar[tid], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  // phase B operands: the message GEMM runs transposed, m^T [32 features x 8 edges] = W_ne^T [32 x 8] . e'^T [8 x 8 edges], as two
  // m16n8k8 tiles.  k-slot t carries edge feature 2t and k-slot t + 4 feature 2t + 1 (a lane reads ONE float2 of the staged row); the
  // slots beyond FE are zero.  Operands are split into their tf32 prefix and the exact rest: hi.hi in tf32, lo.hi + hi.lo in one bf16 mma.
  // The 16 weight fragment registers of a lane are parked in shared memory ([4][32 lanes] float4: conflict-free LDS.128) and
  // fetched at the start of each phase B, so that they are not live (and spilled) across phase A.
  const int g = lane >> 2, t = lane & 3;
  if (NODE && warp == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      float hi4[4], lo4[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = 2 * t + (q >> 1), feat = 16 * i + g + 8 * (q & 1);
        const float w = k < FE ? p.w.w_node[feat * (K + FE) + K + k] : 0.f;
        hi4[q] = tf32_prefix(w); lo4[q] = w - hi4[q];
      }
      s_afrag[2 * i][lane] = make_float4(hi4[0], hi4[1], hi4[2], hi4[3]);
      // correction operand (bf16 m16n8k16): k-slots 0..7 hold W_hi (paired with the rests e'_lo), 8..15 hold W_lo (paired with e'_hi);
      // a register = (k = 2t, 2t + 1) of one row: q = 0, 2 are row g, q = 1, 3 row g + 8
      s_afrag[2 * i + 1][lane] = make_float4(__uint_as_float(pack_bf16(hi4[0], hi4[2])), __uint_as_float(pack_bf16(hi4[1], hi4[3])),
                                             __uint_as_float(pack_bf16(lo4[0], lo4[2])), __uint_as_float(pack_bf16(lo4[1], lo4[3])));
    }
  }
  __syncthreads();

  // A CTA takes a CONTIGUOUS run of tiles: consecutive tiles gather from the same few hundred Q / P / R rows (one frame's nodes, or the
  // neighbourhood of consecutive sources), which then answer from L1 instead of L2
  const int t_per = p.num_tiles / (int)gridDim.x, t_rem = p.num_tiles % (int)gridDim.x;
  const int t_begin = (int)blockIdx.x * t_per + min((int)blockIdx.x, t_rem), t_end = t_begin + t_per + ((int)blockIdx.x < t_rem ? 1 : 0);
  unsigned char* ring = step_dyn + (size_t)warp * NSTAGE * SL::BYTES;      // this warp's slots
  uint64_t* wbar = s_bar + warp * NSTAGE;
  uint32_t parity = 0, isbulk = 0;      // bit per slot
#pragma unroll
  for (int st = 0; st < NSTAGE; ++st) {
    const int64_t t0 = (int64_t)t_begin + st;
    if (t0 < t_end && fill_slot<FE, REATT_E, ATTR>(p, t0, warp * CHUNK, lane, ring + st * SL::BYTES, wbar + st, bulk_ok)) isbulk |= 1u << st;
  }

#ifdef GNNCCA_EXP_STAGGER   // timing experiment: spread the warps of an SM over one tile period instead of starting them in step
  __nanosleep((((blockIdx.x / 148) % 3) * 8 + warp) * GNNCCA_EXP_STAGGER);
#endif
  int it = 0;
  for (int tile = t_begin; tile < t_end; ++tile, ++it) {
    const int base = tile * TILE - p.phase;   // tiles follow the global numbering: same chunking, same bits, on any partition (E < 2^31 - 2 TILE)
    const int sl = it % NSTAGE;
    const unsigned char* slot = ring + sl * SL::BYTES;
    if ((isbulk >> sl) & 1u) { enc::mbar_wait(wbar + sl, (parity >> sl) & 1u); parity ^= 1u << sl; }
    // sources of the edges just before / after the chunk (-1: none): is its first / last piece part of a longer segment?
    const int nb_prev = *reinterpret_cast<const int*>(slot + SL::SRC + (SL::HALO - 1) * 4);
    const int nb_next = *reinterpret_cast<const int*>(slot + SL::SRC + (SL::HALO + CHUNK) * 4);
    // A warp stages and then reduces the SAME edges (its chunk of the tile), so the warps never wait for each other
    const int lo = max(0, -base), hi = min(TILE, n_edges - base);   // valid tile-local edge range
    const int c0 = warp * CHUNK;
    const int a = max(c0, lo), b = min(c0 + CHUNK, hi);
    float r_next[4] = {0.f, 0.f, 0.f, 0.f};   // R row of the chunk's first segment (features g + 8 q), requested before phase A
    if (NODE && p.h_out != nullptr && a < b) {
      const int s_first = *reinterpret_cast<const int*>(slot + SL::SRC + (SL::HALO + a - c0) * 4);
#pragma unroll
      for (int q = 0; q < 4; ++q) r_next[q] = p.R[(int64_t)s_first * FN + g + 8 * q];
    }
    // ------------------------------ phase A: thread per edge pair, one 64-edge sub-chunk after the other ------------------------------
    // Packed arithmetic runs over OUTPUT pairs (o, o + 1) of one edge: the weight pair comes out of shared memory as an aligned
    // register pair and the edge value is the broadcast scalar (pairing the two EDGES instead cost two register moves per FFMA2).
#pragma unroll kUnrollSub
    for (int h = 0; h < SUB; ++h) {
      const int pl = 32 * h + lane;          // this lane's pair of the chunk
      const int te = c0 + 2 * pl;            // its first edge, tile-local
      const int i0 = base + te;
      RawPair<FE, REATT_E, ATTR> cur;
      read_slot<FE, REATT_E, ATTR>(slot, pl, latent, cur);
      const bool va = cur.sa >= 0, vb = cur.sb >= 0;
      float xa[EIN], xb[EIN];
      constexpr int OFF = REATT_E ? FE : 0;
      float2 e0a[ATTR ? FE / 2 : 1], e0b[ATTR ? FE / 2 : 1];
      if (ATTR) {     // e0 = ReLU(W_enc attr + b_enc) (the FMA order of the stand-alone encoder kernel)
        const float ta[FA] = {cur.ta[0].x, cur.ta[0].y, cur.ta[0].z, cur.ta[0].w}, tb[FA] = {cur.tb[0].x, cur.tb[0].y, cur.tb[0].z, cur.tb[0].w};
#pragma unroll
        for (int j = 0; j < FE / 2; ++j) {
          float2 aa = *reinterpret_cast<const float2*>(&s_benc[2 * j]), ab = aa;
#pragma unroll
          for (int k = 0; k < FA; ++k) {
            const float2 w = *reinterpret_cast<const float2*>(&s_wencT[k][2 * j]);
            aa = pk::ffma2(w, pk::bc(ta[k]), aa); ab = pk::ffma2(w, pk::bc(tb[k]), ab);
          }
          e0a[j] = pk::relu2(aa); e0b[j] = pk::relu2(ab);
        }
      }
      if (REATT_E) {  // cat(initial_edge_feats, latent_edge_feats), mpn.py:283
#pragma unroll
        for (int k = 0; k < FE / 2; ++k) {
          if (ATTR) { xa[2 * k] = e0a[k].x; xa[2 * k + 1] = e0a[k].y; xb[2 * k] = e0b[k].x; xb[2 * k + 1] = e0b[k].y; }
          else { xa[2 * k] = cur.xa[k].x; xa[2 * k + 1] = cur.xa[k].y; xb[2 * k] = cur.xb[k].x; xb[2 * k + 1] = cur.xb[k].y; }
        }
      }
#pragma unroll
      for (int k = 0; k < FE / 2; ++k) {
        if (ATTR && p.e_in == nullptr) {   // step 1: the latent features are the encoded ones
          xa[OFF + 2 * k] = e0a[k].x; xa[OFF + 2 * k + 1] = e0a[k].y; xb[OFF + 2 * k] = e0b[k].x; xb[OFF + 2 * k + 1] = e0b[k].y;
        } else {
          xa[OFF + 2 * k] = cur.ea[k].x; xa[OFF + 2 * k + 1] = cur.ea[k].y; xb[OFF + 2 * k] = cur.eb[k].x; xb[OFF + 2 * k + 1] = cur.eb[k].y;
        }
      }
      float2 ea[FE / 2], eb[FE / 2];
      {
        // invalid edges (src = -1, only in the chunks at the ends of the edge range) read node 0's rows: finite values, never stored
#ifdef GNNCCA_EXP_NOGATHER      // timing experiment (tools/variant.sh): every gather hits one row
        const int sa = 0, sb = 0, da = 0, db = 0;
#else
        const int sa = max(cur.sa, 0), sb = max(cur.sb, 0), da = max(cur.da, 0), db = max(cur.db, 0);
#endif
        const float4 pa0 = *reinterpret_cast<const float4*>(p.P + (int64_t)sa * 8), pa1 = *reinterpret_cast<const float4*>(p.P + (int64_t)sa * 8 + 4);
        const float4 qa0 = *reinterpret_cast<const float4*>(p.Q + (int64_t)da * 8), qa1 = *reinterpret_cast<const float4*>(p.Q + (int64_t)da * 8 + 4);
        float4 pb0 = pa0, pb1 = pa1;      // the pair shares its source except across a segment boundary
        if (sb != sa) { pb0 = *reinterpret_cast<const float4*>(p.P + (int64_t)sb * 8); pb1 = *reinterpret_cast<const float4*>(p.P + (int64_t)sb * 8 + 4); }
        const float4 qb0 = *reinterpret_cast<const float4*>(p.Q + (int64_t)db * 8), qb1 = *reinterpret_cast<const float4*>(p.Q + (int64_t)db * 8 + 4);
        const float2 pa[4] = {make_float2(pa0.x, pa0.y), make_float2(pa0.z, pa0.w), make_float2(pa1.x, pa1.y), make_float2(pa1.z, pa1.w)};
        const float2 qa[4] = {make_float2(qa0.x, qa0.y), make_float2(qa0.z, qa0.w), make_float2(qa1.x, qa1.y), make_float2(qa1.z, qa1.w)};
        const float2 pb[4] = {make_float2(pb0.x, pb0.y), make_float2(pb0.z, pb0.w), make_float2(pb1.x, pb1.y), make_float2(pb1.z, pb1.w)};
        const float2 qb[4] = {make_float2(qb0.x, qb0.y), make_float2(qb0.z, qb0.w), make_float2(qb1.x, qb1.y), make_float2(qb1.z, qb1.w)};
#pragma unroll
        for (int j = 0; j < FE / 2; ++j) { ea[j] = pk::fadd2(pa[j], qa[j]); eb[j] = pk::fadd2(pb[j], qb[j]); }
      }
      static_assert((EIN * FE) % 4 == 0, "the weight table is read as float4");
#pragma unroll
      for (int k2 = 0; k2 < EIN; k2 += 2) {          // two k rows = 2 FE floats = whole float4 chunks
        float wr[2 * FE];
#pragma unroll
        for (int q = 0; q < 2 * FE / 4; ++q) {
          const float4 v = *reinterpret_cast<const float4*>(&s_weeT[k2 * FE + 4 * q]);
          wr[4 * q] = v.x; wr[4 * q + 1] = v.y; wr[4 * q + 2] = v.z; wr[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
#pragma unroll
          for (int j = 0; j < FE / 2; ++j) {
            const float2 w = make_float2(wr[kk * FE + 2 * j], wr[kk * FE + 2 * j + 1]);
            ea[j] = pk::ffma2(w, pk::bc(xa[k2 + kk]), ea[j]);
            eb[j] = pk::ffma2(w, pk::bc(xb[k2 + kk]), eb[j]);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < FE / 2; ++j) { ea[j] = pk::relu2(ea[j]); eb[j] = pk::relu2(eb[j]); }
      float row[2 * FE];          // the pair's two rows, contiguous in e_out and in s_e
#pragma unroll
      for (int j = 0; j < FE / 2; ++j) { row[2 * j] = ea[j].x; row[2 * j + 1] = ea[j].y; row[FE + 2 * j] = eb[j].x; row[FE + 2 * j + 1] = eb[j].y; }
      if (p.e_out != nullptr) {       // NULL in the last step of an inference forward: nothing reads e^L (mpn.py:288-299)
        if (vec_out && va && vb) {
#pragma unroll
          for (int q = 0; q < 2 * FE / 4; ++q) st_stream_f4(p.e_out + (int64_t)i0 * FE + 4 * q, make_float4(row[4 * q], row[4 * q + 1], row[4 * q + 2], row[4 * q + 3]));
        } else {
          if (va) {
#pragma unroll
            for (int j = 0; j < FE / 2; ++j) st_stream_f2(p.e_out + (int64_t)i0 * FE + 2 * j, ea[j]);
          }
          if (vb) {
#pragma unroll
            for (int j = 0; j < FE / 2; ++j) st_stream_f2(p.e_out + (int64_t)(i0 + 1) * FE + 2 * j, eb[j]);
          }
        }
      }
      if (NODE) {
#pragma unroll
        for (int q = 0; q < 2 * FE / 4; ++q)
          *reinterpret_cast<float4*>(&s_e[te][0] + 4 * q) = make_float4(row[4 * q], row[4 * q + 1], row[4 * q + 2], row[4 * q + 3]);
        *reinterpret_cast<int2*>(&s_src[te]) = make_int2(cur.sa, cur.sb);
      }
      if (classify) {
        float2 ha[HCS / 2], hb[HCS / 2];
#pragma unroll
        for (int c = 0; c < HCS / 2; ++c) ha[c] = hb[c] = *reinterpret_cast<const float2*>(&s_b1[2 * c]);
#pragma unroll
        for (int k = 0; k < FE; ++k) {
#pragma unroll
          for (int c = 0; c < HCS / 2; ++c) {
            const float2 w = *reinterpret_cast<const float2*>(&s_c1T[k][2 * c]);
            ha[c] = pk::ffma2(w, pk::bc(row[k]), ha[c]);
            hb[c] = pk::ffma2(w, pk::bc(row[FE + k]), hb[c]);
          }
        }
        float la = s_b2, lb = la;
#pragma unroll
        for (int c = 0; c < HCS / 2; ++c) {
          const float2 w = *reinterpret_cast<const float2*>(&s_c2[2 * c]);
          la = fmaf(fmaxf(ha[c].x, 0.f), w.x, la); la = fmaf(fmaxf(ha[c].y, 0.f), w.y, la);
          lb = fmaf(fmaxf(hb[c].x, 0.f), w.x, lb); lb = fmaf(fmaxf(hb[c].y, 0.f), w.y, lb);
        }
        if (vec_out && va && vb && (((uintptr_t)p.logits) & 7) == 0) st_stream_f2(p.logits + i0, make_float2(la, lb));
        else {
          if (va) st_stream_f1(p.logits + i0, la);
          if (vb) st_stream_f1(p.logits + i0 + 1, lb);
        }
      }
    }
    __syncwarp();        // every lane has read its pairs: the slot goes to the tile NSTAGE rounds ahead (and the staged rows are visible)
    {
      const int64_t nt = (int64_t)tile + NSTAGE;
      isbulk &= ~(1u << sl);
      if (nt < t_end && fill_slot<FE, REATT_E, ATTR>(p, nt, warp * CHUNK, lane, ring + sl * SL::BYTES, wbar + sl, bulk_ok)) isbulk |= 1u << sl;
    }
    if (!NODE || p.h_out == nullptr) continue;  // last step: the node update is dead (mpn.py:288-299); no shared state reused
#ifdef GNNCCA_EXP_NOB
    continue;
#endif
    // ------------------------------ phase B: messages of 8 edges per mma ------------------------------
    // The chunk is cut into 8-edge blocks aligned to the global numbering.  A piece (edges of one source inside the chunk) walks
    // its blocks; the accumulator tile starts as the R row of the source (C operand), ReLU and the running aggregate are applied
    // to the D fragment (lane (g, t): features g + 8 q of edges 2t, 2t + 1), edges of a block outside the piece are zeroed (a
    // ReLU output is >= 0, so 0 is neutral for sum and max alike).  Lanes are folded once per piece.
    {
      if (a < b) {
        uint32_t a_c[2][4];      // correction fragments (bf16 pairs)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float4 l4 = s_afrag[2 * i + 1][lane];
          a_c[i][0] = __float_as_uint(l4.x); a_c[i][1] = __float_as_uint(l4.y); a_c[i][2] = __float_as_uint(l4.z); a_c[i][3] = __float_as_uint(l4.w);
        }
        // bit i of m0 (m1): edge c0 + i (c0 + 64 + i) opens a piece; sorted sources: equal ends = one piece
        static_assert(SUB <= 2, "the piece mask is two 64-bit words");
        unsigned long long m0 = (a - c0) < 64 ? 1ull << (a - c0) : 0ull, m1 = (a - c0) >= 64 ? 1ull << (a - c0 - 64) : 0ull;
        if (s_src[a] != s_src[b - 1]) {
#pragma unroll
          for (int w = 0; w < SUB; ++w) {
            const int ie = c0 + 64 * w + 2 * lane;
            const int2 my = *reinterpret_cast<const int2*>(&s_src[ie]);
            const int prev = ie - 1 >= a ? s_src[max(ie - 1, 0)] : nb_prev;
            const bool st_e = ie >= a && ie < b && (ie == a || my.x != prev);
            const bool st_o = ie + 1 >= a && ie + 1 < b && (ie + 1 == a || my.y != my.x);
            const unsigned bits = (st_e ? 1u : 0u) | (st_o ? 2u : 0u);
            const unsigned m_lo = __reduce_or_sync(0xffffffffu, lane < 16 ? bits << (2 * lane) : 0u);
            const unsigned m_hi = __reduce_or_sync(0xffffffffu, lane >= 16 ? bits << (2 * lane - 32) : 0u);
            const unsigned long long m = (unsigned long long)m_lo | ((unsigned long long)m_hi << 32);
            if (w == 0) m0 = m; else m1 = m;
          }
        }
        const bool first_cont = nb_prev == s_src[a];
        const bool last_cont = nb_next == s_src[b - 1];
        // chunk-local index of the first piece start still in the mask, -1 when none is left
#define GNNCCA_LOWEST() (m0 ? __ffsll((long long)m0) - 1 : (m1 ? 64 + __ffsll((long long)m1) - 1 : -1))
        int nxt = GNNCCA_LOWEST();
        while (nxt >= 0) {
          const int j0 = c0 + nxt;
          if (nxt < 64) m0 &= m0 - 1; else m1 &= m1 - 1;
          nxt = GNNCCA_LOWEST();
          const bool more = nxt >= 0;
          const int j1 = more ? c0 + nxt : b;
          const int s = s_src[j0];
          float r[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) r[q] = r_next[q];
          if (more) {   // next segment's row: in flight during this one
            const int sn = s_src[j1];
#pragma unroll
            for (int q = 0; q < 4; ++q) r_next[q] = p.R[(int64_t)sn * FN + g + 8 * q];
          }
          float acc[2][4];
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][q] = 0.f;
          // The R row of the source rides in the two k-slots the FE edge features leave free (k = 3 and 7, the t == 3 lanes): there the A
          // fragment holds R split into tf32 prefix + rest and the B fragment holds 1.0, so hi.hi adds R_hi + R_lo; lo.hi sees a zero A,
          // hi.lo a zero B (1 - prefix(1)).  An edge of the block outside the piece gets an all-zero B column (the 1.0 included): its
          // messages come out as relu(0) = 0, neutral for sum and max.
          float a_hi[2][4];          // whole fragments are selected per piece, so that each lives in one aligned register quad
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const float4 h4 = s_afrag[2 * i][lane];
            const float h0 = tf32_prefix(r[2 * i]), h1 = tf32_prefix(r[2 * i + 1]);
            a_hi[i][0] = t == 3 ? h0 : h4.x; a_hi[i][1] = t == 3 ? h1 : h4.y;
            a_hi[i][2] = t == 3 ? r[2 * i] - h0 : h4.z; a_hi[i][3] = t == 3 ? r[2 * i + 1] - h1 : h4.w;
          }
          const int kb0 = j0 & ~7;
          const unsigned span = (unsigned)(j1 - j0);
          unsigned x = (unsigned)(kb0 + g - j0);                 // edge kb + g is in the piece iff x < span (unsigned)
          const float2* bp = 2 * t < FE ? reinterpret_cast<const float2*>(&s_e[kb0 + g][2 * t]) : reinterpret_cast<const float2*>(s_one);
          int bstep = 2 * t < FE ? 8 * FE * 4 : 0;               // bytes per block
          asm volatile("" : "+r"(bstep));                        // a register stride: one add per block instead of a re-selected constant
#pragma unroll 2
#ifdef GNNCCA_EXP_NOMSG
          if (false)
#endif
          for (int kb = kb0; kb < j1; kb += 8, x += 8, bp = reinterpret_cast<const float2*>(reinterpret_cast<const char*>(bp) + bstep)) {
            float2 v = make_float2(0.f, 0.f);
            if (x < span) v = *bp;
            const float bh0 = tf32_prefix(v.x), bh1 = tf32_prefix(v.y);
            const uint32_t cb0 = pack_bf16(v.x - bh0, v.y - bh1), cb1 = pack_bf16(bh0, bh1);      // (e'_lo pair | e'_hi pair) of features 2t, 2t + 1
            float d[2][4];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
#ifdef GNNCCA_EXP_1MMA
              mma_tf32_dc(d[i], a_hi[i], bh0, bh1, __uint_as_float(cb0), __uint_as_float(cb1), 0.f, 0.f);
#else
              mma_bf16_d(d[i], a_c[i], cb0, cb1);              // W_hi . e'_lo + W_lo . e'_hi
              mma_tf32_acc(d[i], a_hi[i], bh0, bh1);           // + W_hi . e'_hi (+ the R row)
#endif
            }
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              if (IS_MAX) {
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[i][q] = fmaxf(acc[i][q], d[i][q]);    // max(acc, relu(d)) with acc >= 0
              } else {
#ifdef GNNCCA_RELU_FMNMX
                const float2 s01 = pk::fadd2(make_float2(acc[i][0], acc[i][1]), pk::relu2(make_float2(d[i][0], d[i][1])));
                const float2 s23 = pk::fadd2(make_float2(acc[i][2], acc[i][3]), pk::relu2(make_float2(d[i][2], d[i][3])));
#else
                // ReLU on the FMA pipe: d + |d| = 2 relu(d) exactly (FADD with an |.| operand; FMNMX issues on the half-rate ALU pipe, which
                // bounds this loop).  The running sums are therefore 2 x the aggregate; the factor is removed (exactly) when the piece is folded.
                const float2 s01 = pk::fadd2(make_float2(acc[i][0], acc[i][1]), make_float2(d[i][0] + fabsf(d[i][0]), d[i][1] + fabsf(d[i][1])));
                const float2 s23 = pk::fadd2(make_float2(acc[i][2], acc[i][3]), make_float2(d[i][2] + fabsf(d[i][2]), d[i][3] + fabsf(d[i][3])));
#endif
                acc[i][0] = s01.x; acc[i][1] = s01.y; acc[i][2] = s23.x; acc[i][3] = s23.y;
              }
            }
          }
          float v4[4];     // feature g + 8 q: even edges then odd edges, then the four t lanes by two exchanges
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float x = comb<IS_MAX>(acc[q >> 1][2 * (q & 1)], acc[q >> 1][2 * (q & 1) + 1]);
            x = comb<IS_MAX>(x, __shfl_xor_sync(0xffffffffu, x, 1));
            x = comb<IS_MAX>(x, __shfl_xor_sync(0xffffffffu, x, 2));
#ifndef GNNCCA_RELU_FMNMX
            if (!IS_MAX) x *= 0.5f;
#endif
            v4[q] = x;
          }
          const float mine = t == 0 ? v4[0] : t == 1 ? v4[1] : t == 2 ? v4[2] : v4[3];     // lane (g, t) writes feature g + 8 t
          const bool starts = !(j0 == a && first_cont), ends = !(j1 == b && last_cont);
          if (starts && ends) finalize_node_feat(p, s, mine, g + 8 * t);
          else p.chunk_part[(((int64_t)tile * NWARP + warp) * 2 + (starts ? 1 : 0)) * FN + g + 8 * t] = mine;   // folded by mp_step_fold_kernel
        }
      }
    }
    __syncwarp();      // the warp's rows of s_e / s_src are rewritten by its next phase A
  }
}

// Second kernel of a step: one warp per node.  Nodes without out-edges aggregate to 0 (torch_scatter semantics); a node whose
// segment crosses chunk boundaries gets the fold of its pieces in chunk order (first chunk's open tail, then one piece per
// later chunk); everything else was written by the step kernel.
// PROJECT: the warp then computes the per-node projections of the next step from the node's final latents (same arithmetic and FMA
// order as node_project_kernel: acc = bias, k ascending): lane o produces output o and, for o < 12, output 32 + o of P (6) | Q (6) | R (32).
template <bool PROJECT>
__global__ void __launch_bounds__(256) mp_step_fold_kernel(const StepParams p) {
  constexpr int FE = 6, OUT = 2 * FE + FN;
  extern __shared__ float fold_smem[];           // PROJECT: wt [K][OUT] | bias [OUT]
  const int lane = threadIdx.x & 31;
  if (p.guard != nullptr && (p.guard[0] | p.guard[1]) != 0) return;
  const int K = p.nf * FN;
  float* wt = fold_smem;
  float* bias = fold_smem + K * OUT;
  if (PROJECT) {
    const int edge_row = p.edge_row, node_row = K + FE;
    for (int idx = threadIdx.x; idx < K * OUT; idx += blockDim.x) {       // k fastest: coalesced reads of the weight rows
      const int o = idx / K, k = idx - o * K;
      float v;
      if (o < FE) v = p.w.w_edge[o * edge_row + k];                       // source block
      else if (o < 2 * FE) v = p.w.w_edge[(o - FE) * edge_row + K + k];   // target block
      else v = p.w.w_node[(o - 2 * FE) * node_row + k];                   // node block of the node model
      wt[k * OUT + o] = v;
    }
    for (int o = threadIdx.x; o < OUT; o += blockDim.x) bias[o] = (o < FE) ? p.w.b_edge[o] : (o < 2 * FE ? 0.f : p.w.b_node[o - 2 * FE]);
    __syncthreads();
  }
  const bool is_max = p.agg == GNNCCA_AGG_MAX;
  const int64_t wstride = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t v = p.node_lo + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); v < p.node_hi; v += wstride) {
    const int rs = p.rowptr[v], re = p.rowptr[v + 1];
    float out;
    bool written = false;       // does h_out[v] already hold the value (written by the step kernel)?
    if (re <= rs) out = 0.f;
    else {
      const int64_t c0 = ((int64_t)rs + p.phase) / CHUNK, c1 = ((int64_t)re - 1 + p.phase) / CHUNK;
      if (c0 == c1) { written = true; out = PROJECT ? p.h_out[v * FN + lane] : 0.f; }
      else {
        // sums are carried in double across the pieces (a handful of adds per node): the rounding of a long segment is then that
        // of its 64-edge chunks alone
        double tot = p.chunk_part[(c0 * 2 + 1) * FN + lane];
        for (int64_t c = c0 + 1; c <= c1; c += 8) {            // up to eight loads in flight, folded in order
          float pv[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) pv[u] = c + u <= c1 ? p.chunk_part[((c + u) * 2) * FN + lane] : (is_max ? -INFINITY : 0.f);
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            if (c + u <= c1) tot = is_max ? fmax(tot, (double)pv[u]) : tot + pv[u];
          }
        }
        out = (float)tot;
        if (p.agg == GNNCCA_AGG_MEAN) out = out / (float)(re - rs);      // scatter_mean: sum / count, in fp32 like finalize_node
      }
    }
    if (!written) p.h_out[v * FN + lane] = out;
    if (PROJECT) {
      const float h0v = p.nf == 2 ? p.h0[v * FN + lane] : 0.f;           // cat(initial, current), mpn.py:285
      float acc0 = bias[lane], acc1 = lane < OUT - 32 ? bias[32 + lane] : 0.f;
      for (int k = 0; k < K; ++k) {
        const float hk = (p.nf == 2 && k < FN) ? __shfl_sync(0xffffffffu, h0v, k) : __shfl_sync(0xffffffffu, out, k & (FN - 1));
        acc0 = fmaf(hk, wt[k * OUT + lane], acc0);
        if (lane < OUT - 32) acc1 = fmaf(hk, wt[k * OUT + 32 + lane], acc1);
      }
      // outputs 0..5 P, 6..11 Q, 12..43 R; the two padding lanes of the 8-float P / Q rows are zeroed
      if (lane < FE) p.Pn[v * 8 + lane] = acc0;
      else if (lane < 2 * FE) p.Qn[v * 8 + lane - FE] = acc0;
      else p.Rn[v * FN + lane - 2 * FE] = acc0;
      if (lane < OUT - 32) p.Rn[v * FN + 32 - 2 * FE + lane] = acc1;
      if (lane >= 30) { p.Pn[v * 8 + lane - 24] = 0.f; p.Qn[v * 8 + lane - 24] = 0.f; }
    }
  }
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
  const bool node = sp.h_out != nullptr;
  const int grid = (int)min((int64_t)sp.num_tiles, (int64_t)kNumSMs * (node ? STEP_MIN_CTAS : STEP_MIN_CTAS_EDGE));
  const bool is_max = sp.agg == GNNCCA_AGG_MAX;
  const bool attr = sp.attr != nullptr;
#define GNNCCA_STEP(R, M, A) GNNCCA_STEP_N(R, M, A, true)
#define GNNCCA_STEP_N(R, M, A, N)                                                                                           \
  do {                                                                                                                    \
    constexpr int smem = step_dyn_smem<FE, R, A, N>();                                                                     \
    static bool optin[16] = {};          /* static + dynamic shared memory exceed 48 KB: opt in once per device */         \
    int dev = 0; cudaGetDevice(&dev);                                                                                     \
    if (!optin[dev & 15]) {                                                                                               \
      const int rc = check_cuda(cudaFuncSetAttribute(mp_step_fwd_kernel<FE, HC, R, M, A, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
      if (rc) return rc;                                                                                                  \
      optin[dev & 15] = true;                                                                                             \
    }                                                                                                                     \
    mp_step_fwd_kernel<FE, HC, R, M, A, N><<<grid, THREADS, smem, st>>>(sp);                                              \
  } while (0)
  if (!node) {        // the aggregation kind is irrelevant
    if (reatt_e) { if (attr) GNNCCA_STEP_N(true, false, true, false); else GNNCCA_STEP_N(true, false, false, false); }
    else { if (attr) GNNCCA_STEP_N(false, false, true, false); else GNNCCA_STEP_N(false, false, false, false); }
  } else if (reatt_e) {
    if (is_max) { if (attr) GNNCCA_STEP(true, true, true); else GNNCCA_STEP(true, true, false); }
    else { if (attr) GNNCCA_STEP(true, false, true); else GNNCCA_STEP(true, false, false); }
  } else {
    if (is_max) { if (attr) GNNCCA_STEP(false, true, true); else GNNCCA_STEP(false, true, false); }
    else { if (attr) GNNCCA_STEP(false, false, true); else GNNCCA_STEP(false, false, false); }
  }
#undef GNNCCA_STEP
#undef GNNCCA_STEP_N
  if (sp.h_out == nullptr) return check_launch();
  if (sp.Pn != nullptr) {     // fold + the next step's projections: weights staged once per block, blocks loop over the nodes
    const int K = sp.nf * FN;
    const size_t smem = sizeof(float) * (K * (2 * FE + FN) + (2 * FE + FN));
    const unsigned blocks = (unsigned)max((int64_t)1, min(ceil_div(sp.node_hi - sp.node_lo, (int64_t)8), (int64_t)kNumSMs * 8));
    mp_step_fold_kernel<true><<<blocks, 256, smem, st>>>(sp);
  } else {
    mp_step_fold_kernel<false><<<(unsigned)max((int64_t)1, ceil_div(sp.node_hi - sp.node_lo, (int64_t)8)), 256, 0, st>>>(sp);
  }
  return check_launch(2);
}

extern "C" int32_t gnncca_mp_step_tile(void) { return TILE; }

static int mp_step_impl(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr,
                        const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P,
                        const float* Q, const float* R, const float* e0, const float* e_in, float* e_out,
                        float* h_out, float* logits, const int32_t* guard_flags, void* workspace, int64_t workspace_bytes, void* stream,
                        const float* attr, const float* w_enc, const float* b_enc,
                        const float* h0 = nullptr, float* Pn = nullptr, float* Qn = nullptr, float* Rn = nullptr) {
  if (!cfg || !w || n_nodes < 0 || n_edges < 0 || n_edges >= INT32_MAX - 2 * TILE) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const bool next = Pn != nullptr || Qn != nullptr || Rn != nullptr;
  if (next && (!Pn || !Qn || !Rn || !h_out || (cfg->reattach_nodes && !h0) || ((((uintptr_t)Pn | (uintptr_t)Qn) & 15) != 0))) return GNNCCA_ERR_BAD_ARG;
  if (n_edges == 0) {          // nodes without out-edges aggregate to 0 (torch_scatter semantics)
    if (h_out && n_nodes > 0) {
      const int rc = check_cuda(cudaMemsetAsync(h_out, 0, sizeof(float) * n_nodes * FN, st));
      if (rc || !next) return rc;
      return gnncca_node_project(cfg, w, h0, h_out, n_nodes, Pn, Qn, Rn, stream);
    }
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
  sp.Pn = Pn; sp.Qn = Qn; sp.Rn = Rn; sp.h0 = h0;
  sp.w = *w; sp.agg = cfg->agg; sp.nf = cfg->reattach_nodes ? 2 : 1;
  sp.edge_row = 2 * sp.nf * FN + (cfg->reattach_edges ? 2 : 1) * cfg->f_edge;
  if (cfg->edge_phase < 0 || cfg->edge_phase >= TILE) return GNNCCA_ERR_BAD_ARG;
  sp.phase = cfg->edge_phase;
  sp.node_lo = 0; sp.node_hi = n_nodes;
  if (cfg->node_begin != 0 || cfg->node_end != 0) {
    if (cfg->node_begin < 0 || cfg->node_end < cfg->node_begin || cfg->node_end > n_nodes) return GNNCCA_ERR_BAD_ARG;
    sp.node_lo = cfg->node_begin; sp.node_hi = cfg->node_end;
  }
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

extern "C" int gnncca_mp_step_fwd_next(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, const int32_t* src,
                                       const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P, const float* Q, const float* R,
                                       const float* edge_attr, int32_t edge_in_dim, const float* w_enc_edge, const float* b_enc_edge,
                                       const float* e0, const float* e_in, float* e_out, float* h_out, float* logits, const float* h0,
                                       float* P_next, float* Q_next, float* R_next, const int32_t* guard_flags, void* workspace,
                                       int64_t workspace_bytes, void* stream) {
  if (edge_attr != nullptr && edge_in_dim != FA) return GNNCCA_ERR_UNSUPPORTED;
  if (!h_out || !P_next || !Q_next || !R_next) return GNNCCA_ERR_BAD_ARG;
  return mp_step_impl(cfg, w, rowptr, src, dst, n_nodes, n_edges, P, Q, R, edge_attr ? nullptr : e0, e_in, e_out, h_out, logits, guard_flags,
                      workspace, workspace_bytes, stream, edge_attr, w_enc_edge, b_enc_edge, h0, P_next, Q_next, R_next);
}
