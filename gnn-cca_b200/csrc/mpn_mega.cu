// MOTMPNet.forward (reference models/mpn.py:250-299) for batches of small frame graphs as ONE persistent kernel: the
// node encoder (mpn.py:128-142, mlp.py:12-23) streams x through TMA + tcgen05 while, on the same SM, warp groups run
// all message-passing steps (mp_frame.cuh) of the frames whose nodes are already encoded.  The encoder is bound by
// HBM (8 KB per node, no arithmetic on the CUDA cores), message passing by FP32 issue slots (no HBM traffic to
// speak of): run back to back they leave half the machine idle each; interleaved per SM they overlap.
//
// One CTA per SM owns a contiguous range of frames and cuts it into CHUNKS of whole frames with <= 128 nodes (one
// M=128 MMA tile).  Roles:
//   warp 0      TMA producer   x rows of the chunk (one box sized to the chunk) + the W1 k-block, ring of stages
//   warp 1      MMA issuer     tcgen05.mma kind::tf32 into one of two TMEM accumulators
//   warps 2-5   encoder epilogue  ReLU(acc + b1) -> TMEM -> second tcgen05.mma against W2 -> + b2, ReLU -> h rows
//                               (global, L2 resident), then "chunk c encoded" (mbarrier)
//   warps 6..   G message-passing groups of 4 warps: wait for the chunk, take its frames round-robin, run all L steps
//               of a frame out of shared memory, write the logits
// Edge list in RAW form only (caller's int64 edge_index + per-frame edge offsets); violations raise flags[2].
#include "tc_helpers.cuh"
#include "mp_frame.cuh"

namespace gnncca {
namespace mega {

using namespace enc;

constexpr int MAXC = 64;           // chunks per CTA
constexpr int MAXF = 512;          // frames per CTA
constexpr int ENC_THREADS = 192;   // warps 0-5
constexpr int GROUP_THREADS = 128;
constexpr int MAX_GROUPS = 4;
constexpr int MAX_THREADS = ENC_THREADS + MAX_GROUPS * GROUP_THREADS;
constexpr int EPI_BAR = 9;         // named barrier of the four epilogue warps (1..8 belong to the groups)

struct Maps { CUtensorMap x[16]; CUtensorMap w; };   // x with boxes of 8, 16, ..., 128 rows x 32 floats; W1 [H][K]

struct Params {
  mpf::Net net;
  mpf::Batch batch;            // batch.h_enc = h_out
  const float* b1; const float* W2; const float* b2;
  float* h_out;                // [N][32] scratch for the encoded nodes
  int K, stages, groups;
  int frames_base, frames_rem; // CTA c owns frames_base + (c < frames_rem) frames
  unsigned long long* trace;   // development aid, normally NULL
};

// shared-memory layout in bytes from the 1024-aligned base
struct Layout { uint32_t tiles, w2s, b1s, b2s, bars, chunks, gp, mpw, mps, total; };
__host__ __device__ inline Layout layout(int H, int stages, int mp_weight_floats, int mp_state_floats, int groups) {
  Layout l; uint32_t p = 0;
  l.tiles = p; p += (uint32_t)stages * (uint32_t)(BLOCK_M + H) * BLOCK_K * 4;
  l.w2s = p; p += (uint32_t)H * F_OUT * 4;
  l.b1s = p; p += (uint32_t)H * 4;
  l.b2s = p; p += F_OUT * 4;
  l.bars = p; p += 8 * (2 * MAX_STAGES + 2 + 2 + 1 + MAXC) + 16;   // full, empty, tmem_full[2], tmem_empty[2], d2_full, h_ready[MAXC], tmem slot, n_chunks
  p = (p + 15) & ~15u;
  l.chunks = p; p += MAXC * 16;
  l.gp = p; p += (MAXF + 1) * 4;
  p = (p + 15) & ~15u;
  l.mpw = p; p += (uint32_t)mp_weight_floats * 4;
  l.mps = p; p += (uint32_t)mp_state_floats * 4 * (uint32_t)groups;
  l.total = p;
  return l;
}

__device__ __forceinline__ void stamp(const Params& p, int slot) {
#ifdef GNNCCA_MEGA_TRACE
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  if (p.trace) p.trace[(size_t)blockIdx.x * 16 + slot] = t;
#endif
}

template <int H, bool REATT_E, int AGG>
__global__ void __launch_bounds__(MAX_THREADS, 1)
mpn_mega_kernel(const __grid_constant__ Maps maps, const __grid_constant__ Params p) {
  constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 4;
  constexpr uint32_t B_BYTES = H * BLOCK_K * 4;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = 512;                        // 2 x H layer-1 accumulators + 32 layer-2 columns
  static_assert(2 * H + F_OUT <= 512, "hidden width too large");
  constexpr uint32_t IDESC = make_idesc(BLOCK_M, H), IDESC2 = make_idesc(BLOCK_M, F_OUT);
  constexpr uint64_t HINT_EVICT_FIRST = 0x12F0000000000000ull, HINT_EVICT_LAST = 0x14F0000000000000ull;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int K_MP = p.net.reatt_n ? 2 * mpf::FN : mpf::FN, EIN = REATT_E ? 2 * mpf::FE : mpf::FE;
  const mpf::WOff wo = mpf::weight_layout(K_MP, EIN);
  const mpf::SOff so = mpf::state_layout(p.net.cap_n, p.net.cap_e, p.net.reatt_n != 0, REATT_E);
  const Layout lay = layout(H, p.stages, wo.total, so.total, p.groups);
  const int NS = p.stages;
  uint8_t* tiles = smem + lay.tiles;
  float* w2s = (float*)(smem + lay.w2s);
  float* b1s = (float*)(smem + lay.b1s);
  float* b2s = (float*)(smem + lay.b2s);
  uint64_t* full = (uint64_t*)(smem + lay.bars);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tmem_full = empty + MAX_STAGES;      // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint64_t* d2_full = tmem_empty + 2;
  uint64_t* h_ready = d2_full + 1;               // [MAXC]
  uint32_t* tmem_slot = (uint32_t*)(h_ready + MAXC);
  int* n_chunks_s = (int*)(tmem_slot + 1);
  int4* chunk_s = (int4*)(smem + lay.chunks);    // {first row (global), rows, first frame (global), frames}
  int* gp_s = (int*)(smem + lay.gp);             // node offsets of this CTA's frames relative to its first row
  float* mpw = (float*)(smem + lay.mpw);
  float* mps = (float*)(smem + lay.mps);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) stamp(p, 0);
  const int cta = blockIdx.x;
  const int64_t f0 = (int64_t)cta * p.frames_base + min(cta, p.frames_rem);
  const int nf = p.frames_base + (cta < p.frames_rem ? 1 : 0);
  const int KB = p.K / BLOCK_K;

  // ---------------------------------------------------------------- producer state (warp 0, lane 0)
  int pr_s = 0, pr_c = 0, pr_kb = 0; uint32_t pr_ph = 0;
  auto produce = [&](int count, int n_chunks) {
    for (; count > 0 && pr_c < n_chunks; --count) {
      const int4 ch = chunk_s[pr_c];
      const int rows8 = max(8, (ch.y + 7) & ~7);
      mbar_wait(empty + pr_s, pr_ph ^ 1);
      uint8_t* a = tiles + (size_t)pr_s * STAGE_BYTES;
      mbar_expect_tx(full + pr_s, (uint32_t)rows8 * (BLOCK_K * 4) + B_BYTES);
      tma_load_2d(a, &maps.x[rows8 / 8 - 1], full + pr_s, pr_kb * BLOCK_K, ch.x, HINT_EVICT_FIRST);
      tma_load_2d(a + A_BYTES, &maps.w, full + pr_s, pr_kb * BLOCK_K, 0, HINT_EVICT_LAST);
      if (++pr_s == NS) { pr_s = 0; pr_ph ^= 1; }
      if (++pr_kb == KB) { pr_kb = 0; ++pr_c; }
    }
  };

  // ---------------------------------------------------------------- set-up
  if (warp == 0) {
    if (lane < 17) {
      const CUtensorMap* m = lane < 16 ? &maps.x[lane] : &maps.w;
      asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
    }
    const int64_t base = nf > 0 ? p.batch.graph_ptr[f0] : 0;
    bool bad = false;
    for (int i = lane; i <= nf; i += 32) {
      const int64_t d = p.batch.graph_ptr[f0 + i] - base;
      bad |= d < 0 || d > (int64_t)1 << 30;
      gp_s[i] = (int)d;
    }
    bad = __any_sync(0xffffffffu, bad) || base < 0 || base > p.batch.n_nodes;
    __syncwarp();
    if (lane == 0) {
      // chunks: whole frames, <= 128 nodes, about equally many nodes each
      int nc = 0;
      if (!bad && nf > 0) {
        const int total = gp_s[nf];
        const int want = max(1, (total + BLOCK_M - 1) / BLOCK_M);
        const int target = (total + want - 1) / want;
        int cf = 0, cr = 0;
        for (int f = 0; f < nf && !bad; ++f) {
          const int r = gp_s[f + 1] - gp_s[f];
          if (r < 0 || r > BLOCK_M) { bad = true; break; }
          if (f > cf && (cr + r > BLOCK_M || cr >= target)) {
            if (nc == MAXC) { bad = true; break; }
            chunk_s[nc++] = make_int4((int)base + gp_s[cf], cr, (int)f0 + cf, f - cf);
            cf = f; cr = 0;
          }
          cr += r;
        }
        if (!bad) {
          if (nc == MAXC) bad = true;
          else chunk_s[nc++] = make_int4((int)base + gp_s[cf], cr, (int)f0 + cf, nf - cf);
        }
      }
      if (bad) { nc = 0; atomicOr(p.batch.flags + 2, 1); }
      *n_chunks_s = nc;
      for (int s = 0; s < NS; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
      for (int a = 0; a < 2; ++a) { mbar_init(tmem_full + a, 1); mbar_init(tmem_empty + a, 4); }
      mbar_init(d2_full, 1);
      for (int c = 0; c < MAXC; ++c) mbar_init(h_ready + c, 128);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      produce(NS, nc);                       // x is streaming while the other warps set up
    }
  } else if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else if (warp < 6) {   // second encoder layer: coalesced reads of W2 [32][H], written as a swizzled K-major B operand
    const int t = threadIdx.x - 64;
    constexpr int PER = H * F_OUT / 128;
    float wv[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) wv[i] = __ldg(p.W2 + t + i * 128);
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int idx = t + i * 128;
      const int n = idx / H, k = idx - n * H;
      const int kb = k >> 5, kk = k & 31;
      w2s[kb * (F_OUT * 32) + (n >> 3) * 256 + (n & 7) * 32 + (((kk >> 2) ^ (n & 7)) << 2) + (kk & 3)] = tf32_rna(wv[i]);
    }
    for (int idx = t; idx < H; idx += 128) b1s[idx] = p.b1[idx];
    if (t < F_OUT) b2s[t] = p.b2[t];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  } else {
    mpf::load_weights(mpw, wo, p.net, threadIdx.x - ENC_THREADS, p.groups * GROUP_THREADS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_chunks = *n_chunks_s;
  if (threadIdx.x == 0) stamp(p, 1);

  if (warp == 0) {
    if (lane == 0) produce(INT32_MAX, n_chunks);   // ------------------------ TMA producer: the rest of the stream
  } else if (warp == 1) {
    if (lane == 0) {   // ---------------------------------------------------- MMA issuer (encoder layer 1)
      int s = 0; uint32_t ph = 0;
      for (int c = 0; c < n_chunks; ++c) {
        const int buf = c & 1;
        mbar_wait(tmem_empty + buf, (uint32_t)((c >> 1) & 1) ^ 1);     // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(buf * H);
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(full + s, ph);
          if (c == 0 && kb == 0) stamp(p, 2);
          if (c == n_chunks - 1 && kb == KB - 1) stamp(p, 3);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + (size_t)s * STAGE_BYTES);
          const uint64_t adesc = make_smem_desc(a_addr), bdesc = make_smem_desc(a_addr + A_BYTES);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            umma_tf32(d_tmem, adesc + 2 * k, bdesc + 2 * k, IDESC, (uint32_t)((kb | k) != 0));
          umma_commit(empty + s);
          if (kb == KB - 1) umma_commit(tmem_full + buf);
          if (++s == NS) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp < 6) {   // ------------------------------------------------- encoder epilogue: one node row per thread
    const int q = warp & 3;
    const int row_in_tile = q * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int c = 0; c < n_chunks; ++c) {
      const int buf = c & 1;
      const int4 ch = chunk_s[c];
      mbar_wait(tmem_full + buf, (uint32_t)((c >> 1) & 1));
      tc_fence_after();
      // hidden layer: ReLU(acc + b1) (mlp.py:14,18), rounded to tf32, written back over the accumulator
#pragma unroll 1
      for (int cc = 0; cc < H / 32; ++cc) {
        float v[32];
        tmem_ld32(lane_base + (uint32_t)(buf * H + cc * 32), v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = tf32_rna(fmaxf(v[j] + b1s[cc * 32 + j], 0.f));
        tmem_st32(lane_base + (uint32_t)(buf * H + cc * 32), v);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      asm volatile("bar.sync %0, 128;" ::"n"(EPI_BAR) : "memory");
      if (warp == 2 && lane == 0) {
        tc_fence_after();
        const uint32_t w2_addr = smem_u32(w2s);
#pragma unroll
        for (int ks = 0; ks < H / UMMA_K; ++ks) {
          const uint64_t bdesc = make_smem_desc(w2_addr + (uint32_t)(ks >> 2) * (F_OUT * BLOCK_K * 4)) + 2 * (ks & 3);
          umma_tf32_ts(tmem_base + (uint32_t)(2 * H), tmem_base + (uint32_t)(buf * H + ks * UMMA_K), bdesc, IDESC2, (uint32_t)(ks != 0));
        }
        umma_commit(d2_full);
      }
      mbar_wait(d2_full, (uint32_t)(c & 1));
      tc_fence_after();
      float v[32];
      tmem_ld32(lane_base + (uint32_t)(2 * H), v);
      const int64_t row = (int64_t)ch.x + row_in_tile;
      if (row_in_tile < ch.y && row < p.batch.n_nodes) {
        float4* dst = reinterpret_cast<float4*>(p.h_out + row * F_OUT);
#pragma unroll
        for (int o4 = 0; o4 < F_OUT / 4; ++o4)
          dst[o4] = make_float4(fmaxf(v[4 * o4] + b2s[4 * o4], 0.f), fmaxf(v[4 * o4 + 1] + b2s[4 * o4 + 1], 0.f),
                                fmaxf(v[4 * o4 + 2] + b2s[4 * o4 + 2], 0.f), fmaxf(v[4 * o4 + 3] + b2s[4 * o4 + 3], 0.f));
      }
      mbar_arrive(h_ready + c);                       // release: this thread's h row is visible to the groups of this CTA
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty + buf);
      if (c == 0 && threadIdx.x == 64) stamp(p, 9);
      if (c == n_chunks - 1 && threadIdx.x == 64) stamp(p, 4);
    }
  } else {                 // ------------------------------------------------- message passing: groups of four warps
    mpf::Group gr;
    const int gi = (threadIdx.x - ENC_THREADS) / GROUP_THREADS;
    gr.tid = threadIdx.x - ENC_THREADS - gi * GROUP_THREADS; gr.size = GROUP_THREADS;
    gr.warp = gr.tid >> 5; gr.nwarps = GROUP_THREADS / 32; gr.lane = gr.tid & 31; gr.bar = 1 + gi;
    float* ss = mps + (size_t)gi * so.total;
    for (int c = 0; c < n_chunks; ++c) {
      const int4 ch = chunk_s[c];
      mbar_wait(h_ready + c, 0);                      // acquire: the chunk's encoded nodes are in L2
      if (gr.tid == 0 && gi == 0 && c == 0) stamp(p, 7);
      for (int j = gi; j < ch.w; j += p.groups)
        mpf::mp_frame<REATT_E, AGG>(gr, p.net, p.batch, mpw, wo, ss, so, (int64_t)ch.z + j);
      if (gr.tid == 0 && gi == 0 && c == 0) stamp(p, 8);
      if (gr.tid == 0 && gi == 0 && c == n_chunks - 1) stamp(p, 5);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
  }
  if (threadIdx.x == 0) stamp(p, 6);
}

}  // namespace mega
}  // namespace gnncca

using namespace gnncca;

#ifdef GNNCCA_MEGA_TRACE
unsigned long long* g_mega_trace = nullptr;
int g_mega_force_groups = 0, g_mega_force_stages = 0;
#endif

// Host side.  GNNCCA_ERR_UNSUPPORTED = this batch / configuration has no single-kernel path (the caller then runs the
// encoder and the fused message-passing kernel back to back).
int gnncca::mpn_mega_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* enc_W1, const float* enc_b1,
                         const float* enc_W2, const float* enc_b2, int32_t in_dim, int32_t hidden, const float* w_enc_edge,
                         const float* b_enc_edge, int32_t edge_in_dim, const float* x, const int64_t* edge_index,
                         const float* edge_attr, const int64_t* graph_ptr, const int64_t* edge_ptr, int64_t n_nodes,
                         int64_t n_edges, int64_t n_graphs, int32_t num_steps, int32_t first_class_step, int32_t cap_nodes,
                         int32_t cap_edges, float* h_enc, float* logits, int32_t* flags, void* stream) {
  if (hidden != 128 || cfg->f_node != mpf::FN || cfg->f_edge != mpf::FE || in_dim % enc::BLOCK_K != 0 || in_dim < enc::BLOCK_K)
    return GNNCCA_ERR_UNSUPPORTED;
  if (cap_nodes > enc::BLOCK_M || cap_edges > 512 || edge_in_dim < 1 || edge_in_dim > mpf::MAX_FA) return GNNCCA_ERR_UNSUPPORTED;
  if (n_nodes >= ((int64_t)1 << 30) || n_graphs < 1 || n_edges < 1) return GNNCCA_ERR_UNSUPPORTED;
  const bool has_cls = first_class_step <= num_steps;
  if (num_steps < 1 || !has_cls || first_class_step < 1 || cfg->cls_hidden != mpf::HC) return GNNCCA_ERR_UNSUPPORTED;
  if (cfg->agg < GNNCCA_AGG_SUM || cfg->agg > GNNCCA_AGG_MAX) return GNNCCA_ERR_BAD_ARG;
  if (!w->w_edge || !w->b_edge || !w->w_node || !w->b_node || !w->w_c1 || !w->b_c1 || !w->w_c2 || !w->b_c2 || !enc_W1 || !enc_b1 ||
      !enc_W2 || !enc_b2 || !w_enc_edge || !b_enc_edge || !x || !edge_index || !edge_attr || !graph_ptr || !edge_ptr || !h_enc ||
      !logits || !flags)
    return GNNCCA_ERR_BAD_ARG;
  if ((((uintptr_t)x | (uintptr_t)enc_W1 | (uintptr_t)h_enc | (uintptr_t)edge_attr) & 15) != 0) return GNNCCA_ERR_BAD_ARG;

  const int grid = (int)(n_graphs < kNumSMs ? n_graphs : kNumSMs);
  const int64_t frames_base = n_graphs / grid, frames_rem = n_graphs % grid;
  if (frames_base + 1 > mega::MAXF) return GNNCCA_ERR_UNSUPPORTED;
  // chunks a CTA can need: its frames hold at most (frames_base + 1) * cap_nodes nodes; every chunk but the last of a
  // greedy cut holds more than 128 - cap_nodes of them
  const int64_t worst_nodes = (frames_base + 1) * cap_nodes;
  const int64_t worst_chunks = cap_nodes >= enc::BLOCK_M ? frames_base + 1 : ceil_div(worst_nodes, enc::BLOCK_M - cap_nodes + 1) + 1;
  if (worst_chunks > mega::MAXC) return GNNCCA_ERR_UNSUPPORTED;

  mega::Params p{};
  mpf::Net& nt = p.net;
  nt.w = *w; nt.w_enc = w_enc_edge; nt.b_enc = b_enc_edge; nt.fa = edge_in_dim;
  nt.L = num_steps; nt.first_cls = first_class_step; nt.has_cls = 1; nt.agg = cfg->agg;
  nt.reatt_n = cfg->reattach_nodes ? 1 : 0; nt.reatt_e = cfg->reattach_edges ? 1 : 0;
  nt.cap_n = cap_nodes; nt.cap_e = cap_edges;
  mpf::Batch& bt = p.batch;
  bt.graph_ptr = graph_ptr; bt.n_graphs = n_graphs; bt.n_nodes = n_nodes; bt.n_edges = n_edges;
  bt.edge_index = edge_index; bt.edge_ptr = edge_ptr; bt.rowptr = nullptr; bt.src = nullptr; bt.dst = nullptr;
  bt.h_enc = h_enc; bt.edge_attr = edge_attr; bt.logits = logits; bt.h_dump = nullptr; bt.e_dump = nullptr; bt.flags = flags;
  p.b1 = enc_b1; p.W2 = enc_W2; p.b2 = enc_b2; p.h_out = h_enc; p.K = in_dim;
  p.frames_base = (int)frames_base; p.frames_rem = (int)frames_rem;

  const int K_MP = nt.reatt_n ? 64 : 32, EIN = nt.reatt_e ? 12 : 6;
  const int wfl = mpf::weight_layout(K_MP, EIN).total;
  const int sfl = mpf::state_layout(cap_nodes, cap_edges, nt.reatt_n != 0, nt.reatt_e != 0).total;
  // groups first (a chunk of EPFL frames holds 3-5 of them), then as many stages as still fit, at least 3
  const uint32_t budget = 227 * 1024 - 1024;
  int groups = 0, stages = 0;
  for (int g = mega::MAX_GROUPS; g >= 1 && !groups; --g)
    for (int s = enc::MAX_STAGES; s >= 3; --s)
      if (mega::layout(128, s, wfl, sfl, g).total <= budget) { groups = g; stages = s; break; }
  if (!groups) return GNNCCA_ERR_UNSUPPORTED;
#ifdef GNNCCA_MEGA_TRACE
  p.trace = g_mega_trace;
  if (g_mega_force_groups > 0 && g_mega_force_groups < groups) groups = g_mega_force_groups;
  if (g_mega_force_stages >= 2 && g_mega_force_stages < stages) stages = g_mega_force_stages;
#endif
  p.groups = groups; p.stages = stages;
  const size_t smem = (size_t)mega::layout(128, stages, wfl, sfl, groups).total + 1024;

  mega::Maps maps;
  for (int b = 0; b < 16; ++b)
    if (!enc::make_map(&maps.x[b], x, n_nodes, in_dim, 8 * (b + 1))) return GNNCCA_ERR_LAUNCH;
  if (!enc::make_map(&maps.w, enc_W1, 128, in_dim, 128)) return GNNCCA_ERR_LAUNCH;

  void (*kern)(const mega::Maps, const mega::Params) = nullptr;
  if (nt.reatt_e) kern = nt.agg == GNNCCA_AGG_MAX ? mega::mpn_mega_kernel<128, true, GNNCCA_AGG_MAX>
                       : nt.agg == GNNCCA_AGG_MEAN ? mega::mpn_mega_kernel<128, true, GNNCCA_AGG_MEAN> : mega::mpn_mega_kernel<128, true, GNNCCA_AGG_SUM>;
  else kern = nt.agg == GNNCCA_AGG_MAX ? mega::mpn_mega_kernel<128, false, GNNCCA_AGG_MAX>
            : nt.agg == GNNCCA_AGG_MEAN ? mega::mpn_mega_kernel<128, false, GNNCCA_AGG_MEAN> : mega::mpn_mega_kernel<128, false, GNNCCA_AGG_SUM>;
  int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  if (rc) return rc;
  kern<<<grid, mega::ENC_THREADS + groups * mega::GROUP_THREADS, smem, (cudaStream_t)stream>>>(maps, p);
  return check_launch();
}
