// Node encoder of MOTMPNet (reference models/mpn.py:128-142 MLPGraphIndependent.forward, node branch;
// models/mlp.py:12-23 for node_fc_dims=[H]): h = ReLU(W2 ReLU(W1 x + b1) + b2), x [N, K] fp32 -> h [N, 32].
//
// This is the one true dense contraction of the path (N x K x H, K = 2048, H = 128) and, at EPFL sizes, the
// largest HBM stream of the whole forward (8 KB per node).  One persistent, warp-specialised kernel:
//   warp 0      TMA producer : cp.async.bulk.tensor 128-row x 32-float boxes of x (evict-first) and H x 32 boxes
//                              of W1 (evict-last, L2 resident) into a ring of 128B-swizzled shared-memory stages
//   warp 1      MMA issuer   : tcgen05.mma kind::tf32, M=128, N=H, K=8 per instruction, accumulators in TMEM
//                              (double buffered: 2 x H columns) -- fp32 bits are consumed as TF32 in place, so
//                              x is never converted or re-written
//   warps 2..9  epilogue     : tcgen05.ld the accumulator row of "their" node (warps 2-5 sub-tile 0, 6-9 sub-tile 1),
//                              + b1, ReLU, then the 128 -> 32 layer in exact FP32 (packed FFMA2) against W2 held in
//                              shared memory, + b2, ReLU, one 128 B store per node.  The hidden activations never
//                              touch HBM.
// Roofline: HBM.  Algorithmic bytes per launch = N * (4 K + 128) + weights.
#include <cstdlib>
#include "tc_helpers.cuh"

namespace gnncca {
namespace enc {

constexpr int NUM_THREADS = 320;

struct Params {
  const float* b1; const float* W2; const float* b2;
  float* h_out;
  float* h1_out;             // optional [rows][H]: the hidden activations as layer 2 consumed them (training keeps them), else NULL
  int64_t rows;
  int K, H, stages;
  int rows_per_cta;          // multiple of 8: CTA c owns rows [c * rows_per_cta, (c+1) * rows_per_cta)
  uint32_t a_bytes;          // bytes of a stage given to the x rows: min(rows_per_cta, 256) rows, 1024-aligned
  unsigned long long* trace; // development aid (tools/micro/enc_trace.cu): [grid][8] globaltimer stamps, normally NULL
  // split-K form for SMALL inputs (gridDim.y = number of K splits): a CTA contracts k-blocks [blockIdx.y * kb_per_split, +kb_per_split)
  // only and stores its raw layer-1 accumulators to partial[blockIdx.y][row][H]; encoder_finish_kernel sums the splits and applies
  // the rest of the network.  Without it every CTA streams ALL of W1 (1 MB from L2) for its few rows: at N = 2560 that is 40-107 MB
  // of L2 traffic for a 21 MB input and 30 us per forward.
  float* partial; int kb_per_split;

};

#ifdef GNNCCA_ENC_TRACE
__device__ __forceinline__ void stamp(const Params& p, int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  if (p.trace) p.trace[(size_t)blockIdx.x * 12 + slot] = t;
}
#else
__device__ __forceinline__ void stamp(const Params&, int) {}
#endif

struct XMaps { CUtensorMap m[6]; int tail_rows; };   // x viewed with boxes of 8, 16, 32, 64, 128 and tail_rows rows x 32 floats

// Work decomposition.  A CTA owns ONE contiguous row range sized so that all SMs finish together (rows / grid, a
// multiple of 8) and walks it in "super tiles" of up to 256 rows = two M=128 MMA tiles that SHARE every W1 stage:
// per k-block the CTA ingests (rows x 128 B of x) + (H x 128 B of W1), so the L2->SM traffic spent on re-reading
// W1 is amortised over up to 256 rows instead of 128, and partial tiles load only the rows they own (8-row
// granular TMA boxes), never a neighbour's.
template <int H>
__global__ void __launch_bounds__(NUM_THREADS, 1)
node_encoder_tf32_kernel(const __grid_constant__ XMaps xmaps, const __grid_constant__ CUtensorMap tmap_w, const Params p) {
  constexpr uint32_t A_BYTES = BLOCK_M * BLOCK_K * 4;       // 16 KB per sub-tile
  constexpr uint32_t B_BYTES = H * BLOCK_K * 4;
  // TMEM columns: [0, 2H) the two layer-1 accumulators (overwritten in place by ReLU(acc + b1) as the A operand of
  // layer 2), [2H, 2H + 64) the two layer-2 accumulators
  constexpr uint32_t TMEM_NEED = 2 * H + 2 * F_OUT;
  constexpr uint32_t TMEM_COLS = TMEM_NEED <= 256 ? 256 : 512;
  static_assert(TMEM_NEED <= 512, "hidden width too large for one CTA's tensor memory");
  constexpr uint32_t IDESC = make_idesc(BLOCK_M, H);
  constexpr uint32_t IDESC2 = make_idesc(BLOCK_M, F_OUT);
  constexpr uint64_t HINT_EVICT_FIRST = 0x12F0000000000000ull, HINT_EVICT_LAST = 0x14F0000000000000ull;
  const uint32_t STAGE_BYTES = p.a_bytes + B_BYTES;       // sized to the rows this launch gives a CTA: more stages in flight

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);   // 128B swizzle atoms need 1024 B alignment
  const int NS = p.stages;
  uint8_t* tiles = smem;
  float* w2s = (float*)(tiles + (size_t)NS * STAGE_BYTES);     // W2 [32][H] as a K-major 128B-swizzled B operand (tf32, rna)
  float* b1s = w2s + H * F_OUT;
  float* b2s = b1s + H;
  uint64_t* full = (uint64_t*)(b2s + F_OUT);
  uint64_t* empty = full + MAX_STAGES;
  uint64_t* tmem_full = empty + MAX_STAGES;
  uint64_t* tmem_empty = tmem_full + 1;
  uint64_t* d2_full = tmem_empty + 1;               // [2]: layer-2 accumulator of sub-tile t is complete
  uint32_t* tmem_slot = (uint32_t*)(d2_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    stamp(p, 0);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // a dependent launch may take over SMs as CTAs of this grid exit
  }
  const int64_t row_begin = (int64_t)blockIdx.x * p.rows_per_cta;
  const int64_t row_end = min(row_begin + p.rows_per_cta, p.rows);
  const int my_rows = (int)(row_end - row_begin);
  const int n_super = (my_rows + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
  const int KB = p.partial ? p.kb_per_split : p.K / BLOCK_K;
  const int kb0 = p.partial ? (int)blockIdx.y * p.kb_per_split : 0;

  // producer state (warp 0, lane 0): the first ring of stages is issued BEFORE the set-up barrier, so x is already
  // streaming while the other warps allocate tensor memory and stage the small second layer
  int pr_s = 0, pr_st = 0, pr_kb = 0; uint32_t pr_ph = 0;
  auto produce = [&](int count) {
    for (; count > 0 && pr_st < n_super; --count) {
      const int64_t r0 = row_begin + (int64_t)pr_st * 2 * BLOCK_M;
      const int rem = my_rows - pr_st * 2 * BLOCK_M;
      const int rows8 = (min(rem, 2 * BLOCK_M) + 7) & ~7;          // rows of x this super tile loads (8-row boxes)
      mbar_wait(empty + pr_s, pr_ph ^ 1);
      uint8_t* a = tiles + (size_t)pr_s * STAGE_BYTES;
      mbar_expect_tx(full + pr_s, (uint32_t)rows8 * (BLOCK_K * 4) + B_BYTES);
      tma_load_2d(a + p.a_bytes, &tmap_w, full + pr_s, (kb0 + pr_kb) * BLOCK_K, 0, HINT_EVICT_LAST);
      // sub-tile 0 at byte 0, sub-tile 1 at row 128 = byte offset A_BYTES; one box each when the row count is 128
      // or the launch's tail size (every CTA but the last), 8-row granular greedy boxes otherwise
#pragma unroll 1
      for (int sub = 0; sub < 2; ++sub) {
        int need = min(rows8 - sub * BLOCK_M, BLOCK_M), off = sub * BLOCK_M;
        if (need <= 0) break;
        if (need == BLOCK_M) { tma_load_2d(a + (size_t)off * (BLOCK_K * 4), &xmaps.m[4], full + pr_s, (kb0 + pr_kb) * BLOCK_K, (int)(r0 + off), HINT_EVICT_FIRST); continue; }
        if (need == xmaps.tail_rows) { tma_load_2d(a + (size_t)off * (BLOCK_K * 4), &xmaps.m[5], full + pr_s, (kb0 + pr_kb) * BLOCK_K, (int)(r0 + off), HINT_EVICT_FIRST); continue; }
#pragma unroll
        for (int b = 3; b >= 0; --b) {
          const int sz = 8 << b;
          if (need >= sz) {
            tma_load_2d(a + (size_t)off * (BLOCK_K * 4), &xmaps.m[b], full + pr_s, (kb0 + pr_kb) * BLOCK_K, (int)(r0 + off), HINT_EVICT_FIRST);
            off += sz; need -= sz;
          }
        }
      }
      if (++pr_s == NS) { pr_s = 0; pr_ph ^= 1; }
      if (++pr_kb == KB) { pr_kb = 0; ++pr_st; }
    }
  };

  if (warp == 0 && lane < 7) {       // seven descriptors fetched in parallel
    const CUtensorMap* m = lane < 6 ? &xmaps.m[lane] : &tmap_w;
    asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
  }
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(tmem_full, 1); mbar_init(tmem_empty, 8);
    mbar_init(d2_full, 1); mbar_init(d2_full + 1, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    stamp(p, 9);
    produce(NS);
    stamp(p, 10);
  }
  if (warp == 1) {   // TMEM accumulators (this warp also frees them)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    if (lane == 0) stamp(p, 7);
  }
  if (warp >= 2) {   // epilogue warps stage the small second layer: coalesced reads of W2 [32][H], swizzled writes
    const int t = threadIdx.x - 64;
    constexpr int PER = H * F_OUT / 256;             // all loads in flight first (one DRAM round trip), then the stores
    float wv[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) wv[i] = __ldg(p.W2 + t + i * 256);
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const int idx = t + i * 256;
      const int n = idx / H, k = idx - n * H;
      const int kb = k >> 5, kk = k & 31;
      const int off = kb * (F_OUT * 32) + (n >> 3) * 256 + (n & 7) * 32 + (((kk >> 2) ^ (n & 7)) << 2) + (kk & 3);   // floats
      w2s[off] = tf32_rna(wv[i]);
    }
    for (int idx = t; idx < H; idx += 256) b1s[idx] = p.b1[idx];
    if (t < F_OUT) b2s[t] = p.b2[t];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes of w2s -> tensor-core reads
    if (threadIdx.x == 64) stamp(p, 8);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) stamp(p, 1);

  if (warp == 0) {
    if (lane == 0) produce(INT32_MAX);   // ---------------------------------- TMA producer: the rest of the stream
  } else if (warp == 1) {
    if (lane == 0) {   // ------------------------------------------------ MMA issuer (layer 1)
      int s = 0; uint32_t ph = 0;
      for (int st = 0; st < n_super; ++st) {
        const int rem = my_rows - st * 2 * BLOCK_M;
        const int nt = rem > BLOCK_M ? 2 : 1;
        mbar_wait(tmem_empty, (uint32_t)(st & 1) ^ 1);     // epilogue has drained the accumulators
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait(full + s, ph);
          if (st == 0 && kb == 0) stamp(p, 2);
          if (st == n_super - 1 && kb == KB - 1) stamp(p, 3);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + (size_t)s * STAGE_BYTES);
          const uint64_t bdesc = make_smem_desc(a_addr + p.a_bytes);
          for (int t = 0; t < nt; ++t) {
            const uint64_t adesc = make_smem_desc(a_addr + t * A_BYTES);
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)   // +32 bytes along K inside the swizzle atom = +2 in the (>>4) address field
              umma_tf32(tmem_base + (uint32_t)(t * H), adesc + 2 * k, bdesc + 2 * k, IDESC, (uint32_t)((kb | k) != 0));
          }
          umma_commit(empty + s);                  // stage reusable once these MMAs have read it
          if (kb == KB - 1) umma_commit(tmem_full);
          if (++s == NS) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {             // ---------------------------------------------------- epilogue: one node row per thread
    const int q = warp & 3;                         // TMEM lane quarter this warp may access
    const int t = (warp - 2) >> 2;                  // sub-tile: warps 2-5 -> 0, warps 6-9 -> 1
    const int row_in_tile = q * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int st = 0; st < n_super; ++st) {
      const int rem = my_rows - st * 2 * BLOCK_M;
      mbar_wait(tmem_full, (uint32_t)(st & 1));
      if (st == n_super - 1 && threadIdx.x == 64) stamp(p, 4);
      tc_fence_after();
      const int valid = min(rem - t * BLOCK_M, BLOCK_M);               // rows of this sub-tile that exist
      if (valid > 0 && p.partial != nullptr) {                         // split-K: the raw accumulators of this K range, nothing else
#pragma unroll 1
        for (int c = 0; c < H / 32; ++c) {
          float v[32];
          tmem_ld32(lane_base + (uint32_t)(t * H + c * 32), v);
          if (row_in_tile < valid) {
            const int64_t row = row_begin + (int64_t)st * 2 * BLOCK_M + t * BLOCK_M + row_in_tile;
            float4* d1 = reinterpret_cast<float4*>(p.partial + ((int64_t)blockIdx.y * p.rows + row) * H + c * 32);
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) d1[j4] = make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
          }
        }
      } else if (valid > 0) {                                          // uniform over the sub-tile's four warps
        // hidden layer: ReLU(acc + b1) (mlp.py:14,18), rounded to tf32, written back over the accumulator
#pragma unroll 1
        for (int c = 0; c < H / 32; ++c) {
          float v[32];
          tmem_ld32(lane_base + (uint32_t)(t * H + c * 32), v);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = tf32_rna(fmaxf(v[j] + b1s[c * 32 + j], 0.f));
          tmem_st32(lane_base + (uint32_t)(t * H + c * 32), v);
          if (p.h1_out != nullptr && row_in_tile < valid) {            // saved for the backward (uniform branch: NULL in inference)
            const int64_t row = row_begin + (int64_t)st * 2 * BLOCK_M + t * BLOCK_M + row_in_tile;
            float4* d1 = reinterpret_cast<float4*>(p.h1_out + row * H + c * 32);
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) d1[j4] = make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
          }
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        tc_fence_before();
        asm volatile("bar.sync %0, 128;" ::"r"(1 + t) : "memory");     // all 128 rows of the sub-tile are in tensor memory
        if (q == 2 && lane == 0) {                                     // first warp of the group issues layer 2
          tc_fence_after();
          const uint32_t w2_addr = smem_u32(w2s);
#pragma unroll
          for (int ks = 0; ks < H / UMMA_K; ++ks) {
            const uint64_t bdesc = make_smem_desc(w2_addr + (uint32_t)(ks >> 2) * (F_OUT * BLOCK_K * 4)) + 2 * (ks & 3);
            umma_tf32_ts(tmem_base + (uint32_t)(2 * H + t * F_OUT), tmem_base + (uint32_t)(t * H + ks * UMMA_K), bdesc, IDESC2, (uint32_t)(ks != 0));
          }
          umma_commit(d2_full + t);
        }
        mbar_wait(d2_full + t, (uint32_t)(st & 1));
        tc_fence_after();
        float v[32];
        tmem_ld32(lane_base + (uint32_t)(2 * H + t * F_OUT), v);
        if (row_in_tile < valid) {
          const int64_t row = row_begin + (int64_t)st * 2 * BLOCK_M + t * BLOCK_M + row_in_tile;
          float4* dst = reinterpret_cast<float4*>(p.h_out + row * F_OUT);
#pragma unroll
          for (int o4 = 0; o4 < F_OUT / 4; ++o4)
            dst[o4] = make_float4(fmaxf(v[4 * o4] + b2s[4 * o4], 0.f), fmaxf(v[4 * o4 + 1] + b2s[4 * o4 + 1], 0.f),
                                  fmaxf(v[4 * o4 + 2] + b2s[4 * o4 + 2], 0.f), fmaxf(v[4 * o4 + 3] + b2s[4 * o4 + 3], 0.f));
        }
      }
      tc_fence_before();                              // accumulators fully consumed: hand them back to the MMA warp
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty);
      if (st == n_super - 1 && threadIdx.x == 64) stamp(p, 5);
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


// Second launch of the split-K form: h1 = tf32(ReLU(sum_s partial[s] + b1)) (the value layer 2 of the one-launch form consumes, and
// what training keeps), h = ReLU(W2 h1 + b2) with W2 rounded to tf32 like there.  8 rows per 256-thread block; splits summed in order.
template <int H>
__global__ void __launch_bounds__(256) encoder_finish_kernel(const float* __restrict__ partial, int splits, int64_t rows,
                                                             const float* __restrict__ b1, const float* __restrict__ W2,
                                                             const float* __restrict__ b2, float* __restrict__ h1_out, float* __restrict__ h_out) {
  __shared__ float s_w2[F_OUT][H + 1];
  __shared__ float s_h1[8][H];
  const int tid = threadIdx.x, lane = tid & 31, r = tid >> 5;
  for (int i = tid; i < F_OUT * H; i += 256) s_w2[i / H][i % H] = tf32_rna(W2[i]);
  const int64_t row = (int64_t)blockIdx.x * 8 + r;
  if (row < rows) {
    // every partial of the row's H / 32 column groups in flight at once (up to 16 splits x 4 groups of independent loads: one
    // L2 round trip instead of one per split), then summed in split order
    float t[H / 32][16];
#pragma unroll
    for (int c = 0; c < H / 32; ++c)
#pragma unroll
      for (int sp = 0; sp < 16; ++sp) t[c][sp] = sp < splits ? __ldg(partial + ((int64_t)sp * rows + row) * H + c * 32 + lane) : 0.f;
#pragma unroll
    for (int c = 0; c < H / 32; ++c) {
      const int col = c * 32 + lane;
      float a = 0.f;
#pragma unroll
      for (int sp = 0; sp < 16; ++sp) a += t[c][sp];          // the splits beyond `splits` add +0.0
      const float v = tf32_rna(fmaxf(a + b1[col], 0.f));
      s_h1[r][col] = v;
      if (h1_out) h1_out[row * H + col] = v;
    }
  }
  __syncthreads();
  if (row < rows) {
    float a = b2[lane];
#pragma unroll 8
    for (int k = 0; k < H; ++k) a = fmaf(s_h1[r][k], s_w2[lane][k], a);
    h_out[row * F_OUT + lane] = fmaxf(a, 0.f);
  }
}

}  // namespace enc
}  // namespace gnncca

using namespace gnncca;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point: libgnncca has no link-time dependency on libcuda
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)sym;
    (void)cudaGetLastError();
  }
  return fn;
}




}  // namespace

// row-major fp32 matrix [rows][cols] -> boxes of box_rows x 32 floats, 128B swizzle, zero fill outside
bool gnncca::enc::make_map(CUtensorMap* map, const float* base, int64_t rows, int64_t cols, int box_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)enc::BLOCK_K, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

namespace {

#ifdef GNNCCA_ENC_TRACE
unsigned long long* g_enc_trace = nullptr;
int g_enc_force_stages = 0;
#endif

template <int H>
int launch_encoder(const float* x, const float* W1, const enc::Params& p0, cudaStream_t st) {
  enc::Params p = p0;
#ifdef GNNCCA_ENC_TRACE
  p.trace = g_enc_trace;
#endif
  // every SM gets the same number of rows (8-row granules); grids smaller than the machine only for tiny inputs
  int64_t rpc = (ceil_div(p.rows, kNumSMs) + 7) & ~(int64_t)7;
  // small inputs: every CTA streams all of W1 (K x H floats from L2) whatever its row count, so below ~64 rows per CTA the W1
  // re-reads are most of the traffic (N = 2560: 107 CTAs x 1 MB); fewer CTAs with 64 rows each move half the bytes
  static const int64_t min_rows = [] { const char* e = getenv("GNNCCA_ENC_MIN_ROWS"); return e ? (int64_t)atoi(e) : (int64_t)64; }();
  if (rpc < min_rows) rpc = p.rows < min_rows ? ((p.rows + 7) & ~(int64_t)7) : min_rows;
  if (rpc < 8) rpc = 8;
  p.rows_per_cta = (int)rpc;
  const int grid = (int)ceil_div(p.rows, rpc);
  const int64_t tile_rows = rpc < 2 * enc::BLOCK_M ? rpc : 2 * enc::BLOCK_M;
  p.a_bytes = (uint32_t)((tile_rows * enc::BLOCK_K * 4 + 1023) & ~(int64_t)1023);
  const size_t stage = (size_t)p.a_bytes + (size_t)H * enc::BLOCK_K * 4;
  const size_t fixed = 1024 /*alignment slack*/ + sizeof(float) * (H * enc::F_OUT + H + enc::F_OUT) + 8 * (2 * enc::MAX_STAGES + 4) + 32;
  int stages = (int)((227 * 1024 - fixed) / stage);
  if (stages > enc::MAX_STAGES) stages = enc::MAX_STAGES;
  if (stages < 2) return GNNCCA_ERR_UNSUPPORTED;
#ifdef GNNCCA_ENC_TRACE
  if (g_enc_force_stages > 0 && g_enc_force_stages < stages) stages = g_enc_force_stages;
#endif
  p.stages = stages;
  const size_t smem = fixed + stage * stages;
  enc::XMaps xm;
  CUtensorMap mw;
  for (int b = 0; b < 5; ++b)
    if (!enc::make_map(&xm.m[b], x, p.rows, p.K, 8 << b)) return GNNCCA_ERR_LAUNCH;
  {   // the partial sub-tile every full CTA ends with
    const int64_t last = rpc - 2 * enc::BLOCK_M * (ceil_div(rpc, 2 * enc::BLOCK_M) - 1);
    xm.tail_rows = (int)(last > enc::BLOCK_M ? last - enc::BLOCK_M : last);
    if (!enc::make_map(&xm.m[5], x, p.rows, p.K, xm.tail_rows)) return GNNCCA_ERR_LAUNCH;
  }
  if (!enc::make_map(&mw, W1, H, p.K, H)) return GNNCCA_ERR_LAUNCH;
  auto kern = enc::node_encoder_tf32_kernel<H>;
  static bool configured[64] = {};      // the attribute is per device: one flag per device ordinal
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
  if (dev < 0 || !configured[dev]) {
    int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
    if (rc) return rc;
    if (dev >= 0) configured[dev] = true;
  }
  kern<<<grid, enc::NUM_THREADS, smem, st>>>(xm, mw, p);
  return check_launch();
}

// Split-K form (see Params::partial): 128 rows per CTA x `splits` K ranges, then encoder_finish_kernel.
template <int H>
int launch_encoder_splitk(const float* x, const float* W1, const enc::Params& p0, int splits, float* partial, cudaStream_t st) {
  enc::Params p = p0;
  p.rows_per_cta = enc::BLOCK_M;
  p.partial = partial;
  p.kb_per_split = p.K / enc::BLOCK_K / splits;
  const int gx = (int)ceil_div(p.rows, (int64_t)enc::BLOCK_M);
  p.a_bytes = (uint32_t)enc::BLOCK_M * enc::BLOCK_K * 4;
  const size_t stage = (size_t)p.a_bytes + (size_t)H * enc::BLOCK_K * 4;
  const size_t fixed = 1024 + sizeof(float) * (H * enc::F_OUT + H + enc::F_OUT) + 8 * (2 * enc::MAX_STAGES + 4) + 32;
  int stages = (int)((227 * 1024 - fixed) / stage);
  if (stages > enc::MAX_STAGES) stages = enc::MAX_STAGES;
  if (stages > p.kb_per_split) stages = p.kb_per_split < 2 ? 2 : p.kb_per_split;
  if (stages < 2) return GNNCCA_ERR_UNSUPPORTED;
  p.stages = stages;
  const size_t smem = fixed + stage * stages;
  enc::XMaps xm;
  CUtensorMap mw;
  for (int b = 0; b < 5; ++b)
    if (!enc::make_map(&xm.m[b], x, p.rows, p.K, 8 << b)) return GNNCCA_ERR_LAUNCH;
  xm.tail_rows = enc::BLOCK_M;
  if (!enc::make_map(&xm.m[5], x, p.rows, p.K, xm.tail_rows)) return GNNCCA_ERR_LAUNCH;
  if (!enc::make_map(&mw, W1, H, p.K, H)) return GNNCCA_ERR_LAUNCH;
  auto kern = enc::node_encoder_tf32_kernel<H>;
  int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
  if (rc) return rc;
  kern<<<dim3((unsigned)gx, (unsigned)splits), enc::NUM_THREADS, smem, st>>>(xm, mw, p);
  // (launching this one as a programmatic dependent, with its own dependents released at once, measured SLOWER: cfg 1 59 us vs 52 us)
  enc::encoder_finish_kernel<H><<<(unsigned)ceil_div(p.rows, (int64_t)8), 256, 0, st>>>(partial, splits, p.rows, p.b1, p.W2, p.b2, p.h1_out, p.h_out);
  return check_launch(2);
}

}  // namespace

// Number of K splits gnncca_node_encoder_ws uses for `rows` rows (1: the one-launch form).  Small inputs only: as many splits as
// keep (row tiles x splits) within one wave of CTAs.
extern "C" int32_t gnncca_node_encoder_splits(int64_t rows, int32_t in_dim) {
  static const int enabled = [] { const char* e = getenv("GNNCCA_ENC_SPLITK"); return e ? atoi(e) : 1; }();
  if (!enabled || rows <= 0 || in_dim % enc::BLOCK_K != 0) return 1;
  const int64_t tiles = ceil_div(rows, (int64_t)enc::BLOCK_M);
  const int kb = in_dim / enc::BLOCK_K;
  int ks = 1;
  while (ks < 16 && (int64_t)ks * 2 * tiles <= kNumSMs && kb % (ks * 2) == 0 && kb / (ks * 2) >= 2) ks *= 2;
  return ks;
}

extern "C" int64_t gnncca_node_encoder_scratch_bytes(int64_t rows, int32_t in_dim, int32_t hidden) {
  const int ks = gnncca_node_encoder_splits(rows, in_dim);
  return ks > 1 ? (int64_t)ks * rows * hidden * (int64_t)sizeof(float) : 0;
}

// Fused two-layer node encoder on the tensor cores (see the header of this file).  Requirements: in_dim % 32 == 0,
// hidden in {64, 128}, out_dim == 32, x and W1 16-byte aligned.
int gnncca_node_encoder_tf32(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1, float* h,
                             int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* stream) {
  if (out_dim != enc::F_OUT || in_dim % enc::BLOCK_K != 0 || in_dim < enc::BLOCK_K) return GNNCCA_ERR_UNSUPPORTED;
  if ((((uintptr_t)x | (uintptr_t)W1 | (uintptr_t)h | (uintptr_t)h1) & 15) != 0) return GNNCCA_ERR_BAD_ARG;
  if (rows >= ((int64_t)1 << 31)) return GNNCCA_ERR_BAD_ARG;
  enc::Params p{};
  p.b1 = b1; p.W2 = W2; p.b2 = b2; p.h_out = h; p.h1_out = h1; p.rows = rows; p.K = in_dim; p.H = hidden;
  cudaStream_t st = (cudaStream_t)stream;
  switch (hidden) {
    case 64: return launch_encoder<64>(x, W1, p, st);
    case 128: return launch_encoder<128>(x, W1, p, st);
    default: return GNNCCA_ERR_UNSUPPORTED;
  }
}

// The same with caller scratch: small inputs take the split-K form when the scratch holds its partial sums
// (gnncca_node_encoder_scratch_bytes), everything else the one-launch form.
int gnncca_node_encoder_tf32_ws(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1, float* h,
                                int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* scratch, int64_t scratch_bytes,
                                void* stream) {
  const int ks = (hidden == 64 || hidden == 128) && out_dim == enc::F_OUT ? gnncca_node_encoder_splits(rows, in_dim) : 1;
  if (ks <= 1 || !scratch || scratch_bytes < gnncca_node_encoder_scratch_bytes(rows, in_dim, hidden) || (((uintptr_t)scratch) & 15) != 0)
    return gnncca_node_encoder_tf32(x, W1, b1, W2, b2, h1, h, rows, in_dim, hidden, out_dim, stream);
  if (in_dim < enc::BLOCK_K) return GNNCCA_ERR_UNSUPPORTED;
  if ((((uintptr_t)x | (uintptr_t)W1 | (uintptr_t)h | (uintptr_t)h1) & 15) != 0) return GNNCCA_ERR_BAD_ARG;
  enc::Params p{};
  p.b1 = b1; p.W2 = W2; p.b2 = b2; p.h_out = h; p.h1_out = h1; p.rows = rows; p.K = in_dim; p.H = hidden;
  cudaStream_t st = (cudaStream_t)stream;
  return hidden == 64 ? launch_encoder_splitk<64>(x, W1, p, ks, (float*)scratch, st) : launch_encoder_splitk<128>(x, W1, p, ks, (float*)scratch, st);
}
