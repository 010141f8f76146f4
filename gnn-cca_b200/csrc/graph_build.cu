// Cross-camera graph + edge-attribute construction (reference train.py:257-259,286-361,450; same code at
// train.py:582-692 and inference.py:189-279): batch-wide column normalisation, per-frame edge
// enumeration (cameras ascending, sources in row order, targets = every detection of another camera in
// row order), the four edge attributes, edge labels, and the Batch.from_data_list index offsets.
// Integer outputs are bit-exact; geometry is float64 like sklearn's paired_distances, cast to fp32 last.
#include <stdio.h>
#include <atomic>
#include "common.cuh"

namespace gnncca {

// ------------------------------------------------------------------------------------------------
// F.normalize(emb, p=2, dim=0)
// ------------------------------------------------------------------------------------------------
constexpr int CN_ROWS_PER_CHUNK = 128;
constexpr int GT = 32;       // tile edge of the edge-fill kernel (graph_prepare counts tiles of this size)

__global__ void __launch_bounds__(256) colnorm_partial_kernel(const float* __restrict__ x, int64_t rows, int cols, float* __restrict__ partial) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * CN_ROWS_PER_CHUNK;
  const int64_t r1 = min(r0 + CN_ROWS_PER_CHUNK, rows);
  if (c >= cols) return;
  float acc = 0.f;
  for (int64_t r = r0; r < r1; ++r) { const float v = x[r * cols + c]; acc = fmaf(v, v, acc); }
  partial[(int64_t)blockIdx.y * cols + c] = acc;
}
// cols % 4 == 0 and 16-byte aligned rows: a thread owns 4 adjacent columns, 4 rows in flight.  A matrix narrower than the block
// (cols / 4 < 256, e.g. the 256-wide ReID embeddings) spreads the chunk's rows over the RY = 256 / (cols / 4) thread rows the
// columns leave free and folds the RY partial sums in thread-row order: every thread has work, the order is fixed.
__device__ __forceinline__ void colnorm_partial4_body(const float* __restrict__ x, int64_t rows, int cols, float* __restrict__ partial,
                                                      int bx, int chunk, float4* s_red /* [256] */) {
  const int cq = cols >> 2, tid = threadIdx.x;
  int RY = 1, tx = tid, ty = 0;
  if (cq < 256) { RY = 256 / cq; tx = tid % cq; ty = tid / cq; }
  const int c4 = bx * 256 + tx;
  const int64_t r0 = (int64_t)chunk * CN_ROWS_PER_CHUNK;
  const int64_t r1 = min(r0 + CN_ROWS_PER_CHUNK, rows);
  const bool active = ty < RY && c4 < cq;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (active) {
    const float* px = x + (int64_t)c4 * 4;
    int64_t r = r0 + ty;
    for (; r + 3 * RY < r1; r += 4 * RY) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ld_stream_f4(px + (r + (int64_t)u * RY) * cols);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        acc.x = fmaf(v[u].x, v[u].x, acc.x); acc.y = fmaf(v[u].y, v[u].y, acc.y);
        acc.z = fmaf(v[u].z, v[u].z, acc.z); acc.w = fmaf(v[u].w, v[u].w, acc.w);
      }
    }
    for (; r < r1; r += RY) {
      const float4 v = ld_stream_f4(px + r * cols);
      acc.x = fmaf(v.x, v.x, acc.x); acc.y = fmaf(v.y, v.y, acc.y); acc.z = fmaf(v.z, v.z, acc.z); acc.w = fmaf(v.w, v.w, acc.w);
    }
  }
  if (RY > 1) {                                    // block-uniform
    s_red[tid] = acc;
    __syncthreads();
    if (active && ty == 0)
      for (int y = 1; y < RY; ++y) { const float4 o = s_red[y * cq + tx]; acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w; }
  }
  if (active && ty == 0) *reinterpret_cast<float4*>(partial + (int64_t)chunk * cols + c4 * 4) = acc;
}
__global__ void __launch_bounds__(256) colnorm_partial4_kernel(const float* __restrict__ x, int64_t rows, int cols, float* __restrict__ partial) {
  __shared__ float4 s_red[256];
  colnorm_partial4_body(x, rows, cols, partial, blockIdx.x, blockIdx.y, s_red);
}
// 32 columns per block of 32 warps; warp w sums the chunks w, w + 32, ... of its lane's column (up to 8 loads in flight), the 32
// partial sums are folded in warp order
__device__ __forceinline__ void colnorm_final_body(const float* __restrict__ partial, int chunks, int cols, float* __restrict__ norm, int bx,
                                                   float (*s_part)[33] /* [32][33] */, const bool reciprocal = false) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = bx * 32 + lane;
  float acc = 0.f;
  if (c < cols) {
    int k = warp;
    for (; k + 7 * 32 < chunks; k += 8 * 32) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldg(partial + (int64_t)(k + u * 32) * cols + c);
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u];
    }
    for (; k < chunks; k += 32) acc += __ldg(partial + (int64_t)k * cols + c);
  }
  s_part[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < cols) {
    float tot = s_part[0][lane];
#pragma unroll
    for (int w = 1; w < 32; ++w) tot += s_part[w][lane];
    const float nrm = fmaxf(sqrtf(tot), 1e-12f);                             // clamp_min(eps), eps = 1e-12
    norm[c] = reciprocal ? 1.f / nrm : nrm;
  }
}
__global__ void __launch_bounds__(1024) colnorm_final_kernel(const float* __restrict__ partial, int chunks, int cols, float* __restrict__ norm) {
  __shared__ float s_part[32][33];
  colnorm_final_body(partial, chunks, cols, norm, blockIdx.x, s_part);
}
__global__ void __launch_bounds__(256) colnorm_apply_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t rows, int cols,
                                                            const float* __restrict__ norm) {
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) out[t] = x[t] / norm[t % cols];
}
constexpr int CN_APPLY_ROWS = 32;
__global__ void __launch_bounds__(256) colnorm_apply4_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t rows, int cols,
                                                             const float* __restrict__ norm) {
  const int c4 = blockIdx.y * blockDim.x + threadIdx.x;
  if (c4 * 4 >= cols) return;
  const float4 nr = *reinterpret_cast<const float4*>(norm + c4 * 4);
  const int64_t r0 = (int64_t)blockIdx.x * CN_APPLY_ROWS, r1 = min(r0 + CN_APPLY_ROWS, rows);
  const float* px = x + (int64_t)c4 * 4;
  float* po = out + (int64_t)c4 * 4;
  int64_t r = r0;
  for (; r + 4 <= r1; r += 4) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ld_stream_f4(px + (r + u) * cols);
#pragma unroll
    for (int u = 0; u < 4; ++u) *reinterpret_cast<float4*>(po + (r + u) * cols) = make_float4(v[u].x / nr.x, v[u].y / nr.y, v[u].z / nr.z, v[u].w / nr.w);
  }
  for (; r < r1; ++r) {
    const float4 v = ld_stream_f4(px + r * cols);
    *reinterpret_cast<float4*>(po + r * cols) = make_float4(v.x / nr.x, v.y / nr.y, v.z / nr.z, v.w / nr.w);
  }
}

// ------------------------------------------------------------------------------------------------
// Per-frame preparation: one CTA per frame.
//   cam_cnt[i]   = #detections of the frame with the camera of i
//   cam_rank[i]  = #detections j < i with the same camera
//   src_off[i]   = offset (within the frame's edge block) of the edges whose source is i
//   order[k]     = local index of the k-th source in enumeration order (camera asc, row order)
//   cam_first[i] = position in order[] where i's camera block starts
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void graph_prepare_body(const int64_t* __restrict__ graph_ptr, const int32_t* __restrict__ id_cam,
                                                   int32_t* __restrict__ cam_cnt, int32_t* __restrict__ cam_first,
                                                   int64_t* __restrict__ src_off, int32_t* __restrict__ order,
                                                   int64_t* __restrict__ graph_edges, int64_t* __restrict__ graph_tiles,
                                                   int64_t* __restrict__ graph_unsorted, int tile, const int g) {
  const int64_t a = graph_ptr[g];
  const int n = (int)(graph_ptr[g + 1] - a);
  const int32_t* cam = id_cam + a;
  __shared__ unsigned long long s_total;
  if (threadIdx.x == 0) s_total = 0ull;
  __syncthreads();
  unsigned long long mine = 0;
  bool cam_desc = false;       // a row whose camera id is below its predecessor's: the enumeration (camera-major) is then NOT source-major
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int ci = cam[i];
    cam_desc |= i > 0 && ci < cam[i - 1];
    int cnt = 0, rank = 0, lower = 0;
    for (int j = 0; j < n; ++j) {
      const int cj = cam[j];
      cnt += (cj == ci);
      rank += (cj == ci) & (j < i);
      lower += (cj < ci);
    }
    cam_cnt[a + i] = cnt;
    cam_first[a + i] = lower;
    order[a + lower + rank] = i;
    mine += (unsigned long long)(n - cnt);
  }
  atomicAdd(&s_total, mine);   // integer: order-independent
  const int any_desc = __syncthreads_or(cam_desc);
  // src_off needs sum over lower cameras of their degrees: second O(n^2) sweep (degrees are per-camera constants)
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int ci = cam[i];
    int64_t off = 0;
    int cnt = 0, rank = 0;
    for (int j = 0; j < n; ++j) {
      const int cj = cam[j];
      cnt += (cj == ci);
      rank += (cj == ci) & (j < i);
    }
    for (int j = 0; j < n; ++j) {
      const int cj = cam[j];
      if (cj < ci) off += (int64_t)(n - cam_cnt[a + j]);   // written above by this CTA (visible after __syncthreads)
    }
    src_off[a + i] = off + (int64_t)rank * (n - cnt);
  }
  if (threadIdx.x == 0) {
    graph_edges[g] = (int64_t)s_total;
    const int64_t t = (n + tile - 1) / tile;
    graph_tiles[g] = t * t;
    graph_unsorted[g] = any_desc ? 1 : 0;
  }
}
__global__ void __launch_bounds__(256) graph_prepare_kernel(const int64_t* __restrict__ graph_ptr, const int32_t* __restrict__ id_cam,
                                                            int32_t* __restrict__ cam_cnt, int32_t* __restrict__ cam_first,
                                                            int64_t* __restrict__ src_off, int32_t* __restrict__ order,
                                                            int64_t* __restrict__ graph_edges, int64_t* __restrict__ graph_tiles,
                                                            int64_t* __restrict__ graph_unsorted, int tile) {
  graph_prepare_body(graph_ptr, id_cam, cam_cnt, cam_first, src_off, order, graph_edges, graph_tiles, graph_unsorted, tile, blockIdx.x);
}

// exclusive scan of a short int64 array (B+1 outputs), single CTA
// Exclusive scans of the per-frame edge and tile counts in ONE launch (block 0: edges -> edge_ptr, block 1: tiles ->
// tile_ptr) plus the four numbers the host needs to size the outputs, next to each other for a single read back:
// totals = {n_tiles, n_edges, max nodes per frame, max edges per frame, #frames whose rows are not camera-ascending}.
// totals_host (optional): the same words in pinned, device-mapped host memory, written by the kernel itself so that the host
// needs no copy and no stream synchronisation to size the outputs: each of the two blocks writes its values, fences, and then
// its check word ([5] edges block, [6] tiles block) = GNNCCA_TOTALS_MAGIC; the host clears the check words before the launch
// and polls them.
__device__ __forceinline__ void scan_counts_body(const int64_t* __restrict__ graph_edges, const int64_t* __restrict__ graph_tiles,
                                                 const int64_t* __restrict__ graph_unsorted,
                                                 const int64_t* __restrict__ graph_ptr, int64_t n, int64_t* __restrict__ edge_ptr,
                                                 int64_t* __restrict__ tile_ptr, int64_t* __restrict__ totals,
                                                 volatile int64_t* totals_host, const int which, int64_t* s_warp, int64_t* s_max) {
  const bool edges = which == 0;
  const int64_t* in = edges ? graph_edges : graph_tiles;
  int64_t* out = edges ? edge_ptr : tile_ptr;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t per = (n + blockDim.x - 1) / blockDim.x;
  const int64_t b = min((int64_t)t * per, n), e = min(b + per, n);
  int64_t acc = 0, mx = 0;
  int uns = 0;
  for (int64_t i = b; i < e; ++i) {
    const int64_t v = in[i];
    acc += v;
    uns += (int)graph_unsorted[i];
    mx = max(mx, edges ? v : graph_ptr[i + 1] - graph_ptr[i]);     // block 0: max edges, block 1: max nodes per frame
  }
  const int n_uns = __syncthreads_count(uns != 0);   // threads (not frames) that saw one: only "== 0" matters
  int64_t incl = acc;                              // inclusive scan of the per-thread sums: warp shuffles, then the 32 warp totals
  for (int o = 1; o < 32; o <<= 1) { const int64_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 31) s_warp[warp] = incl;
  if (lane == 0) s_max[warp] = mx;
  __syncthreads();
  if (warp == 0) {
    int64_t w = s_warp[lane];
    for (int o = 1; o < 32; o <<= 1) { const int64_t v = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += v; }
    s_warp[lane] = w;                              // inclusive over warps
    int64_t m = s_max[lane];
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_max[0] = m;
  }
  __syncthreads();
  int64_t run = (warp > 0 ? s_warp[warp - 1] : 0) + incl - acc;
  for (int64_t i = b; i < e; ++i) { const int64_t v = in[i]; out[i] = run; run += v; }
  if (t == 0) {
    const int64_t total = s_warp[31];
    out[n] = total;
    totals[edges ? 1 : 0] = total;       // [0] tiles, [1] edges
    totals[edges ? 3 : 2] = s_max[0];    // [2] max nodes, [3] max edges
    if (edges) totals[4] = n_uns;        // [4] != 0: some frame's rows are not camera-ascending
    if (totals_host) {
      totals_host[edges ? 1 : 0] = total;
      totals_host[edges ? 3 : 2] = s_max[0];
      if (edges) totals_host[4] = n_uns;
      __threadfence_system();
      totals_host[edges ? 5 : 6] = GNNCCA_TOTALS_MAGIC;
    }
  }
}
__global__ void __launch_bounds__(1024) scan_counts_kernel(const int64_t* __restrict__ graph_edges, const int64_t* __restrict__ graph_tiles,
                                                           const int64_t* __restrict__ graph_unsorted,
                                                           const int64_t* __restrict__ graph_ptr, int64_t n, int64_t* __restrict__ edge_ptr,
                                                           int64_t* __restrict__ tile_ptr, int64_t* __restrict__ totals) {
  __shared__ int64_t s_warp[32], s_max[32];
  scan_counts_body(graph_edges, graph_tiles, graph_unsorted, graph_ptr, n, edge_ptr, tile_ptr, totals, nullptr, blockIdx.x, s_warp, s_max);
}

// The two launches of gnncca_graph_prepare_all.  (1) blocks [0, reid_blocks): column-norm partial sums of the ReID embeddings,
// the rest: one block per frame (graph_prepare_body).  (2) blocks 0 / 1: the scans, the rest: the ReID norms (32 columns each).
struct PrepAll {
  const int64_t* graph_ptr; const int32_t* id_cam; int64_t n_graphs;
  int32_t* cam_cnt; int32_t* cam_first; int64_t* src_off; int32_t* order;
  int64_t* graph_edges; int64_t* graph_tiles; int64_t* graph_unsorted; int64_t* edge_ptr; int64_t* tile_ptr; int64_t* totals; int64_t* totals_host;
  const float* reid; int64_t rows; int reid_cols; float* reid_partial; float* reid_rnorm; int reid_bx; int reid_chunks;
};
__global__ void __launch_bounds__(256) prep_pass_kernel(const PrepAll p) {
  __shared__ float4 s_red[256];
  const int n_reid = p.reid_bx * p.reid_chunks;
  if ((int)blockIdx.x < n_reid) {
    colnorm_partial4_body(p.reid, p.rows, p.reid_cols, p.reid_partial, blockIdx.x % p.reid_bx, blockIdx.x / p.reid_bx, s_red);
  } else {
    graph_prepare_body(p.graph_ptr, p.id_cam, p.cam_cnt, p.cam_first, p.src_off, p.order, p.graph_edges, p.graph_tiles, p.graph_unsorted, GT,
                       (int)blockIdx.x - n_reid);
  }
}
__global__ void __launch_bounds__(1024) prep_final_kernel(const PrepAll p) {
  __shared__ int64_t s_warp[32], s_max[32];
  __shared__ float s_part[32][33];
  if (blockIdx.x < 2) {
    scan_counts_body(p.graph_edges, p.graph_tiles, p.graph_unsorted, p.graph_ptr, p.n_graphs, p.edge_ptr, p.tile_ptr, p.totals,
                     (volatile int64_t*)p.totals_host, blockIdx.x, s_warp, s_max);
  } else {
    colnorm_final_body(p.reid_partial, p.reid_chunks, p.reid_cols, p.reid_rnorm, (int)blockIdx.x - 2, s_part, true);
  }
}

// ------------------------------------------------------------------------------------------------
// Edge fill: CTA = 32 sources x 32 targets of one frame; reid rows staged in shared memory in chunks.
// ------------------------------------------------------------------------------------------------
constexpr int GCH = 64;      // feature chunk

struct BuildParams {
  const int64_t* graph_ptr; const int64_t* edge_ptr; const int64_t* tile_ptr; int64_t n_graphs;
  const int32_t* id_cam; const int64_t* ids; const double* xw; const double* yw; const double* max_dist;
  const float* reid; const float* reid_rnorm; int reid_dim; int attr_mode;
  const int32_t* cam_cnt; const int32_t* cam_first; const int64_t* src_off; const int32_t* order;
  int64_t* edge_index; int64_t n_edges; float* edge_attr; float* edge_labels; int one_tile_per_frame;
};

#ifndef GNNCCA_BG_PREFETCH
#define GNNCCA_BG_PREFETCH 1
#endif
#ifndef GNNCCA_BG_ADAPT
#define GNNCCA_BG_ADAPT 1
#endif
// NORM: `reid` is raw and is scaled by the reciprocal column norms while staged (one FMUL per element; an IEEE division here
// measured +30 % on the whole kernel)
template <bool NORM>
__global__ void __launch_bounds__(256, 3) build_graphs_kernel(const BuildParams p) {
  const int tid = threadIdx.x, nthr = blockDim.x;   // 32 x (source quads the largest frame needs): 128..256, 256 if any frame > 32 nodes
  const int lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
  // which frame does this tile belong to: tile_ptr[g] <= tb < tile_ptr[g + 1].  One tile per frame (every frame <= 32 nodes,
  // none empty): the tile IS the frame.  Otherwise a 32-ary search, one probe per lane: 3 dependent loads for 4096 frames
  // where a binary search makes 12.
  const int64_t tb = blockIdx.x;
  int64_t g;
  if (p.one_tile_per_frame) {
    g = tb;
  } else {
    int64_t lo = 0, hi = p.n_graphs;  // invariant: tile_ptr[lo] <= tb < tile_ptr[hi]
    while (hi - lo > 1) {
      const int64_t step = (hi - lo + 31) / 32;
      const int64_t probe = lo + (int64_t)(lane + 1) * step;
      const bool le = probe < hi && p.tile_ptr[probe] <= tb;          // monotone in the lane: tile_ptr is non-decreasing
      const int cnt = __popc(__ballot_sync(0xffffffffu, le));
      const int64_t nlo = lo + (int64_t)cnt * step;
      hi = min(hi, nlo + step);
      lo = nlo;
    }
    g = lo;
  }
  const int64_t a = p.graph_ptr[g];
  const int n = (int)(p.graph_ptr[g + 1] - a);
  const int nt = (n + GT - 1) / GT;
  const int local = (int)(tb - (p.one_tile_per_frame ? tb : p.tile_ptr[g]));
  const int i0 = (local / nt) * GT, j0 = (local % nt) * GT;
  const bool need_app = p.attr_mode != GNNCCA_ATTR_ONLY_DIST;
  const bool diag = i0 == j0;     // the only tile of a frame of <= 32 detections: sources and targets are the same rows, staged once

  // ReID rows of the tile, feature-major ([k][row], row stride 36 floats = 16-byte aligned quads): the 4 sources of a thread
  // are one LDS.128 (the same address for the whole warp), the target is lane-consecutive; arithmetic on source PAIRS (FFMA2).
  // reid_rnorm != NULL: `reid` holds the RAW embeddings and the column normalisation (train.py:257-259) happens here, while
  // staging, as a multiplication by 1 / max(||column||, 1e-12): the normalised matrix is never written to memory.
  __shared__ __align__(16) float s_a[GCH][GT + 4];
  __shared__ __align__(16) float s_b[GCH][GT + 4];
  __shared__ float s_rn[2 * GT];      // squared row norms of the 32 source rows | 32 target rows (for the cosine)
  // per-row tables of the tile, [0] sources / [1] targets: loaded once, up front, beside the first ReID chunk (the epilogue
  // used to walk them as chains of dependent global loads)
  __shared__ double s_x[2][GT], s_y[2][GT];
  __shared__ int64_t s_id[2][GT], s_soff[GT];
  __shared__ int32_t s_cam[2][GT], s_first[GT], s_cnt[GT], s_order[GT];
  if (tid < 2 * GT) {
    const int half = tid >> 5, r = tid & 31, row = (half ? j0 : i0) + r;
    const bool in = row < n;
    s_cam[half][r] = in ? p.id_cam[a + row] : -1;
    s_id[half][r] = in ? p.ids[a + row] : 0;
    s_x[half][r] = in ? p.xw[a + row] : 0.0;
    s_y[half][r] = in ? p.yw[a + row] : 0.0;
    if (half == 0) {
      s_first[r] = in ? p.cam_first[a + row] : 0;
      s_cnt[r] = in ? p.cam_cnt[a + row] : 0;
      s_soff[r] = in ? p.src_off[a + row] : 0;
      s_order[r] = (n <= GT && r < n) ? p.order[a + r] : 0;    // a frame of one tile: its whole order[] table
    }
  }
  // distances are scaled by 1 / max_dist in float64 (one division per thread instead of two per edge); the float64 product is
  // within 2 ulp(double) of the quotient: the float32 it is cast to (train.py:344) differs from the reference's only when the
  // quotient lies that close to a float32 rounding boundary (about one value in 2^28)
  const double inv_md = 1.0 / p.max_dist[g];
  const int64_t ebase = p.edge_ptr[g];

  const int tj = lane;            // target within tile
  const int ti0 = warp * 4;       // this thread owns sources ti0 .. ti0 + 3 of the tile
  float2 d2a = make_float2(0.f, 0.f), d2b = d2a, dota = d2a, dotb = d2a;
  // row norms: thread t sums the features k = (t & 3) + 4 m of row t >> 2 (0..31 sources, 32..63 targets), chunk after chunk;
  // the 4 slices of a row are folded at the end in a fixed order (xor-shuffles over adjacent lanes)
  float rn = 0.f;
  const int rrow = tid >> 2, rsl = tid & 3;
  const bool has_rn = tid < (diag ? 4 * GT : 8 * GT);      // warp-uniform
  const float (*sb)[GT + 4] = diag ? s_a : s_b;
  if (need_app) {
    // Staging: a warp owns row quads (4 rows x 8 features per load instruction: whole 32-byte sectors in, 32 different banks
    // out, (4 k + r) mod 32).  Quads 0..7 are the source rows, 8..15 the target rows of an off-diagonal tile; with at least
    // 4 warps (8 off the diagonal) that is at most two quads per warp, 8 loads each per 64-feature chunk, and the chunk after
    // the current one travels in registers while the current one is in the arithmetic.  Quads beyond the frame's rows are
    // zero-filled once and never touched again.
    const int lk = lane & 7, lr = lane >> 3;
    const int nquads = diag ? 8 : 16;
    const float* gp[2];
    float* sp[2];
    bool q_valid[2], q_data[2], l_ok[2];
#pragma unroll
    for (int sl = 0; sl < 2; ++sl) {
      const int q = warp + sl * nwarp, half = q >> 3, rq = q & 7;
      const int base = half ? j0 : i0, row = base + rq * 4 + lr;
      q_valid[sl] = q < nquads;
      q_data[sl] = q_valid[sl] && base + rq * 4 < n;           // warp-uniform
      l_ok[sl] = q_data[sl] && row < n;
      gp[sl] = p.reid + (a + (l_ok[sl] ? row : 0)) * p.reid_dim + lk;
      sp[sl] = (half ? &s_b[0][0] : &s_a[0][0]) + lk * (GT + 4) + rq * 4 + lr;
    }
    float pre[2][8];
    auto fetch = [&](const int k0) {
#pragma unroll
      for (int sl = 0; sl < 2; ++sl) {
        if (!q_data[sl]) continue;
#pragma unroll
        for (int kb = 0; kb < 8; ++kb)
          pre[sl][kb] = (l_ok[sl] && k0 + 8 * kb + lk < p.reid_dim) ? __ldg(gp[sl] + k0 + 8 * kb) : 0.f;
      }
    };
    auto stage = [&](const int k0) {
      float rs[8];
      if (NORM) {
#pragma unroll
        for (int kb = 0; kb < 8; ++kb) rs[kb] = (k0 + 8 * kb + lk < p.reid_dim) ? __ldg(p.reid_rnorm + k0 + 8 * kb + lk) : 0.f;
      }
#pragma unroll
      for (int sl = 0; sl < 2; ++sl) {
        if (q_data[sl]) {
#pragma unroll
          for (int kb = 0; kb < 8; ++kb) sp[sl][8 * kb * (GT + 4)] = NORM ? pre[sl][kb] * rs[kb] : pre[sl][kb];
        } else if (q_valid[sl] && k0 == 0) {
#pragma unroll
          for (int kb = 0; kb < 8; ++kb) sp[sl][8 * kb * (GT + 4)] = 0.f;
        }
      }
    };
    if (GNNCCA_BG_PREFETCH) fetch(0);
    for (int k0 = 0; k0 < p.reid_dim; k0 += GCH) {
      if (!GNNCCA_BG_PREFETCH) fetch(k0);
      stage(k0);
      __syncthreads();
      if (GNNCCA_BG_PREFETCH && k0 + GCH < p.reid_dim) fetch(k0 + GCH);
      if (has_rn) {
        const float* col = rrow < GT ? &s_a[0][rrow] : &s_b[0][rrow - GT];
#pragma unroll
        for (int m = 0; m < GCH / 4; ++m) { const float v = col[(rsl + 4 * m) * (GT + 4)]; rn = fmaf(v, v, rn); }
      }
      const int kmax = (i0 + ti0 < n) ? min(GCH, p.reid_dim - k0) : 0;     // a warp whose four sources lie beyond the frame only stages
#pragma unroll 4
      for (int k = 0; k < kmax; ++k) {
        const float b = sb[k][tj];
        const float4 av = *reinterpret_cast<const float4*>(&s_a[k][ti0]);
        const float2 a01 = make_float2(av.x, av.y), a23 = make_float2(av.z, av.w);
        // F.pairwise_distance: ||a - b + eps||_2, eps = 1e-6 ((a - b) first, then + eps, as torch does)
        const float2 df01 = pk::fadd2(pk::fadd2(a01, pk::bc(-b)), pk::bc(1e-6f));
        const float2 df23 = pk::fadd2(pk::fadd2(a23, pk::bc(-b)), pk::bc(1e-6f));
        d2a = pk::ffma2(df01, df01, d2a); d2b = pk::ffma2(df23, df23, d2b);
        dota = pk::ffma2(a01, pk::bc(b), dota); dotb = pk::ffma2(a23, pk::bc(b), dotb);
      }
      __syncthreads();
    }
    if (has_rn) {
      rn += __shfl_xor_sync(0xffffffffu, rn, 1);
      rn += __shfl_xor_sync(0xffffffffu, rn, 2);
      if (rsl == 0) { s_rn[rrow] = rn; if (diag) s_rn[GT + rrow] = rn; }
    }
  }
  __syncthreads();                // the row tables (and the norms)
  const float d2[4] = {d2a.x, d2a.y, d2b.x, d2b.y}, dot[4] = {dota.x, dota.y, dotb.x, dotb.y};
  const int j = j0 + tj;
  if (j >= n || ti0 >= GT) return;
  const float nb = s_rn[GT + tj];
  const int cj = s_cam[1][tj];
  const double xj = s_x[1][tj], yj = s_y[1][tj];
  const int64_t idj = s_id[1][tj];
  const int width = (p.attr_mode == GNNCCA_ATTR_FULL) ? 4 : 2;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int li = ti0 + q, i = i0 + li;
    if (i >= n) continue;
    const int ci = s_cam[0][li];
    if (ci == cj) continue;
    // targets of source i are the detections of other cameras in row order: j's slot is
    // j - #{j' < j : cam(j') == cam(i)}; that count is a lower_bound inside i's camera block of order[]
    int l = 0, h = s_cnt[li];
    if (n <= GT) {
      const int32_t* blk = s_order + s_first[li];
      while (l < h) { const int m = (l + h) >> 1; if (blk[m] < j) l = m + 1; else h = m; }
    } else {
      const int32_t* blk = p.order + a + s_first[li];
      while (l < h) { const int m = (l + h) >> 1; if (blk[m] < j) l = m + 1; else h = m; }
    }
    const int64_t e = ebase + s_soff[li] + (j - l);
    p.edge_index[e] = a + i;                 // rebased to the frame (train.py:356) + Batch offset (train.py:450)
    p.edge_index[p.n_edges + e] = a + j;
    p.edge_labels[e] = (s_id[0][li] == idj) ? 1.f : 0.f;   // train.py:348-351
    float* out = p.edge_attr + e * width;
    float o[4];
    int col = 0;
    if (p.attr_mode != GNNCCA_ATTR_ONLY_APPEARANCE) {
      const double dx = s_x[0][li] - xj, dy = s_y[0][li] - yj;
      const double l2 = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));   // paired_distances euclidean (float64)
      const double l1 = __dadd_rn(fabs(dx), fabs(dy));                            // manhattan
      o[col++] = (float)(l2 * inv_md);      // train.py:323, cast at :344
      o[col++] = (float)(l1 * inv_md);      // train.py:326
    }
    if (need_app) {
      o[col++] = sqrtf(d2[q]);                                                  // train.py:306
      o[col++] = dot[q] / (fmaxf(sqrtf(s_rn[li]), 1e-8f) * fmaxf(sqrtf(nb), 1e-8f));   // train.py:308, eps = 1e-8
    }
    if (width == 4) *reinterpret_cast<float4*>(out) = make_float4(o[0], o[1], o[2], o[3]);
    else *reinterpret_cast<float2*>(out) = make_float2(o[0], o[1]);
  }
}

}  // namespace gnncca

using namespace gnncca;

extern "C" int64_t gnncca_colnorm_scratch_bytes(int64_t rows, int32_t cols) {
  return (int64_t)sizeof(float) * (ceil_div(rows > 0 ? rows : 1, CN_ROWS_PER_CHUNK) + 1) * cols;
}

static bool colnorm_vec_ok(const float* emb, const float* out, const float* scratch, int64_t rows, int32_t cols) {
  return cols % 4 == 0 && ((((uintptr_t)emb | (uintptr_t)out | (uintptr_t)scratch) & 15) == 0);
}

extern "C" int gnncca_colnorm_norms(const float* emb, int64_t rows, int32_t cols, float* scratch, float* norm, void* stream) {
  if (rows < 0 || cols <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!emb || !scratch || !norm) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int chunks = (int)ceil_div(rows, CN_ROWS_PER_CHUNK);
  if (chunks > 65535) return GNNCCA_ERR_UNSUPPORTED;
  if (colnorm_vec_ok(emb, emb, scratch, rows, cols)) {
    dim3 g1((unsigned)ceil_div(cols / 4, 256), (unsigned)chunks);
    colnorm_partial4_kernel<<<g1, 256, 0, st>>>(emb, rows, cols, scratch);
  } else {
    dim3 g1((unsigned)ceil_div(cols, 256), (unsigned)chunks);
    colnorm_partial_kernel<<<g1, 256, 0, st>>>(emb, rows, cols, scratch);
  }
  colnorm_final_kernel<<<(unsigned)ceil_div(cols, 32), 1024, 0, st>>>(scratch, chunks, cols, norm);
  return check_launch(2);
}

extern "C" int gnncca_colnorm_apply(const float* emb, float* out, int64_t rows, int32_t cols, const float* norm, void* stream) {
  if (rows < 0 || cols <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!emb || !out || !norm) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  // (rows of 32) x (columns of 1024) blocks; the row blocks ride on grid.x (2^31 - 1 of them), so any batch size takes the vector path
  if (colnorm_vec_ok(emb, out, norm, rows, cols) && ceil_div(cols / 4, 256) <= 65535) {
    dim3 g3((unsigned)ceil_div(rows, CN_APPLY_ROWS), (unsigned)ceil_div(cols / 4, 256));
    colnorm_apply4_kernel<<<g3, 256, 0, st>>>(emb, out, rows, cols, norm);
  } else {
    colnorm_apply_kernel<<<grid_for(rows * cols, 256), 256, 0, st>>>(emb, out, rows, cols, norm);
  }
  return check_launch();
}

extern "C" int gnncca_colnorm(const float* emb, float* out, int64_t rows, int32_t cols, float* scratch, void* stream) {
  if (rows < 0 || cols <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!emb || !out || !scratch) return GNNCCA_ERR_BAD_ARG;
  const int chunks = (int)ceil_div(rows, CN_ROWS_PER_CHUNK);
  float* norm = scratch + (int64_t)chunks * cols;
  int rc = gnncca_colnorm_norms(emb, rows, cols, scratch, norm, stream);
  if (rc) return rc;
  return gnncca_colnorm_apply(emb, out, rows, cols, norm, stream);
}

// ------------------------------------------------------------------------------------------------
// gnncca_colnorm without caller scratch and without any allocation: the partial sums and the norms live in a file-scope
// device array.  Every GPU that loads the library holds GNNCCA_COLNORM_STATIC_MAX_FLOATS * 4 bytes of it until exit.
// In host code `g_colnorm_static` is only a handle of the symbol: the device address comes from cudaGetSymbolAddress, once
// per device with that device current (the array exists once per device, at a different address on each).
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_colnorm_static[GNNCCA_COLNORM_STATIC_MAX_FLOATS];

static int colnorm_static_ptr(float** out) {
  constexpr int MAX_DEV = 64;
  static std::atomic<float*> s_ptr[MAX_DEV];          // zero-initialised; a race between two first calls stores the same value
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) { set_error_text("gnncca_colnorm_static: no current device"); return GNNCCA_ERR_LAUNCH; }
  float* p = s_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    void* sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_colnorm_static);
    if (e != cudaSuccess || !sym) {
      char msg[200];
      snprintf(msg, sizeof(msg), "gnncca_colnorm_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      set_error_text(msg);
      (void)cudaGetLastError();
      return GNNCCA_ERR_LAUNCH;
    }
    p = static_cast<float*>(sym);
    s_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return GNNCCA_OK;
}

extern "C" int gnncca_colnorm_static(const float* emb, float* out, int64_t rows, int32_t cols, void* stream) {
  if (rows < 0 || cols <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!emb || !out) return GNNCCA_ERR_BAD_ARG;
  const int64_t need = gnncca_colnorm_scratch_bytes(rows, cols) / (int64_t)sizeof(float);
  if (need > (int64_t)GNNCCA_COLNORM_STATIC_MAX_FLOATS) {
    char msg[240];
    snprintf(msg, sizeof(msg), "gnncca_colnorm_static: [%lld, %d] needs %lld floats of scratch, the built-in array holds %lld "
             "(use gnncca_colnorm with caller scratch)", (long long)rows, (int)cols, (long long)need,
             (long long)GNNCCA_COLNORM_STATIC_MAX_FLOATS);
    set_error_text(msg);
    return GNNCCA_ERR_WORKSPACE;
  }
  float* scratch = nullptr;
  const int rc = colnorm_static_ptr(&scratch);
  if (rc) return rc;
  // nothing in the array has to start at zero: the first launch writes every partial sum [chunk][col] the second one reads,
  // and the second writes every norm the third reads (what the previous call left behind is never read)
  return gnncca_colnorm(emb, out, rows, cols, scratch, stream);
}

// out[r][c] = W[r][c] / norm[c]: the column scaling of F.normalize(x, dim=0) moved onto the weights of the Linear that reads x
// (x_normalised W^T = x (W diag(1 / norm))^T), so that a 200 MB activation matrix is never rewritten for the sake of 2048 scalars
__global__ void __launch_bounds__(256) scale_columns_kernel(const float* __restrict__ W, float* __restrict__ out, int64_t total, int cols,
                                                            const float* __restrict__ norm) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) out[t] = W[t] / norm[t % cols];
}

extern "C" int gnncca_scale_columns(const float* W, float* out, int64_t rows, int32_t cols, const float* norm, void* stream) {
  if (rows < 0 || cols <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!W || !out || !norm) return GNNCCA_ERR_BAD_ARG;
  scale_columns_kernel<<<grid_for(rows * cols, 256, 4), 256, 0, (cudaStream_t)stream>>>(W, out, rows * cols, cols, norm);
  return check_launch();
}

extern "C" int64_t gnncca_graph_prepare_scratch_bytes(int64_t n_graphs, int64_t n_nodes) {
  // cam_cnt, cam_first, order (int32 x N each) + src_off (int64 x N) + graph_edges, graph_tiles (int64 x B) + tile_ptr (B+1)
  return (n_nodes + 8) * (3 * 4 + 8) + (4 * n_graphs + 16) * 8 + 1024;
}

struct PrepLayout { int32_t* cam_cnt; int32_t* cam_first; int32_t* order; int64_t* src_off; int64_t* graph_edges; int64_t* graph_tiles; int64_t* graph_unsorted; int64_t* tile_ptr; int64_t* totals; };
static PrepLayout prep_layout(void* scratch, int64_t n_graphs, int64_t n_nodes) {
  char* p = (char*)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
  PrepLayout L;
  const int64_t n8 = (n_nodes + 7) & ~(int64_t)7;
  L.src_off = (int64_t*)p; p += 8 * n8;
  L.graph_edges = (int64_t*)p; p += 8 * (n_graphs + 1);
  L.graph_tiles = (int64_t*)p; p += 8 * (n_graphs + 1);
  L.graph_unsorted = (int64_t*)p; p += 8 * (n_graphs + 1);
  L.tile_ptr = (int64_t*)p; p += 8 * (n_graphs + 2);
  L.totals = (int64_t*)p; p += 8 * 8;
  L.cam_cnt = (int32_t*)p; p += 4 * n8;
  L.cam_first = (int32_t*)p; p += 4 * n8;
  L.order = (int32_t*)p;
  return L;
}

extern "C" int gnncca_graph_edge_counts(const int64_t* graph_ptr, const int32_t* id_cam, int64_t n_graphs, int64_t n_nodes,
                                        int64_t* edge_ptr, void* scratch, int64_t scratch_bytes, void* stream) {
  if (n_graphs < 0 || n_nodes < 0 || !edge_ptr) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (n_graphs == 0) return check_cuda(cudaMemsetAsync(edge_ptr, 0, sizeof(int64_t), st));
  if (!graph_ptr || !id_cam || !scratch) return GNNCCA_ERR_BAD_ARG;
  if (scratch_bytes < gnncca_graph_prepare_scratch_bytes(n_graphs, n_nodes)) return GNNCCA_ERR_WORKSPACE;
  if (n_graphs > INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  PrepLayout L = prep_layout(scratch, n_graphs, n_nodes);
  graph_prepare_kernel<<<(unsigned)n_graphs, 256, 0, st>>>(graph_ptr, id_cam, L.cam_cnt, L.cam_first, L.src_off, L.order,
                                                           L.graph_edges, L.graph_tiles, L.graph_unsorted, GT);
  scan_counts_kernel<<<2, 1024, 0, st>>>(L.graph_edges, L.graph_tiles, L.graph_unsorted, graph_ptr, n_graphs, edge_ptr, L.tile_ptr, L.totals);
  return check_launch(2);
}

extern "C" int gnncca_build_graphs(const int64_t* graph_ptr, const int64_t* edge_ptr, int64_t n_graphs, int64_t n_nodes,
                                   int64_t n_edges, int64_t n_tiles, const int32_t* id_cam, const int64_t* ids, const double* xw,
                                   const double* yw, const double* max_dist, const float* reid, const float* reid_rnorm, int32_t reid_dim,
                                   int32_t attr_mode, int32_t max_frame_nodes, const void* scratch, int64_t* edge_index,
                                   float* edge_attr, float* edge_labels, void* stream) {
  if (n_graphs < 0 || n_nodes < 0 || n_edges < 0 || n_tiles < 0 || max_frame_nodes < 0) return GNNCCA_ERR_BAD_ARG;
  if (attr_mode < GNNCCA_ATTR_FULL || attr_mode > GNNCCA_ATTR_ONLY_DIST) return GNNCCA_ERR_BAD_ARG;
  if (n_edges == 0 || n_tiles == 0) return GNNCCA_OK;
  if (!graph_ptr || !edge_ptr || !id_cam || !ids || !xw || !yw || !max_dist || !scratch || !edge_index || !edge_attr || !edge_labels)
    return GNNCCA_ERR_BAD_ARG;
  if (attr_mode != GNNCCA_ATTR_ONLY_DIST && (!reid || reid_dim <= 0)) return GNNCCA_ERR_BAD_ARG;
  if (n_tiles > INT32_MAX) return GNNCCA_ERR_UNSUPPORTED;
  PrepLayout L = prep_layout(const_cast<void*>(scratch), n_graphs, n_nodes);
  BuildParams p{graph_ptr, edge_ptr, L.tile_ptr, n_graphs, id_cam, ids, xw, yw, max_dist, reid, reid_rnorm, reid_dim, attr_mode,
                L.cam_cnt, L.cam_first, L.src_off, L.order, edge_index, n_edges, edge_attr, edge_labels, 0};
  // a warp per source quad: frames of at most 32 detections are one tile each, and the quads beyond the largest frame never work
  int threads = 256;
  if (GNNCCA_BG_ADAPT && max_frame_nodes > 0 && max_frame_nodes <= GT) threads = 32 * (int)ceil_div(max_frame_nodes, 4);
  if (threads < 128) threads = 128;                   // the staging keeps <= 16 elements per thread in registers
  // every frame non-empty and <= 32 nodes <=> one tile each: the kernel then skips the search for the tile's frame
  p.one_tile_per_frame = (max_frame_nodes > 0 && max_frame_nodes <= GT && n_tiles == n_graphs) ? 1 : 0;
  if (reid_rnorm) build_graphs_kernel<true><<<(unsigned)n_tiles, threads, 0, (cudaStream_t)stream>>>(p);
  else build_graphs_kernel<false><<<(unsigned)n_tiles, threads, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

// ------------------------------------------------------------------------------------------------
// Step 1 of graph construction and both column-norm reductions in one call.  The frame tables, the scans and the ReID norms
// are two launches on `stream` (prep_pass_kernel, prep_final_kernel); the norms of the (8 x larger) node embeddings, which
// nothing in graph construction reads, are forked onto a library-owned side stream so that their HBM-bound pass runs beside the
// issue-bound edge fill.  gnncca_graph_join makes `stream` wait for them.
// ------------------------------------------------------------------------------------------------
struct SideStream { cudaStream_t st; cudaEvent_t fork, join; bool ok; };
static SideStream* side_stream() {
  constexpr int MAX_DEV = 64;
  static SideStream s_side[MAX_DEV];
  static std::atomic<int> s_state[MAX_DEV];      // 0: not created, 1: being created, 2: ready
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) return nullptr;
  int expect = 0;
  if (s_state[dev].compare_exchange_strong(expect, 1)) {
    SideStream& s = s_side[dev];
    s.ok = cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking) == cudaSuccess &&
           cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming) == cudaSuccess;
    s_state[dev].store(2, std::memory_order_release);
  }
  while (s_state[dev].load(std::memory_order_acquire) != 2) { }
  return s_side[dev].ok ? &s_side[dev] : nullptr;
}

extern "C" int gnncca_graph_prepare_all(const int64_t* graph_ptr, const int32_t* id_cam, int64_t n_graphs, int64_t n_nodes,
                                        int64_t* edge_ptr, void* scratch, int64_t scratch_bytes,
                                        const float* node, int32_t node_cols, float* node_norm, float* node_scratch,
                                        const float* reid, int32_t reid_cols, float* reid_rnorm, float* reid_scratch,
                                        int64_t* totals_host, void* stream) {
  if (node && (node_cols <= 0 || !node_norm || !node_scratch)) return GNNCCA_ERR_BAD_ARG;
  if (n_graphs <= 0 || n_nodes <= 0 || n_graphs > INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  if (!graph_ptr || !id_cam || !edge_ptr || !scratch) return GNNCCA_ERR_BAD_ARG;
  if (scratch_bytes < gnncca_graph_prepare_scratch_bytes(n_graphs, n_nodes)) return GNNCCA_ERR_WORKSPACE;
  if (reid && (reid_cols <= 0 || !reid_rnorm || !reid_scratch)) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  int launches = 0;
  SideStream* side = nullptr;
  if (node) {                                           // fork point: the node norms depend on nothing this call launches
    side = side_stream();
    if (!side) { set_error_text("gnncca_graph_prepare_all: could not create the side stream"); return GNNCCA_ERR_LAUNCH; }
    const int rc = check_cuda(cudaEventRecord(side->fork, st));
    if (rc) return rc;
  }
  PrepLayout L = prep_layout(scratch, n_graphs, n_nodes);
  int64_t* host_dev = nullptr;                          // the device's address of the pinned totals words
  if (totals_host) {
    void* dp = nullptr;
    if (cudaHostGetDevicePointer(&dp, totals_host, 0) != cudaSuccess || !dp) {
      (void)cudaGetLastError();
      set_error_text("gnncca_graph_prepare_all: totals_host is not pinned, device-mapped host memory");
      return GNNCCA_ERR_BAD_ARG;
    }
    host_dev = (int64_t*)dp;
  }
  const bool reid_fused = reid && reid_cols % 4 == 0 && reid_cols / 4 <= 256 && ((((uintptr_t)reid | (uintptr_t)reid_scratch) & 15) == 0) &&
                          ceil_div(n_nodes, CN_ROWS_PER_CHUNK) <= 65535;
  if (reid && !reid_fused) {
    set_error_text("gnncca_graph_prepare_all: reid must be 16-byte aligned with reid_cols % 4 == 0 and <= 1024");
    return GNNCCA_ERR_UNSUPPORTED;
  }
  PrepAll p{graph_ptr, id_cam, n_graphs, L.cam_cnt, L.cam_first, L.src_off, L.order, L.graph_edges, L.graph_tiles, L.graph_unsorted,
            edge_ptr, L.tile_ptr, L.totals, host_dev, reid, n_nodes, reid_cols, reid_scratch, reid_rnorm,
            reid_fused ? 1 : 0, reid_fused ? (int)ceil_div(n_nodes, CN_ROWS_PER_CHUNK) : 0};
  const int n_reid = p.reid_bx * p.reid_chunks;
  prep_pass_kernel<<<(unsigned)(n_reid + n_graphs), 256, 0, st>>>(p);
  prep_final_kernel<<<(unsigned)(2 + (reid_fused ? ceil_div(reid_cols, 32) : 0)), 1024, 0, st>>>(p);
  launches += 2;
  int rc = check_launch(launches);
  if (rc || !node) return rc;
  // The side stream's launches come second (what the host waits for, the totals, is in the two launches above) but BEFORE the
  // edge fill is enqueued: the pass then runs while the host sizes and launches the fill.  Launched after the fill it finds
  // every SM's registers taken by the fill's CTAs and runs behind it (measured: 152 us instead of 131 us per build at cfg 2).
  rc = check_cuda(cudaStreamWaitEvent(side->st, side->fork, 0));
  if (!rc) rc = gnncca_colnorm_norms(node, n_nodes, node_cols, node_scratch, node_norm, side->st);
  if (!rc) rc = check_cuda(cudaEventRecord(side->join, side->st));
  return rc;
}

extern "C" int gnncca_graph_join(void* stream) {
  SideStream* side = side_stream();
  if (!side) return GNNCCA_ERR_LAUNCH;
  return check_cuda(cudaStreamWaitEvent((cudaStream_t)stream, side->join, 0));
}


// {n_tiles, n_edges, max nodes per frame, max edges per frame, camera-order flag} of the last gnncca_graph_edge_counts live in the
// scratch, next to each other (5 x int64): the host reads them with ONE copy to size step 2 (and the capacity class of the fused
// message-passing kernel).  [4] != 0 when some frame's rows are not in ascending camera order: the enumeration of train.py:292-299
// (camera by camera) is then not source-major, and the batch must take the checked (sorting) path of the forward.
extern "C" const int64_t* gnncca_graph_tile_total_ptr(const void* scratch, int64_t n_graphs, int64_t n_nodes) {
  PrepLayout L = prep_layout(const_cast<void*>(scratch), n_graphs, n_nodes);
  return L.totals;
}
