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

__global__ void __launch_bounds__(256) colnorm_partial_kernel(const float* __restrict__ x, int64_t rows, int cols, float* __restrict__ partial) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * CN_ROWS_PER_CHUNK;
  const int64_t r1 = min(r0 + CN_ROWS_PER_CHUNK, rows);
  if (c >= cols) return;
  float acc = 0.f;
  for (int64_t r = r0; r < r1; ++r) { const float v = x[r * cols + c]; acc = fmaf(v, v, acc); }
  partial[(int64_t)blockIdx.y * cols + c] = acc;
}
// cols % 4 == 0 and 16-byte aligned rows: a thread owns 4 adjacent columns, 4 rows in flight (same per-column row order)
__global__ void __launch_bounds__(256) colnorm_partial4_kernel(const float* __restrict__ x, int64_t rows, int cols, float* __restrict__ partial) {
  const int c4 = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * CN_ROWS_PER_CHUNK;
  const int64_t r1 = min(r0 + CN_ROWS_PER_CHUNK, rows);
  if (c4 * 4 >= cols) return;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const float* px = x + (int64_t)c4 * 4;
  int64_t r = r0;
  for (; r + 4 <= r1; r += 4) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ld_stream_f4(px + (r + u) * cols);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      acc.x = fmaf(v[u].x, v[u].x, acc.x); acc.y = fmaf(v[u].y, v[u].y, acc.y);
      acc.z = fmaf(v[u].z, v[u].z, acc.z); acc.w = fmaf(v[u].w, v[u].w, acc.w);
    }
  }
  for (; r < r1; ++r) {
    const float4 v = ld_stream_f4(px + r * cols);
    acc.x = fmaf(v.x, v.x, acc.x); acc.y = fmaf(v.y, v.y, acc.y); acc.z = fmaf(v.z, v.z, acc.z); acc.w = fmaf(v.w, v.w, acc.w);
  }
  *reinterpret_cast<float4*>(partial + (int64_t)blockIdx.y * cols + c4 * 4) = acc;
}
// 32 columns per block; warp w sums the chunks w, w + 8, ... of its lane's column, the 8 partial sums are folded in warp order
__global__ void __launch_bounds__(256) colnorm_final_kernel(const float* __restrict__ partial, int chunks, int cols, float* __restrict__ norm) {
  __shared__ float s_part[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (c < cols) {
#pragma unroll 4
    for (int k = warp; k < chunks; k += 8) acc += __ldg(partial + (int64_t)k * cols + c);
  }
  s_part[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < cols) {
    float tot = s_part[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) tot += s_part[w][lane];
    norm[c] = fmaxf(sqrtf(tot), 1e-12f);                                     // clamp_min(eps), eps = 1e-12
  }
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
__global__ void __launch_bounds__(256) graph_prepare_kernel(const int64_t* __restrict__ graph_ptr, const int32_t* __restrict__ id_cam,
                                                            int32_t* __restrict__ cam_cnt, int32_t* __restrict__ cam_first,
                                                            int64_t* __restrict__ src_off, int32_t* __restrict__ order,
                                                            int64_t* __restrict__ graph_edges, int64_t* __restrict__ graph_tiles,
                                                            int64_t* __restrict__ graph_unsorted, int tile) {
  const int g = blockIdx.x;
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

// exclusive scan of a short int64 array (B+1 outputs), single CTA
// Exclusive scans of the per-frame edge and tile counts in ONE launch (block 0: edges -> edge_ptr, block 1: tiles ->
// tile_ptr) plus the four numbers the host needs to size the outputs, next to each other for a single read back:
// totals = {n_tiles, n_edges, max nodes per frame, max edges per frame, #frames whose rows are not camera-ascending}.
__global__ void __launch_bounds__(1024) scan_counts_kernel(const int64_t* __restrict__ graph_edges, const int64_t* __restrict__ graph_tiles,
                                                           const int64_t* __restrict__ graph_unsorted,
                                                           const int64_t* __restrict__ graph_ptr, int64_t n, int64_t* __restrict__ edge_ptr,
                                                           int64_t* __restrict__ tile_ptr, int64_t* __restrict__ totals) {
  __shared__ int64_t s_warp[32], s_max[32];
  const int64_t* in = blockIdx.x == 0 ? graph_edges : graph_tiles;
  int64_t* out = blockIdx.x == 0 ? edge_ptr : tile_ptr;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t per = (n + blockDim.x - 1) / blockDim.x;
  const int64_t b = min((int64_t)t * per, n), e = min(b + per, n);
  int64_t acc = 0, mx = 0;
  int uns = 0;
  for (int64_t i = b; i < e; ++i) {
    const int64_t v = in[i];
    acc += v;
    uns += (int)graph_unsorted[i];
    mx = max(mx, blockIdx.x == 0 ? v : graph_ptr[i + 1] - graph_ptr[i]);     // block 0: max edges, block 1: max nodes per frame
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
    totals[blockIdx.x == 0 ? 1 : 0] = total;       // [0] tiles, [1] edges
    totals[blockIdx.x == 0 ? 3 : 2] = s_max[0];    // [2] max nodes, [3] max edges
    if (blockIdx.x == 0) totals[4] = n_uns;        // [4] != 0: some frame's rows are not camera-ascending
  }
}

// ------------------------------------------------------------------------------------------------
// Edge fill: CTA = 32 sources x 32 targets of one frame; reid rows staged in shared memory in chunks.
// ------------------------------------------------------------------------------------------------
constexpr int GT = 32;       // tile edge
constexpr int GCH = 64;      // feature chunk

struct BuildParams {
  const int64_t* graph_ptr; const int64_t* edge_ptr; const int64_t* tile_ptr; int64_t n_graphs;
  const int32_t* id_cam; const int64_t* ids; const double* xw; const double* yw; const double* max_dist;
  const float* reid; int reid_dim; int attr_mode;
  const int32_t* cam_cnt; const int32_t* cam_first; const int64_t* src_off; const int32_t* order;
  int64_t* edge_index; int64_t n_edges; float* edge_attr; float* edge_labels;
};

__global__ void __launch_bounds__(256, 3) build_graphs_kernel(const BuildParams p) {
  // which frame does this tile belong to: binary search in tile_ptr
  const int64_t tb = blockIdx.x;
  int64_t lo = 0, hi = p.n_graphs;  // invariant: tile_ptr[lo] <= tb < tile_ptr[hi]
  while (hi - lo > 1) { const int64_t mid = (lo + hi) >> 1; if (p.tile_ptr[mid] <= tb) lo = mid; else hi = mid; }
  const int64_t g = lo;
  const int64_t a = p.graph_ptr[g];
  const int n = (int)(p.graph_ptr[g + 1] - a);
  const int nt = (n + GT - 1) / GT;
  const int local = (int)(tb - p.tile_ptr[g]);
  const int i0 = (local / nt) * GT, j0 = (local % nt) * GT;
  const bool need_app = p.attr_mode != GNNCCA_ATTR_ONLY_DIST;

  // ReID rows of the tile, feature-major ([k][row], row stride 36 floats = 16-byte aligned quads): the 4 sources of a thread
  // are one LDS.128 (the same address for the whole warp), the target is lane-consecutive; arithmetic on source PAIRS (FFMA2)
  __shared__ __align__(16) float s_a[GCH][GT + 4];
  __shared__ __align__(16) float s_b[GCH][GT + 4];
  const int tid = threadIdx.x;
  const int tj = tid & 31;        // target within tile
  const int ti0 = (tid >> 5) * 4; // this thread owns sources ti0 .. ti0 + 3 of the tile
  float2 d2a = make_float2(0.f, 0.f), d2b = d2a, dota = d2a, dotb = d2a, naa = d2a, nab = d2a;
  float nb = 0.f;
  if (need_app) {
    for (int k0 = 0; k0 < p.reid_dim; k0 += GCH) {
      for (int idx = tid; idx < GT * GCH; idx += 256) {
        const int r = idx / GCH, k = idx % GCH;
        const bool kin = (k0 + k) < p.reid_dim;
        s_a[k][r] = (i0 + r < n && kin) ? p.reid[(a + i0 + r) * p.reid_dim + k0 + k] : 0.f;
        s_b[k][r] = (j0 + r < n && kin) ? p.reid[(a + j0 + r) * p.reid_dim + k0 + k] : 0.f;
      }
      __syncthreads();
      const int kmax = (i0 + ti0 < n) ? min(GCH, p.reid_dim - k0) : 0;     // a warp whose four sources lie beyond the frame only stages
#pragma unroll 4
      for (int k = 0; k < kmax; ++k) {
        const float b = s_b[k][tj];
        const float4 av = *reinterpret_cast<const float4*>(&s_a[k][ti0]);
        const float2 a01 = make_float2(av.x, av.y), a23 = make_float2(av.z, av.w);
        nb = fmaf(b, b, nb);
        // F.pairwise_distance: ||a - b + eps||_2, eps = 1e-6 ((a - b) first, then + eps, as torch does)
        const float2 df01 = pk::fadd2(pk::fadd2(a01, pk::bc(-b)), pk::bc(1e-6f));
        const float2 df23 = pk::fadd2(pk::fadd2(a23, pk::bc(-b)), pk::bc(1e-6f));
        d2a = pk::ffma2(df01, df01, d2a); d2b = pk::ffma2(df23, df23, d2b);
        dota = pk::ffma2(a01, pk::bc(b), dota); dotb = pk::ffma2(a23, pk::bc(b), dotb);
        naa = pk::ffma2(a01, a01, naa); nab = pk::ffma2(a23, a23, nab);
      }
      __syncthreads();
    }
  }
  const float d2[4] = {d2a.x, d2a.y, d2b.x, d2b.y}, dot[4] = {dota.x, dota.y, dotb.x, dotb.y}, na[4] = {naa.x, naa.y, nab.x, nab.y};
  const int j = j0 + tj;
  if (j >= n) return;
  const int cj = p.id_cam[a + j];
  const double xj = p.xw[a + j], yj = p.yw[a + j];
  const int64_t idj = p.ids[a + j];
  const double md = p.max_dist[g];
  const int64_t ebase = p.edge_ptr[g];
  const int width = (p.attr_mode == GNNCCA_ATTR_FULL) ? 4 : 2;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int i = i0 + ti0 + q;
    if (i >= n) continue;
    const int ci = p.id_cam[a + i];
    if (ci == cj) continue;
    // targets of source i are the detections of other cameras in row order: j's slot is
    // j - #{j' < j : cam(j') == cam(i)}; that count is a lower_bound inside i's camera block of order[]
    const int32_t* blk = p.order + a + p.cam_first[a + i];
    int l = 0, h = p.cam_cnt[a + i];
    while (l < h) { const int m = (l + h) >> 1; if (blk[m] < j) l = m + 1; else h = m; }
    const int64_t e = ebase + p.src_off[a + i] + (j - l);
    p.edge_index[e] = a + i;                 // rebased to the frame (train.py:356) + Batch offset (train.py:450)
    p.edge_index[p.n_edges + e] = a + j;
    p.edge_labels[e] = (p.ids[a + i] == idj) ? 1.f : 0.f;   // train.py:348-351
    float* out = p.edge_attr + e * width;
    int col = 0;
    if (p.attr_mode != GNNCCA_ATTR_ONLY_APPEARANCE) {
      const double dx = p.xw[a + i] - xj, dy = p.yw[a + i] - yj;
      const double l2 = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));   // paired_distances euclidean (float64)
      const double l1 = __dadd_rn(fabs(dx), fabs(dy));                            // manhattan
      out[col++] = (float)(l2 / md);      // train.py:323, cast at :344
      out[col++] = (float)(l1 / md);      // train.py:326
    }
    if (need_app) {
      out[col++] = sqrtf(d2[q]);                                                  // train.py:306
      out[col++] = dot[q] / (fmaxf(sqrtf(na[q]), 1e-8f) * fmaxf(sqrtf(nb), 1e-8f));   // train.py:308, eps = 1e-8
    }
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
  colnorm_final_kernel<<<(unsigned)ceil_div(cols, 32), 256, 0, st>>>(scratch, chunks, cols, norm);
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
                                   const double* yw, const double* max_dist, const float* reid, int32_t reid_dim,
                                   int32_t attr_mode, const void* scratch, int64_t* edge_index, float* edge_attr,
                                   float* edge_labels, void* stream) {
  if (n_graphs < 0 || n_nodes < 0 || n_edges < 0 || n_tiles < 0) return GNNCCA_ERR_BAD_ARG;
  if (attr_mode < GNNCCA_ATTR_FULL || attr_mode > GNNCCA_ATTR_ONLY_DIST) return GNNCCA_ERR_BAD_ARG;
  if (n_edges == 0 || n_tiles == 0) return GNNCCA_OK;
  if (!graph_ptr || !edge_ptr || !id_cam || !ids || !xw || !yw || !max_dist || !scratch || !edge_index || !edge_attr || !edge_labels)
    return GNNCCA_ERR_BAD_ARG;
  if (attr_mode != GNNCCA_ATTR_ONLY_DIST && (!reid || reid_dim <= 0)) return GNNCCA_ERR_BAD_ARG;
  if (n_tiles > INT32_MAX) return GNNCCA_ERR_UNSUPPORTED;
  PrepLayout L = prep_layout(const_cast<void*>(scratch), n_graphs, n_nodes);
  BuildParams p{graph_ptr, edge_ptr, L.tile_ptr, n_graphs, id_cam, ids, xw, yw, max_dist, reid, reid_dim, attr_mode,
                L.cam_cnt, L.cam_first, L.src_off, L.order, edge_index, n_edges, edge_attr, edge_labels};
  build_graphs_kernel<<<(unsigned)n_tiles, 256, 0, (cudaStream_t)stream>>>(p);
  return check_launch();
}

// {n_tiles, n_edges, max nodes per frame, max edges per frame, camera-order flag} of the last gnncca_graph_edge_counts live in the
// scratch, next to each other (5 x int64): the host reads them with ONE copy to size step 2 (and the capacity class of the fused
// message-passing kernel).  [4] != 0 when some frame's rows are not in ascending camera order: the enumeration of train.py:292-299
// (camera by camera) is then not source-major, and the batch must take the checked (sorting) path of the forward.
extern "C" const int64_t* gnncca_graph_tile_total_ptr(const void* scratch, int64_t n_graphs, int64_t n_nodes) {
  PrepLayout L = prep_layout(const_cast<void*>(scratch), n_graphs, n_nodes);
  return L.totals;
}
