// Cross-camera graph + edge-attribute construction (reference train.py:257-259,286-361,450; same code at
// train.py:582-692 and inference.py:189-279): batch-wide column normalisation, per-frame edge
// enumeration (cameras ascending, sources in row order, targets = every detection of another camera in
// row order), the four edge attributes, edge labels, and the Batch.from_data_list index offsets.
// Integer outputs are bit-exact; geometry is float64 like sklearn's paired_distances, cast to fp32 last.
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
__global__ void __launch_bounds__(256) colnorm_final_kernel(const float* __restrict__ partial, int chunks, int cols, float* __restrict__ norm) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float acc = 0.f;
  for (int k = 0; k < chunks; ++k) acc += partial[(int64_t)k * cols + c];   // fixed order
  norm[c] = fmaxf(sqrtf(acc), 1e-12f);                                       // clamp_min(eps), eps = 1e-12
}
__global__ void __launch_bounds__(256) colnorm_apply_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t rows, int cols,
                                                            const float* __restrict__ norm) {
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) out[t] = x[t] / norm[t % cols];
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
                                                            int64_t* __restrict__ graph_edges, int64_t* __restrict__ graph_tiles, int tile) {
  const int g = blockIdx.x;
  const int64_t a = graph_ptr[g];
  const int n = (int)(graph_ptr[g + 1] - a);
  const int32_t* cam = id_cam + a;
  __shared__ unsigned long long s_total;
  if (threadIdx.x == 0) s_total = 0ull;
  __syncthreads();
  unsigned long long mine = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int ci = cam[i];
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
  __syncthreads();
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
  }
}

// exclusive scan of a short int64 array (B+1 outputs), single CTA
__global__ void __launch_bounds__(1024) scan_i64_kernel(const int64_t* __restrict__ in, int64_t n, int64_t* __restrict__ out) {
  __shared__ int64_t s_part[1024];
  const int t = threadIdx.x;
  const int64_t per = (n + blockDim.x - 1) / blockDim.x;
  const int64_t b = min((int64_t)t * per, n), e = min(b + per, n);
  int64_t acc = 0;
  for (int64_t i = b; i < e; ++i) acc += in[i];
  s_part[t] = acc;
  __syncthreads();
  if (t == 0) {
    int64_t run = 0;
    for (int k = 0; k < (int)blockDim.x; ++k) { const int64_t v = s_part[k]; s_part[k] = run; run += v; }
    out[n] = run;
  }
  __syncthreads();
  int64_t run = s_part[t];
  for (int64_t i = b; i < e; ++i) { const int64_t v = in[i]; out[i] = run; run += v; }
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

__global__ void __launch_bounds__(256) build_graphs_kernel(const BuildParams p) {
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

  __shared__ float s_a[GT][GCH + 1];
  __shared__ float s_b[GT][GCH + 1];
  const int tid = threadIdx.x;
  const int tj = tid & 31;        // target within tile
  const int ti0 = tid >> 5;       // first source within tile; this thread owns sources ti0 + 8*q, q = 0..3
  float d2[4] = {0, 0, 0, 0}, dot[4] = {0, 0, 0, 0}, na[4] = {0, 0, 0, 0}, nb = 0.f;
  if (need_app) {
    for (int k0 = 0; k0 < p.reid_dim; k0 += GCH) {
      for (int idx = tid; idx < GT * GCH; idx += 256) {
        const int r = idx / GCH, k = idx % GCH;
        const bool kin = (k0 + k) < p.reid_dim;
        s_a[r][k] = (i0 + r < n && kin) ? p.reid[(a + i0 + r) * p.reid_dim + k0 + k] : 0.f;
        s_b[r][k] = (j0 + r < n && kin) ? p.reid[(a + j0 + r) * p.reid_dim + k0 + k] : 0.f;
      }
      __syncthreads();
      const int kmax = min(GCH, p.reid_dim - k0);
      for (int k = 0; k < kmax; ++k) {
        const float b = s_b[tj][k];
        nb = fmaf(b, b, nb);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float av = s_a[ti0 + 8 * q][k];
          const float df = (av - b) + 1e-6f;          // F.pairwise_distance: ||a - b + eps||_2, eps = 1e-6
          d2[q] = fmaf(df, df, d2[q]);
          dot[q] = fmaf(av, b, dot[q]);
          na[q] = fmaf(av, av, na[q]);
        }
      }
      __syncthreads();
    }
  }
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
    const int i = i0 + ti0 + 8 * q;
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

extern "C" int gnncca_colnorm(const float* emb, float* out, int64_t rows, int32_t cols, float* scratch, void* stream) {
  if (rows < 0 || cols <= 0) return GNNCCA_ERR_BAD_ARG;
  if (rows == 0) return GNNCCA_OK;
  if (!emb || !out || !scratch) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int chunks = (int)ceil_div(rows, CN_ROWS_PER_CHUNK);
  if (chunks > 65535) return GNNCCA_ERR_UNSUPPORTED;
  float* norm = scratch + (int64_t)chunks * cols;
  dim3 g1((unsigned)ceil_div(cols, 256), (unsigned)chunks);
  colnorm_partial_kernel<<<g1, 256, 0, st>>>(emb, rows, cols, scratch);
  colnorm_final_kernel<<<(unsigned)ceil_div(cols, 256), 256, 0, st>>>(scratch, chunks, cols, norm);
  colnorm_apply_kernel<<<grid_for(rows * cols, 256), 256, 0, st>>>(emb, out, rows, cols, norm);
  return check_launch(3);
}

extern "C" int64_t gnncca_graph_prepare_scratch_bytes(int64_t n_graphs, int64_t n_nodes) {
  // cam_cnt, cam_first, order (int32 x N each) + src_off (int64 x N) + graph_edges, graph_tiles (int64 x B) + tile_ptr (B+1)
  return (n_nodes + 8) * (3 * 4 + 8) + (3 * n_graphs + 4) * 8 + 1024;
}

struct PrepLayout { int32_t* cam_cnt; int32_t* cam_first; int32_t* order; int64_t* src_off; int64_t* graph_edges; int64_t* graph_tiles; int64_t* tile_ptr; };
static PrepLayout prep_layout(void* scratch, int64_t n_graphs, int64_t n_nodes) {
  char* p = (char*)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
  PrepLayout L;
  const int64_t n8 = (n_nodes + 7) & ~(int64_t)7;
  L.src_off = (int64_t*)p; p += 8 * n8;
  L.graph_edges = (int64_t*)p; p += 8 * (n_graphs + 1);
  L.graph_tiles = (int64_t*)p; p += 8 * (n_graphs + 1);
  L.tile_ptr = (int64_t*)p; p += 8 * (n_graphs + 2);
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
                                                           L.graph_edges, L.graph_tiles, GT);
  scan_i64_kernel<<<1, 1024, 0, st>>>(L.graph_edges, n_graphs, edge_ptr);
  scan_i64_kernel<<<1, 1024, 0, st>>>(L.graph_tiles, n_graphs, L.tile_ptr);
  return check_launch(3);
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

// tile_ptr[B] (total number of 32x32 tiles) lives in the scratch; expose where so the host can read it with edge_ptr[B]
extern "C" const int64_t* gnncca_graph_tile_total_ptr(const void* scratch, int64_t n_graphs, int64_t n_nodes) {
  PrepLayout L = prep_layout(const_cast<void*>(scratch), n_graphs, n_nodes);
  return L.tile_ptr + n_graphs;
}
