// CSR / index work for the message-passing step: bit-exact integer kernels.
// Replaces the implicit x[row] / x[col] indexing and torch_scatter's index handling
// (reference models/mpn.py:44,48,97-99) with an int32 source-sorted CSR built once per forward.
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"

namespace gnncca {

// One pass over edge_index: narrow to int32, range check, order check and -- valid only when the rows
// turn out to be non-decreasing -- rowptr by boundary detection (edge i opens every row in
// (src[i-1], src[i]]; the last edge closes the rows after it).
// `row` / `col` are the two rows of the caller's edge_index (or a column range of them: a block of whole frames of a larger
// batch); `node_base` is subtracted from every id, so that a block's CSR is local to its own node range.
__global__ void __launch_bounds__(256) csr_narrow_kernel(const int64_t* __restrict__ row, const int64_t* __restrict__ col, int64_t node_base,
                                                         int64_t n_nodes, int64_t n_edges,
                                                         int32_t* __restrict__ src, int32_t* __restrict__ dst,
                                                         int32_t* __restrict__ rowptr, int32_t* __restrict__ flags) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  bool unsorted = false, bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_edges; i += stride) {
    const int64_t r = row[i] - node_base, c = col[i] - node_base;
    const int64_t prev = (i == 0) ? -1 : row[i - 1] - node_base;
    const bool ok = (r >= 0) & (r < n_nodes) & (c >= 0) & (c < n_nodes);
    src[i] = ok ? (int32_t)r : 0;   // out-of-range entries are reported through flags[1]; never let them index memory
    dst[i] = ok ? (int32_t)c : 0;
    bad |= !ok;
    unsorted |= (r < prev);
    if (ok && prev < r && prev >= -1) {
      for (int64_t v = prev + 1; v <= r; ++v) rowptr[v] = (int32_t)i;
    }
    if (ok && i == n_edges - 1) {
      for (int64_t v = r + 1; v <= n_nodes; ++v) rowptr[v] = (int32_t)n_edges;
    }
  }
  if (__syncthreads_or(unsorted) && threadIdx.x == 0) flags[0] = 1;
  if (__syncthreads_or(bad) && threadIdx.x == 0) flags[1] = 1;
}

__global__ void __launch_bounds__(256) rowptr_from_sorted_kernel(const int32_t* __restrict__ src, int64_t n_nodes, int64_t n_edges,
                                                                 int32_t* __restrict__ rowptr) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_edges; i += stride) {
    const int32_t r = src[i];
    const int32_t prev = (i == 0) ? -1 : src[i - 1];
    for (int64_t v = (int64_t)prev + 1; v <= r; ++v) rowptr[v] = (int32_t)i;
    if (i == n_edges - 1)
      for (int64_t v = (int64_t)r + 1; v <= n_nodes; ++v) rowptr[v] = (int32_t)n_edges;
  }
}

__global__ void __launch_bounds__(256) iota_kernel(int32_t* __restrict__ p, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = (int32_t)i;
}

__global__ void __launch_bounds__(256) gather_i32_kernel(const int32_t* __restrict__ in, const int32_t* __restrict__ perm,
                                                         int32_t* __restrict__ out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = in[perm[i]];
}

__global__ void __launch_bounds__(256) permute_rows_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                           const int32_t* __restrict__ perm, int64_t n_rows, int width, int scatter) {
  const int64_t total = n_rows * width;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int64_t i = t / width;
    const int k = (int)(t - i * width);
    const int64_t j = perm[i];
    if (scatter) out[j * width + k] = in[t];
    else out[t] = in[j * width + k];
  }
}

// torch_scatter scatter_add / scatter_mean / scatter_max(...)[0] over source-sorted rows (models/mpn.py:196-202, dim 0): one warp per
// segment, lane = feature (two passes when width > 32), rows folded in edge order -- no atomics, bit-reproducible.  Empty segments give
// 0; mean divides by max(count, 1).
__global__ void __launch_bounds__(256) segment_agg_kernel(const float* __restrict__ rows, const int32_t* __restrict__ rowptr, int64_t n_nodes,
                                                          int width, int agg, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t v = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (v >= n_nodes) return;
  const int j0 = rowptr[v], j1 = rowptr[v + 1];
  for (int c = lane; c < width; c += 32) {
    float acc = agg == GNNCCA_AGG_MAX ? -INFINITY : 0.f;
    for (int j = j0; j < j1; ++j) {
      const float m = rows[(int64_t)j * width + c];
      acc = agg == GNNCCA_AGG_MAX ? fmaxf(acc, m) : acc + m;
    }
    if (j1 <= j0) acc = 0.f;
    else if (agg == GNNCCA_AGG_MEAN) acc = acc / (float)(j1 - j0);
    out[v * width + c] = acc;
  }
}

static int key_bits(int64_t n_nodes) {
  int b = 1;
  while (((int64_t)1 << b) < n_nodes && b < 31) ++b;
  return b;
}
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace gnncca

using namespace gnncca;

extern "C" int gnncca_csr_from_edge_index(const int64_t* edge_index, int64_t n_nodes, int64_t n_edges, int32_t* src,
                                          int32_t* dst, int32_t* rowptr, int32_t* flags, void* stream) {
  if (n_nodes < 0 || n_edges < 0 || n_nodes >= INT32_MAX || n_edges >= INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  if (!rowptr || !flags || (n_edges > 0 && (!edge_index || !src || !dst))) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  int rc = check_cuda(cudaMemsetAsync(flags, 0, sizeof(int32_t) * GNNCCA_CSR_NFLAGS, st));
  if (rc) return rc;
  if (n_edges == 0) return check_cuda(cudaMemsetAsync(rowptr, 0, sizeof(int32_t) * (n_nodes + 1), st));
  csr_narrow_kernel<<<grid_for(n_edges, 256), 256, 0, st>>>(edge_index, edge_index + n_edges, 0, n_nodes, n_edges, src, dst, rowptr, flags);
  return check_launch();
}

extern "C" int gnncca_csr_from_edge_rows(const int64_t* row, const int64_t* col, int64_t node_base, int64_t n_nodes, int64_t n_edges,
                                         int32_t* src, int32_t* dst, int32_t* rowptr, int32_t* flags, void* stream) {
  if (n_nodes < 0 || n_edges < 0 || n_nodes >= INT32_MAX || n_edges >= INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  if (!rowptr || !flags || (n_edges > 0 && (!row || !col || !src || !dst))) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  int rc = check_cuda(cudaMemsetAsync(flags, 0, sizeof(int32_t) * GNNCCA_CSR_NFLAGS, st));
  if (rc) return rc;
  if (n_edges == 0) return check_cuda(cudaMemsetAsync(rowptr, 0, sizeof(int32_t) * (n_nodes + 1), st));
  csr_narrow_kernel<<<grid_for(n_edges, 256), 256, 0, st>>>(row, col, node_base, n_nodes, n_edges, src, dst, rowptr, flags);
  return check_launch();
}

// Reverse-edge table of a SYMMETRIC source-major graph whose segments list their targets in ascending order (what the
// camera-by-camera enumeration of train.py:292-299 gives for camera-major rows): position q holds the edge (j -> i); its
// reverse (i -> j) sits in i's segment at the lower bound of j.  That table IS the target-major permutation a stable sort by
// dst would produce (the in-edges of j in ascending source order are the reverses of j's out-edges in order), and the
// target-major offsets equal rowptr.  flags[3] is raised when some edge has no reverse (the table entry is then q itself).
__global__ void __launch_bounds__(256) csr_reverse_edges_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src,
                                                                const int32_t* __restrict__ dst, int64_t n_edges, int32_t* __restrict__ tperm,
                                                                int32_t* __restrict__ flags) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_edges; q += stride) {
    const int j = src[q], i = dst[q];
    int lo = rowptr[i], hi = rowptr[i + 1];
    const int end = hi;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (dst[mid] < j) lo = mid + 1; else hi = mid; }
    if (lo < end && dst[lo] == j) {
      tperm[q] = lo;
    } else {
      tperm[q] = (int32_t)q;
      flags[3] = 1;
    }
  }
}

extern "C" int gnncca_csr_reverse_edges(const int32_t* rowptr, const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges,
                                        int32_t* tperm, int32_t* flags, void* stream) {
  if (n_nodes < 0 || n_edges < 0 || n_edges >= INT32_MAX) return GNNCCA_ERR_BAD_ARG;
  if (n_edges == 0) return GNNCCA_OK;
  if (!rowptr || !src || !dst || !tperm || !flags) return GNNCCA_ERR_BAD_ARG;
  csr_reverse_edges_kernel<<<grid_for(n_edges, 256), 256, 0, (cudaStream_t)stream>>>(rowptr, src, dst, n_edges, tperm, flags);
  return check_launch();
}

extern "C" int64_t gnncca_csr_sort_workspace_bytes(int64_t n_nodes, int64_t n_edges) {
  if (n_edges <= 0) return 256;
  size_t temp = 0;
  cub::DoubleBuffer<int32_t> k(nullptr, nullptr), v(nullptr, nullptr);
  cub::DeviceRadixSort::SortPairs(nullptr, temp, k, v, (int)n_edges, 0, key_bits(n_nodes));
  return (int64_t)(align256(temp) + 3 * align256(sizeof(int32_t) * n_edges) + 256);
}

extern "C" int gnncca_csr_sort(int64_t n_nodes, int64_t n_edges, int32_t* src, int32_t* dst, int32_t* perm,
                               int32_t* rowptr, void* workspace, int64_t workspace_bytes, void* stream) {
  if (n_nodes < 0 || n_edges < 0 || !rowptr) return GNNCCA_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (n_edges == 0) return check_cuda(cudaMemsetAsync(rowptr, 0, sizeof(int32_t) * (n_nodes + 1), st));
  if (!src || !dst || !perm || !workspace) return GNNCCA_ERR_BAD_ARG;
  if (workspace_bytes < gnncca_csr_sort_workspace_bytes(n_nodes, n_edges)) return GNNCCA_ERR_WORKSPACE;
  const size_t ebytes = align256(sizeof(int32_t) * n_edges);
  char* ws = (char*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  int32_t* key_alt = (int32_t*)ws;
  int32_t* val_a = (int32_t*)(ws + ebytes);
  int32_t* val_b = (int32_t*)(ws + 2 * ebytes);
  void* temp = ws + 3 * ebytes;
  size_t temp_bytes = 0;
  cub::DoubleBuffer<int32_t> k(src, key_alt), v(val_a, val_b);
  cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, k, v, (int)n_edges, 0, key_bits(n_nodes));
  const int g = grid_for(n_edges, 256);
  iota_kernel<<<g, 256, 0, st>>>(val_a, n_edges);
  int rc = check_cuda(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, k, v, (int)n_edges, 0, key_bits(n_nodes), st));
  if (rc) return rc;
  // results may sit in either half of the double buffers
  rc = check_cuda(cudaMemcpyAsync(perm, v.Current(), sizeof(int32_t) * n_edges, cudaMemcpyDeviceToDevice, st));
  if (rc) return rc;
  if (k.Current() != src)
    rc = check_cuda(cudaMemcpyAsync(src, k.Current(), sizeof(int32_t) * n_edges, cudaMemcpyDeviceToDevice, st));
  if (rc) return rc;
  // dst in sorted order (val_a/val_b are free again once perm is copied; use whichever is not Current)
  int32_t* tmp = (v.Current() == val_a) ? val_b : val_a;
  gather_i32_kernel<<<g, 256, 0, st>>>(dst, perm, tmp, n_edges);
  rc = check_cuda(cudaMemcpyAsync(dst, tmp, sizeof(int32_t) * n_edges, cudaMemcpyDeviceToDevice, st));
  if (rc) return rc;
  rowptr_from_sorted_kernel<<<g, 256, 0, st>>>(src, n_nodes, n_edges, rowptr);
  return check_launch(3);
}

extern "C" int gnncca_segment_agg(const float* rows, const int32_t* rowptr, int64_t n_nodes, int32_t width, int32_t agg, float* out,
                                  void* stream) {
  if (n_nodes < 0 || width <= 0 || agg < GNNCCA_AGG_SUM || agg > GNNCCA_AGG_MAX) return GNNCCA_ERR_BAD_ARG;
  if (n_nodes == 0) return GNNCCA_OK;
  if (!rowptr || !out) return GNNCCA_ERR_BAD_ARG;
  segment_agg_kernel<<<(unsigned)ceil_div(n_nodes, (int64_t)8), 256, 0, (cudaStream_t)stream>>>(rows, rowptr, n_nodes, width, agg, out);
  return check_launch();
}

extern "C" int gnncca_permute_rows(const float* in, float* out, const int32_t* perm, int64_t n_rows, int32_t width,
                                   int32_t scatter, void* stream) {
  if (n_rows < 0 || width <= 0) return GNNCCA_ERR_BAD_ARG;
  if (n_rows == 0) return GNNCCA_OK;
  if (!in || !out || !perm || in == out) return GNNCCA_ERR_BAD_ARG;
  permute_rows_kernel<<<grid_for(n_rows * width, 256), 256, 0, (cudaStream_t)stream>>>(in, out, perm, n_rows, width, scatter);
  return check_launch();
}
