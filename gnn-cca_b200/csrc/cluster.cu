// Edge decisions, pruning of one-directional edges and strongly connected components per frame graph, on the device:
// the consumer of the MPN's logits in inference (inference.py:286-345; utils.py:295-317 compute_SCC_and_Clusters,
// utils.py:387-404 remove_edges_single_direction).  SURVEY.md section 8(f) rank 2.  The reference does this with
// Python list comprehensions + networkx per batch (O(E^2) tuple-membership tests in the pruning).
//
// One warp per frame graph (frames are disjoint blocks of the batch graph, <= 64 nodes: EPFL frames have <= 40):
//   A[u]  = 64-bit adjacency row of the ACTIVE edges u -> v (integer atomicOr in shared memory: order-independent),
//   prune : keep u -> v iff v -> u is active too (A[u] & transpose(A)[u]),
//   R     = reflexive-transitive closure of A by Warshall over bit rows (n rounds, one broadcast row per round),
//   SCC(u) = R[u] & transpose(R)[u]; its lowest set bit is the component's smallest node = its representative.
// Cluster ids are dense over the batch in order of the representatives' node index (the reference's numbering comes
// out of networkx's traversal order and is only ever consumed by permutation-invariant clustering metrics,
// inference.py:347-352, and by the count of clusters).  Integer / bit arithmetic only: results are exact.
#include "common.cuh"

namespace gnncca {
namespace cluster {

constexpr int MAX_N = 64;
constexpr int WARPS = 8;
enum { KIND_LOGITS = 0, KIND_LABELS = 1, KIND_FLAGS = 2 };

struct Params {
  const void* decision; int kind;
  const int64_t* edge_index; int64_t n_edges;
  const int64_t* graph_ptr; const int64_t* edge_ptr; int64_t n_graphs; int64_t n_nodes;
  int prune;
  int64_t* pred; int32_t* local_id; int32_t* counts; int32_t* flags;
};

__device__ __forceinline__ bool active_edge(const Params& p, int64_t e) {
  if (p.kind == KIND_LOGITS) {   // (sigmoid(x) >= 0.5) * 1, inference.py:289-291, with fp32 sigmoid as torch computes it
    const float x = ((const float*)p.decision)[e];
    return 1.f / (1.f + expf(-x)) >= 0.5f;
  }
  if (p.kind == KIND_LABELS) return ((const float*)p.decision)[e] == 1.f;      // `if p == 1`, inference.py:297
  return ((const int64_t*)p.decision)[e] == 1;
}

__global__ void __launch_bounds__(WARPS * 32) frames_kernel(const Params p) {
  __shared__ unsigned long long A[WARPS][MAX_N], R[WARPS][MAX_N];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t g = (int64_t)blockIdx.x * WARPS + warp;
  if (g >= p.n_graphs) return;
  const int64_t n0 = p.graph_ptr[g], n1 = p.graph_ptr[g + 1], e0 = p.edge_ptr[g], e1 = p.edge_ptr[g + 1];
  const int n = (int)(n1 - n0);
  if (n < 0 || n > MAX_N || e1 < e0 || n1 > p.n_nodes || e1 > p.n_edges) {      // uniform per warp
    if (lane == 0) atomicOr(&p.flags[n > MAX_N ? 1 : 0], 1);
    return;
  }
  unsigned long long* a = A[warp];
  unsigned long long* r = R[warp];
  a[lane] = 0ull; a[lane + 32] = 0ull;
  __syncwarp();
  const int64_t* src = p.edge_index;
  const int64_t* dst = p.edge_index + p.n_edges;
  bool bad = false;
  for (int64_t e = e0 + lane; e < e1; e += 32) {
    const int64_t u = src[e] - n0, v = dst[e] - n0;
    if (u < 0 || u >= n || v < 0 || v >= n) { bad = true; continue; }
    if (active_edge(p, e)) atomicOr(&a[u], 1ull << v);
  }
  if (__any_sync(0xffffffffu, bad)) {
    if (lane == 0) atomicOr(&p.flags[0], 1);
    return;
  }
  __syncwarp();
  // transpose of A for this lane's two rows: at[i] bit j = A[j] bit i
  unsigned long long at0 = 0ull, at1 = 0ull;
  for (int j = 0; j < n; ++j) {
    const unsigned long long row = a[j];
    at0 |= ((row >> lane) & 1ull) << j;
    at1 |= ((row >> (lane + 32)) & 1ull) << j;
  }
  __syncwarp();
  if (p.prune) {            // utils.py:388: an active edge whose reverse is not active is removed
    a[lane] &= at0; a[lane + 32] &= at1;
    __syncwarp();
  }
  if (p.pred) {
    for (int64_t e = e0 + lane; e < e1; e += 32) {
      const int u = (int)(src[e] - n0), v = (int)(dst[e] - n0);
      // a[u] holds the kept (u, v) pairs; the edge's own decision is and-ed in so that an inactive duplicate stays 0
      p.pred[e] = (active_edge(p, e) && ((a[u] >> v) & 1ull)) ? 1 : 0;
    }
  }
  // closure (Warshall): after round k, R[i] holds the nodes reachable from i through intermediates < = k
  unsigned long long r0 = a[lane] | (1ull << lane), r1 = a[lane + 32] | (1ull << (lane + 32));
  r[lane] = r0; r[lane + 32] = r1;
  __syncwarp();
  for (int k = 0; k < n; ++k) {
    const unsigned long long rk = r[k];
    if ((r0 >> k) & 1ull) r0 |= rk;
    if ((r1 >> k) & 1ull) r1 |= rk;
    __syncwarp();
    r[lane] = r0; r[lane + 32] = r1;
    __syncwarp();
  }
  unsigned long long t0 = 0ull, t1 = 0ull;
  for (int j = 0; j < n; ++j) {
    const unsigned long long row = r[j];
    t0 |= ((row >> lane) & 1ull) << j;
    t1 |= ((row >> (lane + 32)) & 1ull) << j;
  }
  const int root0 = __ffsll((long long)(r0 & t0)) - 1, root1 = __ffsll((long long)(r1 & t1)) - 1;   // >= 0: i is in its own SCC
  const unsigned lo = __ballot_sync(0xffffffffu, lane < n && root0 == lane);
  const unsigned hi = __ballot_sync(0xffffffffu, lane + 32 < n && root1 == lane + 32);
  const unsigned long long roots = (unsigned long long)lo | ((unsigned long long)hi << 32);
  if (lane < n) p.local_id[n0 + lane] = __popcll(roots & ((1ull << root0) - 1ull));
  if (lane + 32 < n) p.local_id[n0 + lane + 32] = __popcll(roots & ((1ull << root1) - 1ull));
  if (lane == 0) p.counts[g] = __popcll(roots);
}

// ids[v] = (number of clusters in the frames before v's) + local_id[v]; n_clusters[0] = total.  A block owns WARPS
// consecutive frames and first sums the counts of all earlier frames (G ints, coalesced: G is a few thousand).
__global__ void __launch_bounds__(WARPS * 32) number_kernel(const int64_t* __restrict__ graph_ptr, int64_t n_graphs, const int32_t* __restrict__ counts,
                                                            const int32_t* __restrict__ local_id, int64_t* __restrict__ ids, int64_t* __restrict__ n_clusters) {
  __shared__ long long warp_sum[WARPS];
  __shared__ long long base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t g0 = (int64_t)blockIdx.x * WARPS;
  long long s = 0;
  for (int64_t i = threadIdx.x; i < g0; i += blockDim.x) s += counts[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) warp_sum[warp] = s;
  __syncthreads();
  if (threadIdx.x == 0) { long long b = 0; for (int w = 0; w < WARPS; ++w) b += warp_sum[w]; base_s = b; }
  __syncthreads();
  long long base = base_s;
  const int64_t g = g0 + warp;
  for (int64_t i = g0; i < g && i < n_graphs; ++i) base += counts[i];
  if (g >= n_graphs) return;
  const int64_t n0 = graph_ptr[g], n1 = graph_ptr[g + 1];
  for (int64_t v = n0 + lane; v < n1; v += 32) ids[v] = base + local_id[v];
  if (g == n_graphs - 1 && lane == 0) n_clusters[0] = base + counts[g];
}

}  // namespace cluster
}  // namespace gnncca

using namespace gnncca;

extern "C" int gnncca_frame_clusters(const void* decision, int32_t kind, const int64_t* edge_index, int64_t n_edges, const int64_t* graph_ptr,
                                     const int64_t* edge_ptr, int64_t n_graphs, int64_t n_nodes, int32_t prune, int64_t* pred,
                                     int64_t* ids, int64_t* n_clusters, int32_t* scratch, int32_t* flags, void* stream) {
  if (n_graphs < 0 || n_nodes < 0 || n_edges < 0 || kind < 0 || kind > 2 || !flags || !n_clusters) return GNNCCA_ERR_BAD_ARG;
  if (n_graphs == 0) return n_nodes == 0 ? GNNCCA_OK : GNNCCA_ERR_BAD_ARG;
  if (!graph_ptr || !edge_ptr || !ids || !scratch || (n_edges > 0 && (!decision || !edge_index))) return GNNCCA_ERR_BAD_ARG;
  cluster::Params p{};
  p.decision = decision; p.kind = kind; p.edge_index = edge_index; p.n_edges = n_edges; p.graph_ptr = graph_ptr; p.edge_ptr = edge_ptr;
  p.n_graphs = n_graphs; p.n_nodes = n_nodes; p.prune = prune; p.pred = pred; p.local_id = scratch; p.counts = scratch + n_nodes; p.flags = flags;
  cudaStream_t st = (cudaStream_t)stream;
  const int blocks = (int)ceil_div(n_graphs, cluster::WARPS);
  cluster::frames_kernel<<<blocks, cluster::WARPS * 32, 0, st>>>(p);
  cluster::number_kernel<<<blocks, cluster::WARPS * 32, 0, st>>>(graph_ptr, n_graphs, p.counts, p.local_id, ids, n_clusters);
  return check_launch(2);
}
