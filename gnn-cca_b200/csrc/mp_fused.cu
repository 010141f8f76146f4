// All message-passing steps of MOTMPNet.forward (reference models/mpn.py:270-299) for batches of SMALL graphs,
// as ONE kernel.  One persistent CTA per SM runs up to 8 independent warp GROUPS; a group takes one frame graph at
// a time and keeps its node and edge latents in shared memory across the L steps (mp_frame.cuh has the per-frame
// routine and its arithmetic).  EPFL frames are 2..40 detections / <= 1200 directed edges (~30 KB of state), so
// nothing of size E is ever written between steps: compulsory traffic is 44 B/edge instead of 268 B/edge.
//
// Two ways in: RAW (the caller's int64 edge_index + per-frame edge offsets; source order, range and closedness
// of every frame are verified on the fly) or CSR (the int32 arrays of gnncca_csr_from_edge_index).  A violation
// raises flags[2] and the host re-runs the batch through the general per-step kernels.
#include "mp_frame.cuh"

namespace gnncca {
namespace fused {

struct Params {
  mpf::Net net;
  mpf::Batch batch;
  int groups, group_threads;
};

template <bool REATT_E, int AGG>
__global__ void __launch_bounds__(896, 1) mp_fused_fwd_kernel(const __grid_constant__ Params p) {
  extern __shared__ __align__(16) float smem_f[];
  if (!p.batch.edge_index && (p.batch.flags[0] | p.batch.flags[1]) != 0) return;   // CSR not valid: uniform exit
  const int K = p.net.reatt_n ? 2 * mpf::FN : mpf::FN, EIN = REATT_E ? 2 * mpf::FE : mpf::FE;
  const mpf::WOff wo = mpf::weight_layout(K, EIN);
  const mpf::SOff so = mpf::state_layout(p.net.cap_n, p.net.cap_e, p.net.reatt_n != 0, REATT_E);
  mpf::load_weights(smem_f, wo, p.net, threadIdx.x, blockDim.x);
  __syncthreads();
  // Launched as a programmatic dependent of the node encoder (gnncca_mpn_forward_frames): the weights above AND the edge side
  // of every group's first frame (edge list check, local CSR, edge encoder: inputs of the forward, not products of the
  // encoder) overlap the encoder's tail; mp_frame waits for the encoder right before it reads the encoded nodes.
  // griddepcontrol.wait is a no-op for ordinary launches.
  mpf::Group gr;
  const int gi = threadIdx.x / p.group_threads;
  gr.tid = threadIdx.x - gi * p.group_threads; gr.size = p.group_threads;
  gr.warp = gr.tid >> 5; gr.nwarps = gr.size >> 5; gr.lane = gr.tid & 31; gr.bar = 1 + gi;
  float* ss = smem_f + wo.total + gi * so.total;
  const int64_t stride = (int64_t)gridDim.x * p.groups;
  bool first = true;
  for (int64_t g = (int64_t)gi * gridDim.x + blockIdx.x; g < p.batch.n_graphs; g += stride, first = false)
    mpf::mp_frame<REATT_E, AGG>(gr, p.net, p.batch, smem_f, wo, ss, so, g, first);
}

}  // namespace fused
}  // namespace gnncca

using namespace gnncca;

extern "C" int gnncca_mp_fused_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* w_enc_edge,
                                   const float* b_enc_edge, int32_t edge_in_dim, const int64_t* graph_ptr, int64_t n_graphs,
                                   const int64_t* edge_index, const int64_t* edge_ptr,
                                   const int32_t* rowptr, const int32_t* src, const int32_t* dst, int64_t n_nodes,
                                   int64_t n_edges, const float* h_enc, const float* edge_attr, int32_t num_steps,
                                   int32_t first_class_step, int32_t cap_nodes, int32_t cap_edges, float* logits,
                                   float* h_dump, float* e_dump, int32_t* flags, void* stream) {
  if (!cfg || !w || n_graphs < 0 || n_nodes < 0 || n_edges < 0 || num_steps < 1 || !flags) return GNNCCA_ERR_BAD_ARG;
  if (n_graphs == 0 || n_nodes == 0) return GNNCCA_OK;
  const bool raw = edge_index != nullptr;
  if (!graph_ptr || !h_enc || !w->w_edge || !w->b_edge || !w->w_node || !w->b_node || !w_enc_edge || !b_enc_edge)
    return GNNCCA_ERR_BAD_ARG;
  if (raw ? !edge_ptr : !rowptr) return GNNCCA_ERR_BAD_ARG;
  if (n_edges > 0 && ((!raw && (!src || !dst)) || !edge_attr)) return GNNCCA_ERR_BAD_ARG;
  if (cfg->f_node != mpf::FN || cfg->f_edge != mpf::FE) return GNNCCA_ERR_UNSUPPORTED;
  if (edge_in_dim < 1 || edge_in_dim > mpf::MAX_FA) return GNNCCA_ERR_UNSUPPORTED;
  if (cfg->agg < GNNCCA_AGG_SUM || cfg->agg > GNNCCA_AGG_MAX) return GNNCCA_ERR_BAD_ARG;
  if (cap_nodes < 1 || cap_nodes > 65535 || cap_edges < 1) return GNNCCA_ERR_BAD_ARG;
  const bool has_cls = first_class_step <= num_steps;
  if (has_cls && (cfg->cls_hidden != mpf::HC || !w->w_c1 || !w->b_c1 || !w->w_c2 || !w->b_c2 || !logits || first_class_step < 1))
    return cfg->cls_hidden != mpf::HC ? GNNCCA_ERR_UNSUPPORTED : GNNCCA_ERR_BAD_ARG;
  if ((((uintptr_t)h_enc | (uintptr_t)edge_attr) & 15) != 0) return GNNCCA_ERR_BAD_ARG;

  fused::Params p{};
  mpf::Net& nt = p.net;
  nt.w = *w; nt.w_enc = w_enc_edge; nt.b_enc = b_enc_edge; nt.fa = edge_in_dim;
  nt.L = num_steps; nt.first_cls = first_class_step; nt.has_cls = has_cls ? 1 : 0; nt.agg = cfg->agg;
  nt.reatt_n = cfg->reattach_nodes ? 1 : 0; nt.reatt_e = cfg->reattach_edges ? 1 : 0;
  nt.cap_n = cap_nodes; nt.cap_e = cap_edges;
  mpf::Batch& bt = p.batch;
  bt.graph_ptr = graph_ptr; bt.n_graphs = n_graphs; bt.n_nodes = n_nodes; bt.n_edges = n_edges;
  bt.edge_index = edge_index; bt.edge_ptr = edge_ptr; bt.rowptr = rowptr; bt.src = src; bt.dst = dst;
  bt.h_enc = h_enc; bt.edge_attr = edge_attr; bt.logits = logits; bt.h_dump = h_dump; bt.e_dump = e_dump; bt.flags = flags;

  const int K = nt.reatt_n ? 64 : 32, EIN = nt.reatt_e ? 12 : 6;
  const size_t wbytes = (size_t)mpf::weight_layout(K, EIN).total * sizeof(float);
  const size_t sbytes = (size_t)mpf::state_layout(cap_nodes, cap_edges, nt.reatt_n != 0, nt.reatt_e != 0).total * sizeof(float);
  const size_t budget = 227 * 1024 - 16;
  if (wbytes + sbytes > budget) return GNNCCA_ERR_UNSUPPORTED;   // frame class too large for one CTA: use the per-step kernels
  // One CTA per SM; as many groups as fit, then trimmed so that every group of the machine gets the same number
  // of frames (1024 frames on 148 SMs: 7 groups x 1 round instead of 8 groups of which one idles)
  const int gt = cap_edges <= 512 ? 128 : 256;
  int gmax = (int)((budget - wbytes) / sbytes);
  if (gmax > 896 / gt) gmax = 896 / gt;       // __launch_bounds__(896, 1)
  const int64_t per_sm = ceil_div(n_graphs, kNumSMs);
  const int64_t rounds = ceil_div(per_sm, gmax);
  const int groups = (int)ceil_div(per_sm, rounds);
  const int grid = (int)(n_graphs < kNumSMs ? n_graphs : kNumSMs);
  p.groups = groups; p.group_threads = gt;
  const size_t smem = wbytes + sbytes * groups + 16;
  cudaStream_t st = (cudaStream_t)stream;
  void (*kern)(const fused::Params) = nullptr;
  if (nt.reatt_e) kern = nt.agg == GNNCCA_AGG_MAX ? fused::mp_fused_fwd_kernel<true, GNNCCA_AGG_MAX>
                       : nt.agg == GNNCCA_AGG_MEAN ? fused::mp_fused_fwd_kernel<true, GNNCCA_AGG_MEAN> : fused::mp_fused_fwd_kernel<true, GNNCCA_AGG_SUM>;
  else kern = nt.agg == GNNCCA_AGG_MAX ? fused::mp_fused_fwd_kernel<false, GNNCCA_AGG_MAX>
            : nt.agg == GNNCCA_AGG_MEAN ? fused::mp_fused_fwd_kernel<false, GNNCCA_AGG_MEAN> : fused::mp_fused_fwd_kernel<false, GNNCCA_AGG_SUM>;
  if (smem > 48 * 1024) {
    int rc = check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    if (rc) return rc;
  }
  if (tl_launch_as_dependent && raw) {
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3(grid); lc.blockDim = dim3(groups * gt); lc.dynamicSmemBytes = smem; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = 1;
    int rc = check_cuda(cudaLaunchKernelEx(&lc, kern, p));
    if (rc) return rc;
  } else {
    kern<<<grid, groups * gt, smem, st>>>(p);
  }
  return check_launch();
}
