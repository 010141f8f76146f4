// MOTMPNet.forward (reference models/mpn.py:250-299) for batches of small frame graphs as one C-ABI call: the host
// side of the hot path (two launches: the node encoder and, as its programmatic dependent, the fused all-steps kernel; the entry point the drop-in module binds).
// A single persistent kernel doing both (round 1's mpn_mega.cu) measured 101 us against 85 us for the two launches and was removed in round 2..
#include "common.cuh"

extern "C" int gnncca_mpn_forward_frames(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w,
                                         const float* enc_W1, const float* enc_b1, const float* enc_W2, const float* enc_b2,
                                         int32_t in_dim, int32_t hidden,
                                         const float* w_enc_edge, const float* b_enc_edge, int32_t edge_in_dim,
                                         const float* x, const int64_t* edge_index, const float* edge_attr,
                                         const int64_t* graph_ptr, const int64_t* edge_ptr,
                                         int64_t n_nodes, int64_t n_edges, int64_t n_graphs,
                                         int32_t num_steps, int32_t first_class_step, int32_t cap_nodes, int32_t cap_edges,
                                         float* h_enc, float* logits, int32_t* flags, void* enc_scratch, int64_t enc_scratch_bytes,
                                         void* stream) {
  if (!cfg || !w || !x || !h_enc || !edge_index || !edge_ptr) return GNNCCA_ERR_BAD_ARG;
  if (n_nodes == 0 || n_graphs == 0) return GNNCCA_OK;
  // small batches: the encoder's split-K form when the caller brought its scratch (gnncca_node_encoder_scratch_bytes)
  int rc = gnncca_node_encoder_ws(x, enc_W1, enc_b1, enc_W2, enc_b2, nullptr, h_enc, n_nodes, in_dim, hidden, cfg->f_node, enc_scratch,
                                  enc_scratch_bytes, stream);
  if (rc) return rc;
  gnncca::tl_launch_as_dependent = true;
  rc = gnncca_mp_fused_fwd(cfg, w, w_enc_edge, b_enc_edge, edge_in_dim, graph_ptr, n_graphs, edge_index, edge_ptr, nullptr, nullptr,
                           nullptr, n_nodes, n_edges, h_enc, edge_attr, num_steps, first_class_step, cap_nodes, cap_edges, logits,
                           nullptr, nullptr, flags, stream);
  gnncca::tl_launch_as_dependent = false;
  return rc;
}
