// Development aid: runs the single-kernel forward (mpn_mega.cu) on a synthetic cfg-2 shaped batch and prints, per CTA
// event, when it happened relative to the first CTA's start (globaltimer).  Build: see tools/micro/README in DESIGN.md.
#define GNNCCA_MEGA_TRACE 1
#include "../../gnn-cca_b200/csrc/encoder_tc.cu"
#include "../../gnn-cca_b200/csrc/mpn_mega.cu"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
namespace gnncca { void set_cuda_error(cudaError_t e) { printf("cuda error: %s\n", cudaGetErrorString(e)); } void count_launches(int) {} }

int main(int argc, char** argv) {
  const int B = argc > 1 ? atoi(argv[1]) : 1024, C = 4, D = 6, n = C * D, e = n * n - C * D * D;
  const int64_t N = (int64_t)B * n, E = (int64_t)B * e;
  const int K = 2048, H = 128, NSET = 3;
  std::vector<int64_t> gp(B + 1), ep(B + 1), ei(2 * E);
  for (int g = 0; g <= B; ++g) { gp[g] = (int64_t)g * n; ep[g] = (int64_t)g * e; }
  for (int g = 0; g < B; ++g) {
    int64_t k = ep[g];
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) if (i / D != j / D) { ei[k] = gp[g] + i; ei[E + k] = gp[g] + j; ++k; }
  }
  auto dev = [](size_t bytes, const void* src) { void* d; cudaMalloc(&d, bytes); if (src) cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice); else cudaMemset(d, 0, bytes); return d; };
  std::vector<float*> xs(NSET);
  for (auto& x : xs) x = (float*)dev((size_t)N * K * 4, nullptr);
  int64_t* d_gp = (int64_t*)dev((B + 1) * 8, gp.data()); int64_t* d_ep = (int64_t*)dev((B + 1) * 8, ep.data());
  int64_t* d_ei = (int64_t*)dev(2 * E * 8, ei.data());
  float* ea = (float*)dev(E * 16, nullptr);
  float* W1 = (float*)dev(H * K * 4, nullptr); float* b1 = (float*)dev(H * 4, nullptr);
  float* W2 = (float*)dev(32 * H * 4, nullptr); float* b2 = (float*)dev(32 * 4, nullptr);
  float* small = (float*)dev(4096 * 4, nullptr);
  float* h = (float*)dev((size_t)N * 32 * 4, nullptr); float* logits = (float*)dev((size_t)3 * E * 4, nullptr);
  int32_t* flags = (int32_t*)dev(16, nullptr);
  unsigned long long* trace = (unsigned long long*)dev(148 * 16 * 8, nullptr);
  gnncca_mp_cfg cfg{}; cfg.f_node = 32; cfg.f_edge = 6; cfg.cls_hidden = 4; cfg.agg = GNNCCA_AGG_SUM;
  gnncca_mp_weights w{}; w.w_edge = small; w.b_edge = small; w.w_node = small; w.b_node = small; w.w_c1 = small; w.b_c1 = small; w.w_c2 = small; w.b_c2 = small;
  auto run = [&](int it) {
    return gnncca::mpn_mega_fwd(&cfg, &w, W1, b1, W2, b2, K, H, small, small, 4, xs[it % NSET], d_ei, ea, d_gp, d_ep, N, E, B, 4, 2, 24, 448, h, logits, flags, 0);
  };
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int cfgs[][2] = {{0, 0}, {3, 0}, {2, 0}, {1, 0}, {0, 3}};
  for (auto& c : cfgs) {
    g_mega_force_groups = c[0]; g_mega_force_stages = c[1]; g_mega_trace = nullptr;
    int rc = 0;
    for (int it = 0; it < 3; ++it) rc |= run(it);
    cudaEventRecord(e0);
    for (int it = 0; it < 12; ++it) rc |= run(it);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("force groups %d stages %d: rc %d  %.1f us per launch (%s)\n", c[0], c[1], rc, ms / 12 * 1e3, cudaGetErrorString(cudaGetLastError()));
    g_mega_trace = trace; cudaMemset(trace, 0, 148 * 16 * 8);
    run(0); cudaDeviceSynchronize();
    std::vector<unsigned long long> t(148 * 16);
    cudaMemcpy(t.data(), trace, 148 * 16 * 8, cudaMemcpyDeviceToHost);
    unsigned long long t0 = ~0ull;
    for (int c2 = 0; c2 < 148; ++c2) if (t[c2 * 16]) t0 = std::min(t0, t[c2 * 16]);
    const char* names[10] = {"cta start", "setup done", "first stage landed", "last stage landed", "last chunk encoded", "group 0 done", "cta end",
                             "group 0 starts chunk 0", "group 0 ends chunk 0", "chunk 0 encoded"};
    const int order[10] = {0, 1, 2, 9, 7, 8, 3, 4, 5, 6};
    for (int oi = 0; oi < 10; ++oi) {
      const int ev = order[oi];
      std::vector<double> v;
      for (int c2 = 0; c2 < 148; ++c2) if (t[c2 * 16 + ev]) v.push_back((double)(t[c2 * 16 + ev] - t0) * 1e-3);
      if (v.empty()) continue;
      std::sort(v.begin(), v.end());
      printf("  %-24s min %6.2f  med %6.2f  max %6.2f us\n", names[ev], v.front(), v[v.size() / 2], v.back());
    }
  }
  int32_t fl[4]; cudaMemcpy(fl, flags, 16, cudaMemcpyDeviceToHost);
  printf("flags %d %d %d %d\n", fl[0], fl[1], fl[2], fl[3]);
  return 0;
}
