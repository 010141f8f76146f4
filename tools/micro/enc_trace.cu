// Development aid: runs the node encoder kernel stand-alone and prints, per CTA event, when it happened relative to the
// first CTA's start (globaltimer, ns).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17
//   --expt-relaxed-constexpr -o enc_trace enc_trace.cu
#define GNNCCA_ENC_TRACE 1
#include "../../gnn-cca_b200/csrc/encoder_tc.cu"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
namespace gnncca { void set_cuda_error(cudaError_t) {} void count_launches(int) {} }

int main(int argc, char** argv) {
  const int64_t N = argc > 1 ? atoll(argv[1]) : 24576;
  const int K = 2048, H = 128, NSET = 3;
  std::vector<float*> xs(NSET);
  for (auto& x : xs) { cudaMalloc(&x, (size_t)N * K * 4); cudaMemset(x, 0, (size_t)N * K * 4); }
  float *W1, *b1, *W2, *b2, *h;
  cudaMalloc(&W1, H * K * 4); cudaMemset(W1, 0, H * K * 4);
  cudaMalloc(&b1, H * 4); cudaMemset(b1, 0, H * 4);
  cudaMalloc(&W2, 32 * H * 4); cudaMemset(W2, 0, 32 * H * 4);
  cudaMalloc(&b2, 32 * 4); cudaMemset(b2, 0, 32 * 4);
  cudaMalloc(&h, (size_t)N * 32 * 4);
  unsigned long long* trace; cudaMalloc(&trace, 148 * 12 * 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int force : {0}) {
    g_enc_force_stages = force;
    g_enc_trace = nullptr;
    for (int it = 0; it < 3; ++it) gnncca_node_encoder_tf32(xs[it % NSET], W1, b1, W2, b2, h, N, K, H, 32, 0);
    cudaEventRecord(e0);
    for (int it = 0; it < 12; ++it) gnncca_node_encoder_tf32(xs[it % NSET], W1, b1, W2, b2, h, N, K, H, 32, 0);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("stages cap %d: %.1f us per launch (%s)\n", force, ms / 12 * 1e3, cudaGetErrorString(cudaGetLastError()));
    g_enc_trace = trace;
    cudaMemset(trace, 0, 148 * 12 * 8);
    gnncca_node_encoder_tf32(xs[0], W1, b1, W2, b2, h, N, K, H, 32, 0);
    cudaDeviceSynchronize();
    std::vector<unsigned long long> t(148 * 12);
    cudaMemcpy(t.data(), trace, 148 * 12 * 8, cudaMemcpyDeviceToHost);
    unsigned long long t0 = ~0ull;
    for (int c = 0; c < 148; ++c) if (t[c * 12]) t0 = std::min(t0, t[c * 12]);
    const char* names[11] = {"cta start", "setup done", "first stage landed", "last stage landed", "accumulator ready", "epilogue done", "cta end", "tmem allocated", "W2 staged", "barriers ready", "first ring issued"};
    for (int e = 0; e < 11; ++e) {
      std::vector<double> v;
      for (int c = 0; c < 148; ++c) if (t[c * 12 + e]) v.push_back((double)(t[c * 12 + e] - t0) * 1e-3);
      if (v.empty()) continue;
      std::sort(v.begin(), v.end());
      printf("  %-20s min %6.2f  med %6.2f  max %6.2f us  (%zu ctas)\n", names[e], v.front(), v[v.size() / 2], v.back(), v.size());
    }
  }
  return 0;
}
