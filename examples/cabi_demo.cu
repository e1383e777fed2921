// Torch-free host calling libb200sft.so through its C ABI: plain cudaMalloc buffers, a cudaStream_t, and
// a CPU check of the result.  Build:  nvcc -arch=sm_100a -o cabi_demo examples/cabi_demo.cu \
//     -Iinclude -Lllm_fine_tune_distributed_b200 -lb200sft -Xlinker -rpath=llm_fine_tune_distributed_b200
// What it shows: the boundary needs nothing but device pointers, sizes and a stream.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "b200sft.h"

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e = (x);                                                       \
    if (e != cudaSuccess) {                                                    \
      fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e));                  \
      return 2;                                                                \
    }                                                                          \
  } while (0)

static float frand(unsigned& s) {  // deterministic uniform in [-1, 1)
  s = s * 1664525u + 1013904223u;
  return ((s >> 8) & 0xFFFF) / 32768.0f - 1.0f;
}

int main() {
  const int N = 300, H = 64, V = 1000, IGNORE = -100;
  std::vector<__nv_bfloat16> x(N * H), w(V * H);
  std::vector<int> labels(N);
  unsigned seed = 7;
  for (auto& v : x) v = __float2bfloat16(frand(seed));
  for (auto& v : w) v = __float2bfloat16(0.3f * frand(seed));
  int n_valid_ref = 0;
  for (int i = 0; i < N; ++i) {
    seed = seed * 1664525u + 1013904223u;
    labels[i] = (i % 5 == 0) ? IGNORE : static_cast<int>((seed >> 8) % V);
    n_valid_ref += labels[i] != IGNORE;
  }
  // CPU reference: mean over valid rows of logsumexp(z) - z[label], z = x . w^T in double
  double ref = 0.0;
  for (int i = 0; i < N; ++i) {
    if (labels[i] == IGNORE) continue;
    double m = -1e300, tgt = 0.0;
    std::vector<double> z(V);
    for (int v = 0; v < V; ++v) {
      double acc = 0.0;
      for (int k = 0; k < H; ++k) acc += double(__bfloat162float(x[i * H + k])) * double(__bfloat162float(w[v * H + k]));
      z[v] = acc;
      if (acc > m) m = acc;
      if (v == labels[i]) tgt = acc;
    }
    double s = 0.0;
    for (int v = 0; v < V; ++v) s += exp(z[v] - m);
    ref += m + log(s) - tgt;
  }
  ref /= n_valid_ref;

  void *dx, *dw, *ws;
  int* dl;
  float *dloss, *dlse;
  long long* dnv;
  cudaStream_t stream;
  CK(cudaStreamCreate(&stream));
  CK(cudaMalloc(&dx, x.size() * 2));
  CK(cudaMalloc(&dw, w.size() * 2));
  CK(cudaMalloc(&dl, N * sizeof(int)));
  CK(cudaMalloc(&dloss, sizeof(float)));
  CK(cudaMalloc(&dlse, N * sizeof(float)));
  CK(cudaMalloc(&dnv, sizeof(long long)));
  const size_t ws_bytes = b200sft_flce_fwd_workspace_bytes(N, H, V);
  CK(cudaMalloc(&ws, ws_bytes));
  CK(cudaMemcpyAsync(dx, x.data(), x.size() * 2, cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(dw, w.data(), w.size() * 2, cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(dl, labels.data(), N * sizeof(int), cudaMemcpyHostToDevice, stream));

  int rc = b200sft_flce_fwd(dx, dw, dl, N, H, V, IGNORE, 1.0f, B200SFT_DENOM_VALID, nullptr, 0, nullptr, dloss, dlse,
                            reinterpret_cast<int64_t*>(dnv), ws, ws_bytes, stream);
  if (rc != 0) {
    fprintf(stderr, "b200sft_flce_fwd failed (%d): %s\n", rc, b200sft_last_error());
    return 3;
  }
  float loss = 0.f;
  long long nv = 0;
  CK(cudaMemcpyAsync(&loss, dloss, sizeof(float), cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(&nv, dnv, sizeof(long long), cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  const double rel = fabs(loss - ref) / fabs(ref);
  printf("loss %.6f  cpu reference %.6f  rel err %.2e  valid rows %lld (expected %d)\n", loss, ref, rel, nv, n_valid_ref);
  return (rel <= 1e-3 && nv == n_valid_ref) ? 0 : 1;
}
