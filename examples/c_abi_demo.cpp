// Using libb200dp from plain C++ (no Python, no PyTorch): one training step of the reference's
// 2-3-1 MLP (ref.py:41-45) on 16 rows with the whole-network kernel and the local SGD kernel.
//
//   g++ -std=c++17 examples/c_abi_demo.cpp -Iinclude -I/usr/local/cuda/include \
//       -Lneural-networks-parallel-training-with-mpi_b200/lib -lb200dp -L/usr/local/cuda/lib64 -lcudart \
//       -Wl,-rpath,$PWD/neural-networks-parallel-training-with-mpi_b200/lib -o /tmp/c_abi_demo
#include <cuda_runtime.h>

#include <cstdio>
#include <vector>

#include "b200dp.h"

#define CHECK(x)                                                              \
  do {                                                                        \
    if ((x) != 0) {                                                           \
      std::fprintf(stderr, "%s failed: %s\n", #x, b200dp_last_error());       \
      return 1;                                                               \
    }                                                                         \
  } while (0)

int main() {
  std::printf("libb200dp version %d\n", b200dp_version());
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    std::printf("no CUDA device: ABI link check only\n");
    return 0;
  }
  // flat layout: every segment padded to 128 floats (see models/mlp.py)
  b200dp_mlp_desc m = {};
  m.num_layers = 2;
  m.sizes[0] = 2; m.sizes[1] = 3; m.sizes[2] = 1;
  m.w_off[0] = 0; m.b_off[0] = 128; m.w_off[1] = 256; m.b_off[1] = 384;
  m.flat_numel = 512;
  const int rows = 16;
  std::vector<float> hp(512, 0.f), hx(rows * 2), hy(rows);
  for (int i = 0; i < 6; ++i) hp[i] = 0.1f * (i + 1);
  for (int i = 0; i < 3; ++i) hp[256 + i] = 0.2f;
  for (int i = 0; i < rows; ++i) { hx[2 * i] = 0.1f * i; hx[2 * i + 1] = 1.f - 0.05f * i; hy[i] = 0.3f * i; }
  float *p, *g, *mom, *x, *y, *loss;
  cudaMalloc(&p, 512 * 4); cudaMalloc(&g, 512 * 4); cudaMalloc(&mom, 512 * 4);
  cudaMalloc(&x, rows * 2 * 4); cudaMalloc(&y, rows * 4); cudaMalloc(&loss, 4);
  cudaMemcpy(p, hp.data(), 512 * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(x, hx.data(), rows * 2 * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(y, hy.data(), rows * 4, cudaMemcpyHostToDevice);
  cudaMemset(mom, 0, 512 * 4);
  for (int step = 0; step < 3; ++step) {
    // the allocation-free entry point: inputs and outputs only
    CHECK(b200dp_tiny_mlp_fwd_bwd_noalloc(&m, p, x, y, rows, g, loss, nullptr));
    CHECK(b200dp_sgd_momentum(p, g, 0, mom, nullptr, 512, 0.001f, 0.9f, 1.0f, 0, nullptr));
    float h = 0.f;
    cudaMemcpy(&h, loss, 4, cudaMemcpyDeviceToHost);
    std::printf("step %d loss %.6f\n", step, h);
  }
  return 0;
}
