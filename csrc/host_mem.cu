// Pinned host staging memory for input batches.
//
// The reference never leaves the CPU (its batches are DataLoader tensors, ref.py:146,155); here every
// step's inputs cross PCIe, and for the wide MLP that copy (201 MB per step per GPU) is as long as the
// step itself, so the end-to-end rate IS the H2D rate.  Measured on the B200 boxes of this pool: pinned
// buffers from cudaHostAlloc / torch's pin_memory() copy at 45-55 GB/s depending on the allocation
// (4 KB pages; one buffer in four came out ~20 % slow), anonymous memory advised to 2 MB huge pages and
// then registered with cudaHostRegister copies at 55.2-55.5 GB/s every time.  This is that allocator.
#include <errno.h>
#include <string.h>
#include <sys/mman.h>

#include "common.cuh"

namespace {
constexpr size_t kHuge = size_t(1) << 21;
}

using namespace b200dp;

// *ptr: 2 MB-aligned host pointer, page-locked and mapped for DMA; *mapped_bytes: what to pass to
// b200dp_host_free.  Pages are touched here so the first copy does not pay for faults.
extern "C" int b200dp_host_alloc(long long bytes, void** ptr, long long* mapped_bytes) {
  B200DP_REQUIRE(bytes > 0 && ptr && mapped_bytes, "host_alloc: bad arguments");
  const size_t size = (size_t(bytes) + kHuge - 1) & ~(kHuge - 1);
  const size_t span = size + kHuge;                         // room to align the start to 2 MB
  void* raw = mmap(nullptr, span, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
  B200DP_REQUIRE(raw != MAP_FAILED, "host_alloc: mmap of %zu bytes failed: %s", span, strerror(errno));
  const uintptr_t al = (reinterpret_cast<uintptr_t>(raw) + kHuge - 1) & ~uintptr_t(kHuge - 1);
  // give back the unaligned head and tail so the mapping is exactly [al, al + size)
  if (al > reinterpret_cast<uintptr_t>(raw)) munmap(raw, al - reinterpret_cast<uintptr_t>(raw));
  const uintptr_t end = reinterpret_cast<uintptr_t>(raw) + span;
  if (end > al + size) munmap(reinterpret_cast<void*>(al + size), end - (al + size));
  void* p = reinterpret_cast<void*>(al);
  madvise(p, size, MADV_HUGEPAGE);                          // best effort: falls back to 4 KB pages
  memset(p, 0, size);
  cudaError_t e = cudaHostRegister(p, size, cudaHostRegisterDefault);
  if (e != cudaSuccess) {
    munmap(p, size);
    return fail("host_alloc: cudaHostRegister failed: %s", cudaGetErrorString(e));
  }
  *ptr = p;
  *mapped_bytes = (long long)size;
  return B200DP_OK;
}

extern "C" int b200dp_host_free(void* ptr, long long mapped_bytes) {
  if (!ptr) return B200DP_OK;
  cudaError_t e = cudaHostUnregister(ptr);
  munmap(ptr, size_t(mapped_bytes));
  B200DP_REQUIRE(e == cudaSuccess, "host_free: cudaHostUnregister failed: %s", cudaGetErrorString(e));
  return B200DP_OK;
}
