// Symmetric (peer-mapped) memory and NVLS multicast bootstrap - the transport that stands where
// the reference has MPI.COMM_WORLD (ref.py:61-63).  Host code only; the kernels that use these
// mappings are in allreduce_sgd.cu.
//
//   ipc flavour : cudaMalloc + cudaIpcGetMemHandle / cudaIpcOpenMemHandle.
//   vmm flavour : cuMemCreate (POSIX-fd shareable) + cuMemMap; a peer duplicates the owner's fd
//                 with pidfd_open + pidfd_getfd (same user, same node) and imports it.  The same
//                 allocations can be bound into a multicast object (cuMulticast*), which gives the
//                 address multimem.ld_reduce / multimem.st operate on.
// Driver entry points are resolved at run time (cudaGetDriverEntryPoint) so the library has no
// link-time dependency on libcuda and loads on a box without a GPU.
#include <sys/syscall.h>
#include <unistd.h>

#include <map>
#include <mutex>

#include "common.cuh"

namespace b200dp {
namespace {

template <typename Fn>
Fn driver_fn(const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<Fn>(sym);
}

#define DRV(name, ...)                                                         \
  static auto p_##name = driver_fn<CUresult (*)(__VA_ARGS__)>(#name);          \
  if (!p_##name) return fail("driver entry point %s not available", #name)

#define CU_OK(call)                                                                        \
  do {                                                                                     \
    CUresult _r = (call);                                                                  \
    if (_r != CUDA_SUCCESS) return fail("%s failed with CUresult %d", #call, int(_r));      \
  } while (0)

struct VmmAlloc {
  CUmemGenericAllocationHandle handle;
  size_t bytes;
};
std::mutex g_mu;
std::map<void*, VmmAlloc> g_vmm;     // va -> allocation (own or imported)

int current_device(CUdevice* dev, int* ordinal) {
  int d = 0;
  B200DP_CUDA_OK(cudaGetDevice(&d));
  B200DP_CUDA_OK(cudaFree(0));       // make sure the primary context exists
  DRV(cuDeviceGet, CUdevice*, int);
  CU_OK(p_cuDeviceGet(dev, d));
  *ordinal = d;
  return B200DP_OK;
}

int dup_peer_fd(int pid, int fd) {
  if (pid == getpid()) return dup(fd);
  const int pidfd = int(syscall(SYS_pidfd_open, pid, 0));
  if (pidfd < 0) return -1;
  const int out = int(syscall(SYS_pidfd_getfd, pidfd, fd, 0));
  close(pidfd);
  return out;
}

int map_handle(CUmemGenericAllocationHandle h, size_t bytes, int ordinal, void** ptr) {
  DRV(cuMemAddressReserve, CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  DRV(cuMemMap, CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  DRV(cuMemSetAccess, CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUdeviceptr va = 0;
  CU_OK(p_cuMemAddressReserve(&va, bytes, 0, 0, 0));
  CU_OK(p_cuMemMap(va, bytes, 0, h, 0));
  CUmemAccessDesc acc = {};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ordinal;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CU_OK(p_cuMemSetAccess(va, bytes, &acc, 1));
  *ptr = reinterpret_cast<void*>(va);
  return B200DP_OK;
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

// ------------------------------------------------------------------------------------------ ipc
extern "C" int b200dp_ipc_alloc(long long bytes, void** ptr, unsigned char* handle64) {
  B200DP_REQUIRE(bytes > 0 && ptr && handle64, "ipc_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  B200DP_CUDA_OK(cudaMalloc(ptr, size_t(bytes)));
  B200DP_CUDA_OK(cudaMemset(*ptr, 0, size_t(bytes)));
  cudaIpcMemHandle_t h;
  B200DP_CUDA_OK(cudaIpcGetMemHandle(&h, *ptr));
  memcpy(handle64, &h, 64);
  return B200DP_OK;
}

extern "C" int b200dp_ipc_open(const unsigned char* handle64, void** ptr) {
  B200DP_REQUIRE(handle64 && ptr, "ipc_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200DP_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200DP_OK;
}

extern "C" int b200dp_ipc_close(void* ptr) {
  B200DP_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return B200DP_OK;
}

extern "C" int b200dp_ipc_free(void* ptr) {
  B200DP_CUDA_OK(cudaFree(ptr));
  return B200DP_OK;
}

// ------------------------------------------------------------------------------------------ vmm
extern "C" int b200dp_vmm_supported(int* fd_export, int* multicast) {
  CUdevice dev;
  int ord;
  if (int rc = current_device(&dev, &ord)) return rc;
  DRV(cuDeviceGetAttribute, int*, CUdevice_attribute, CUdevice);
  int a = 0, b = 0;
  CU_OK(p_cuDeviceGetAttribute(&a, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev));
  CU_OK(p_cuDeviceGetAttribute(&b, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  if (fd_export) *fd_export = a;
  if (multicast) *multicast = b;
  return B200DP_OK;
}

static int vmm_props(CUmemAllocationProp* prop, int ordinal) {
  memset(prop, 0, sizeof(*prop));
  prop->type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop->location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop->location.id = ordinal;
  prop->requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return 0;
}

extern "C" int b200dp_vmm_alloc(long long bytes, void** ptr, int* fd, long long* padded_bytes) {
  B200DP_REQUIRE(bytes > 0 && ptr && fd && padded_bytes, "vmm_alloc: bad arguments");
  CUdevice dev;
  int ord;
  if (int rc = current_device(&dev, &ord)) return rc;
  DRV(cuMemGetAllocationGranularity, size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  DRV(cuMemCreate, CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  DRV(cuMemExportToShareableHandle, void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
      unsigned long long);
  DRV(cuMulticastGetGranularity, size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUmemAllocationProp prop;
  vmm_props(&prop, ord);
  size_t gran = 0;
  CU_OK(p_cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  // keep the size bindable to a multicast object as well (its granularity can be larger)
  CUmulticastObjectProp mp = {};
  mp.numDevices = 2;
  mp.size = size_t(bytes);
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t mgran = 0;
  if (p_cuMulticastGetGranularity(&mgran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mgran > gran)
    gran = mgran;
  const size_t padded = (size_t(bytes) + gran - 1) / gran * gran;
  CUmemGenericAllocationHandle h;
  CU_OK(p_cuMemCreate(&h, padded, &prop, 0));
  int out_fd = -1;
  CU_OK(p_cuMemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  if (int rc = map_handle(h, padded, ord, ptr)) return rc;
  B200DP_CUDA_OK(cudaMemset(*ptr, 0, padded));
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_vmm[*ptr] = VmmAlloc{h, padded};
  }
  *fd = out_fd;
  *padded_bytes = (long long)padded;
  return B200DP_OK;
}

extern "C" int b200dp_vmm_import(int owner_pid, int owner_fd, long long padded_bytes, void** ptr) {
  B200DP_REQUIRE(ptr && padded_bytes > 0, "vmm_import: bad arguments");
  CUdevice dev;
  int ord;
  if (int rc = current_device(&dev, &ord)) return rc;
  DRV(cuMemImportFromShareableHandle, CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  const int fd = dup_peer_fd(owner_pid, owner_fd);
  B200DP_REQUIRE(fd >= 0, "vmm_import: pidfd_getfd(pid %d, fd %d) failed (errno %d)", owner_pid, owner_fd, errno);
  CUmemGenericAllocationHandle h;
  CUresult r = p_cuMemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);
  if (r != CUDA_SUCCESS) return fail("cuMemImportFromShareableHandle failed with CUresult %d", int(r));
  if (int rc = map_handle(h, size_t(padded_bytes), ord, ptr)) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  g_vmm[*ptr] = VmmAlloc{h, size_t(padded_bytes)};
  return B200DP_OK;
}

extern "C" int b200dp_vmm_free(void* ptr, long long padded_bytes) {
  DRV(cuMemUnmap, CUdeviceptr, size_t);
  DRV(cuMemRelease, CUmemGenericAllocationHandle);
  DRV(cuMemAddressFree, CUdeviceptr, size_t);
  VmmAlloc a;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_vmm.find(ptr);
    B200DP_REQUIRE(it != g_vmm.end(), "vmm_free: unknown pointer");
    a = it->second;
    g_vmm.erase(it);
  }
  (void)padded_bytes;
  CU_OK(p_cuMemUnmap(reinterpret_cast<CUdeviceptr>(ptr), a.bytes));
  CU_OK(p_cuMemRelease(a.handle));
  CU_OK(p_cuMemAddressFree(reinterpret_cast<CUdeviceptr>(ptr), a.bytes));
  return B200DP_OK;
}

// ------------------------------------------------------------------------------------ multicast
extern "C" int b200dp_mc_create(int world, long long padded_bytes, int* fd, unsigned long long* mc_handle) {
  B200DP_REQUIRE(world >= 2 && fd && mc_handle, "mc_create: bad arguments");
  DRV(cuMulticastCreate, CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  DRV(cuMemExportToShareableHandle, void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
      unsigned long long);
  CUmulticastObjectProp mp = {};
  mp.numDevices = unsigned(world);
  mp.size = size_t(padded_bytes);
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  CU_OK(p_cuMulticastCreate(&h, &mp));
  int out_fd = -1;
  CU_OK(p_cuMemExportToShareableHandle(&out_fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = out_fd;
  *mc_handle = (unsigned long long)h;
  return B200DP_OK;
}

extern "C" int b200dp_mc_import(int owner_pid, int owner_fd, unsigned long long* mc_handle) {
  B200DP_REQUIRE(mc_handle, "mc_import: bad arguments");
  DRV(cuMemImportFromShareableHandle, CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  const int fd = dup_peer_fd(owner_pid, owner_fd);
  B200DP_REQUIRE(fd >= 0, "mc_import: pidfd_getfd(pid %d, fd %d) failed (errno %d)", owner_pid, owner_fd, errno);
  CUmemGenericAllocationHandle h;
  CUresult r = p_cuMemImportFromShareableHandle(&h, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  close(fd);
  if (r != CUDA_SUCCESS) return fail("cuMemImportFromShareableHandle(multicast) failed with CUresult %d", int(r));
  *mc_handle = (unsigned long long)h;
  return B200DP_OK;
}

extern "C" int b200dp_mc_add_device(unsigned long long mc_handle) {
  CUdevice dev;
  int ord;
  if (int rc = current_device(&dev, &ord)) return rc;
  DRV(cuMulticastAddDevice, CUmemGenericAllocationHandle, CUdevice);
  CU_OK(p_cuMulticastAddDevice((CUmemGenericAllocationHandle)mc_handle, dev));
  return B200DP_OK;
}

// Binds this rank's vmm allocation into the multicast object (every rank must have called
// mc_add_device first) and maps the multicast object into this process.
extern "C" int b200dp_mc_bind(unsigned long long mc_handle, void* local_ptr, long long padded_bytes, void** mc_ptr) {
  B200DP_REQUIRE(local_ptr && mc_ptr, "mc_bind: bad arguments");
  CUdevice dev;
  int ord;
  if (int rc = current_device(&dev, &ord)) return rc;
  DRV(cuMulticastBindMem, CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
      unsigned long long);
  VmmAlloc a;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_vmm.find(local_ptr);
    B200DP_REQUIRE(it != g_vmm.end(), "mc_bind: local_ptr is not a vmm allocation");
    a = it->second;
  }
  B200DP_REQUIRE(size_t(padded_bytes) == a.bytes, "mc_bind: size mismatch");
  CU_OK(p_cuMulticastBindMem((CUmemGenericAllocationHandle)mc_handle, 0, a.handle, 0, a.bytes, 0));
  return map_handle((CUmemGenericAllocationHandle)mc_handle, a.bytes, ord, mc_ptr);
}

extern "C" int b200dp_mc_release(unsigned long long mc_handle, void* mc_ptr, long long padded_bytes) {
  DRV(cuMemUnmap, CUdeviceptr, size_t);
  DRV(cuMemRelease, CUmemGenericAllocationHandle);
  DRV(cuMemAddressFree, CUdeviceptr, size_t);
  if (mc_ptr) {
    CU_OK(p_cuMemUnmap(reinterpret_cast<CUdeviceptr>(mc_ptr), size_t(padded_bytes)));
    CU_OK(p_cuMemAddressFree(reinterpret_cast<CUdeviceptr>(mc_ptr), size_t(padded_bytes)));
  }
  CU_OK(p_cuMemRelease((CUmemGenericAllocationHandle)mc_handle));
  return B200DP_OK;
}
