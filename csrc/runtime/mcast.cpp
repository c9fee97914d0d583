// NVLS multicast images: one buffer per GPU plus one multicast address that, when stored to,
// writes all of them through the NVSwitch (BASELINE.json north star: "Gather ... through NVLS
// multicast where the host exposes it"). The reference's gather ends on rank 0 only
// (kernel.cu:223-225); this is the variant where every GPU ends up with the whole result, at the
// link cost of one copy per sender instead of N-1.
//
// Driver API (virtual memory management + multicast objects), reached through
// cudaGetDriverEntryPoint like the tensor-map encoder, so the library still links against the
// runtime only. One process, all devices: no handle export.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>

#include <cstring>
#include <vector>

#include "../core/error.hpp"
#include "b200img.h"

namespace {

struct Drv {
  PFN_cuDeviceGet_v2000 DeviceGet = nullptr;
  PFN_cuDeviceGetAttribute_v2000 DeviceGetAttribute = nullptr;
  PFN_cuMulticastCreate_v12010 MulticastCreate = nullptr;
  PFN_cuMulticastAddDevice_v12010 MulticastAddDevice = nullptr;
  PFN_cuMulticastBindMem_v12010 MulticastBindMem = nullptr;
  PFN_cuMulticastGetGranularity_v12010 MulticastGetGranularity = nullptr;
  PFN_cuMulticastUnbind_v12010 MulticastUnbind = nullptr;
  PFN_cuMemCreate_v10020 MemCreate = nullptr;
  PFN_cuMemGetAllocationGranularity_v10020 MemGetAllocationGranularity = nullptr;
  PFN_cuMemAddressReserve_v10020 MemAddressReserve = nullptr;
  PFN_cuMemMap_v10020 MemMap = nullptr;
  PFN_cuMemSetAccess_v10020 MemSetAccess = nullptr;
  PFN_cuMemUnmap_v10020 MemUnmap = nullptr;
  PFN_cuMemRelease_v10020 MemRelease = nullptr;
  PFN_cuMemAddressFree_v10020 MemAddressFree = nullptr;
  bool ok = false;
};

template <class F>
bool sym(const char* name, F* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
    cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<F>(p);
  return true;
}

const Drv& drv() {
  static Drv d = [] {
    Drv x;
    x.ok = sym("cuDeviceGet", &x.DeviceGet) && sym("cuDeviceGetAttribute", &x.DeviceGetAttribute) && sym("cuMulticastCreate", &x.MulticastCreate) &&
           sym("cuMulticastAddDevice", &x.MulticastAddDevice) && sym("cuMulticastBindMem", &x.MulticastBindMem) &&
           sym("cuMulticastGetGranularity", &x.MulticastGetGranularity) &&
           sym("cuMulticastUnbind", &x.MulticastUnbind) && sym("cuMemCreate", &x.MemCreate) &&
           sym("cuMemGetAllocationGranularity", &x.MemGetAllocationGranularity) &&
           sym("cuMemAddressReserve", &x.MemAddressReserve) && sym("cuMemMap", &x.MemMap) &&
           sym("cuMemSetAccess", &x.MemSetAccess) && sym("cuMemUnmap", &x.MemUnmap) &&
           sym("cuMemRelease", &x.MemRelease) && sym("cuMemAddressFree", &x.MemAddressFree);
    return x;
  }();
  return d;
}

}  // namespace

struct b2i_mcast {
  int n = 0;
  int devs[B2I_MAX_GPUS];
  size_t size = 0;
  bool multicast = false;  // false: one device, a plain allocation stands in
  CUmemGenericAllocationHandle mc = 0;
  CUmemGenericAllocationHandle mem[B2I_MAX_GPUS];
  CUdeviceptr uc[B2I_MAX_GPUS];
  bool bound[B2I_MAX_GPUS], mapped[B2I_MAX_GPUS];
  CUdeviceptr mc_ptr = 0;
  bool mc_mapped = false;
  void* plain = nullptr;
};

#define B2I_DRV(expr)                                                                              \
  do {                                                                                             \
    CUresult r_ = (expr);                                                                          \
    if (r_ != CUDA_SUCCESS) {                                                                      \
      b2i_mcast_destroy(m);                                                                        \
      B2I_FAIL(r_ == CUDA_ERROR_NOT_SUPPORTED ? B2I_ERR_UNSUPPORTED : B2I_ERR_CUDA, "%s: CUresult %d", \
               #expr, (int)r_);                                                                    \
    }                                                                                              \
  } while (0)

extern "C" void b2i_mcast_destroy(b2i_mcast_t* m) {
  if (!m) return;
  const Drv& D = drv();
  if (m->plain) {
    cudaSetDevice(m->devs[0]);
    cudaFree(m->plain);
  }
  if (m->multicast && D.ok) {
    if (m->mc_mapped) D.MemUnmap(m->mc_ptr, m->size);
    if (m->mc_ptr) D.MemAddressFree(m->mc_ptr, m->size);
    for (int i = 0; i < m->n; ++i) {
      CUdevice dev;
      if (m->bound[i] && D.DeviceGet(&dev, m->devs[i]) == CUDA_SUCCESS) D.MulticastUnbind(m->mc, dev, 0, m->size);
      if (m->mapped[i]) D.MemUnmap(m->uc[i], m->size);
      if (m->uc[i]) D.MemAddressFree(m->uc[i], m->size);
      if (m->mem[i]) D.MemRelease(m->mem[i]);
    }
    if (m->mc) D.MemRelease(m->mc);
  }
  delete m;
}

// `n` distinct devices, `bytes` per replica. With n == 1 the object degenerates to one plain
// allocation whose "multicast" address is the allocation itself.
extern "C" int b2i_mcast_create(int n, const int* devices, size_t bytes, b2i_mcast_t** out) {
  B2I_CHECK_ARG(out && devices && n >= 1 && n <= B2I_MAX_GPUS && bytes > 0, "bad multicast request");
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < i; ++j) B2I_CHECK_ARG(devices[i] != devices[j], "multicast needs distinct devices");
  b2i_mcast* m = new b2i_mcast();
  m->n = n;
  memset(m->mem, 0, sizeof m->mem);
  memset(m->uc, 0, sizeof m->uc);
  memset(m->bound, 0, sizeof m->bound);
  memset(m->mapped, 0, sizeof m->mapped);
  for (int i = 0; i < n; ++i) m->devs[i] = devices[i];
  if (n == 1) {
    cudaError_t e = cudaSetDevice(devices[0]);
    if (e == cudaSuccess) e = cudaMalloc(&m->plain, bytes);
    if (e != cudaSuccess) {
      delete m;
      B2I_FAIL(B2I_ERR_CUDA, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
    }
    m->size = bytes;
    *out = m;
    return B2I_OK;
  }
  for (int i = 0; i < n; ++i) {  // primary contexts exist before the driver API is used
    if (cudaSetDevice(devices[i]) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) {
      delete m;
      B2I_FAIL(B2I_ERR_CUDA, "cannot initialise GPU %d", devices[i]);
    }
  }
  const Drv& D = drv();
  if (!D.ok) {
    delete m;
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "this driver has no multicast object API");
  }
  for (int i = 0; i < n; ++i) {
    CUdevice dev;
    int sup = 0;
    if (D.DeviceGet(&dev, devices[i]) != CUDA_SUCCESS ||
        D.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS || !sup) {
      delete m;
      B2I_FAIL(B2I_ERR_UNSUPPORTED, "GPU %d does not support multicast objects (no NVSwitch / NVLS)", devices[i]);
    }
  }
  m->multicast = true;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof mp);
  mp.numDevices = (unsigned)n;
  mp.size = bytes;
  mp.handleTypes = 0;
  size_t gran = 0;
  B2I_DRV(D.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  // the physical allocations must satisfy their own granularity as well
  for (int i = 0; i < n; ++i) {
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof ap);
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = devices[i];
    size_t g2 = 0;
    B2I_DRV(D.MemGetAllocationGranularity(&g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (g2 > gran) gran = g2;
  }
  m->size = (bytes + gran - 1) / gran * gran;
  mp.size = m->size;
  B2I_DRV(D.MulticastCreate(&m->mc, &mp));
  for (int i = 0; i < n; ++i) {  // every device joins before any memory is bound
    CUdevice dev;
    B2I_DRV(D.DeviceGet(&dev, devices[i]));
    B2I_DRV(D.MulticastAddDevice(m->mc, dev));
  }
  std::vector<CUmemAccessDesc> acc(n);
  for (int i = 0; i < n; ++i) {
    acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc[i].location.id = devices[i];
    acc[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  for (int i = 0; i < n; ++i) {
    if (cudaSetDevice(devices[i]) != cudaSuccess) {
      b2i_mcast_destroy(m);
      B2I_FAIL(B2I_ERR_CUDA, "cudaSetDevice(%d) failed", devices[i]);
    }
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof ap);
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = devices[i];
    B2I_DRV(D.MemCreate(&m->mem[i], m->size, &ap, 0));
    B2I_DRV(D.MulticastBindMem(m->mc, 0, m->mem[i], 0, m->size, 0));
    m->bound[i] = true;
    B2I_DRV(D.MemAddressReserve(&m->uc[i], m->size, gran, 0, 0));
    B2I_DRV(D.MemMap(m->uc[i], m->size, 0, m->mem[i], 0));
    m->mapped[i] = true;
    B2I_DRV(D.MemSetAccess(m->uc[i], m->size, acc.data(), (size_t)n));
  }
  B2I_DRV(D.MemAddressReserve(&m->mc_ptr, m->size, gran, 0, 0));
  B2I_DRV(D.MemMap(m->mc_ptr, m->size, 0, m->mc, 0));
  m->mc_mapped = true;
  B2I_DRV(D.MemSetAccess(m->mc_ptr, m->size, acc.data(), (size_t)n));
  *out = m;
  return B2I_OK;
}

// The address whose stores land in every replica.
extern "C" uint64_t b2i_mcast_ptr(const b2i_mcast_t* m) {
  if (!m) return 0;
  return m->multicast ? (uint64_t)m->mc_ptr : (uint64_t)(uintptr_t)m->plain;
}

// The replica of device number `i` of the creation list (an ordinary device pointer).
extern "C" uint64_t b2i_mcast_replica(const b2i_mcast_t* m, int i) {
  if (!m || i < 0 || i >= m->n) return 0;
  return m->multicast ? (uint64_t)m->uc[i] : (uint64_t)(uintptr_t)m->plain;
}

extern "C" size_t b2i_mcast_bytes(const b2i_mcast_t* m) { return m ? m->size : 0; }
extern "C" int b2i_mcast_is_multicast(const b2i_mcast_t* m) { return m && m->multicast; }
