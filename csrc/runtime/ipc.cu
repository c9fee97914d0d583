// Multi-process plumbing: CUDA IPC buffers, pinned host registration and flag kernels.
//
// `torchrun` starts one process per GPU (the shape of `mpirun -np N`, kernel.cu:104-107). The
// filter kernels still read and write peer memory directly; what changes is how a rank obtains a
// pointer into another rank's GPU: through cudaIpcGetMemHandle / cudaIpcOpenMemHandle instead of
// cudaDeviceEnablePeerAccess. The 64-byte handles travel over torch.distributed; no image data
// ever does. MPI_Barrier (kernel.cu:127,133) becomes device-side epoch flags.
#include <cuda_runtime.h>

#include <algorithm>

#include <cstring>

#include "../core/error.hpp"
#include "../kernels/device_utils.cuh"

#define B2I_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) B2I_FAIL(B2I_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

namespace {

// Flag addresses travel by value: no device buffer to allocate or upload.
struct FlagList {
  uint64_t p[B2I_MAX_GPUS];
};
// Waits until every flag has reached `epoch`; records a timeout in *status instead of hanging.
__global__ void wait_flags_kernel_v(FlagList flags, int n, uint32_t epoch, uint32_t* status,
                                    unsigned long long timeout_ns) {
  const int i = threadIdx.x;
  if (i < n && flags.p[i]) {
    if (!b2i::dev::wait_flag_ge(reinterpret_cast<const uint32_t*>(flags.p[i]), epoch, timeout_ns))
      atomicExch(status, 100u + i);
  }
}

__global__ void set_flags_kernel(FlagList flags, int n, uint32_t epoch) {
  const int i = threadIdx.x;
  if (i < n && flags.p[i]) {
    __threadfence_system();
    b2i::dev::st_release_sys(reinterpret_cast<uint32_t*>(flags.p[i]), epoch);
  }
}

}  // namespace

// Device memory that other processes can map. handle_out receives the 64-byte IPC handle.
extern "C" int b2i_ipc_alloc(size_t bytes, uint64_t* ptr, uint8_t* handle_out /*64*/) {
  B2I_CHECK_ARG(bytes > 0 && ptr && handle_out, "bad arguments");
  void* p = nullptr;
  B2I_CUDA(cudaMalloc(&p, bytes));
  B2I_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    B2I_FAIL(B2I_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof h == 64, "IPC handle size");
  memcpy(handle_out, &h, 64);
  *ptr = (uint64_t)(size_t)p;
  return B2I_OK;
}

extern "C" int b2i_ipc_open(const uint8_t* handle /*64*/, uint64_t* ptr) {
  B2I_CHECK_ARG(handle && ptr, "null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  B2I_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = (uint64_t)(size_t)p;
  return B2I_OK;
}

extern "C" int b2i_ipc_close(uint64_t ptr) {
  if (ptr) B2I_CUDA(cudaIpcCloseMemHandle(reinterpret_cast<void*>(ptr)));
  return B2I_OK;
}

extern "C" int b2i_dev_free(uint64_t ptr) {
  if (ptr) B2I_CUDA(cudaFree(reinterpret_cast<void*>(ptr)));
  return B2I_OK;
}

// Pins an existing host range (e.g. a shared-memory mapping all ranks see) so that every rank
// can cudaMemcpyAsync its band straight out of / into it.
extern "C" int b2i_host_register(void* p, size_t bytes) {
  B2I_CHECK_ARG(p && bytes, "bad arguments");
  B2I_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
  return B2I_OK;
}
extern "C" int b2i_host_unregister(void* p) {
  if (p) B2I_CUDA(cudaHostUnregister(p));
  return B2I_OK;
}

// Stream-ordered: the stream continues once all `n` flags (device addresses, 0 = skip) have
// reached `epoch`. `status` (u32, device) receives 100+i if flag i timed out.
extern "C" int b2i_cuda_wait_flags(const uint64_t* flags, int n, uint32_t epoch, uint64_t status,
                                   uint64_t timeout_ns, void* stream) {
  B2I_CHECK_ARG(flags && n >= 0 && n <= B2I_MAX_GPUS && status, "bad arguments");
  if (n == 0) return B2I_OK;
  FlagList fl;
  memset(&fl, 0, sizeof fl);
  for (int i = 0; i < n; ++i) fl.p[i] = flags[i];
  wait_flags_kernel_v<<<1, 32, 0, (cudaStream_t)stream>>>(fl, n, epoch, reinterpret_cast<uint32_t*>(status),
                                                         timeout_ns ? timeout_ns : 2000000000ull);
  B2I_CUDA(cudaGetLastError());
  return B2I_OK;
}

// Stream-ordered: after everything before it on the stream, write `epoch` to each flag.
extern "C" int b2i_cuda_set_flags(const uint64_t* flags, int n, uint32_t epoch, void* stream) {
  B2I_CHECK_ARG(flags && n >= 0 && n <= B2I_MAX_GPUS, "bad arguments");
  if (n == 0) return B2I_OK;
  FlagList fl;
  memset(&fl, 0, sizeof fl);
  for (int i = 0; i < n; ++i) fl.p[i] = flags[i];
  set_flags_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(fl, n, epoch);
  B2I_CUDA(cudaGetLastError());
  return B2I_OK;
}

// K back-to-back steps of one band launch with rising epochs, driven from C so that the host
// cost per step is one kernel launch (the Python/ctypes call overhead would otherwise exceed the
// kernel time). Step i signals epoch_base+i+1 through the launch's done/signal flags; when
// n_wait > 0 (the root rank) a flag-wait kernel follows each step, which makes "step i complete"
// mean "every rank's rows of image i are in the root image". A non-root rank passes the address
// of its done flag on the root as signal_flag.
extern "C" int b2i_cuda_run_steps(b2i_plan_t* plan, const b2i_launch_t* launch, int steps,
                                  uint32_t epoch_base, const uint64_t* wait_flags, int n_wait,
                                  uint64_t signal_flag) {
  B2I_CHECK_ARG(plan && launch && steps >= 0, "bad arguments");
  b2i_launch_t L = *launch;
  // The flags ride in the launch: the sweep kernels publish / wait in their last warp, any other
  // kernel is followed by one-thread flag kernels (b2i_cuda_run). The fused wait wants the flags
  // as one u32 array; anything else keeps the separate wait kernel.
  const bool empty = L.y0 == L.y1;  // no rows: no kernel to carry the flags
  L.done_flag = empty ? 0 : signal_flag;
  bool fused_wait = false;
  if (n_wait > 0 && !empty) {
    uint64_t base = wait_flags[0];
    for (int k = 1; k < n_wait; ++k) base = std::min(base, wait_flags[k]);
    uint32_t mask = 0;
    bool ok = true;
    for (int k = 0; k < n_wait; ++k) {
      const uint64_t off = wait_flags[k] - base;
      if ((off & 3) || off / 4 >= B2I_MAX_GPUS) ok = false;
      else mask |= 1u << (off / 4);
    }
    if (ok) {
      L.wait_done_flags = base;
      L.wait_done_mask = mask;
      fused_wait = true;
    }
  }
  for (int i = 0; i < steps; ++i) {
    const uint32_t epoch = epoch_base + (uint32_t)i + 1;
    L.signal_epoch = epoch;
    B2I_TRY(b2i_cuda_run(plan, &L));
    if (empty && signal_flag) B2I_TRY(b2i_cuda_set_flags(&signal_flag, 1, epoch, L.stream));
    if (n_wait > 0 && !fused_wait) {
      B2I_CHECK_ARG(L.status, "waiting needs the status word");
      B2I_TRY(b2i_cuda_wait_flags(wait_flags, n_wait, epoch, L.status, L.timeout_ns, L.stream));
    }
  }
  return B2I_OK;
}

// `passes` applications of a chain on a resident band: even passes read buffer A and write B,
// odd passes the reverse (the two launch descriptions are prepared by the caller). Pass p waits
// for its neighbours' epoch_base+p halo flags (p > 0) and signals epoch_base+p+1.
extern "C" int b2i_cuda_run_resident_steps(b2i_plan_t* plan, const b2i_launch_t* even,
                                           const b2i_launch_t* odd, int passes, uint32_t epoch_base) {
  B2I_CHECK_ARG(plan && even && odd && passes >= 0, "bad arguments");
  for (int p = 0; p < passes; ++p) {
    b2i_launch_t L = (p & 1) ? *odd : *even;
    const bool more = p + 1 < passes;
    if (p == 0) L.wait_up_flag = L.wait_dn_flag = 0;
    L.wait_epoch = epoch_base + (uint32_t)p;
    L.signal_epoch = epoch_base + (uint32_t)p + 1;
    if (!more) {
      L.push_up.ptr = L.push_dn.ptr = 0;
      L.push_rows = 0;
      L.signal_up_flag = L.signal_dn_flag = 0;
    }
    B2I_TRY(b2i_cuda_run(plan, &L));
  }
  return B2I_OK;
}

// The same with a cycle of `period` launch descriptions (pass p runs descs[p % period]): the caller
// exchanges halo rows only every few passes and lets the passes between compute a shrinking margin of
// ghost rows themselves (parallel/dist.py: run_resident(cycle=...)). Flags are waited and signalled
// in every pass as above; which passes push rows is up to the descriptions.
extern "C" int b2i_cuda_run_resident_cycle(b2i_plan_t* plan, const b2i_launch_t* descs, int period, int passes,
                                           uint32_t epoch_base) {
  B2I_CHECK_ARG(plan && descs && period >= 1 && passes >= 0, "bad arguments");
  for (int p = 0; p < passes; ++p) {
    b2i_launch_t L = descs[p % period];
    const bool more = p + 1 < passes;
    if (p == 0) L.wait_up_flag = L.wait_dn_flag = 0;
    L.wait_epoch = epoch_base + (uint32_t)p;
    L.signal_epoch = epoch_base + (uint32_t)p + 1;
    if (!more) {
      L.push_up.ptr = L.push_dn.ptr = 0;
      L.push_rows = 0;
      L.signal_up_flag = L.signal_dn_flag = 0;
    }
    // A pass that only waits: a one-warp kernel waits for the flags, then the pass runs as an ordinary
    // single-GPU launch (measured on a 540-row band: 14.6 us with the wait inside the filter kernel's
    // multi-GPU instantiation, 11 us for the plain kernel).
    if ((L.wait_up_flag || L.wait_dn_flag) && !L.push_up.ptr && !L.push_dn.ptr && !L.signal_up_flag && !L.signal_dn_flag) {
      uint64_t fl[2];
      int n = 0;
      if (L.wait_up_flag) fl[n++] = L.wait_up_flag;
      if (L.wait_dn_flag) fl[n++] = L.wait_dn_flag;
      B2I_CHECK_ARG(L.status, "waiting needs the status word");
      B2I_TRY(b2i_cuda_wait_flags(fl, n, L.wait_epoch, L.status, L.timeout_ns, L.stream));
      L.wait_up_flag = L.wait_dn_flag = 0;
      L.push_rows = 0;
    }
    B2I_TRY(b2i_cuda_run(plan, &L));
  }
  return B2I_OK;
}

// 2-D copy between any two of {host, local device, peer/IPC-mapped device}; kind: 0 H2D, 1 D2H,
// 2 D2D. Asynchronous on `stream` (host memory should be pinned).
extern "C" int b2i_cuda_memcpy2d(uint64_t dst, size_t dpitch, uint64_t src, size_t spitch, size_t width,
                                 size_t height, int kind, void* stream) {
  const cudaMemcpyKind k = kind == 0 ? cudaMemcpyHostToDevice : kind == 1 ? cudaMemcpyDeviceToHost
                                                                          : cudaMemcpyDeviceToDevice;
  if (width == 0 || height == 0) return B2I_OK;
  B2I_CUDA(cudaMemcpy2DAsync(reinterpret_cast<void*>(dst), dpitch, reinterpret_cast<const void*>(src), spitch,
                             width, height, k, (cudaStream_t)stream));
  return B2I_OK;
}
