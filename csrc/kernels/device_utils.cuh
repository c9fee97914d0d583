// sm_100a device helpers: mbarrier + TMA (cp.async.bulk.tensor) wrappers, system-scope
// release/acquire flags for cross-GPU ordering, and the byte-SIMD arithmetic the filters share.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2i {
namespace dev {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- mbarrier -------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
// Waits for the phase with the given parity. A TMA transfer that never completes (bad tensor
// map, faulting peer) would otherwise spin forever and hang the GPU. After ~4 s the wait gives up:
// with a status word (multi-GPU launches pass theirs) it records code 5 there and the thread leaves
// the kernel, so the launch ends and the context stays usable; without one it traps, which surfaces
// as a launch failure on the host.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, uint32_t* status = nullptr) {
  const uint32_t addr = smem_u32(bar);
  unsigned long long t0 = 0;
  for (uint32_t spins = 0;; ++spins) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) return;
    if ((spins & 0xFFF) == 0xFFF) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t0 == 0) t0 = t;
      else if (t - t0 > 4000000000ull) {
        if (!status) __trap();
        atomicExch(status, 5u);
        asm volatile("exit;");
      }
    }
  }
}

// ---- programmatic dependent launch ---------------------------------------------------------------------
// A kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start while the kernel
// before it in the stream is still running, once every CTA of that kernel has called
// pdl_launch_dependents() (or exited). Everything before pdl_wait() must not touch memory the earlier
// kernel writes: the sweeps run their prologue there (barrier init, table copy, descriptor prefetch).
// pdl_wait() returns when the earlier kernel has completed and its writes are visible. Both are no-ops in
// a launch without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- TMA ------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// 2-D tile global -> shared; completes `bytes` on the mbarrier. Out-of-range elements read 0.
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* m, int x, int y,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(x), "r"(y), "r"(smem_u32(bar))
      : "memory");
}
// 2-D tile shared -> global; out-of-range elements are not written.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, int x, int y, const void* src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::
                   "l"(reinterpret_cast<uint64_t>(m)),
               "r"(x), "r"(y), "r"(smem_u32(src))
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// Source shared memory may be overwritten again once this returns.
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// The stores are complete (globally performed) once this returns.
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// Orders generic-proxy shared-memory writes before async-proxy (TMA) reads of them.
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- NVLS multicast stores ------------------------------------------------------------------------
// Stores to a multicast address (b2i_mcast_ptr): the NVSwitch replicates them into every bound
// GPU's copy. On an ordinary address they are ordinary stores (same SASS: STG.E.128).

__device__ __forceinline__ void st_mc_v4(uint64_t addr, uint4 v) {
  asm volatile("multimem.st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
// A finished box of `rows` staged rows (row pitch `row_bytes`, a multiple of 16) leaves for an NVLS
// multicast image in coalesced 16-byte stores by the whole warp; `valid_bytes` (a multiple of 16)
// clips the last strip of a row.
__device__ __forceinline__ void mc_store_box(const uint8_t* buf, int rows, int row_bytes, int valid_bytes,
                                             uint64_t g, int64_t pitch, int lane) {
  __syncwarp();  // the staging writes of every lane
  for (int r = 0; r < rows; ++r)
    for (int k = lane * 16; k < valid_bytes; k += 32 * 16)
      st_mc_v4(g + (int64_t)r * pitch + k, *reinterpret_cast<const uint4*>(buf + r * row_bytes + k));
  __syncwarp();  // the buffer may be written again
}

// ---- cross-GPU flags (peer-mapped or IPC-mapped global memory) ---------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// A flag that is kAbortBias or more ahead of any epoch asked for is poison: its writer gave up
// (timed out itself, or saw poison) and says so to whoever waits on it, so that one lost GPU ends
// every GPU's kernel within one timeout instead of one timeout per neighbour in a row.
constexpr uint32_t kAbortBias = 1u << 30;
__device__ __forceinline__ void flag_poison(uint32_t* p, uint32_t epoch) { st_release_sys(p, epoch + kAbortBias); }
// Spins until *p >= want (epochs only grow). Gives up after `timeout_ns`, when the flag is poison,
// or when this GPU's status word (if given) is already nonzero, so that a lost peer shows up as an
// error code instead of a hung GPU (the reference's peers hang, kernel.cu:111-114,127).
// Returns false when it gave up.
__device__ __forceinline__ bool wait_flag_ge(const uint32_t* p, uint32_t want, unsigned long long timeout_ns,
                                             const uint32_t* status = nullptr) {
  uint32_t d = ld_acquire_sys(p) - want;
  if ((int32_t)d >= 0) return d < (kAbortBias >> 1);
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    d = ld_acquire_sys(p) - want;
    if ((int32_t)d >= 0) return d < (kAbortBias >> 1);
    if (status && *reinterpret_cast<const volatile uint32_t*>(status) != 0) return false;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) return false;
    __nanosleep(64);
  }
}

// The same at GPU scope, for flags written by other SMs of this GPU (task-to-task ordering inside
// a resident multi-pass kernel).
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool wait_flag_ge_gpu(const uint32_t* p, uint32_t want, unsigned long long timeout_ns,
                                                 const uint32_t* status = nullptr) {
  if ((int32_t)(ld_acquire_gpu(p) - want) >= 0) return true;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    if ((int32_t)(ld_acquire_gpu(p) - want) >= 0) return true;
    if (status && *reinterpret_cast<const volatile uint32_t*>(status) != 0) return false;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) return false;
  }
}

// End of a sweep kernel (lane 0 of every warp that had work, after its stores have completed): the
// last warp to arrive waits for the other GPUs' done flags (the root's side) and / or publishes this
// GPU's done flag (everyone else's side) -- MPI_Gather's completion (kernel.cu:223) without a
// separate kernel. counter: a zeroed u32 that is left zeroed.
__device__ __forceinline__ void sweep_finish(uint32_t* counter, uint32_t n_warps, uint64_t done_flag, uint32_t epoch,
                                             uint64_t wait_flags, uint32_t wait_mask, uint32_t* status,
                                             unsigned long long timeout_ns) {
  asm volatile("fence.proxy.async;" ::: "memory");  // TMA stores (async proxy) before the flag (generic proxy)
  __threadfence_system();
  if (atomicAdd(counter, 1u) + 1 != n_warps) return;
  atomicExch(counter, 0u);
  for (uint32_t m = wait_mask; m; m &= m - 1) {
    const int r = __ffs(m) - 1;
    if (!wait_flag_ge(reinterpret_cast<const uint32_t*>(wait_flags) + r, epoch, timeout_ns)) atomicExch(status, 3u);
  }
  if (done_flag) st_release_sys(reinterpret_cast<uint32_t*>(done_flag), epoch);
}

// ---- streaming vector loads/stores (local HBM or NVLink peer memory) --------------------------
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream_v4(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// ---- byte arithmetic ----------------------------------------------------------------------------

// kernel.cu:39-42 numerics without FP64 or integer division: floor(c*v) for an 8-bit v is
// one FFMA with round-down when v rides in the mantissa of 2^23+v and c = n/2^17 (n chosen so
// floor(n*v/2^17) == floor(c*v) for every v in 0..255; checked exhaustively by the tests).
// Three chained FFMA.RM give floor(.11B)+floor(.59G)+floor(.3R) in the low byte of the result:
// every partial sum stays in [2^23, 2^24) where the fp32 grid is exactly the integers.
struct GrayRefCoef {
  float c0, c1, c2;  // coefficient of byte 0, 1, 2 of a pixel (BGR: .11,.59,.3; RGB: .3,.59,.11)
};
__host__ __device__ inline GrayRefCoef gray_ref_coef(int rgb_order) {
  const float cb = 14418.0f / 131072.0f, cg = 77333.0f / 131072.0f, cr = 39322.0f / 131072.0f;
  GrayRefCoef k;
  k.c0 = rgb_order ? cr : cb;
  k.c1 = cg;
  k.c2 = rgb_order ? cb : cr;
  return k;
}
// Byte IDX (0..3) of w as the float 2^23 + v: one PRMT against the constant 0x4B000000.
template <int IDX>
__device__ __forceinline__ float byte_to_f23(uint32_t w) {
  return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540 + IDX));
}
// f0,f1,f2 = byte_to_f23 of the pixel's bytes 0,1,2. Returns a word whose LOW BYTE is the gray
// value (the upper bytes hold 0x4B74 00 and must be masked or dropped by the caller).
__device__ __forceinline__ uint32_t gray_ref_f23(float f0, float f1, float f2,
                                                 const GrayRefCoef& k) {
  // 7602112 = 15990784 - 2^23*(c0+c1+c2); the three scaled constants are integers.
  float y = __fmaf_rd(f0, k.c0, 7602112.0f);
  y = __fmaf_rd(f1, k.c1, y);
  y = __fmaf_rd(f2, k.c2, y);
  return __float_as_uint(y);
}

// cv::cvtColor(BGR2GRAY) for 8-bit input (kern.cpp:73): 15-bit fixed point, round to nearest.
__device__ __forceinline__ uint32_t gray_cv_px(uint32_t b0, uint32_t b1, uint32_t b2,
                                               int rgb_order) {
  const uint32_t k0 = rgb_order ? 9798u : 3735u, k2 = rgb_order ? 3735u : 9798u;
  return (b0 * k0 + b1 * 19235u + b2 * k2 + 16384u) >> 15;
}

// Four unsigned pixel bytes times four signed weight bytes, accumulated: IDP.4A.U8.S8.
__device__ __forceinline__ int dp4a_us(uint32_t px, uint32_t w, int acc) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px), "r"(w), "r"(acc));
  return d;
}

// Packs the low bytes of four words into one word.
__device__ __forceinline__ uint32_t pack4_lo(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  return __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
}

// clamp to [0,255] four int32 and pack: two cvt.pack.sat.u8.s32.
__device__ __forceinline__ uint32_t sat_pack4(int a, int b, int c, int d) {
  uint32_t r;
  asm("{\n\t.reg .b32 t;\n\t"
      "cvt.pack.sat.u8.s32.b32 t, %4, %3, 0;\n\t"
      "cvt.pack.sat.u8.s32.b32 %0, %2, %1, t;\n\t}"
      : "=r"(r)
      : "r"(a), "r"(b), "r"(c), "r"(d));
  return r;
}

}  // namespace dev
}  // namespace b2i
