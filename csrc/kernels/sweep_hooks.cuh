// Multi-GPU hooks shared by the sweep kernels (sweep5.cu, sweep_sobel.cu; sweep3.cu uses the kernel
// end only). Every kernel's parameter struct carries the same fields (sweep5.hpp, sweep_sobel.hpp):
// neighbour flags and epochs, push pointers, counters, the status word and the timeout.
#pragma once
#include "device_utils.cuh"
#include "sweep5.hpp"

#include <type_traits>

namespace b2i {
namespace dev {

// Lane 0, before a task's first TMA load. A task that reads ghost rows, or writes the neighbour's,
// first sees the neighbour's flag of the previous pass: its rows have landed here and it has
// finished reading what this task overwrites. `reach` = rows of halo the stencil reads beyond the
// task's own rows.
// This GPU gives up (a wait timed out, or a neighbour's flag was poison): record why, and poison the
// flags this launch would have released, so that the neighbours stop waiting too.
template <class Params>
__device__ __forceinline__ void sweep_abort(const Params& P, uint32_t code, uint32_t epoch, int n_flags = 1) {
  atomicExch(reinterpret_cast<uint32_t*>(P.status), code);
  for (int i = 0; i < n_flags; ++i) {
    if (P.signal_up_flag) flag_poison(reinterpret_cast<uint32_t*>(P.signal_up_flag) + i, epoch);
    if (P.signal_dn_flag) flag_poison(reinterpret_cast<uint32_t*>(P.signal_dn_flag) + i, epoch);
  }
}

template <class Params>
__device__ __forceinline__ void sweep_wait_neighbours(const Params& P, int ya, int yb, int reach, bool sig_up,
                                                      bool sig_dn, uint32_t pass = 0) {
  const uint32_t* st = reinterpret_cast<const uint32_t*>(P.status);
  if (P.wait_up_flag && (ya - reach < P.y0 || sig_up)) {
    if (!wait_flag_ge(reinterpret_cast<const uint32_t*>(P.wait_up_flag), P.wait_epoch + pass, P.timeout_ns, st))
      sweep_abort(P, 1u, P.signal_epoch + pass);
    asm volatile("fence.proxy.async;" ::: "memory");  // the flag (generic proxy) before TMA loads of the ghost rows
  }
  if (P.wait_dn_flag && (yb + reach > P.y1 || sig_dn)) {
    if (!wait_flag_ge(reinterpret_cast<const uint32_t*>(P.wait_dn_flag), P.wait_epoch + pass, P.timeout_ns, st))
      sweep_abort(P, 2u, P.signal_epoch + pass);
    asm volatile("fence.proxy.async;" ::: "memory");
  }
}

// Output row oy of a 3-channel strip, 24 bytes per lane in r[0..5]: the first / last push_rows rows
// of the band also go straight from registers into the neighbours' ghost rows (peer or IPC memory).
// `col` = byte offset of the lane's pixels in the row.
template <class Params>
__device__ __forceinline__ void sweep_push_row(const Params& P, int oy, int64_t col, const uint32_t (&r)[6],
                                               bool odd = false) {
  // (resident multi-pass launches push into the other band buffer on odd passes: sweep5.hpp)
  uint64_t up = P.push_up_ptr, dn = P.push_dn_ptr;
  if constexpr (std::is_same<Params, Sweep5Params>::value) {
    if (odd) {
      up = reinterpret_cast<const Sweep5Params&>(P).push_up_ptr_odd;
      dn = reinterpret_cast<const Sweep5Params&>(P).push_dn_ptr_odd;
    }
  }
  if (up && oy < P.y0 + P.push_rows) {
    uint2* g = reinterpret_cast<uint2*>(up + (int64_t)(oy - P.y0) * P.push_pitch + col);
    g[0] = make_uint2(r[0], r[1]);
    g[1] = make_uint2(r[2], r[3]);
    g[2] = make_uint2(r[4], r[5]);
  }
  if (dn && oy >= P.y1 - P.push_rows) {
    uint2* g = reinterpret_cast<uint2*>(dn + (int64_t)(oy - (P.y1 - P.push_rows)) * P.push_pitch + col);
    g[0] = make_uint2(r[0], r[1]);
    g[1] = make_uint2(r[2], r[3]);
    g[2] = make_uint2(r[4], r[5]);
  }
}

// Whole warp, after a task that pushed rows: count the task in; the last task of each side releases
// the neighbour's flag (the counters are left zeroed for the next pass).
// `pass`: resident multi-pass launches count even and odd passes apart (boundary tasks of one side
// are never more than one pass apart: each waits for the neighbour, which waits for all of them),
// and the epoch they publish grows with the pass.
template <class Params>
__device__ __forceinline__ void sweep_signal_neighbours(const Params& P, bool sig_up, bool sig_dn, int lane,
                                                        uint32_t pass = 0) {
  __threadfence_system();  // every lane's pushes
  __syncwarp();
  if (lane != 0) return;
  uint32_t* counters = reinterpret_cast<uint32_t*>(P.counters) + (pass & 1) * 4;
  if (sig_up && atomicAdd(&counters[0], 1u) + 1 == (uint32_t)P.n_up_tasks) {
    atomicExch(&counters[0], 0u);
    st_release_sys(reinterpret_cast<uint32_t*>(P.signal_up_flag), P.signal_epoch + pass);
  }
  if (sig_dn && atomicAdd(&counters[1], 1u) + 1 == (uint32_t)P.n_dn_tasks) {
    atomicExch(&counters[1], 0u);
    st_release_sys(reinterpret_cast<uint32_t*>(P.signal_dn_flag), P.signal_epoch + pass);
  }
}

// Whole warp, end of the kernel: the warp's stores complete; with ROOT mode's completion flags the
// last warp that had work waits for the other GPUs' done flags and / or publishes this GPU's.
template <class Params>
__device__ __forceinline__ void sweep_kernel_end(const Params& P, int global_warp, int total_warps, int lane) {
  if (P.done_flag | P.wait_done_mask) __threadfence_system();  // every lane's plain stores before the flag
  __syncwarp();
  if (lane != 0) return;
  tma_store_wait_all();
  if ((P.done_flag | P.wait_done_mask) && global_warp < P.n_tasks)
    sweep_finish(reinterpret_cast<uint32_t*>(P.counters) + 2, (uint32_t)min(total_warps, P.n_tasks), P.done_flag,
                 P.signal_epoch, P.wait_done_flags, P.wait_done_mask, reinterpret_cast<uint32_t*>(P.status),
                 P.timeout_ns);
}

}  // namespace dev
}  // namespace b2i
