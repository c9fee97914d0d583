// In-process multi-GPU runtime: topology discovery, peer access, band buffers with ghost rows,
// and the three ways an image crosses GPUs.
//
// This layer stands where the reference has MPI (kernel.cu:104-107,127-137,223-225): instead of
// N OS processes that all call cudaSetDevice(0) (kernel.cu:147) and move bands through host
// memory with MPI_Scatter / MPI_Gather, ONE process drives N GPUs, enables peer access between
// all of them and lets the filter kernels themselves read and write peer memory:
//
//   ROOT mode      root-owned input and output like the reference. GPU i's kernel loads its band
//                  (+halo) straight from the root GPU's image and stores its rows straight into
//                  the root GPU's output: scatter and gather are fused into the filter pass.
//   RESIDENT mode  bands (+ghost rows) live on their GPUs across passes; every pass stores its
//                  boundary rows into the neighbours' ghost rows as part of its output phase, so
//                  only halos cross NVLink. Ordering between passes: cross-device stream events,
//                  or device-side release/acquire flags (B200IMG_SYNC=flags).
//   DIRECT mode    host-resident images: every GPU copies its own band over its own PCIe link
//                  (pinned memory, async) — no NVLink hop, no root bottleneck.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../core/error.hpp"
#include "../core/trace.hpp"
#include "../kernels/tile_plan.hpp"

#define B2I_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess)                                                              \
      B2I_FAIL(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver             \
                   ? B2I_ERR_NO_DEVICE                                                  \
                   : B2I_ERR_CUDA,                                                      \
               "%s: %s", #expr, cudaGetErrorString(e_));                                \
  } while (0)

using b2i::kGhostRows;

// ---- topology -----------------------------------------------------------------------------------

extern "C" int b2i_topology_json(char* buf, size_t len) {
  B2I_CHECK_ARG(buf && len > 0, "null buffer");
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  std::string s = "{\"devices\": [";
  PFN_cuDeviceGetAttribute_v2000 get_attr = nullptr;
  {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (n > 0 && cudaGetDriverEntryPoint("cuDeviceGetAttribute", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      get_attr = reinterpret_cast<PFN_cuDeviceGetAttribute_v2000>(p);
  }
  char tmp[512];
  for (int d = 0; d < n; ++d) {
    cudaDeviceProp pr;
    B2I_CUDA(cudaGetDeviceProperties(&pr, d));
    int mc = -1;
    if (get_attr) {
      int v = 0;
      if (get_attr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) == CUDA_SUCCESS) mc = v;
    }
    snprintf(tmp, sizeof tmp,
             "%s{\"id\": %d, \"name\": \"%s\", \"cc\": \"%d.%d\", \"sms\": %d, \"mem_gb\": %.1f, "
             "\"l2_mb\": %.1f, \"smem_per_sm_kb\": %d, \"multicast\": %d}",
             d ? ", " : "", d, pr.name, pr.major, pr.minor, pr.multiProcessorCount,
             pr.totalGlobalMem / 1e9, pr.l2CacheSize / 1048576.0,
             (int)(pr.sharedMemPerMultiprocessor / 1024), mc);
    s += tmp;
  }
  s += "], \"p2p\": [";
  for (int a = 0; a < n; ++a) {
    s += a ? ", [" : "[";
    for (int b = 0; b < n; ++b) {
      int access = a == b, atomics = a == b, rank = 0;
      if (a != b) {
        cudaDeviceCanAccessPeer(&access, a, b);
        cudaDeviceGetP2PAttribute(&atomics, cudaDevP2PAttrNativeAtomicSupported, a, b);
        cudaDeviceGetP2PAttribute(&rank, cudaDevP2PAttrPerformanceRank, a, b);
      }
      snprintf(tmp, sizeof tmp, "%s{\"access\": %d, \"atomics\": %d, \"perf_rank\": %d}",
               b ? ", " : "", access, atomics, rank);
      s += tmp;
    }
    s += "]";
  }
  s += "]}";
  B2I_CHECK_ARG(s.size() + 1 <= len, "topology report needs %zu bytes", s.size() + 1);
  memcpy(buf, s.c_str(), s.size() + 1);
  return B2I_OK;
}

// ---- context --------------------------------------------------------------------------------------

namespace {

struct Dev {
  int id = -1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
  std::vector<cudaEvent_t> ev_pass;  // one per resident pass, recycled
  uint8_t* band[2] = {nullptr, nullptr};  // resident ping-pong band buffers (with ghost rows)
  uint8_t* tmp[2] = {nullptr, nullptr};   // scratch for multi-segment plans
  uint32_t* words = nullptr;  // [0..3] counters, [4] status, [8] flag_up, [9] flag_dn, [10] done
  size_t band_bytes = 0, tmp_bytes = 0;
  int y0 = 0, y1 = 0;
};

}  // namespace

struct b2i_ctx {
  int n = 0, root = 0;
  std::vector<Dev> devs;
  // root-owned images (device memory of the root GPU)
  uint8_t *root_src = nullptr, *root_dst = nullptr;
  size_t root_src_bytes = 0, root_dst_bytes = 0;
  int W = 0, H = 0, C = 0, out_c = 0;
  size_t src_pitch = 0, dst_pitch = 0, band_pitch = 0;
  cudaEvent_t ev_src_ready = nullptr;
  uint32_t epoch = 0;
  bool peer_ok = true;
  double root_share = 0;  // ROOT mode: fraction of the rows the root GPU keeps (0 = equal bands)
  // replicated output (b2i_ctx_set_replicated): one copy of the result per physical GPU behind one
  // NVLS multicast address
  bool replicated = false;
  b2i_mcast_t* rep = nullptr;
  std::vector<int> rep_devs;  // distinct physical devices, in order of first use
};

namespace {

size_t align16(size_t v) { return (v + 15) & ~size_t(15); }

int ensure(uint8_t** p, size_t* have, size_t want) {
  if (*have >= want) return B2I_OK;
  if (*p) B2I_CUDA(cudaFree(*p));
  *p = nullptr;
  *have = 0;
  B2I_CUDA(cudaMalloc(p, want));
  *have = want;
  return B2I_OK;
}

}  // namespace

extern "C" void b2i_ctx_destroy(b2i_ctx* c) {
  if (!c) return;
  for (Dev& d : c->devs) {
    if (d.id < 0) continue;
    cudaSetDevice(d.id);
    for (int k = 0; k < 2; ++k) {
      if (d.band[k]) cudaFree(d.band[k]);
      if (d.tmp[k]) cudaFree(d.tmp[k]);
    }
    if (d.words) cudaFree(d.words);
    for (cudaEvent_t e : d.ev_pass) cudaEventDestroy(e);
    if (d.ev_start) cudaEventDestroy(d.ev_start);
    if (d.ev_stop) cudaEventDestroy(d.ev_stop);
    if (d.stream) cudaStreamDestroy(d.stream);
  }
  if (c->rep) b2i_mcast_destroy(c->rep);
  if (c->n > 0) {
    cudaSetDevice(c->devs[c->root].id);
    if (c->root_src) cudaFree(c->root_src);
    if (c->root_dst) cudaFree(c->root_dst);
    if (c->ev_src_ready) cudaEventDestroy(c->ev_src_ready);
  }
  delete c;
}

// Discovers the devices (MPI_Comm_size/rank in the reference, kernel.cu:106-107), enables peer
// access all-to-all and creates one stream and a few words of flag memory per GPU.
extern "C" int b2i_ctx_create(int n_gpus, const int* devices, int root, b2i_ctx** out) {
  B2I_CHECK_ARG(out, "null argument");
  int avail = 0;
  B2I_CUDA(cudaGetDeviceCount(&avail));
  if (avail == 0) B2I_FAIL(B2I_ERR_NO_DEVICE, "no CUDA device");
  B2I_CHECK_ARG(n_gpus >= 1 && n_gpus <= B2I_MAX_GPUS, "gpus must be in [1,%d]", B2I_MAX_GPUS);
  B2I_CHECK_ARG(root >= 0 && root < n_gpus, "root %d is not one of the %d GPUs", root, n_gpus);
  b2i_ctx* c = new b2i_ctx();
  c->n = n_gpus;
  c->root = root;
  c->devs.resize(n_gpus);
  // "virtual" GPUs: with fewer physical devices than requested, bands share devices. Peer
  // pointers are then local pointers and launches serialise, which is how the multi-GPU paths
  // are tested on one GPU.
  for (int i = 0; i < n_gpus; ++i) c->devs[i].id = devices ? devices[i] : i % avail;
  auto bail = [&](int rc) {
    std::string msg = b2i_last_error();
    b2i_ctx_destroy(c);
    b2i::last_error_slot() = msg;
    return rc;
  };
  for (int i = 0; i < n_gpus; ++i) {
    Dev& d = c->devs[i];
    if (d.id < 0 || d.id >= avail) return bail(b2i::fail(B2I_ERR_ARG, __FILE__, __LINE__, "device %d does not exist", d.id));
    if (cudaSetDevice(d.id) != cudaSuccess) return bail(b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "cudaSetDevice(%d) failed", d.id));
    for (int j = 0; j < n_gpus; ++j) {
      const int pj = c->devs[j].id >= 0 ? c->devs[j].id : (devices ? devices[j] : j % avail);
      if (pj == d.id) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, d.id, pj);
      if (!can) {
        c->peer_ok = false;
        continue;
      }
      cudaError_t e = cudaDeviceEnablePeerAccess(pj, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) c->peer_ok = false;
      cudaGetLastError();
    }
    if (cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&d.ev_start) != cudaSuccess || cudaEventCreate(&d.ev_stop) != cudaSuccess ||
        cudaMalloc(&d.words, 256) != cudaSuccess || cudaMemset(d.words, 0, 256) != cudaSuccess)
      return bail(b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "per-device setup failed on GPU %d: %s", d.id,
                            cudaGetErrorString(cudaGetLastError())));
  }
  if (!c->peer_ok && n_gpus > 1)
    return bail(b2i::fail(B2I_ERR_UNSUPPORTED, __FILE__, __LINE__,
                          "peer access is not available between all requested GPUs"));
  cudaSetDevice(c->devs[root].id);
  if (cudaEventCreateWithFlags(&c->ev_src_ready, cudaEventDisableTiming) != cudaSuccess)
    return bail(b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "event creation failed"));
  *out = c;
  return B2I_OK;
}

extern "C" int b2i_ctx_gpus(const b2i_ctx* c) { return c ? c->n : 0; }

// Uploads a host image into the root GPU's source image (the reference's rank-0 imread +
// cudaMemcpy, kernel.cu:110,163) and sizes the root-owned output for `out_c` channels.
extern "C" int b2i_ctx_load_root(b2i_ctx* c, const uint8_t* src, int W, int H, int C, size_t pitch,
                                 int out_c) {
  B2I_CHECK_ARG(c && src && W > 0 && H > 0 && (C == 1 || C == 3) && (out_c == 1 || out_c == 3),
                "bad arguments");
  Dev& r = c->devs[c->root];
  B2I_CUDA(cudaSetDevice(r.id));
  c->W = W;
  c->H = H;
  c->C = C;
  c->out_c = out_c;
  c->src_pitch = align16((size_t)W * C);
  c->dst_pitch = align16((size_t)W * out_c);
  B2I_TRY(ensure(&c->root_src, &c->root_src_bytes, c->src_pitch * H));
  B2I_TRY(ensure(&c->root_dst, &c->root_dst_bytes, c->dst_pitch * H));
  if (c->replicated && (!c->rep || b2i_mcast_bytes(c->rep) < c->dst_pitch * H)) {
    if (c->rep) b2i_mcast_destroy(c->rep);
    c->rep = nullptr;
    c->rep_devs.clear();
    for (const Dev& d : c->devs)
      if (std::find(c->rep_devs.begin(), c->rep_devs.end(), d.id) == c->rep_devs.end()) c->rep_devs.push_back(d.id);
    B2I_TRY(b2i_mcast_create((int)c->rep_devs.size(), c->rep_devs.data(), c->dst_pitch * H, &c->rep));
    B2I_CUDA(cudaSetDevice(c->devs[c->root].id));
  }
  B2I_CUDA(cudaMemcpy2DAsync(c->root_src, c->src_pitch, src, pitch, (size_t)W * C, H,
                             cudaMemcpyHostToDevice, r.stream));
  B2I_CUDA(cudaEventRecord(c->ev_src_ready, r.stream));
  for (int i = 0; i < c->n; ++i) b2i_band_rows(H, c->n, i, &c->devs[i].y0, &c->devs[i].y1);
  if (c->root_share > 0 && c->n > 1) {
    // link-aware plan: the root works out of its own HBM, everyone else through the root's one
    // NVLink port, so the root keeps `root_share` of the rows and the others split the rest
    const int cut = std::min(H, std::max(0, (int)(H * c->root_share + 0.5)));
    int next = 0;
    for (int k = 0; k < c->n; ++k) {
      const int i = (c->root + k) % c->n;  // root first, then the others in rank order
      Dev& d = c->devs[i];
      if (k == 0) {
        d.y0 = 0;
        d.y1 = cut;
        next = cut;
      } else {
        int a, b;
        b2i_band_rows(H - cut, c->n - 1, k - 1, &a, &b);
        d.y0 = cut + a;
        d.y1 = cut + b;
        next = d.y1;
      }
    }
    (void)next;
  }
  return B2I_OK;
}

extern "C" int b2i_ctx_set_root_share(b2i_ctx* c, double share) {
  B2I_CHECK_ARG(c && share >= 0 && share <= 1, "share must be in [0,1]");
  c->root_share = share;
  return B2I_OK;
}

extern "C" int b2i_ctx_fetch_root(b2i_ctx* c, uint8_t* dst, size_t pitch) {
  B2I_CHECK_ARG(c && dst && c->root_dst, "nothing to fetch");
  if (c->replicated && c->rep) return b2i_ctx_fetch_replica(c, c->root, dst, pitch);
  Dev& r = c->devs[c->root];
  B2I_CUDA(cudaSetDevice(r.id));
  B2I_CUDA(cudaMemcpy2DAsync(dst, pitch, c->root_dst, c->dst_pitch, (size_t)c->W * c->out_c, c->H,
                             cudaMemcpyDeviceToHost, r.stream));
  B2I_CUDA(cudaStreamSynchronize(r.stream));
  return B2I_OK;
}

extern "C" int b2i_ctx_set_replicated(b2i_ctx* c, int on) {
  B2I_CHECK_ARG(c, "null context");
  if (on) {  // fail here, not at the next load, when the GPUs cannot do it
    std::vector<int> devs;
    for (const Dev& d : c->devs)
      if (std::find(devs.begin(), devs.end(), d.id) == devs.end()) devs.push_back(d.id);
    b2i_mcast_t* probe = nullptr;
    B2I_TRY(b2i_mcast_create((int)devs.size(), devs.data(), 4096, &probe));
    b2i_mcast_destroy(probe);
    B2I_CUDA(cudaSetDevice(c->devs[c->root].id));
  }
  c->replicated = on != 0;
  if (!on && c->rep) {
    b2i_mcast_destroy(c->rep);
    c->rep = nullptr;
  }
  return B2I_OK;
}

extern "C" int b2i_ctx_replicated(const b2i_ctx* c) {
  if (!c || !c->replicated) return 0;
  return c->rep && b2i_mcast_is_multicast(c->rep) ? 2 : c->rep ? 1 : (int)c->replicated;
}

extern "C" int b2i_ctx_fetch_replica(b2i_ctx* c, int gpu, uint8_t* dst, size_t pitch) {
  B2I_CHECK_ARG(c && dst && c->rep && gpu >= 0 && gpu < c->n, "no replicated output to fetch");
  Dev& d = c->devs[gpu];
  const int k = (int)(std::find(c->rep_devs.begin(), c->rep_devs.end(), d.id) - c->rep_devs.begin());
  B2I_CUDA(cudaSetDevice(d.id));
  B2I_CUDA(cudaMemcpy2DAsync(dst, pitch, reinterpret_cast<const void*>(b2i_mcast_replica(c->rep, k)), c->dst_pitch,
                             (size_t)c->W * c->out_c, c->H, cudaMemcpyDeviceToHost, d.stream));
  B2I_CUDA(cudaStreamSynchronize(d.stream));
  return B2I_OK;
}

namespace {

int ensure_tmp(b2i_ctx* c, Dev& d, int n_segments, int rows) {
  if (n_segments <= 1) return B2I_OK;
  const size_t want = (size_t)c->W * 3 * rows;
  for (int k = 0; k < 2; ++k) {
    size_t have = d.tmp[k] ? d.tmp_bytes : 0;
    B2I_TRY(ensure(&d.tmp[k], &have, want));
  }
  d.tmp_bytes = std::max(d.tmp_bytes, want);
  return B2I_OK;
}

}  // namespace

// ROOT mode: `iters` times, every GPU produces its band of chain(root_src) into root_dst.
// kernel_ms = the slowest GPU's device-event time per iteration (max over "ranks").
extern "C" int b2i_ctx_run_root(b2i_ctx* c, b2i_plan_t* plan, int iters, int flags, double* kernel_ms,
                                double* per_gpu_ms /* n entries or null */) {
  b2i::NvtxRange range("b2i_ctx_run_root");
  B2I_CHECK_ARG(c && plan && c->root_src, "load an image first");
  if (iters < 1) iters = 1;
  b2i_plan_info_t info;
  B2I_TRY(b2i_plan_info(plan, &info));
  B2I_CHECK_ARG(info.cin == c->C && info.cout == c->out_c, "plan/channel mismatch");
  const int halo = info.halo;
  Dev& root = c->devs[c->root];
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    if (d.y1 == d.y0) continue;
    B2I_CUDA(cudaSetDevice(d.id));
    const int ra = std::max(0, d.y0 - halo), rb = std::min(c->H, d.y1 + halo);
    B2I_TRY(ensure_tmp(c, d, info.n_segments, rb - ra));
    B2I_CUDA(cudaStreamWaitEvent(d.stream, c->ev_src_ready, 0));
    B2I_CUDA(cudaEventRecord(d.ev_start, d.stream));
    for (int it = 0; it < iters; ++it) {
      b2i_launch_t L;
      memset(&L, 0, sizeof L);
      L.W = c->W;
      L.H = c->H;
      L.y0 = d.y0;
      L.y1 = d.y1;
      L.src = {(uint64_t)c->root_src, (int64_t)c->src_pitch, 0, c->H};  // peer loads: the scatter
      L.dst = {(uint64_t)c->root_dst, (int64_t)c->dst_pitch, 0, c->H};  // peer stores: the gather
      if (c->replicated && c->rep)  // one store per byte, replicated into every GPU by the switch
        L.dst = {b2i_mcast_ptr(c->rep), (int64_t)c->dst_pitch, 0, c->H};
      if (info.n_segments > 1) {
        L.tmp[0] = {(uint64_t)d.tmp[0], (int64_t)c->W * 3, ra, rb};
        L.tmp[1] = {(uint64_t)d.tmp[1], (int64_t)c->W * 3, ra, rb};
      }
      L.stream = d.stream;
      L.flags = flags | (c->replicated && c->rep ? B2I_LAUNCH_DST_MCAST : 0);
      B2I_TRY(b2i_cuda_run(plan, &L));
    }
    B2I_CUDA(cudaEventRecord(d.ev_stop, d.stream));
  }
  // the root learns that the gathered image is complete (MPI_Gather's return, kernel.cu:223)
  B2I_CUDA(cudaSetDevice(root.id));
  for (int i = 0; i < c->n; ++i)
    if (i != c->root && c->devs[i].y1 > c->devs[i].y0)
      B2I_CUDA(cudaStreamWaitEvent(root.stream, c->devs[i].ev_stop, 0));
  double worst = 0;
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    float ms = 0;
    if (d.y1 > d.y0) {
      B2I_CUDA(cudaSetDevice(d.id));
      B2I_CUDA(cudaEventSynchronize(d.ev_stop));
      B2I_CUDA(cudaEventElapsedTime(&ms, d.ev_start, d.ev_stop));
    }
    if (per_gpu_ms) per_gpu_ms[i] = ms / iters;
    worst = std::max(worst, (double)ms / iters);
  }
  if (kernel_ms) *kernel_ms = worst;
  return B2I_OK;
}

// RESIDENT mode. Bands are copied once from the root image into per-GPU buffers that carry
// kGhostRows ghost rows on each side; then `passes` applications of the chain run with the
// bands staying where they are. Each pass stores its first/last rows into the neighbours'
// ghost rows from inside the filter kernel (tile_chain.cu, "halo push"); the result is copied
// back into the root image at the end. use_flags selects device-side release/acquire flags
// instead of cross-device stream events for the pass-to-pass ordering.
extern "C" int b2i_ctx_run_resident(b2i_ctx* c, b2i_plan_t* plan, int passes, int flags, int use_flags,
                                    double* kernel_ms, double* per_gpu_ms) {
  b2i::NvtxRange range("b2i_ctx_run_resident");
  B2I_CHECK_ARG(c && plan && c->root_src, "load an image first");
  if (passes < 1) passes = 1;
  for (int i = 0; i < c->n; ++i)  // bands sharing one physical GPU must not spin on each other
    for (int j = 0; j < i; ++j)
      if (c->devs[i].id == c->devs[j].id) use_flags = 0;
  b2i_plan_info_t info;
  B2I_TRY(b2i_plan_info(plan, &info));
  B2I_CHECK_ARG(info.cin == c->C && info.cout == c->out_c, "plan/channel mismatch");
  B2I_CHECK_ARG(passes == 1 || info.cin == info.cout, "repeated passes need a chain that keeps the channel count");
  B2I_CHECK_ARG(info.halo <= kGhostRows, "resident passes support a halo of at most %d rows per pass (chain needs %d)",
                kGhostRows, info.halo);
  const int G = kGhostRows, halo = info.halo, cmax = std::max(c->C, c->out_c);
  c->band_pitch = align16((size_t)c->W * cmax);
  for (int i = 0; i < c->n; ++i) {
    const Dev& d = c->devs[i];
    B2I_CHECK_ARG(c->n == 1 || d.y1 - d.y0 >= std::max(halo, 1), "band %d has %d rows, fewer than the halo %d", i,
                  d.y1 - d.y0, halo);
  }
  // allocate band buffers and scatter (band + ghost rows) from the root image
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    B2I_CUDA(cudaSetDevice(d.id));
    const size_t want = c->band_pitch * (size_t)(d.y1 - d.y0 + 2 * G);
    for (int k = 0; k < 2; ++k) {
      size_t have = d.band[k] ? d.band_bytes : 0;
      B2I_TRY(ensure(&d.band[k], &have, want));
    }
    d.band_bytes = std::max(d.band_bytes, want);
    const int ra = std::max(0, d.y0 - G), rb = std::min(c->H, d.y1 + G);
    B2I_TRY(ensure_tmp(c, d, info.n_segments, rb - ra));
    B2I_CUDA(cudaStreamWaitEvent(d.stream, c->ev_src_ready, 0));
    // rows [ra, rb) of the root image land at buffer row (ra - (y0 - G))
    B2I_CUDA(cudaMemcpy2DAsync(d.band[0] + c->band_pitch * (size_t)(ra - (d.y0 - G)), c->band_pitch,
                               c->root_src + c->src_pitch * (size_t)ra, c->src_pitch, (size_t)c->W * c->C,
                               rb - ra, cudaMemcpyDeviceToDevice, d.stream));
    B2I_CUDA(cudaMemsetAsync(d.words, 0, 256, d.stream));
    while ((int)d.ev_pass.size() < passes + 1) {
      cudaEvent_t e;
      B2I_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      d.ev_pass.push_back(e);
    }
    B2I_CUDA(cudaEventRecord(d.ev_pass[0], d.stream));  // "scatter done, flags cleared"
  }
  // everyone's scatter must be complete before anyone pushes halos or flags into a neighbour
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    B2I_CUDA(cudaSetDevice(d.id));
    for (int j = std::max(0, i - 1); j <= std::min(c->n - 1, i + 1); ++j)
      if (j != i) B2I_CUDA(cudaStreamWaitEvent(d.stream, c->devs[j].ev_pass[0], 0));
    B2I_CUDA(cudaEventRecord(d.ev_start, d.stream));
  }
  const uint32_t epoch0 = c->epoch;
  // one (device, pass) launch
  auto launch_pass = [&](int i, int p, b2i_plan_t* plan) -> int {
    const int in = p & 1, outb = in ^ 1;
    Dev& d = c->devs[i];
    if (d.y1 == d.y0) return B2I_OK;
    B2I_CUDA(cudaSetDevice(d.id));
    Dev* up = i > 0 ? &c->devs[i - 1] : nullptr;
    Dev* dn = i + 1 < c->n ? &c->devs[i + 1] : nullptr;
    if (p > 0 && !use_flags) {  // neighbours' previous pass (and its pushes) must be done
      if (up) B2I_CUDA(cudaStreamWaitEvent(d.stream, up->ev_pass[p], 0));
      if (dn) B2I_CUDA(cudaStreamWaitEvent(d.stream, dn->ev_pass[p], 0));
    }
    b2i_launch_t L;
    memset(&L, 0, sizeof L);
    L.W = c->W;
    L.H = c->H;
    L.y0 = d.y0;
    L.y1 = d.y1;
    const int va = std::max(0, d.y0 - G), vb = std::min(c->H, d.y1 + G);
    L.src = {(uint64_t)(d.band[in] + c->band_pitch * (size_t)(va - (d.y0 - G))), (int64_t)c->band_pitch, va, vb};
    L.dst = {(uint64_t)(d.band[outb] + c->band_pitch * (size_t)G), (int64_t)c->band_pitch, d.y0, d.y1};
    const bool more = p + 1 < passes;
    if (more && halo > 0) {
      // my first rows -> the upper neighbour's lower ghost rows (its buffer rows G+h .. G+h+halo)
      L.push_rows = halo;
      if (up)
        L.push_up = {(uint64_t)(up->band[outb] + c->band_pitch * (size_t)(G + (up->y1 - up->y0))),
                     (int64_t)c->band_pitch, d.y0, d.y0 + G};
      if (dn)  // my last rows -> the lower neighbour's upper ghost rows (buffer rows G-halo .. G)
        L.push_dn = {(uint64_t)(dn->band[outb] + c->band_pitch * (size_t)(G - halo)), (int64_t)c->band_pitch,
                     d.y1 - halo, d.y1};
    }
    if (use_flags && halo > 0) {
      L.counters = (uint64_t)d.words;
      L.status = (uint64_t)(d.words + 4);
      if (p > 0) {
        L.wait_epoch = epoch0 + p;
        if (up) L.wait_up_flag = (uint64_t)(d.words + 8);
        if (dn) L.wait_dn_flag = (uint64_t)(d.words + 9);
      }
      if (more) {
        L.signal_epoch = epoch0 + p + 1;
        if (up) L.signal_up_flag = (uint64_t)(up->words + 9);  // I am its lower neighbour
        if (dn) L.signal_dn_flag = (uint64_t)(dn->words + 8);  // I am its upper neighbour
      }
    }
    if (info.n_segments > 1) {
      L.tmp[0] = {(uint64_t)d.tmp[0], (int64_t)c->W * 3, va, vb};
      L.tmp[1] = {(uint64_t)d.tmp[1], (int64_t)c->W * 3, va, vb};
    }
    L.stream = d.stream;
    L.flags = flags;
    B2I_TRY(b2i_cuda_run(plan, &L));
    B2I_CUDA(cudaEventRecord(d.ev_pass[p + 1], d.stream));
    return B2I_OK;
  };
  if (use_flags && c->n > 1) {
    // Device-side flags order the passes, so the devices need no host-side coordination: one
    // host thread per GPU keeps every GPU's launch queue full (a single thread launching for 8
    // GPUs in turn was host-bound: 75 us per pass at N = 8 against 36 us with a thread per GPU).
    std::vector<std::thread> workers;
    std::vector<int> rcs(c->n, B2I_OK);
    std::vector<std::string> errs(c->n);
    for (int i = 0; i < c->n; ++i)
      workers.emplace_back([&, i] {
        b2i_plan_t* mine = nullptr;   // the plan records its last launch: one copy per launching thread
        rcs[i] = b2i_plan_clone(plan, &mine);
        for (int p = 0; p < passes && rcs[i] == B2I_OK; ++p) {
          rcs[i] = launch_pass(i, p, mine);
          if (rcs[i] != B2I_OK) errs[i] = b2i_last_error();
        }
        if (mine) b2i_plan_destroy(mine);
      });
    for (auto& w : workers) w.join();
    for (int i = 0; i < c->n; ++i)
      if (rcs[i] != B2I_OK) {
        b2i::last_error_slot() = "GPU " + std::to_string(i) + ": " + errs[i];
        return rcs[i];
      }
  } else {
    for (int p = 0; p < passes; ++p)
      for (int i = 0; i < c->n; ++i) B2I_TRY(launch_pass(i, p, plan));
  }
  c->epoch = epoch0 + passes + 1;
  // gather: each GPU copies its rows of the last output buffer into the root image
  Dev& root = c->devs[c->root];
  const int last = passes & 1;
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    B2I_CUDA(cudaSetDevice(d.id));
    B2I_CUDA(cudaEventRecord(d.ev_stop, d.stream));
    if (d.y1 > d.y0)
      B2I_CUDA(cudaMemcpy2DAsync(c->root_dst + c->dst_pitch * (size_t)d.y0, c->dst_pitch,
                                 d.band[last] + c->band_pitch * (size_t)G, c->band_pitch,
                                 (size_t)c->W * c->out_c, d.y1 - d.y0, cudaMemcpyDeviceToDevice, d.stream));
    B2I_CUDA(cudaEventRecord(d.ev_pass[0], d.stream));
  }
  B2I_CUDA(cudaSetDevice(root.id));
  for (int i = 0; i < c->n; ++i)
    if (i != c->root) B2I_CUDA(cudaStreamWaitEvent(root.stream, c->devs[i].ev_pass[0], 0));
  double worst = 0;
  int timeout = 0;
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    B2I_CUDA(cudaSetDevice(d.id));
    B2I_CUDA(cudaStreamSynchronize(d.stream));
    float ms = 0;
    B2I_CUDA(cudaEventElapsedTime(&ms, d.ev_start, d.ev_stop));
    if (per_gpu_ms) per_gpu_ms[i] = ms / passes;
    worst = std::max(worst, (double)ms / passes);
    if (use_flags) {
      uint32_t st = 0;
      B2I_CUDA(cudaMemcpy(&st, d.words + 4, 4, cudaMemcpyDeviceToHost));
      if (st) timeout = i + 1;
    }
  }
  if (kernel_ms) *kernel_ms = worst;
  if (timeout) B2I_FAIL(B2I_ERR_TIMEOUT, "GPU %d timed out waiting for a neighbour's halo flag", timeout - 1);
  return B2I_OK;
}

// DIRECT mode: host image in, host image out, every GPU moving its own band (+halo) over its
// own PCIe link. src/dst should be pinned (cudaHostRegister / cudaMallocHost) for the copies to
// be asynchronous. total_ms is the wall time from the first copy to the last byte back.
extern "C" int b2i_ctx_run_direct(b2i_ctx* c, b2i_plan_t* plan, const uint8_t* src, int W, int H, int C,
                                  size_t src_pitch, uint8_t* dst, size_t dst_pitch, int flags,
                                  double* total_ms) {
  b2i::NvtxRange range("b2i_ctx_run_direct");
  B2I_CHECK_ARG(c && plan && src && dst, "bad arguments");
  b2i_plan_info_t info;
  B2I_TRY(b2i_plan_info(plan, &info));
  B2I_CHECK_ARG(info.cin == C, "plan/channel mismatch");
  const int halo = info.halo, out_c = info.cout;
  c->W = W;
  c->H = H;
  c->C = C;
  c->out_c = out_c;
  const size_t bp = align16((size_t)W * 3);
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < c->n; ++i) {
    Dev& d = c->devs[i];
    b2i_band_rows(H, c->n, i, &d.y0, &d.y1);
    if (d.y1 == d.y0) continue;
    B2I_CUDA(cudaSetDevice(d.id));
    const int ra = std::max(0, d.y0 - halo), rb = std::min(H, d.y1 + halo);
    const size_t want = bp * (size_t)(rb - ra);
    for (int k = 0; k < 2; ++k) {
      size_t have = d.band[k] ? d.band_bytes : 0;
      B2I_TRY(ensure(&d.band[k], &have, want));
    }
    d.band_bytes = std::max(d.band_bytes, want);
    B2I_TRY(ensure_tmp(c, d, info.n_segments, rb - ra));
    B2I_CUDA(cudaMemcpy2DAsync(d.band[0], bp, src + src_pitch * (size_t)ra, src_pitch, (size_t)W * C, rb - ra,
                               cudaMemcpyHostToDevice, d.stream));
    b2i_launch_t L;
    memset(&L, 0, sizeof L);
    L.W = W;
    L.H = H;
    L.y0 = d.y0;
    L.y1 = d.y1;
    L.src = {(uint64_t)d.band[0], (int64_t)bp, ra, rb};
    L.dst = {(uint64_t)d.band[1], (int64_t)bp, d.y0, d.y1};
    if (info.n_segments > 1) {
      L.tmp[0] = {(uint64_t)d.tmp[0], (int64_t)W * 3, ra, rb};
      L.tmp[1] = {(uint64_t)d.tmp[1], (int64_t)W * 3, ra, rb};
    }
    L.stream = d.stream;
    L.flags = flags;
    B2I_TRY(b2i_cuda_run(plan, &L));
    B2I_CUDA(cudaMemcpy2DAsync(dst + dst_pitch * (size_t)d.y0, dst_pitch, d.band[1], bp, (size_t)W * out_c,
                               d.y1 - d.y0, cudaMemcpyDeviceToHost, d.stream));
  }
  for (int i = 0; i < c->n; ++i) {
    B2I_CUDA(cudaSetDevice(c->devs[i].id));
    B2I_CUDA(cudaStreamSynchronize(c->devs[i].stream));
  }
  if (total_ms)
    *total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return B2I_OK;
}

// What the CLI calls for --gpus N: load on the root, run, fetch, report.
extern "C" int b2i_ctx_run_cli(int gpus, int root, const b2i_chain_t* chain, const uint8_t* src, int W, int H,
                               int C, uint8_t* dst, int iters, int flags, int resident, double* kernel_ms,
                               double* total_ms, char* report, size_t report_len) {
  B2I_CHECK_ARG(chain && src && dst, "null argument");
  const int out_c = b2i_chain_out_channels(chain, C);
  if (out_c < 0) return out_c;
  const auto t0 = std::chrono::steady_clock::now();
  b2i_ctx* c = nullptr;
  B2I_TRY(b2i_ctx_create(gpus, nullptr, root, &c));
  b2i_plan_t* plan = nullptr;
  int rc = b2i_plan_create(chain, C, &plan);
  std::vector<double> per(gpus, 0.0);
  const bool replicate = resident == 2;  // ROOT mode with the result on every GPU (NVLS multicast)
  if (replicate) resident = 0;
  if (rc == B2I_OK && replicate) rc = b2i_ctx_set_replicated(c, 1);
  if (rc == B2I_OK) rc = b2i_ctx_load_root(c, src, W, H, C, (size_t)W * C, out_c);
  const char* sync_env = getenv("B200IMG_SYNC");
  const int use_flags = sync_env && !strcmp(sync_env, "flags");
  if (rc == B2I_OK) {
    if (resident) {
      rc = b2i_ctx_run_resident(c, plan, iters, flags, use_flags, kernel_ms, per.data());
    } else {
      if (iters > 1) rc = b2i_ctx_run_root(c, plan, 1, flags, nullptr, nullptr);  // warm-up
      if (rc == B2I_OK) rc = b2i_ctx_run_root(c, plan, iters, flags, kernel_ms, per.data());
    }
  }
  if (rc == B2I_OK) rc = b2i_ctx_fetch_root(c, dst, (size_t)W * out_c);
  int replicas_equal = -1;
  if (rc == B2I_OK && replicate) {  // every GPU's copy must be the root's
    std::vector<uint8_t> other((size_t)W * out_c * H);
    replicas_equal = 1;
    for (int g = 0; g < gpus && rc == B2I_OK; ++g) {
      rc = b2i_ctx_fetch_replica(c, g, other.data(), (size_t)W * out_c);
      if (rc == B2I_OK && memcmp(other.data(), dst, other.size()) != 0) replicas_equal = 0;
    }
  }
  if (report && report_len > 0) {
    std::string s = "{\"mode\": \"";
    s += resident ? (use_flags ? "resident/flags" : "resident/events")
                  : replicate ? (b2i_ctx_replicated(c) == 2 ? "root/replicated-nvls" : "root/replicated-plain") : "root";
    if (replicate) s += std::string("\", \"replicas_identical\": \"") + (replicas_equal == 1 ? "yes" : "no");
    s += "\", \"gpus\": " + std::to_string(gpus) + ", \"per_gpu_ms\": [";
    char tmp[64];
    for (int i = 0; i < gpus; ++i) {
      snprintf(tmp, sizeof tmp, "%s%.6f", i ? ", " : "", per[i]);
      s += tmp;
    }
    s += "]}";
    if (s.size() + 1 <= report_len) memcpy(report, s.c_str(), s.size() + 1);
  }
  std::string err = rc == B2I_OK ? "" : b2i_last_error();
  if (plan) b2i_plan_destroy(plan);
  b2i_ctx_destroy(c);
  if (rc != B2I_OK) b2i::last_error_slot() = err;
  if (total_ms)
    *total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return rc;
}
