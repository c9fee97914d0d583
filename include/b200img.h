/*
 * b200img — Blackwell-native distributed image filtering: public C ABI.
 *
 * Everything a caller needs is in this one header. The library
 * (mpi_cuda_imagemanipulation_b200/lib/libb200img.so) is plain C ABI so that the C++ CLI,
 * the Python/ctypes front end and any foreign-language binding share one surface.
 *
 * Images are 8-bit, interleaved, row-major (HWC), 1 or 3 channels, `pitch` bytes per row.
 * That is the in-memory / on-wire layout the reference uses (`kernel.cu:116-119,135,142`).
 * Three-channel data is BGR by default (OpenCV order, `kernel.cu:39-42`); PPM files are RGB and
 * are swapped at the I/O boundary or handled through `b2i_chain_t::order`.
 *
 * All functions return B2I_OK (0) or a negative b2i_status; the failing call leaves a
 * human-readable message retrievable with b2i_last_error() (thread-local).
 */
#ifndef B200IMG_H_
#define B200IMG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2I_MAX_OPS 16
#define B2I_MAX_GPUS 16
#define B2I_VERSION "0.1.0"

typedef enum b2i_status {
  B2I_OK = 0,
  B2I_ERR_ARG = -1,
  B2I_ERR_IO = -2,
  B2I_ERR_CUDA = -3,
  B2I_ERR_UNSUPPORTED = -4,
  B2I_ERR_NO_DEVICE = -5,
  B2I_ERR_TIMEOUT = -6,
  B2I_ERR_NOMEM = -7
} b2i_status;

/* Filter kinds. GRAY_* map 3 channels to 1, EXPAND maps 1 to 3, the rest keep the channel count
 * and act per byte (LUT) or per channel (STENCIL, SOBEL). */
typedef enum b2i_opkind {
  B2I_OP_GRAY_REF = 1, /* floor(.11B)+floor(.59G)+floor(.3R): kernel.cu:39-42 numerics */
  B2I_OP_GRAY_CV = 2,  /* (3735B+19235G+9798R+16384)>>15: cv::cvtColor(BGR2GRAY), kern.cpp:73 */
  B2I_OP_LUT = 3,      /* out = lut[in]: contrast (kernel.cu:49-58, kern.cpp:74), invert, ... */
  B2I_OP_STENCIL = 4,  /* out = clamp((sum w*in + bias) >> shift), k = 3 or 5: kernel.cu:64-94 */
  B2I_OP_SOBEL = 5,    /* out = clamp(|gx|+|gy|), 3x3 Sobel pair */
  B2I_OP_EXPAND = 6    /* gray -> 3 equal channels: cv::cvtColor(GRAY2BGR), kernel.cu:210 */
} b2i_opkind;

/* Border rule of a stencil.
 * REF: the defined subset of the reference's embossKernel guard (kernel.cu:83): pixels with
 *      x in [r+1, W-r-1] and y in [r+1, H-r-1] are filtered, every other pixel passes through.
 *      That includes column W-r and row H-r, which the reference does filter, but from memory that
 *      wraps into the next row or lies past the end of the band (kernel.cu:83): undefined there, so
 *      passed through here.
 * REFLECT101: cv::filter2D default (kern.cpp:75). REPLICATE / ZERO: the usual meanings. */
typedef enum b2i_border {
  B2I_BORDER_REF = 0,
  B2I_BORDER_REFLECT101 = 1,
  B2I_BORDER_REPLICATE = 2,
  B2I_BORDER_ZERO = 3
} b2i_border;

typedef enum b2i_order { B2I_ORDER_BGR = 0, B2I_ORDER_RGB = 1 } b2i_order;

typedef struct b2i_op {
  int32_t kind;   /* b2i_opkind */
  int32_t border; /* b2i_border (STENCIL, SOBEL) */
  int32_t ksize;  /* 3 or 5 (STENCIL) */
  int32_t shift;  /* STENCIL: arithmetic right shift applied to (sum + bias) */
  int32_t bias;   /* STENCIL: added before the shift (rounding term) */
  int32_t reserved;
  int8_t weights[25 + 7]; /* STENCIL: row-major ksize*ksize, weights[row*ksize + col] */
  uint8_t lut[256];       /* LUT */
  char name[24];          /* token this op was parsed from, for reports */
} b2i_op_t;

typedef struct b2i_chain {
  int32_t n_ops;
  int32_t order; /* b2i_order of 3-channel input; only the GRAY_* ops look at it */
  b2i_op_t ops[B2I_MAX_OPS];
} b2i_chain_t;

const char* b2i_last_error(void);
const char* b2i_version(void);

/* ---- chain construction (csrc/core/chain.cpp) ------------------------------------------- */

/* Parses "gray,contrast:3.5,emboss3" style specs. Tokens:
 *   gray | gray_ref | gray_cv | invert | contrast:F | contrast_cv:F | brightness:D | threshold:T
 *   | emboss3 | emboss5 | gauss5 | gauss3 | box3 | sharpen | sobel | expand | bgr
 *   | filter3:w0/w1/.../w8[/shift] | filter5:...
 * A stencil token may carry "@ref", "@reflect101", "@replicate" or "@zero"; otherwise the
 * stencil takes `default_border`, or, when that is negative, its own default
 * (emboss*: REF like kernel.cu, everything else REFLECT101 like OpenCV). */
int b2i_chain_parse(const char* spec, int default_border, int order, b2i_chain_t* out);
/* "kernel_cu": gray_ref,contrast:3.5,emboss3@ref,expand   (kernel.cu:192-195,210)
 * "kern_cpp" : gray_cv,contrast_cv:3,emboss3@reflect101,expand (kern.cpp:73-77) */
int b2i_chain_preset(const char* name, int order, b2i_chain_t* out);
/* Channel count after the chain for `in_channels` input, or a negative status if some op
 * cannot take the channel count it would be given. */
int b2i_chain_out_channels(const b2i_chain_t* chain, int in_channels);
/* Sum of the stencil radii: the number of halo rows a band needs above and below. */
int b2i_chain_halo(const b2i_chain_t* chain);
/* Writes a normalised description ("gray_ref|lut|stencil3@ref|expand") into buf. */
int b2i_chain_describe(const b2i_chain_t* chain, char* buf, size_t buflen);
void b2i_lut_contrast_ref(float factor, uint8_t* lut256); /* kernel.cu:56, fp32 + truncation */
void b2i_lut_contrast_cv(double factor, uint8_t* lut256); /* kern.cpp:74, saturate_cast */

/* ---- band planner (b2i_band_rows in csrc/host_ref/cpu_filters.cpp) ------------------------ */

/* Rows [y0, y1) of band i of n over H rows: y0 = floor(i*H/n). Unlike kernel.cu:117 the
 * remainder rows are distributed, never dropped. */
int b2i_band_rows(int H, int n, int i, int* y0, int* y1);

/* ---- host reference + host "ranks" backend (csrc/host_ref) ------------------------------ */

/* Whole image on the calling thread. dst must hold H rows of out_channels*W bytes. */
int b2i_host_run(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C,
                 size_t src_pitch, uint8_t* dst, size_t dst_pitch);
/* Rows [y0, y1) of the result of the chain applied to the whole image; dst points at row y0.
 * Only source rows [y0-halo, y1+halo) are read. */
int b2i_host_run_rows(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C,
                      size_t src_pitch, int y0, int y1, uint8_t* dst_row_y0, size_t dst_pitch);
/* The reference's scatter -> filter -> gather flow (kernel.cu:137-225) with `ranks` host
 * workers: rank 0 owns src and dst, every worker receives a private copy of its band plus
 * halo rows, filters it and hands its rows back. seams=1 reproduces the reference's missing
 * halo exchange (each band filtered in isolation, kern.cpp:75). */
int b2i_host_run_ranks(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C,
                       size_t src_pitch, uint8_t* dst, size_t dst_pitch, int ranks, int seams,
                       double* elapsed_ms);

/* ---- file I/O (csrc/core/io.cpp) --------------------------------------------------------- */

/* Reads PPM (P6, RGB on disk) or PGM (P5). 3-channel data is returned in `order`.
 * *data is malloc'ed; free with b2i_free. */
int b2i_read_pnm(const char* path, int order, uint8_t** data, int* W, int* H, int* C);
int b2i_write_pnm(const char* path, int order, const uint8_t* data, int W, int H, int C,
                  size_t pitch);
/* JPEG, the reference's own file format (cv::imread / cv::imwrite on rank 0: kernel.cu:110,236;
 * kern.cpp:33,92), through nvJPEG on the current CUDA device (csrc/core/jpeg.cpp; the library is
 * loaded on first use, B2I_ERR_UNSUPPORTED without it). A grayscale file reads as 1 channel, like
 * IMREAD_UNCHANGED. Writing is baseline 4:4:4 at `quality` (1..100, 0 = 95): lossy, so results
 * read back close to, not equal to, what was written. File I/O is outside every timed region. */
int b2i_jpeg_available(void);
int b2i_read_jpeg(const char* path, int order, uint8_t** data, int* W, int* H, int* C);
int b2i_write_jpeg(const char* path, int order, const uint8_t* data, int W, int H, int C, size_t pitch,
                   int quality);
/* Any supported file: JPEG by content, else PPM / PGM; writing picks the format by extension
 * (.jpg / .jpeg -> JPEG, anything else -> PPM / PGM). */
int b2i_read_image(const char* path, int order, uint8_t** data, int* W, int* H, int* C);
int b2i_write_image(const char* path, int order, const uint8_t* data, int W, int H, int C, size_t pitch,
                    int quality);
/* Headerless interleaved bytes, exactly the reference's scatter payload (kernel.cu:135-137). */
int b2i_read_raw(const char* path, int W, int H, int C, uint8_t** data);
int b2i_write_raw(const char* path, const uint8_t* data, int W, int H, int C, size_t pitch);
void b2i_free(void* p);
/* Deterministic synthetic test card (gradients + edges + hashed noise), any size. */
int b2i_synth_image(uint8_t* dst, int W, int H, int C, size_t pitch, uint32_t seed);

/* ---- CUDA path: plans and band launches (csrc/runtime/plan.cpp, csrc/kernels) ------------- */

/* A plan is a chain compiled for a given input channel count: LUTs folded, ops grouped into
 * segments (one kernel launch each), every stencil lowered to packed dp4a weights. */
typedef struct b2i_plan b2i_plan_t;

/* Rows [y_begin, y_end) of an image, addressable from the device that launches: local memory,
 * a peer GPU's memory (cudaDeviceEnablePeerAccess) or an IPC mapping. ptr is row y_begin. */
typedef struct b2i_view {
  uint64_t ptr;
  int64_t pitch;
  int32_t y_begin, y_end;
} b2i_view_t;

#define B2I_LAUNCH_NO_TMA 1   /* force the plain ld/st tile loaders (debug, unaligned images) */
#define B2I_LAUNCH_NAIVE 2    /* reference-style one-launch-per-op kernels (baseline / oracle) */
#define B2I_LAUNCH_DST_MCAST 4 /* dst is an NVLS multicast address (b2i_mcast_ptr): result rows leave by
                                  plain stores from registers / shared memory, which the switch
                                  replicates into every GPU's copy; TMA stores are not used on it */
#define B2I_LAUNCH_STRIP_FLAGS 8 /* b2i_cuda_run_resident_fused: wait_*_flag / signal_*_flag are arrays of one u32
                                    per 240-pixel strip (B2I_MAX_STRIPS each) instead of one flag per side: a
                                    boundary task waits for the three strips above / below it and releases its own,
                                    so no pass-wide count stands between a pushed halo row and its reader */
#define B2I_MAX_STRIPS 256

/* One band launch: produce rows [y0, y1) of chain(image) into dst, reading src.
 * src must cover rows [y0-halo, y1+halo) clipped to the image (halo = b2i_plan_halo); when src
 * is another GPU's image this IS the scatter, when dst is the root GPU's image this IS the
 * gather. Nothing is allocated and nothing synchronises with the host. */
typedef struct b2i_launch {
  int32_t W, H;          /* whole-image geometry */
  int32_t y0, y1;        /* rows to produce */
  b2i_view_t src, dst;
  b2i_view_t tmp[2];     /* scratch images for plans with more than one segment (3 bytes/pixel,
                            rows [y0-halo, y1+halo) clipped); ptr 0 when the plan has one segment */
  /* multi-GPU hooks, all optional (0) */
  b2i_view_t push_up, push_dn; /* neighbours' ghost rows: rows [y0,y0+push_rows) resp.
                                  [y1-push_rows,y1) of the result are stored there as well */
  int32_t push_rows;
  uint32_t wait_epoch, signal_epoch;
  uint64_t wait_up_flag, wait_dn_flag;     /* u32 flags in local memory */
  uint64_t signal_up_flag, signal_dn_flag; /* u32 flags in the neighbours' memory */
  uint64_t done_flag;                      /* u32 flag (root GPU): set to signal_epoch when every row is stored
                                              (by the last warp of a sweep kernel, else by a one-thread kernel) */
  uint64_t counters;                       /* 16 zeroed bytes of local device memory (flag bookkeeping); with
                                              B2I_LAUNCH_NAIVE or images under 17 pixels: 256*B2I_MAX_OPS bytes
                                              that receive the LUTs of the per-op kernels */
  uint64_t status;                         /* u32 in local device memory, nonzero = wait timed out */
  uint64_t timeout_ns;                     /* 0 = 2 s */
  void* stream;                            /* cudaStream_t */
  int32_t tile_h;                          /* 0 = choose */
  int32_t flags;                           /* B2I_LAUNCH_* */
  /* The root's side of done_flag: u32 flags [0..B2I_MAX_GPUS) in local memory, one per GPU; the
   * launch returns (in stream order) only after every flag selected by the mask has reached
   * signal_epoch. The sweep kernels do this wait in their last warp, other kernels are followed
   * by a one-thread wait kernel. 0 = off. */
  uint64_t wait_done_flags;
  uint32_t wait_done_mask, reserved0;
} b2i_launch_t;

/* A plan may keep two cuts of its chain: n_segments and seg_* describe the one with
 * more launches, which is what callers size tmp images by. last_tile_h > 0: tile height of the
 * tile-chain kernel; < 0: minus the rows per segment of a sweep kernel; 0: streaming pointwise. */
typedef struct b2i_plan_info {
  int32_t cin, cout, halo, n_segments;
  int32_t seg_pointwise[B2I_MAX_OPS], seg_halo[B2I_MAX_OPS], seg_phases[B2I_MAX_OPS];
  int32_t seg_cout[B2I_MAX_OPS];
  /* filled by the last b2i_cuda_run on this plan (not thread-safe, for reports) */
  int32_t last_tile_h, last_smem_bytes, last_grid, last_blocks_per_sm, last_tma_load,
      last_tma_store, last_launches;
} b2i_plan_info_t;

int b2i_plan_create(const b2i_chain_t* chain, int in_channels, b2i_plan_t** out);
void b2i_plan_destroy(b2i_plan_t* plan);
/* A copy for another host thread: b2i_cuda_run records the figures of its last launch in the plan
 * (b2i_plan_info last_*), so one plan is launched from one thread at a time. */
int b2i_plan_clone(const b2i_plan_t* plan, b2i_plan_t** out);
int b2i_plan_halo(const b2i_plan_t* plan);
int b2i_plan_info(const b2i_plan_t* plan, b2i_plan_info_t* info);
int b2i_cuda_run(b2i_plan_t* plan, const b2i_launch_t* launch);
/* Whole image, device pointers, current device, no bands: the call the torch front end makes.
 * tmp0/tmp1: W*H*3 bytes each, needed for multi-segment plans and B2I_LAUNCH_NAIVE (with
 * lut_scratch: 256*B2I_MAX_OPS bytes); may be 0 otherwise. */
int b2i_cuda_run_image(b2i_plan_t* plan, uint64_t src, int64_t src_pitch, uint64_t dst,
                       int64_t dst_pitch, int W, int H, uint64_t tmp0, uint64_t tmp1,
                       uint64_t lut_scratch, void* stream, int flags);
/* No counterpart in the reference (it cudaMallocs every buffer, kernel.cu:156,173).
 * Same result as b2i_cuda_run_image, for callers that cannot allocate device memory at the point
 * of the call: code running inside another framework's execution callback where the host
 * framework owns the allocator, processes whose device memory is already fully reserved by a
 * caching allocator, latency-bound loops that cannot take cudaMalloc's device-wide
 * synchronisation. The two scratch images and the table block that b2i_cuda_run_image takes
 * from the caller come instead from a fixed array inside the library's device image; the caller
 * passes only the input and the output. This function calls no cudaMalloc, cudaFree or other
 * allocating routine.
 *  - Limit: W*H <= B2I_NOALLOC_MAX_PIXELS (2 Mi pixels, e.g. 1920x1080); larger images are
 *    rejected with B2I_ERR_UNSUPPORTED and a message naming the limit. The array is
 *    B2I_NOALLOC_SCRATCH_BYTES (about 12 MiB) and every GPU that loads the library holds it
 *    until the process exits.
 *  - The first call on each device loads the module that holds the array (and performs the
 *    kernels' one-time attribute set-up); make that first call before any stream capture. The
 *    array's device address is looked up on every call (no allocation), so a cudaDeviceReset
 *    between calls is harmless.
 *  - All calls on a device share the one array. A call never depends on what an earlier call
 *    left in it (the part that must start clean is cleared on the stream, the rest is written
 *    before it is read), but two calls that overlap on the GPU - different streams, or
 *    different threads - race on it: serialise them, e.g. by using one stream. */
#define B2I_NOALLOC_MAX_PIXELS (2 * 1024 * 1024)
#define B2I_NOALLOC_SCRATCH_BYTES (2 * 3 * B2I_NOALLOC_MAX_PIXELS + 256 * B2I_MAX_OPS)
int b2i_cuda_run_image_noalloc(b2i_plan_t* plan, uint64_t src, int64_t src_pitch, uint64_t dst,
                               int64_t dst_pitch, int W, int H, void* stream, int flags);
int b2i_cuda_device_count(void);
/* Host image in, host image out on one GPU: upload, `iters` runs (one warm-up first when
 * iters > 1), download. kernel_ms is the device-event time of one run, total_ms the wall time
 * of the whole call. The reference's kernel.cu:147-206 sequence in one call. */
int b2i_cuda_run_host_image(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C,
                            size_t src_pitch, uint8_t* dst, size_t dst_pitch, int device,
                            int iters, int flags, double* kernel_ms, double* total_ms);

/* ---- multi-GPU runtime (csrc/runtime/context.cpp): what stands where the reference has MPI -- */

/* JSON report: devices (name, SMs, memory, NVLS multicast support) and the P2P matrix
 * (access, native atomics, performance rank). MPI_Comm_size/rank become this discovery. */
int b2i_topology_json(char* buf, size_t len);

typedef struct b2i_ctx b2i_ctx_t;
/* One process drives n_gpus devices (devices == NULL: 0..n-1, wrapping onto the devices that
 * exist, so multi-band runs can be tested on one GPU). Peer access is enabled all-to-all. */
int b2i_ctx_create(int n_gpus, const int* devices, int root, b2i_ctx_t** out);
void b2i_ctx_destroy(b2i_ctx_t* ctx);
int b2i_ctx_gpus(const b2i_ctx_t* ctx);
/* Host image -> the root GPU's source image (kernel.cu:110,163). */
int b2i_ctx_load_root(b2i_ctx_t* ctx, const uint8_t* src, int W, int H, int C, size_t pitch,
                      int out_channels);
/* ROOT mode band plan: the root GPU keeps `share` of the rows, the others split the rest (0 =
 * equal bands). The root works out of its own HBM, every other GPU through the root's one NVLink
 * port, so equal bands leave the root idle most of the step. Takes effect at the next load. */
int b2i_ctx_set_root_share(b2i_ctx_t* ctx, double share);
/* Root GPU's output image -> host (kernel.cu:202,223). Synchronises. */
int b2i_ctx_fetch_root(b2i_ctx_t* ctx, uint8_t* dst, size_t pitch);
/* ROOT mode: scatter and gather fused into the filter kernels (peer loads from / peer stores
 * to the root GPU). kernel_ms: slowest GPU, device events, per iteration. */
int b2i_ctx_run_root(b2i_ctx_t* ctx, b2i_plan_t* plan, int iters, int flags, double* kernel_ms,
                     double* per_gpu_ms);
/* RESIDENT mode: bands stay on their GPUs for `passes` applications of the chain; halos are
 * pushed into the neighbours' ghost rows by the filter kernel itself. use_flags: device-side
 * release/acquire flags instead of cross-device stream events between passes. */
int b2i_ctx_run_resident(b2i_ctx_t* ctx, b2i_plan_t* plan, int passes, int flags, int use_flags,
                         double* kernel_ms, double* per_gpu_ms);
/* DIRECT mode: host image in, host image out; every GPU moves its own band over its own PCIe
 * link (pin src/dst for asynchronous copies). */
int b2i_ctx_run_direct(b2i_ctx_t* ctx, b2i_plan_t* plan, const uint8_t* src, int W, int H, int C,
                       size_t src_pitch, uint8_t* dst, size_t dst_pitch, int flags,
                       double* total_ms);
/* The CLI's multi-GPU run. resident: 0 = ROOT mode, 1 = RESIDENT mode (iters = passes), 2 = ROOT
 * mode with replicated output (every GPU's copy is compared with the root's). */
int b2i_ctx_run_cli(int gpus, int root, const b2i_chain_t* chain, const uint8_t* src, int W, int H,
                    int C, uint8_t* dst, int iters, int flags, int resident, double* kernel_ms,
                    double* total_ms, char* report, size_t report_len);

/* Replicated output: ROOT mode ends with the whole result on EVERY GPU instead of on the root
 * only. The last kernel of each GPU stores its band once, to an NVLS multicast address, and the
 * NVSwitch writes all copies (one plain buffer stands in when the context has one physical GPU).
 * Takes effect at the next b2i_ctx_load_root. Returns B2I_ERR_UNSUPPORTED without multicast. */
int b2i_ctx_set_replicated(b2i_ctx_t* ctx, int on);
/* 0 = off, 1 = on with a plain buffer (one physical GPU), 2 = on through an NVLS multicast object. */
int b2i_ctx_replicated(const b2i_ctx_t* ctx);
/* GPU `gpu`'s copy of the replicated output -> host. Synchronises. */
int b2i_ctx_fetch_replica(b2i_ctx_t* ctx, int gpu, uint8_t* dst, size_t pitch);

/* ---- NVLS multicast images (csrc/runtime/mcast.cpp); no counterpart in the reference ---------- */

/* `bytes` on each of n distinct devices (cuMemCreate) bound to one multicast object
 * (cuMulticastCreate / cuMulticastBindMem): a store to b2i_mcast_ptr lands in every replica. */
typedef struct b2i_mcast b2i_mcast_t;
int b2i_mcast_create(int n, const int* devices, size_t bytes, b2i_mcast_t** out);
void b2i_mcast_destroy(b2i_mcast_t* m);
uint64_t b2i_mcast_ptr(const b2i_mcast_t* m);             /* the multicast address (write-only use) */
uint64_t b2i_mcast_replica(const b2i_mcast_t* m, int i);  /* device i's copy, an ordinary pointer */
size_t b2i_mcast_bytes(const b2i_mcast_t* m);
int b2i_mcast_is_multicast(const b2i_mcast_t* m);          /* 0: n == 1, a plain allocation */

/* ---- multi-process plumbing (csrc/runtime/ipc.cu): one process per GPU, as under torchrun ---- */

/* Device memory other processes can map; the 64-byte handle goes over any host channel
 * (torch.distributed here) and is opened by the peers. Image bytes never take that channel. */
int b2i_ipc_alloc(size_t bytes, uint64_t* ptr, uint8_t* handle64);
int b2i_ipc_open(const uint8_t* handle64, uint64_t* ptr);
int b2i_ipc_close(uint64_t ptr);
int b2i_dev_free(uint64_t ptr);
/* Pins a host range (e.g. a POSIX shared-memory mapping every rank sees). */
int b2i_host_register(void* p, size_t bytes);
int b2i_host_unregister(void* p);
/* Stream-ordered epoch flags, the device-side replacement of MPI_Barrier (kernel.cu:127,133):
 * wait until n flags (device addresses, 0 = skip) reach `epoch` / set n flags to `epoch`. */
int b2i_cuda_wait_flags(const uint64_t* flags, int n, uint32_t epoch, uint64_t status,
                        uint64_t timeout_ns, void* stream);
int b2i_cuda_set_flags(const uint64_t* flags, int n, uint32_t epoch, void* stream);
/* K steps of one band launch from C (one kernel launch of host cost per step), epochs rising
 * from epoch_base+1; the root rank passes the other ranks' done flags to wait on each step,
 * every other rank the address of its own done flag on the root (signal_flag). */
int b2i_cuda_run_steps(b2i_plan_t* plan, const b2i_launch_t* launch, int steps, uint32_t epoch_base,
                       const uint64_t* wait_flags, int n_wait, uint64_t signal_flag);
/* `passes` applications of a chain on a resident band, ping-ponging between the two launch
 * descriptions, halo flags waited/signalled with epochs epoch_base+p / epoch_base+p+1. */
int b2i_cuda_run_resident_steps(b2i_plan_t* plan, const b2i_launch_t* even, const b2i_launch_t* odd,
                                int passes, uint32_t epoch_base);
/* The same with a cycle of `period` launch descriptions, pass p running descs[p % period]: halo rows are
 * exchanged every few passes only, the passes between compute a shrinking margin of ghost rows themselves
 * (parallel/dist.py, run_resident(cycle=...)). */
int b2i_cuda_run_resident_cycle(b2i_plan_t* plan, const b2i_launch_t* descs, int period, int passes,
                                uint32_t epoch_base);
/* The same passes in ONE cooperative kernel launch (one separable 3-channel stencil, e.g. gauss5):
 * the kernel walks the passes itself, tasks wait for the tasks around them through `task_flags`
 * (n_task_flags u32 of local device memory, zeroed once when allocated and only ever growing:
 * at least tasks + 8, 16384 is plenty) and for the neighbour GPUs through the epoch flags.
 * B2I_ERR_UNSUPPORTED if the plan or the buffers do not fit (use b2i_cuda_run_resident_steps).
 * The halo exchange the reference never had (kernel.cu:195) without its MPI_Barriers (kernel.cu:127,133). */
int b2i_cuda_run_resident_fused(b2i_plan_t* plan, const b2i_launch_t* even, const b2i_launch_t* odd, int passes,
                                uint32_t epoch_base, uint64_t task_flags, int n_task_flags);
int b2i_cuda_memcpy2d(uint64_t dst, size_t dpitch, uint64_t src, size_t spitch, size_t width,
                      size_t height, int kind /*0 H2D, 1 D2H, 2 D2D*/, void* stream);
/* Experiment (kernels/cluster_probe.cu): neighbouring stencil tiles of T rows load the 2R rows they
 * share twice (one TMA box per CTA) or once (a two-CTA cluster, the shared rows by one multicast TMA load
 * into both CTAs' shared memory). Runs both on a W3-byte x H-row image it allocates on the current GPU;
 * time per launch of each and the number of wrong output bytes (0). */
int b2i_cuda_cluster_probe(int W3, int H, int T, int R, int iters, float* us_plain, float* us_cluster,
                           long long* mismatches);

#ifdef __cplusplus
}
#endif
#endif /* B200IMG_H_ */
