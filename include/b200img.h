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

/* ---- band planner (csrc/runtime/band_plan.cpp) ------------------------------------------ */

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
/* Headerless interleaved bytes, exactly the reference's scatter payload (kernel.cu:135-137). */
int b2i_read_raw(const char* path, int W, int H, int C, uint8_t** data);
int b2i_write_raw(const char* path, const uint8_t* data, int W, int H, int C, size_t pitch);
void b2i_free(void* p);
/* Deterministic synthetic test card (gradients + edges + hashed noise), any size. */
int b2i_synth_image(uint8_t* dst, int W, int H, int C, size_t pitch, uint32_t seed);

#ifdef __cplusplus
}
#endif
#endif /* B200IMG_H_ */
