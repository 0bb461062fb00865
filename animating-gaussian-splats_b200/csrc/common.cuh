// common.cuh -- shared device helpers for libgsb200 (sm_100a only).
//
// Bit-exactness note: radii / tiles_touched / depth key bits / means2D must match the reference
// rasterizer bit for bit (SURVEY.md section 7 "hard parts" #1).  The reference is compiled without
// fast-math, and nvcc's FMA contraction decides the rounding sequence.  Every FP32 operation on
// such a path is therefore written with an explicit round-to-nearest intrinsic below, which nvcc
// neither contracts nor reassociates.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/gsb200.h"

#define GSB_BLOCK_X 16
#define GSB_BLOCK_Y 16
#define GSB_BLOCK_SIZE (GSB_BLOCK_X * GSB_BLOCK_Y)
#define GSB_NCH 3
#define GSB_MAX_CAMS 320   // cameras per batch: the backward keeps [n_cams][32] matrix floats in 40 KB of shared memory

namespace gsb {

__device__ __forceinline__ float FMA(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float MUL(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float ADD(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float SUB(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float DIV(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float SQRT(float a) { return __fsqrt_rn(a); }

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// Error plumbing shared by the host side of every translation unit.
void set_last_error(const char* what, cudaError_t e);
#define GSB_CUDA_CHECK(expr)                                  \
  do {                                                        \
    cudaError_t _e = (expr);                                  \
    if (_e != cudaSuccess) {                                  \
      gsb::set_last_error(#expr, _e);                         \
      return GSB_ECUDA;                                       \
    }                                                         \
  } while (0)
#define GSB_LAUNCH_CHECK() GSB_CUDA_CHECK(cudaGetLastError())

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Host launchers implemented across the .cu files (all asynchronous on `st`).
struct CamParams {  // per-entry scalars that live in a small device table
  float tanfovx, tanfovy, focal_x, focal_y;
  int row0, row1;          // tile rows the entry renders (tile-band sharding); whole camera: 0, tiles_y
  int split_slot;          // >= 0: the camera's rows are shared with another rank (gsb_raster_args.entries)
  int count_vis;           // this entry counts the camera in the visibility / radius statistics
};
static_assert(sizeof(CamParams) == 32, "kernels copy a CamParams as eight 32-bit words / read its first 16 bytes");

int launch_preprocess(const gsb_raster_args& a, const CamParams* cams_dev, const gsb_raster_ctx& c,
                      int32_t* radii_out, uint32_t* gauss_keys_in, int* err_flag_dev, cudaStream_t st);
int launch_cam_params(int n_cams, int W, int H, const float* tanfov_dev, const int32_t* entries, CamParams* cams_dev,
                      cudaStream_t st);
int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st);
// flat != nullptr: the exchange form (gsb_raster_backward_flat): rows of GSB_FLAT_WIDTH floats instead of the
// separate tables; dL_dopacity / dL_dcolor are then the blend's camera-summed tables (inputs), flat_mode 0 | 1
int launch_preprocess_bwd(const gsb_raster_args& a, const CamParams* cams_dev, const gsb_raster_ctx& c,
                          float* dL_dmean2D, const float* dL_dconic, float* dL_dmean3D, float* dL_dcov3D,
                          float* dL_dscale, float* dL_drot, cudaStream_t st, float* flat = nullptr, int flat_mode = 0,
                          const float* dL_dopacity = nullptr, const float* dL_dcolor = nullptr,
                          float* dL_dmean2D_out = nullptr, int flat_stride = GSB_FLAT_WIDTH,
                          const float* dL_dcolor2 = nullptr, const float* dL_dmean2D_first = nullptr);
int launch_flat_apply_stats(int P, const float* flat, int flat_stride, float* grad_accum, float* vis_count,
                            float* max_2d_radii, cudaStream_t st);
int launch_allreduce_flat(int rank, int world, void* const* peers, size_t payload_bytes, int64_t n_float4, int row_float4,
                          int max_word, int max_comp, cudaStream_t st);
size_t comm_bytes(size_t payload_bytes);

size_t scan_scratch_bytes(int64_t n);
int launch_inclusive_scan_u32(int64_t n, const uint32_t* in, uint32_t* out, void* scratch, cudaStream_t st,
                              const uint32_t* perm = nullptr, int perm_block = 0);
int launch_duplicate_with_keys(const gsb_raster_ctx& c, const CamParams* cams_dev, cudaStream_t st);
size_t sort_temp_bytes(int64_t n);
int launch_sort_pairs(int64_t n, const uint64_t* kin, uint64_t* kout, const uint32_t* vin, uint32_t* vout,
                      int end_bit, void* temp, cudaStream_t st);

// segsort.cu: segmented wide-digit onesweep (the product path's sorts) and tile ranges
size_t seg_sort_temp_bytes(int64_t n_upper, int n_seg, int key_bytes, int log_nb);
int launch_seg_setup(int n_seg, int mode, int len, int key_bytes, const uint32_t* offsets, uint32_t* seg_start,
                     uint32_t* seg_tile0, cudaStream_t st, int64_t capacity = -1, int* status = nullptr);
int launch_seg_sort_u16(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                        const uint16_t* kin, uint16_t* k_out, uint16_t* k_alt, const uint32_t* vin, uint32_t* v_out,
                        uint32_t* v_alt, int end_bit, int log_nb, bool want_keys, void* temp, cudaStream_t st,
                        const uint32_t** hist_out, int val_bits = 32);
int launch_seg_sort_u32(int64_t n_upper, int n_seg, const uint32_t* seg_start, const uint32_t* seg_tile0,
                        const uint32_t* kin, uint32_t* k_out, uint32_t* k_alt, const uint32_t* vin, uint32_t* v_out,
                        uint32_t* v_alt, int end_bit, int log_nb, bool want_keys, void* temp, cudaStream_t st,
                        const uint32_t** hist_out);
int launch_ranges_from_hist(int n_cams, int tiles_per_cam, int log_nb, const uint32_t* seg_start, const uint32_t* hist,
                            uint32_t* ranges, cudaStream_t st);
int launch_tile_ranges_seg(int64_t R, int n_cams, int tiles_per_cam, const uint32_t* seg_start, const uint16_t* keys,
                           uint32_t* ranges, cudaStream_t st);

int launch_tile_order(int tiles_total, const uint32_t* ranges, uint32_t* order, cudaStream_t st);
int launch_render_fwd(const gsb_raster_args& a, const gsb_raster_ctx& c, float* out_color, float* out_depth,
                      cudaStream_t st);
int launch_count_pairs(const gsb_raster_args& a, const gsb_raster_ctx& c, unsigned long long* counts, cudaStream_t st);
int launch_render_bwd(const gsb_raster_args& a, const gsb_raster_ctx& c, const float* dL_dpix, float* dL_dmean2D,
                      float* dL_dconic, float* dL_dopacity, float* dL_dcolor, cudaStream_t st,
                      const float* dL_dpix2 = nullptr, float* dL_dcolor2 = nullptr, float* dL_dmean2D_first = nullptr);

}  // namespace gsb
