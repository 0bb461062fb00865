// preprocess.cu -- per-Gaussian kernels: forward preprocess, markVisible, fused backward preprocess.
//
// Forward replaces preprocessCUDA + computeCov3D + computeCov2D + in_frustum + getRect + ndc2Pix
// (cr/forward.cu:74-256, cr/auxiliary.h:41-56,139-164).  One thread per (camera, Gaussian); the
// AoS [P][3] tables are staged through shared memory with 16-byte loads, the two 4x4 matrices of the
// camera live in shared memory (the reference re-reads them from global in every thread).
// The arithmetic sequence is pinned with explicit round-to-nearest intrinsics so that radii, tile
// counts, depths and pixel centres equal the reference's bit for bit (see common.cuh).
//
// Backward replaces computeCov2DCUDA + preprocessCUDA(bwd) + computeCov3D(bwd)
// (cr/backward.cu:144-396) as ONE kernel, one thread per Gaussian looping over the cameras of the
// batch, so the per-camera gradients are summed in registers and each output row is written once.
#include "common.cuh"

namespace gsb {

// Stage a block's 256 x 3 floats of an AoS [P][3] table into shared memory with float4 loads.
__device__ __forceinline__ void stage3(const float* __restrict__ src, int P, int block_first, float* smem) {
  const long long first = 3ll * block_first;
  const long long total = 3ll * P;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src + first) & 15u) == 0);
  if (threadIdx.x < 192) {
    const long long e = first + 4ll * threadIdx.x;
    if (aligned && e + 3 < total) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(src + e));
      reinterpret_cast<float4*>(smem)[threadIdx.x] = v;
    } else {
#pragma unroll
      for (int k = 0; k < 4; k++) smem[4 * threadIdx.x + k] = (e + k < total) ? __ldg(src + e + k) : 0.f;
    }
  }
}

// x' = (fma(z, m8, fma(x, m0, y*m4))) + m12   -- transformPoint4x3/4x4 as contracted by nvcc in the reference
__device__ __forceinline__ float xform_row(const float* m, int r, float x, float y, float z) {
  return ADD(m[12 + r], FMA(z, m[8 + r], FMA(x, m[r], MUL(y, m[4 + r]))));
}

// computeCov3D (cr/forward.cu:118-152): Sigma = (S R^T)^T (S R^T), six unique entries.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float4 q, float* c) {
  const float qr = q.x, qx = q.y, qy = q.z, qz = q.w;
  const float yy = MUL(qy, qy);
  const float r22 = SUB(1.0f, ADD(FMA(qx, qx, yy), FMA(qx, qx, yy)));
  const float msz = MUL(mod, sz), msx = MUL(mod, sx), msy = MUL(mod, sy);
  const float xz = MUL(qx, qz);
  const float a7t = FMA(qy, -qr, xz);
  const float a7 = ADD(a7t, a7t);   // 2(xz - ry)
  const float xr = MUL(qx, qr);
  const float a10t = FMA(qz, qy, xr);
  const float a10 = ADD(a10t, a10t);  // 2(yz + rx)
  const float z10 = MUL(0.0f, a10);
  const float m13 = FMA(r22, msz, FMA(0.0f, a7, z10));
  const float m16 = FMA(0.0f, r22, FMA(a7, msx, z10));
  const float m20 = FMA(0.0f, r22, FMA(0.0f, a7, MUL(a10, msy)));
  const float c5 = FMA(m13, m13, FMA(m16, m16, MUL(m20, m20)));
  const float a25t = FMA(qr, qy, xz);
  const float a25 = ADD(a25t, a25t);  // 2(xz + ry)
  const float zr = MUL(qz, qr);
  const float a28t = FMA(qx, qy, -zr);
  const float a28 = ADD(a28t, a28t);  // 2(xy - rz)
  const float zz = MUL(qz, qz);
  const float s30 = ADD(yy, zz);
  const float r00 = SUB(1.0f, ADD(s30, s30));
  const float z00 = MUL(0.0f, r00);
  const float m35 = FMA(a25, msz, FMA(0.0f, a28, z00));
  const float m38 = FMA(0.0f, a25, FMA(0.0f, a28, MUL(msx, r00)));
  const float m40 = FMA(0.0f, a25, FMA(msy, a28, z00));
  const float c0 = FMA(m35, m35, FMA(m38, m38, MUL(m40, m40)));
  const float a45t = FMA(qz, qy, -xr);
  const float a45 = ADD(a45t, a45t);  // 2(yz - rx)
  const float a47t = FMA(qx, qy, zr);
  const float a47 = ADD(a47t, a47t);  // 2(xy + rz)
  const float s48 = FMA(qx, qx, zz);
  const float r11 = SUB(1.0f, ADD(s48, s48));
  const float z11 = MUL(0.0f, r11);
  const float m53 = FMA(a45, msz, FMA(0.0f, a47, z11));
  const float m55 = FMA(0.0f, a45, FMA(msx, a47, z11));
  const float m58 = FMA(0.0f, a45, FMA(0.0f, a47, MUL(msy, r11)));
  const float c1 = FMA(m53, m35, FMA(m38, m55, MUL(m40, m58)));
  const float c2 = FMA(m13, m35, FMA(m16, m38, MUL(m40, m20)));
  const float c3 = FMA(m53, m53, FMA(m55, m55, MUL(m58, m58)));
  const float c4 = FMA(m13, m53, FMA(m16, m55, MUL(m58, m20)));
  c[0] = c0; c[1] = c1; c[2] = c2; c[3] = c3; c[4] = c4; c[5] = c5;
}

// getRect (cr/auxiliary.h:46-56): the int radius goes back to float, /16 is folded to *0.0625,
// float -> int conversions truncate (cvt.rzi), clamps are unsigned-after-max(0).
__device__ __forceinline__ void tile_rect(float px, float py, int rad, int gx, int gy, uint32_t& x0, uint32_t& y0,
                                          uint32_t& x1, uint32_t& y1) {
  const float rf = (float)rad;
  const int ix0 = __float2int_rz(MUL(SUB(px, rf), 0.0625f));
  const int iy0 = __float2int_rz(MUL(SUB(py, rf), 0.0625f));
  const int ix1 = __float2int_rz(MUL(ADD(ADD(ADD(px, rf), 16.0f), -1.0f), 0.0625f));
  const int iy1 = __float2int_rz(MUL(ADD(ADD(ADD(py, rf), 16.0f), -1.0f), 0.0625f));
  x0 = min((uint32_t)max(ix0, 0), (uint32_t)gx);
  y0 = min((uint32_t)max(iy0, 0), (uint32_t)gy);
  x1 = min((uint32_t)max(ix1, 0), (uint32_t)gx);
  y1 = min((uint32_t)max(iy1, 0), (uint32_t)gy);
}

__global__ void __launch_bounds__(256)
preprocess_fwd_kernel(int P, int W, int H, int gx, int gy, float mod, int prefiltered,
                      const float* __restrict__ means3D, const float* __restrict__ scales,
                      const float* __restrict__ rotations, const float* __restrict__ opacities,
                      const float* __restrict__ cov3D_precomp, const float* __restrict__ viewmatrix,
                      const float* __restrict__ projmatrix, const CamParams* __restrict__ cams,
                      float* __restrict__ depths, int32_t* __restrict__ radii_state, int32_t* __restrict__ radii_out,
                      float2* __restrict__ means2D, float* __restrict__ cov3D, float4* __restrict__ conic_opacity,
                      uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ gauss_keys,
                      int* __restrict__ err_flag) {
  __shared__ __align__(16) float s_mean[768];
  __shared__ __align__(16) float s_scale[768];
  __shared__ float s_V[16], s_Pm[16];
  __shared__ CamParams s_cp;   // the block's camera: one load per block instead of 32 bytes per thread
  const int cam = blockIdx.y;
  const int block_first = blockIdx.x * 256;
  const int idx = block_first + threadIdx.x;
  if (threadIdx.x < 16) s_V[threadIdx.x] = viewmatrix[16 * cam + threadIdx.x];
  else if (threadIdx.x < 32) s_Pm[threadIdx.x - 16] = projmatrix[16 * cam + threadIdx.x - 16];
  else if (threadIdx.x < 40)
    reinterpret_cast<int*>(&s_cp)[threadIdx.x - 32] = reinterpret_cast<const int*>(cams + cam)[threadIdx.x - 32];
  stage3(means3D, P, block_first, s_mean);
  if (scales) stage3(scales, P, block_first, s_scale);
  __syncthreads();
  if (idx >= P) return;
  const size_t o = (size_t)cam * P + idx;

  // Culled Gaussians: radii = tiles_touched = 0, nothing else written (cr/forward.cu:179-194).
  int rad_w = 0;
  uint32_t tt_w = 0;
  uint32_t depth_key = 0xFFFFFFFFu;  // culled Gaussians sort behind every visible one of their camera
  do {
    const float mx = s_mean[3 * threadIdx.x], my = s_mean[3 * threadIdx.x + 1], mz = s_mean[3 * threadIdx.x + 2];
    const float* V = s_V;
    const float* Pm = s_Pm;
    const float tz = xform_row(V, 2, mx, my, mz);
    if (tz <= 0.2f) {  // in_frustum (cr/auxiliary.h:152-162); NaN passes, like the reference
      if (prefiltered) atomicExch(err_flag, GSB_EPREFILTER);
      break;
    }
    const float hx = xform_row(Pm, 0, mx, my, mz);
    const float hy = xform_row(Pm, 1, mx, my, mz);
    const float hw = ADD(Pm[15], FMA(mz, Pm[11], FMA(Pm[3], mx, MUL(Pm[7], my))));
    const float p_w = DIV(1.0f, ADD(hw, 0.0000001f));
    const float ppx = MUL(hx, p_w);
    const float ppy = MUL(hy, p_w);

    float c[6];
    if (cov3D_precomp) {
#pragma unroll
      for (int k = 0; k < 6; k++) c[k] = __ldg(cov3D_precomp + 6 * (size_t)idx + k);
    } else {
      const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
      cov3d_from_scale_rot(s_scale[3 * threadIdx.x], s_scale[3 * threadIdx.x + 1], s_scale[3 * threadIdx.x + 2],
                           mod, q, c);
      // one table for the whole batch: the covariance does not depend on the camera (every camera that sees the
      // Gaussian writes the same six values)
      float2* dst = reinterpret_cast<float2*>(cov3D + 6 * (size_t)idx);
      dst[0] = make_float2(c[0], c[1]);
      dst[1] = make_float2(c[2], c[3]);
      dst[2] = make_float2(c[4], c[5]);
    }

    // computeCov2D (cr/forward.cu:74-113)
    const CamParams cp = s_cp;
    const float tx_ = xform_row(V, 0, mx, my, mz);
    const float ty_ = xform_row(V, 1, mx, my, mz);
    const float limx = MUL(cp.tanfovx, 1.3f);
    const float limy = MUL(cp.tanfovy, 1.3f);
    const float txtz = DIV(tx_, tz);
    const float tytz = DIV(ty_, tz);
    const float cx = fminf(limx, fmaxf(txtz, -limx));
    const float cy = fminf(limy, fmaxf(tytz, -limy));
    const float ntx = MUL(tz, -cx);
    const float nty = MUL(tz, -cy);
    const float tz2 = MUL(tz, tz);
    const float J02 = DIV(MUL(cp.focal_x, ntx), tz2);
    const float J12 = DIV(MUL(cp.focal_y, nty), tz2);
    const float J00 = DIV(cp.focal_x, tz);
    const float J11 = DIV(cp.focal_y, tz);
    const float T02 = FMA(V[10], J02, FMA(V[8], J00, MUL(0.0f, V[9])));
    const float T00 = FMA(V[2], J02, FMA(V[0], J00, MUL(0.0f, V[1])));
    const float T01 = FMA(V[6], J02, FMA(V[4], J00, MUL(0.0f, V[5])));
    const float T12 = FMA(V[10], J12, FMA(0.0f, V[8], MUL(V[9], J11)));
    const float T10 = FMA(V[2], J12, FMA(0.0f, V[0], MUL(V[1], J11)));
    const float T11 = FMA(V[6], J12, FMA(0.0f, V[4], MUL(V[5], J11)));
    const float A20 = FMA(c[5], T02, FMA(c[2], T00, MUL(c[4], T01)));
    const float A00 = FMA(c[2], T02, FMA(c[0], T00, MUL(c[1], T01)));
    const float A10 = FMA(c[4], T02, FMA(c[1], T00, MUL(c[3], T01)));
    const float cov00 = FMA(T02, A20, FMA(T00, A00, MUL(T01, A10)));
    const float a = ADD(cov00, 0.3f);
    const float A21 = FMA(c[5], T12, FMA(c[2], T10, MUL(c[4], T11)));
    const float A01 = FMA(c[2], T12, FMA(c[0], T10, MUL(c[1], T11)));
    const float A11 = FMA(c[4], T12, FMA(c[1], T10, MUL(c[3], T11)));
    const float cov11 = FMA(T12, A21, FMA(T10, A01, MUL(T11, A11)));
    const float cc = ADD(cov11, 0.3f);
    const float b = FMA(T02, A21, FMA(T00, A01, MUL(T01, A11)));

    // conic + radius (cr/forward.cu:218-232)
    const float det = FMA(a, cc, -MUL(b, b));
    if (det == 0.0f) break;
    const float det_inv = DIV(1.0f, det);
    const float con_x = MUL(cc, det_inv);
    const float con_y = MUL(b, -det_inv);
    const float con_z = MUL(a, det_inv);
    const float mid = MUL(ADD(a, cc), 0.5f);
    const float disc = fmaxf(FMA(mid, mid, -det), 0.1f);
    const float sq = SQRT(disc);
    const float lam = fmaxf(ADD(mid, sq), SUB(mid, sq));
    const int rad = __float2int_ru(MUL(SQRT(lam), 3.0f));

    // ndc2Pix in FP64 (cr/auxiliary.h:41-44)
    const float pix_x = (float)(__dmul_rn(__fma_rn(__dadd_rn((double)ppx, 1.0), (double)W, -1.0), 0.5));
    const float pix_y = (float)(__dmul_rn(__fma_rn(__dadd_rn((double)ppy, 1.0), (double)H, -1.0), 0.5));

    uint32_t x0, y0, x1, y1;
    tile_rect(pix_x, pix_y, rad, gx, gy, x0, y0, x1, y1);
    uint32_t area = (x1 - x0) * (y1 - y0);
    if (area == 0) break;
    // tile-band sharding: the Gaussian stays visible (its state is the whole camera's) but instances are only
    // emitted for the tile rows this entry renders
    y0 = min(max(y0, (uint32_t)cp.row0), (uint32_t)cp.row1);
    y1 = min(max(y1, (uint32_t)cp.row0), (uint32_t)cp.row1);
    area = (x1 - x0) * (y1 - y0);

    depths[o] = tz;
    means2D[o] = make_float2(pix_x, pix_y);
    conic_opacity[o] = make_float4(con_x, con_y, con_z, __ldg(opacities + idx));
    rad_w = rad;
    tt_w = area;
    depth_key = __float_as_uint(tz);
  } while (false);
  gauss_keys[o] = depth_key;   // the camera's segment of the depth sort; the payload is the index itself
  radii_state[o] = rad_w;
  if (radii_out) radii_out[o] = rad_w;
  tiles_touched[o] = tt_w;
}

int launch_preprocess(const gsb_raster_args& a, const CamParams* cams_dev, const gsb_raster_ctx& c,
                      int32_t* radii_out, uint32_t* gauss_keys_in, int* err_flag_dev, cudaStream_t st) {
  if (a.P <= 0) return GSB_OK;
  dim3 grid((a.P + 255) / 256, a.n_cams);
  preprocess_fwd_kernel<<<grid, 256, 0, st>>>(
      a.P, a.width, a.height, c.tiles_x, c.tiles_y, a.scale_modifier, a.prefiltered, a.means3D, a.scales,
      a.rotations, a.opacities, a.cov3D_precomp, a.viewmatrix, a.projmatrix, cams_dev, c.depths, c.radii,
      radii_out, reinterpret_cast<float2*>(c.means2D), c.cov3D, reinterpret_cast<float4*>(c.conic_opacity),
      c.tiles_touched, gauss_keys_in, err_flag_dev);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// checkFrustum (cr/rasterizer_impl.cu:54-66): present = p_view.z > 0.2
__global__ void __launch_bounds__(256)
mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                    uint8_t* __restrict__ present) {
  __shared__ __align__(16) float s_mean[768];
  stage3(means3D, P, blockIdx.x * 256, s_mean);
  __syncthreads();
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= P) return;
  const float z = xform_row(view, 2, s_mean[3 * threadIdx.x], s_mean[3 * threadIdx.x + 1], s_mean[3 * threadIdx.x + 2]);
  present[idx] = (z <= 0.2f) ? 0 : 1;
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t st) {
  if (P <= 0) return GSB_OK;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, view, present);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// ------------------------------------------------------------------------------------------------
// Fused backward preprocess.  glm::mat3 is column-major; below m[c][r] mirrors that convention so the
// formulas can be checked against cr/backward.cu line by line.  Tolerance-level parity (1e-4 rel).
// ------------------------------------------------------------------------------------------------
struct M3 { float m[3][3]; };  // m[col][row]
__device__ __forceinline__ M3 m3_mul(const M3& A, const M3& B) {
  M3 R;
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int i = 0; i < 3; i++) R.m[j][i] = A.m[0][i] * B.m[j][0] + A.m[1][i] * B.m[j][1] + A.m[2][i] * B.m[j][2];
  return R;
}
__device__ __forceinline__ M3 m3_t(const M3& A) {
  M3 R;
#pragma unroll
  for (int j = 0; j < 3; j++)
#pragma unroll
    for (int i = 0; i < 3; i++) R.m[j][i] = A.m[i][j];
  return R;
}

// FLAT: the exchange form of the step's backward (gsb_raster_backward_flat): the thread's results go into one
// GSB_FLAT_WIDTH-float row (five 16-byte stores, rows contiguous across the warp) together with the blend's
// camera-summed opacity / colour gradients and the densification-statistic increments of this batch
// (per-camera ||dL/dmean2D.xy|| summed over the visible cameras, their count, the largest radius), so that no
// pack pass, no statistics launch and no separate gradient tables exist on the multi-GPU path.
// TWO (FLAT only): the batch was rendered with two colour tables in one walk: the second table's colour gradient fills
// cols 17-19 and the statistic's norm is taken of the first image's screen-space gradient alone.
template <bool FLAT, bool TWO>
__global__ void __launch_bounds__(256)
preprocess_bwd_kernel(int P, int n_cams, int W, int H, float mod, const float* __restrict__ means3D,
                      const float* __restrict__ scales, const float* __restrict__ rotations,
                      const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                      const CamParams* __restrict__ cams, const int32_t* __restrict__ radii,
                      const float* __restrict__ cov3Ds, size_t cov_cam_stride,
                      const float4* __restrict__ conic_opacity, float* __restrict__ dL_dmean2D,
                      const float* __restrict__ dL_dconic, float* __restrict__ dL_dmean3D,
                      float* __restrict__ dL_dcov3D, float* __restrict__ dL_dscale, float* __restrict__ dL_drot,
                      float* __restrict__ flat, int flat_mode, const float* __restrict__ dL_dopacity,
                      const float* __restrict__ dL_dcolor, float* __restrict__ dL_dmean2D_out, int flat_stride,
                      const float* __restrict__ dL_dcolor2, const float* __restrict__ dL_dmean2D_first) {
  extern __shared__ float s_mats[];  // [n_cams][32]: view then proj; FLAT: then int [n_cams][2] = split slot, count flag
  for (int i = threadIdx.x; i < n_cams * 32; i += blockDim.x) {
    const int cam = i >> 5, k = i & 31;
    s_mats[i] = (k < 16) ? viewmatrix[16 * cam + k] : projmatrix[16 * cam + k - 16];
  }
  int* const s_entry = reinterpret_cast<int*>(s_mats + 32 * n_cams);
  if (FLAT)
    for (int i = threadIdx.x; i < n_cams; i += blockDim.x) {
      s_entry[2 * i] = cams[i].split_slot;
      s_entry[2 * i + 1] = cams[i].count_vis;
    }
  __shared__ __align__(16) float s_mean[768];
  stage3(means3D, P, blockIdx.x * 256, s_mean);
  __syncthreads();
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= P) return;
  const float mx = s_mean[3 * threadIdx.x], my = s_mean[3 * threadIdx.x + 1], mz = s_mean[3 * threadIdx.x + 2];

  float gm[3] = {0.f, 0.f, 0.f};
  float gc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  bool any = false;
  float st_norm = 0.f, st_cnt = 0.f, st_rad = 0.f;   // FLAT: statistic increments of this batch
  float4* const frow = FLAT ? reinterpret_cast<float4*>(flat) + (size_t)idx * (flat_stride / 4) : nullptr;
  if (FLAT && flat_mode == 0)   // split-camera slots: zero unless one of this batch's entries owns the slot
    for (int k = GSB_FLAT_WIDTH / 4; k < flat_stride / 4; k++) frow[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  // the 3D covariance does not depend on the camera: one row per Gaussian (cov_cam_stride is 0)
  float cov3D[6];
  {
    const float* cp6 = cov3Ds + 6 * (size_t)idx;
#pragma unroll
    for (int k = 0; k < 6; k++) cov3D[k] = __ldg(cp6 + k);
  }
  (void)cov_cam_stride;
  for (int cam = 0; cam < n_cams; cam++) {
    const size_t o = (size_t)cam * P + idx;
    const int rad = radii[o];
    if (!(rad > 0)) {
      if (FLAT && dL_dmean2D_out) { dL_dmean2D_out[3 * o] = 0.f; dL_dmean2D_out[3 * o + 1] = 0.f; dL_dmean2D_out[3 * o + 2] = 0.f; }
      continue;
    }
    any = true;
    const float* view = s_mats + 32 * cam;
    const float* proj = view + 16;
    // the four scalars only (one 16-byte load): the entry fields of the FLAT form sit in shared memory
    const float4 cp4 = __ldg(reinterpret_cast<const float4*>(cams + cam));
    struct { float tanfovx, tanfovy, focal_x, focal_y; } cp = {cp4.x, cp4.y, cp4.z, cp4.w};
    const float h_x = cp.focal_x, h_y = cp.focal_y;
    // the blend kernel leaves raw pixel moments of w = dL/dG * G: (S1, S2) in dL_dmean2D, (S1 dx, S1 dy, -, S2 dy)
    // in dL_dconic; their linear maps to the gradients (cr/backward.cu:536-548) are applied here, once per Gaussian
    const float4 dcon = __ldg(reinterpret_cast<const float4*>(dL_dconic) + o);
    const float4 con = __ldg(conic_opacity + o);
    const float dA = -0.5f * dcon.x, dB = -0.5f * dcon.y, dC = -0.5f * dcon.w;  // slot z is never written (cr/backward.cu:165)
    const float rs1 = dL_dmean2D[3 * o], rs2 = dL_dmean2D[3 * o + 1];
    const float g2x = -(0.5f * W) * (con.x * rs1 + con.y * rs2);
    const float g2y = -(0.5f * H) * (con.z * rs2 + con.y * rs1);
    if (FLAT) {
      float n2x = g2x, n2y = g2y;   // the gradient whose norm is the statistic: the first (colour) image's alone
      if (TWO) {
        const float f1 = dL_dmean2D_first[2 * o], f2 = dL_dmean2D_first[2 * o + 1];
        n2x = -(0.5f * W) * (con.x * f1 + con.y * f2);
        n2y = -(0.5f * H) * (con.z * f2 + con.y * f1);
      }
      const int split_slot = s_entry[2 * cam];
      if (split_slot >= 0) {
        // rows of this camera are rendered elsewhere too: the norm belongs to the SUMMED gradient (after the exchange)
        float* slot = flat + (size_t)idx * flat_stride + GSB_FLAT_WIDTH + 2 * split_slot;
        if (flat_mode == 0) { slot[0] = n2x; slot[1] = n2y; }
        else if (flat_mode == 2) { slot[0] += n2x; slot[1] += n2y; }
      } else {
        st_norm += sqrtf(n2x * n2x + n2y * n2y);   // torch.norm of the first two components, one term per camera
      }
      if (s_entry[2 * cam + 1]) {
        st_cnt += 1.0f;
        st_rad = fmaxf(st_rad, (float)rad);
      }
      if (dL_dmean2D_out) { dL_dmean2D_out[3 * o] = n2x; dL_dmean2D_out[3 * o + 1] = n2y; dL_dmean2D_out[3 * o + 2] = 0.f; }
    } else {
      dL_dmean2D[3 * o] = g2x;
      dL_dmean2D[3 * o + 1] = g2y;
    }

    // computeCov2DCUDA (cr/backward.cu:144-274)
    float tx = view[0] * mx + view[4] * my + view[8] * mz + view[12];
    float ty = view[1] * mx + view[5] * my + view[9] * mz + view[13];
    const float tz = view[2] * mx + view[6] * my + view[10] * mz + view[14];
    const float limx = 1.3f * cp.tanfovx, limy = 1.3f * cp.tanfovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
    M3 J = {{{h_x / tz, 0.0f, -(h_x * tx) / (tz * tz)}, {0.0f, h_y / tz, -(h_y * ty) / (tz * tz)}, {0.f, 0.f, 0.f}}};
    M3 Wm = {{{view[0], view[4], view[8]}, {view[1], view[5], view[9]}, {view[2], view[6], view[10]}}};
    M3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
    const M3 T = m3_mul(Wm, J);
    const M3 cov2D = m3_mul(m3_mul(m3_t(T), m3_t(Vrk)), T);
    const float a = cov2D.m[0][0] + 0.3f, b = cov2D.m[0][1], c = cov2D.m[1][1] + 0.3f;
    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
#define TT(i, j) T.m[i][j]
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * dA + 2 * b * c * dB + (denom - a * c) * dC);
      dL_dc = denom2inv * (-a * a * dC + 2 * a * b * dB + (denom - a * c) * dA);
      dL_db = denom2inv * 2 * (b * c * dA - (denom + 2 * b * b) * dB + a * b * dC);
      gc[0] += (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(1, 0) * TT(1, 0) * dL_dc);
      gc[3] += (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(1, 1) * TT(1, 1) * dL_dc);
      gc[5] += (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(1, 2) * TT(1, 2) * dL_dc);
      gc[1] += 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
      gc[2] += 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
      gc[4] += 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
    }
#define VV(i, j) Vrk.m[i][j]
    const float dL_dT00 = 2 * (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_da + (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_db;
    const float dL_dT01 = 2 * (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_da + (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_db;
    const float dL_dT02 = 2 * (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_da + (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_db;
    const float dL_dT10 = 2 * (TT(1, 0) * VV(0, 0) + TT(1, 1) * VV(0, 1) + TT(1, 2) * VV(0, 2)) * dL_dc + (TT(0, 0) * VV(0, 0) + TT(0, 1) * VV(0, 1) + TT(0, 2) * VV(0, 2)) * dL_db;
    const float dL_dT11 = 2 * (TT(1, 0) * VV(1, 0) + TT(1, 1) * VV(1, 1) + TT(1, 2) * VV(1, 2)) * dL_dc + (TT(0, 0) * VV(1, 0) + TT(0, 1) * VV(1, 1) + TT(0, 2) * VV(1, 2)) * dL_db;
    const float dL_dT12 = 2 * (TT(1, 0) * VV(2, 0) + TT(1, 1) * VV(2, 1) + TT(1, 2) * VV(2, 2)) * dL_dc + (TT(0, 0) * VV(2, 0) + TT(0, 1) * VV(2, 1) + TT(0, 2) * VV(2, 2)) * dL_db;
#undef VV
#undef TT
#define WW(i, j) Wm.m[i][j]
    const float dL_dJ00 = WW(0, 0) * dL_dT00 + WW(0, 1) * dL_dT01 + WW(0, 2) * dL_dT02;
    const float dL_dJ02 = WW(2, 0) * dL_dT00 + WW(2, 1) * dL_dT01 + WW(2, 2) * dL_dT02;
    const float dL_dJ11 = WW(1, 0) * dL_dT10 + WW(1, 1) * dL_dT11 + WW(1, 2) * dL_dT12;
    const float dL_dJ12 = WW(2, 0) * dL_dT10 + WW(2, 1) * dL_dT11 + WW(2, 2) * dL_dT12;
#undef WW
    const float itz = 1.f / tz, tz2 = itz * itz, tz3 = tz2 * itz;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * tx) * tz3 * dL_dJ02 + (2 * h_y * ty) * tz3 * dL_dJ12;
    // transformVec4x3Transpose (cr/auxiliary.h:89-97)
    float dmx = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
    float dmy = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
    float dmz = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;

    // preprocessCUDA bwd (cr/backward.cu:370-387): mean2D gradient through the projection
    const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
    const float m_w = 1.0f / (m_hw + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    dmx += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
    dmy += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
    dmz += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
    gm[0] += dmx; gm[1] += dmy; gm[2] += dmz;
  }

  if (!FLAT) {
    dL_dmean3D[3 * (size_t)idx] = gm[0];
    dL_dmean3D[3 * (size_t)idx + 1] = gm[1];
    dL_dmean3D[3 * (size_t)idx + 2] = gm[2];
    if (dL_dcov3D) {
#pragma unroll
      for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = gc[k];
    }
    if (!dL_dscale || !dL_drot) return;
  }
  float gs[3] = {0.f, 0.f, 0.f};
  float4 gq = make_float4(0.f, 0.f, 0.f, 0.f);
  if (any && scales && rotations) {
    // computeCov3D bwd (cr/backward.cu:278-341)
    const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    M3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    const float sx = mod * __ldg(scales + 3 * (size_t)idx), sy = mod * __ldg(scales + 3 * (size_t)idx + 1),
                sz = mod * __ldg(scales + 3 * (size_t)idx + 2);
    M3 S = {{{sx, 0, 0}, {0, sy, 0}, {0, 0, sz}}};
    const M3 M = m3_mul(S, R);
    M3 dSig = {{{gc[0], 0.5f * gc[1], 0.5f * gc[2]}, {0.5f * gc[1], gc[3], 0.5f * gc[4]}, {0.5f * gc[2], 0.5f * gc[4], gc[5]}}};
    M3 M2;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
      for (int i = 0; i < 3; i++) M2.m[j][i] = 2.0f * M.m[j][i];
    const M3 dL_dM = m3_mul(M2, dSig);
    const M3 Rt = m3_t(R);
    M3 dMt = m3_t(dL_dM);
#pragma unroll
    for (int k = 0; k < 3; k++) gs[k] = Rt.m[k][0] * dMt.m[k][0] + Rt.m[k][1] * dMt.m[k][1] + Rt.m[k][2] * dMt.m[k][2];
#pragma unroll
    for (int i = 0; i < 3; i++) { dMt.m[0][i] *= sx; dMt.m[1][i] *= sy; dMt.m[2][i] *= sz; }
#define D(i, j) dMt.m[i][j]
    gq.x = 2 * z * (D(0, 1) - D(1, 0)) + 2 * y * (D(2, 0) - D(0, 2)) + 2 * x * (D(1, 2) - D(2, 1));
    gq.y = 2 * y * (D(1, 0) + D(0, 1)) + 2 * z * (D(2, 0) + D(0, 2)) + 2 * r * (D(1, 2) - D(2, 1)) - 4 * x * (D(2, 2) + D(1, 1));
    gq.z = 2 * x * (D(1, 0) + D(0, 1)) + 2 * r * (D(2, 0) - D(0, 2)) + 2 * z * (D(1, 2) + D(2, 1)) - 4 * y * (D(2, 2) + D(0, 0));
    gq.w = 2 * r * (D(0, 1) - D(1, 0)) + 2 * x * (D(2, 0) + D(0, 2)) + 2 * y * (D(1, 2) + D(2, 1)) - 4 * z * (D(1, 1) + D(0, 0));
#undef D
  }
  if (FLAT) {
    float4* row = frow;
    const float dop = dL_dopacity[idx];
    const float c0 = dL_dcolor[3 * (size_t)idx], c1 = dL_dcolor[3 * (size_t)idx + 1], c2 = dL_dcolor[3 * (size_t)idx + 2];
    if (flat_mode == 0) {
      row[0] = make_float4(gm[0], gm[1], gm[2], c0);
      row[1] = make_float4(c1, c2, gq.x, gq.y);
      row[2] = make_float4(gq.z, gq.w, dop, gs[0]);
      row[3] = make_float4(gs[1], gs[2], st_norm, st_cnt);
      if (TWO)   // two-table backward: the second colour table's gradient arrived in the same walk
        row[4] = make_float4(st_rad, dL_dcolor2[3 * (size_t)idx], dL_dcolor2[3 * (size_t)idx + 1], dL_dcolor2[3 * (size_t)idx + 2]);
      else
        row[4] = make_float4(st_rad, 0.f, 0.f, 0.f);
    } else if (flat_mode == 2) {
      // another camera group of the same step (rendered concurrently on its own stream): everything adds up
      float4 r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3], r4 = row[4];
      r0.x += gm[0]; r0.y += gm[1]; r0.z += gm[2]; r0.w += c0;
      r1.x += c1; r1.y += c2; r1.z += gq.x; r1.w += gq.y;
      r2.x += gq.z; r2.y += gq.w; r2.z += dop; r2.w += gs[0];
      r3.x += gs[1]; r3.y += gs[2]; r3.z += st_norm; r3.w += st_cnt;
      r4.x = fmaxf(r4.x, st_rad);
      row[0] = r0; row[1] = r1; row[2] = r2; row[3] = r3; row[4] = r4;
    } else {
      // second render of the iteration (another colour table): geometry gradients add up, colours go to 17-19
      float4 r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3];
      r0.x += gm[0]; r0.y += gm[1]; r0.z += gm[2];
      r1.z += gq.x; r1.w += gq.y;
      r2.x += gq.z; r2.y += gq.w; r2.z += dop; r2.w += gs[0];
      r3.x += gs[1]; r3.y += gs[2];
      row[0] = r0; row[1] = r1; row[2] = r2; row[3] = r3;
      row[4] = make_float4(row[4].x, c0, c1, c2);
    }
    return;
  }
  dL_dscale[3 * (size_t)idx] = gs[0];
  dL_dscale[3 * (size_t)idx + 1] = gs[1];
  dL_dscale[3 * (size_t)idx + 2] = gs[2];
  reinterpret_cast<float4*>(dL_drot)[idx] = gq;
}

// DensificationVariables += the statistic columns of a (reduced) flat buffer
__global__ void __launch_bounds__(256)
flat_apply_stats_kernel(int P, const float* __restrict__ flat, int flat_stride, float* __restrict__ grad_accum,
                        float* __restrict__ vis_count, float* __restrict__ max_2d_radii) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const float* row = flat + (size_t)i * flat_stride;
  float2 nc = *reinterpret_cast<const float2*>(row + GSB_FLAT_COL_GRADNORM);
  const float mr = row[GSB_FLAT_COL_MAXRADIUS];
  // split cameras (tile-band sharding): the norm of the camera's summed screen-space gradient
  for (int k = GSB_FLAT_WIDTH; k + 1 < flat_stride; k += 2) {
    const float2 g = *reinterpret_cast<const float2*>(row + k);
    nc.x += sqrtf(g.x * g.x + g.y * g.y);
  }
  if (nc.y > 0.f) {
    grad_accum[i] += nc.x;
    vis_count[i] += nc.y;
  }
  if (max_2d_radii) max_2d_radii[i] = fmaxf(max_2d_radii[i], mr);
}

int launch_flat_apply_stats(int P, const float* flat, int flat_stride, float* grad_accum, float* vis_count,
                            float* max_2d_radii, cudaStream_t st) {
  if (P <= 0) return GSB_OK;
  flat_apply_stats_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, flat, flat_stride, grad_accum, vis_count, max_2d_radii);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

// per-camera scalars on the device (asynchronous forward): focal = dim / (2 tanfov), the same two FP32 operations
// the host performs on the blocking path (cr/rasterizer_impl.cu:222-223), so the bits agree
// tanfov == nullptr: the host already wrote the four scalars (blocking path); only the entry fields are filled in
__global__ void cam_params_kernel(int n_cams, int W, int H, const float* __restrict__ tanfov,
                                  const int32_t* __restrict__ entries, CamParams* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_cams) return;
  CamParams c = out[i];
  if (tanfov) {
    c.tanfovx = tanfov[2 * i];
    c.tanfovy = tanfov[2 * i + 1];
    c.focal_y = __fdiv_rn((float)H, __fmul_rn(2.0f, c.tanfovy));
    c.focal_x = __fdiv_rn((float)W, __fmul_rn(2.0f, c.tanfovx));
  }
  const int gy = (H + GSB_TILE - 1) / GSB_TILE;
  c.row0 = 0; c.row1 = gy; c.split_slot = -1; c.count_vis = 1;
  if (entries) {
    c.row0 = min(max(entries[4 * i], 0), gy);
    c.row1 = min(max(entries[4 * i + 1], c.row0), gy);
    c.split_slot = entries[4 * i + 2];
    c.count_vis = entries[4 * i + 3] != 0;
  }
  out[i] = c;
}

int launch_cam_params(int n_cams, int W, int H, const float* tanfov_dev, const int32_t* entries, CamParams* cams_dev,
                      cudaStream_t st) {
  cam_params_kernel<<<(n_cams + 127) / 128, 128, 0, st>>>(n_cams, W, H, tanfov_dev, entries, cams_dev);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int launch_preprocess_bwd(const gsb_raster_args& a, const CamParams* cams_dev, const gsb_raster_ctx& c,
                          float* dL_dmean2D, const float* dL_dconic, float* dL_dmean3D, float* dL_dcov3D,
                          float* dL_dscale, float* dL_drot, cudaStream_t st, float* flat, int flat_mode,
                          const float* dL_dopacity, const float* dL_dcolor, float* dL_dmean2D_out, int flat_stride,
                          const float* dL_dcolor2, const float* dL_dmean2D_first) {
  if (a.P <= 0) return GSB_OK;
  const size_t smem = (size_t)a.n_cams * (32 * sizeof(float) + (flat ? 2 * sizeof(int) : 0));
  if (a.n_cams > GSB_MAX_CAMS) return GSB_EINVAL;   // gsb_raster_forward refuses such batches as well
  if (flat) {
    if (!dL_dopacity || !dL_dcolor || !a.scales || !a.rotations) return GSB_EINVAL;
    if ((dL_dcolor2 != nullptr) != (dL_dmean2D_first != nullptr)) return GSB_EINVAL;
    auto kern = dL_dcolor2 ? preprocess_bwd_kernel<true, true> : preprocess_bwd_kernel<true, false>;
    kern<<<(a.P + 255) / 256, 256, smem, st>>>(
        a.P, a.n_cams, a.width, a.height, a.scale_modifier, a.means3D, a.scales, a.rotations, a.viewmatrix,
        a.projmatrix, cams_dev, c.radii, c.cov3D, (size_t)0, reinterpret_cast<const float4*>(c.conic_opacity),
        dL_dmean2D, dL_dconic, nullptr, nullptr, nullptr, nullptr, flat, flat_mode, dL_dopacity, dL_dcolor,
        dL_dmean2D_out, flat_stride, dL_dcolor2, dL_dmean2D_first);
  } else {
    preprocess_bwd_kernel<false, false><<<(a.P + 255) / 256, 256, smem, st>>>(
        a.P, a.n_cams, a.width, a.height, a.scale_modifier, a.means3D, a.scales, a.rotations, a.viewmatrix,
        a.projmatrix, cams_dev, c.radii, a.cov3D_precomp ? a.cov3D_precomp : c.cov3D, (size_t)0,
        reinterpret_cast<const float4*>(c.conic_opacity), dL_dmean2D, dL_dconic, dL_dmean3D,
        dL_dcov3D, dL_dscale, dL_drot, nullptr, 0, nullptr, nullptr, nullptr, GSB_FLAT_WIDTH, nullptr, nullptr);
  }
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // namespace gsb
