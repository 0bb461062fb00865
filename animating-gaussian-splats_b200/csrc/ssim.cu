// ssim.cu -- fused L1 + SSIM loss, forward and backward.
//
// Replaces external.calc_ssim / _ssim / create_window / gaussian (external.py:49-110: five grouped
// 11x11 conv2d calls plus ~15 elementwise kernels, window rebuilt on the CPU every call), torch's
// l1_loss and the 0.8/0.2 mix (densify.py:127-129, train.py:362-364) with two kernels:
//   fwd : per 16x16 tile, stage the 26x26 halo of both images in shared memory, run the separable
//         11-tap Gaussian (sigma 1.5, zero padding 5) over x, y, x^2, y^2, xy, evaluate the SSIM map,
//         reduce sum(ssim) and sum|x-y| per image, and (if a gradient is wanted) store the three
//         partial-derivative maps d ssim/d(mu1, E[x^2], E[xy]).
//   bwd : the adjoint of a symmetric zero-padded correlation is the same correlation, so the three
//         maps are filtered once more and combined with the L1 sign term into dL/dimg.
// HBM-bound: 24 B/pixel read (+36 B/pixel of maps) forward, 60 B/pixel read + 12 B/pixel written backward.
#include "common.cuh"

namespace gsb {

constexpr int SS_TW = 32, SS_TH = 32;            // output tile (one block of 256 threads, 4 outputs each)
constexpr int SS_R = 5;                          // window radius
constexpr int SS_HW = SS_TW + 2 * SS_R;          // 42 halo columns
constexpr int SS_HH = SS_TH + 2 * SS_R;          // 42 halo rows
constexpr int SS_PX = SS_HW + 2;                 // halo row pitch: 44 floats, rows stay 16-byte aligned
constexpr int SS_PH = SS_TW + 4;                 // row-filtered pitch: 36 floats
constexpr int SS_THREADS = 256;
constexpr int SS_ROW_ITEMS = SS_HH * (SS_TW / 4);  // (halo row, group of 4 output columns)

struct SsimWindow { float g[11]; };

// gaussian(11, 1.5) as external.py:49-56 builds it: python doubles -> float32 tensor -> / sum() in float32
static SsimWindow make_window() {
  SsimWindow w;
  float s = 0.f;
  for (int i = 0; i < 11; i++) {
    const double d = (double)(i - 5);
    w.g[i] = (float)exp(-(d * d) / (2.0 * 1.5 * 1.5));
  }
  for (int i = 0; i < 11; i++) s += w.g[i];
  for (int i = 0; i < 11; i++) w.g[i] = w.g[i] / s;
  return w;
}

// Zero-padded halos of N planes that share their geometry (the two images forward, the three derivative
// maps backward).  A warp owns halo rows warp, warp + 8, ...; a lane the columns lane and lane + 32, so there
// is no index arithmetic per element and all global loads of the block are in flight before the first store.
// The two pad columns of each row are zeroed (they are loaded by the 16-byte row loads, never used).
// Rows outside [ylo, yhi) read as zero (the image border, or the edge of a band's derivative maps).
template <int N>
__device__ __forceinline__ void load_halos(const float* const (&src)[N], int ylo, int yhi, int W, int x0, int y0,
                                           float (*const (&dst)[N])[SS_PX], unsigned tid) {
  const int lane = tid & 31, warp = tid >> 5;
  const int xa = x0 - SS_R + lane, xb = xa + 32;
  const bool va = xa >= 0 && xa < W;
  const bool vb = lane + 32 < SS_HW && xb >= 0 && xb < W;
  constexpr int ROWS = (SS_HH + 7) / 8;
  float a[N][ROWS], b[N][ROWS];
#pragma unroll
  for (int i = 0; i < ROWS; i++) {
    const int r = warp + 8 * i;
    const int y = y0 + r - SS_R;
    const bool rok = r < SS_HH && y >= ylo && y < yhi;
    const size_t ro = (size_t)(rok ? y : 0) * W;
#pragma unroll
    for (int n = 0; n < N; n++) {
      a[n][i] = (rok && va) ? __ldg(src[n] + ro + xa) : 0.f;
      b[n][i] = (rok && vb) ? __ldg(src[n] + ro + xb) : 0.f;
    }
  }
#pragma unroll
  for (int i = 0; i < ROWS; i++) {
    const int r = warp + 8 * i;
    if (r < SS_HH) {
#pragma unroll
      for (int n = 0; n < N; n++) {
        dst[n][r][lane] = a[n][i];
        if (lane + 32 < SS_PX) dst[n][r][lane + 32] = b[n][i];
      }
    }
  }
}

// 16 consecutive halo values of one row starting at a column that is a multiple of 4 (four 16-byte loads)
__device__ __forceinline__ void load16(const float* row, float (&v)[16]) {
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const float4 t = *reinterpret_cast<const float4*>(row + 4 * q);
    v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
  }
}

// Register-blocked separable filter on a 32x32 output tile.  Row pass: a work item filters 4 adjacent
// outputs of one halo row from 14 inputs fetched with four 16-byte shared loads per image (336 items over
// 256 threads); column pass: a thread filters 4 vertically adjacent outputs of one column from 14 row-filtered
// values.  Per output ~2.5 shared loads per filtered quantity instead of 22 (one per tap and image).
__global__ void __launch_bounds__(SS_THREADS)
ssim_fwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ target, SsimWindow win,
                double* __restrict__ accum /* [n][2] */, float* __restrict__ maps /* [3][planes][H][W] or null */,
                size_t plane_count, const int32_t* __restrict__ rows /* [n][4] or null: bands */) {
  __shared__ __align__(16) float s_x[SS_HH][SS_PX];
  __shared__ __align__(16) float s_y[SS_HH][SS_PX];
  __shared__ __align__(16) float s_h[5][SS_HH][SS_PH];
  __shared__ float s_red[2][SS_THREADS / 32];
  const unsigned tid = threadIdx.x;
  const size_t plane = blockIdx.z;  // image * 3 + channel
  const size_t HW = (size_t)H * W;
  const int x0 = blockIdx.x * SS_TW, y0 = blockIdx.y * SS_TH;
  int ly0 = 0, ly1 = H;   // pixel rows that enter the loss
  if (rows) { ly0 = max(rows[4 * (plane / 3)], 0); ly1 = min(rows[4 * (plane / 3) + 1], H); }
  if (y0 >= ly1 || y0 + SS_TH <= ly0) return;   // block-uniform
  {
    const float* const src[2] = {img + plane * HW, target + plane * HW};
    float (*const dst[2])[SS_PX] = {s_x, s_y};
    load_halos<2>(src, 0, H, W, x0, y0, dst, tid);
  }
  __syncthreads();
  for (int item = tid; item < SS_ROW_ITEMS; item += SS_THREADS) {
    const int r = item / (SS_TW / 4), c0 = (item % (SS_TW / 4)) * 4;
    float xv[16], yv[16];
    load16(&s_x[r][c0], xv);
    load16(&s_y[r][c0], yv);
    float a[5][4];
#pragma unroll
    for (int q = 0; q < 5; q++)
#pragma unroll
      for (int o = 0; o < 4; o++) a[q][o] = 0.f;
#pragma unroll
    for (int k = 0; k < 14; k++) {
      const float x = xv[k], y = yv[k];
      const float xx = x * x, yy = y * y, xy = x * y;
#pragma unroll
      for (int o = 0; o < 4; o++) {
        const int t = k - o;  // tap index for output o
        if (t >= 0 && t < 11) {
          const float g = win.g[t];
          a[0][o] = fmaf(g, x, a[0][o]);
          a[1][o] = fmaf(g, y, a[1][o]);
          a[2][o] = fmaf(g, xx, a[2][o]);
          a[3][o] = fmaf(g, yy, a[3][o]);
          a[4][o] = fmaf(g, xy, a[4][o]);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < 5; q++)
      *reinterpret_cast<float4*>(&s_h[q][r][c0]) = make_float4(a[q][0], a[q][1], a[q][2], a[q][3]);
  }
  __syncthreads();
  const int tx = tid & 31, ty0 = (tid >> 5) * 4;
  float f[5][4];
#pragma unroll
  for (int q = 0; q < 5; q++) {
    float v[14];
#pragma unroll
    for (int k = 0; k < 14; k++) v[k] = s_h[q][ty0 + k][tx];
#pragma unroll
    for (int o = 0; o < 4; o++) {
      float b = 0.f;
#pragma unroll
      for (int k = 0; k < 11; k++) b = fmaf(win.g[k], v[k + o], b);
      f[q][o] = b;
    }
  }
  float ssim_sum = 0.f, l1_sum = 0.f;
  const int x = x0 + tx;
#pragma unroll
  for (int o = 0; o < 4; o++) {
    const int y = y0 + ty0 + o;
    if (x < W && y >= ly0 && y < ly1) {
      const float mu1 = f[0][o], mu2 = f[1][o], e11 = f[2][o], e22 = f[3][o], e12 = f[4][o];
      const float c1 = 0.0001f, c2 = 0.0009f;  // 0.01^2, 0.03^2 (external.py:96-97)
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
      const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
      const float An = 2.f * mu12 + c1, Bn = 2.f * s12 + c2, Cd = mu1_sq + mu2_sq + c1, Dd = s1 + s2 + c2;
      const float inv = 1.0f / (Cd * Dd);
      const float ssim = An * Bn * inv;
      ssim_sum += ssim;
      l1_sum += fabsf(s_x[ty0 + o + SS_R][tx + SS_R] - s_y[ty0 + o + SS_R][tx + SS_R]);
      if (maps) {
        // d ssim / d mu1, d E[x^2], d E[xy]   (sigma1^2 = E[x^2] - mu1^2, sigma12 = E[xy] - mu1 mu2)
        const float d_mu1 = (2.f * mu2 * Bn - 2.f * mu2 * An) * inv - ssim * (2.f * mu1 * Dd - 2.f * mu1 * Cd) * inv;
        const float d_e11 = -ssim * Cd * inv;
        const float d_e12 = 2.f * An * inv;
        const size_t off = plane * HW + (size_t)y * W + x;
        maps[off] = d_mu1;
        maps[plane_count * HW + off] = d_e11;
        maps[2 * plane_count * HW + off] = d_e12;
      }
    }
  }
  // block reduction -> one double atomic per quantity per block
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    ssim_sum += __shfl_xor_sync(0xffffffffu, ssim_sum, s);
    l1_sum += __shfl_xor_sync(0xffffffffu, l1_sum, s);
  }
  if ((tid & 31u) == 0) { s_red[0][tid >> 5] = l1_sum; s_red[1][tid >> 5] = ssim_sum; }
  __syncthreads();
  if (tid < 2) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < SS_THREADS / 32; w++) t += (double)s_red[tid][w];
    atomicAdd(accum + 2 * (plane / 3) + tid, t);
  }
}

__global__ void ssim_finalize_kernel(int n, int H, double inv_count, float w_l1, float w_ssim,
                                     const double* __restrict__ accum, const int32_t* __restrict__ rows,
                                     float* __restrict__ loss_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double l1 = accum[2 * i] * inv_count, ss = accum[2 * i + 1] * inv_count;
  double frac = 1.0;   // a band's share of the constant in (1 - ssim), so that the bands of an image add up
  if (rows) frac = (double)max(min(rows[4 * i + 1], H) - max(rows[4 * i], 0), 0) / (double)H;
  loss_out[3 * i] = (float)l1;
  loss_out[3 * i + 1] = (float)ss;
  loss_out[3 * i + 2] = (float)((double)w_l1 * l1 + (double)w_ssim * (frac - ss));
}

__global__ void __launch_bounds__(SS_THREADS)
ssim_bwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ target, SsimWindow win,
                const float* __restrict__ maps, size_t plane_count, float k_l1, float k_ssim,
                float* __restrict__ dL_dimg, const int32_t* __restrict__ rows) {
  __shared__ __align__(16) float s_m[3][SS_HH][SS_PX];
  __shared__ __align__(16) float s_h[3][SS_HH][SS_PH];
  const unsigned tid = threadIdx.x;
  const size_t plane = blockIdx.z;
  const size_t HW = (size_t)H * W;
  const int x0 = blockIdx.x * SS_TW, y0 = blockIdx.y * SS_TH;
  int ly0 = 0, ly1 = H, wy0 = 0, wy1 = H;   // loss rows (the maps exist there), rows of dL/dimg to write
  if (rows) {
    const int32_t* r = rows + 4 * (plane / 3);
    ly0 = max(r[0], 0); ly1 = min(r[1], H); wy0 = max(r[2], 0); wy1 = min(r[3], H);
  }
  if (y0 >= wy1 || y0 + SS_TH <= wy0) return;   // block-uniform
  {
    const float* const src[3] = {maps + plane * HW, maps + (plane_count + plane) * HW, maps + (2 * plane_count + plane) * HW};
    float (*const dst[3])[SS_PX] = {s_m[0], s_m[1], s_m[2]};
    load_halos<3>(src, ly0, ly1, W, x0, y0, dst, tid);
  }
  __syncthreads();
  for (int item = tid; item < SS_ROW_ITEMS; item += SS_THREADS) {
    const int r = item / (SS_TW / 4), c0 = (item % (SS_TW / 4)) * 4;
#pragma unroll
    for (int m = 0; m < 3; m++) {
      float v[16];
      load16(&s_m[m][r][c0], v);
      float acc[4];
#pragma unroll
      for (int o = 0; o < 4; o++) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) t = fmaf(win.g[k], v[o + k], t);
        acc[o] = t;
      }
      *reinterpret_cast<float4*>(&s_h[m][r][c0]) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    }
  }
  __syncthreads();
  const int tx = tid & 31, ty0 = (tid >> 5) * 4;
  float b[3][4];
#pragma unroll
  for (int m = 0; m < 3; m++) {
    float v[14];
#pragma unroll
    for (int k = 0; k < 14; k++) v[k] = s_h[m][ty0 + k][tx];
#pragma unroll
    for (int o = 0; o < 4; o++) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < 11; k++) t = fmaf(win.g[k], v[k + o], t);
      b[m][o] = t;
    }
  }
  const int x = x0 + tx;
#pragma unroll
  for (int o = 0; o < 4; o++) {
    const int y = y0 + ty0 + o;
    if (x >= W || y < wy0 || y >= wy1) continue;
    const size_t off = plane * HW + (size_t)y * W + x;
    const float xv = __ldg(img + off), yv = __ldg(target + off);
    const float d = xv - yv;
    float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);  // torch l1_loss backward: sign(0) = 0
    if (y < ly0 || y >= ly1) sgn = 0.f;                       // outside the band's loss rows
    const float g = b[0][o] + 2.f * xv * b[1][o] + yv * b[2][o];
    dL_dimg[off] = k_l1 * sgn - k_ssim * g;
  }
}

size_t ssim_scratch_bytes(int n, int H, int W) {
  return 256 + align_up((size_t)n * 2 * sizeof(double), 256) + (size_t)3 * n * 3 * H * W * sizeof(float);
}

int launch_ssim_l1(int n, int H, int W, const float* img, const float* target, float w_l1, float w_ssim,
                   float grad_scale, float* loss_out, float* dL_dimg, void* scratch, cudaStream_t st,
                   const int32_t* rows = nullptr) {
  static const SsimWindow win = make_window();
  double* accum = reinterpret_cast<double*>(scratch);
  float* maps = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + align_up((size_t)n * 2 * sizeof(double), 256));
  GSB_CUDA_CHECK(cudaMemsetAsync(accum, 0, (size_t)n * 2 * sizeof(double), st));
  dim3 grid((W + SS_TW - 1) / SS_TW, (H + SS_TH - 1) / SS_TH, n * 3), block(SS_THREADS);
  const size_t planes = (size_t)n * 3;
  ssim_fwd_kernel<<<grid, block, 0, st>>>(H, W, img, target, win, accum, dL_dimg ? maps : nullptr, planes, rows);
  GSB_LAUNCH_CHECK();
  const double inv_count = 1.0 / ((double)3 * H * W);
  if (loss_out) {
    ssim_finalize_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, H, inv_count, w_l1, w_ssim, accum, rows, loss_out);
    GSB_LAUNCH_CHECK();
  }
  if (dL_dimg) {
    ssim_bwd_kernel<<<grid, block, 0, st>>>(H, W, img, target, win, maps, planes,
                                            (float)(grad_scale * w_l1 * inv_count),
                                            (float)(grad_scale * w_ssim * inv_count), dL_dimg, rows);
    GSB_LAUNCH_CHECK();
  }
  return GSB_OK;
}

}  // namespace gsb

extern "C" {
size_t gsb_ssim_scratch_bytes(int32_t n, int32_t height, int32_t width) {
  return gsb::ssim_scratch_bytes(n, height, width);
}
int gsb_ssim_l1(int32_t n, int32_t height, int32_t width, const float* img, const float* target, float w_l1,
                float w_ssim, float grad_scale, float* loss_out, float* dL_dimg, void* scratch, void* stream) {
  if (n <= 0 || height <= 0 || width <= 0 || !img || !target || !scratch) return GSB_EINVAL;
  return gsb::launch_ssim_l1(n, height, width, img, target, w_l1, w_ssim, grad_scale, loss_out, dL_dimg, scratch,
                             reinterpret_cast<cudaStream_t>(stream));
}
int gsb_ssim_l1_rows(int32_t n, int32_t height, int32_t width, const float* img, const float* target, const int32_t* rows,
                     float w_l1, float w_ssim, float grad_scale, float* loss_out, float* dL_dimg, void* scratch,
                     void* stream) {
  if (n <= 0 || height <= 0 || width <= 0 || !img || !target || !scratch) return GSB_EINVAL;
  return gsb::launch_ssim_l1(n, height, width, img, target, w_l1, w_ssim, grad_scale, loss_out, dL_dimg, scratch,
                             reinterpret_cast<cudaStream_t>(stream), rows);
}
}
