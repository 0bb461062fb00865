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

constexpr int SS_T = 16;          // tile edge
constexpr int SS_R = 5;           // window radius
constexpr int SS_H = SS_T + 2 * SS_R;  // 26
constexpr int SS_P = SS_H + 1;    // padded row pitch

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

__device__ __forceinline__ void load_halo(const float* __restrict__ img, int H, int W, int x0, int y0,
                                          float (*s)[SS_P], unsigned tid) {
  for (int e = tid; e < SS_H * SS_H; e += SS_T * SS_T) {
    const int r = e / SS_H, c = e - r * SS_H;
    const int y = y0 + r - SS_R, x = x0 + c - SS_R;
    s[r][c] = (y >= 0 && y < H && x >= 0 && x < W) ? __ldg(img + (size_t)y * W + x) : 0.f;
  }
}

__global__ void __launch_bounds__(SS_T * SS_T)
ssim_fwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ target, SsimWindow win,
                double* __restrict__ accum /* [n][2] */, float* __restrict__ maps /* [3][planes][H][W] or null */,
                size_t plane_count) {
  __shared__ float s_x[SS_H][SS_P];
  __shared__ float s_y[SS_H][SS_P];
  __shared__ float s_h[5][SS_H][SS_T + 1];
  __shared__ float s_red[2][SS_T * SS_T / 32];
  const unsigned tid = threadIdx.y * SS_T + threadIdx.x;
  const size_t plane = blockIdx.z;  // image * 3 + channel
  const size_t HW = (size_t)H * W;
  const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
  load_halo(img + plane * HW, H, W, x0, y0, s_x, tid);
  load_halo(target + plane * HW, H, W, x0, y0, s_y, tid);
  __syncthreads();
  // horizontal pass: SS_H rows x SS_T columns
  for (int e = tid; e < SS_H * SS_T; e += SS_T * SS_T) {
    const int r = e / SS_T, c = e - r * SS_T;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, a4 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
      const float g = win.g[k];
      const float x = s_x[r][c + k], y = s_y[r][c + k];
      a0 = fmaf(g, x, a0);
      a1 = fmaf(g, y, a1);
      a2 = fmaf(g, x * x, a2);
      a3 = fmaf(g, y * y, a3);
      a4 = fmaf(g, x * y, a4);
    }
    s_h[0][r][c] = a0; s_h[1][r][c] = a1; s_h[2][r][c] = a2; s_h[3][r][c] = a3; s_h[4][r][c] = a4;
  }
  __syncthreads();
  const int tx = threadIdx.x, ty = threadIdx.y;
  float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
  for (int k = 0; k < 11; k++) {
    const float g = win.g[k];
    mu1 = fmaf(g, s_h[0][ty + k][tx], mu1);
    mu2 = fmaf(g, s_h[1][ty + k][tx], mu2);
    e11 = fmaf(g, s_h[2][ty + k][tx], e11);
    e22 = fmaf(g, s_h[3][ty + k][tx], e22);
    e12 = fmaf(g, s_h[4][ty + k][tx], e12);
  }
  const int x = x0 + tx, y = y0 + ty;
  const bool inside = x < W && y < H;
  float ssim = 0.f, l1 = 0.f;
  if (inside) {
    const float c1 = 0.0001f, c2 = 0.0009f;  // 0.01^2, 0.03^2 (external.py:96-97)
    const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
    const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
    const float An = 2.f * mu12 + c1, Bn = 2.f * s12 + c2, Cd = mu1_sq + mu2_sq + c1, Dd = s1 + s2 + c2;
    const float inv = 1.0f / (Cd * Dd);
    ssim = An * Bn * inv;
    l1 = fabsf(s_x[ty + SS_R][tx + SS_R] - s_y[ty + SS_R][tx + SS_R]);
    if (maps) {
      // d ssim / d mu1, d E[x^2], d E[xy]   (sigma1^2 = E[x^2] - mu1^2, sigma12 = E[xy] - mu1 mu2)
      const float d_mu1 = (2.f * mu2 * Bn - 2.f * mu2 * An) * inv - ssim * (2.f * mu1 * Dd - 2.f * mu1 * Cd) * inv;
      const float d_e11 = -ssim * Cd * inv;
      const float d_e12 = 2.f * An * inv;
      const size_t o = plane * HW + (size_t)y * W + x;
      maps[o] = d_mu1;
      maps[plane_count * HW + o] = d_e11;
      maps[2 * plane_count * HW + o] = d_e12;
    }
  }
  // block reduction -> one double atomic per quantity per block
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    ssim += __shfl_xor_sync(0xffffffffu, ssim, s);
    l1 += __shfl_xor_sync(0xffffffffu, l1, s);
  }
  if ((tid & 31u) == 0) { s_red[0][tid >> 5] = l1; s_red[1][tid >> 5] = ssim; }
  __syncthreads();
  if (tid < 2) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < SS_T * SS_T / 32; w++) t += (double)s_red[tid][w];
    atomicAdd(accum + 2 * (plane / 3) + tid, t);
  }
}

__global__ void ssim_finalize_kernel(int n, double inv_count, float w_l1, float w_ssim,
                                     const double* __restrict__ accum, float* __restrict__ loss_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double l1 = accum[2 * i] * inv_count, ss = accum[2 * i + 1] * inv_count;
  loss_out[3 * i] = (float)l1;
  loss_out[3 * i + 1] = (float)ss;
  loss_out[3 * i + 2] = (float)((double)w_l1 * l1 + (double)w_ssim * (1.0 - ss));
}

__global__ void __launch_bounds__(SS_T * SS_T)
ssim_bwd_kernel(int H, int W, const float* __restrict__ img, const float* __restrict__ target, SsimWindow win,
                const float* __restrict__ maps, size_t plane_count, float k_l1, float k_ssim,
                float* __restrict__ dL_dimg) {
  __shared__ float s_m[3][SS_H][SS_P];
  __shared__ float s_h[3][SS_H][SS_T + 1];
  const unsigned tid = threadIdx.y * SS_T + threadIdx.x;
  const size_t plane = blockIdx.z;
  const size_t HW = (size_t)H * W;
  const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
#pragma unroll
  for (int m = 0; m < 3; m++) load_halo(maps + (m * plane_count + plane) * HW, H, W, x0, y0, s_m[m], tid);
  __syncthreads();
  for (int e = tid; e < SS_H * SS_T; e += SS_T * SS_T) {
    const int r = e / SS_T, c = e - r * SS_T;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
      const float g = win.g[k];
      a0 = fmaf(g, s_m[0][r][c + k], a0);
      a1 = fmaf(g, s_m[1][r][c + k], a1);
      a2 = fmaf(g, s_m[2][r][c + k], a2);
    }
    s_h[0][r][c] = a0; s_h[1][r][c] = a1; s_h[2][r][c] = a2;
  }
  __syncthreads();
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int x = x0 + tx, y = y0 + ty;
  if (x >= W || y >= H) return;
  float b0 = 0.f, b1 = 0.f, b2 = 0.f;
#pragma unroll
  for (int k = 0; k < 11; k++) {
    const float g = win.g[k];
    b0 = fmaf(g, s_h[0][ty + k][tx], b0);
    b1 = fmaf(g, s_h[1][ty + k][tx], b1);
    b2 = fmaf(g, s_h[2][ty + k][tx], b2);
  }
  const size_t o = plane * HW + (size_t)y * W + x;
  const float xv = __ldg(img + o), yv = __ldg(target + o);
  const float d = xv - yv;
  const float sgn = (d > 0.f) ? 1.f : ((d < 0.f) ? -1.f : 0.f);  // torch l1_loss backward: sign(0) = 0
  const float g = b0 + 2.f * xv * b1 + yv * b2;
  dL_dimg[o] = k_l1 * sgn - k_ssim * g;
}

size_t ssim_scratch_bytes(int n, int H, int W) {
  return 256 + align_up((size_t)n * 2 * sizeof(double), 256) + (size_t)3 * n * 3 * H * W * sizeof(float);
}

int launch_ssim_l1(int n, int H, int W, const float* img, const float* target, float w_l1, float w_ssim,
                   float grad_scale, float* loss_out, float* dL_dimg, void* scratch, cudaStream_t st) {
  static const SsimWindow win = make_window();
  double* accum = reinterpret_cast<double*>(scratch);
  float* maps = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + align_up((size_t)n * 2 * sizeof(double), 256));
  GSB_CUDA_CHECK(cudaMemsetAsync(accum, 0, (size_t)n * 2 * sizeof(double), st));
  dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, n * 3), block(SS_T, SS_T);
  const size_t planes = (size_t)n * 3;
  ssim_fwd_kernel<<<grid, block, 0, st>>>(H, W, img, target, win, accum, dL_dimg ? maps : nullptr, planes);
  GSB_LAUNCH_CHECK();
  const double inv_count = 1.0 / ((double)3 * H * W);
  if (loss_out) {
    ssim_finalize_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, inv_count, w_l1, w_ssim, accum, loss_out);
    GSB_LAUNCH_CHECK();
  }
  if (dL_dimg) {
    ssim_bwd_kernel<<<grid, block, 0, st>>>(H, W, img, target, win, maps, planes,
                                            (float)(grad_scale * w_l1 * inv_count),
                                            (float)(grad_scale * w_ssim * inv_count), dL_dimg);
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
}
