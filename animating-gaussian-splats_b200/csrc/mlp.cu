// mlp.cu -- the element-wise stages of the DeformationNetwork around the tcgen05 dense layers:
//   * BatchNorm1d in TRAINING mode (the reference never calls .eval(): train.py:57-78 and inference :550-640 both
//     normalise with the batch statistics of the current N Gaussians) + optional residual add + exact GELU, one
//     pass: y = gelu(gamma (z - mean) / sqrt(var + eps) + beta [+ identity]).  mean / var come from the column
//     sums the GEMM epilogue accumulated (gsb_linear_tf32 col_stats), so Z is read exactly once.  Block 0 also
//     performs the running-statistics update (momentum 0.1, unbiased variance), like torch.
//   * PositionalEncoding (train.py:113-127) of min/max-normalised inputs (train.py:186-205) written straight into
//     a column block of the network's [N, 192] input matrix.  The reference's second assignment reads the ALREADY
//     overwritten even entries, so odd entries are cos(sin(f x)), not cos(f x); that is reproduced.
// HBM-bound: 4 B read + 4 B written per element (+ 4 B for the residual).
#include "common.cuh"

namespace gsb {

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

__global__ void __launch_bounds__(256)
bn_gelu_kernel(long long rows, int N, const float* __restrict__ Z, long long ldz, const double* __restrict__ stats,
               const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
               const float* __restrict__ residual, long long ldr, float* __restrict__ Y, long long ldy,
               float* __restrict__ running_mean, float* __restrict__ running_var, float momentum, int apply_gelu) {
  __shared__ float s_scale[256], s_shift[256];
  for (int c = threadIdx.x; c < N; c += 256) {
    const double mean = stats[c] / (double)rows;
    double var = stats[N + c] / (double)rows - mean * mean;
    var = var < 0.0 ? 0.0 : var;
    const float inv = (float)(1.0 / sqrt(var + (double)eps));
    const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    s_scale[c] = g * inv;
    s_shift[c] = b - (float)mean * g * inv;
    if (blockIdx.x == 0 && running_mean) {
      const double unbiased = rows > 1 ? var * (double)rows / (double)(rows - 1) : var;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
  }
  __syncthreads();
  const int n4 = N >> 2;   // N % 4 == 0 (checked by the caller)
  const long long total = rows * n4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long r = i / n4;
    const int c = (int)(i - r * n4) << 2;
    float4 z = *reinterpret_cast<const float4*>(Z + r * ldz + c);
    float4 o;
    o.x = fmaf(z.x, s_scale[c], s_shift[c]);
    o.y = fmaf(z.y, s_scale[c + 1], s_shift[c + 1]);
    o.z = fmaf(z.z, s_scale[c + 2], s_shift[c + 2]);
    o.w = fmaf(z.w, s_scale[c + 3], s_shift[c + 3]);
    if (residual) {
      const float4 q = *reinterpret_cast<const float4*>(residual + r * ldr + c);
      o.x += q.x; o.y += q.y; o.z += q.z; o.w += q.w;
    }
    if (apply_gelu) { o.x = gelu_erf(o.x); o.y = gelu_erf(o.y); o.z = gelu_erf(o.z); o.w = gelu_erf(o.w); }
    *reinterpret_cast<float4*>(Y + r * ldy + c) = o;
  }
}

// out[r, col0 + j * D + d] for j in [0, 2F): j even -> sin(2^(j/2) pi u), j odd -> cos(sin(2^(j/2) pi u)),
// u = 2 (x[r, d] - lo[d]) / span[d] - 1 when lo/span are given (normalize_and_encode, train.py:186-205), else x.
__global__ void __launch_bounds__(256)
posenc_kernel(long long rows, int D, int F, const float* __restrict__ x, long long ldx, const float* __restrict__ lo,
              const float* __restrict__ span, float* __restrict__ out, long long ldo, int col0) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= rows * D) return;
  const long long r = i / D;
  const int d = (int)(i - r * D);
  float u = x[r * ldx + d];
  if (lo) u = 2.0f * (u - lo[d]) / span[d] - 1.0f;
  float* o = out + r * ldo + col0 + d;
  float f = 3.14159265358979323846f;   // torch.pi in float32
  for (int k = 0; k < F; k++) {
    const float s = sinf(f * u);
    o[(2 * k) * D] = s;
    o[(2 * k + 1) * D] = cosf(s);
    f *= 2.0f;
  }
}

}  // namespace gsb

using namespace gsb;

extern "C" {

int gsb_bn_gelu(int64_t rows, int32_t N, const float* Z, int64_t ldz, const double* col_stats, const float* gamma,
                const float* beta, float eps, const float* residual, int64_t ldr, float* Y, int64_t ldy,
                float* running_mean, float* running_var, float momentum, int32_t apply_gelu, void* stream) {
  if (rows < 0 || N <= 0 || N > 256 || (N & 3)) return GSB_EINVAL;
  if (rows == 0) return GSB_OK;
  if (!Z || !col_stats || !Y || (ldz & 3) || (ldy & 3) || (residual && (ldr & 3))) return GSB_EINVAL;
  if ((running_mean == nullptr) != (running_var == nullptr)) return GSB_EINVAL;
  const long long total = rows * (N >> 2);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  bn_gelu_kernel<<<(unsigned)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      rows, N, Z, ldz, col_stats, gamma, beta, eps, residual, ldr, Y, ldy, running_mean, running_var, momentum,
      apply_gelu);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_positional_encoding(int64_t rows, int32_t D, int32_t F, const float* x, int64_t ldx, const float* lo,
                            const float* span, float* out, int64_t ldo, int32_t col0, void* stream) {
  if (rows < 0 || D <= 0 || F <= 0 || F > 30) return GSB_EINVAL;
  if (rows == 0) return GSB_OK;
  if (!x || !out || ((lo == nullptr) != (span == nullptr))) return GSB_EINVAL;
  const long long total = rows * D;
  posenc_kernel<<<(unsigned)((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      rows, D, F, x, ldx, lo, span, out, ldo, col0);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // extern "C"
