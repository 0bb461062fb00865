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
bn_gelu_kernel(long long rows, long long batch_rows, int N, const float* __restrict__ Z, long long ldz, const double* __restrict__ stats,
               const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
               const float* __restrict__ residual, long long ldr, float* __restrict__ Y, long long ldy,
               float* __restrict__ running_mean, float* __restrict__ running_var, float momentum, int apply_gelu) {
  __shared__ float s_scale[256], s_shift[256];
  for (int c = threadIdx.x; c < N; c += 256) {
    // the statistics cover batch_rows rows (>= rows: the rows of all ranks when the batch is sharded by rows)
    const double mean = stats[c] / (double)batch_rows;
    double var = stats[N + c] / (double)batch_rows - mean * mean;
    var = var < 0.0 ? 0.0 : var;
    const float inv = (float)(1.0 / sqrt(var + (double)eps));
    const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
    s_scale[c] = g * inv;
    s_shift[c] = b - (float)mean * g * inv;
    if (blockIdx.x == 0 && running_mean) {
      const double unbiased = batch_rows > 1 ? var * (double)batch_rows / (double)(batch_rows - 1) : var;
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

// ------------------------------------------------------------------------------------------------
// Backward of the BatchNorm (training mode) + residual + GELU stage, two passes over [rows, N]:
//   pass 1 (bn_act_bwd_reduce):  g = dy * act'(s z + t [+ r]);  writes g, accumulates per column sum(g) and
//                                sum(g xhat), xhat = (z - mean) inv_std  -> dbeta, dgamma
//   pass 2 (bn_bwd_apply):       dz = gamma inv_std (g - dbeta / rows - xhat dgamma / rows)     (in place over g)
// g is also the gradient of the residual input.  Threads keep a fixed column group (4 columns) and stride over
// rows, so the column sums are per-thread registers, one shared-memory reduction and 2 N double atomics per block.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float gelu_erf_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  return cdf + x * 0.39894228040143267794f * expf(-0.5f * x * x);
}

struct BnCols { float mean, inv_std, scale, shift; };

__device__ __forceinline__ BnCols bn_col(const double* stats, int N, int c, long long rows, const float* gamma,
                                         const float* beta, float eps) {
  const double mean = stats[c] / (double)rows;
  double var = stats[N + c] / (double)rows - mean * mean;
  var = var < 0.0 ? 0.0 : var;
  BnCols b;
  b.mean = (float)mean;
  b.inv_std = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[c] : 1.f;
  b.scale = g * b.inv_std;
  b.shift = (beta ? beta[c] : 0.f) - b.mean * b.scale;
  return b;
}

__global__ void __launch_bounds__(256)
bn_act_bwd_reduce_kernel(long long rows, long long batch_rows, int N, const float* __restrict__ dY, long long ldd, const float* __restrict__ Z,
                         long long ldz, const double* __restrict__ stats, const float* __restrict__ gamma,
                         const float* __restrict__ beta, float eps, const float* __restrict__ residual, long long ldr,
                         int apply_gelu, float* __restrict__ G, long long ldg, double* __restrict__ sums) {
  __shared__ float s_red[2][256][4];
  const int n4 = N >> 2;
  const int row_lanes = blockDim.x / n4;              // blockDim.x = n4 * row_lanes
  const int cg = threadIdx.x % n4, rl = threadIdx.x / n4;
  const int c = cg << 2;
  BnCols b[4];
#pragma unroll
  for (int e = 0; e < 4; e++) b[e] = bn_col(stats, N, c + e, batch_rows, gamma, beta, eps);
  float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
  for (long long r = (long long)blockIdx.x * row_lanes + rl; r < rows; r += (long long)gridDim.x * row_lanes) {
    const float4 z = *reinterpret_cast<const float4*>(Z + r * ldz + c);
    const float4 dy = *reinterpret_cast<const float4*>(dY + r * ldd + c);
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (residual) q = *reinterpret_cast<const float4*>(residual + r * ldr + c);
    float g[4];
    const float zz[4] = {z.x, z.y, z.z, z.w}, dd[4] = {dy.x, dy.y, dy.z, dy.w}, qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const float pre = fmaf(zz[e], b[e].scale, b[e].shift) + qq[e];
      g[e] = apply_gelu ? dd[e] * gelu_erf_grad(pre) : dd[e];
      s1[e] += g[e];
      s2[e] = fmaf(g[e], (zz[e] - b[e].mean) * b[e].inv_std, s2[e]);
    }
    *reinterpret_cast<float4*>(G + r * ldg + c) = make_float4(g[0], g[1], g[2], g[3]);
  }
#pragma unroll
  for (int e = 0; e < 4; e++) { s_red[0][threadIdx.x][e] = s1[e]; s_red[1][threadIdx.x][e] = s2[e]; }
  __syncthreads();
  if (rl == 0) {
#pragma unroll
    for (int e = 0; e < 4; e++) {
      double a1 = 0.0, a2 = 0.0;
      for (int k = 0; k < row_lanes; k++) { a1 += (double)s_red[0][k * n4 + cg][e]; a2 += (double)s_red[1][k * n4 + cg][e]; }
      atomicAdd(sums + c + e, a1);
      atomicAdd(sums + N + c + e, a2);
    }
  }
}

__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(long long rows, long long batch_rows, int N, const float* G, long long ldg, float* dZ, long long lddz,
                    const float* __restrict__ Z, long long ldz, const double* __restrict__ stats, const float* __restrict__ gamma, float eps,
                    const double* __restrict__ sums, float* __restrict__ d_gamma, float* __restrict__ d_beta) {
  __shared__ float s_k0[256], s_k1[256], s_k2[256];   // dz = k0 g + k1 + k2 z
  for (int c = threadIdx.x; c < N; c += 256) {
    const BnCols b = bn_col(stats, N, c, batch_rows, gamma, nullptr, eps);
    const float db = (float)(sums[c] / (double)batch_rows), dg = (float)(sums[N + c] / (double)batch_rows);
    // dz = scale (g - db - xhat dg), xhat = (z - mean) inv_std
    s_k0[c] = b.scale;
    s_k2[c] = -b.scale * dg * b.inv_std;
    s_k1[c] = -b.scale * db - s_k2[c] * b.mean;
    if (blockIdx.x == 0) {
      if (d_gamma) d_gamma[c] = (float)sums[N + c];
      if (d_beta) d_beta[c] = (float)sums[c];
    }
  }
  __syncthreads();
  const int n4 = N >> 2;
  const long long total = rows * n4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long r = i / n4;
    const int c = (int)(i - r * n4) << 2;
    float4 g = *reinterpret_cast<const float4*>(G + r * ldg + c);
    const float4 z = *reinterpret_cast<const float4*>(Z + r * ldz + c);
    g.x = fmaf(s_k0[c], g.x, fmaf(s_k2[c], z.x, s_k1[c]));
    g.y = fmaf(s_k0[c + 1], g.y, fmaf(s_k2[c + 1], z.y, s_k1[c + 1]));
    g.z = fmaf(s_k0[c + 2], g.z, fmaf(s_k2[c + 2], z.z, s_k1[c + 2]));
    g.w = fmaf(s_k0[c + 3], g.w, fmaf(s_k2[c + 3], z.w, s_k1[c + 3]));
    *reinterpret_cast<float4*>(dZ + r * lddz + c) = g;   // dZ may alias G (same element, read before written)
  }
}


// ------------------------------------------------------------------------------------------------
// The two backward products of a THIN dense layer (the network's output layer: 7 columns, train.py:80-110).  Seven is
// not a tensor-core tile and both products are pure streaming: FP32 FMAs, every operand read once.
//   thin_dx:    dX[rows, H] = G[rows, M] . W[M, H]            (M <= 8): writes rows x H x 4 B
//   thin_wgrad: dW[M, H] = G^T . X[rows, H], db[M] = sum_rows G: reads rows x H x 4 B; per-block partial sums over a
//               contiguous row range (FP32), summed over the blocks by a second kernel in double
// ------------------------------------------------------------------------------------------------
constexpr int THIN_MAX_M = 8;

__global__ void __launch_bounds__(256)
thin_dx_kernel(long long rows, int M, int H, const float* __restrict__ G, long long ldg, const float* __restrict__ W,
               long long ldw, float* __restrict__ dX, long long ldx) {
  extern __shared__ float s_w[];   // [M][H]
  for (int i = threadIdx.x; i < M * H; i += 256) s_w[i] = W[(long long)(i / H) * ldw + (i % H)];
  __syncthreads();
  // a thread owns one group of 4 columns; 256 / (H / 4) rows per block pass (H <= 1024)
  const int h4 = H >> 2, lanes = 256 / h4;
  const int cg = threadIdx.x % h4, rl = threadIdx.x / h4;
  if (rl >= lanes) return;
  float4 w[THIN_MAX_M];
#pragma unroll
  for (int m = 0; m < THIN_MAX_M; m++)
    w[m] = m < M ? *reinterpret_cast<const float4*>(s_w + m * H + 4 * cg) : make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long r = (long long)blockIdx.x * lanes + rl; r < rows; r += (long long)gridDim.x * lanes) {
    const float* g = G + r * ldg;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int m = 0; m < THIN_MAX_M; m++) {
      if (m < M) {
        const float gm = __ldg(g + m);
        acc.x = fmaf(gm, w[m].x, acc.x); acc.y = fmaf(gm, w[m].y, acc.y); acc.z = fmaf(gm, w[m].z, acc.z); acc.w = fmaf(gm, w[m].w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(dX + r * ldx + 4 * cg) = acc;
  }
}

constexpr int THIN_WG_ROWS = 64;   // rows of G staged per round

__global__ void __launch_bounds__(256)
thin_wgrad_kernel(long long rows, int M, int H, const float* __restrict__ G, long long ldg, const float* __restrict__ X,
                  long long ldx, float* __restrict__ partial /* [gridDim.x][M + 1][H] */) {
  __shared__ __align__(16) float s_g[THIN_WG_ROWS][THIN_MAX_M];
  extern __shared__ float s_acc[];   // [M][H] + [THIN_MAX_M]: the block's sums over its row lanes
  const int parts = gridDim.x;
  const long long r0 = rows * blockIdx.x / parts, r1 = rows * (blockIdx.x + 1) / parts;
  // a thread owns one group of 4 columns and every (256 / (H / 4))-th row of the block's range
  const int h4 = H >> 2, lanes = 256 / h4;
  const int cg = threadIdx.x % h4, rl = threadIdx.x / h4;
  const bool active = rl < lanes;
  for (int i = threadIdx.x; i < M * H + THIN_MAX_M; i += 256) s_acc[i] = 0.f;
  float4 acc[THIN_MAX_M];
#pragma unroll
  for (int m = 0; m < THIN_MAX_M; m++) acc[m] = make_float4(0.f, 0.f, 0.f, 0.f);
  float gsum = 0.f;   // column threadIdx.x % 8 of G over the elements this thread stages (the bias gradient)
  for (long long rb = r0; rb < r1; rb += THIN_WG_ROWS) {
    const int nr = (int)((r1 - rb) < THIN_WG_ROWS ? (r1 - rb) : THIN_WG_ROWS);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < THIN_WG_ROWS * THIN_MAX_M / 256; k++) {
      const int i = threadIdx.x + 256 * k;
      const int r = i / THIN_MAX_M, m = i % THIN_MAX_M;
      const float v = (r < nr && m < M) ? G[(rb + r) * ldg + m] : 0.f;
      s_g[r][m] = v;
      gsum += v;
    }
    __syncthreads();
    if (active) {
      const float* xp = X + rb * ldx + 4 * cg;
#pragma unroll 4
      for (int r = rl; r < nr; r += lanes) {
        const float4 x = *reinterpret_cast<const float4*>(xp + (long long)r * ldx);
        const float4 g0 = *reinterpret_cast<const float4*>(&s_g[r][0]);
        const float4 g1 = *reinterpret_cast<const float4*>(&s_g[r][4]);
        const float gm[THIN_MAX_M] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
        for (int m = 0; m < THIN_MAX_M; m++) {
          acc[m].x = fmaf(gm[m], x.x, acc[m].x); acc[m].y = fmaf(gm[m], x.y, acc[m].y);
          acc[m].z = fmaf(gm[m], x.z, acc[m].z); acc[m].w = fmaf(gm[m], x.w, acc[m].w);
        }
      }
    }
  }
  if (active) {
#pragma unroll
    for (int m = 0; m < THIN_MAX_M; m++)
      if (m < M) {
        float* d = s_acc + m * H + 4 * cg;
        atomicAdd(d, acc[m].x); atomicAdd(d + 1, acc[m].y); atomicAdd(d + 2, acc[m].z); atomicAdd(d + 3, acc[m].w);
      }
  }
  gsum += __shfl_xor_sync(0xffffffffu, gsum, 8);
  gsum += __shfl_xor_sync(0xffffffffu, gsum, 16);
  if ((threadIdx.x & 31u) < THIN_MAX_M) atomicAdd(s_acc + M * H + (threadIdx.x & 7u), gsum);
  __syncthreads();
  float* out = partial + (size_t)blockIdx.x * (M + 1) * H;
  for (int i = threadIdx.x; i < M * H; i += 256) out[i] = s_acc[i];
  if (threadIdx.x < M) out[(size_t)M * H + threadIdx.x] = s_acc[M * H + threadIdx.x];
}

// 32 consecutive output elements per block; the 8 thread rows each sum every 8th partial (coalesced 128-byte reads,
// independent chains), then meet in shared memory
__global__ void __launch_bounds__(256)
thin_wgrad_reduce_kernel(int parts, int M, int H, const float* __restrict__ partial, float* __restrict__ dW,
                         float* __restrict__ db) {
  __shared__ double s_sum[8][33];
  const int n = M * H + M;
  const int x = threadIdx.x & 31, y = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + x;
  double s = 0.0;
  if (i < n)
    for (int c = y; c < parts; c += 8) s += (double)partial[(size_t)c * (M + 1) * H + i];
  s_sum[y][x] = s;
  __syncthreads();
  if (y == 0 && i < n) {
#pragma unroll
    for (int k = 1; k < 8; k++) s += s_sum[k][x];
    if (i < M * H) dW[i] = (float)s;
    else if (db) db[i - M * H] = (float)s;
  }
}

}  // namespace gsb

using namespace gsb;

extern "C" {

int gsb_bn_gelu(int64_t rows, int64_t batch_rows, int32_t N, const float* Z, int64_t ldz, const double* col_stats,
                const float* gamma, const float* beta, float eps, const float* residual, int64_t ldr, float* Y, int64_t ldy,
                float* running_mean, float* running_var, float momentum, int32_t apply_gelu, void* stream) {
  if (batch_rows == 0) batch_rows = rows;
  if (rows < 0 || batch_rows < rows || N <= 0 || N > 256 || (N & 3)) return GSB_EINVAL;
  if (rows == 0) return GSB_OK;
  if (!Z || !col_stats || !Y || (ldz & 3) || (ldy & 3) || (residual && (ldr & 3))) return GSB_EINVAL;
  if ((running_mean == nullptr) != (running_var == nullptr)) return GSB_EINVAL;
  const long long total = rows * (N >> 2);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  bn_gelu_kernel<<<(unsigned)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      rows, batch_rows, N, Z, ldz, col_stats, gamma, beta, eps, residual, ldr, Y, ldy, running_mean, running_var, momentum,
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

int gsb_bn_gelu_backward(int64_t rows, int64_t batch_rows, int32_t phase, int32_t N, const float* dY, int64_t ldd,
                         const float* Z, int64_t ldz, const double* col_stats, const float* gamma, const float* beta,
                         float eps, const float* residual, int64_t ldr, int32_t apply_gelu, float* G, int64_t ldg,
                         float* dZ, int64_t lddz, double* sums, float* d_gamma, float* d_beta, void* stream) {
  if (batch_rows == 0) batch_rows = rows;
  if (rows < 0 || batch_rows < rows || phase < 0 || phase > 2 || N <= 0 || N > 256 || (N & 3)) return GSB_EINVAL;
  if (!sums) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (phase != 2) GSB_CUDA_CHECK(cudaMemsetAsync(sums, 0, 2 * (size_t)N * sizeof(double), st));
  if (rows == 0 && batch_rows == 0) return GSB_OK;
  if (rows > 0 && (!dY || !Z || !col_stats || !G || !dZ)) return GSB_EINVAL;
  if ((ldd & 3) || (ldz & 3) || (ldg & 3) || (lddz & 3) || (residual && (ldr & 3))) return GSB_EINVAL;
  const int n4 = N >> 2;
  const int row_lanes = 256 / n4;
  if (phase != 2 && rows > 0) {
    bn_act_bwd_reduce_kernel<<<148 * 4, n4 * row_lanes, 0, st>>>(rows, batch_rows, N, dY, ldd, Z, ldz, col_stats, gamma,
                                                                  beta, eps, residual, ldr, apply_gelu, G, ldg, sums);
    GSB_LAUNCH_CHECK();
  }
  if (phase != 1) {
    // a rank without rows still runs one block: it writes d_gamma / d_beta from the (exchanged) sums
    const long long total = rows * n4;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    if (!col_stats) return GSB_EINVAL;
    bn_bwd_apply_kernel<<<(unsigned)blocks, 256, 0, st>>>(rows, batch_rows, N, G, ldg, dZ, lddz, Z, ldz, col_stats, gamma,
                                                          eps, sums, d_gamma, d_beta);
    GSB_LAUNCH_CHECK();
  }
  return GSB_OK;
}

int gsb_linear_thin_dx(int64_t rows, int32_t M, int32_t H, const float* G, int64_t ldg, const float* W, int64_t ldw,
                       float* dX, int64_t ldx, void* stream) {
  if (rows < 0 || M <= 0 || M > THIN_MAX_M || H <= 0 || (H & 3) || H > 1024) return GSB_EINVAL;
  if (rows == 0) return GSB_OK;
  if (!G || !W || !dX || ldg < M || ldw < H || ldx < H || (ldx & 3) || (reinterpret_cast<uintptr_t>(dX) & 15u)) return GSB_EINVAL;
  const long long total = rows * (H >> 2);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  thin_dx_kernel<<<(unsigned)blocks, 256, (size_t)M * H * sizeof(float), reinterpret_cast<cudaStream_t>(stream)>>>(
      rows, M, H, G, ldg, W, ldw, dX, ldx);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

static int thin_wgrad_parts(int64_t rows) {
  const int64_t want = (rows + THIN_WG_ROWS - 1) / THIN_WG_ROWS;
  return (int)(want < 148 * 4 ? (want > 0 ? want : 1) : 148 * 4);
}
size_t gsb_linear_thin_wgrad_scratch_bytes(int32_t M, int32_t H) {
  if (M <= 0 || H <= 0) return 0;
  return (size_t)148 * 4 * (M + 1) * H * sizeof(float);
}
int gsb_linear_thin_wgrad(int64_t rows, int32_t M, int32_t H, const float* G, int64_t ldg, const float* X, int64_t ldx,
                          float* dW, float* db, void* scratch, void* stream) {
  if (rows < 0 || M <= 0 || M > THIN_MAX_M || H < THIN_MAX_M || (H & 3) || H > 1024) return GSB_EINVAL;
  if (!dW || !scratch) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (rows == 0) {
    GSB_CUDA_CHECK(cudaMemsetAsync(dW, 0, (size_t)M * H * sizeof(float), st));
    if (db) GSB_CUDA_CHECK(cudaMemsetAsync(db, 0, (size_t)M * sizeof(float), st));
    return GSB_OK;
  }
  if (!G || !X || ldg < M || ldx < H || (ldx & 3) || (reinterpret_cast<uintptr_t>(X) & 15u)) return GSB_EINVAL;
  const int parts = thin_wgrad_parts(rows);
  thin_wgrad_kernel<<<parts, 256, ((size_t)M * H + THIN_MAX_M) * sizeof(float), st>>>(rows, M, H, G, ldg, X, ldx,
                                                                                    reinterpret_cast<float*>(scratch));
  GSB_LAUNCH_CHECK();
  const int n = M * H + M;
  thin_wgrad_reduce_kernel<<<(n + 31) / 32, 256, 0, st>>>(parts, M, H, reinterpret_cast<const float*>(scratch), dW, db);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // extern "C"
