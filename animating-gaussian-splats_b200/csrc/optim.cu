// optim.cu -- the two steps either side of the rasterizer in every training iteration:
//   * create_render_arguments (shared.py:29-42): rotations = normalize(q), opacities = sigmoid(logit),
//     scales = exp(log_scale) -- forward and backward as ONE kernel each (the reference runs three
//     elementwise torch kernels plus their autograd nodes per render, twice per iteration);
//   * the Adam update of create_optimizer (densify.py:68-86: torch.optim.Adam, per-table learning rates,
//     eps 1e-15, betas (0.9, 0.999), no weight decay / amsgrad) over all parameter tables in one launch.
// HBM-bound elementwise work: activation fwd 32 B in + 32 B out per Gaussian, bwd 64 B in + 32 B out;
// Adam 16 B read + 12 B written per parameter element.
#include <cmath>
#include "common.cuh"

namespace gsb {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// torch.nn.functional.normalize(q, p=2, dim=-1, eps=1e-12): q / max(||q||, eps)
__global__ void __launch_bounds__(256)
activate_fwd_kernel(int P, const float4* __restrict__ quats, const float* __restrict__ logits,
                    const float* __restrict__ log_scales, float4* __restrict__ rotations,
                    float* __restrict__ opacities, float* __restrict__ scales) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const float4 q = __ldg(quats + i);
  const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  const float d = fmaxf(n, 1e-12f);
  rotations[i] = make_float4(q.x / d, q.y / d, q.z / d, q.w / d);
  opacities[i] = sigmoidf_(__ldg(logits + i));
#pragma unroll
  for (int k = 0; k < 3; k++) scales[3 * (size_t)i + k] = expf(__ldg(log_scales + 3 * (size_t)i + k));
}

// Adjoint of the above.  y = q / d, d = max(||q||, eps):
//   ||q|| > eps : dq = (dy - y (y . dy)) / d        ||q|| <= eps : dq = dy / eps
//   dlogit = dop * s (1 - s),  dlog_scale = dscale * scale.
__global__ void __launch_bounds__(256)
activate_bwd_kernel(int P, const float4* __restrict__ quats, const float* __restrict__ opacities,
                    const float* __restrict__ scales, const float4* __restrict__ d_rot,
                    const float* __restrict__ d_opa, const float* __restrict__ d_sca, float4* __restrict__ d_quats,
                    float* __restrict__ d_logits, float* __restrict__ d_log_scales) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  if (d_quats) {
    const float4 q = __ldg(quats + i);
    const float4 g = d_rot ? __ldg(d_rot + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    float4 r;
    if (n > 1e-12f) {
      const float inv = 1.0f / n;
      const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
      const float dot = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
      r = make_float4((g.x - y.x * dot) * inv, (g.y - y.y * dot) * inv, (g.z - y.z * dot) * inv,
                      (g.w - y.w * dot) * inv);
    } else {
      r = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
    }
    d_quats[i] = r;
  }
  if (d_logits) {
    const float s = __ldg(opacities + i);
    d_logits[i] = d_opa ? __ldg(d_opa + i) * s * (1.0f - s) : 0.f;
  }
  if (d_log_scales) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const size_t o = 3 * (size_t)i + k;
      d_log_scales[o] = d_sca ? __ldg(d_sca + o) * __ldg(scales + o) : 0.f;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Multi-tensor Adam, the arithmetic of torch.optim.Adam's single-tensor path (torch/optim/adam.py):
//   m = m + (g - m)(1 - b1)            [exp_avg.lerp_(grad, 1 - beta1)]
//   v = v b2 + (1 - b2) g g            [exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)]
//   p = p - (lr / (1 - b1^t)) m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The bias corrections are computed on the host in double like Python does.
// ------------------------------------------------------------------------------------------------
struct AdamGroupDev {
  float* p; const float* g; float* m; float* v;
  long long n;
  float step_size, bc2_sqrt;
};
constexpr int ADAM_MAX_GROUPS = 16;
struct AdamArgs { AdamGroupDev grp[ADAM_MAX_GROUPS]; };

__global__ void __launch_bounds__(256)
adam_kernel(AdamArgs a, float w1, float beta2, float w2, float eps) {
  const AdamGroupDev G = a.grp[blockIdx.y];
  const long long n4 = G.n / 4;
  const bool vec = ((reinterpret_cast<uintptr_t>(G.p) | reinterpret_cast<uintptr_t>(G.g) |
                     reinterpret_cast<uintptr_t>(G.m) | reinterpret_cast<uintptr_t>(G.v)) & 15u) == 0;
  const long long stride = (long long)gridDim.x * 256;
  auto upd = [&](float& p, float g, float& m, float& v) {
    m = m + (g - m) * w1;
    v = v * beta2 + w2 * g * g;
    const float denom = sqrtf(v) / G.bc2_sqrt + eps;
    p = p - G.step_size * (m / denom);
  };
  if (vec) {
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += stride) {
      float4 p = reinterpret_cast<float4*>(G.p)[i];
      const float4 g = __ldg(reinterpret_cast<const float4*>(G.g) + i);
      float4 m = reinterpret_cast<float4*>(G.m)[i];
      float4 v = reinterpret_cast<float4*>(G.v)[i];
      upd(p.x, g.x, m.x, v.x); upd(p.y, g.y, m.y, v.y); upd(p.z, g.z, m.z, v.z); upd(p.w, g.w, m.w, v.w);
      reinterpret_cast<float4*>(G.p)[i] = p;
      reinterpret_cast<float4*>(G.m)[i] = m;
      reinterpret_cast<float4*>(G.v)[i] = v;
    }
    for (long long i = n4 * 4 + (long long)blockIdx.x * 256 + threadIdx.x; i < G.n; i += stride)
      upd(G.p[i], G.g[i], G.m[i], G.v[i]);
  } else {
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < G.n; i += stride)
      upd(G.p[i], G.g[i], G.m[i], G.v[i]);
  }
}

// ------------------------------------------------------------------------------------------------
// The optimiser step of the first-frame loop straight from the (exchanged) flat step record: activation adjoint
// (normalize / sigmoid / exp, the arithmetic of activate_bwd_kernel) + Adam (the arithmetic of adam_kernel) over
// the six per-Gaussian tables in ONE launch, thread per Gaussian.  The step count lives on the device (graph
// replays advance it; adam_tick_kernel), the bias corrections are evaluated in double per block.
// ------------------------------------------------------------------------------------------------
struct FlatAdamDev {
  float* p[6]; float* m[6]; float* v[6];   // means, colors, segmentation colours, quaternions, opacity logits, log scales
  double lr[6];
};

__global__ void adam_tick_kernel(long long* step) { *step += 1; }

__global__ void __launch_bounds__(256)
adam_flat_kernel(int P, FlatAdamDev a, const float* __restrict__ flat, int flat_stride, const long long* __restrict__ step,
                 double beta1, double beta2, float w1, float fbeta2, float w2, float eps) {
  __shared__ float s_step_size[6], s_bc2_sqrt;
  if (threadIdx.x == 0) {
    const double t = (double)*step;
    const double bc1 = 1.0 - pow(beta1, t), bc2 = 1.0 - pow(beta2, t);
    for (int k = 0; k < 6; k++) s_step_size[k] = (float)(a.lr[k] / bc1);
    s_bc2_sqrt = (float)sqrt(bc2);
  }
  __syncthreads();
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const float4* row = reinterpret_cast<const float4*>(flat) + (size_t)i * (flat_stride / 4);
  const float4 r0 = __ldg(row), r1 = __ldg(row + 1), r2 = __ldg(row + 2), r3 = __ldg(row + 3), r4 = __ldg(row + 4);
  const float bc2s = s_bc2_sqrt;
  auto upd = [&](int t, size_t o, float g) {
    float m = a.m[t][o], v = a.v[t][o];
    m = m + (g - m) * w1;
    v = v * fbeta2 + w2 * g * g;
    const float denom = sqrtf(v) / bc2s + eps;
    a.p[t][o] = a.p[t][o] - s_step_size[t] * (m / denom);
    a.m[t][o] = m;
    a.v[t][o] = v;
  };
  const size_t i3 = 3 * (size_t)i, i4 = 4 * (size_t)i;
  // activation adjoints first (they read the parameters the updates below overwrite)
  const float4 q = reinterpret_cast<const float4*>(a.p[3])[i];
  const float4 g = make_float4(r1.z, r1.w, r2.x, r2.y);
  const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  float4 dq;
  if (n > 1e-12f) {
    const float inv = 1.0f / n;
    const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
    const float dot = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
    dq = make_float4((g.x - y.x * dot) * inv, (g.y - y.y * dot) * inv, (g.z - y.z * dot) * inv, (g.w - y.w * dot) * inv);
  } else {
    dq = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
  }
  const float sg = sigmoidf_(a.p[4][i]);
  const float dlogit = r2.z * sg * (1.0f - sg);
  const float dls0 = r2.w * expf(a.p[5][i3]), dls1 = r3.x * expf(a.p[5][i3 + 1]), dls2 = r3.y * expf(a.p[5][i3 + 2]);
  upd(0, i3, r0.x); upd(0, i3 + 1, r0.y); upd(0, i3 + 2, r0.z);
  upd(1, i3, r0.w); upd(1, i3 + 1, r1.x); upd(1, i3 + 2, r1.y);
  upd(2, i3, r4.y); upd(2, i3 + 1, r4.z); upd(2, i3 + 2, r4.w);
  upd(3, i4, dq.x); upd(3, i4 + 1, dq.y); upd(3, i4 + 2, dq.z); upd(3, i4 + 3, dq.w);
  upd(4, (size_t)i, dlogit);
  upd(5, i3, dls0); upd(5, i3 + 1, dls1); upd(5, i3 + 2, dls2);
}

}  // namespace gsb

using namespace gsb;

extern "C" {

int gsb_adam_flat(const gsb_flat_adam_args* a, const float* flat, int32_t flat_stride, int64_t* step_dev, double beta1,
                  double beta2, double eps, void* stream) {
  if (!a || a->P < 0 || flat_stride < GSB_FLAT_WIDTH || (flat_stride & 3)) return GSB_EINVAL;
  if (a->P == 0) return GSB_OK;
  if (!flat || !step_dev || ((uintptr_t)flat & 15u)) return GSB_EINVAL;
  FlatAdamDev d;
  for (int k = 0; k < 6; k++) {
    if (!a->param[k] || !a->exp_avg[k] || !a->exp_avg_sq[k]) return GSB_EINVAL;
    d.p[k] = a->param[k]; d.m[k] = a->exp_avg[k]; d.v[k] = a->exp_avg_sq[k]; d.lr[k] = a->lr[k];
  }
  if ((uintptr_t)d.p[3] & 15u) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  adam_tick_kernel<<<1, 1, 0, st>>>(reinterpret_cast<long long*>(step_dev));
  GSB_LAUNCH_CHECK();
  adam_flat_kernel<<<(a->P + 255) / 256, 256, 0, st>>>(a->P, d, flat, flat_stride, reinterpret_cast<const long long*>(step_dev), beta1,
                                                        beta2, (float)(1.0 - beta1), (float)beta2, (float)(1.0 - beta2),
                                                        (float)eps);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_activate(int32_t P, const float* quats, const float* opacity_logits, const float* log_scales,
                 float* rotations, float* opacities, float* scales, void* stream) {
  if (P < 0) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  if (!quats || !opacity_logits || !log_scales || !rotations || !opacities || !scales) return GSB_EINVAL;
  activate_fwd_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, reinterpret_cast<const float4*>(quats), opacity_logits, log_scales, reinterpret_cast<float4*>(rotations),
      opacities, scales);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_activate_backward(int32_t P, const float* quats, const float* opacities, const float* scales,
                          const float* d_rotations, const float* d_opacities, const float* d_scales,
                          float* d_quats, float* d_opacity_logits, float* d_log_scales, void* stream) {
  if (P < 0) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  if ((d_quats && !quats) || (d_opacity_logits && !opacities) || (d_log_scales && !scales)) return GSB_EINVAL;
  activate_bwd_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, reinterpret_cast<const float4*>(quats), opacities, scales, reinterpret_cast<const float4*>(d_rotations),
      d_opacities, d_scales, reinterpret_cast<float4*>(d_quats), d_opacity_logits, d_log_scales);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_adam_step(int32_t n_groups, const gsb_adam_group* groups_host, double beta1, double beta2, double eps,
                  void* stream) {
  if (n_groups < 0 || n_groups > ADAM_MAX_GROUPS || (n_groups > 0 && !groups_host)) return GSB_EINVAL;
  AdamArgs a;
  int used = 0;
  long long max_n = 0;
  for (int i = 0; i < n_groups; i++) {
    const gsb_adam_group& h = groups_host[i];
    if (h.numel < 0 || h.step < 1) return GSB_EINVAL;
    if (h.numel == 0 || !h.grad) continue;   // torch skips parameters without a gradient
    if (!h.param || !h.exp_avg || !h.exp_avg_sq) return GSB_EINVAL;
    const double bc1 = 1.0 - pow(beta1, (double)h.step);
    const double bc2 = 1.0 - pow(beta2, (double)h.step);
    AdamGroupDev& d = a.grp[used++];
    d.p = h.param; d.g = h.grad; d.m = h.exp_avg; d.v = h.exp_avg_sq; d.n = h.numel;
    d.step_size = (float)(h.lr / bc1);
    d.bc2_sqrt = (float)sqrt(bc2);
    max_n = h.numel > max_n ? h.numel : max_n;
  }
  if (used == 0) return GSB_OK;
  const long long want = (max_n / 4 + 255) / 256;
  dim3 grid((unsigned)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want)), used);
  // Python evaluates 1 - beta in double and hands the result to the kernels as a float scalar
  adam_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(a, (float)(1.0 - beta1), (float)beta2,
                                                                         (float)(1.0 - beta2), (float)eps);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // extern "C"
