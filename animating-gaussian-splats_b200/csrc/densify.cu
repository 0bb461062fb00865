// densify.cu -- densification statistics and stable stream compaction of the parameter tables.
//
// gsb_densify_accumulate fuses update_max_2d_radii_and_visibility_mask (densify.py:154-162, three
// masked-index kernels each hiding a nonzero() sync) with accumulate_mean_2d_gradients
// (external.py:113-124) into one pass: 24 B/Gaussian.
// gsb_mask_scan + gsb_gather_rows replace the boolean-mask indexing / torch.cat re-allocation of
// remove_points and cat_params_to_optimizer (external.py:127-204): an exclusive rank of the kept rows
// (the single-pass scan of binning.cu) followed by an order-preserving row gather, which is what makes
// the compaction stable.
#include "common.cuh"

namespace gsb {

__global__ void __launch_bounds__(256)
densify_accumulate_kernel(int P, const int32_t* __restrict__ radii, const float* __restrict__ grad_means2D,
                          float* __restrict__ max_2d_radii, float* __restrict__ grad_accum,
                          float* __restrict__ vis_count, uint8_t* __restrict__ visible) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const int r = radii[i];
  const bool vis = r > 0;
  if (visible) visible[i] = vis ? 1 : 0;
  if (!vis) return;
  const float rf = (float)r;
  if (max_2d_radii) max_2d_radii[i] = fmaxf(rf, max_2d_radii[i]);
  const float gx = grad_means2D[3 * (size_t)i], gy = grad_means2D[3 * (size_t)i + 1];
  grad_accum[i] += sqrtf(gx * gx + gy * gy);  // torch.norm(dim=-1) of the first two components
  vis_count[i] += 1.0f;
}

// The same over all cameras of a batch in one launch: thread per Gaussian, cameras in order (so the
// float accumulation order equals n_cams successive single-camera calls).
__global__ void __launch_bounds__(256)
densify_accumulate_batch_kernel(int P, int n_cams, const int32_t* __restrict__ radii,
                                const float* __restrict__ grad_means2D, float* __restrict__ max_2d_radii,
                                float* __restrict__ grad_accum, float* __restrict__ vis_count,
                                uint8_t* __restrict__ visible) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  float mr = max_2d_radii ? max_2d_radii[i] : 0.f, acc = grad_accum[i], cnt = vis_count[i];
  bool vis = false;
  for (int c = 0; c < n_cams; c++) {
    const size_t o = (size_t)c * P + i;
    const int r = radii[o];
    vis = r > 0;
    if (!vis) continue;
    mr = fmaxf((float)r, mr);
    const float gx = grad_means2D[3 * o], gy = grad_means2D[3 * o + 1];
    acc += sqrtf(gx * gx + gy * gy);
    cnt += 1.0f;
  }
  if (max_2d_radii) max_2d_radii[i] = mr;
  grad_accum[i] = acc;
  vis_count[i] = cnt;
  if (visible) visible[i] = vis ? 1 : 0;  // mask of the last camera, like repeated single-camera calls
}

__global__ void __launch_bounds__(256)
mask_to_u32_kernel(int P, const uint8_t* __restrict__ keep, uint32_t* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < P) out[i] = keep[i] ? 1u : 0u;
}

__global__ void __launch_bounds__(256)
mask_rank_finish_kernel(int P, const uint8_t* __restrict__ keep, int32_t* __restrict__ idx /* in: inclusive */,
                        int32_t* __restrict__ count_out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const int32_t inc = idx[i];
  if (i == P - 1 && count_out) *count_out = inc;
  idx[i] = inc - (keep[i] ? 1 : 0);
}

__global__ void __launch_bounds__(256)
gather_rows_kernel(long long total, int row_floats, const uint8_t* __restrict__ keep,
                   const int32_t* __restrict__ dst_index, const float* __restrict__ src, float* __restrict__ dst) {
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  if (e >= total) return;
  const long long row = e / row_floats;
  const int k = (int)(e - row * row_floats);
  if (!keep[row]) return;
  dst[(long long)dst_index[row] * row_floats + k] = src[e];
}

// accumulate_mean_2d_gradients (external.py:113-124) with the visibility mask as the reference passes it
__global__ void __launch_bounds__(256)
densify_accumulate_masked_kernel(int P, const uint8_t* __restrict__ visible, const float* __restrict__ grad_means2D,
                                 float* __restrict__ grad_accum, float* __restrict__ vis_count) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P || !visible[i]) return;
  const float gx = grad_means2D[3 * (size_t)i], gy = grad_means2D[3 * (size_t)i + 1];
  grad_accum[i] += sqrtf(gx * gx + gy * gy);
  vis_count[i] += 1.0f;
}

// Split (external.py:251-271): selected row i with rank r produces rows r and S + r ([sel; sel], repeat(2,1)):
//   mean' = mean + R(normalize(q)) * sample,   log_scale' = log(exp(log_scale) / 1.6)
__global__ void __launch_bounds__(256)
split_rows_kernel(int P, int S, const uint8_t* __restrict__ sel, const int32_t* __restrict__ rank,
                  const float* __restrict__ means, const float* __restrict__ quats,
                  const float* __restrict__ log_scales, const float* __restrict__ samples,
                  float* __restrict__ out_means, float* __restrict__ out_log_scales) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P || !sel[i]) return;
  const int r = rank[i];
  float4 q = reinterpret_cast<const float4*>(quats)[i];
  const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);  // build_rotation, external.py:27-46
  q.x /= norm; q.y /= norm; q.z /= norm; q.w /= norm;
  const float rr = q.x, x = q.y, y = q.z, z = q.w;
  const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - rr * z), R02 = 2.f * (x * z + rr * y);
  const float R10 = 2.f * (x * y + rr * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - rr * x);
  const float R20 = 2.f * (x * z - rr * y), R21 = 2.f * (y * z + rr * x), R22 = 1.f - 2.f * (x * x + y * y);
  const float mx = means[3 * (size_t)i], my = means[3 * (size_t)i + 1], mz = means[3 * (size_t)i + 2];
  float ls[3];
#pragma unroll
  for (int k = 0; k < 3; k++) ls[k] = logf(expf(log_scales[3 * (size_t)i + k]) * (1.0f / 1.6f));
#pragma unroll
  for (int rep = 0; rep < 2; rep++) {
    const size_t j = (size_t)rep * S + r;
    const float sx = samples[3 * j], sy = samples[3 * j + 1], sz = samples[3 * j + 2];
    out_means[3 * j] = mx + (R00 * sx + R01 * sy + R02 * sz);
    out_means[3 * j + 1] = my + (R10 * sx + R11 * sy + R12 * sz);
    out_means[3 * j + 2] = mz + (R20 * sx + R21 * sy + R22 * sz);
#pragma unroll
    for (int k = 0; k < 3; k++) out_log_scales[3 * j + k] = ls[k];
  }
}

// Packs the step's exchange buffer [P][16] in one pass: means3D(3) colours(3) rotations(4) opacity(1)
// scales(3) grad-norm accumulator(1) visibility count(1)  (gsb200/dist.py COLS).
__global__ void __launch_bounds__(256)
pack_flat_kernel(int P, const float* __restrict__ g_means, const float* __restrict__ g_colors,
                 const float* __restrict__ g_rot, const float* __restrict__ g_opac, const float* __restrict__ g_scales,
                 const float* __restrict__ accum, const float* __restrict__ count, float4* __restrict__ flat) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const size_t j = (size_t)i;
  const float4 q = reinterpret_cast<const float4*>(g_rot)[j];
  flat[4 * j + 0] = make_float4(g_means[3 * j], g_means[3 * j + 1], g_means[3 * j + 2], g_colors[3 * j]);
  flat[4 * j + 1] = make_float4(g_colors[3 * j + 1], g_colors[3 * j + 2], q.x, q.y);
  flat[4 * j + 2] = make_float4(q.z, q.w, g_opac[j], g_scales[3 * j]);
  flat[4 * j + 3] = make_float4(g_scales[3 * j + 1], g_scales[3 * j + 2], accum ? accum[j] : 0.f, count ? count[j] : 0.f);
}

}  // namespace gsb

using namespace gsb;

extern "C" {

int gsb_densify_accumulate(int32_t P, const int32_t* radii, const float* grad_means2D, float* max_2d_radii,
                           float* grad_accum, float* vis_count, uint8_t* visible, void* stream) {
  if (P < 0) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  if (!radii || !grad_means2D || !grad_accum || !vis_count) return GSB_EINVAL;
  densify_accumulate_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, radii, grad_means2D, max_2d_radii, grad_accum, vis_count, visible);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_densify_accumulate_batch(int32_t P, int32_t n_cams, const int32_t* radii, const float* grad_means2D,
                                 float* max_2d_radii, float* grad_accum, float* vis_count, uint8_t* visible,
                                 void* stream) {
  if (P < 0 || n_cams < 0) return GSB_EINVAL;
  if (P == 0 || n_cams == 0) return GSB_OK;
  if (!radii || !grad_means2D || !grad_accum || !vis_count) return GSB_EINVAL;
  densify_accumulate_batch_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, n_cams, radii, grad_means2D, max_2d_radii, grad_accum, vis_count, visible);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_pack_flat(int32_t P, const float* g_means3D, const float* g_colors, const float* g_rotations,
                  const float* g_opacities, const float* g_scales, const float* grad_accum, const float* vis_count,
                  float* flat, void* stream) {
  if (P < 0) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  if (!g_means3D || !g_colors || !g_rotations || !g_opacities || !g_scales || !flat) return GSB_EINVAL;
  pack_flat_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, g_means3D, g_colors, g_rotations, g_opacities, g_scales, grad_accum, vis_count,
      reinterpret_cast<float4*>(flat));
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_densify_accumulate_masked(int32_t P, const uint8_t* visible, const float* grad_means2D, float* grad_accum,
                                  float* vis_count, void* stream) {
  if (P < 0) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  if (!visible || !grad_means2D || !grad_accum || !vis_count) return GSB_EINVAL;
  densify_accumulate_masked_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, visible, grad_means2D, grad_accum, vis_count);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_split_rows(int32_t P, int32_t S, const uint8_t* sel, const int32_t* rank, const float* means,
                   const float* quats, const float* log_scales, const float* samples, float* out_means,
                   float* out_log_scales, void* stream) {
  if (P < 0 || S < 0) return GSB_EINVAL;
  if (P == 0 || S == 0) return GSB_OK;
  if (!sel || !rank || !means || !quats || !log_scales || !samples || !out_means || !out_log_scales) return GSB_EINVAL;
  split_rows_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, S, sel, rank, means, quats, log_scales, samples, out_means, out_log_scales);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_mask_scan(int32_t P, const uint8_t* keep, int32_t* dst_index, int32_t* count_out, void* scratch,
                  void* stream) {
  if (P < 0) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (P == 0) {
    if (count_out) GSB_CUDA_CHECK(cudaMemsetAsync(count_out, 0, 4, st));
    return GSB_OK;
  }
  if (!keep || !dst_index || !scratch) return GSB_EINVAL;
  uint32_t* u = reinterpret_cast<uint32_t*>(dst_index);
  mask_to_u32_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, keep, u);
  GSB_LAUNCH_CHECK();
  int rc = launch_inclusive_scan_u32(P, u, u, scratch, st);  // in place: a tile reads its slice before writing it
  if (rc) return rc;
  mask_rank_finish_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, keep, dst_index, count_out);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_gather_rows(int32_t P, const uint8_t* keep, const int32_t* dst_index, const float* src, float* dst,
                    int32_t row_floats, void* stream) {
  if (P < 0 || row_floats <= 0) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  if (!keep || !dst_index || !src || !dst) return GSB_EINVAL;
  const long long total = (long long)P * row_floats;
  gather_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      total, row_floats, keep, dst_index, src, dst);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // extern "C"
