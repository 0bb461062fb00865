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

// ------------------------------------------------------------------------------------------------
// One densify event as one pass over the Gaussians (external.py:211-314).  The reference appends clones, appends
// split children, compacts away the split parents and compacts again for the prune -- every step a torch.cat /
// boolean-mask re-allocation of all six tables and both Adam moments.  Every row of the FINAL tables is a function of
// exactly one source row, and whether it survives can be decided from that source row alone:
//   original i     survives iff not split and not pruned                       -> [0, K0)
//   clone of i     survives iff cloned and not pruned (a copy of the row)      -> [K0, K0 + K1)
//   children of i  survive iff split and not pruned with log_scale - log 1.6   -> [K0 + K1, +K2) and [.., +K2)
// (order inside each group = source order, which is what cat + stable mask indexing leave).  plan: flags + four
// 0/1 columns; their inclusive scans are the ranks; apply: thread per source row writes its 0-4 output rows of all
// 18 tables.  SURVEY.md 8(d): <= 2 x 204 B per surviving Gaussian.
// ------------------------------------------------------------------------------------------------
enum { DF_ORIG = 1, DF_CLONE = 2, DF_CHILD = 4, DF_SPLIT = 8 };

__global__ void __launch_bounds__(256)
densify_plan_kernel(int P, const float* __restrict__ grad_accum, const float* __restrict__ vis_count,
                    const float* __restrict__ opacity_logits, const float* __restrict__ log_scales, float grad_thr,
                    float clone_max_scale, float prune_opacity, int prune_big, float big_scale,
                    uint8_t* __restrict__ flags, uint32_t* __restrict__ cols /* [4][P] */) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  float avg = grad_accum[i] / vis_count[i];          // external.py:224-225: NaN (0 / 0) -> 0
  if (avg != avg) avg = 0.f;
  const float s0 = expf(log_scales[3 * (size_t)i]), s1 = expf(log_scales[3 * (size_t)i + 1]), s2 = expf(log_scales[3 * (size_t)i + 2]);
  const float max_s = fmaxf(s0, fmaxf(s1, s2));
  const bool hot = avg >= grad_thr;
  const bool clone = hot && max_s <= clone_max_scale;   // external.py:232-238
  const bool split = hot && max_s > clone_max_scale;    // external.py:241-250 (clones carry a padded gradient of 0)
  const float op = 1.0f / (1.0f + expf(-opacity_logits[i]));
  const bool low = op < prune_opacity;                  // external.py:290-296
  const bool prune_row = low || (prune_big && max_s > big_scale);
  // the children's scales as the reference forms them: exp(log(exp(ls) / 1.6))
  const float c0 = expf(logf(s0 * (1.0f / 1.6f))), c1 = expf(logf(s1 * (1.0f / 1.6f))), c2 = expf(logf(s2 * (1.0f / 1.6f)));
  const bool prune_child = low || (prune_big && fmaxf(c0, fmaxf(c1, c2)) > big_scale);
  const unsigned f = ((!split && !prune_row) ? DF_ORIG : 0) | ((clone && !prune_row) ? DF_CLONE : 0) |
                     ((split && !prune_child) ? DF_CHILD : 0) | (split ? DF_SPLIT : 0);
  flags[i] = (uint8_t)f;
  cols[i] = (f & DF_ORIG) ? 1u : 0u;
  cols[(size_t)P + i] = (f & DF_CLONE) ? 1u : 0u;
  cols[2 * (size_t)P + i] = (f & DF_CHILD) ? 1u : 0u;
  cols[3 * (size_t)P + i] = (f & DF_SPLIT) ? 1u : 0u;
}

__global__ void densify_totals_kernel(int P, const uint32_t* __restrict__ cols, int32_t* __restrict__ totals) {
  if (threadIdx.x < 4) totals[threadIdx.x] = (int32_t)cols[(size_t)threadIdx.x * P + P - 1];   // inclusive scans
}

struct DensifyTables {
  const float* src[18];   // means, colors, segmentation colours, quaternions, opacity logits, log scales; then their
  float* dst[18];         // exp_avg (6) and exp_avg_sq (6) in the same order
};
__constant__ const int kRowFloats[6] = {3, 3, 3, 4, 1, 3};

__global__ void __launch_bounds__(256)
densify_apply_kernel(int P, const uint8_t* __restrict__ flags, const uint32_t* __restrict__ cols, int K0, int K1, int K2, int S,
                     DensifyTables t, const float* __restrict__ samples) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const unsigned f = flags[i];
  if (!(f & (DF_ORIG | DF_CLONE | DF_CHILD))) return;
  auto copy_row = [&](int table, size_t dst_row, bool zero) {
    const int w = kRowFloats[table % 6];
    const float* s = t.src[table] + (size_t)i * w;
    float* d = t.dst[table] + dst_row * w;
    for (int k = 0; k < w; k++) d[k] = zero ? 0.f : s[k];
  };
  if (f & DF_ORIG) {
    const size_t r = cols[i] - 1u;
    for (int tb = 0; tb < 18; tb++) copy_row(tb, r, false);
  }
  if (f & DF_CLONE) {   // cat_params_to_optimizer: the row is appended, its Adam moments start at zero
    const size_t r = (size_t)K0 + cols[(size_t)P + i] - 1u;
    for (int tb = 0; tb < 18; tb++) copy_row(tb, r, tb >= 6);
  }
  if (f & DF_CHILD) {   // external.py:251-271
    const size_t ra = (size_t)K0 + K1 + cols[2 * (size_t)P + i] - 1u, rb = ra + K2;
    const size_t sj = cols[3 * (size_t)P + i] - 1u;   // row of the split draw: rank among ALL split rows
    for (int tb = 0; tb < 18; tb++) {
      copy_row(tb, ra, tb >= 6);
      copy_row(tb, rb, tb >= 6);
    }
    float4 q = reinterpret_cast<const float4*>(t.src[3])[i];
    const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);  // build_rotation, external.py:27-46
    q.x /= norm; q.y /= norm; q.z /= norm; q.w /= norm;
    const float rr = q.x, x = q.y, y = q.z, z = q.w;
    const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - rr * z), R02 = 2.f * (x * z + rr * y);
    const float R10 = 2.f * (x * y + rr * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - rr * x);
    const float R20 = 2.f * (x * z - rr * y), R21 = 2.f * (y * z + rr * x), R22 = 1.f - 2.f * (x * x + y * y);
    const float mx = t.src[0][3 * (size_t)i], my = t.src[0][3 * (size_t)i + 1], mz = t.src[0][3 * (size_t)i + 2];
    float ls[3];
#pragma unroll
    for (int k = 0; k < 3; k++) ls[k] = logf(expf(t.src[5][3 * (size_t)i + k]) * (1.0f / 1.6f));
#pragma unroll
    for (int rep = 0; rep < 2; rep++) {
      const size_t j = (size_t)rep * S + sj, r = rep ? rb : ra;
      const float sx = samples[3 * j], sy = samples[3 * j + 1], sz = samples[3 * j + 2];
      t.dst[0][3 * r] = mx + (R00 * sx + R01 * sy + R02 * sz);
      t.dst[0][3 * r + 1] = my + (R10 * sx + R11 * sy + R12 * sz);
      t.dst[0][3 * r + 2] = mz + (R20 * sx + R21 * sy + R22 * sz);
#pragma unroll
      for (int k = 0; k < 3; k++) t.dst[5][3 * r + k] = ls[k];
    }
  }
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

size_t gsb_densify_plan_scratch_bytes(int32_t P) {
  // flags u8 [P] | four 0/1 -> rank columns u32 [4][P] | scan scratch
  const size_t n = (size_t)(P > 0 ? P : 1);
  return align_up(n, 256) + align_up(16 * n, 256) + scan_scratch_bytes((int64_t)n) + 256;
}

int gsb_densify_plan(int32_t P, const float* grad_accum, const float* vis_count, const float* opacity_logits,
                     const float* log_scales, float grad_threshold, float clone_max_scale, float prune_opacity,
                     int32_t prune_big, float big_scale, void* scratch, int32_t* totals_dev, void* stream) {
  if (P < 0 || !scratch || !totals_dev) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (P == 0) { GSB_CUDA_CHECK(cudaMemsetAsync(totals_dev, 0, 16, st)); return GSB_OK; }
  if (!grad_accum || !vis_count || !opacity_logits || !log_scales) return GSB_EINVAL;
  char* base = reinterpret_cast<char*>(scratch);
  uint8_t* flags = reinterpret_cast<uint8_t*>(base);
  uint32_t* cols = reinterpret_cast<uint32_t*>(base + align_up((size_t)P, 256));
  void* scan_scratch = base + align_up((size_t)P, 256) + align_up(16 * (size_t)P, 256);
  densify_plan_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, grad_accum, vis_count, opacity_logits, log_scales, grad_threshold,
                                                       clone_max_scale, prune_opacity, prune_big, big_scale, flags, cols);
  GSB_LAUNCH_CHECK();
  for (int k = 0; k < 4; k++) {
    uint32_t* c = cols + (size_t)k * P;
    int rc = launch_inclusive_scan_u32(P, c, c, scan_scratch, st);
    if (rc) return rc;
  }
  densify_totals_kernel<<<1, 32, 0, st>>>(P, cols, totals_dev);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_densify_apply(int32_t P, const void* scratch, const int32_t* totals_host, const float* const* src18,
                      float* const* dst18, const float* samples, void* stream) {
  if (P < 0 || !scratch || !totals_host || !src18 || !dst18) return GSB_EINVAL;
  if (P == 0) return GSB_OK;
  const int K0 = totals_host[0], K1 = totals_host[1], K2 = totals_host[2], S = totals_host[3];
  if (K0 < 0 || K1 < 0 || K2 < 0 || S < K2 || (S > 0 && !samples)) return GSB_EINVAL;
  if (K0 + K1 + 2 * K2 == 0) return GSB_OK;
  DensifyTables t;
  for (int k = 0; k < 18; k++) {
    if (!src18[k] || !dst18[k]) return GSB_EINVAL;
    t.src[k] = src18[k]; t.dst[k] = dst18[k];
  }
  if ((uintptr_t)t.src[3] & 15u) return GSB_EINVAL;
  const char* base = reinterpret_cast<const char*>(scratch);
  const uint8_t* flags = reinterpret_cast<const uint8_t*>(base);
  const uint32_t* cols = reinterpret_cast<const uint32_t*>(base + align_up((size_t)P, 256));
  densify_apply_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(P, flags, cols, K0, K1, K2, S, t,
                                                                                         samples);
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
