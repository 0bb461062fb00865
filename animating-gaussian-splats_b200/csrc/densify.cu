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
