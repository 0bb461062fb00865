// rigidity.cu -- the rigidity loss of the per-timestep training step as ONE kernel that
// returns the loss and its gradient, plus the kernel that snapshots the previous timestep's foreground state.
//
// Reference: calculate_rigidity_loss (train.py:317-351), quat_mult (train.py:301-308), weighted_l2_loss_v2
// (train.py:311-315), build_rotation (external.py:27-47), create_foreground_info (train.py:226-247).  There the loss
// is ~30 torch kernels with [F,K,3,3] temporaries, evaluated once per sampled VIEW although it does not depend
// on the view (train.py:373-377 inside :402-412); here it is evaluated once per step.
//
//   q   = normalize(quats[fg])                      rel = q (x) prev_inv_rot          R = rotation(rel / |rel|)
//   off = means[fg][nbr] - means[fg]                rot = R^T off
//   loss = mean_ij sqrt(w_ij |rot_ij - prev_off_ij|^2 + 1e-20)
//
// One thread per foreground Gaussian walks its K neighbours: R, dL/dR (9 sums) and dL/dmean_i stay in
// registers, each neighbour receives its share with 3 float atomics, and the chain dR -> rel -> q -> raw
// quaternion is applied once per Gaussian.  The per-pair tables are addressed through (row, column) strides so
// that callers can hand over the constant ones (indices, weights) transposed ([K][F]: coalesced for this
// mapping) without this file caring.  HBM/L2-bound gather work: per pair 4 B index + 4 B weight + 12 B
// previous offset + 12 B gathered mean read, 12 B of atomics.
#include "common.cuh"

namespace gsb {

struct Quat { float w, x, y, z; };

__device__ __forceinline__ Quat quat_mul(const Quat a, const Quat b) {
  return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}

__device__ __forceinline__ void rotation_of(const Quat u, float (&R)[9]) {
  const float r = u.w, x = u.x, y = u.y, z = u.z;
  R[0] = 1.f - 2.f * (y * y + z * z); R[1] = 2.f * (x * y - r * z); R[2] = 2.f * (x * z + r * y);
  R[3] = 2.f * (x * y + r * z); R[4] = 1.f - 2.f * (x * x + z * z); R[5] = 2.f * (y * z - r * x);
  R[6] = 2.f * (x * z - r * y); R[7] = 2.f * (y * z + r * x); R[8] = 1.f - 2.f * (x * x + y * y);
}

struct PairTables {
  const int32_t* nbr; long long nbr_si, nbr_sj;            // neighbour index (foreground space)
  const float* w; long long w_si, w_sj;                    // weight
  const float* prev; long long prev_si, prev_sj, prev_sc;  // previous offsets
};

template <bool GRAD>
__global__ void __launch_bounds__(128)
rigidity_kernel(int F, int K, const int32_t* __restrict__ fg_index, const float* __restrict__ means,
                const float4* __restrict__ quats, const float4* __restrict__ prev_inv_rot, PairTables t,
                float inv_count, double* __restrict__ loss_sum, float* __restrict__ d_means,
                float4* __restrict__ d_quats) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  float part = 0.f;
  if (i < F) {
    const int row = fg_index ? fg_index[i] : i;
    const float4 qr = __ldg(quats + row);
    const float n = fmaxf(sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w), 1e-12f);
    const float inv_n = 1.f / n;
    const Quat q = {qr.x * inv_n, qr.y * inv_n, qr.z * inv_n, qr.w * inv_n};
    const float4 pr = __ldg(prev_inv_rot + i);
    const Quat p = {pr.x, pr.y, pr.z, pr.w};
    const Quat rel = quat_mul(q, p);
    const float rel_n = sqrtf(rel.w * rel.w + rel.x * rel.x + rel.y * rel.y + rel.z * rel.z);
    const float inv_rel = 1.f / rel_n;
    const Quat u = {rel.w * inv_rel, rel.x * inv_rel, rel.y * inv_rel, rel.z * inv_rel};
    float R[9];
    rotation_of(u, R);
    const float mx = __ldg(means + 3 * (size_t)row), my = __ldg(means + 3 * (size_t)row + 1),
                mz = __ldg(means + 3 * (size_t)row + 2);
    float dR[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dmx = 0.f, dmy = 0.f, dmz = 0.f;
    for (int j = 0; j < K; j++) {
      const int nb = __ldg(t.nbr + i * t.nbr_si + j * t.nbr_sj);
      const float w = __ldg(t.w + i * t.w_si + j * t.w_sj);
      const float* pv = t.prev + i * t.prev_si + j * t.prev_sj;
      const float p0 = __ldg(pv), p1 = __ldg(pv + t.prev_sc), p2 = __ldg(pv + 2 * t.prev_sc);
      const int nrow = fg_index ? fg_index[nb] : nb;
      const float ox = __ldg(means + 3 * (size_t)nrow) - mx, oy = __ldg(means + 3 * (size_t)nrow + 1) - my,
                  oz = __ldg(means + 3 * (size_t)nrow + 2) - mz;
      // rot = R^T off
      const float d0 = R[0] * ox + R[3] * oy + R[6] * oz - p0;
      const float d1 = R[1] * ox + R[4] * oy + R[7] * oz - p1;
      const float d2 = R[2] * ox + R[5] * oy + R[8] * oz - p2;
      const float root = sqrtf((d0 * d0 + d1 * d1 + d2 * d2) * w + 1e-20f);
      part += root;
      if (GRAD) {
        const float c = w / root * inv_count;
        const float g0 = c * d0, g1 = c * d1, g2 = c * d2;
        // d off = R g
        const float fx = R[0] * g0 + R[1] * g1 + R[2] * g2;
        const float fy = R[3] * g0 + R[4] * g1 + R[5] * g2;
        const float fz = R[6] * g0 + R[7] * g1 + R[8] * g2;
        dmx -= fx; dmy -= fy; dmz -= fz;
        atomicAdd(d_means + 3 * (size_t)nrow, fx);
        atomicAdd(d_means + 3 * (size_t)nrow + 1, fy);
        atomicAdd(d_means + 3 * (size_t)nrow + 2, fz);
        dR[0] += ox * g0; dR[1] += ox * g1; dR[2] += ox * g2;
        dR[3] += oy * g0; dR[4] += oy * g1; dR[5] += oy * g2;
        dR[6] += oz * g0; dR[7] += oz * g1; dR[8] += oz * g2;
      }
    }
    if (GRAD) {
      atomicAdd(d_means + 3 * (size_t)row, dmx);
      atomicAdd(d_means + 3 * (size_t)row + 1, dmy);
      atomicAdd(d_means + 3 * (size_t)row + 2, dmz);
      const float r = u.w, x = u.x, y = u.y, z = u.z;
      Quat du;
      du.w = 2.f * (-z * dR[1] + y * dR[2] + z * dR[3] - x * dR[5] - y * dR[6] + x * dR[7]);
      du.x = 2.f * (y * dR[1] + z * dR[2] + y * dR[3] - 2.f * x * dR[4] - r * dR[5] + z * dR[6] + r * dR[7] -
                    2.f * x * dR[8]);
      du.y = 2.f * (-2.f * y * dR[0] + x * dR[1] + r * dR[2] + x * dR[3] + z * dR[5] - r * dR[6] + z * dR[7] -
                    2.f * y * dR[8]);
      du.z = 2.f * (-2.f * z * dR[0] - r * dR[1] + x * dR[2] + r * dR[3] - 2.f * z * dR[4] + y * dR[5] + x * dR[6] +
                    y * dR[7]);
      // through rel / |rel|
      const float dot_u = u.w * du.w + u.x * du.x + u.y * du.y + u.z * du.z;
      const Quat dr = {(du.w - u.w * dot_u) * inv_rel, (du.x - u.x * dot_u) * inv_rel, (du.y - u.y * dot_u) * inv_rel,
                       (du.z - u.z * dot_u) * inv_rel};
      // through rel = q (x) p  (q is the left operand)
      const Quat dq = {dr.w * p.w + dr.x * p.x + dr.y * p.y + dr.z * p.z,
                       -dr.w * p.x + dr.x * p.w - dr.y * p.z + dr.z * p.y,
                       -dr.w * p.y + dr.x * p.z + dr.y * p.w - dr.z * p.x,
                       -dr.w * p.z - dr.x * p.y + dr.y * p.x + dr.z * p.w};
      // through q = raw / max(|raw|, eps)   (component order of the raw quaternion is (w, x, y, z) = float4 xyzw)
      float4 out;
      if (n > 1e-12f) {
        const float dot_q = q.w * dq.w + q.x * dq.x + q.y * dq.y + q.z * dq.z;
        out = make_float4((dq.w - q.w * dot_q) * inv_n, (dq.x - q.x * dot_q) * inv_n, (dq.y - q.y * dot_q) * inv_n,
                          (dq.z - q.z * dot_q) * inv_n);
      } else {
        out = make_float4(dq.w * 1e12f, dq.x * 1e12f, dq.y * 1e12f, dq.z * 1e12f);
      }
      // foreground rows are distinct, so a plain read-modify-write of the caller's accumulator is race-free
      float4 acc = d_quats[row];
      acc.x += out.x; acc.y += out.y; acc.z += out.z; acc.w += out.w;
      d_quats[row] = acc;
    }
  }
  // block sum of the per-thread partial losses, one double atomic per block
  __shared__ double s_part[4];
  double v = (double)part;
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) atomicAdd(loss_sum, (s_part[0] + s_part[1] + s_part[2] + s_part[3]) * (double)inv_count);
}

// create_foreground_info (train.py:226-247): inverted normalised rotation and offsets to the neighbours.
__global__ void __launch_bounds__(128)
foreground_info_kernel(int F, int K, const int32_t* __restrict__ fg_index, const float* __restrict__ means,
                       const float4* __restrict__ quats, const int32_t* __restrict__ nbr, long long nbr_si,
                       long long nbr_sj, float4* __restrict__ inv_rot, float* __restrict__ offsets, long long off_si,
                       long long off_sj, long long off_sc) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= F) return;
  const int row = fg_index ? fg_index[i] : i;
  const float4 qr = __ldg(quats + row);
  const float d = fmaxf(sqrtf(qr.x * qr.x + qr.y * qr.y + qr.z * qr.z + qr.w * qr.w), 1e-12f);
  inv_rot[i] = make_float4(qr.x / d, -(qr.y / d), -(qr.z / d), -(qr.w / d));
  const float mx = __ldg(means + 3 * (size_t)row), my = __ldg(means + 3 * (size_t)row + 1),
              mz = __ldg(means + 3 * (size_t)row + 2);
  for (int j = 0; j < K; j++) {
    const int nb = __ldg(nbr + i * nbr_si + j * nbr_sj);
    const int nrow = fg_index ? fg_index[nb] : nb;
    float* o = offsets + i * off_si + j * off_sj;
    o[0] = __ldg(means + 3 * (size_t)nrow) - mx;
    o[off_sc] = __ldg(means + 3 * (size_t)nrow + 1) - my;
    o[2 * off_sc] = __ldg(means + 3 * (size_t)nrow + 2) - mz;
  }
}

}  // namespace gsb

using namespace gsb;

extern "C" {

int gsb_rigidity_loss(int32_t F, int32_t K, const int32_t* fg_index, const float* means, const float* quats,
                      const float* prev_inv_rot, const int32_t* nbr, int64_t nbr_stride_i, int64_t nbr_stride_j,
                      const float* weights, int64_t w_stride_i, int64_t w_stride_j, const float* prev_offsets,
                      int64_t prev_stride_i, int64_t prev_stride_j, int64_t prev_stride_c, double* loss,
                      float* d_means, float* d_quats, void* stream) {
  if (F < 0 || K < 0) return GSB_EINVAL;
  if (!loss) return GSB_EINVAL;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (cudaMemsetAsync(loss, 0, sizeof(double), st) != cudaSuccess) return GSB_ECUDA;
  if (F == 0 || K == 0) return GSB_OK;   // mean over an empty set: callers treat 0 pairs as loss 0
  if (!means || !quats || !prev_inv_rot || !nbr || !weights || !prev_offsets) return GSB_EINVAL;
  if ((d_means == nullptr) != (d_quats == nullptr)) return GSB_EINVAL;
  PairTables t = {nbr, nbr_stride_i, nbr_stride_j, weights, w_stride_i, w_stride_j,
                  prev_offsets, prev_stride_i, prev_stride_j, prev_stride_c};
  const float inv_count = (float)(1.0 / ((double)F * (double)K));
  const int blocks = (F + 127) / 128;
  if (d_means)
    rigidity_kernel<true><<<blocks, 128, 0, st>>>(F, K, fg_index, means, reinterpret_cast<const float4*>(quats),
                                                  reinterpret_cast<const float4*>(prev_inv_rot), t, inv_count, loss,
                                                  d_means, reinterpret_cast<float4*>(d_quats));
  else
    rigidity_kernel<false><<<blocks, 128, 0, st>>>(F, K, fg_index, means, reinterpret_cast<const float4*>(quats),
                                                   reinterpret_cast<const float4*>(prev_inv_rot), t, inv_count, loss,
                                                   nullptr, nullptr);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

int gsb_foreground_info(int32_t F, int32_t K, const int32_t* fg_index, const float* means, const float* quats,
                        const int32_t* nbr, int64_t nbr_stride_i, int64_t nbr_stride_j, float* inv_rot,
                        float* offsets, int64_t off_stride_i, int64_t off_stride_j, int64_t off_stride_c,
                        void* stream) {
  if (F < 0 || K < 0) return GSB_EINVAL;
  if (F == 0) return GSB_OK;
  if (!means || !quats || !inv_rot || (K > 0 && (!nbr || !offsets))) return GSB_EINVAL;
  foreground_info_kernel<<<(F + 127) / 128, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      F, K, fg_index, means, reinterpret_cast<const float4*>(quats), nbr, nbr_stride_i, nbr_stride_j,
      reinterpret_cast<float4*>(inv_rot), offsets, off_stride_i, off_stride_j, off_stride_c);
  GSB_LAUNCH_CHECK();
  return GSB_OK;
}

}  // extern "C"
