// probtalk_b200 -- pose post-processing: normalised 6-D network output -> SMPL-X axis-angle parameters (sm_100a).
//
// One pass over the final poses: denormalise, 6-D -> rotation matrix (Gram-Schmidt) -> quaternion ->
// axis-angle per joint, copy the expression channels, splice in the fixed lower-body joints.  HBM-bound elementwise
// work: channel-first reads are coalesced along time, the [T, out_dim] rows leave through a shared-memory tile.
#include "pt_common.cuh"
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace {

constexpr int kTileT = 32;
constexpr int kLower = 27;                 // fixed lower-body values spliced in by part2full

// F.normalize(v, dim=-1): v / max(||v||, 1e-12)
__device__ __forceinline__ void normalize3(float& x, float& y, float& z) {
  const float n = fmaxf(sqrtf(x * x + y * y + z * z), 1e-12f);
  x = x / n; y = y / n; z = z / n;
}
__device__ __forceinline__ float sqrt_pos(float v) { return v > 0.f ? sqrtf(v) : 0.f; }
// rotation_conversion._copysign: flip a when the signs of a and b differ ("a < 0" vs "b < 0", so -0 / NaN as there)
__device__ __forceinline__ float copysign_ref(float a, float b) { return ((a < 0.f) != (b < 0.f)) ? -a : a; }

// index of part2full's output column `o` in the [3J + E] vector it is given, or -(k+1) for lower-body constant k
__device__ __forceinline__ int full_source(int o) {
  if (o < 3) return o;
  if (o < 18) return -(o - 3 + 1);            // lp[0:15]
  if (o < 21) return o - 18 + 3;              // in[3:6]
  if (o < 27) return -(o - 21 + 15 + 1);      // lp[15:21]
  if (o < 30) return o - 27 + 6;              // in[6:9]
  if (o < 36) return -(o - 30 + 21 + 1);      // lp[21:27]
  return o - 36 + 9;                          // in[9:]
}

__global__ void __launch_bounds__(256) pose_post_kernel(const float* __restrict__ poses, const float* __restrict__ mean,
                                                        const float* __restrict__ stdv, const float* __restrict__ lower,
                                                        float* __restrict__ out, int T, int J, int E) {
  extern __shared__ float rowbuf[];           // [kTileT][W + 1], W = 3J + E (the vector part2full receives)
  const int W = 3 * J + E, C = 6 * J + E, ld = W + 1, OUT = W + kLower;
  const int b = blockIdx.y, t0 = blockIdx.x * kTileT;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // tx = time step within the tile, ty = 0..7
  const int t = t0 + tx;
  const float* src = poses + (size_t)b * C * T;
  auto load = [&](int c) {                    // denormalised channel c at time t (nets/utils.py:27-43)
    float v = src[(size_t)c * T + t];
    if (mean) v = v * stdv[c] + mean[c];
    return v;
  };
  if (t < T) {
    for (int j = ty; j < J; j += 8) {
      float a1x = load(6 * j), a1y = load(6 * j + 1), a1z = load(6 * j + 2);
      float a2x = load(6 * j + 3), a2y = load(6 * j + 4), a2z = load(6 * j + 5);
      // rotation_6d_to_matrix (rotation_conversion.py:512-534): rows b1, b2, b3
      normalize3(a1x, a1y, a1z);
      const float d = a1x * a2x + a1y * a2y + a1z * a2z;
      float b2x = a2x - d * a1x, b2y = a2y - d * a1y, b2z = a2z - d * a1z;
      normalize3(b2x, b2y, b2z);
      const float b3x = a1y * b2z - a1z * b2y, b3y = a1z * b2x - a1x * b2z, b3z = a1x * b2y - a1y * b2x;
      // matrix_to_quaternion (:97-119): m00 = b1.x, m11 = b2.y, m22 = b3.z
      const float m00 = a1x, m11 = b2y, m22 = b3z;
      const float o0 = 0.5f * sqrt_pos(1.f + m00 + m11 + m22);
      float qx = 0.5f * sqrt_pos(1.f + m00 - m11 - m22);
      float qy = 0.5f * sqrt_pos(1.f - m00 + m11 - m22);
      float qz = 0.5f * sqrt_pos(1.f - m00 - m11 + m22);
      qx = copysign_ref(qx, b3y - b2z);       // m21 - m12
      qy = copysign_ref(qy, a1z - b3x);       // m02 - m20
      qz = copysign_ref(qz, b2x - a1y);       // m10 - m01
      // quaternion_to_axis_angle (:481-509)
      const float nrm = sqrtf(qx * qx + qy * qy + qz * qz);
      const float half = atan2f(nrm, o0);
      const float ang = 2.f * half;
      const float s = (fabsf(ang) < 1e-6f) ? (0.5f - ang * ang / 48.f) : (sinf(half) / ang);
      rowbuf[tx * ld + 3 * j] = qx / s;
      rowbuf[tx * ld + 3 * j + 1] = qy / s;
      rowbuf[tx * ld + 3 * j + 2] = qz / s;
    }
    for (int e = ty; e < E; e += 8) rowbuf[tx * ld + 3 * J + e] = load(6 * J + e);   // expression channels pass through
  }
  __syncthreads();
  // part2full (lower_body.py:66-83) while storing: consecutive threads write consecutive output columns of a row
  for (int r = ty; r < kTileT; r += 8) {
    if (t0 + r >= T) break;
    float* dst = out + ((size_t)b * T + t0 + r) * OUT;
    for (int o = tx; o < OUT; o += 32) {
      const int sidx = full_source(o);
      dst[o] = sidx >= 0 ? rowbuf[r * ld + sidx] : lower[-sidx - 1];
    }
  }
}

}  // namespace

extern "C" int pt_pose_postprocess(const float* poses, const float* mean, const float* stdv, const float* lower27,
                                   float* out, int B, int T, int n_joints, int n_expr, void* stream) {
  if (B <= 0 || T <= 0) return 0;
  if (!poses || !lower27 || !out) return pt::fail("pt_pose_postprocess: null pointer");
  if ((mean == nullptr) != (stdv == nullptr)) return pt::fail("pt_pose_postprocess: mean and std go together");
  if (n_joints < 3 || n_expr < 0) return pt::fail("pt_pose_postprocess: need at least 3 joints (got %d)", n_joints);
  if (B > 65535) return pt::fail("pt_pose_postprocess: at most 65535 samples per call (got %d)", B);
  const size_t smem = (size_t)kTileT * (3 * n_joints + n_expr + 1) * sizeof(float);
  if (smem > 200 * 1024) return pt::fail("pt_pose_postprocess: %d joints + %d expression channels do not fit", n_joints, n_expr);
  if (smem > 48 * 1024) PT_SMEM_OPT_IN(pose_post_kernel, smem);
  pose_post_kernel<<<dim3((T + kTileT - 1) / kTileT, B), 256, smem, (cudaStream_t)stream>>>(poses, mean, stdv, lower27, out, T,
                                                                                             n_joints, n_expr);
  PT_CUDA(cudaGetLastError());
  return 0;
}
