// Camera geometry right after the camera head and the stand-alone adaLN `modulate`.
//
//   sr_pose_to_extri_intri   sailrecon/utils/pose_enc.py:68-135 (pose_encoding_to_extri_intri, "absT_quaR_FoV") with
//                            utils/rotation.py:14-44 (quat_to_mat, scalar-last quaternion) fused: one thread per camera
//   sr_unproject_depth       sailrecon/utils/geometry.py:19-128 (unproject_depth_map_to_point_map ->
//                            depth_to_world_coords_points -> depth_to_cam_coords_points + closed_form_inverse_se3).  The
//                            reference does this in numpy on the host, which forces a device->host sync of every depth
//                            map inside SailRecon.forward (sail_recon.py:139-141); here the maps never leave HBM.
//   sr_modulate_f32          heads/camera_head.py:189-194 modulate(x, shift, scale) = x * (1 + scale) + shift
//
// All fp32 (the reference disables autocast around the heads, sail_recon.py:119).  HBM-bound, trivially small except the
// unprojection (S x H x W pixels: 4 B read, 12 B written per pixel, float4-free because of the 3-float stride; one thread
// per pixel, consecutive threads on consecutive pixels -> coalesced 4 B loads / 12 B contiguous stores per warp).
#include "common.cuh"
#include "sr_internal.h"

namespace sr {
extern int g_launch_count;

__global__ void pose_to_extri_intri_kernel(const float* __restrict__ pose, float* __restrict__ extri,
                                           float* __restrict__ intri, int64_t n, float H, float W) {
  const int64_t r = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (r >= n) return;
  const float* p = pose + r * 9;
  const float tx = p[0], ty = p[1], tz = p[2];
  const float i = p[3], j = p[4], k = p[5], w = p[6];       // XYZW, real part last (rotation.py:16)
  const float two_s = 2.0f / (i * i + j * j + k * k + w * w);
  float* e = extri + r * 12;                                // [R | t], row major 3 x 4
  e[0] = 1.0f - two_s * (j * j + k * k);
  e[1] = two_s * (i * j - k * w);
  e[2] = two_s * (i * k + j * w);
  e[3] = tx;
  e[4] = two_s * (i * j + k * w);
  e[5] = 1.0f - two_s * (i * i + k * k);
  e[6] = two_s * (j * k - i * w);
  e[7] = ty;
  e[8] = two_s * (i * k - j * w);
  e[9] = two_s * (j * k + i * w);
  e[10] = 1.0f - two_s * (i * i + j * j);
  e[11] = tz;
  if (intri != nullptr) {                                   // pose_enc.py:122-132
    const float fy = (H / 2.0f) / tanf(p[7] / 2.0f);
    const float fx = (W / 2.0f) / tanf(p[8] / 2.0f);
    float* m = intri + r * 9;
    m[0] = fx;   m[1] = 0.0f; m[2] = W / 2.0f;
    m[3] = 0.0f; m[4] = fy;   m[5] = H / 2.0f;
    m[6] = 0.0f; m[7] = 0.0f; m[8] = 1.0f;
  }
}

// Backward of pose_to_extri_intri_kernel.  R = I + s M(q), s = 2 / |q|^2:  dL/dq_m = s <G, dM/dq_m> - s^2 q_m <G, M>;
// d fov = d f * -(size / 4) / sin^2(fov / 2)  (f = (size / 2) / tan(fov / 2)).
__global__ void pose_to_extri_intri_bwd_kernel(const float* __restrict__ pose, const float* __restrict__ d_extri,
                                               const float* __restrict__ d_intri, float* __restrict__ d_pose, int64_t n, float H,
                                               float W) {
  const int64_t r = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (r >= n) return;
  const float* p = pose + r * 9;
  const float* g = d_extri + r * 12;
  const float i = p[3], j = p[4], k = p[5], w = p[6];
  const float s = 2.0f / (i * i + j * j + k * k + w * w);
  const float g00 = g[0], g01 = g[1], g02 = g[2], g10 = g[4], g11 = g[5], g12 = g[6], g20 = g[8], g21 = g[9], g22 = g[10];
  const float gm = g00 * -(j * j + k * k) + g01 * (i * j - k * w) + g02 * (i * k + j * w) + g10 * (i * j + k * w) +
                   g11 * -(i * i + k * k) + g12 * (j * k - i * w) + g20 * (i * k - j * w) + g21 * (j * k + i * w) +
                   g22 * -(i * i + j * j);
  const float di = g01 * j + g02 * k + g10 * j - 2.0f * g11 * i - g12 * w + g20 * k + g21 * w - 2.0f * g22 * i;
  const float dj = -2.0f * g00 * j + g01 * i + g02 * w + g10 * i + g12 * k - g20 * w + g21 * k - 2.0f * g22 * j;
  const float dk = -2.0f * g00 * k - g01 * w + g02 * i + g10 * w - 2.0f * g11 * k + g12 * j + g20 * i + g21 * j;
  const float dw = -g01 * k + g02 * j + g10 * k - g12 * i - g20 * j + g21 * i;
  float* o = d_pose + r * 9;
  o[0] = g[3];
  o[1] = g[7];
  o[2] = g[11];
  o[3] = s * di - s * s * i * gm;
  o[4] = s * dj - s * s * j * gm;
  o[5] = s * dk - s * s * k * gm;
  o[6] = s * dw - s * s * w * gm;
  o[7] = 0.f;
  o[8] = 0.f;
  if (d_intri != nullptr) {
    const float sh = sinf(p[7] / 2.0f), sw = sinf(p[8] / 2.0f);
    o[7] = d_intri[r * 9 + 4] * (-(H / 4.0f) / (sh * sh));
    o[8] = d_intri[r * 9 + 0] * (-(W / 4.0f) / (sw * sw));
  }
}

// world = R^T (cam - t)  with cam = ((u - cu) d / fu, (v - cv) d / fv, d); (R | t) is camera-from-world, so the inverse of
// geometry.py:closed_form_inverse_se3 is R^T and -R^T t.  Arithmetic order follows the reference: cam @ R_inv^T + t_inv.
__global__ void unproject_depth_kernel(const float* __restrict__ depth, const float* __restrict__ extri,
                                       const float* __restrict__ intri, float* __restrict__ out, int S, int H, int W) {
  const int64_t hw = static_cast<int64_t>(H) * W;
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= hw * S) return;
  const int s = static_cast<int>(idx / hw);
  const int64_t pix = idx - s * hw;
  const int v = static_cast<int>(pix / W), u = static_cast<int>(pix - static_cast<int64_t>(v) * W);
  const float* e = extri + s * 12;
  const float* m = intri + s * 9;
  const float fu = m[0], fv = m[4], cu = m[2], cv = m[5];
  const float d = depth[idx];
  const float x = (u - cu) * d / fu, y = (v - cv) * d / fv, z = d;
  // R_inv = R^T, t_inv = -R^T t ; world_i = sum_j cam_j * R_inv[i][j] + t_inv[i] = sum_j cam_j * R[j][i] + t_inv[i]
  const float t0 = -(e[0] * e[3] + e[4] * e[7] + e[8] * e[11]);
  const float t1 = -(e[1] * e[3] + e[5] * e[7] + e[9] * e[11]);
  const float t2 = -(e[2] * e[3] + e[6] * e[7] + e[10] * e[11]);
  float* o = out + idx * 3;
  o[0] = x * e[0] + y * e[4] + z * e[8] + t0;
  o[1] = x * e[1] + y * e[5] + z * e[9] + t1;
  o[2] = x * e[2] + y * e[6] + z * e[10] + t2;
}

__global__ void modulate_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ shift, int64_t lds,
                                const float* __restrict__ scale, int64_t ldc, float* __restrict__ out, int64_t ldo,
                                int64_t rows, int C) {
  const int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (idx >= rows * C) return;
  const int64_t r = idx / C;
  const int c = static_cast<int>(idx - r * C);
  out[r * ldo + c] = x[r * ldx + c] * (1.0f + scale[r * ldc + c]) + shift[r * lds + c];
}

}  // namespace sr

extern "C" int sr_pose_to_extri_intri(const float* pose, float* extrinsics, float* intrinsics, int64_t n, int32_t image_h,
                                      int32_t image_w, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(pose && extrinsics, "sr_pose_to_extri_intri: null pointer");
  SR_REQUIRE(intrinsics == nullptr || (image_h > 0 && image_w > 0), "sr_pose_to_extri_intri: image size required for intrinsics");
  if (n == 0) return SR_OK;
  pose_to_extri_intri_kernel<<<static_cast<unsigned>((n + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      pose, extrinsics, intrinsics, n, static_cast<float>(image_h), static_cast<float>(image_w));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_pose_to_extri_intri_bwd(const float* pose, const float* d_extrinsics, const float* d_intrinsics, float* d_pose,
                                          int64_t n, int32_t image_h, int32_t image_w, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  if (n == 0) return SR_OK;
  SR_REQUIRE(pose && d_extrinsics && d_pose && n > 0, "sr_pose_to_extri_intri_bwd: null pointer");
  SR_REQUIRE(d_intrinsics == nullptr || (image_h > 0 && image_w > 0), "sr_pose_to_extri_intri_bwd: image size required for intrinsics");
  pose_to_extri_intri_bwd_kernel<<<static_cast<unsigned>((n + 127) / 128), 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      pose, d_extrinsics, d_intrinsics, d_pose, n, static_cast<float>(image_h), static_cast<float>(image_w));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_unproject_depth(const float* depth, const float* extrinsics, const float* intrinsics, float* points, int64_t S,
                                  int64_t H, int64_t W, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(depth && extrinsics && intrinsics && points, "sr_unproject_depth: null pointer");
  SR_REQUIRE(S >= 0 && H > 0 && W > 0 && H < (1 << 20) && W < (1 << 20) && S < (1 << 20), "sr_unproject_depth: bad shape");
  const int64_t total = S * H * W;
  if (total == 0) return SR_OK;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_unproject_depth: too many pixels");
  unproject_depth_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      depth, extrinsics, intrinsics, points, static_cast<int>(S), static_cast<int>(H), static_cast<int>(W));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}

extern "C" int sr_modulate_f32(const float* x, int64_t ldx, const float* shift, int64_t lds, const float* scale, int64_t ldc,
                               float* out, int64_t ldo, int64_t rows, int64_t C, void* stream) {
  using namespace sr;
  if (int rc = require_sm100()) return rc;
  SR_REQUIRE(x && shift && scale && out, "sr_modulate_f32: null pointer");
  SR_REQUIRE(C > 0 && C < (1 << 30) && ldx >= C && lds >= C && ldc >= C && ldo >= C, "sr_modulate_f32: bad strides");
  const int64_t total = rows * C;
  if (total == 0) return SR_OK;
  SR_REQUIRE((total + 255) / 256 < (1ll << 31), "sr_modulate_f32: too many elements");
  modulate_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, ldx, shift, lds, scale, ldc, out, ldo, rows, static_cast<int>(C));
  ++g_launch_count;
  SR_CHECK_CUDA(cudaGetLastError());
  return SR_OK;
}
