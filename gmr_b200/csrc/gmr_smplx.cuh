// gmr_smplx.cuh - SMPL-X model output -> human frame stream on the device.
//
// Replaces get_smplx_data_offline_fast (general_motion_retargeting/utils/smpl.py:109-198), which loops per
// frame and per joint in Python over scipy Rotation objects:
//   :124-171  fps alignment - for every output time t: the two neighbouring source frames, SLERP of each joint's
//             rotation (smpl.py:78-107 on scipy quaternions, back to a rotation vector), linear interpolation of
//             the joint positions (scipy interp1d);
//   :178-196  global orientation chain R_i = R_parent(i) * R(rotvec_i), output quaternion scalar first.
// scipy's Rotation arithmetic is restated from its published formulas (from_rotvec / as_rotvec with their
// small-angle series at 1e-3, Hamilton product followed by normalisation, canonical sign before as_rotvec).
// One kernel, thread = output frame, float64; the joint chain lives in a per-thread array (J <= 64).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gmr {

#define SMPLX_MAX_JOINTS 64

struct SmplxArgs {
  const float* pose;        // [N,J,3] rotation vectors, joint 0 = global_orient
  const float* joints;      // [N,Jall,3]
  int Jall;
  const int* src_start;     // [C+1] first source frame of each clip
  const int* clip_of;       // [N_out]
  const double* time;       // [N_out] source-frame time of each output frame, relative to its clip
  const int* aligned;       // [N_out] 1: interpolate (even when the time is an integer, as the reference does)
  int N, N_out, J;
  const int* parents;       // [J]
  void* out;                // [N_out,J,7]
};

// xyzw, scipy from_rotvec
__device__ __forceinline__ void sx_from_rotvec(const double* v, double* q) {
  const double angle = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  const double a2 = angle * angle;
  double s, c;
  sincos(angle / 2, &s, &c);
  const double scale = angle <= 1e-3 ? 0.5 - a2 / 48 + a2 * a2 / 3840 : s / angle;
  q[0] = v[0] * scale; q[1] = v[1] * scale; q[2] = v[2] * scale; q[3] = c;
}
__device__ __forceinline__ void sx_normalize(double* q) {
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}
// scipy compose_quat(p, q) + the normalisation of Rotation.__mul__
__device__ __forceinline__ void sx_compose(const double* p, const double* q, double* r) {
  const double cx = p[1] * q[2] - p[2] * q[1], cy = p[2] * q[0] - p[0] * q[2], cz = p[0] * q[1] - p[1] * q[0];
  r[0] = p[3] * q[0] + q[3] * p[0] + cx;
  r[1] = p[3] * q[1] + q[3] * p[1] + cy;
  r[2] = p[3] * q[2] + q[3] * p[2] + cz;
  r[3] = p[3] * q[3] - p[0] * q[0] - p[1] * q[1] - p[2] * q[2];
  sx_normalize(r);
}
// scipy as_rotvec (canonical sign first)
__device__ __forceinline__ void sx_as_rotvec(const double* qin, double* v) {
  double q[4] = {qin[0], qin[1], qin[2], qin[3]};
  const bool neg = q[3] < 0 || (q[3] == 0 && (q[0] < 0 || (q[0] == 0 && (q[1] < 0 || (q[1] == 0 && q[2] < 0)))));
  if (neg) { q[0] = -q[0]; q[1] = -q[1]; q[2] = -q[2]; q[3] = -q[3]; }
  const double axn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  const double angle = 2 * atan2(axn, q[3]);
  const double a2 = angle * angle;
  const double scale = angle <= 1e-3 ? 2 + a2 / 12 + 7 * a2 * a2 / 2880 : angle / sin(angle / 2.0);
  v[0] = scale * q[0]; v[1] = scale * q[1]; v[2] = scale * q[2];
}
// smpl.py:78-107
__device__ __forceinline__ void sx_slerp(const double* q1in, const double* q2in, double t, double* q) {
  double q1[4] = {q1in[0], q1in[1], q1in[2], q1in[3]}, q2[4] = {q2in[0], q2in[1], q2in[2], q2in[3]};
  sx_normalize(q1);
  sx_normalize(q2);
  double dot = q1[0] * q2[0] + q1[1] * q2[1] + q1[2] * q2[2] + q1[3] * q2[3];
  if (dot < 0.0) { q2[0] = -q2[0]; q2[1] = -q2[1]; q2[2] = -q2[2]; q2[3] = -q2[3]; dot = -dot; }
  if (dot > 0.9995) {
    for (int k = 0; k < 4; k++) q[k] = q1[k] + t * (q2[k] - q1[k]);
  } else {
    const double th0 = acos(dot), th = th0 * t;
    double s, c;
    sincos(th, &s, &c);
    const double s00 = sin(th0);
    const double s0 = c - dot * s / s00, s1 = s / s00;
    for (int k = 0; k < 4; k++) q[k] = s0 * q1[k] + s1 * q2[k];
  }
  sx_normalize(q);                      // R.from_quat
}

template <typename TOut>
__global__ void smplx_stream_kernel(const SmplxArgs a) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= a.N_out) return;
  const int c = a.clip_of[n];
  const int base = a.src_start[c], T = a.src_start[c + 1] - base;
  const double tt = a.time[n];
  const bool al = a.aligned[n] != 0;
  const int i1 = (int)floor(tt), i2 = min(i1 + 1, T - 1);
  const double alpha = tt - (double)i1;
  // scipy interp1d, linear: bracket [lo, hi] = searchsorted(arange(T), t) clipped to [1, T-1]
  int hi = (int)ceil(tt);
  hi = max(1, min(hi, T - 1));
  const int lo = hi - 1;
  double gq[SMPLX_MAX_JOINTS][4];
  TOut* out = (TOut*)a.out + (size_t)n * a.J * 7;
  for (int i = 0; i < a.J; i++) {
    double rv[3], lq[4];
    const float* p1 = a.pose + ((size_t)(base + i1) * a.J + i) * 3;
    if (al) {
      const float* p2 = a.pose + ((size_t)(base + i2) * a.J + i) * 3;
      const double v1[3] = {p1[0], p1[1], p1[2]}, v2[3] = {p2[0], p2[1], p2[2]};
      double q1[4], q2[4], qs[4];
      sx_from_rotvec(v1, q1);
      sx_from_rotvec(v2, q2);
      sx_slerp(q1, q2, alpha, qs);
      sx_as_rotvec(qs, rv);
    } else {
      rv[0] = p1[0]; rv[1] = p1[1]; rv[2] = p1[2];
    }
    sx_from_rotvec(rv, lq);
    if (i == 0) { gq[0][0] = lq[0]; gq[0][1] = lq[1]; gq[0][2] = lq[2]; gq[0][3] = lq[3]; }
    else sx_compose(gq[a.parents[i]], lq, gq[i]);
    double pos[3];
    if (al && T > 1) {
      const float* y0 = a.joints + ((size_t)(base + lo) * a.Jall + i) * 3;
      const float* y1 = a.joints + ((size_t)(base + hi) * a.Jall + i) * 3;
      for (int k = 0; k < 3; k++) pos[k] = ((double)y1[k] - (double)y0[k]) / (double)(hi - lo) * (tt - (double)lo) + (double)y0[k];
    } else {
      const float* y = a.joints + ((size_t)(base + i1) * a.Jall + i) * 3;
      pos[0] = y[0]; pos[1] = y[1]; pos[2] = y[2];
    }
    TOut* o = out + (size_t)i * 7;
    o[0] = (TOut)pos[0]; o[1] = (TOut)pos[1]; o[2] = (TOut)pos[2];
    o[3] = (TOut)gq[i][3]; o[4] = (TOut)gq[i][0]; o[5] = (TOut)gq[i][1]; o[6] = (TOut)gq[i][2];      // wxyz (smpl.py:194)
  }
}

}  // namespace gmr
