// gmr_bvh.cuh - BVH motion -> human frame stream on the device.
//
// Replaces, for all frames of all clips of a folder at once, what the reference does per file in
// numpy and per frame / per bone in Python after the text parse:
//   utils/lafan_vendor/extract.py:162-163   Euler (degrees, channel order) -> quaternion, sign continuity
//   utils/lafan_vendor/utils.py:56-75       euler_to_quat: q = q0 * (q1 * q2), q_k = [cos a/2, sin a/2 * axis_k]
//   utils/lafan_vendor/utils.py:251-269     remove_quat_discontinuities (sequential in time, per bone)
//   utils/lafan_vendor/utils.py:83-103      quat_fk
//   utils/lafan1.py:20-39                   Y-up cm -> Z-up m, orientation = R_quat * q, Left/RightFootMod rows
//   utils/lafan1.py:42-69                   height heuristic on the last frame
//
// Three kernels, float64 like the reference (products and sums are kept unfused and in the reference's operand
// order, so results differ from numpy only through sin/cos):
//   bvh_sign_kernel   thread = (clip, bone): walks the clip in time and records, per frame, whether the
//                     reference would have negated the bone's local quaternion.  Negating a local quaternion
//                     negates the global quaternions of the whole subtree and leaves positions untouched (the
//                     products are bilinear, negation is exact), so only the sign bits are stored.
//   bvh_fk_kernel     thread = frame: local quaternions again from the angles, FK down the bone list (parents
//                     precede children), cumulative signs; globals go to a [N,J,8] scratch.
//   bvh_emit_kernel   thread = (frame, output row): axis change, cm -> m, FootMod rows (position of one bone,
//                     orientation of another), cast to the stream dtype; clip heights by thread = clip.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gmr {

struct BvhArgs {
  const double* lpos;       // [N,J,3]
  const double* euler;      // [N,J,3] degrees
  const int* clip_start;    // [C+1]
  int C, N, J;
  const int* parents;       // [J]
  int ax0, ax1, ax2;        // axis of each rotation channel
  signed char* sign;        // [N,J] scratch: -1 where the reference flips the local quaternion
  double* G;                // [N,J,8] scratch: global quat wxyz, global pos xyz, pad
  const int* src_pos;       // [h]
  const int* src_rot;       // [h]
  int h;
  void* out;                // [N,h,7]
  int head;                 // output row of "Head" or -1
  int feet[4];              // candidate foot rows, -1 = absent
  double* height;           // [C]
};

// numpy operand order, no contraction into FMAs
__device__ __forceinline__ double bmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double badd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double bsub(double a, double b) { return __dsub_rn(a, b); }

// utils.py:132-148 quat_mul(x, y)
__device__ __forceinline__ void bvh_qmul(const double* x, const double* y, double* r) {
  const double r0 = bsub(bsub(bsub(bmul(y[0], x[0]), bmul(y[1], x[1])), bmul(y[2], x[2])), bmul(y[3], x[3]));
  const double r1 = badd(bsub(badd(bmul(y[0], x[1]), bmul(y[1], x[0])), bmul(y[2], x[3])), bmul(y[3], x[2]));
  const double r2 = bsub(badd(badd(bmul(y[0], x[2]), bmul(y[1], x[3])), bmul(y[2], x[0])), bmul(y[3], x[1]));
  const double r3 = badd(badd(bsub(bmul(y[0], x[3]), bmul(y[1], x[2])), bmul(y[2], x[1])), bmul(y[3], x[0]));
  r[0] = r0; r[1] = r1; r[2] = r2; r[3] = r3;
}
// numpy.cross(a, b)
__device__ __forceinline__ void bvh_cross(const double* a, const double* b, double* c) {
  const double c0 = bsub(bmul(a[1], b[2]), bmul(a[2], b[1]));
  const double c1 = bsub(bmul(a[2], b[0]), bmul(a[0], b[2]));
  const double c2 = bsub(bmul(a[0], b[1]), bmul(a[1], b[0]));
  c[0] = c0; c[1] = c1; c[2] = c2;
}
// utils.py:151-163 quat_mul_vec
__device__ __forceinline__ void bvh_qrot(const double* q, const double* v, double* r) {
  double t[3], u[3];
  bvh_cross(q + 1, v, t);
  t[0] = bmul(2.0, t[0]); t[1] = bmul(2.0, t[1]); t[2] = bmul(2.0, t[2]);
  bvh_cross(q + 1, t, u);
  for (int k = 0; k < 3; k++) r[k] = badd(badd(v[k], bmul(q[0], t[k])), u[k]);
}
// utils.py:40-75 with e in degrees (extract.py:162 np.radians first)
__device__ __forceinline__ void bvh_local_quat(const double* e_deg, int ax0, int ax1, int ax2, double* q) {
  const double d2r = 3.14159265358979323846 / 180.0;
  double q0[4], q1[4], q2[4], t[4];
  double s, c;
  sincos(bmul(e_deg[0], d2r) / 2.0, &s, &c);
  q0[0] = c; q0[1] = ax0 == 0 ? s : 0.0; q0[2] = ax0 == 1 ? s : 0.0; q0[3] = ax0 == 2 ? s : 0.0;
  sincos(bmul(e_deg[1], d2r) / 2.0, &s, &c);
  q1[0] = c; q1[1] = ax1 == 0 ? s : 0.0; q1[2] = ax1 == 1 ? s : 0.0; q1[3] = ax1 == 2 ? s : 0.0;
  sincos(bmul(e_deg[2], d2r) / 2.0, &s, &c);
  q2[0] = c; q2[1] = ax2 == 0 ? s : 0.0; q2[2] = ax2 == 1 ? s : 0.0; q2[3] = ax2 == 2 ? s : 0.0;
  bvh_qmul(q1, q2, t);
  bvh_qmul(q0, t, q);
}

__global__ void bvh_sign_kernel(const BvhArgs a) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= a.C * a.J) return;
  const int c = id / a.J, j = id - c * a.J;
  const int n0 = a.clip_start[c], n1 = a.clip_start[c + 1];
  double prev[4] = {0, 0, 0, 0};
  for (int n = n0; n < n1; n++) {
    double q[4];
    bvh_local_quat(a.euler + ((size_t)n * a.J + j) * 3, a.ax0, a.ax1, a.ax2, q);
    signed char sg = 1;
    if (n > n0) {
      // np.sum over 4 products, left to right; the reference compares it with the sum for -q, which is its negation
      const double d = badd(badd(badd(bmul(prev[0], q[0]), bmul(prev[1], q[1])), bmul(prev[2], q[2])), bmul(prev[3], q[3]));
      if (d < -d) sg = -1;
    }
    a.sign[(size_t)n * a.J + j] = sg;
    for (int k = 0; k < 4; k++) prev[k] = sg < 0 ? -q[k] : q[k];
  }
}

__global__ void bvh_fk_kernel(const BvhArgs a) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= a.N) return;
  double* G = a.G + (size_t)n * a.J * 8;
  for (int j = 0; j < a.J; j++) {
    double lq[4], gq[4], gp[3];
    bvh_local_quat(a.euler + ((size_t)n * a.J + j) * 3, a.ax0, a.ax1, a.ax2, lq);
    if (a.sign[(size_t)n * a.J + j] < 0) { lq[0] = -lq[0]; lq[1] = -lq[1]; lq[2] = -lq[2]; lq[3] = -lq[3]; }
    const double* lp = a.lpos + ((size_t)n * a.J + j) * 3;
    const int p = a.parents[j];
    if (j == 0 || p < 0) {
      for (int k = 0; k < 4; k++) gq[k] = lq[k];
      for (int k = 0; k < 3; k++) gp[k] = lp[k];
    } else {
      const double* P = G + (size_t)p * 8;
      const double pq[4] = {P[0], P[1], P[2], P[3]}, v[3] = {lp[0], lp[1], lp[2]};
      double r[3];
      bvh_qrot(pq, v, r);
      for (int k = 0; k < 3; k++) gp[k] = badd(r[k], P[4 + k]);
      bvh_qmul(pq, lq, gq);
    }
    double* o = G + (size_t)j * 8;
    o[0] = gq[0]; o[1] = gq[1]; o[2] = gq[2]; o[3] = gq[3]; o[4] = gp[0]; o[5] = gp[1]; o[6] = gp[2]; o[7] = 0.0;
  }
}

#define BVH_ROT_W 0x1.6a09e667f3bccp-1     // scipy Rotation.from_matrix(...).as_quat(scalar_first=True)[0] (lafan1.py:20-21)

template <typename TOut>
__global__ void bvh_emit_kernel(const BvhArgs a) {
  const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= (long long)a.N * a.h) return;
  const int n = (int)(id / a.h), k = (int)(id - (long long)n * a.h);
  const double* Gp = a.G + ((size_t)n * a.J + a.src_pos[k]) * 8;
  const double* Gq = a.G + ((size_t)n * a.J + a.src_rot[k]) * 8;
  const double rq[4] = {BVH_ROT_W, BVH_ROT_W, 0.0, 0.0}, gq[4] = {Gq[0], Gq[1], Gq[2], Gq[3]};
  double q[4];
  bvh_qmul(rq, gq, q);                                          // lafan1.py:31
  TOut* o = (TOut*)a.out + (size_t)id * 7;
  o[0] = (TOut)(Gp[4] / 100.0);                                 // lafan1.py:32: p @ R^T / 100 = (x, -z, y) / 100
  o[1] = (TOut)(-Gp[6] / 100.0);
  o[2] = (TOut)(Gp[5] / 100.0);
  o[3] = (TOut)q[0]; o[4] = (TOut)q[1]; o[5] = (TOut)q[2]; o[6] = (TOut)q[3];
}

__global__ void bvh_height_kernel(const BvhArgs a) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= a.C) return;
  const int n0 = a.clip_start[c], n1 = a.clip_start[c + 1];
  double hgt = 1.75;
  if (n1 > n0) {
    const double* G = a.G + (size_t)(n1 - 1) * a.J * 8;
    double zmin = INFINITY, zmax = -INFINITY;
    for (int k = 0; k < a.h; k++) {
      const double z = G[(size_t)a.src_pos[k] * 8 + 5] / 100.0;
      zmin = fmin(zmin, z); zmax = fmax(zmax, z);
    }
    if (a.head >= 0) {
      const double head_z = G[(size_t)a.src_pos[a.head] * 8 + 5] / 100.0;
      double fmin_ = INFINITY;
      bool any = false;
      for (int f = 0; f < 4; f++)
        if (a.feet[f] >= 0) { fmin_ = fmin(fmin_, G[(size_t)a.src_pos[a.feet[f]] * 8 + 5] / 100.0); any = true; }
      hgt = head_z - (any ? fmin_ : zmin);
    } else {
      hgt = zmax - zmin;
    }
    if (!isfinite(hgt) || hgt < 0.9 || hgt > 2.3) hgt = 1.75;
  }
  a.height[c] = hgt;
}

}  // namespace gmr
