// gmr_stages.cuh - the warp-level stages of one retarget() call.
//
//   stage_prep      Kernel 1  scale + offset (+ground)        motion_retarget.py:209-270
//   stage_fk        Kernel 2  mj_kinematics + spatial axes    (mink Configuration.update)
//   stage_errors    Kernel 2  FrameTask.compute_error         motion_retarget.py:188-200
//   stage_stiffness Kernel 2  FrameTask.compute_jacobian, folded into a 6x6 stiffness
//   stage_assemble  Kernel 3  mink build_ik: H, c, box
//   qp_solve_aba    Kernel 3  exact box QP on the kinematic tree, articulated-body sweeps (stands in for daqp)
//   stage_integrate Kernel 4  mj_integratePos
// Kernel 4's convergence control lives in gmr_clip.cuh.
//
// Every function is called by all 32 lanes of the owning warp with warp-uniform arguments; `sm` is
// the warp's private shared-memory block laid out by `Lay`; `mv` gives the model tables (staged in
// shared memory or read from global memory, see ModelView).  Everything is force-inlined into ONE
// call site per stage of the tick loop, so that the compiler sees every pointer derive from the
// kernel's __shared__ array (LDS/STS with 32-bit addresses) and the instruction stream holds one
// copy of each stage.
#pragma once
#include "gmr_device.cuh"

namespace gmr {

#define LIE_EPS 1e-10   // mink get_epsilon(float64)

// ------------------------------------------------------------------ Kernel 1: target preparation
// lane i < s handles scaled body i (entry 0 = human root).  Reads one frame of the human
// stream (7 numbers per body, float or double), writes sm[scaled + 7*i .. +7] = pos, quat wxyz.
// Returns whether every number the frame contributed is finite (warp-uniform).
template <typename TIn, class MV>
__device__ __forceinline__ bool stage_prep(const MV& mv, const Lay& lay, double* sm, const TIn* frame,
                                           const double* scale, bool to_ground, int lane) {
  const DevModel& M = mv.M();
  const int s = M.n_scaled;
  double p[3] = {0, 0, 0}, q[4] = {1, 0, 0, 0};
  // the whole frame comes in with coalesced streaming loads (it may live in pinned host memory: one
  // burst of full lines instead of per-body gathers), staged in the K/g scratch, which is dead here
  TIn* fbuf = (TIn*)(sm + lay.K);
  const int ne = 7 * M.n_human;
  for (int i = lane; i < ne; i += 32) fbuf[i] = __ldcs(frame + i);
  __syncwarp();
  if (lane < s) {
    const TIn* b = fbuf + 7 * mv(M.scaled_index)[lane];
    p[0] = (double)b[0]; p[1] = (double)b[1]; p[2] = (double)b[2];
    q[0] = (double)b[3]; q[1] = (double)b[4]; q[2] = (double)b[5]; q[3] = (double)b[6];
  }
  // root position to every lane
  double r0 = __shfl_sync(GMR_FULL, p[0], 0), r1 = __shfl_sync(GMR_FULL, p[1], 0), r2 = __shfl_sync(GMR_FULL, p[2], 0);
  double out[7] = {0, 0, 0, 1, 0, 0, 0};
  if (lane < s) {
    double sc = scale[lane], s0 = scale[0];
    double pos[3];
    if (lane == 0) { pos[0] = s0 * r0; pos[1] = s0 * r1; pos[2] = s0 * r2; }
    else { pos[0] = (p[0] - r0) * sc + s0 * r0; pos[1] = (p[1] - r1) * sc + s0 * r1; pos[2] = (p[2] - r2) * sc + s0 * r2; }
    // scipy Rotation normalises on from_quat and the product is unit again
    double n = rsqrt_pos(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    q[0] *= n; q[1] *= n; q[2] *= n; q[3] *= n;
    const double* rop = mv(M.rot_off) + 4 * lane;
    double ro[4] = {rop[0], rop[1], rop[2], rop[3]};
    double uq[4];
    qmul(q, ro, uq);
    n = rsqrt_pos(uq[0] * uq[0] + uq[1] * uq[1] + uq[2] * uq[2] + uq[3] * uq[3]);
    uq[0] *= n; uq[1] *= n; uq[2] *= n; uq[3] *= n;
    const double* pop = mv(M.pos_off) + 3 * lane;
    double po[3] = {pop[0], pop[1], pop[2]}, off[3];
    qrot(uq, po, off);
    out[0] = pos[0] + off[0]; out[1] = pos[1] + off[1]; out[2] = pos[2] + off[2];
    out[3] = uq[0]; out[4] = uq[1]; out[5] = uq[2]; out[6] = uq[3];
  }
  if (to_ground) {   // motion_retarget.py:252-270
    double z = (lane < s && mv(M.is_foot)[lane]) ? out[2] : INFINITY;
    double lowest = warp_min(z);
    out[2] = out[2] - lowest + 0.1;
  }
  if (lane < s) {
    double* d = sm + lay.scaled + 7 * lane;
#pragma unroll
    for (int k = 0; k < 7; k++) d[k] = out[k];
  }
  const bool fin = isfinite(((out[0] + out[1]) + (out[2] + out[3])) + ((out[4] + out[5]) + out[6]));
  __syncwarp();
  return __all_sync(GMR_FULL, fin);
}

// ------------------------------------------------------------------ Kernel 2: forward kinematics
// mj_kinematics in three passes:
//   A (lane = body, all bodies at once): the body's LOCAL transform, joints included -
//       tl = body_pos + R(body_quat) sum(slide axis * value),  ql = body_quat (x) prod(hinge rotations),
//       and each joint axis expressed in the PARENT frame (al, parked in S[k][3..5]);
//   B (pointer jumping, lane = body): compose every body's transform with that of its 2^r-th ancestor,
//       r = 0, 1, ...: ceil(log2(tree depth)) rounds over all bodies instead of one pass per level;
//   C (lane = dof): spatial axis S_k = (v_O, w): twist of the dof expressed at the reference point O =
//       origin of the floating base (world origin if there is none), so that the geometric Jacobian
//       column of ANY descendant frame at x is [v_O + w x (x - O); w].
// Joint anchors are the body origin (jnt_pos = 0, checked at model creation) and slides precede
// hinges inside a body, which is what makes the local transform independent of the parent.
// ALL = false: only the bodies the IK needs are swept (M.fk_body: bodies of scheduled joints, task frames, their
// ancestors - 30 of 39 for unitree_g1; at most 32 for every robot of the pack, so a lane keeps its body's transform
// in registers from pass A to the normalisation); the poses of the other bodies are then stale.
// ALL = true (gmr_fk, diagnostics): every body.
template <bool ALL, class MV>
__device__ __forceinline__ void stage_fk(const MV& mv, const Lay& lay, double* sm, int lane) {
  const DevModel& M = mv.M();
  const double* q = sm + lay.q;
  double* xpos = sm + lay.xpos;
  double* xquat = sm + lay.xquat;
  double* S = sm + lay.S;
  double* tl = sm + lay.tl;
  double* ql = sm + lay.ql;
  const int nb = M.nbody;
  // pass B ping-pongs between (tl, ql) and (xpos, xquat); start where an even number of swaps ends in x
  const int R = M.fk_rounds;
  double* const T_even = (R & 1) ? tl : xpos;
  double* const T_odd = (R & 1) ? xpos : tl;
  double* const Q_even = (R & 1) ? ql : xquat;
  double* const Q_odd = (R & 1) ? xquat : ql;
  const int nfk = ALL ? nb : M.n_fk;
  const short* fkb = mv(M.fk_body);
  // local transform of body b (relative to its parent) with its joints folded in; the joint axes in the parent frame go to S
  auto local_tf = [&](const int b, double (&t)[3], double (&qq)[4]) {
    if (b > 0) {
      const int jadr = mv(M.body_jntadr)[b], jnum = mv(M.body_jntnum)[b];
      if (jnum == 1 && mv(M.jnt_type)[jadr] == 0) {                     // free joint: pose straight from q
        const int qa = mv(M.jnt_qposadr)[jadr];
        t[0] = q[qa]; t[1] = q[qa + 1]; t[2] = q[qa + 2];
        qq[0] = q[qa + 3]; qq[1] = q[qa + 4]; qq[2] = q[qa + 5]; qq[3] = q[qa + 6];
        normalize4(qq);            // qpos may hold any quaternion; everything below needs a rotation
      } else {
        const double* bpp = mv(M.body_pos) + 3 * b;
        const double* bqp = mv(M.body_quat) + GMR_QS * b;
        t[0] = bpp[0]; t[1] = bpp[1]; t[2] = bpp[2];
        qq[0] = bqp[0]; qq[1] = bqp[1]; qq[2] = bqp[2]; qq[3] = bqp[3];
        for (int j = jadr; j < jadr + jnum; j++) {
          const int qa = mv(M.jnt_qposadr)[j], da = mv(M.jnt_dofadr)[j];
          const double* axp = mv(M.jnt_axis) + 3 * j;
          const double ax[3] = {axp[0], axp[1], axp[2]};
          double al[3];
          qrot(qq, ax, al);                                             // axis in the parent frame
          S[GMR_SS * da + 3] = al[0]; S[GMR_SS * da + 4] = al[1]; S[GMR_SS * da + 5] = al[2];
          const double val = q[qa] - mv(M.qpos0)[qa];
          if (mv(M.jnt_type)[j] == 2) {                                 // slide
            t[0] += al[0] * val; t[1] += al[1] * val; t[2] += al[2] * val;
          } else {                                                      // hinge about the body origin
            double sh, ch;
            sincos(0.5 * val, &sh, &ch);
            const double qj[4] = {ch, sh * ax[0], sh * ax[1], sh * ax[2]};
            double nq[4];
            qmul(qq, qj, nq);
            qq[0] = nq[0]; qq[1] = nq[1]; qq[2] = nq[2]; qq[3] = nq[3];
          }
        }
      }
    }
  };
  const short* jump = mv(M.fk_jump);
  if (!ALL && nfk <= 32) {
    // ---- one body per lane (every robot of the pack): the body's transform stays in registers from pass A through the
    // rounds of pass B to the normalisation; shared memory only carries it to the lanes that compose with it
    const bool on = lane < nfk;
    const int b = on ? fkb[lane] : 0;
    double t[3] = {0, 0, 0}, qq[4] = {1, 0, 0, 0};
    if (on) {
      local_tf(b, t, qq);
      if (R == 0) normalize4(qq);
      double* T0 = T_even + 3 * b;
      double* Q0 = Q_even + GMR_QS * b;
      T0[0] = t[0]; T0[1] = t[1]; T0[2] = t[2];
      Q0[0] = qq[0]; Q0[1] = qq[1]; Q0[2] = qq[2]; Q0[3] = qq[3];
    }
    __syncwarp();
    for (int r = 0; r < R; r++) {
      const double* Ti = (r & 1) ? T_odd : T_even;
      const double* Qi = (r & 1) ? Q_odd : Q_even;
      double* To = (r & 1) ? T_even : T_odd;
      double* Qo = (r & 1) ? Q_even : Q_odd;
      if (on) {
        const int anc2 = jump[r * nb + b];                              // 2^r-th ancestor, -1: already in the world frame
        if (anc2 >= 0) {
          const double aq[4] = {Qi[GMR_QS * anc2], Qi[GMR_QS * anc2 + 1], Qi[GMR_QS * anc2 + 2], Qi[GMR_QS * anc2 + 3]};
          double rt[3], nq[4];
          qrot(aq, t, rt);
          t[0] = Ti[3 * anc2] + rt[0]; t[1] = Ti[3 * anc2 + 1] + rt[1]; t[2] = Ti[3 * anc2 + 2] + rt[2];
          qmul(aq, qq, nq);
          qq[0] = nq[0]; qq[1] = nq[1]; qq[2] = nq[2]; qq[3] = nq[3];
        }
        if (r == R - 1) normalize4(qq);                                 // mju_normalize4, once, after the last composition
        To[3 * b] = t[0]; To[3 * b + 1] = t[1]; To[3 * b + 2] = t[2];
        Qo[GMR_QS * b] = qq[0]; Qo[GMR_QS * b + 1] = qq[1]; Qo[GMR_QS * b + 2] = qq[2]; Qo[GMR_QS * b + 3] = qq[3];
      }
      __syncwarp();
    }
    if (lane == 31) { xpos[0] = xpos[1] = xpos[2] = 0.0; xquat[0] = 1.0; xquat[1] = xquat[2] = xquat[3] = 0.0; }      // the world
    __syncwarp();
  } else {
    // ---- pass A: local transform of every body (relative to its parent)
    for (int i = lane; i < nfk; i += 32) {
      const int b = ALL ? i : fkb[i];
      double t[3] = {0, 0, 0}, qq[4] = {1, 0, 0, 0};
      local_tf(b, t, qq);
      double* T0 = T_even + 3 * b;
      double* Q0 = Q_even + GMR_QS * b;
      T0[0] = t[0]; T0[1] = t[1]; T0[2] = t[2];
      Q0[0] = qq[0]; Q0[1] = qq[1]; Q0[2] = qq[2]; Q0[3] = qq[3];
    }
    __syncwarp();
    // ---- pass B: pointer jumping.  Before round r a body's transform is relative to its 2^r-th
    // ancestor (or already to the world); composing it with that ancestor's transform doubles the
    // reach, so ceil(log2(depth)) rounds over ALL bodies replace `depth` sequential levels.
    for (int r = 0; r < R; r++) {
      const double* Ti = (r & 1) ? T_odd : T_even;
      const double* Qi = (r & 1) ? Q_odd : Q_even;
      double* To = (r & 1) ? T_even : T_odd;
      double* Qo = (r & 1) ? Q_even : Q_odd;
      for (int i = lane; i < nfk; i += 32) {
        const int b = ALL ? i : fkb[i];
        const int anc2 = jump[r * nb + b];                                // 2^r-th ancestor, -1: already in the world frame
        double t[3] = {Ti[3 * b], Ti[3 * b + 1], Ti[3 * b + 2]};
        double qq[4] = {Qi[GMR_QS * b], Qi[GMR_QS * b + 1], Qi[GMR_QS * b + 2], Qi[GMR_QS * b + 3]};
        if (anc2 >= 0) {
          const double aq[4] = {Qi[GMR_QS * anc2], Qi[GMR_QS * anc2 + 1], Qi[GMR_QS * anc2 + 2], Qi[GMR_QS * anc2 + 3]};
          double rt[3], nq[4];
          qrot(aq, t, rt);
          t[0] = Ti[3 * anc2] + rt[0]; t[1] = Ti[3 * anc2 + 1] + rt[1]; t[2] = Ti[3 * anc2 + 2] + rt[2];
          qmul(aq, qq, nq);
          qq[0] = nq[0]; qq[1] = nq[1]; qq[2] = nq[2]; qq[3] = nq[3];
        }
        To[3 * b] = t[0]; To[3 * b + 1] = t[1]; To[3 * b + 2] = t[2];
        Qo[GMR_QS * b] = qq[0]; Qo[GMR_QS * b + 1] = qq[1]; Qo[GMR_QS * b + 2] = qq[2]; Qo[GMR_QS * b + 3] = qq[3];
      }
      __syncwarp();
    }
    // mju_normalize4 of every body quaternion, all bodies at once (the products above keep unit norm
    // to rounding, so normalising after the sweep instead of inside it changes results by ~1e-16)
    if (ALL) { for (int i = 1 + lane; i < nfk; i += 32) normalize4(xquat + GMR_QS * i); }          // body 0 is the world
    else {
      for (int i = lane; i < nfk; i += 32) normalize4(xquat + GMR_QS * fkb[i]);                    // the list leaves the world out
      if (lane == 31) { xpos[0] = xpos[1] = xpos[2] = 0.0; xquat[0] = 1.0; xquat[1] = xquat[2] = xquat[3] = 0.0; }
    }
    __syncwarp();
  }
  // ---- pass C
  const int ref = M.ref_body;
  const double O[3] = {xpos[3 * ref], xpos[3 * ref + 1], xpos[3 * ref + 2]};    // world (body 0) has xpos = 0
  const int skip = M.fk_skip;          // a floating base's translations: S = unit vectors, written by three lanes below
  if (lane < skip) {
    double* sk = S + GMR_SS * lane;
    sk[0] = (lane == 0); sk[1] = (lane == 1); sk[2] = (lane == 2); sk[3] = sk[4] = sk[5] = 0;
  }
  for (int i = lane; i < 6 * M.n_udof; i += 32) S[GMR_SS * mv(M.udof)[i / 6] + i % 6] = 0.0;   // never scheduled: no axis
  const short* cdof = mv(M.cdof);
  for (int ci = lane; ci < M.n_cdof; ci += 32) {
    const int k = cdof[ci];
    const int kind = mv(M.dof_kind)[k], b = mv(M.dof_body)[k];
    double* sk = S + GMR_SS * k;
    if (kind == DOF_FREE_T) {
      const int i = k - mv(M.jnt_dofadr)[mv(M.body_jntadr)[b]];
      sk[0] = (i == 0); sk[1] = (i == 1); sk[2] = (i == 2); sk[3] = sk[4] = sk[5] = 0;
    } else {
      double aw[3];
      if (kind == DOF_FREE_R) {
        const int i = k - mv(M.jnt_dofadr)[mv(M.body_jntadr)[b]] - 3;
        const double ei[3] = {i == 0 ? 1.0 : 0.0, i == 1 ? 1.0 : 0.0, i == 2 ? 1.0 : 0.0};
        qrot(xquat + GMR_QS * b, ei, aw);                               // column i of R(Q_root)
      } else {
        const int p = mv(M.body_parent)[b];
        const double al[3] = {sk[3], sk[4], sk[5]};
        qrot(xquat + GMR_QS * p, al, aw);
      }
      if (kind == DOF_SLIDE) {
        sk[0] = aw[0]; sk[1] = aw[1]; sk[2] = aw[2]; sk[3] = sk[4] = sk[5] = 0;
      } else {                                                          // rotation about the body origin
        const double r[3] = {O[0] - xpos[3 * b], O[1] - xpos[3 * b + 1], O[2] - xpos[3 * b + 2]};
        double v[3];
        cross3(aw, r, v);
        sk[0] = v[0]; sk[1] = v[1]; sk[2] = v[2]; sk[3] = aw[0]; sk[4] = aw[1]; sk[5] = aw[2];
      }
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------------ mink SO3/SE3 log
// Also returns sin/cos of the rotation angle th = |w| (read straight off the quaternion: with phi = th/2,
// sin phi : cos phi = |xyz| : |qw|), which spares the sincos of the V^-1 coefficient here and the one of
// the Jlog coefficients in stage_stiffness.
__device__ __forceinline__ void se3_log_dev(const double* q, const double* p, double* out, double& sin_th, double& cos_th) {
  const double qw = q[0], n2 = q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  const double aw = fabs(qw);
  const double rn = n2 > 1e-30 ? rsqrt_pos(n2) : 0.0, n = n2 * rn;
  double f, k;
  if (n2 < LIE_EPS) {                       // mink SO3.log: Taylor branch; th^2 = f^2 n2 < 4.1e-10, where
    f = 2.0 / qw - 2.0 / 3.0 * n2 / (qw * qw * qw);       // (1 - th/2 cot(th/2)) / th^2 = 1/12 + th^2/720 to 1e-21
    k = 1.0 / 12.0 + (f * f * n2) * (1.0 / 720.0);
  } else {
    // phi = atan2(n, |qw|) in [0, pi/2] (half the rotation angle); mink: 2 atan2(qw < 0 ? -n : n, |qw|) / n
    double phi;
    if (aw < LIE_EPS) phi = 1.5707963267948966;
    else phi = n <= aw ? atan_ratio(n, aw) : 1.5707963267948966 - atan_ratio(aw, n);
    f = (qw < 0 ? -2.0 : 2.0) * phi * rn;
    // |w| = 2 phi, so V^-1's coefficient (1 - th/2 cot(th/2)) / th^2 needs no second root: cot(phi) = |qw| / n
    k = (1.0 - phi * (aw * rn)) * rcp_pos(4.0 * phi * phi);
  }
  const double in2 = rcp_pos(n2 + qw * qw);
  sin_th = 2.0 * n * aw * in2;
  cos_th = (qw * qw - n2) * in2;
  const double w0 = f * q[1], w1 = f * q[2], w2 = f * q[3];
  // V^-1 p = p - 1/2 w x p + k w x (w x p)
  double w[3] = {w0, w1, w2}, c1[3], c2[3];
  cross3(w, p, c1);
  cross3(w, c1, c2);
  out[0] = p[0] - 0.5 * c1[0] + k * c2[0];
  out[1] = p[1] - 0.5 * c1[1] + k * c2[1];
  out[2] = p[2] - 0.5 * c1[2] + k * c2[2];
  out[3] = w0; out[4] = w1; out[5] = w2;
}

// ------------------------------------------------------------------ Kernel 2: task errors
// lane t < m: e_t = log(T_b^-1 T_t) -> sm[e + GMR_ES t]; returns ||concat e|| (unweighted, :188-200),
// bitwise identical on all lanes.
template <class MV>
__device__ __forceinline__ double stage_errors(const MV& mv, const Lay& lay, double* sm, int stage, int lane) {
  const DevModel& M = mv.M();
  const int m = M.n_task[stage];
  double ss = 0;
  if (lane < m) {
    const int body = mv(M.task_body)[stage * lay.maxt + lane];
    const double* tgt = sm + lay.scaled + 7 * mv(M.task_src)[stage * lay.maxt + lane];
    const double* xp = sm + lay.xpos + 3 * body;
    const double* xq = sm + lay.xquat + GMR_QS * body;
    double qi[4] = {xq[0], -xq[1], -xq[2], -xq[3]}, qbt[4], d[3] = {tgt[0] - xp[0], tgt[1] - xp[1], tgt[2] - xp[2]}, pbt[3], e[6];
    qmul(qi, tgt + 3, qbt);
    qrot(qi, d, pbt);
    double sn, cs;
    se3_log_dev(qbt, pbt, e, sn, cs);
    double* eo = sm + lay.e + GMR_ES * lane;
#pragma unroll
    for (int k = 0; k < 6; k++) { eo[k] = e[k]; ss += e[k] * e[k]; }
    eo[6] = sn; eo[7] = cs;               // for stage_stiffness
  }
  ss = warp_sum(ss);
  __syncwarp();
  return sqrt(ss);
}

// ------------------------------------------------------------------ Kernel 2: task Jacobians as stiffness
// mink: J_t = -Jl^-1(e_t) [R^T Jp; R^T Jr]  (FrameTask.compute_jacobian) and
//       H += (W J_t)^T (W J_t), c += (W e_t)^T (W J_t)   (Task.compute_qp_objective).
// The geometric Jacobian column of dof k at frame t is T(r_t) S_k with r_t = x_t - O and
// T(r) = [[I, -[r]x],[0, I]], so W J_t[:,k] = N_t S_k with the 6x6
//       N_t = -[[wp P, wp Pr],[0, wr P]],  P = Ji R^T,  Pr = -P [r]x + B R^T,
// where Jl^-1(e) = [[Ji, B],[0, Ji]].  Lane t stores K_t = N_t^T N_t (21 packed) and
// g_t = N_t^T W e_t (6) as one block of GMR_KS = 27 doubles; the assembly then needs only S_k^T (sum K_t) S_l - no 6 x nv Jacobian.
// Returns sum_t mu_t = sum_t ||W e_t||^2 (all lanes).
template <class MV>
__device__ __forceinline__ double stage_stiffness(const MV& mv, const Lay& lay, double* sm, int stage, int lane) {
  const DevModel& M = mv.M();
  const int m = M.n_task[stage];
  double mu = 0;
  if (lane < m) {
    const int body = mv(M.task_body)[stage * lay.maxt + lane];
    const double wp = mv(M.task_wpos)[stage * lay.maxt + lane], wr = mv(M.task_wrot)[stage * lay.maxt + lane];
    const double a = wp * wp, b = wr * wr;
    const double* e = sm + lay.e + GMR_ES * lane;
    const double ev[3] = {e[0], e[1], e[2]}, w[3] = {e[3], e[4], e[5]};
    mu = a * (ev[0] * ev[0] + ev[1] * ev[1] + ev[2] * ev[2]) + b * (w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    double R[9], Rt[9];
    q2mat(sm + lay.xquat + GMR_QS * body, R);
    Rt[0] = R[0]; Rt[1] = R[3]; Rt[2] = R[6]; Rt[3] = R[1]; Rt[4] = R[4]; Rt[5] = R[7]; Rt[6] = R[2]; Rt[7] = R[5]; Rt[8] = R[8];
    double P[9], Pr[9];
    if (th2 < LIE_EPS) {            // mink SE3.ljacinv returns the identity here
#pragma unroll
      for (int i = 0; i < 9; i++) { P[i] = Rt[i]; Pr[i] = 0; }
    } else {
      const double ith = rsqrt_pos(th2), th = th2 * ith, ith2 = ith * ith;       // th2 >= LIE_EPS here
      const double s = e[6], c = e[7];                                         // sin th, cos th from stage_errors
      const double A = ith2 * (1.0 - (th * s * 0.5 * rcp_pos(1.0 - c)));
      // products of skew matrices in closed form: [a]x[b]x = b a' - (a.b) I, so with cw = w.v, u = v x w:
      //   W^2 = w w' - th2 I,  WV = v w' - cw I,  VW = w v' - cw I,  WVW = -cw W,
      //   VWW = u w' - th2 V,  WWV = -(VWW)',  WVWW + WWVW = -2 cw W^2
      double Wm[9], V[9], Ji[9], Q[9], u[3];
      skew3(w, Wm);
      skew3(ev, V);
      cross3(ev, w, u);
      const double cw = w[0] * ev[0] + w[1] * ev[1] + w[2] * ev[2];
      // Barfoot Q(v, w)
      const double ith3 = ith2 * ith;
      const double Bc = (th - s) * ith3, Cc = (1.0 - 0.5 * th2 - c) * (ith2 * ith2),
                   Dc = (2.0 * th - 3.0 * s + th * c) * (0.5 * ith3 * ith2);
      const double kV = 0.5 + 2.0 * Cc * th2, kW = -cw * (Bc + 3.0 * Cc), kI = -2.0 * cw * (Bc - Dc * th2), kWW = -2.0 * Dc * cw;
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
          const int ij = 3 * i + j;
          const double dlt = i == j ? 1.0 : 0.0;
          Ji[ij] = dlt * (1.0 - A * th2) - 0.5 * Wm[ij] + A * (w[i] * w[j]);
          Q[ij] = kV * V[ij] + kW * Wm[ij] + dlt * kI + Bc * (ev[i] * w[j] + w[i] * ev[j]) - Cc * (u[i] * w[j] - w[i] * u[j]) + kWW * (w[i] * w[j]);
        }
      double T[9], TP[9];
      mm3(Ji, Q, T);
      mm3(Ji, Rt, P);
      mm3(T, P, TP);                // B R' = -Ji Q Ji R' = -T P
#pragma unroll
      for (int i = 0; i < 9; i++) Pr[i] = -TP[i];
    }
    // Pr += -P [r]x
    const int ref = M.ref_body;
    const double* xp = sm + lay.xpos + 3 * body;
    const double* xo = sm + lay.xpos + 3 * ref;     // world (body 0) has xpos = 0
    const double r[3] = {xp[0] - xo[0], xp[1] - xo[1], xp[2] - xo[2]};
#pragma unroll
    for (int i = 0; i < 3; i++) {     // row i of P [r]x is p_i x r
      double pr[3];
      cross3(P + 3 * i, r, pr);
      Pr[3 * i] -= pr[0]; Pr[3 * i + 1] -= pr[1]; Pr[3 * i + 2] -= pr[2];
    }
    // K = [[a P'P, a P'Pr],[., a Pr'Pr + b P'P]]
    double PtP[9], PtPr[9], PrtPr[9];
    mtm3(P, P, PtP);
    mtm3(P, Pr, PtPr);
    mtm3(Pr, Pr, PrtPr);
    double* K = sm + lay.K + GMR_KS * lane;      // block of the task: 21 packed K, then g
    // packed lower 6x6: (i,j) -> i(i+1)/2 + j
    K[0] = a * PtP[0];
    K[1] = a * PtP[3]; K[2] = a * PtP[4];
    K[3] = a * PtP[6]; K[4] = a * PtP[7]; K[5] = a * PtP[8];
#pragma unroll
    for (int i = 0; i < 3; i++) {       // rows 3..5: [ (a P'Pr)^T | a Pr'Pr + b P'P ]
      const int base = (3 + i) * (4 + i) / 2;
#pragma unroll
      for (int j = 0; j < 3; j++) K[base + j] = a * PtPr[3 * j + i];
#pragma unroll
      for (int j = 0; j <= i; j++) K[base + 3 + j] = a * PrtPr[3 * i + j] + b * PtP[3 * i + j];
    }
    // g = -[a P' ev ; a Pr' ev + b P' w]
    double* g = K + 21;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      g[i] = -a * (P[i] * ev[0] + P[3 + i] * ev[1] + P[6 + i] * ev[2]);
      g[3 + i] = -(a * (Pr[i] * ev[0] + Pr[3 + i] * ev[1] + Pr[6 + i] * ev[2]) + b * (P[i] * w[0] + P[3 + i] * w[1] + P[6 + i] * w[2]));
    }
  }
  mu = warp_sum(mu);
  __syncwarp();
  return mu;
}

// ------------------------------------------------------------------ Kernel 3: assembly (mink build_ik)
//   H(l,k) = S_k' Kc_l S_l  for k an ancestor-or-self of l,   c_l = S_l . gc_l,
// where Kc_l / gc_l sum K_t / g_t over the tasks in the subtree of dof l; everything else is
// structurally zero and not stored.  diag += damping + lm_damping * sum mu.  Three passes:
//   1. subtree sums on the TASK tree, in place: bottom-up, K_t += K_child for every task whose body
//      hangs below t's body (lane = (parent task, one of the 27 numbers); a handful of passes);
//   2. lane = dof l: Kc_l = sum over the few top-most tasks of l's subtree (1-3 for a humanoid),
//      F_l = Kc_l S_l -> shared memory, c_l, and the ConfigurationLimit box (registers);
//   3. lane = stored entry (l,k) of the tree layout: H = S_k . F_l (+ diag) - all nnz entries in
//      nnz/32 passes, whatever the depth of their row.
template <class MV>
__device__ __forceinline__ void stage_assemble(const MV& mv, const Lay& lay, double* sm, int stage, double mu_sum, int lane,
                                               double (&c_out)[2], double (&lo_out)[2], double (&hi_out)[2]) {
  const DevModel& M = mv.M();
  const int n = M.nv;
  double* H = sm + lay.H;
  double* Kt = sm + lay.K;                  // blocks [K_t (21) | g_t (6)]
  double* F = sm + lay.F;
  const double diag = M.damping + M.lm_damping * mu_sum;
  const double* q = sm + lay.q;
  // ---- 1. subtree sums over the task tree (parents sorted by height, children as bit masks)
  const int* tl_start = mv(M.tasklev_start) + stage * (lay.maxt + 1);
  const int* tl_task = mv(M.tasklev_task) + stage * lay.maxt;
  const unsigned* child_mask = mv(M.task_childmask) + stage * lay.maxt;
  const int nlev = M.n_tasklev[stage];
  for (int lv = 0; lv < nlev; lv++) {
    const int beg = tl_start[lv], cnt = (tl_start[lv + 1] - beg) * 27;
    for (int it = lane; it < cnt; it += 32) {
      const int t = tl_task[beg + it / 27], e = it % 27;
      double* dst = Kt + GMR_KS * t + e;
      double acc = *dst;
      for (unsigned cm = child_mask[t]; cm; cm &= cm - 1) acc += Kt[GMR_KS * (__ffs(cm) - 1) + e];
      *dst = acc;
    }
    __syncwarp();
  }
  // ---- 2. per dof: F_l, c_l, box (c, lo, hi stay in registers: lane l / l+32 is the lane that owns
  //         dof l in the QP)
#pragma unroll
  for (int r = 0; r < 2; r++) {
    const int l = lane + 32 * r;
    c_out[r] = 0; lo_out[r] = -INFINITY; hi_out[r] = INFINITY;
    if (l >= n) continue;
    unsigned mask = mv(M.dof_topmask)[stage * n + l];
    double Kc[21], gc[6];
#pragma unroll
    for (int i = 0; i < 21; i++) Kc[i] = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) gc[i] = 0;
    while (mask) {
      const int t = __ffs(mask) - 1;
      mask &= mask - 1;
      const double* K = Kt + GMR_KS * t;
      const double* g = K + 21;
#pragma unroll
      for (int i = 0; i < 21; i++) Kc[i] += K[i];
#pragma unroll
      for (int i = 0; i < 6; i++) gc[i] += g[i];
    }
    const double* Sl = sm + lay.S + GMR_SS * l;
    double s[6];
#pragma unroll
    for (int i = 0; i < 6; i++) s[i] = Sl[i];
#pragma unroll
    for (int i = 0; i < 6; i++) {
      double acc = 0;
#pragma unroll
      for (int j = 0; j < 6; j++) acc += Kc[i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i] * s[j];
      F[GMR_SS * l + i] = acc;
    }
    double cl = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) cl += s[i] * gc[i];
    c_out[r] = cl;
    // mink ConfigurationLimit: -gain (q - lo) <= dq <= gain (hi - q) on limited hinge/slide dofs
    double lo = -INFINITY, hi = INFINITY;
    if (mv(M.dof_limited)[l]) {
      const double qv = q[mv(M.dof_qadr)[l]];
      hi = M.limit_gain * (mv(M.dof_hi)[l] - qv);
      lo = -M.limit_gain * (qv - mv(M.dof_lo)[l]);
    }
    lo_out[r] = lo;
    hi_out[r] = hi;
  }
  __syncwarp();
  // ---- 3. every stored entry of H
  const short* ent_row = mv(M.ent_row);
  const short* anc = mv(M.dof_anc);
  for (int e = lane; e < M.nnz; e += 32) {
    const int l = ent_row[e], k = anc[e];
    const double* Sk = sm + lay.S + GMR_SS * k;
    const double* Fl = F + GMR_SS * l;
    double h = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) h += Sk[i] * Fl[i];
    H[e] = h + (k == l ? diag : 0.0);
  }
  __syncwarp();
}

// ------------------------------------------------------------------ Kernel 3: exact box QP
//   min 1/2 x'Hx + c'x   s.t. lo <= x <= hi,   H SPD (>= damping I).
// Block principal pivoting (Judice-Pires) with Murty's single-pivot fallback; each pass is one
// Cholesky of H with the rows/columns of the active set replaced by identity, two triangular
// solves and a KKT check.  The minimiser is unique, so any exact method (daqp in the reference)
// returns the same x up to rounding, whatever active set the search starts from.
//
// Storage: H and L are packed lower triangles in COLUMN-major order, element (i,j), i >= j, at
// colbase(n,j) + i: the lanes of a warp walk down a column, so every shared-memory access of the
// factorisation is either contiguous across lanes or a broadcast (no bank conflicts).  The
// factorisation is left-looking by columns; at step k lane l owns row k + l, so all rows still
// to be computed sit in the low lanes and a second pass is only needed while n - k > 32.
// Uses sm[L] (diagonal holds 1/L_kk), writes sm[x].  `act_lo`/`act_hi` (warp-uniform bit masks of
// the dofs held at their lower/upper bound) seed the search and return the final active set, so
// consecutive solves of one frame start from the previous answer.  Returns #factorisations
// (negative if H was not positive definite).
#define QP_PTOL 1e-10
#define QP_DTOL 1e-9

// symmetric read of a packed column-major lower triangle
__device__ __forceinline__ double sym_at(const double* H, int n, int i, int j) {
  return i >= j ? H[colbase(n, j) + i] : H[colbase(n, i) + j];
}

__device__ __noinline__ int qp_solve(int n, const double* __restrict__ H, const double* __restrict__ c,
                                     const double* __restrict__ lo, const double* __restrict__ hi,
                                     double* __restrict__ L, double* __restrict__ x, int lane,
                                     unsigned long long& act_lo, unsigned long long& act_hi) {
  const int i0 = lane, i1 = lane + 32;
  const bool has0 = i0 < n, has1 = i1 < n, two = n > 32;
  const double lo0 = has0 ? lo[i0] : 0, hi0 = has0 ? hi[i0] : 0, lo1 = has1 ? lo[i1] : 0, hi1 = has1 ? hi[i1] : 0;
  const double c0 = has0 ? c[i0] : 0, c1 = has1 ? c[i1] : 0;
  const int cb0 = colbase(n, i0), cb1 = colbase(n, i1);
  // 0 free, -1 at lower, +1 at upper; a seed is only honoured where the bound is finite
  int st0 = 0, st1 = 0;
  if (has0) st0 = ((act_lo >> i0) & 1ull) && lo0 > -INFINITY ? -1 : (((act_hi >> i0) & 1ull) && hi0 < INFINITY ? 1 : 0);
  if (has1) st1 = ((act_lo >> i1) & 1ull) && lo1 > -INFINITY ? -1 : (((act_hi >> i1) & 1ull) && hi1 < INFINITY ? 1 : 0);
  unsigned long long act = (unsigned long long)__ballot_sync(GMR_FULL, st0 != 0) | ((unsigned long long)__ballot_sync(GMR_FULL, st1 != 0) << 32);
  int best = n + 1, p = 3, nfact = 0;
  bool spd = true;
  double r0 = 0, r1 = 0;
#pragma unroll 1
  for (int it = 0; it < 400; it++) {
    // ---- right-hand side: active rows carry their bound, free rows -c_F - H_FA x_A
    const double xa0 = st0 < 0 ? lo0 : (st0 > 0 ? hi0 : 0.0), xa1 = st1 < 0 ? lo1 : (st1 > 0 ? hi1 : 0.0);
    r0 = st0 ? xa0 : -c0;
    r1 = st1 ? xa1 : -c1;
    if (act) {
      if (has0) x[i0] = xa0;
      if (has1) x[i1] = xa1;
      __syncwarp();
#pragma unroll 1
      for (unsigned long long mk = act; mk; mk &= mk - 1) {
        const int j = __ffsll((long long)mk) - 1;
        const double xj = x[j];
        if (has0 && !st0) r0 -= sym_at(H, n, i0, j) * xj;
        if (has1 && !st1) r1 -= sym_at(H, n, i1, j) * xj;
      }
      __syncwarp();
    }
    // ---- Cholesky, left-looking by columns, identity on the active set
#pragma unroll 1
    for (int k = 0; k < n; k++) {
      const bool ak = (act >> k) & 1ull;
      double* Lk = L + colbase(n, k);
      const int npass = (n - k > 32) ? 2 : 1;
      if (ak) {                                    // warp-uniform
#pragma unroll 1
        for (int ps = 0; ps < npass; ps++) {
          const int i = k + lane + 32 * ps;
          if (i < n) Lk[i] = (i == k) ? 1.0 : 0.0;
        }
      } else {
        double inv = 0;
#pragma unroll 1
        for (int ps = 0; ps < npass; ps++) {
          const int i = k + lane + 32 * ps;
          const bool in = i < n;
          const int ic = in ? i : k;               // clamp: idle lanes recompute the pivot row
          double a = H[colbase(n, k) + ic], b = 0;
          const double* Lj = L;
          int j = 0, step = n - 1;
#pragma unroll 2
          for (; j + 1 < k; j += 2) {
            const double* Lj1 = Lj + step;
            a -= Lj[ic] * Lj[k];
            b -= Lj1[ic] * Lj1[k];
            Lj = Lj1 + step - 1;
            step -= 2;
          }
          if (j < k) a -= Lj[ic] * Lj[k];
          const double s = a + b;
          if (ps == 0) {
            double piv = __shfl_sync(GMR_FULL, s, 0);
            if (!(piv > 0.0)) { spd = false; piv = 1.0; }
            inv = rsqrt(piv);
          }
          if (in) Lk[i] = (i == k) ? inv : (((act >> i) & 1ull) ? 0.0 : s * inv);   // diagonal stores 1/L_kk
        }
      }
      __syncwarp();
    }
    nfact++;
    // ---- forward solve L y = r (column oriented), then L' x = y
#pragma unroll 1
    for (int k = 0; k < n; k++) {
      const double* Lk = L + colbase(n, k);
      const double yk = __shfl_sync(GMR_FULL, (k < 32) ? r0 : r1, k & 31) * Lk[k];
      if (i0 == k) r0 = yk; else if (has0 && i0 > k) r0 -= Lk[i0] * yk;
      if (two) { if (i1 == k) r1 = yk; else if (has1 && i1 > k) r1 -= Lk[i1] * yk; }
    }
#pragma unroll 1
    for (int k = n - 1; k >= 0; k--) {
      const double xk = __shfl_sync(GMR_FULL, (k < 32) ? r0 : r1, k & 31) * L[colbase(n, k) + k];
      if (i0 == k) r0 = xk; else if (i0 < k) r0 -= L[cb0 + k] * xk;
      if (two) { if (i1 == k) r1 = xk; else if (has1 && i1 < k) r1 -= L[cb1 + k] * xk; }
    }
    // exact bounds on the active set (the solve reproduces them up to rounding)
    if (st0) r0 = xa0;
    if (st1) r1 = xa1;
    if (has0) x[i0] = r0;
    if (has1) x[i1] = r1;
    __syncwarp();
    // ---- KKT check
    bool bad0 = false, bad1 = false;
    if (has0 && !st0) bad0 = (r0 < lo0 - QP_PTOL) || (r0 > hi0 + QP_PTOL);
    if (has1 && !st1) bad1 = (r1 < lo1 - QP_PTOL) || (r1 > hi1 + QP_PTOL);
    if (act) {
      if (has0 && st0) {
        double g = c0;
#pragma unroll 1
        for (int j = 0; j < n; j++) g += sym_at(H, n, i0, j) * x[j];
        bad0 = st0 < 0 ? (g < -QP_DTOL) : (g > QP_DTOL);
      }
      if (has1 && st1) {
        double g = c1;
#pragma unroll 1
        for (int j = 0; j < n; j++) g += sym_at(H, n, i1, j) * x[j];
        bad1 = st1 < 0 ? (g < -QP_DTOL) : (g > QP_DTOL);
      }
    }
    const unsigned b0 = __ballot_sync(GMR_FULL, bad0), b1 = __ballot_sync(GMR_FULL, bad1);
    const int nbad = __popc(b0) + __popc(b1);
    if (nbad == 0) break;
    bool all;
    if (nbad < best) { best = nbad; p = 3; all = true; }
    else if (p > 0) { p--; all = true; }
    else all = false;
    const int last = b1 ? 63 - __clz(b1) : 31 - __clz(b0);
    if (bad0 && (all || i0 == last)) st0 = st0 ? 0 : (r0 < lo0 ? -1 : 1);
    if (bad1 && (all || i1 == last)) st1 = st1 ? 0 : (r1 < lo1 ? -1 : 1);
    act = (unsigned long long)__ballot_sync(GMR_FULL, st0 != 0) | ((unsigned long long)__ballot_sync(GMR_FULL, st1 != 0) << 32);
  }
  act_lo = (unsigned long long)__ballot_sync(GMR_FULL, st0 < 0) | ((unsigned long long)__ballot_sync(GMR_FULL, st1 < 0) << 32);
  act_hi = (unsigned long long)__ballot_sync(GMR_FULL, st0 > 0) | ((unsigned long long)__ballot_sync(GMR_FULL, st1 > 0) << 32);
  __syncwarp();
  return spd ? nfact : -nfact;
}

// ------------------------------------------------------------------ Kernel 3: exact box QP on the kinematic tree
// Same problem and same pivoting rules as qp_solve above, but neither H nor c is ever formed.  With v_i the
// "spatial velocity" of dof i's link (v_i = v_parent + S_i x_i), the objective of mink's build_ik is
//   1/2 sum_t v' K_t v + sum_t g_t' v  (v at the link task t hangs on)  +  1/2 d |x|^2,   d = damping + lm * sum mu,
// i.e. the forward dynamics of a tree whose link "inertias" are the task stiffnesses K_t, whose bias forces are g_t
// and whose joints carry an armature d - so the stationary point is what the articulated-body algorithm computes,
// in O(n) 6x6 operations instead of the O(sum depth^2) of a factorisation of H:
//   up   (leaves to root):  X_i = [I_i | p_i] = own tasks + sum_children X^a_c,
//                            U_c = I_c S_c,  D_c = S_c'U_c + d,  alpha_c = 1/D_c,  beta_c = S_c'p_c / D_c,
//                            X^a_c = [I_c - U_c U_c' alpha_c | p_c - U_c beta_c];
//   down (root to leaves):  x_c = -(U_c'v_parent) alpha_c - beta_c,  v_c = v_parent + S_c x_c.
// A dof held at a bound b_c is a locked joint: alpha_c = 0, beta_c = -b_c (then X^a_c = [I_c | p_c + U_c b_c] and
// the same downward formula returns x_c = b_c).  Its multiplier is (Hx + c)_c = S_c' sum_{t below c} f_t + d x_c with
// f_t = K_t v + g_t.  All spatial quantities are expressed about the one reference point O of stage_fk, so nothing is
// transformed between links.
// Schedule (DevModel.aba_*, static, built at model creation per stage): dofs without a task below them are left out
// (x = 0 or their bound); the others are cut into chains (a dof continues into its tallest child).  A pass handles
// one dof on each of the warp's four 8-lane groups; a chain stays on its group, and lane i < 6 of the group keeps
// row i of the chain's X = [I | p] in REGISTERS from one dof to the next: per dof it subtracts the previous dof's
// rank-1 term (U, alpha, beta come back from Ud with four 128-bit loads), adds the dof's own task block (packed) and
// the accumulators of side chains that end here (shared memory, stored by their tops), forms U_i = X_i S and the
// partial sums of D and S'p (three shuffle rounds inside the group); lanes 6 and 7 store alpha and beta.  Only a
// chain's top writes X to shared memory.  The downward sweep runs the same chains top-down with v in registers.
// One __syncwarp per pass.  `act_lo`/`act_hi` seed the search and return the final active set; returns #sweeps of the
// pivoting loop, negative if the search hit its cap (the caller flags the frame; the last iterate is returned).
#define QP_MAX_SWEEPS 100
template <class MV>
__device__ __forceinline__ int qp_solve_aba(const MV& mv, const Lay& lay, double* __restrict__ sm, int stage, double mu_sum,
                                            int lane, unsigned long long* __restrict__ actp) {
  const DevModel& M = mv.M();
  const int n = M.nv, m = M.n_task[stage];
  const double diag = M.damping + M.lm_damping * mu_sum;
  double* XA = sm + lay.K;
  const double* S = sm + lay.S;
  double* Ud = sm + lay.Ud;
  double* v = sm + lay.v;
  double* x = sm + lay.x;
  double* fv = sm + lay.e;                  // task forces of the KKT check (e is dead during the solve)
  const double* q = sm + lay.q;
  const int nup = M.aba_nup[stage], ndn = M.aba_ndn[stage];
  const int g8 = lane >> 3, i8 = lane & 7;
  const int4* const up0 = mv(M.aba_up) + 8 * M.aba_up0[stage] + 2 * g8;
  const int4* const dn0 = mv(M.aba_dn) + 4 * M.aba_dn0[stage] + g8;
  const int2* chl = mv(M.aba_ch);
  // row i8 of a packed symmetric block: byte offsets of K(i8, 0..5), one byte each
  unsigned row_lo = 0, row_hi = 0;
  {
    const int r = i8 < 6 ? i8 : 0;
#pragma unroll
    for (int j = 0; j < 6; j++) {
      const unsigned off = 8u * (unsigned)(r >= j ? r * (r + 1) / 2 + j : j * (j + 1) / 2 + r);
      if (j < 4) row_lo |= off << (8 * j); else row_hi |= off << (8 * (j - 4));
    }
  }
  // everything below addresses shared memory as base + byte offset (the schedule tables hold byte offsets)
  char* const XAc = (char*)XA;
  const char* const Sb = (const char*)S;
  char* const Udb = (char*)Ud;
  char* const vb = (char*)v;
  const int i8b = 8 * i8;
  const int rootX = M.aba_root[stage], rootdof = M.aba_rootdof[stage];
  int ra = 0, rb = 0;                       // lane e of the root block: M(ra, rb) for e < 21, r(ra) for 21 <= e < 27
  if (lane < 21) { ra = (int)((0x5b6db2491b6d2448ull >> (3 * lane)) & 7ull); rb = lane - ra * (ra + 1) / 2; }      // row of packed entry `lane`, 3 bits each
  else if (lane < GMR_KS) ra = lane - 21;
  // lane l owns the box of slots l and l + 32 (slot = index among the dofs that can carry a box: everything but a
  // free joint's six; 29 slots for unitree_g1, so the second slot is skipped by a warp-uniform branch)
  const int nbox = M.n_box;
  const bool two = nbox > 32;
  const bool has0 = lane < nbox, has1 = two && lane + 32 < nbox;
  const int i0 = has0 ? mv(M.box_dof)[lane] : 0, i1 = has1 ? mv(M.box_dof)[lane + 32] : 0;
  // mink ConfigurationLimit: -gain (q - lo) <= dq <= gain (hi - q) on limited hinge/slide dofs, intersected with the
  // optional velocity box |dq| <= vmax dt (off by default: the reference passes no velocity limit, motion_retarget.py:147-149)
  // (the box of slot `lane` is computed once per solve and kept in registers; the second slot of robots with more than 32 slots is re-derived where needed)
  auto box = [&](int i, bool has, double& lo, double& hi) {
    lo = -INFINITY; hi = INFINITY;
    if (!has) return;
    if (mv(M.dof_limited)[i]) {
      const double qv = q[mv(M.dof_qadr)[i]];
      hi = M.limit_gain * (mv(M.dof_hi)[i] - qv); lo = -M.limit_gain * (qv - mv(M.dof_lo)[i]);
    }
    if (M.has_vmax) { const double bnd = mv(M.dof_vmax)[i] * M.timestep; hi = fmin(hi, bnd); lo = fmax(lo, -bnd); }
  };
  // tasks that hang on the same dof are summed into the first of them
  {
    const int2* mg = mv(M.aba_merge) + stage * GMR_MAX_TASKS;
    for (int k = 0; k < M.aba_nmerge[stage]; k++) {
      const int2 pr = mg[k];
      if (lane < GMR_KS) XA[GMR_KS * pr.x + lane] += XA[GMR_KS * pr.y + lane];
    }
  }
  int st0 = 0, st1 = 0;      // 0 free, -1 at lower, +1 at upper; a seed is only honoured where the bound is finite
  double blo0, bhi0;         // the box of slot `lane`, kept for the whole solve (q does not move during it)
  {
    const unsigned long long act_lo = actp[0], act_hi = actp[1];
    double lo, hi;
    box(i0, has0, lo, hi);
    blo0 = lo; bhi0 = hi;
    if (has0) st0 = ((act_lo >> lane) & 1ull) && lo > -INFINITY ? -1 : (((act_hi >> lane) & 1ull) && hi < INFINITY ? 1 : 0);
    if (two) {
      box(i1, has1, lo, hi);
      if (has1) st1 = ((act_lo >> (lane + 32)) & 1ull) && lo > -INFINITY ? -1 : (((act_hi >> (lane + 32)) & 1ull) && hi < INFINITY ? 1 : 0);
    }
  }
  unsigned long long act = (unsigned long long)__ballot_sync(GMR_FULL, st0 != 0);
  if (two) act |= (unsigned long long)__ballot_sync(GMR_FULL, st1 != 0) << 32;
  if (rootX < 0) {                          // no floating-base block: a free joint's dofs (if any) are not swept and stay at 0
    for (int i = lane; i < n; i += 32) x[i] = 0.0;
    __syncwarp();
  }
  int best = n + 1, p = 3, nfact = 0;
  bool restarted = act == 0, capped = false;      // a warm-started search that hits its cap starts over cold once
  double r0 = 0, r1 = 0;
#pragma unroll 1
  for (int it = 0;; it++) {
    // dofs at a bound carry it; free dofs without a task below them stay at 0 (the sweeps skip them)
    {
      if (has0) x[i0] = st0 < 0 ? blo0 : (st0 > 0 ? bhi0 : 0.0);
      if (two) { double lo, hi; box(i1, has1, lo, hi); if (has1) x[i1] = st1 < 0 ? lo : (st1 > 0 ? hi : 0.0); }
    }
    __syncwarp();
    // ---- upward sweep.  A pass has two halves around one __syncwarp: (1) lane i < 6 completes row i of X_c (the chain's
    // registers + own task block + side chains from shared memory) and writes U_i = X_i S and p_i; (2) every lane of
    // the group reads the whole U and p back (six 128-bit loads), forms D = S'U + d, S'p, alpha, beta for itself - no
    // shuffle rounds - and lanes < 6 subtract the dof's rank-1 term at once, so the registers carry the ARTICULATED
    // X^a = [I - U U' alpha | p - U beta] into the next dof of the chain.
    const unsigned actw0 = (unsigned)act, actw1 = (unsigned)(act >> 32);
    {
      double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0, a5 = 0, pi = 0;      // row i8 of the chain's X^a
      double* const pscr = fv + 8 * g8;                                  // the group's p, for the second half
      const int4* e = up0;
#pragma unroll 1
      for (int ps = 0; ps < nup; ps++, e += 8) {
        // {dof (-1: idle), flags, S_c, Ud_c} {own task block (-1: none), side children [y, z), X_c}
        const int4 en = e[0], eb = e[1];
        const bool valid = en.x >= 0;
        double U = 0.0;
        const double2* Sd = (const double2*)(Sb + en.z);       // an idle entry points at dof 0: loaded, not used
        const double2 s01 = Sd[0], s23 = Sd[1], s45 = Sd[2];
        if (valid && i8 < 6) {
          if (!(en.y & 1)) a0 = a1 = a2 = a3 = a4 = a5 = pi = 0.0;      // a chain's bottom
          if (eb.x >= 0) {                                     // own task block, packed symmetric
            const char* own = XAc + eb.x;
            a0 += *(const double*)(own + __byte_perm(row_lo, 0, 0x4440)); a1 += *(const double*)(own + __byte_perm(row_lo, 0, 0x4441));
            a2 += *(const double*)(own + __byte_perm(row_lo, 0, 0x4442)); a3 += *(const double*)(own + __byte_perm(row_lo, 0, 0x4443));
            a4 += *(const double*)(own + __byte_perm(row_hi, 0, 0x4440)); a5 += *(const double*)(own + __byte_perm(row_hi, 0, 0x4441));
            pi += *(const double*)(own + 168 + i8b);
          }
#pragma unroll 1
          for (int ch = eb.y; ch < eb.z; ch++) {               // side chains that end here: their X^a from shared memory
            const char* Xc = XAc + chl[ch].x;
            const double2* Xr = (const double2*)(Xc + 6 * i8b);
            const double2 x01 = Xr[0], x23 = Xr[1], x45 = Xr[2];
            a0 += x01.x; a1 += x01.y; a2 += x23.x; a3 += x23.y; a4 += x45.x; a5 += x45.y;
            pi += *(const double*)(Xc + 288 + i8b);
          }
          if (en.y & 4) {                                      // the floating base's link: X as it is, for the 6x6 block
            double2* Xd = (double2*)(XAc + eb.w + 6 * i8b);
            Xd[0] = make_double2(a0, a1); Xd[1] = make_double2(a2, a3); Xd[2] = make_double2(a4, a5);
            *(double*)(XAc + eb.w + 288 + i8b) = pi;
          }
          U = fma(a4, s45.x, fma(a2, s23.x, a0 * s01.x)) + fma(a5, s45.y, fma(a3, s23.y, a1 * s01.y));
          *(double*)(Udb + en.w + i8b) = U;
          pscr[i8] = pi;
        }
        __syncwarp();
        if (valid) {
          const double2* ud = (const double2*)(Udb + en.w);
          const double2* pp = (const double2*)pscr;
          const double2 u01 = ud[0], u23 = ud[1], u45 = ud[2], p01 = pp[0], p23 = pp[1], p45 = pp[2];
          const double D = fma(u45.x, s45.x, fma(u23.x, s23.x, u01.x * s01.x)) + fma(u45.y, s45.y, fma(u23.y, s23.y, fma(u01.y, s01.y, diag)));
          const double sp = fma(p45.x, s45.x, fma(p23.x, s23.x, p01.x * s01.x)) + fma(p45.y, s45.y, fma(p23.y, s23.y, p01.y * s01.y));
          const int slot = (en.y >> 16) - 1;                       // box slot of the dof (-1: a free joint's dof)
          const bool isact = slot >= 0 && (((slot < 32 ? actw0 : actw1) >> (slot & 31)) & 1u);
          const double inv = rcp_pos(D);
          // a dof held at its bound is a locked joint: alpha = 0, beta = -bound (x[c] holds the bound)
          const double alpha = isact ? 0.0 : inv, beta = isact ? -x[en.x] : sp * inv;
          // every lane of the group holds the same alpha, beta: all eight store them (one word, one value) and all
          // eight run the update (lanes 6, 7 carry U = 0 and rows nobody reads) - branches cost more than the work
          *(double2*)(Udb + en.w + 48) = make_double2(alpha, beta);
          const double ua = -U * alpha;
          a0 = fma(ua, u01.x, a0); a1 = fma(ua, u01.y, a1); a2 = fma(ua, u23.x, a2);
          a3 = fma(ua, u23.y, a3); a4 = fma(ua, u45.x, a4); a5 = fma(ua, u45.y, a5);
          pi = fma(-U, beta, pi);
          if ((en.y & 2) && i8 < 6) {                            // a side chain's top: X^a to shared memory
            double2* Xd = (double2*)(XAc + eb.w + 6 * i8b);
            Xd[0] = make_double2(a0, a1); Xd[1] = make_double2(a2, a3); Xd[2] = make_double2(a4, a5);
            *(double*)(XAc + eb.w + 288 + i8b) = pi;
          }
        }
        __syncwarp();
      }
    }
    // ---- floating base: its six dofs span R^6 and are never limited, so they are solved as one 6x6 block
    // (M = S_b' I S_b + d, M x_b = -S_b' p) instead of six more levels up and six down
    if (rootX >= 0) {
      const char* Xr = XAc + rootX;
      if (i8 < 6) {                                             // Y_b = I S_b, group g: b = g, g + 4 -> Ud of the base dofs
        const double* row = (const double*)(Xr + 6 * i8b);
        const double r0 = row[0], r1 = row[1], r2 = row[2], r3 = row[3], r4 = row[4], r5 = row[5];
        for (int b = g8; b < 6; b += 4) {
          const double* Sd = S + GMR_SS * (rootdof + b);
          Ud[8 * (rootdof + b) + i8] = (r0 * Sd[0] + r2 * Sd[2] + r4 * Sd[4]) + (r1 * Sd[1] + r3 * Sd[3] + r5 * Sd[5]);
        }
      }
      __syncwarp();
      if (lane < GMR_KS) {                                      // M(a,b), a >= b (lane < 21); r_a (lanes 21..26)
        const double* Sa = S + GMR_SS * (rootdof + ra);
        const double* Y = lane < 21 ? Ud + 8 * (rootdof + rb) : (const double*)(Xr + 288);
        double acc = (Sa[0] * Y[0] + Sa[2] * Y[2] + Sa[4] * Y[4]) + (Sa[1] * Y[1] + Sa[3] * Y[3] + Sa[5] * Y[5]);
        if (lane < 21 && ra == rb) acc += diag;
        fv[lane] = acc;
      }
      __syncwarp();
      if (lane < 6) {                                           // lanes 0..5 each solve the block (L D L'); lane i forms v_i
        double Lm[21], dinv[6], y[6];
#pragma unroll
        for (int k = 0; k < 21; k++) Lm[k] = fv[k];
#pragma unroll
        for (int j = 0; j < 6; j++) {
          double tmp[6];
          double dj = Lm[j * (j + 1) / 2 + j];
#pragma unroll
          for (int k = 0; k < j; k++) { tmp[k] = Lm[j * (j + 1) / 2 + k] * Lm[k * (k + 1) / 2 + k]; dj -= Lm[j * (j + 1) / 2 + k] * tmp[k]; }
          Lm[j * (j + 1) / 2 + j] = dj;
          dinv[j] = rcp_pos(dj);
#pragma unroll
          for (int i = j + 1; i < 6; i++) {
            double sacc = Lm[i * (i + 1) / 2 + j];
#pragma unroll
            for (int k = 0; k < j; k++) sacc -= Lm[i * (i + 1) / 2 + k] * tmp[k];
            Lm[i * (i + 1) / 2 + j] = sacc * dinv[j];
          }
        }
#pragma unroll
        for (int i = 0; i < 6; i++) {
          double sacc = -fv[21 + i];
#pragma unroll
          for (int k = 0; k < i; k++) sacc -= Lm[i * (i + 1) / 2 + k] * y[k];
          y[i] = sacc;
        }
#pragma unroll
        for (int i = 0; i < 6; i++) y[i] *= dinv[i];
#pragma unroll
        for (int i = 5; i >= 0; i--) {
          double sacc = y[i];
#pragma unroll
          for (int k = i + 1; k < 6; k++) sacc -= Lm[k * (k + 1) / 2 + i] * y[k];
          y[i] = sacc;
        }
        double vi = 0.0;
#pragma unroll
        for (int b = 0; b < 6; b++) vi += S[GMR_SS * (rootdof + b) + lane] * y[b];
        v[6 * (rootdof + 5) + lane] = vi;                       // velocity of the base link: the parent of everything below
        if (lane == 0) {
#pragma unroll
          for (int b = 0; b < 6; b++) x[rootdof + b] = y[b];
        }
      }
      __syncwarp();
    }
    nfact++;
    // ---- downward sweep: the same chains top-down.  Every lane of a group carries the whole link velocity v (six
    // registers) and forms U'v itself from three 128-bit loads of U: no shuffle rounds, a third of the dependent
    // latency of a per-lane split - the sweep is a chain of depth `ndn`, so latency is what it costs.
    {
      double v0 = 0, v1 = 0, v2 = 0, v3 = 0, v4 = 0, v5 = 0;
      const int4* e = dn0;
#pragma unroll 1
      for (int ps = 0; ps < ndn; ps++, e += 4) {
        const int4 en = *e;                                      // dof (-1: idle), flags, byte offset of v_parent (-1: none)
        if (en.x >= 0) {
          if (!(en.y & 1)) {                                     // a chain's top: v_parent from shared memory (0 at a tree root)
            v0 = v1 = v2 = v3 = v4 = v5 = 0.0;
            if (en.z >= 0) {
              const double2* vp = (const double2*)(vb + en.z);
              const double2 p01 = vp[0], p23 = vp[1], p45 = vp[2];
              v0 = p01.x; v1 = p01.y; v2 = p23.x; v3 = p23.y; v4 = p45.x; v5 = p45.y;
            }
          }
          const double2* ud = (const double2*)(Udb + 64 * en.x);
          const double2* Sd = (const double2*)(Sb + GMR_SS * 8 * en.x);
          const double2 u01 = ud[0], u23 = ud[1], u45 = ud[2], ab = ud[3], s01 = Sd[0], s23 = Sd[1], s45 = Sd[2];
          const double dot = fma(u45.x, v4, fma(u23.x, v2, u01.x * v0)) + fma(u45.y, v5, fma(u23.y, v3, u01.y * v1));
          const double xc = fma(-dot, ab.x, -ab.y);
          v0 = fma(s01.x, xc, v0); v1 = fma(s01.y, xc, v1); v2 = fma(s23.x, xc, v2);
          v3 = fma(s23.y, xc, v3); v4 = fma(s45.x, xc, v4); v5 = fma(s45.y, xc, v5);
          x[en.x] = xc;                                          // all eight lanes hold the same xc and v: same word, same value
          if (en.y & 2) {
            double2* vd = (double2*)(vb + 48 * en.x);
            vd[0] = make_double2(v0, v1); vd[1] = make_double2(v2, v3); vd[2] = make_double2(v4, v5);
          }
        }
        __syncwarp();
      }
    }
    double lo0, hi0, lo1 = -INFINITY, hi1 = INFINITY;
    lo0 = blo0; hi0 = bhi0;
    if (two) box(i1, has1, lo1, hi1);
    // the sweeps reproduce the bounds of the active set up to rounding: x[] still holds them exactly for the dofs
    // the sweeps skip, and the swept ones are overwritten with the exact bound here
    r0 = has0 ? (st0 ? (st0 < 0 ? lo0 : hi0) : x[i0]) : 0.0;
    r1 = (two && has1) ? (st1 ? (st1 < 0 ? lo1 : hi1) : x[i1]) : 0.0;
    // ---- KKT check
    bool bad0 = false, bad1 = false;
    if (has0 && !st0) bad0 = (r0 < lo0 - QP_PTOL) || (r0 > hi0 + QP_PTOL);
    if (two && has1 && !st1) bad1 = (r1 < lo1 - QP_PTOL) || (r1 > hi1 + QP_PTOL);
    if (act) {
      const int* tdof = mv(M.task_dof) + stage * GMR_MAX_TASKS;
      for (int k = lane; k < 6 * m; k += 32) {                   // f_t = K_t v + g_t
        const int t = k / 6, row = k - 6 * t, a = tdof[t];
        const double* Kt = XA + GMR_KS * t;
        double acc = Kt[21 + row];
        if (a >= 0) {
          const int ii = row * (row + 1) / 2;
#pragma unroll
          for (int j = 0; j < 6; j++) acc += Kt[row >= j ? ii + j : j * (j + 1) / 2 + row] * v[6 * a + j];
        }
        fv[GMR_ES * t + row] = acc;
      }
      __syncwarp();
      const unsigned* tmask = mv(M.dof_taskmask) + stage * n;
      const unsigned keep = M.aba_kktmask[stage];
#pragma unroll
      for (int r = 0; r < 2; r++) {
        if (r && !two) break;
        const int i = r ? i1 : i0;
        const int st = r ? st1 : st0;
        if ((r ? has1 : has0) && st) {
          double F[6] = {0, 0, 0, 0, 0, 0};
          for (unsigned mk = tmask[i] & keep; mk; mk &= mk - 1) {
            const double* f = fv + GMR_ES * (__ffs(mk) - 1);
#pragma unroll
            for (int k = 0; k < 6; k++) F[k] += f[k];
          }
          double g = diag * (r ? r1 : r0);
#pragma unroll
          for (int k = 0; k < 6; k++) g += S[GMR_SS * i + k] * F[k];
          const bool bad = st < 0 ? (g < -QP_DTOL) : (g > QP_DTOL);
          if (r) bad1 = bad; else bad0 = bad;
        }
      }
      __syncwarp();
    }
    const unsigned b0 = __ballot_sync(GMR_FULL, bad0), b1 = two ? __ballot_sync(GMR_FULL, bad1) : 0u;
    const int nbad = __popc(b0) + __popc(b1);
    if (nbad == 0) break;
    if (it + 1 >= QP_MAX_SWEEPS) {
      if (restarted) { capped = true; break; }
      restarted = true;                                          // cycling on a stale seed: once more from the empty set
      it = -1; st0 = st1 = 0; act = 0; best = n + 1; p = 3;
      continue;
    }
    bool all;
    if (nbad < best) { best = nbad; p = 3; all = true; }
    else if (p > 0) { p--; all = true; }
    else all = false;
    const int last = b1 ? 63 - __clz(b1) : 31 - __clz(b0);
    if (bad0 && (all || lane == last)) st0 = st0 ? 0 : (r0 < lo0 ? -1 : 1);           // `last` is a slot: slots ascend with the dofs
    if (bad1 && (all || lane + 32 == last)) st1 = st1 ? 0 : (r1 < lo1 ? -1 : 1);
    act = (unsigned long long)__ballot_sync(GMR_FULL, st0 != 0);
    if (two) act |= (unsigned long long)__ballot_sync(GMR_FULL, st1 != 0) << 32;
  }
  if (has0) x[i0] = r0;
  if (two && has1) x[i1] = r1;
  unsigned long long act_lo = (unsigned long long)__ballot_sync(GMR_FULL, st0 < 0);
  unsigned long long act_hi = (unsigned long long)__ballot_sync(GMR_FULL, st0 > 0);
  if (two) {
    act_lo |= (unsigned long long)__ballot_sync(GMR_FULL, st1 < 0) << 32;
    act_hi |= (unsigned long long)__ballot_sync(GMR_FULL, st1 > 0) << 32;
  }
  if (lane == 0) { actp[0] = act_lo; actp[1] = act_hi; }
  __syncwarp();
  return capped ? -nfact : nfact;
}

// H(i,j) of the tree layout (0 if neither dof is an ancestor of the other)
template <class MV>
__device__ __forceinline__ double tree_at(const MV& mv, const double* H, int i, int j) {
  const DevModel& M = mv.M();
  const int di = mv(M.dof_depth)[i], dj = mv(M.dof_depth)[j];
  if (dj <= di) { const int e = mv(M.dof_rowptr)[i] + dj; return mv(M.dof_anc)[e] == j ? H[e] : 0.0; }
  const int e = mv(M.dof_rowptr)[j] + di;
  return mv(M.dof_anc)[e] == i ? H[e] : 0.0;
}

// ------------------------------------------------------------------ Kernel 4: integrate
// solve_ik returns v = dq/dt and integrate_inplace applies q <- q (+) v dt (mj_integratePos); v is formed with the
// host-computed 1/dt (one rounding away from the reference's division, far below the parity tolerances).
template <class MV>
__device__ __forceinline__ void stage_integrate(const MV& mv, const Lay& lay, double* sm, int lane) {
  const DevModel& M = mv.M();
  double* q = sm + lay.q;
  const double* x = sm + lay.x;
  const double dt = M.timestep, idt = M.inv_timestep;
  const short* idof = mv(M.integ_dof);
  for (int j = lane; j < M.n_integ; j += 32) {       // hinges, slides, and one lane per free joint
    const int i = idof[j];
    const int kind = mv(M.dof_kind)[i], qa = mv(M.dof_qadr)[i];
    if (kind == DOF_FREE_R) {
      {                  // first rotational dof of the free joint: its three translations, then the quaternion
#pragma unroll
        for (int k = 1; k <= 3; k++) q[qa - k] += dt * (x[i - k] * idt);
        double w[3] = {x[i] * idt, x[i + 1] * idt, x[i + 2] * idt};
        const double n2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
        double quat[4] = {q[qa], q[qa + 1], q[qa + 2], q[qa + 3]};
        normalize4(quat);
        if (n2 >= 1e-30) {                                   // mju_quatIntegrate: nothing to do for |w| < 1e-15
          const double in = rsqrt_pos(n2), ang = dt * (n2 * in);
          w[0] *= in; w[1] *= in; w[2] *= in;
          double sh, ch;
          sincos(0.5 * ang, &sh, &ch);
          double qr[4] = {ch, sh * w[0], sh * w[1], sh * w[2]}, nq[4];
          qmul(quat, qr, nq);
          quat[0] = nq[0]; quat[1] = nq[1]; quat[2] = nq[2]; quat[3] = nq[3];
        }
        q[qa] = quat[0]; q[qa + 1] = quat[1]; q[qa + 2] = quat[2]; q[qa + 3] = quat[3];
      }
    } else {
      q[qa] += dt * (x[i] * idt);
    }
  }
  __syncwarp();
}

}  // namespace gmr
