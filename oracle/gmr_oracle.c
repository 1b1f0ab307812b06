/* TEST INFRASTRUCTURE (oracle) - not part of the product path.
 *
 * Plain-C float64 restatement of the reference hot path
 *   GeneralMotionRetargeting.retarget()   general_motion_retargeting/motion_retarget.py:139-185
 * with the arithmetic the reference delegates to third-party packages restated from
 * their published algorithms:
 *   mujoco (unpinned, setup.py:17): mj_kinematics, mj_jacBody, mj_integratePos
 *   mink   (unpinned, setup.py:16): FrameTask.compute_error/compute_jacobian, build_ik,
 *                                   ConfigurationLimit, SE3/SO3 log and jlog
 *   daqp   (not listed; default solver, motion_retarget.py:18): exact dense QP
 * PARITY UNPINNED for the IK part: none of these is installable here and the reference
 * holds no golden vector for this path.  This file is cross-checked against the
 * independent numpy restatement (oracle/gmr_oracle_np.py), finite differences and
 * scipy BVLS in tests/.
 *
 * It is deliberately the "mink way": explicit dense 6 x nv task Jacobians, dense
 * H = sum (WJ)'(WJ), and a reduced-system active-set QP - a different route from the
 * CUDA kernels, so that agreement means something.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include "../include/gmr_b200.h"

#define MAXV GMR_MAX_NV
#define MAXB GMR_MAX_NBODY
#define MAXT GMR_MAX_TASKS
#define MAXS GMR_MAX_SCALED
#define LIE_EPS 1e-10

typedef struct {
  double xpos[MAXB][3], xquat[MAXB][4], xmat[MAXB][9];
  double dof_axis[MAXV][3], dof_anchor[MAXV][3];
} kin_t;

/* ---------- small helpers ---------- */
static void qmul(const double* a, const double* b, double* o) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  o[0] = w; o[1] = x; o[2] = y; o[3] = z;
}
static void q2mat(const double* q, double* m) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = 1 - 2 * (y * y + z * z); m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = 1 - 2 * (x * x + z * z); m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = 1 - 2 * (x * x + y * y);
}
static void mv3(const double* m, const double* v, double* o) {
  double a = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double b = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double c = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  o[0] = a; o[1] = b; o[2] = c;
}
static void mtv3(const double* m, const double* v, double* o) { /* m^T v */
  double a = m[0] * v[0] + m[3] * v[1] + m[6] * v[2];
  double b = m[1] * v[0] + m[4] * v[1] + m[7] * v[2];
  double c = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  o[0] = a; o[1] = b; o[2] = c;
}
static void cross3(const double* a, const double* b, double* o) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
static void normalize4(double* q) { /* mju_normalize4 */
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < 1e-15) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}
static void skew(const double* w, double* S) {
  S[0] = 0; S[1] = -w[2]; S[2] = w[1]; S[3] = w[2]; S[4] = 0; S[5] = -w[0]; S[6] = -w[1]; S[7] = w[0]; S[8] = 0;
}
static void mm3(const double* A, const double* B, double* C) {
  double t[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(C, t, sizeof t);
}
static void tr3(const double* A, double* T) {
  double t[9] = {A[0], A[3], A[6], A[1], A[4], A[7], A[2], A[5], A[8]};
  memcpy(T, t, sizeof t);
}

/* ---------- mink SO3/SE3 ---------- */
static void so3_log(const double* q, double* w) {
  double qw = q[0], n2 = q[1] * q[1] + q[2] * q[2] + q[3] * q[3], f;
  if (n2 < LIE_EPS) {
    f = 2.0 / qw - 2.0 / 3.0 * n2 / (qw * qw * qw);
  } else {
    double n = sqrt(n2);
    if (fabs(qw) < LIE_EPS) f = (qw > 0 ? 1.0 : -1.0) * M_PI / n;
    else f = 2.0 * atan2(qw < 0 ? -n : n, fabs(qw)) / n;
  }
  w[0] = f * q[1]; w[1] = f * q[2]; w[2] = f * q[3];
}
static void se3_log(const double* q, const double* p, double* out) {
  double w[3], W[9], W2[9], Vi[9];
  so3_log(q, w);
  double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], k;
  skew(w, W); mm3(W, W, W2);
  if (th2 < LIE_EPS) k = 1.0 / 12.0;
  else { double th = sqrt(th2), h = 0.5 * th; k = (1.0 - 0.5 * th * cos(h) / sin(h)) / th2; }
  for (int i = 0; i < 9; i++) Vi[i] = (i % 4 == 0 ? 1.0 : 0.0) - 0.5 * W[i] + k * W2[i];
  mv3(Vi, p, out);
  out[3] = w[0]; out[4] = w[1]; out[5] = w[2];
}
static void so3_ljacinv(const double* w, double* J) {
  double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]), A, W[9], W2[9];
  if (th < LIE_EPS) { double t2 = th * th; A = (1.0 / 12.0) * (1.0 + t2 / 60.0 * (1.0 + t2 / 42.0 * (1.0 + t2 / 40.0))); }
  else A = (1.0 / (th * th)) * (1.0 - (th * sin(th) / (2.0 * (1.0 - cos(th)))));
  skew(w, W); mm3(W, W, W2);
  for (int i = 0; i < 9; i++) J[i] = (i % 4 == 0 ? 1.0 : 0.0) - 0.5 * W[i] + A * W2[i];
}
static void se3_Q(const double* c, double* Q) {
  const double* w = c + 3;
  double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], A = 0.5, B, C, D;
  if (th2 < LIE_EPS) { B = 1.0 / 6.0 + th2 / 120.0; C = -1.0 / 24.0 + th2 / 720.0; D = -1.0 / 60.0; }
  else {
    double th = sqrt(th2), s = sin(th), co = cos(th);
    B = (th - s) / (th2 * th); C = (1.0 - th2 / 2.0 - co) / (th2 * th2);
    D = (2.0 * th - 3.0 * s + th * co) / (2.0 * th2 * th2 * th);
  }
  double V[9], W[9], VW[9], WV[9], WVW[9], VWW[9], VWWt[9], t1[9], t2[9];
  skew(c, V); skew(w, W);
  mm3(V, W, VW); tr3(VW, WV); mm3(WV, W, WVW); mm3(VW, W, VWW); tr3(VWW, VWWt);
  mm3(WVW, W, t1); mm3(W, WVW, t2);
  for (int i = 0; i < 9; i++)
    Q[i] = A * V[i] + B * (WV[i] + VW[i] + WVW[i]) - C * (VWW[i] - VWWt[i] - 3.0 * WVW[i]) + D * (t1[i] + t2[i]);
}
/* 6x6 inverse left Jacobian; identity when |w|^2 < eps (as mink) */
static void se3_ljacinv(const double* c, double* J) {
  const double* w = c + 3;
  memset(J, 0, 36 * sizeof(double));
  if (w[0] * w[0] + w[1] * w[1] + w[2] * w[2] < LIE_EPS) { for (int i = 0; i < 6; i++) J[7 * i] = 1; return; }
  double Q[9], Ji[9], T[9];
  se3_Q(c, Q); so3_ljacinv(w, Ji);
  mm3(Ji, Q, T); mm3(T, Ji, T);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      J[6 * i + j] = Ji[3 * i + j]; J[6 * i + 3 + j] = -T[3 * i + j]; J[6 * (i + 3) + 3 + j] = Ji[3 * i + j];
    }
}

/* ---------- MuJoCo restatement ---------- */
void oracle_fk(const gmr_model_desc* m, const double* q, kin_t* k) {
  memset(k, 0, sizeof *k);
  k->xquat[0][0] = 1; k->xmat[0][0] = k->xmat[0][4] = k->xmat[0][8] = 1;
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parentid[b], jadr = m->body_jntadr[b], jnum = m->body_jntnum[b];
    double xpos[3], xquat[4], R[9], t[3];
    if (jnum == 1 && m->jnt_type[jadr] == GMR_JNT_FREE) {
      int qa = m->jnt_qposadr[jadr], da = m->jnt_dofadr[jadr];
      memcpy(xpos, q + qa, 3 * sizeof(double));
      memcpy(xquat, q + qa + 3, 4 * sizeof(double));
      normalize4(xquat);
      q2mat(xquat, R);
      for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) {
          k->dof_axis[da + i][j] = (i == j);
          k->dof_axis[da + 3 + i][j] = R[3 * j + i];   /* column i of R */
        }
        memcpy(k->dof_anchor[da + i], xpos, sizeof xpos);
        memcpy(k->dof_anchor[da + 3 + i], xpos, sizeof xpos);
      }
    } else {
      mv3(k->xmat[p], m->body_pos + 3 * b, t);
      for (int i = 0; i < 3; i++) xpos[i] = k->xpos[p][i] + t[i];
      qmul(k->xquat[p], m->body_quat + 4 * b, xquat);
      for (int j = jadr; j < jadr + jnum; j++) {
        int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
        double axw[3], anc[3], val = q[qa] - m->qpos0[qa];
        q2mat(xquat, R);
        mv3(R, m->jnt_axis + 3 * j, axw);
        mv3(R, m->jnt_pos + 3 * j, t);
        for (int i = 0; i < 3; i++) anc[i] = xpos[i] + t[i];
        memcpy(k->dof_axis[da], axw, sizeof axw);
        memcpy(k->dof_anchor[da], anc, sizeof anc);
        if (m->jnt_type[j] == GMR_JNT_SLIDE) {
          for (int i = 0; i < 3; i++) xpos[i] += axw[i] * val;
        } else { /* hinge */
          double h = 0.5 * val, s = sin(h), ql[4] = {cos(h), s * m->jnt_axis[3 * j], s * m->jnt_axis[3 * j + 1], s * m->jnt_axis[3 * j + 2]};
          qmul(xquat, ql, xquat);
          q2mat(xquat, R);
          mv3(R, m->jnt_pos + 3 * j, t);
          for (int i = 0; i < 3; i++) xpos[i] = anc[i] - t[i];
        }
      }
      normalize4(xquat);
    }
    memcpy(k->xpos[b], xpos, sizeof xpos);
    memcpy(k->xquat[b], xquat, sizeof xquat);
    q2mat(xquat, k->xmat[b]);
  }
}

/* mj_jacBody: world-aligned, at body origin. jacp/jacr are [3][MAXV] */
static void jac_body(const gmr_model_desc* m, const kin_t* k, int body, double jacp[3][MAXV], double jacr[3][MAXV]) {
  memset(jacp, 0, 3 * MAXV * sizeof(double));
  memset(jacr, 0, 3 * MAXV * sizeof(double));
  const double* pt = k->xpos[body];
  for (int b = body; b != 0; b = m->body_parentid[b]) {
    int jadr = m->body_jntadr[b], jnum = m->body_jntnum[b];
    for (int j = jadr; j < jadr + jnum; j++) {
      int da = m->jnt_dofadr[j], t = m->jnt_type[j];
      if (t == GMR_JNT_FREE) {
        for (int i = 0; i < 3; i++) jacp[i][da + i] = 1.0;
        for (int i = 0; i < 3; i++) {
          const double* a = k->dof_axis[da + 3 + i];
          double r[3] = {pt[0] - k->dof_anchor[da + 3 + i][0], pt[1] - k->dof_anchor[da + 3 + i][1], pt[2] - k->dof_anchor[da + 3 + i][2]}, c[3];
          cross3(a, r, c);
          for (int x = 0; x < 3; x++) { jacr[x][da + 3 + i] = a[x]; jacp[x][da + 3 + i] = c[x]; }
        }
      } else if (t == GMR_JNT_SLIDE) {
        for (int x = 0; x < 3; x++) jacp[x][da] = k->dof_axis[da][x];
      } else {
        const double* a = k->dof_axis[da];
        double r[3] = {pt[0] - k->dof_anchor[da][0], pt[1] - k->dof_anchor[da][1], pt[2] - k->dof_anchor[da][2]}, c[3];
        cross3(a, r, c);
        for (int x = 0; x < 3; x++) { jacr[x][da] = a[x]; jacp[x][da] = c[x]; }
      }
    }
  }
}

static void integrate_pos(const gmr_model_desc* m, double* q, const double* v, double dt) {
  for (int j = 0; j < m->njnt; j++) {
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    if (m->jnt_type[j] == GMR_JNT_FREE) {
      for (int i = 0; i < 3; i++) q[qa + i] += dt * v[da + i];
      double w[3] = {v[da + 3], v[da + 4], v[da + 5]}, n = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
      normalize4(q + qa + 3);
      if (n < 1e-15) { w[0] = 1; w[1] = w[2] = 0; n = 0; } else { w[0] /= n; w[1] /= n; w[2] /= n; }
      double ang = dt * n;
      if (ang != 0.0) {
        double s = sin(0.5 * ang), qr[4] = {cos(0.5 * ang), s * w[0], s * w[1], s * w[2]};
        qmul(q + qa + 3, qr, q + qa + 3);
      }
    } else {
      q[qa] += dt * v[da];
    }
  }
}

/* ---------- mink FrameTask ---------- */
/* e = log(T_b^-1 T_t) ; target = (tp, tq) */
static void task_error(const kin_t* k, int body, const double* tq, const double* tp, double* e) {
  const double* bq = k->xquat[body];
  double qi[4] = {bq[0], -bq[1], -bq[2], -bq[3]}, qbt[4], d[3], pbt[3];
  qmul(qi, tq, qbt);
  for (int i = 0; i < 3; i++) d[i] = tp[i] - k->xpos[body][i];
  mtv3(k->xmat[body], d, pbt);
  se3_log(qbt, pbt, e);
}
/* J[6][MAXV] = -jlog(T_t^-1 T_b) * [R^T jacp; R^T jacr] */
static void task_jacobian(const gmr_model_desc* m, const kin_t* k, int body, const double* tq, const double* tp, double J[6][MAXV]) {
  double jacp[3][MAXV], jacr[3][MAXV], jb[6][MAXV];
  jac_body(m, k, body, jacp, jacr);
  const double* R = k->xmat[body];
  for (int c = 0; c < m->nv; c++) {
    double a[3] = {jacp[0][c], jacp[1][c], jacp[2][c]}, b[3] = {jacr[0][c], jacr[1][c], jacr[2][c]}, o[3];
    mtv3(R, a, o); jb[0][c] = o[0]; jb[1][c] = o[1]; jb[2][c] = o[2];
    mtv3(R, b, o); jb[3][c] = o[0]; jb[4][c] = o[1]; jb[5][c] = o[2];
  }
  /* T_tb = T_t^-1 T_b */
  double qti[4] = {tq[0], -tq[1], -tq[2], -tq[3]}, qtb[4], Rt[9], d[3], ptb[3], ltb[6], mlog[6], JL[36];
  qmul(qti, k->xquat[body], qtb);
  q2mat(tq, Rt);
  for (int i = 0; i < 3; i++) d[i] = k->xpos[body][i] - tp[i];
  mtv3(Rt, d, ptb);
  se3_log(qtb, ptb, ltb);
  for (int i = 0; i < 6; i++) mlog[i] = -ltb[i];
  se3_ljacinv(mlog, JL);      /* jlog(T) = rjacinv(log T) = ljacinv(-log T) */
  for (int r = 0; r < 6; r++)
    for (int c = 0; c < m->nv; c++) {
      double s = 0;
      for (int x = 0; x < 6; x++) s += JL[6 * r + x] * jb[x][c];
      J[r][c] = -s;
    }
}

/* ---------- mink build_ik ---------- */
/* H [nv*nv] row-major, c, lo, hi [nv], e_out [m*6] (may be NULL) */
void oracle_build_qp(const gmr_model_desc* m, const double* q, const kin_t* k, int stage,
                     const double* scaled /*[s,7]*/, double* H, double* c, double* lo, double* hi, double* e_out) {
  int nv = m->nv;
  for (int i = 0; i < nv * nv; i++) H[i] = 0;
  for (int i = 0; i < nv; i++) { H[i * nv + i] = m->damping; c[i] = 0; }
  for (int t = 0; t < m->n_task[stage]; t++) {
    int body = m->task_body[stage][t];
    const double* tgt = scaled + 7 * m->task_src[stage][t];
    double J[6][MAXV], e[6], W[6], We[6], mu = 0;
    task_jacobian(m, k, body, tgt + 3, tgt, J);
    task_error(k, body, tgt + 3, tgt, e);
    if (e_out) memcpy(e_out + 6 * t, e, sizeof e);
    for (int r = 0; r < 6; r++) W[r] = r < 3 ? m->task_wpos[stage][t] : m->task_wrot[stage][t];
    for (int r = 0; r < 6; r++) { We[r] = W[r] * (-1.0 * e[r]); mu += We[r] * We[r]; }
    mu *= m->lm_damping;
    for (int r = 0; r < 6; r++)
      for (int a = 0; a < nv; a++) J[r][a] *= W[r];
    for (int a = 0; a < nv; a++) {
      for (int b = 0; b < nv; b++) {
        double s = 0;
        for (int r = 0; r < 6; r++) s += J[r][a] * J[r][b];
        H[a * nv + b] += s;
      }
      H[a * nv + a] += mu;
      double s = 0;
      for (int r = 0; r < 6; r++) s += We[r] * J[r][a];
      c[a] += -s;
    }
  }
  /* ConfigurationLimit */
  for (int i = 0; i < nv; i++) { lo[i] = -INFINITY; hi[i] = INFINITY; }
  for (int j = 0; j < m->njnt; j++) {
    if (m->jnt_type[j] == GMR_JNT_FREE || !m->jnt_limited[j]) continue;
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    hi[da] = m->limit_gain * (m->jnt_range[2 * j + 1] - q[qa]);
    lo[da] = -m->limit_gain * (q[qa] - m->jnt_range[2 * j]);
  }
}

/* ---------- exact box QP (stands in for daqp) ----------
 * block principal pivoting on the reduced (free-variable) system, Murty fallback */
int oracle_box_qp(int n, const double* H, const double* c, const double* lo, const double* hi, double* x, int* nfact) {
  signed char st[MAXV];
  int idx[MAXV];
  double A[MAXV * MAXV], r[MAXV], g[MAXV];
  const double ptol = 1e-10, dtol = 1e-9;
  int best = n + 1, p = 3;
  memset(st, 0, sizeof st);
  for (int it = 0; it < 400; it++) {
    int nf = 0;
    for (int i = 0; i < n; i++) { if (st[i] == 0) idx[nf++] = i; x[i] = st[i] < 0 ? lo[i] : (st[i] > 0 ? hi[i] : 0.0); }
    for (int a = 0; a < nf; a++) {
      int i = idx[a];
      double s = -c[i];
      for (int j = 0; j < n; j++) if (st[j] != 0) s -= H[i * n + j] * x[j];
      r[a] = s;
      for (int b = 0; b < nf; b++) A[a * nf + b] = H[i * n + idx[b]];
    }
    /* Cholesky A = L L^T (lower, in place) */
    for (int j = 0; j < nf; j++) {
      double d = A[j * nf + j];
      for (int k2 = 0; k2 < j; k2++) d -= A[j * nf + k2] * A[j * nf + k2];
      if (d <= 0) return -1;
      d = sqrt(d); A[j * nf + j] = d;
      for (int i = j + 1; i < nf; i++) {
        double s = A[i * nf + j];
        for (int k2 = 0; k2 < j; k2++) s -= A[i * nf + k2] * A[j * nf + k2];
        A[i * nf + j] = s / d;
      }
    }
    for (int i = 0; i < nf; i++) { double s = r[i]; for (int k2 = 0; k2 < i; k2++) s -= A[i * nf + k2] * r[k2]; r[i] = s / A[i * nf + i]; }
    for (int i = nf - 1; i >= 0; i--) { double s = r[i]; for (int k2 = i + 1; k2 < nf; k2++) s -= A[k2 * nf + i] * r[k2]; r[i] = s / A[i * nf + i]; }
    for (int a = 0; a < nf; a++) x[idx[a]] = r[a];
    if (nfact) *nfact = it + 1;
    int nbad = 0, last = -1;
    for (int i = 0; i < n; i++) {
      double s = c[i];
      for (int j = 0; j < n; j++) s += H[i * n + j] * x[j];
      g[i] = s;
      int bad = 0;
      if (st[i] == 0) bad = (x[i] < lo[i] - ptol) || (x[i] > hi[i] + ptol);
      else if (st[i] < 0) bad = g[i] < -dtol;
      else bad = g[i] > dtol;
      if (bad) { nbad++; last = i; }
      idx[i] = bad;  /* reuse as flag */
    }
    if (nbad == 0) return 0;
    int all;
    if (nbad < best) { best = nbad; p = 3; all = 1; }
    else if (p > 0) { p--; all = 1; }
    else all = 0;
    for (int i = 0; i < n; i++) {
      if (!idx[i] || (!all && i != last)) continue;
      if (st[i] == 0) st[i] = x[i] < lo[i] ? -1 : 1; else st[i] = 0;
    }
  }
  return -2;
}

/* ---------- target preparation (motion_retarget.py:209-270) ---------- */
void oracle_prepare(const gmr_model_desc* m, const double* scale, const double* frame /*[h,7]*/, int to_ground, double* out /*[s,7]*/) {
  const double* root = frame + 7 * m->scaled_human_index[0];
  double sroot[3] = {scale[0] * root[0], scale[0] * root[1], scale[0] * root[2]};
  for (int i = 0; i < m->n_scaled; i++) {
    const double* b = frame + 7 * m->scaled_human_index[i];
    double pos[3], qn[4] = {b[3], b[4], b[5], b[6]}, uq[4], R[9], off[3];
    for (int x = 0; x < 3; x++) pos[x] = i == 0 ? sroot[x] : (b[x] - root[x]) * scale[i] + sroot[x];
    double n = sqrt(qn[0] * qn[0] + qn[1] * qn[1] + qn[2] * qn[2] + qn[3] * qn[3]);
    for (int x = 0; x < 4; x++) qn[x] /= n;
    qmul(qn, m->rot_off + 4 * i, uq);
    n = sqrt(uq[0] * uq[0] + uq[1] * uq[1] + uq[2] * uq[2] + uq[3] * uq[3]);
    for (int x = 0; x < 4; x++) uq[x] /= n;
    q2mat(uq, R); mv3(R, m->pos_off + 3 * i, off);
    for (int x = 0; x < 3; x++) out[7 * i + x] = pos[x] + off[x];
    for (int x = 0; x < 4; x++) out[7 * i + 3 + x] = uq[x];
  }
  if (to_ground) {
    double lowest = INFINITY;
    for (int i = 0; i < m->n_scaled; i++) if (m->is_foot[i] && out[7 * i + 2] < lowest) lowest = out[7 * i + 2];
    for (int i = 0; i < m->n_scaled; i++) out[7 * i + 2] = out[7 * i + 2] - lowest + 0.1;
  }
}

static double stage_error(const gmr_model_desc* m, const kin_t* k, int stage, const double* scaled) {
  double s = 0;
  for (int t = 0; t < m->n_task[stage]; t++) {
    const double* tgt = scaled + 7 * m->task_src[stage][t];
    double e[6];
    task_error(k, m->task_body[stage][t], tgt + 3, tgt, e);
    for (int r = 0; r < 6; r++) s += e[r] * e[r];
  }
  return sqrt(s);
}

typedef struct { double q[MAXV + 1]; kin_t kin; } state_t;

static int ik_step(const gmr_model_desc* m, state_t* st, int stage, const double* scaled, int* nfact_acc) {
  double H[MAXV * MAXV], c[MAXV], lo[MAXV], hi[MAXV], dq[MAXV], v[MAXV];
  int nf = 0;
  oracle_build_qp(m, st->q, &st->kin, stage, scaled, H, c, lo, hi, NULL);
  int rc = oracle_box_qp(m->nv, H, c, lo, hi, dq, &nf);
  if (rc) return rc;
  if (nfact_acc) *nfact_acc += nf;
  for (int i = 0; i < m->nv; i++) v[i] = dq[i] / m->timestep;     /* solve_ik returns dq/dt */
  integrate_pos(m, st->q, v, m->timestep);                        /* integrate_inplace(v, dt) */
  oracle_fk(m, st->q, &st->kin);
  return 0;
}

/* one retarget() call; returns 0 or <0 */
static int retarget_frame(const gmr_model_desc* m, state_t* st, const double* scale, const double* frame, int to_ground,
                          double* err2, int* it2, int* nfact) {
  double scaled[MAXS * 7];
  oracle_prepare(m, scale, frame, to_ground, scaled);
  if (nfact) *nfact = 0;
  for (int stage = 0; stage < 2; stage++) {
    if (err2) err2[stage] = 0;
    if (it2) it2[stage] = 0;
    if (!m->use_table[stage]) continue;
    double curr = stage_error(m, &st->kin, stage, scaled);
    int rc = ik_step(m, st, stage, scaled, nfact);
    if (rc) return rc;
    double next = stage_error(m, &st->kin, stage, scaled);
    int it = 0;
    while (curr - next > m->converge_eps && it < m->max_iter) {
      curr = next;
      rc = ik_step(m, st, stage, scaled, nfact);
      if (rc) return rc;
      next = stage_error(m, &st->kin, stage, scaled);
      it++;
    }
    if (err2) err2[stage] = next;
    if (it2) it2[stage] = it + 1;
  }
  return 0;
}

/* ---------- exported batch driver ----------
 * human [B,T_max,h,7] (dtype GMR_F32/GMR_F64), T_of [B] or NULL, scale [B,s] or NULL,
 * q_state [B,nq] in/out or NULL (cold start from qpos0), qpos_out [B,T_max,nq],
 * err_out [B,T_max,2], iters_out [B,T_max,2], nfact_out [B,T_max] (each may be NULL) */
typedef struct {
  const gmr_model_desc* m; int B, T_max; const int32_t* T_of; const void* human; int dtype;
  const double* scale; int to_ground; double* q_state; double* qpos_out; double* err_out;
  int32_t* iters_out; int32_t* nfact_out; int next; int fail;
} job_t;

static void* worker(void* arg) {
  job_t* j = (job_t*)arg;
  const gmr_model_desc* m = j->m;
  state_t* st = (state_t*)malloc(sizeof(state_t));
  double* frame = (double*)malloc(sizeof(double) * 7 * m->n_human);
  for (;;) {
    int b = __atomic_fetch_add(&j->next, 1, __ATOMIC_RELAXED);
    if (b >= j->B) break;
    const double* sc = j->scale ? j->scale + (size_t)b * m->n_scaled : m->scale;
    memcpy(st->q, j->q_state ? j->q_state + (size_t)b * m->nq : m->qpos0, sizeof(double) * m->nq);
    oracle_fk(m, st->q, &st->kin);
    int T = j->T_of ? j->T_of[b] : j->T_max;
    for (int t = 0; t < T; t++) {
      size_t fo = ((size_t)b * j->T_max + t) * m->n_human * 7;
      if (j->dtype == GMR_F32) { const float* h = (const float*)j->human + fo; for (int i = 0; i < 7 * m->n_human; i++) frame[i] = h[i]; }
      else memcpy(frame, (const double*)j->human + fo, sizeof(double) * 7 * m->n_human);
      size_t o = (size_t)b * j->T_max + t;
      int rc = retarget_frame(m, st, sc, frame, j->to_ground, j->err_out ? j->err_out + 2 * o : NULL,
                              j->iters_out ? j->iters_out + 2 * o : NULL, j->nfact_out ? j->nfact_out + o : NULL);
      if (rc) { __atomic_store_n(&j->fail, rc, __ATOMIC_RELAXED); break; }
      memcpy(j->qpos_out + o * m->nq, st->q, sizeof(double) * m->nq);
    }
    if (j->q_state) memcpy(j->q_state + (size_t)b * m->nq, st->q, sizeof(double) * m->nq);
  }
  free(st); free(frame);
  return NULL;
}

int oracle_run_clips(const gmr_model_desc* m, int B, int T_max, const int32_t* T_of, const void* human, int dtype,
                     const double* scale, int to_ground, double* q_state, double* qpos_out,
                     double* err_out, int32_t* iters_out, int32_t* nfact_out, int nthreads) {
  if (m->nv > MAXV || m->nbody > MAXB || m->n_scaled > MAXS || m->n_task[0] > MAXT || m->n_task[1] > MAXT) return -10;
  job_t j = {m, B, T_max, T_of, human, dtype, scale, to_ground, q_state, qpos_out, err_out, iters_out, nfact_out, 0, 0};
  if (nthreads < 1) nthreads = 1;
  if (nthreads > B) nthreads = B > 0 ? B : 1;
  if (nthreads == 1) { worker(&j); return j.fail; }
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
  pthread_attr_t attr;
  pthread_attr_init(&attr);
  pthread_attr_setstacksize(&attr, 8u << 20);     /* QP scratch lives on the stack */
  for (int i = 0; i < nthreads; i++) pthread_create(&th[i], &attr, worker, &j);
  for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
  pthread_attr_destroy(&attr);
  free(th);
  return j.fail;
}

/* ---------- exported stage functions (per-stage parity) ---------- */
int oracle_fk_export(const gmr_model_desc* m, const double* q, double* xpos /*[nbody,3]*/, double* xquat /*[nbody,4]*/) {
  kin_t* k = (kin_t*)malloc(sizeof(kin_t));
  oracle_fk(m, q, k);
  for (int b = 0; b < m->nbody; b++) { memcpy(xpos + 3 * b, k->xpos[b], 24); memcpy(xquat + 4 * b, k->xquat[b], 32); }
  free(k);
  return 0;
}
int oracle_build_qp_export(const gmr_model_desc* m, const double* q, int stage, const double* scaled,
                           double* H, double* c, double* lo, double* hi, double* e_out) {
  kin_t* k = (kin_t*)malloc(sizeof(kin_t));
  oracle_fk(m, q, k);
  oracle_build_qp(m, q, k, stage, scaled, H, c, lo, hi, e_out);
  free(k);
  return 0;
}
int oracle_task_jacobian_export(const gmr_model_desc* m, const double* q, int body, const double* tgt /*pos3 quat4*/, double* J /*[6,nv]*/, double* e) {
  kin_t* k = (kin_t*)malloc(sizeof(kin_t));
  double Jb[6][MAXV];
  oracle_fk(m, q, k);
  task_jacobian(m, k, body, tgt + 3, tgt, Jb);
  task_error(k, body, tgt + 3, tgt, e);
  for (int r = 0; r < 6; r++) memcpy(J + r * m->nv, Jb[r], sizeof(double) * m->nv);
  free(k);
  return 0;
}
int oracle_integrate_export(const gmr_model_desc* m, double* q, const double* dq) {
  double v[MAXV];
  for (int i = 0; i < m->nv; i++) v[i] = dq[i] / m->timestep;
  integrate_pos(m, q, v, m->timestep);
  return 0;
}
