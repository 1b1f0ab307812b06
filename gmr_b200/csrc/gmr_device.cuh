// gmr_device.cuh - device-side data structures and small math helpers (sm_100a).
//
// One warp owns one IK instance (clip x robot).  All arithmetic is float64: the reference
// path is float64 end to end (numpy / mjtNum) and its convergence rule
// `curr - next > 0.001` (motion_retarget.py:153,172) is a discontinuous decision on an
// error norm, so reduced precision would flip iterations; B200 issues FP64 FMAs at half
// the FP32 rate, which is plenty for a latency/issue-bound kernel.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define GMR_FULL 0xffffffffu
#ifndef GMR_MAX_TASKS
#define GMR_MAX_TASKS 32
#endif

namespace gmr {

// dof kinds (derived from the joint type at flatten time)
enum : int { DOF_FREE_T = 0, DOF_FREE_R = 1, DOF_SLIDE = 2, DOF_HINGE = 3 };

// Flattened (source, robot) constants, immutable, in global memory (read through L1).
// Kinematic tree replaces mj.MjModel (motion_retarget.py:27); task tables replace the
// FrameTask lists (:78-114); prep tables replace pos/rot_offsets1 + human_scale_table.
struct DevModel {
  const char* blob;          // the one device allocation every table pointer below points into
  int blob_bytes;            // leading bytes of the blob that a CTA stages in shared memory (ModelView)
  int nbody, njnt, nq, nv, n_human, n_scaled, nlevel, ref_body;
  int use[2], n_task[2];
  int max_iter;
  double damping, converge_eps, limit_gain, lm_damping, timestep, inv_timestep;
  // bodies
  const int* body_parent;    // [nbody]
  const double* body_pos;    // [nbody*3]
  const double* body_quat;   // [nbody*GMR_QS] wxyz, padded like the pose quaternions (lane = body reads)
  const int* body_jntadr;    // [nbody]
  const int* body_jntnum;    // [nbody]
  int n_fk;                  // bodies the IK needs: those with joints, task frames, their ancestors, the world
  const short* fk_body;      // [n_fk] increasing body ids, entry 0 = world
  int fk_rounds;             // ceil(log2(deepest body))
  const short* fk_jump;      // [fk_rounds*nbody] 2^r-th ancestor of each body, -1 once that is the world
  const int* level_start;    // [nlevel+1] into level_body (level 1 = children of world); kept for tools, no kernel reads it
  const int* level_body;     // [nbody-1] body ids sorted by depth
  // joints
  const int* jnt_type;       // [njnt]
  const int* jnt_qposadr;    // [njnt]
  const int* jnt_dofadr;     // [njnt]
  const double* jnt_axis;    // [njnt*3]
  const double* qpos0;       // [nq]
  // dofs
  const int* dof_kind;       // [nv]
  const int* dof_qadr;       // [nv] qpos index (free rot dofs: index of the quaternion w)
  const int* dof_parent;     // [nv] parent dof in the kinematic tree, -1 at the top
  const int* dof_body;       // [nv] body that owns the dof's joint
  const int* dof_limited;    // [nv]
  const double* dof_lo;      // [nv] joint range lower
  const double* dof_hi;      // [nv]
  int fk_skip;               // leading translational dofs of a floating base (their spatial axes are constants): 3 or 0
  int n_cdof, n_udof;        // FK pass C: scheduled dofs (a task below them) past fk_skip; unscheduled dofs (axes kept at zero)
  const short* cdof;         // [n_cdof]
  const short* udof;         // [n_udof]
  int n_integ;               // dofs stage_integrate visits: hinges, slides and the first rotational dof of a free joint
  const short* integ_dof;    // [n_integ]
  int n_box;                 // dofs that can carry a box constraint (everything but a free joint's six), ascending
  const short* box_dof;      // [n_box]: slot -> dof; the active-set masks of the box QP are indexed by slot
  int has_vmax;              // optional velocity box |dq| <= dof_vmax * timestep (off: the reference passes no velocity limit)
  const double* dof_vmax;    // [nv] (+inf: none)
  const unsigned* dof_taskmask;  // [2*nv] bit t set: task t of the stage is in the dof's subtree
  const unsigned* dof_topmask;   // [2*nv] the top-most of those tasks (no other task of the subtree above them)
  // task tree of each stage (task t's parent = the task on the nearest ancestor body that has one)
  int n_tasklev[2];              // heights 1..n of parent tasks
  const int* tasklev_start;      // [2*(MAXT+1)] into tasklev_task, per height
  const int* tasklev_task;       // [2*MAXT] parent tasks sorted by height (children first)
  const unsigned* task_childmask;// [2*MAXT] bit c set: task c is a child of this task
  // branch-induced sparsity of H = J'WJ + mu I (Featherstone): H(i,k) != 0 only if k is an
  // ancestor-or-self of i (or vice versa).  Row i of the packed lower triangle holds its
  // depth[i]+1 entries in chain order: entry a (0 = topmost ancestor, depth[i] = i itself) is
  // column anc[rowptr[i]+a], stored at rowptr[i]+a.  The ancestor at chain position a has depth a,
  // so H(anc_a, anc_b), b <= a, sits at rowptr[anc_a] + b = ancrow[rowptr[i]+a] + b.
  int nnz, max_depth;
  const int* dof_depth;          // [nv]
  const int* dof_rowptr;         // [nv+1]
  const short* dof_anc;          // [nnz]
  const short* dof_ancrow;       // [nnz] rowptr of that ancestor
  const short* ent_row;          // [nnz] row (dof) of each stored entry
  // static schedule of the articulated-body box-QP solver (qp_solve_aba), per stage.  Byte offsets are relative to
  // the per-warp block array (Lay::K): n_task packed task blocks [K_t | g_t] of GMR_KS doubles, then the dofs'
  // accumulators X = [I (6x6 rows) | p] of GMR_XS doubles
  int aba_nup[2], aba_ndn[2], aba_up0[2], aba_dn0[2], aba_nmerge[2];   // passes of each sweep, first pass of the stage
  int aba_root[2], aba_rootdof[2];   // floating base solved as one block: byte offset of the base link's accumulator (-1: none), first base dof
  unsigned aba_kktmask[2];       // tasks whose block is not merged into another task's (same dof)
  int aba_doubles;               // doubles of the block array a warp needs (max over stages)
  // upward sweep, per (pass, 8-lane group) two words:
  //   {dof (-1: idle), flags, S_dof, Ud_dof}  flags: 1 = X^a carried in registers from the previous dof of the chain,
  //                                                  2 = store X^a (a side chain's top), 4 = store X before the dof's own
  //                                                  rank-1 term (the floating base's link), bits 16..23: box slot + 1
  //   {own task block (-1: none), side children [y,z) of aba_ch, X_dof (where a top stores)}
  const int4* aba_up;
  const int2* aba_ch;            // per side child: its accumulator (X^a, stored by its top), its Ud
  // downward sweep, per (pass, group): {dof (-1: idle), flags, v_parent (-1: none), -}
  //   flags: 1 = v carried in registers from the parent (previous dof of the chain), 2 = store v
  const int4* aba_dn;
  const int2* aba_merge;         // [2*GMR_MAX_TASKS] (dst task, src task) hanging on the same dof
  const int* task_dof;           // [2*GMR_MAX_TASKS] dof the task's body hangs on (-1: none, the task moves nothing)
  // target preparation
  const int* scaled_index;   // [s] body index inside a human frame
  const double* scale;       // [s]
  const double* pos_off;     // [s*3]
  const double* rot_off;     // [s*4]
  const int* is_foot;        // [s]
  // tasks, stage-major: [2*MAXT]
  const int* task_body;
  const int* task_src;
  const double* task_wpos;
  const double* task_wrot;
};

// Table access of the kernels.  SM = true: the CTA staged the model blob into shared memory at
// `sblob` (single-model batches); a table pointer is rebased into that copy, and because `sblob`
// derives from the kernel's __shared__ array the compiler emits LDS with 32-bit addresses.
// SM = false (heterogeneous launches, stage kernels): the tables are read from global memory.
template <bool SM>
struct ModelView {
  const DevModel* m;
  const char* sblob;
  __device__ __forceinline__ const DevModel& M() const { return *m; }
  template <typename T>
  __device__ __forceinline__ const T* operator()(const T* g) const {
    if (SM) return (const T*)(sblob + ((const char*)g - m->blob));
    return g;
  }
};

// Body quaternions are stored with a stride of 5 doubles: lane = body accesses with a stride of 4 doubles
// (32 bytes) hit every bank four times over, 40 bytes spread a half-warp over all 32 banks.
#define GMR_QS 5
// Spatial axes S_k and the assembly's F_l are 6-vectors stored with a stride of 6 doubles = 48 bytes: 16-byte
// aligned, so the solver's sweeps fetch an axis with three 128-bit loads (the lanes of an 8-lane group read the same
// axis: a broadcast); lane = dof accesses (FK pass C) pay a 2-way bank conflict for it.
#define GMR_SS 6
// Task error e_t is stored with a stride of 8: 6 components, then sin and cos of its rotation angle.
#define GMR_ES 8
// Task stiffness block: K_t (21, packed lower 6x6) followed by g_t (6).
#define GMR_KS 27
// Accumulator of the articulated-body solver: I as 6 full rows, then p.
#define GMR_XS 42

// Per-warp shared-memory layout, offsets in doubles (sized for the largest model of a batch).
#define GMR_ACT_WORDS 8
struct Lay {
  int q, scaled, x;                       // per-frame state: configuration, targets, the last solve's dq
  int st;                                 // control state that would otherwise sit in registers: curr_error, then GMR_ACT_WORDS words of
                                          // remembered active sets: [stage][first solve of a frame / later solves][lower, upper]
  int H;                                  // gmr_build_qp only: nnz packed tree rows; aliases e/K (written once they are folded into F)
  int xpos, xquat, S;                     // kinematics: body poses, spatial axes (stride GMR_SS)
  int e, K;                               // task errors (stride GMR_ES), task blocks [K_t | g_t] (stride GMR_KS) + solver accumulators
  int tl, ql;                             // FK local transforms (alias K: dead before the stiffness pass)
  int F;                                  // F_l = Kc_l S_l of the assembly (aliases xpos/xquat: dead after the stiffness pass)
  int Ud, v;                              // qp_solve_aba: per dof [U (6), alpha, beta]; link velocities (6 per dof, alias xpos/xquat)
  int total;                              // doubles per warp
  int maxt;                               // stride of the stage-major task tables
};

// packed lower triangle, COLUMN-major: element (i,j), i >= j, lives at colbase(n,j) + i
__device__ __forceinline__ int colbase(int n, int j) { return j * n - ((j * (j + 1)) >> 1); }
__device__ __forceinline__ int tri(int i) { return i * (i + 1) / 2; }

// ---------------------------------------------------------------- short reciprocal / root / arctangent
// The library's double-precision 1/x, sqrt and atan2 carry slow paths for denormals, infinities and signed zeros
// (20-150 instructions each, inlined at every call site of a kernel whose instruction footprint is what stalls it).
// The arguments here are plain normal numbers, so the hardware's approximate seed (MUFU.RCP64H / RSQ64H, ~20 bits)
// plus two Newton steps (quadratic: 2^-20 -> 2^-40 -> 2^-80) gives the same result to the last bit or two in 5-7
// instructions.
__device__ __forceinline__ double rcp_pos(double D) {            // 1 / D, D > 0
  double inv;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(inv) : "d"(D));        // MUFU.RCP64H: ~20 bits
  inv = fma(inv, fma(-D, inv, 1.0), inv);
  return fma(inv, fma(-D, inv, 1.0), inv);
}
__device__ __forceinline__ double rsqrt_pos(double x) {          // 1 / sqrt(x), x > 0
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));        // MUFU.RSQ64H
  const double hx = 0.5 * x;
  y = fma(y, fma(-hx * y, y, 0.5), y);
  return fma(y, fma(-hx * y, y, 0.5), y);
}
// atan(lo / hi) for 0 <= lo <= hi, hi > 0: odd polynomial in t = lo / hi on [0, 1] (interpolant at 20 Chebyshev
// nodes of atan(sqrt(s)) / sqrt(s), s = t^2; error 2.2e-16 = one ulp at pi/4)
__device__ __forceinline__ double atan_ratio(double lo, double hi) {
  const double t = lo * rcp_pos(hi), s = t * t;
  double p = -1.93423475928923e-05;
  p = fma(p, s, 0.00021423810738603946); p = fma(p, s, -0.0011252544302234645); p = fma(p, s, 0.003751138483965141);
  p = fma(p, s, -0.00899108054265826); p = fma(p, s, 0.01671959606350739); p = fma(p, s, -0.02556862364437174);
  p = fma(p, s, 0.03387126702700675); p = fma(p, s, -0.040811247503178855); p = fma(p, s, 0.04668745304848529);
  p = fma(p, s, -0.052374234719188166); p = fma(p, s, 0.058768281144872724); p = fma(p, s, -0.06665764910689723);
  p = fma(p, s, 0.07692198997458294); p = fma(p, s, -0.09090899793217341); p = fma(p, s, 0.11111110578002083);
  p = fma(p, s, -0.14285714266926733); p = fma(p, s, 0.1999999999964796); p = fma(p, s, -0.333333333333307);
  p = fma(p, s, 1.0);
  return p * t;
}

// ---------------------------------------------------------------- quaternion / vector helpers (wxyz)
__device__ __forceinline__ void qmul(const double* a, const double* b, double* o) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  o[0] = w; o[1] = x; o[2] = y; o[3] = z;
}
// v' = R(q) v
__device__ __forceinline__ void qrot(const double* q, const double* v, double* o) {
  double tx = 2.0 * (q[2] * v[2] - q[3] * v[1]);
  double ty = 2.0 * (q[3] * v[0] - q[1] * v[2]);
  double tz = 2.0 * (q[1] * v[1] - q[2] * v[0]);
  double ox = v[0] + q[0] * tx + (q[2] * tz - q[3] * ty);
  double oy = v[1] + q[0] * ty + (q[3] * tx - q[1] * tz);
  double oz = v[2] + q[0] * tz + (q[1] * ty - q[2] * tx);
  o[0] = ox; o[1] = oy; o[2] = oz;
}
// v' = R(q)^T v
__device__ __forceinline__ void qrot_inv(const double* q, const double* v, double* o) {
  double c[4] = {q[0], -q[1], -q[2], -q[3]};
  qrot(c, v, o);
}
__device__ __forceinline__ void q2mat(const double* q, double* m) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = 1 - 2 * (y * y + z * z); m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = 1 - 2 * (x * x + z * z); m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = 1 - 2 * (x * x + y * y);
}
__device__ __forceinline__ void normalize4(double* q) {   // mju_normalize4
  double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  if (n2 < 1e-30) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  double inv = rsqrt_pos(n2);
  q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv;
}
__device__ __forceinline__ void cross3(const double* a, const double* b, double* o) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
// C = A B (3x3 row-major); C may alias neither input
__device__ __forceinline__ void mm3(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
// C = A^T B
__device__ __forceinline__ void mtm3(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) C[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
}
__device__ __forceinline__ void skew3(const double* w, double* S) {
  S[0] = 0; S[1] = -w[2]; S[2] = w[1]; S[3] = w[2]; S[4] = 0; S[5] = -w[0]; S[6] = -w[1]; S[7] = w[0]; S[8] = 0;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(GMR_FULL, v, o);
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(GMR_FULL, v, o));
  return v;
}

}  // namespace gmr
