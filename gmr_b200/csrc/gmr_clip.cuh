// gmr_clip.cuh - the fused clip kernel (Kernels 1-4 of the north star in one launch) and the
// per-stage kernels that expose the same device code for parity tests.
//
// Work decomposition: one warp = one IK instance (clip x robot) at a time; frames are strictly
// sequential inside a clip (warm start, motion_retarget.py:185), clips are independent.  Warps pull
// work units (32 consecutive frames of one clip, range-major, optionally in a host-given clip order,
// longest first) from a global atomic queue; the warm start crosses units through the batch state.
// A launch is therefore bounded below by its slowest clip (a sequential chain of solves): see
// DESIGN.md section 4 for the measured batch sizes from which the throughput bound takes over.
#pragma once
#include "gmr_stages.cuh"
#include "gmr_layouts.cuh"

namespace gmr {

// per-frame status bits (include/gmr_b200.h: GMR_STATUS_*)
#define GMR_STATUS_MAX_ITER1 1
#define GMR_STATUS_MAX_ITER2 2
#define GMR_STATUS_QP_CAP 4
#define GMR_STATUS_BAD_INPUT 8
#define GMR_STATUS_BAD_OUTPUT 16

// One CTA of up to 16 warps per SM, 128 registers per thread.  (Registers are allocated to a CTA in groups of
// four warps, so the next step up is 20 warps at 96 registers: measured 17 % slower on the bench workload.)
#define GMR_CTA_THREADS 512

struct RunArgs {
  const DevModel* models;        // [n_models]
  const int* model_of;           // [B] or null (all 0)
  const void* human;             // dense [B, T_max, h_max, 7], or CSR [N, h_max, 7] (clip_start != null)
  long long human_frame_stride;  // elements between frames (rows)
  const long long* clip_start;   // CSR layout: [B+1] first row of every clip (null: dense, clip i starts at row i * T_max)
  const int* T_of;               // dense layout: [B] clip lengths or null (= T_max)
  int T_max;                     // dense: frames per clip slot; CSR: the longest clip
  int t_begin, t_end;            // frames [t_begin, min(T, t_end)) of every clip are processed by this launch
  int to_ground;
  const double* scale;           // [B, s_max]
  int s_max;
  double* q_state;               // [B, nq_max] in/out
  unsigned long long* act_state; // [B, GMR_ACT_WORDS] in/out: remembered active sets (first guesses of the next solves)
  int nq_max;
  double* qpos_out;              // [B, T_max, nq_max] or null
  double* scaled_out;            // [B, T_max, s_max, 7] or null
  double* err_out;               // [B, T_max, 2] or null
  int* iters_out;                // [B, T_max, 2] or null
  int* nfact_out;                // [B, T_max] or null
  int* status_out;               // [B, T_max] or null: GMR_STATUS_* bits of every frame
  int* counter;                  // work queue head
  const int* order;              // [B] or null
  int B;
  int unit_frames;               // frames per work unit (a clip is handed out in ranges of this many frames)
  int n_ranges;                  // ceil((t_end - t_begin) / unit_frames)
  int* done;                     // [B] ranges of each clip completed in this launch (zeroed by the host)
  Lay lay;
};

// Per-warp control state of the tick loop (all fields warp-uniform, kept in registers).
struct Slot {
  int inst;                      // instance being retargeted, -1 = none
  int t, T;                      // current frame, clip length
  int stage, it;                 // stage being solved, loop counter of motion_retarget.py:152/171
  int nfact;                     // sweeps of the box QP in this frame (1 per solve_ik + 1 per change of the active set)
  int flags;                     // GMR_STATUS_* bits of the frame
  long long row0;                // row of the clip's frame 0 in the stream / the outputs
  bool new_frame, fresh, need_fk;
  // curr_error of the reference loop and the active set of the instance's previous solve (whatever its frame or stage:
  // a joint at its limit usually stays there; kept in RunArgs::act_state between units and launches) live in the warp's
  // shared-memory block (Lay::st): the solver's sweeps need every register
};

__device__ __forceinline__ int next_stage(const DevModel& M, int stage) {
  stage++;
  while (stage < 2 && !M.use[stage]) stage++;
  return stage;
}

// The fused kernel.  A "tick" is: [FK if q changed] -> control (Kernel 4) -> one solve_ik +
// integrate_inplace.  The control block is what the reference does between two solve_ik calls
// (motion_retarget.py:143-185), as a state machine: evaluate the stage error; first evaluation of a
// stage -> it is curr_error, solve; otherwise compare with curr_error -> solve again, or close the
// stage (next stage / next frame: Kernel 1 on the new frame / next clip from the queue).  Every warp
// of the CTA therefore executes the same solver code in every tick, whatever frame, stage or
// iteration its own instance is in, and each stage function has exactly one call site.
// SM = true: single-model batch, model tables staged in shared memory.
// LAYOUT >= 0: the warp block's offsets are those of fixed_layout(LAYOUT) (gmr_layouts.cuh, generated for every
// constructible (source, robot) pair) and fold into the addresses as immediates; -1: the layout comes in RunArgs.
template <typename TIn, bool SM, int LAYOUT>
__global__ void __launch_bounds__(GMR_CTA_THREADS, 1) clip_kernel(const RunArgs a) {
  extern __shared__ __align__(256) double smem[];
  __shared__ DevModel smodel;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const Lay lay = LAYOUT >= 0 ? fixed_layout(LAYOUT < 0 ? 0 : LAYOUT) : a.lay;
  double* sm = smem + (size_t)warp * lay.total;
  ModelView<SM> mv;
  mv.m = a.models;
  mv.sblob = nullptr;
  if (SM) {   // stage the model: struct + table blob (16-byte words) behind the per-warp blocks
    char* sb = (char*)(smem + (size_t)(blockDim.x >> 5) * lay.total);
    const int4* src = (const int4*)a.models[0].blob;
    int4* dst = (int4*)sb;
    const int nw = (a.models[0].blob_bytes + 15) >> 4;
    for (int i = threadIdx.x; i < nw; i += blockDim.x) dst[i] = src[i];
    const int* ms = (const int*)a.models;
    int* md = (int*)&smodel;
    for (int i = threadIdx.x; i < (int)(sizeof(DevModel) / 4); i += blockDim.x) md[i] = ms[i];
    mv.m = &smodel;
    mv.sblob = sb;
  }
  __syncthreads();
  Slot s;
  s.inst = -1; s.t = s.T = 0; s.stage = 2; s.it = 0; s.nfact = 0; s.flags = 0; s.row0 = 0;
  s.new_frame = s.fresh = s.need_fk = false;
  int range = 0;
  const int n_units = a.n_ranges * a.B;
  for (;;) {
    // ---- a warp without work pulls the next unit from the queue.  A unit is one range of
    // `unit_frames` frames of one clip; units are handed out range-major (range 0 of every clip, then
    // range 1, ...), so the queue balances the SMs at the granularity of a few frames and the tail of
    // the launch is a unit, not a clip.  The warm start crosses units through q_state: the warp that
    // takes range k of a clip waits (normally not at all: range k-1 was handed out B units earlier)
    // until range k-1 is flagged done.  Warps never wait on one another in any other way.
    while (s.inst < 0) {
      int slot = 0;
      if (lane == 0) slot = atomicAdd(a.counter, 1);
      slot = __shfl_sync(GMR_FULL, slot, 0);
      if (slot >= n_units) return;
      range = slot / a.B;
      const int c = slot - range * a.B;
      const int inst = a.order ? a.order[c] : c;
      const long long r0 = a.clip_start ? a.clip_start[inst] : (long long)inst * a.T_max;
      const int T_len = a.clip_start ? (int)(a.clip_start[inst + 1] - r0) : (a.T_of ? min(a.T_of[inst], a.T_max) : a.T_max);
      const int T_inst = min(T_len, a.t_end);
      const int t0 = a.t_begin + range * a.unit_frames;
      if (t0 >= T_inst) continue;                      // the clip ended before this range
      if (!SM) mv.m = a.models + (a.model_of ? a.model_of[inst] : 0);
      if (range > 0) {
        if (lane == 0) { while (*((volatile int*)a.done + inst) < range) { } }
        __syncwarp();
        __threadfence();
      }
      s.inst = inst; s.t = t0; s.T = min(T_inst, t0 + a.unit_frames); s.row0 = r0;
      const double* qs = a.q_state + (size_t)inst * a.nq_max;
      for (int i = lane; i < mv.M().nq; i += 32) sm[lay.q + i] = __ldcg(qs + i);     // written by another SM: bypass L1
      if (lane < GMR_ACT_WORDS) ((unsigned long long*)(sm + lay.st + 1))[lane] = __ldcg(a.act_state + GMR_ACT_WORDS * (size_t)inst + lane);
      __syncwarp();
      s.new_frame = true; s.need_fk = true;
    }
    const DevModel& M = mv.M();
    double* const curr = sm + lay.st;                                             // [0] curr_error
    unsigned long long* const actm = (unsigned long long*)(sm + lay.st + 1);      // remembered active sets
    // ---- Kernel 2: FK (q changed: integrate of the previous tick, or a new clip)
    if (s.need_fk) { stage_fk<false>(mv, lay, sm, lane); s.need_fk = false; }
    // ---- Kernel 4: control
    bool solve = false;
    for (;;) {
      const size_t o = (size_t)(s.row0 + s.t);
      if (s.new_frame) {      // Kernel 1 on frame t
        const TIn* human = (const TIn*)a.human + o * a.human_frame_stride;
        const bool finite = stage_prep<TIn>(mv, lay, sm, human, a.scale + (size_t)s.inst * a.s_max, a.to_ground != 0, lane);
        s.flags = finite ? 0 : GMR_STATUS_BAD_INPUT;
        if (a.scaled_out) {
          double* so = a.scaled_out + o * a.s_max * 7;
          for (int i = lane; i < M.n_scaled * 7; i += 32) so[i] = sm[lay.scaled + i];
        }
        s.new_frame = false;
        s.stage = next_stage(M, -1);
        s.fresh = true; s.it = 0; s.nfact = 0;
        if (lane == 0) {      // stages that are switched off report zeros
          for (int k = 0; k < 2; k++)
            if (!M.use[k]) {
              if (a.err_out) a.err_out[2 * o + k] = 0;
              if (a.iters_out) a.iters_out[2 * o + k] = 0;
            }
        }
      }
      if (s.stage < 2) {
        const double err = stage_errors(mv, lay, sm, s.stage, lane);
        if (s.fresh) { if (lane == 0) *curr = err; s.fresh = false; solve = true; break; }              // :145 / :165
        if (*curr - err > M.converge_eps) {                                                                   // :153 / :172
          if (s.it < M.max_iter) { __syncwarp(); if (lane == 0) *curr = err; s.it++; solve = true; break; }
          s.flags |= GMR_STATUS_MAX_ITER1 << s.stage;        // the loop ended on its counter, not on the error test
        }
        if (lane == 0) {
          if (a.err_out) a.err_out[2 * o + s.stage] = err;
          if (a.iters_out) a.iters_out[2 * o + s.stage] = s.it + 1;
        }
        s.stage = next_stage(M, s.stage);
        s.it = 0;
        if (s.stage < 2) { s.fresh = true; continue; }
      }
      // frame finished (:185)
      if (lane == 0 && a.nfact_out) a.nfact_out[o] = s.nfact;
      {
        bool fin = true;
        double* qo = a.qpos_out ? a.qpos_out + o * a.nq_max : nullptr;
        for (int i = lane; i < M.nq; i += 32) {
          const double qi = sm[lay.q + i];
          fin = fin && isfinite(qi);
          if (qo) qo[i] = qi;
        }
        if (a.status_out) {
          if (!__all_sync(GMR_FULL, fin)) s.flags |= GMR_STATUS_BAD_OUTPUT;
          if (lane == 0) a.status_out[o] = s.flags;
        }
      }
      if (++s.t < s.T) { s.new_frame = true; continue; }
      double* qs = a.q_state + (size_t)s.inst * a.nq_max;
      for (int i = lane; i < M.nq; i += 32) qs[i] = sm[lay.q + i];
      if (lane < GMR_ACT_WORDS) a.act_state[GMR_ACT_WORDS * (size_t)s.inst + lane] = actm[lane];
      __threadfence();
      __syncwarp();
      if (lane == 0) *((volatile int*)a.done + s.inst) = range + 1;      // release the next range of this clip
      s.inst = -1;
      break;
    }
    if (!solve) continue;
    // ---- Kernel 2+3: task stiffness, box QP of mink build_ik (never assembled: articulated-body sweeps over
    //      the task stiffnesses); Kernel 4: integrate
    const double mu = stage_stiffness(mv, lay, sm, s.stage, lane);
    // first guess of the active set: what the same kind of solve ended with last time - the first solve of a frame
    // (the targets just moved: joints run into their limits) or a later one (corrections: they come off again)
    unsigned long long* const actp = actm + 4 * s.stage + 2 * (s.it > 0);
    const int nf = qp_solve_aba(mv, lay, sm, s.stage, mu, lane, actp);
    if (nf < 0) s.flags |= GMR_STATUS_QP_CAP;                // the active-set search hit its cap: last iterate used
    s.nfact += abs(nf);
    stage_integrate(mv, lay, sm, lane);
    s.need_fk = true;
  }
}

// ---------------------------------------------------------------- per-stage kernels (parity tests)
struct StageArgs {
  const DevModel* models;
  const int* model_of;
  int B;
  Lay lay;
  int nq_max, s_max, nbody_max, nv_max, m_max;
  int stage, to_ground;
  const void* human;      // [B,h,7]
  long long human_inst_stride;
  const double* scale;    // [B,s_max]
  const double* q;        // [B,nq_max]
  const double* scaled;   // [B,s_max,7]
  double *scaled_out, *xpos_out, *xquat_out, *e_out, *H_out, *c_out, *lo_out, *hi_out;
  double* dq_out;         // [B,nv_max]
  int* nfact_out;         // [B]
};

template <typename TIn>
__global__ void prep_kernel(const StageArgs a) {
  extern __shared__ __align__(256) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + warp;
  if (inst >= a.B) return;
  double* sm = smem + (size_t)warp * a.lay.total;
  ModelView<false> mv{a.models + (a.model_of ? a.model_of[inst] : 0), nullptr};
  const DevModel& M = mv.M();
  stage_prep<TIn>(mv, a.lay, sm, (const TIn*)a.human + (size_t)inst * a.human_inst_stride,
                  a.scale + (size_t)inst * a.s_max, a.to_ground != 0, lane);
  double* so = a.scaled_out + (size_t)inst * a.s_max * 7;
  for (int i = lane; i < M.n_scaled * 7; i += 32) so[i] = sm[a.lay.scaled + i];
}

__global__ void fk_kernel(const StageArgs a) {
  extern __shared__ __align__(256) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + warp;
  if (inst >= a.B) return;
  double* sm = smem + (size_t)warp * a.lay.total;
  ModelView<false> mv{a.models + (a.model_of ? a.model_of[inst] : 0), nullptr};
  const DevModel& M = mv.M();
  for (int i = lane; i < M.nq; i += 32) sm[a.lay.q + i] = a.q[(size_t)inst * a.nq_max + i];
  __syncwarp();
  stage_fk<true>(mv, a.lay, sm, lane);
  for (int i = lane; i < M.nbody * 3; i += 32) a.xpos_out[(size_t)inst * a.nbody_max * 3 + i] = sm[a.lay.xpos + i];
  for (int i = lane; i < M.nbody * 4; i += 32) a.xquat_out[(size_t)inst * a.nbody_max * 4 + i] = sm[a.lay.xquat + GMR_QS * (i >> 2) + (i & 3)];
}

__global__ void build_qp_kernel(const StageArgs a) {
  extern __shared__ __align__(256) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + warp;
  if (inst >= a.B) return;
  const Lay& lay = a.lay;
  double* sm = smem + (size_t)warp * lay.total;
  ModelView<false> mv{a.models + (a.model_of ? a.model_of[inst] : 0), nullptr};
  const DevModel& M = mv.M();
  for (int i = lane; i < M.nq; i += 32) sm[lay.q + i] = a.q[(size_t)inst * a.nq_max + i];
  for (int i = lane; i < M.n_scaled * 7; i += 32) sm[lay.scaled + i] = a.scaled[(size_t)inst * a.s_max * 7 + i];
  __syncwarp();
  stage_fk<false>(mv, lay, sm, lane);
  stage_errors(mv, lay, sm, a.stage, lane);
  const double mu = stage_stiffness(mv, lay, sm, a.stage, lane);
  const int n = M.nv, m = M.n_task[a.stage];
  if (a.e_out) for (int i = lane; i < 6 * m; i += 32) a.e_out[(size_t)inst * a.m_max * 6 + i] = sm[lay.e + GMR_ES * (i / 6) + i % 6];   // H overwrites e
  __syncwarp();
  double qc[2], qlo[2], qhi[2];
  stage_assemble(mv, lay, sm, a.stage, mu, lane, qc, qlo, qhi);
  double* Ho = a.H_out + (size_t)inst * a.nv_max * a.nv_max;
  for (int idx = lane; idx < n * n; idx += 32) {
    const int i = idx / n, j = idx % n;
    Ho[i * a.nv_max + j] = tree_at(mv, sm + lay.H, i, j);
  }
  for (int r = 0; r < 2; r++) {
    const int i = lane + 32 * r;
    if (i < n) {
      a.c_out[(size_t)inst * a.nv_max + i] = qc[r];
      a.lo_out[(size_t)inst * a.nv_max + i] = qlo[r];
      a.hi_out[(size_t)inst * a.nv_max + i] = qhi[r];
    }
  }
}

// one mink.solve_ik call on the batch state: dq = argmin of the stage's QP (articulated-body solver)
__global__ void solve_ik_kernel(const StageArgs a) {
  extern __shared__ __align__(256) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + warp;
  if (inst >= a.B) return;
  const Lay& lay = a.lay;
  double* sm = smem + (size_t)warp * lay.total;
  ModelView<false> mv{a.models + (a.model_of ? a.model_of[inst] : 0), nullptr};
  const DevModel& M = mv.M();
  for (int i = lane; i < M.nq; i += 32) sm[lay.q + i] = a.q[(size_t)inst * a.nq_max + i];
  for (int i = lane; i < M.n_scaled * 7; i += 32) sm[lay.scaled + i] = a.scaled[(size_t)inst * a.s_max * 7 + i];
  __syncwarp();
  stage_fk<false>(mv, lay, sm, lane);
  stage_errors(mv, lay, sm, a.stage, lane);
  const double mu = stage_stiffness(mv, lay, sm, a.stage, lane);
  unsigned long long* actp = (unsigned long long*)(sm + lay.st + 1);
  if (lane < 2) actp[lane] = 0ull;
  __syncwarp();
  const int nf = qp_solve_aba(mv, lay, sm, a.stage, mu, lane, actp);
  for (int i = lane; i < M.nv; i += 32) a.dq_out[(size_t)inst * a.nv_max + i] = sm[lay.x + i];
  if (lane == 0 && a.nfact_out) a.nfact_out[inst] = nf;
}

// generic box QP: H [n_qp, n, n] full symmetric
__global__ void qp_kernel(int n_qp, int n, const double* H, const double* c, const double* lo, const double* hi,
                          double* x_out, int* nfact_out) {
  extern __shared__ __align__(256) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int inst = blockIdx.x * (blockDim.x >> 5) + warp;
  if (inst >= n_qp) return;
  const int nH = n * (n + 1) / 2;
  double* sm = smem + (size_t)warp * (2 * nH + 4 * n);
  double *sH = sm, *sL = sm + nH, *sc = sL + nH, *slo = sc + n, *shi = slo + n, *sx = shi + n;
  const double* Hi = H + (size_t)inst * n * n;
  for (int idx = lane; idx < n * n; idx += 32) {
    const int i = idx / n, j = idx % n;
    if (j <= i) sH[colbase(n, j) + i] = Hi[idx];
  }
  for (int i = lane; i < n; i += 32) { sc[i] = c[(size_t)inst * n + i]; slo[i] = lo[(size_t)inst * n + i]; shi[i] = hi[(size_t)inst * n + i]; }
  __syncwarp();
  unsigned long long act_lo = 0, act_hi = 0;      // cold start: the oracle's pivoting sequence
  const int nf = qp_solve(n, sH, sc, slo, shi, sL, sx, lane, act_lo, act_hi);
  for (int i = lane; i < n; i += 32) x_out[(size_t)inst * n + i] = sx[i];
  if (lane == 0 && nfact_out) nfact_out[inst] = nf;
}

// ---------------------------------------------------------------- FK tail, float32 (drivers' local_body_pos)
// KinematicsModel.forward_kinematics (kinematics_model.py:213-246): xyzw quaternions, float32,
// body 0 = the root body (given pose), one thread per sample walking the bodies in order.
struct FkTailModel {
  int nbody;                 // bodies excluding world
  const int* parent;         // [nbody] index among these bodies, -1 for the root
  const float* trans;        // [nbody*3]
  const float* rot_xyzw;     // [nbody*4]
  const float* axis;         // [nbody*3] normalised joint axis (0 if none)
  const int* dof;            // [nbody] index into dof_pos or -1
  int ndof;
};

__device__ __forceinline__ void quat_mul_xyzw(const float* a, const float* b, float* o) {
  // same factorisation as torch_utils.quat_mul (torch_utils.py:118-138)
  const float x1 = a[0], y1 = a[1], z1 = a[2], w1 = a[3], x2 = b[0], y2 = b[1], z2 = b[2], w2 = b[3];
  const float ww = (z1 + x1) * (x2 + y2), yy = (w1 - y1) * (w2 + z2), zz = (w1 + y1) * (w2 - z2);
  const float xx = ww + yy + zz;
  const float qq = 0.5f * (xx + (z1 - x1) * (x2 - y2));
  o[3] = qq - ww + (z1 - y1) * (y2 - z2);
  o[0] = qq - xx + (x1 + w1) * (x2 + w2);
  o[1] = qq - yy + (w1 - x1) * (y2 + z2);
  o[2] = qq - zz + (z1 + y1) * (w2 - x2);
}
__device__ __forceinline__ void quat_rotate_xyzw(const float* q, const float* v, float* o) {
  // torch_utils.quat_rotate (torch_utils.py:66-75): a + b + c
  const float w = q[3];
  const float s = 2.0f * w * w - 1.0f;
  const float cx = q[1] * v[2] - q[2] * v[1], cy = q[2] * v[0] - q[0] * v[2], cz = q[0] * v[1] - q[1] * v[0];
  const float d = 2.0f * (q[0] * v[0] + q[1] * v[1] + q[2] * v[2]);
  o[0] = v[0] * s + cx * w * 2.0f + q[0] * d;
  o[1] = v[1] * s + cy * w * 2.0f + q[1] * d;
  o[2] = v[2] * s + cz * w * 2.0f + q[2] * d;
}

__global__ void fk_tail_kernel(FkTailModel m, int N, const float* root_pos, const float* root_rot, const float* dof_pos,
                               float* body_pos, float* body_rot_or_null, float* rot_scratch) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float* bp = body_pos + (size_t)n * m.nbody * 3;
  float* br = (body_rot_or_null ? body_rot_or_null : rot_scratch) + (size_t)n * m.nbody * 4;
  for (int k = 0; k < 3; k++) bp[k] = root_pos[3 * n + k];
  for (int k = 0; k < 4; k++) br[k] = root_rot[4 * n + k];
  for (int j = 1; j < m.nbody; j++) {
    const int p = m.parent[j];
    float jr[4] = {0.f, 0.f, 0.f, 1.f};
    const int d = m.dof[j];
    if (d >= 0) {
      // axis_angle_to_quat (torch_utils.py:354-359): normalize(axis) * sin(theta/2), cos(theta/2), then quat_unit
      const float th = dof_pos[(size_t)n * m.ndof + d] / 2.0f;
      const float s = sinf(th), c = cosf(th);
      jr[0] = m.axis[3 * j] * s; jr[1] = m.axis[3 * j + 1] * s; jr[2] = m.axis[3 * j + 2] * s; jr[3] = c;
      const float nn = fmaxf(sqrtf(jr[0] * jr[0] + jr[1] * jr[1] + jr[2] * jr[2] + jr[3] * jr[3]), 1e-9f);
      jr[0] /= nn; jr[1] /= nn; jr[2] /= nn; jr[3] /= nn;
    }
    float pr[4] = {br[4 * p], br[4 * p + 1], br[4 * p + 2], br[4 * p + 3]}, wt[3], lr[4], cr[4];
    quat_rotate_xyzw(pr, m.trans + 3 * j, wt);
    for (int k = 0; k < 3; k++) bp[3 * j + k] = bp[3 * p + k] + wt[k];
    quat_mul_xyzw(m.rot_xyzw + 4 * j, jr, lr);
    quat_mul_xyzw(pr, lr, cr);
    for (int k = 0; k < 4; k++) br[4 * j + k] = cr[k];
  }
}

}  // namespace gmr
