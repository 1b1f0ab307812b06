/* gmr_b200.h - C ABI of the B200-native batched retargeting engine.
 *
 * Drop-in boundary for the hot path of Zudva/GMR (pure Python reference):
 *   GeneralMotionRetargeting(src_human, tgt_robot).retarget(human_data)
 *       general_motion_retargeting/motion_retarget.py:13-21 (ctor), :139-185 (retarget)
 *   scripts/bvh_to_robot_dataset.py:84-124 and scripts/smplx_to_robot_dataset.py:79-112
 *       (per-file construct -> per-frame retarget -> FK tail)
 * The reference has no FFI of its own (it calls mink / mujoco / daqp from Python); these
 * entry points are what a ctypes binding inside the reference's class would call
 * (INTEGRATION.md shows that binding).
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error (gmr_last_error() has the text);
 *     nothing throws, nothing is printed;
 *   - the caller owns every buffer; the library owns only the opaque handles;
 *   - pointers named *_dev are CUDA device pointers on the handle's device, pointers in
 *     gmr_model_desc are host pointers (copied at create time);
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); all work of a
 *     batch is enqueued on its stream, calls return without synchronising unless stated;
 *   - quaternions are wxyz (scalar first) as in the reference; human frames are
 *     [pos xyz, quat wxyz] = 7 numbers per body;
 *   - a handle is bound to one device and one stream and is not thread-safe (the
 *     reference's object is not either: one object per worker, smplx_to_robot_dataset.py:79).
 */
#ifndef GMR_B200_H
#define GMR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GMR_ABI_VERSION 2

/* hard limits of the kernels (largest shipped robot: nv = 49, nbody = 53, 15 tasks) */
#define GMR_MAX_NV     64
#define GMR_MAX_NBODY  64     /* including world */
#define GMR_MAX_TASKS  32     /* per stage */
#define GMR_MAX_SCALED 32

/* joint types = MuJoCo mjtJoint */
#define GMR_JNT_FREE  0
#define GMR_JNT_SLIDE 2
#define GMR_JNT_HINGE 3

/* dtype of the human stream */
#define GMR_F32 0
#define GMR_F64 1

/* Flattened (source, robot) description: kinematic tree + IK-config tables.
 * Replaces mj.MjModel.from_xml_path (motion_retarget.py:27) and the JSON handling of
 * __init__/setup_retarget_configuration (motion_retarget.py:29-114).  Host pointers. */
typedef struct gmr_model_desc {
  int32_t abi_version;              /* GMR_ABI_VERSION */
  /* kinematic tree, MuJoCo naming; body 0 = world */
  int32_t nbody, njnt, nq, nv;
  const int32_t* body_parentid;     /* [nbody] */
  const double*  body_pos;          /* [nbody,3] */
  const double*  body_quat;         /* [nbody,4] unit wxyz */
  const int32_t* body_jntadr;       /* [nbody] first joint or -1 */
  const int32_t* body_jntnum;       /* [nbody] */
  const int32_t* jnt_type;          /* [njnt] */
  const int32_t* jnt_qposadr;       /* [njnt] */
  const int32_t* jnt_dofadr;        /* [njnt] */
  const double*  jnt_axis;          /* [njnt,3] unit, body frame */
  const double*  jnt_pos;           /* [njnt,3] must be 0 for the CUDA path */
  const int32_t* jnt_limited;       /* [njnt] */
  const double*  jnt_range;         /* [njnt,2] */
  const double*  qpos0;             /* [nq] cold-start state (mink.Configuration(model), :75) */
  double         timestep;          /* model.opt.timestep (:146) */
  /* target preparation (motion_retarget.py:117-136, 209-270) */
  int32_t n_human;                  /* h: bodies per frame in the human stream */
  int32_t n_scaled;                 /* s: bodies of human_scale_table, entry 0 = human root */
  const int32_t* scaled_human_index;/* [s] index into the h bodies of a frame */
  const double*  scale;             /* [s] human_scale_table (ratio not folded in) */
  const double*  pos_off;           /* [s,3] table-1 pos offset - ground_height*z (:91) */
  const double*  rot_off;           /* [s,4] table-1 rot offset wxyz unit (:92) */
  const int32_t* is_foot;           /* [s] name contains "Foot"/"foot" (:261) */
  /* tasks of stage 1 and 2 (ik_match_table1/2, :78-114) */
  int32_t use_table[2];
  int32_t n_task[2];
  const int32_t* task_body[2];      /* [m] robot body id (FrameTask frame_name) */
  const int32_t* task_src[2];       /* [m] index into the scaled list */
  const double*  task_wpos[2];      /* [m] position_cost */
  const double*  task_wrot[2];      /* [m] orientation_cost */
  /* solver constants */
  double  damping;                  /* ctor damping, default 0.5 (:19) */
  int32_t max_iter;                 /* 10 (:56) */
  double  converge_eps;             /* 0.001 (:153,172) */
  double  limit_gain;               /* mink ConfigurationLimit gain 0.95 */
  double  lm_damping;               /* FrameTask lm_damping = 1 (:88,106) */
  /* optional velocity limit (north_star; mink.VelocityLimit): per dof |dq| <= dof_vmax * timestep, intersected with
   * the position box of every solve.  NULL = off = the reference's behaviour (it passes no limits to solve_ik,
   * :147-149, so only mink's default ConfigurationLimit applies).  [nv], +inf = unlimited; the six dofs of a free
   * joint must be +inf. */
  const double*  dof_vmax;
} gmr_model_desc;

/* per-frame status bits (gmr_diag.status_dev).  The reference has no such report: its loops end silently on
 * max_iter (:153,172), and a frame mink/daqp cannot solve raises and the driver skips the file
 * (scripts/bvh_to_robot_dataset.py:75-80 pattern); the batch cannot raise per clip, so it flags. */
#define GMR_STATUS_MAX_ITER1  1   /* stage 1 left its loop on the iteration counter, not on the error test */
#define GMR_STATUS_MAX_ITER2  2   /* same for stage 2 */
#define GMR_STATUS_QP_CAP     4   /* the box QP's active-set search hit its cap (even after a cold restart) */
#define GMR_STATUS_BAD_INPUT  8   /* a non-finite number in the frame's targets */
#define GMR_STATUS_BAD_OUTPUT 16  /* a non-finite number in the frame's qpos (poisons the rest of the clip through the warm start) */

/* per-frame diagnostics (reference: error1()/error2() :188-200, loop counters :152,171); rows are [B,T] for the
 * dense calls and [N] (one per frame of the CSR stream) for the *_csr calls */
typedef struct gmr_diag {
  double*  err_dev;      /* [B,T,2] final error1/error2 of the frame, or NULL */
  int32_t* iters_dev;    /* [B,T,2] QP solves per stage (1..max_iter+1), or NULL */
  int32_t* nfact_dev;    /* [B,T]   sweeps of the box-QP solver in the frame (one per solve_ik + one per change of the active set), or NULL */
  int32_t* status_dev;   /* [B,T]   GMR_STATUS_* bits, or NULL */
} gmr_diag;

typedef struct gmr_model gmr_model;
typedef struct gmr_batch gmr_batch;

const char* gmr_last_error(void);
int gmr_abi_version(void);

/* device = a CUDA device ordinal, or GMR_HOST_ONLY: the flattened tables stay in host memory and no CUDA call is
 * made (such a model cannot back a batch; it exists so that the flattener can be tested without a GPU). */
#define GMR_HOST_ONLY (-1)
int gmr_model_create(const gmr_model_desc* desc, int device, gmr_model** out);
int gmr_model_destroy(gmr_model* m);
/* Static schedule of the fused kernel's box-QP solver for one stage (host-only models): copies table `which`
 * into out[0..cap) and returns the number of int32 it holds.  Byte offsets in the tables are relative to a warp's
 * block array / Ud / v / S arrays as laid out in csrc/gmr_device.cuh (DevModel.aba_*). */
#define GMR_TABLE_INFO       0   /* nv, n_task, n_up_passes, n_down_passes, n_merge, root block offset (-1: none), first base dof,
                                    kkt task mask, doubles of the block array, GMR_KS, GMR_XS, GMR_SS */
#define GMR_TABLE_UP         2   /* per (pass, 8-lane group) 8 ints: dof (-1 idle), flags (1: X^a carried in registers from the
                                    chain's previous dof, 2: store X^a, 4: store X before the dof's rank-1 term, bits 16..23: box
                                    slot + 1), S, Ud offsets, own task block (-1), side children [b, e), X offset */
#define GMR_TABLE_CHILD      3   /* per side child: accumulator offset, Ud offset */
#define GMR_TABLE_DN         5   /* per (pass, group) 4 ints: dof (-1 idle), flags (1: v carried, 2: store v), v_parent offset (-1: none), 0 */
#define GMR_TABLE_MERGE      6   /* pairs (dst task, src task) hanging on one dof */
#define GMR_TABLE_TASK_DOF   7
#define GMR_TABLE_DOF_PARENT 8
#define GMR_TABLE_TASKMASK   9
#define GMR_TABLE_FK_BODY   10   /* bodies the fused kernel's forward kinematics sweeps (the world left out: its pose is a constant) */
int gmr_model_table(const gmr_model* m, int which, int stage, int32_t* out, int cap);
/* Offsets (in doubles) of the per-instance shared-memory block a single-model batch of `m` uses, in the field order of
 * csrc/gmr_device.cuh::Lay; returns the number of ints, copies at most `cap`.  tools/gen_layouts.py turns them into
 * csrc/gmr_layouts.cuh (the fused kernel is compiled once per known layout, offsets as immediates); a batch whose layout
 * is not in that file runs the generic kernel.  Host only, no CUDA call.  No counterpart in the reference. */
int gmr_model_layout(const gmr_model* m, int32_t* out, int cap);

/* B independent IK instances (clip x robot).  n_models > 1 = heterogeneous launch;
 * model_of_instance (host, [B]) may be NULL when n_models == 1.  State starts at qpos0. */
int gmr_batch_create(gmr_model* const* models, int n_models, const int32_t* model_of_instance,
                     int B, void* stream, gmr_batch** out);
int gmr_batch_destroy(gmr_batch* b);
/* q0_dev: [B,nq_max] or NULL for qpos0 (the reference's cold start; it has no reset). */
int gmr_batch_reset(gmr_batch* b, const double* q0_dev);
/* per-instance scale table [B,s_max] (human_scale_table * actual_height/assumption, :36-43);
 * NULL restores the model's table. */
int gmr_batch_set_scale(gmr_batch* b, const double* scale_dev);
/* The state that persists between calls is q plus the box QP's remembered active sets (the first guesses of the next
 * solves, csrc/gmr_clip.cuh).  The sets only decide how many sweeps a solve takes, not its solution, so a clip that is
 * carried over through get_state / reset(q0) (which clears the sets) continues with the same results to rounding. */
int gmr_batch_get_state(gmr_batch* b, double* q_dev /* [B,nq_max] */);
/* processing order of the work queue (device, [B], a permutation; longest clips first balances
 * ragged batches); NULL = instance order. */
int gmr_batch_set_order(gmr_batch* b, const int32_t* order_dev);
/* launch geometry chosen at create time (for reporting) */
int gmr_batch_launch_info(const gmr_batch* b, int32_t* grid, int32_t* warps_per_cta,
                          int32_t* smem_bytes_per_cta, int32_t* smem_bytes_per_instance);
/* index of the compile-time layout (csrc/gmr_layouts.cuh) the batch's fused kernel was compiled for; -1 = the generic
 * kernel with the layout passed at run time (heterogeneous batch, unknown robot, or GMR_DEBUG_RUNTIME_LAYOUT set). */
int gmr_batch_layout_id(const gmr_batch* b);
int gmr_batch_dims(const gmr_batch* b, int32_t* B, int32_t* nq_max, int32_t* s_max,
                   int32_t* h_max, int32_t* nbody_max);

/* retarget(): one frame for every instance, warm-started from the batch state.
 * human_dev [B,h,7]; qpos_out_dev [B,nq_max]; scaled_out_dev [B,s_max,7] or NULL
 * (= .scaled_human_data, :124). */
int gmr_batch_step(gmr_batch* b, const void* human_dev, int human_dtype, int offset_to_ground,
                   double* qpos_out_dev, double* scaled_out_dev, const gmr_diag* diag);

/* whole clips in one launch: instance i retargets frames 0..T_of_instance[i]-1 in order.
 * human_dev [B,T_max,h,7]; T_of_instance_dev [B] or NULL (= T_max for all);
 * qpos_out_dev [B,T_max,nq_max]. */
int gmr_batch_run(gmr_batch* b, const void* human_dev, int human_dtype,
                  const int32_t* T_of_instance_dev, int T_max, int offset_to_ground,
                  double* qpos_out_dev, const gmr_diag* diag);

/* gmr_batch_run on a CSR stream: the clips lie back to back, clip i owns rows [clip_start[i], clip_start[i+1]) of
 * human_dev [N,h,7] and of qpos_out_dev [N,nq_max] (and of the diag arrays); nothing is padded to the longest clip,
 * so a dataset with lengths from 60 to 3000 frames costs its frames, not B x 3000.  clip_start_dev: int64 [B+1]
 * (device); T_max = the longest clip (the caller knows the lengths). */
int gmr_batch_run_csr(gmr_batch* b, const void* human_dev, int human_dtype, const int64_t* clip_start_dev,
                      int T_max, int offset_to_ground, double* qpos_out_dev, const gmr_diag* diag);

/* gmr_batch_run_csr for HOST buffers: human_host [N,h,7], clip_start_host int64 [B+1], qpos_out_host [N,nq_max],
 * status_out_host int32 [N] or NULL.  Zero-copy when all buffers are page-locked (the kernel reads and writes them
 * over the bus), else one H2D copy, the launch, one D2H copy through device staging.  cold_start as below.
 * Asynchronous on the batch's stream. */
int gmr_batch_run_host_csr(gmr_batch* b, const void* human_host, int human_dtype, const int64_t* clip_start_host,
                           int offset_to_ground, int cold_start, double* qpos_out_host, int32_t* status_out_host);

/* one retarget() for every instance straight from / to page-locked HOST buffers (the B = 1 shim's call: one
 * launch, no staging copies): human_host [B,h,7], qpos_out_host [B,nq_max], scaled_out_host [B,s_max,7] or NULL.
 * sync != 0: returns after the batch's stream has finished. */
int gmr_batch_step_host(gmr_batch* b, const void* human_host, int human_dtype, int offset_to_ground,
                        double* qpos_out_host, double* scaled_out_host, int sync);

/* gmr_batch_run for clips held in HOST memory (pinned for full overlap): what the dataset drivers
 * call per folder (scripts/bvh_to_robot_dataset.py:96-113, smplx_to_robot_dataset.py:83-104).
 * human_host [B,T_max,h,7], T_of_host [B] or NULL, qpos_out_host [B,T_max,nq_max].
 * n_chunks >= 1: the frames are cut into n_chunks ranges; H2D of range k+1, retargeting of range k and
 * D2H of range k-1 overlap on three streams through device staging buffers (any host memory).
 * n_chunks == 0: zero-copy - one launch; each warp reads its frame from human_host and writes its qpos row
 * to qpos_out_host over the bus directly (both must be page-locked: cudaHostAlloc / cudaHostRegister;
 * -6 otherwise).  Same result bit for bit either way.
 * cold_start != 0 resets every instance to qpos0 first (a new retargeter object per clip).
 * Asynchronous: synchronise the batch's stream before reading qpos_out_host. */
int gmr_batch_run_host(gmr_batch* b, const void* human_host, int human_dtype, const int32_t* T_of_host,
                       int T_max, int offset_to_ground, int cold_start, int n_chunks,
                       double* qpos_out_host);

/* --- stage entry points (same device code as the fused path; used for per-stage parity) --- */
/* Kernel 1: scale + offset (+ground) -> scaled_out_dev [B,s,7] */
int gmr_target_prep(gmr_batch* b, const void* human_dev, int human_dtype, int offset_to_ground,
                    double* scaled_out_dev);
/* Kernel 2 (FK half) on the batch state or on q_dev if not NULL:
 * xpos_out_dev [B,nbody_max,3], xquat_out_dev [B,nbody_max,4] */
int gmr_fk(gmr_batch* b, const double* q_dev, double* xpos_out_dev, double* xquat_out_dev);
/* Kernel 2+3 (assembly): task errors e [B,m_max,6] and the QP of one solve_ik call:
 * H [B,nv_max,nv_max] (full symmetric), c, lo, hi [B,nv_max]; targets = scaled_dev [B,s,7] */
int gmr_build_qp(gmr_batch* b, int stage, const double* scaled_dev, double* e_out_dev,
                 double* H_out_dev, double* c_out_dev, double* lo_out_dev, double* hi_out_dev);
/* Kernel 2+3: one mink.solve_ik on the batch state (motion_retarget.py:147-149): task stiffness + the
 * tree-structured box QP of the fused path (articulated-body sweeps, H never formed);
 * dq_out_dev [B,nv_max] (= v*dt), nfact [B] (solver sweeps) or NULL */
int gmr_solve_ik(gmr_batch* b, int stage, const double* scaled_dev, double* dq_out_dev,
                 int32_t* nfact_out_dev);
/* Kernel 3 (dense variant): box QP for arbitrary SPD H: x [n_qp,n] ; nfact [n_qp] or NULL */
int gmr_qp_solve(int device, void* stream, int n_qp, int n, const double* H_dev,
                 const double* c_dev, const double* lo_dev, const double* hi_dev,
                 double* x_out_dev, int32_t* nfact_out_dev);

/* FK tail of the dataset drivers (KinematicsModel.forward_kinematics,
 * kinematics_model.py:213-246): float32, xyzw root quaternion, bodies 1..nbody-1.
 * root_pos_dev [N,3], root_rot_xyzw_dev [N,4], dof_pos_dev [N,ndof] ->
 * body_pos_out_dev [N,nbody-1,3], body_rot_out_dev [N,nbody-1,4] (xyzw) or NULL. */
int gmr_fk_bodies_f32(gmr_model* m, void* stream, int N, const float* root_pos_dev,
                      const float* root_rot_xyzw_dev, const float* dof_pos_dev,
                      float* body_pos_out_dev, float* body_rot_out_dev);

/* BVH motion -> human frame stream, for all frames of all clips of a folder at once: what
 * load_lafan1_file does after the text parse (general_motion_retargeting/utils/lafan1.py:8-71 with
 * utils/lafan_vendor/extract.py:162-163 and utils/lafan_vendor/utils.py:40-103,251-269): Euler angles
 * (degrees, in channel order; channel_axis[k] = 0/1/2 for X/Y/Z) -> local quaternions with the
 * reference's sign continuity over time, FK down the bone list (parents precede children), Y-up cm ->
 * Z-up m, orientation = R_quat * q.  Clips are concatenated along the frame axis: clip c owns frames
 * [clip_start[c], clip_start[c+1]).  Output row k takes the position of bone src_pos[k] and the
 * orientation of bone src_rot[k] (LeftFootMod = LeftFoot position + LeftToe orientation, :35-39);
 * stream_out [n_frames, n_rows, 7] = pos xyz, quat wxyz in out_dtype - the layout gmr_batch_run reads.
 * height_out [n_clips] (or NULL): the loader's height guess from each clip's last frame (:42-69) with
 * head_row = output row of "Head" (-1: absent) and foot_rows[4] = rows of LeftFootMod, RightFootMod,
 * LeftFoot, RightFoot (-1: absent).  All pointers except channel_axis / foot_rows are device memory;
 * asynchronous on `stream`. */
int gmr_bvh_to_stream(int device, void* stream, const double* lpos_dev, const double* euler_deg_dev,
                      const int32_t* clip_start_dev, int n_clips, int n_frames, int n_bones,
                      const int32_t* parents_dev, const int32_t* channel_axis, const int32_t* src_pos_dev,
                      const int32_t* src_rot_dev, int n_rows, int out_dtype, void* stream_out_dev,
                      int head_row, const int32_t* foot_rows, double* height_out_dev);

/* gmr_bvh_to_stream for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same kernels, same results; the per-(frame, bone) scratch (global transforms + sign bits, 65 bytes
 * per element) lives in a fixed array in the library's device image instead of a stream-ordered allocation, so the
 * call makes no cudaMalloc / cudaFree / cudaMallocAsync and calls no allocating library routine.
 *   limit: n_frames * n_bones <= GMR_BVH_STATIC_MAX_ELEMS (22-bone LAFAN1 skeleton: 23 831 frames per call);
 *          larger calls return -4 with a message in gmr_last_error() and launch nothing - split them or use
 *          gmr_bvh_to_stream.  The array (32.5 MiB) is held by every device that loads the library, until exit.
 *   sharing: every call on a device uses the same array, which keeps the previous call's contents (the part a call
 *          reads without writing is cleared first).  Calls that overlap on the GPU - different streams, or graph
 *          launches in flight at once - race on it; serialise them.
 *   first call per device: looks the array's device address up (cudaGetSymbolAddress); make one call outside a
 *          capture before capturing. */
#define GMR_BVH_STATIC_MAX_ELEMS (1 << 19)
int gmr_bvh_to_stream_static(int device, void* stream, const double* lpos_dev, const double* euler_deg_dev,
                             const int32_t* clip_start_dev, int n_clips, int n_frames, int n_bones,
                             const int32_t* parents_dev, const int32_t* channel_axis, const int32_t* src_pos_dev,
                             const int32_t* src_rot_dev, int n_rows, int out_dtype, void* stream_out_dev,
                             int head_row, const int32_t* foot_rows, double* height_out_dev);

/* SMPL-X model output -> human frame stream, for all frames of all clips at once: what
 * get_smplx_data_offline_fast does (general_motion_retargeting/utils/smpl.py:109-198) - fps alignment
 * (SLERP of every joint's rotation vector between the two neighbouring source frames :78-107,133-160, linear
 * interpolation of the joint positions :163-168) and the global-orientation chain R_i = R_parent(i) *
 * R(rotvec_i) (:178-196).  full_pose [n_frames_in, n_joints, 3] float32 rotation vectors (joint 0 =
 * global_orient) and joints [n_frames_in, n_joints_in, 3] float32 are what the body model returned (the model
 * forward itself is outside: licensed files), clips concatenated, clip c = source frames
 * [src_start[c], src_start[c+1]).  Output frame n belongs to clip clip_of_out[n] and samples it at source time
 * time[n] (frames, relative to the clip; np.linspace(0, T-1, T // int(src_fps / tgt_fps)) in the reference);
 * aligned[n] = 0 copies frame (int)time[n] as is (the reference's tgt_fps >= src_fps branch).
 * stream_out [n_frames_out, n_joints, 7] = pos xyz, quat wxyz in out_dtype.  Device pointers; asynchronous. */
int gmr_smplx_to_stream(int device, void* stream, const float* full_pose_dev, const float* joints_dev,
                        int n_joints_in, const int32_t* src_start_dev, const int32_t* clip_of_out_dev,
                        const double* time_dev, const int32_t* aligned_dev, int n_frames_in, int n_frames_out,
                        int n_joints, const int32_t* parents_dev, int out_dtype, void* stream_out_dev);

/* Measurement helper (no reference counterpart): measured peak of the CUDA-core FMA pipe in
 * TFLOP/s (fp64 != 0: DFMA, else FFMA) - the compute roofline denominator of bench.py.
 * Synchronous, default stream. */
int gmr_probe_fma_peak(int device, int fp64, double* tflops_out);
/* Test helper: the kernels' short reciprocal, reciprocal root and arctangent (csrc/gmr_device.cuh: hardware seed +
 * two Newton steps; odd polynomial) evaluated on x, y > 0: out[3i] = 1/x, out[3i+1] = 1/sqrt(x),
 * out[3i+2] = atan(min(x,y)/max(x,y)).  Device pointers, default stream. */
int gmr_probe_math(int device, int n, const double* x_dev, const double* y_dev, double* out_dev);

#ifdef __cplusplus
}
#endif
#endif /* GMR_B200_H */
