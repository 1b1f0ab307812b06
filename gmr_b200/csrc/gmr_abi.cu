// gmr_abi.cu - C ABI of include/gmr_b200.h: model flattening to device tables, batch state,
// kernel launches.  No torch types, no exceptions across the boundary.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <string>
#include <vector>
#include <algorithm>
#include <map>
#include <mutex>

#include "../../include/gmr_b200.h"
#include "gmr_clip.cuh"
#include "gmr_bvh.cuh"
#include "gmr_smplx.cuh"

using namespace gmr;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(-100, std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)

struct gmr_model {
  int device = 0;
  DevModel host{};           // device pointers inside
  char* blob = nullptr;      // one device allocation holding every table (device < 0: = host_blob.data())
  std::vector<char> host_blob;   // host-only model (device = GMR_HOST_ONLY): the tables, for gmr_model_table
  size_t blob_bytes = 0;
  size_t staged_bytes = 0;   // leading part of the blob that the fused kernel stages in shared memory
  int max_task = 0;
  // FK tail tables
  FkTailModel tail{};
  char* tail_blob = nullptr;
};

struct gmr_batch {
  int device = 0;
  cudaStream_t stream = nullptr;
  int B = 0, n_models = 0;
  std::vector<gmr_model*> models;
  DevModel* d_models = nullptr;
  int* d_model_of = nullptr;
  double* d_q = nullptr;        // [B, nq_max]
  double* d_scale = nullptr;    // [B, s_max]
  int* d_counter = nullptr;
  int* d_done = nullptr;        // [B] per-clip range flags of the current launch
  unsigned long long* d_act = nullptr;   // [B,GMR_ACT_WORDS] remembered active sets of each instance (warm starts of its next solves)
  int* d_order = nullptr;
  int nq_max = 0, s_max = 0, h_max = 0, nbody_max = 0, nv_max = 0, m_max = 0, nnz_max = 0, blocks_max = 1;
  Lay lay{};
  int warps_per_cta = 4, grid = 0;
  size_t smem_bytes = 0;          // per-warp blocks only (stage kernels)
  size_t clip_smem_bytes = 0;     // + the staged model of the fused kernel
  // host-buffer pipeline (gmr_batch_run_host)
  void* stage_in = nullptr; size_t stage_in_bytes = 0;
  void* stage_out = nullptr; size_t stage_out_bytes = 0;
  int* stage_T = nullptr;
  int64_t* stage_cs = nullptr;   // clip_start of the CSR host call
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  bool smem_model = false;
  int layout_id = -1;            // index into gmr_layouts.cuh when the batch's layout is one of the compiled ones
};

extern "C" const char* gmr_last_error(void) { return g_err.c_str(); }
extern "C" int gmr_abi_version(void) { return GMR_ABI_VERSION; }

namespace {
struct Blob {
  std::vector<char> bytes;
  template <typename T> size_t push(const std::vector<T>& v) {
    size_t off = (bytes.size() + 15) & ~size_t(15);
    bytes.resize(off + std::max<size_t>(v.size(), 1) * sizeof(T));
    if (!v.empty()) memcpy(bytes.data() + off, v.data(), v.size() * sizeof(T));
    return off;
  }
};
template <typename T> std::vector<T> vec(const T* p, size_t n) { return std::vector<T>(p, p + n); }
}  // namespace

extern "C" int gmr_model_create(const gmr_model_desc* d, int device, gmr_model** out) {
  if (!d || !out) return fail(-1, "null argument");
  if (d->abi_version != GMR_ABI_VERSION) return fail(-2, "abi_version mismatch");
  if (d->nbody < 2 || d->nbody > GMR_MAX_NBODY || d->nv < 1 || d->nv > GMR_MAX_NV || d->nq > GMR_MAX_NV + 1)
    return fail(-3, "model exceeds kernel limits (nbody<=64, nv<=64)");
  if (d->n_scaled < 1 || d->n_scaled > GMR_MAX_SCALED) return fail(-3, "n_scaled out of range (1..32)");
  for (int k = 0; k < 2; k++)
    if (d->n_task[k] < 0 || d->n_task[k] > GMR_MAX_TASKS) return fail(-3, "n_task out of range (0..32)");
  if (!(d->timestep > 0)) return fail(-3, "timestep must be > 0");
  const int nb = d->nbody, nj = d->njnt, nv = d->nv;

  // ---- validate the tree and derive dof tables
  std::vector<int> depth(nb, 0), dof_kind(nv, 0), dof_qadr(nv, -1), dof_parent(nv, -1), dof_limited(nv, 0), dof_body(nv, 0);
  std::vector<double> dof_lo(nv, 0.0), dof_hi(nv, 0.0);
  std::vector<int> body_lastdof(nb, -1);   // last dof of the nearest ancestor-or-self with dofs
  int ref_body = 0, nlevel = 0;
  for (int b = 1; b < nb; b++) {
    int p = d->body_parentid[b];
    if (p < 0 || p >= b) return fail(-4, "bodies must be ordered parent-before-child");
    depth[b] = depth[p] + 1;
    nlevel = std::max(nlevel, depth[b]);
    int last = body_lastdof[p];
    int jadr = d->body_jntadr[b], jnum = d->body_jntnum[b];
    bool seen_hinge = false;
    for (int j = jadr; j < jadr + jnum; j++) {
      if (j < 0 || j >= nj) return fail(-4, "body_jntadr out of range");
      int type = d->jnt_type[j], da = d->jnt_dofadr[j], qa = d->jnt_qposadr[j];
      if (type == GMR_JNT_FREE) {
        if (jnum != 1 || p != 0) return fail(-4, "free joint must be the only joint of a top-level body");
        if (ref_body == 0) ref_body = b;
        for (int i = 0; i < 6; i++) {
          dof_kind[da + i] = i < 3 ? DOF_FREE_T : DOF_FREE_R;
          dof_qadr[da + i] = i < 3 ? qa + i : (i == 3 ? qa + 3 : -1);
          dof_parent[da + i] = i == 0 ? last : da + i - 1;
          dof_body[da + i] = b;
        }
        last = da + 5;
      } else if (type == GMR_JNT_SLIDE || type == GMR_JNT_HINGE) {
        if (type == GMR_JNT_SLIDE && seen_hinge) return fail(-4, "slide after hinge in one body is unsupported");
        if (type == GMR_JNT_HINGE) {
          seen_hinge = true;
          const double* jp = d->jnt_pos + 3 * j;
          if (jp[0] != 0 || jp[1] != 0 || jp[2] != 0) return fail(-4, "hinge joints with pos != 0 are unsupported");
        }
        dof_kind[da] = type == GMR_JNT_SLIDE ? DOF_SLIDE : DOF_HINGE;
        dof_qadr[da] = qa;
        dof_parent[da] = last;
        dof_body[da] = b;
        dof_limited[da] = d->jnt_limited[j] ? 1 : 0;
        dof_lo[da] = d->jnt_range[2 * j];
        dof_hi[da] = d->jnt_range[2 * j + 1];
        last = da;
      } else {
        return fail(-4, "unsupported joint type");
      }
    }
    body_lastdof[b] = last;
  }
  for (int i = 0; i < nv; i++)
    if (dof_parent[i] >= i) return fail(-4, "dofs must be ordered parent-before-child");
  // branch-induced sparsity tables (see DevModel)
  std::vector<int> dof_depth(nv, 0), dof_rowptr(nv + 1, 0);
  int max_depth = 0;
  for (int i = 0; i < nv; i++) {
    dof_depth[i] = dof_parent[i] < 0 ? 0 : dof_depth[dof_parent[i]] + 1;
    max_depth = std::max(max_depth, dof_depth[i]);
    dof_rowptr[i + 1] = dof_rowptr[i] + dof_depth[i] + 1;
  }
  if (max_depth > 31) return fail(-4, "kinematic chains longer than 32 dofs are unsupported");
  const int nnz = dof_rowptr[nv];
  std::vector<short> dof_anc(nnz, 0), dof_ancrow(nnz, 0), ent_row(nnz, 0);
  for (int i = 0; i < nv; i++) {
    int k = i;
    for (int a = dof_depth[i]; a >= 0; a--) {
      ent_row[dof_rowptr[i] + a] = (short)i;
      dof_anc[dof_rowptr[i] + a] = (short)k;
      dof_ancrow[dof_rowptr[i] + a] = (short)dof_rowptr[k];
      k = dof_parent[k];
    }
  }
  // pointer-jumping table of the FK sweep: 2^r-th ancestor of every body (-1 = the world is reached)
  int fk_rounds = 0;
  while ((1 << fk_rounds) < nlevel) fk_rounds++;
  std::vector<short> fk_jump((size_t)std::max(fk_rounds, 1) * nb, -1);
  for (int b = 1; b < nb; b++) {
    for (int r = 0; r < fk_rounds; r++) {
      int a2 = b;
      for (int k2 = 0; k2 < (1 << r) && a2 > 0; k2++) a2 = d->body_parentid[a2];
      fk_jump[(size_t)r * nb + b] = (short)(a2 > 0 ? a2 : -1);
    }
  }
  // levels (level 1 first)
  std::vector<int> level_start(nlevel + 1, 0), level_body;
  for (int lv = 1; lv <= nlevel; lv++) {
    level_start[lv - 1] = (int)level_body.size();
    for (int b = 1; b < nb; b++) if (depth[b] == lv) level_body.push_back(b);
  }
  level_start[nlevel] = (int)level_body.size();

  // ---- tasks, stage-major with stride GMR_MAX_TASKS, and per-dof subtree masks
  const int MT = GMR_MAX_TASKS;
  std::vector<int> task_body(2 * MT, 0), task_src(2 * MT, 0);
  std::vector<double> task_wpos(2 * MT, 0.0), task_wrot(2 * MT, 0.0);
  std::vector<unsigned> taskmask(2 * nv, 0u);
  for (int k = 0; k < 2; k++) {
    if (!d->use_table[k]) continue;
    for (int t = 0; t < d->n_task[k]; t++) {
      int body = d->task_body[k][t], src = d->task_src[k][t];
      if (body < 1 || body >= nb) return fail(-5, "task body id out of range");
      if (src < 0 || src >= d->n_scaled) return fail(-5, "task source index out of range");
      task_body[k * MT + t] = body; task_src[k * MT + t] = src;
      task_wpos[k * MT + t] = d->task_wpos[k][t]; task_wrot[k * MT + t] = d->task_wrot[k][t];
      for (int i = body_lastdof[body]; i >= 0; i = dof_parent[i]) taskmask[k * nv + i] |= 1u << t;
    }
  }
  // task tree per stage: parent = the task on the nearest proper-ancestor body that carries one
  std::vector<unsigned> topmask(2 * nv, 0u), childmask(2 * MT, 0u);
  std::vector<int> tasklev_start(2 * (MT + 1), 0), tasklev_task(2 * MT, 0);
  int n_tasklev[2] = {0, 0};
  for (int k = 0; k < 2; k++) {
    if (!d->use_table[k]) continue;
    const int m_k = d->n_task[k];
    std::vector<int> task_of_body(nb, -1), tparent(m_k, -1), height(m_k, 0);
    for (int t = 0; t < m_k; t++) {
      if (task_of_body[task_body[k * MT + t]] >= 0) return fail(-5, "two tasks of one stage on the same robot body");
      task_of_body[task_body[k * MT + t]] = t;
    }
    for (int t = 0; t < m_k; t++)
      for (int p = d->body_parentid[task_body[k * MT + t]]; p > 0; p = d->body_parentid[p])
        if (task_of_body[p] >= 0) { tparent[t] = task_of_body[p]; break; }
    // heights: process tasks by decreasing body depth so children come first
    std::vector<int> order(m_k);
    for (int t = 0; t < m_k; t++) order[t] = t;
    std::sort(order.begin(), order.end(), [&](int x, int y) { return depth[task_body[k * MT + x]] > depth[task_body[k * MT + y]]; });
    int hmax = 0;
    for (int t : order)
      if (tparent[t] >= 0) {
        childmask[k * MT + tparent[t]] |= 1u << t;
        height[tparent[t]] = std::max(height[tparent[t]], height[t] + 1);
        hmax = std::max(hmax, height[tparent[t]]);
      }
    n_tasklev[k] = hmax;
    int pos = 0;
    for (int h = 1; h <= hmax; h++) {
      tasklev_start[k * (MT + 1) + h - 1] = pos;
      for (int t = 0; t < m_k; t++) if (height[t] == h) tasklev_task[k * MT + pos++] = t;
    }
    tasklev_start[k * (MT + 1) + hmax] = pos;
    for (int l = 0; l < nv; l++) {
      const unsigned mk = taskmask[k * nv + l];
      for (int t = 0; t < m_k; t++)
        if (((mk >> t) & 1u) && (tparent[t] < 0 || !((mk >> tparent[t]) & 1u))) topmask[k * nv + l] |= 1u << t;
    }
  }
  // ---- dofs stage_integrate visits, box slots of the QP, constant spatial axes of a floating base
  std::vector<short> integ_dof, box_dof;
  std::vector<int> box_slot(nv, -1);
  for (int i = 0; i < nv; i++) {
    if (dof_kind[i] == DOF_SLIDE || dof_kind[i] == DOF_HINGE) {
      integ_dof.push_back((short)i);
      box_slot[i] = (int)box_dof.size();
      box_dof.push_back((short)i);
    } else if (dof_kind[i] == DOF_FREE_R && dof_qadr[i] >= 0) {
      integ_dof.push_back((short)i);
    }
  }
  const int fk_skip = (nv >= 6 && dof_kind[0] == DOF_FREE_T && dof_kind[1] == DOF_FREE_T && dof_kind[2] == DOF_FREE_T) ? 3 : 0;
  // ---- static schedule of the articulated-body solver (qp_solve_aba), per stage.
  // The scheduled dofs (those with a task below them) are cut into CHAINS: every dof continues into its tallest
  // child, its other children are the tops of side chains.  A chain runs bottom-up on one 8-lane group with the
  // accumulator X = [I | p] carried in REGISTERS from dof to dof; only a chain's top stores X in shared memory
  // (for the dof that gathers it as a side child, or for the floating-base block).  A pass handles one dof per
  // group; a chain idles (registers kept) while a side chain it gathers is still running.  The downward sweep
  // uses the same chains top-down with the link velocity v carried in registers.
  std::vector<int> aba_up, aba_ch, aba_dn;            // passes of stage 0, then of stage 1
  int aba_up0[2] = {0, 0}, aba_dn0[2] = {0, 0};
  std::vector<int> aba_merge(2 * MT * 2, 0), task_dof(2 * MT, -1);
  int aba_nup[2] = {0, 0}, aba_ndn[2] = {0, 0}, aba_nmerge[2] = {0, 0}, aba_root[2] = {-1, -1}, aba_rootdof[2] = {0, 0}, aba_doubles = GMR_KS;
  unsigned aba_kktmask[2] = {0u, 0u};
  for (int k = 0; k < 2; k++) {
    if (!d->use_table[k]) continue;
    const int m_k = d->n_task[k];
    std::vector<int> own(nv, -1);                       // first task hanging on the dof
    for (int t = 0; t < m_k; t++) {
      const int a = body_lastdof[task_body[k * MT + t]];
      task_dof[k * MT + t] = a;
      if (a < 0) continue;                              // a task on a body no dof moves
      if (own[a] < 0) { own[a] = t; aba_kktmask[k] |= 1u << t; }
      else { aba_merge[2 * (k * MT + aba_nmerge[k])] = own[a]; aba_merge[2 * (k * MT + aba_nmerge[k]) + 1] = t; aba_nmerge[k]++; }
    }
    std::vector<char> in(nv, 0);
    for (int i = 0; i < nv; i++) in[i] = taskmask[k * nv + i] != 0;
    // floating base (six unlimited dofs at the top of the tree): solved as one block by the kernel, so dofs
    // fb .. fb+4 leave the upward schedule (fb+5, the base link, still gathers its X) and all six the downward one
    int fb = -1;
    for (int i = 0; i + 5 < nv && fb < 0; i++)
      if (dof_kind[i] == DOF_FREE_T && dof_parent[i] < 0 && in[i + 5]) fb = i;
    auto skip_up = [&](int i) { return fb >= 0 && i >= fb && i < fb + 5; };
    auto skip_dn = [&](int i) { return fb >= 0 && i >= fb && i <= fb + 5; };
    std::vector<int> hgt(nv, 0);
    std::vector<std::vector<int>> kids(nv);
    for (int i = nv - 1; i >= 0; i--) {
      if (!in[i] || skip_up(i)) continue;
      const int pr = dof_parent[i];
      if (pr >= 0 && !skip_up(pr)) { hgt[pr] = std::max(hgt[pr], hgt[i] + 1); kids[pr].push_back(i); }
    }
    for (int i = 0; i < nv; i++) std::sort(kids[i].begin(), kids[i].end());
    // chains: cont[c] = the child c continues into (tallest, lowest index on ties)
    std::vector<int> cont(nv, -1), chain_of(nv, -1);
    for (int i = 0; i < nv; i++)
      for (int c : kids[i]) if (cont[i] < 0 || hgt[c] > hgt[cont[i]]) cont[i] = c;
    std::vector<std::vector<int>> chains;               // bottom -> top
    for (int i = nv - 1; i >= 0; i--) {
      if (!in[i] || skip_up(i) || cont[i] >= 0) continue;     // chain bottoms = scheduled dofs without children
      std::vector<int> ch{i};
      for (int c = i;;) {
        const int pr = dof_parent[c];
        if (pr < 0 || skip_up(pr) || cont[pr] != c) break;
        ch.push_back(pr); c = pr;
      }
      for (int c : ch) chain_of[c] = (int)chains.size();
      chains.push_back(ch);
    }
    const int nchain = (int)chains.size();
    // accumulators behind the task blocks: one per chain (its top's X)
    const int accbase = (m_k * GMR_KS * 8 + 15) & ~15;
    auto accoff = [&](int chain) { return accbase + chain * GMR_XS * 8; };
    aba_doubles = std::max(aba_doubles, accbase / 8 + nchain * GMR_XS);
    aba_root[k] = fb >= 0 ? accoff(chain_of[fb + 5]) : -1;
    aba_rootdof[k] = std::max(fb, 0);
    // ---- upward passes: chains start in order of their top's height (what a chain gathers always tops out lower)
    {
      std::vector<int> order(nchain);
      for (int c = 0; c < nchain; c++) order[c] = c;
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return hgt[chains[a].back()] < hgt[chains[b].back()]; });
      std::vector<int> done(nv, -1);
      int next = 0, finished = 0, cur[4] = {-1, -1, -1, -1}, pos[4] = {0, 0, 0, 0};
      aba_up0[k] = (int)aba_up.size() / 32;
      for (int t = 0; finished < nchain; t++) {
        if (t > 4 * nv + 8) return fail(-4, "internal: upward schedule does not terminate");
        for (int g = 0; g < 4; g++) if (cur[g] < 0 && next < nchain) { cur[g] = order[next++]; pos[g] = 0; }
        for (int g = 0; g < 4; g++) {
          int c = -1;
          if (cur[g] >= 0) {
            c = chains[cur[g]][pos[g]];
            bool ready = true;                            // every side chain it gathers finished in an earlier pass
            for (int kid : kids[c]) if (kid != cont[c] && !(done[kid] >= 0 && done[kid] < t)) ready = false;
            if (!ready) c = -1;
          }
          if (c < 0) { for (int z = 0; z < 8; z++) aba_up.push_back(z == 0 ? -1 : 0); continue; }
          const bool top = pos[g] + 1 == (int)chains[cur[g]].size();
          const int cb = (int)aba_ch.size() / 2;
          for (int kid : kids[c]) if (kid != cont[c]) { aba_ch.push_back(accoff(chain_of[kid])); aba_ch.push_back(kid * 64); }
          const bool is_root = fb >= 0 && c == fb + 5;       // the floating base's link: raw X for the 6x6 block
          const bool side = top && !is_root && dof_parent[c] >= 0 && !skip_up(dof_parent[c]);
          const int flags = (pos[g] > 0 ? 1 : 0) | (side ? 2 : 0) | (is_root ? 4 : 0) | ((box_slot[c] + 1) << 16);
          aba_up.push_back(c); aba_up.push_back(flags); aba_up.push_back(c * GMR_SS * 8); aba_up.push_back(c * 64);
          aba_up.push_back(own[c] >= 0 ? own[c] * GMR_KS * 8 : -1);
          aba_up.push_back(cb); aba_up.push_back((int)aba_ch.size() / 2); aba_up.push_back(accoff(cur[g]));
          done[c] = t;
          if (top) { cur[g] = -1; finished++; } else pos[g]++;
        }
        aba_nup[k]++;
      }
    }
    // ---- downward passes: the same chains top-down; a chain starts once its top's parent has stored its velocity
    {
      std::vector<std::vector<int>> dch;                 // chains without the dofs of the floating-base block
      for (auto& ch : chains) {
        std::vector<int> r;
        for (int i = (int)ch.size() - 1; i >= 0; i--) if (!skip_dn(ch[i])) r.push_back(ch[i]);
        if (!r.empty()) dch.push_back(r);
      }
      const int nd = (int)dch.size();
      std::vector<int> dep(nv, 0);
      for (int i = 0; i < nv; i++) if (in[i]) dep[i] = dof_parent[i] < 0 ? 0 : dep[dof_parent[i]] + 1;
      std::vector<int> order(nd);
      for (int c = 0; c < nd; c++) order[c] = c;
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return dep[dch[a][0]] < dep[dch[b][0]]; });
      // v is stored for dofs another chain starts from and for dofs that carry a task (the KKT check reads it)
      std::vector<char> store(nv, 0);
      for (auto& ch : dch) { const int pr = dof_parent[ch[0]]; if (pr >= 0) store[pr] = 1; }
      for (int i = 0; i < nv; i++) if (own[i] >= 0) store[i] = 1;
      std::vector<int> done(nv, -1);
      int next = 0, finished = 0, cur[4] = {-1, -1, -1, -1}, pos[4] = {0, 0, 0, 0};
      aba_dn0[k] = (int)aba_dn.size() / 16;
      for (int t = 0; finished < nd; t++) {
        if (t > 4 * nv + 8) return fail(-4, "internal: downward schedule does not terminate");
        for (int g = 0; g < 4; g++) if (cur[g] < 0 && next < nd) { cur[g] = order[next++]; pos[g] = 0; }
        for (int g = 0; g < 4; g++) {
          int c = -1;
          if (cur[g] >= 0) {
            c = dch[cur[g]][pos[g]];
            const int pr = dof_parent[c];
            if (pos[g] == 0 && pr >= 0 && !skip_dn(pr) && !(done[pr] >= 0 && done[pr] < t)) c = -1;   // parent's v not stored yet
          }
          if (c < 0) { aba_dn.push_back(-1); aba_dn.push_back(0); aba_dn.push_back(-1); aba_dn.push_back(0); continue; }
          const int pr = dof_parent[c];
          aba_dn.push_back(c); aba_dn.push_back((pos[g] > 0 ? 1 : 0) | (store[c] ? 2 : 0));
          aba_dn.push_back(pr >= 0 ? pr * 48 : -1); aba_dn.push_back(0);
          done[c] = t;
          if (pos[g] + 1 == (int)dch[cur[g]].size()) { cur[g] = -1; finished++; } else pos[g]++;
        }
        aba_ndn[k]++;
      }
    }
  }
  std::vector<double> dof_vmax(nv, INFINITY);
  if (d->dof_vmax)
    for (int i = 0; i < nv; i++) {
      dof_vmax[i] = d->dof_vmax[i];
      if (!(dof_vmax[i] >= 0)) return fail(-5, "dof_vmax must be >= 0 (+inf: unlimited)");
      if ((dof_kind[i] == DOF_FREE_T || dof_kind[i] == DOF_FREE_R) && std::isfinite(dof_vmax[i]))
        return fail(-5, "dof_vmax of a free joint's dofs must be +inf (the floating base is solved as an unconstrained block)");
    }
  for (int i = 0; i < d->n_scaled; i++)
    if (d->scaled_human_index[i] < 0 || d->scaled_human_index[i] >= d->n_human) return fail(-5, "scaled_human_index out of range");

  // ---- pack and upload
  std::vector<double> body_quat_padded((size_t)GMR_QS * nb, 0.0);
  for (int b = 0; b < nb; b++)
    for (int i = 0; i < 4; i++) body_quat_padded[(size_t)GMR_QS * b + i] = d->body_quat[4 * b + i];
  // bodies whose pose the IK reads: owners of scheduled joints, task frames, every ancestor of those, and the world
  std::vector<char> fk_need(nb, 0);
  fk_need[0] = 1;
  // a dof is "scheduled" when some task of some stage hangs below it; the others (fingers of a robot whose tasks
  // end at the wrists) never move and nothing reads their axes, so their bodies stay out of the sweep
  std::vector<short> cdof, udof;
  for (int i = 0; i < nv; i++) {
    const bool sched = (taskmask[i] | taskmask[nv + i]) != 0u;
    if (sched) fk_need[dof_body[i]] = 1;
    if (i < fk_skip) continue;                 // a floating base's translations: constant axes, written apart
    if (sched) cdof.push_back((short)i); else udof.push_back((short)i);
  }
  for (int v : task_body) if (v > 0 && v < nb) fk_need[v] = 1;
  for (int b = nb - 1; b > 0; b--) if (fk_need[b]) fk_need[d->body_parentid[b]] = 1;
  std::vector<short> fk_body;
  for (int b = 1; b < nb; b++) if (fk_need[b]) fk_body.push_back((short)b);      // the world's pose is a constant (stage_fk writes it apart)
  Blob bl;
  // tables of the fused kernel first: a CTA stages exactly this leading part of the blob in shared memory
  size_t o_parent = bl.push(vec(d->body_parentid, nb)), o_bpos = bl.push(vec(d->body_pos, 3 * nb)),
         o_bquat = bl.push(body_quat_padded), o_jadr = bl.push(vec(d->body_jntadr, nb)),
         o_jnum = bl.push(vec(d->body_jntnum, nb)),
         o_jtype = bl.push(vec(d->jnt_type, nj)), o_jq = bl.push(vec(d->jnt_qposadr, nj)),
         o_jd = bl.push(vec(d->jnt_dofadr, nj)), o_jaxis = bl.push(vec(d->jnt_axis, 3 * nj)),
         o_q0 = bl.push(vec(d->qpos0, d->nq)), o_dk = bl.push(dof_kind), o_dq = bl.push(dof_qadr),
         o_db = bl.push(dof_body), o_dl = bl.push(dof_limited), o_dlo = bl.push(dof_lo), o_dhi = bl.push(dof_hi),
         o_tm = bl.push(taskmask), o_si = bl.push(vec(d->scaled_human_index, d->n_scaled)),
         o_sc = bl.push(vec(d->scale, d->n_scaled)), o_po = bl.push(vec(d->pos_off, 3 * d->n_scaled)),
         o_ro = bl.push(vec(d->rot_off, 4 * d->n_scaled)), o_ft = bl.push(vec(d->is_foot, d->n_scaled)),
         o_tb = bl.push(task_body), o_ts = bl.push(task_src), o_wp = bl.push(task_wpos), o_wr = bl.push(task_wrot),
         o_fj = bl.push(fk_jump), o_fb = bl.push(fk_body), o_ap1 = bl.push(aba_up), o_ach = bl.push(aba_ch), o_adn = bl.push(aba_dn), o_amg = bl.push(aba_merge),
         o_atd = bl.push(task_dof), o_vm = bl.push(dof_vmax),
         o_ig = bl.push(integ_dof), o_bx = bl.push(box_dof), o_cd = bl.push(cdof), o_ud = bl.push(udof);
  const size_t staged_bytes = (bl.bytes.size() + 15) & ~size_t(15);
  // tables only the stage kernels read (gmr_build_qp: the assembly of H), from global memory
  size_t o_lstart = bl.push(level_start), o_lbody = bl.push(level_body), o_dp = bl.push(dof_parent),
         o_dd = bl.push(dof_depth), o_rp = bl.push(dof_rowptr), o_an = bl.push(dof_anc), o_ar = bl.push(dof_ancrow),
         o_er = bl.push(ent_row), o_tpm = bl.push(topmask), o_chm = bl.push(childmask), o_tls = bl.push(tasklev_start),
         o_tlt = bl.push(tasklev_task);

  // FK tail tables (KinematicsModel semantics: bodies under the first top-level body, float32, xyzw)
  const int nbt = nb - 1;
  std::vector<int> t_parent(nbt, -1), t_dof(nbt, -1);
  std::vector<float> t_trans(3 * nbt, 0.f), t_rot(4 * nbt, 0.f), t_axis(3 * nbt, 0.f);
  int root_dofs = 0;
  if (nb > 1) {
    int jadr = d->body_jntadr[1], jnum = d->body_jntnum[1];
    for (int j = jadr; j < jadr + jnum; j++) root_dofs += d->jnt_type[j] == GMR_JNT_FREE ? 6 : 1;
  }
  for (int b = 1; b < nb; b++) {
    int k = b - 1;
    t_parent[k] = d->body_parentid[b] - 1;
    for (int i = 0; i < 3; i++) t_trans[3 * k + i] = (float)d->body_pos[3 * b + i];
    t_rot[4 * k + 0] = (float)d->body_quat[4 * b + 1]; t_rot[4 * k + 1] = (float)d->body_quat[4 * b + 2];
    t_rot[4 * k + 2] = (float)d->body_quat[4 * b + 3]; t_rot[4 * k + 3] = (float)d->body_quat[4 * b + 0];
    if (b > 1 && d->body_jntnum[b] == 1) {
      int j = d->body_jntadr[b];
      t_dof[k] = d->jnt_dofadr[j] - root_dofs;
      for (int i = 0; i < 3; i++) t_axis[3 * k + i] = (float)d->jnt_axis[3 * j + i];
    }
  }

  gmr_model* m = new gmr_model();
  m->device = device;
  Blob tb;
  size_t t_p = tb.push(t_parent), t_t = tb.push(t_trans), t_r = tb.push(t_rot), t_a = tb.push(t_axis), t_d = tb.push(t_dof);
  cudaError_t e = cudaSuccess;
  if (device < 0) {          // host-only: no CUDA call at all; the table pointers below point into host memory
    m->host_blob = bl.bytes;
    m->blob = m->host_blob.data();
  } else {
    int prev = 0;
    cudaGetDevice(&prev);
    e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaMalloc(&m->blob, bl.bytes.size());
    if (e == cudaSuccess) e = cudaMemcpy(m->blob, bl.bytes.data(), bl.bytes.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&m->tail_blob, tb.bytes.size());
    if (e == cudaSuccess) e = cudaMemcpy(m->tail_blob, tb.bytes.data(), tb.bytes.size(), cudaMemcpyHostToDevice);
    cudaSetDevice(prev);
  }
  if (e != cudaSuccess) {
    if (m->blob) cudaFree(m->blob);
    if (m->tail_blob) cudaFree(m->tail_blob);
    delete m;
    return fail(-100, std::string("model upload: ") + cudaGetErrorString(e));
  }
  m->blob_bytes = bl.bytes.size();
  m->staged_bytes = staged_bytes;
  DevModel& h = m->host;
  h.nbody = nb; h.njnt = nj; h.nq = d->nq; h.nv = nv; h.n_human = d->n_human; h.n_scaled = d->n_scaled;
  h.nlevel = nlevel; h.ref_body = ref_body;
  for (int k = 0; k < 2; k++) { h.use[k] = d->use_table[k] ? 1 : 0; h.n_task[k] = d->use_table[k] ? d->n_task[k] : 0; }
  h.max_iter = d->max_iter; h.damping = d->damping; h.converge_eps = d->converge_eps;
  h.limit_gain = d->limit_gain; h.lm_damping = d->lm_damping; h.timestep = d->timestep; h.inv_timestep = 1.0 / d->timestep;
  char* B0 = m->blob;
  h.body_parent = (const int*)(B0 + o_parent); h.body_pos = (const double*)(B0 + o_bpos);
  h.body_quat = (const double*)(B0 + o_bquat); h.body_jntadr = (const int*)(B0 + o_jadr);
  h.body_jntnum = (const int*)(B0 + o_jnum); h.level_start = (const int*)(B0 + o_lstart);
  h.level_body = (const int*)(B0 + o_lbody); h.jnt_type = (const int*)(B0 + o_jtype);
  h.jnt_qposadr = (const int*)(B0 + o_jq); h.jnt_dofadr = (const int*)(B0 + o_jd);
  h.jnt_axis = (const double*)(B0 + o_jaxis); h.qpos0 = (const double*)(B0 + o_q0);
  h.dof_kind = (const int*)(B0 + o_dk); h.dof_qadr = (const int*)(B0 + o_dq); h.dof_parent = (const int*)(B0 + o_dp);
  h.dof_body = (const int*)(B0 + o_db); h.blob = B0; h.blob_bytes = (int)staged_bytes;
  h.fk_skip = fk_skip; h.n_integ = (int)integ_dof.size(); h.integ_dof = (const short*)(B0 + o_ig);
  h.n_box = (int)box_dof.size(); h.box_dof = (const short*)(B0 + o_bx);
  h.has_vmax = d->dof_vmax ? 1 : 0; h.dof_vmax = (const double*)(B0 + o_vm);
  h.dof_limited = (const int*)(B0 + o_dl); h.dof_lo = (const double*)(B0 + o_dlo); h.dof_hi = (const double*)(B0 + o_dhi);
  h.dof_taskmask = (const unsigned*)(B0 + o_tm); h.scaled_index = (const int*)(B0 + o_si);
  h.scale = (const double*)(B0 + o_sc); h.pos_off = (const double*)(B0 + o_po); h.rot_off = (const double*)(B0 + o_ro);
  h.is_foot = (const int*)(B0 + o_ft); h.task_body = (const int*)(B0 + o_tb); h.task_src = (const int*)(B0 + o_ts);
  h.task_wpos = (const double*)(B0 + o_wp); h.task_wrot = (const double*)(B0 + o_wr);
  h.nnz = nnz; h.max_depth = max_depth;
  h.dof_depth = (const int*)(B0 + o_dd); h.dof_rowptr = (const int*)(B0 + o_rp);
  h.dof_anc = (const short*)(B0 + o_an); h.dof_ancrow = (const short*)(B0 + o_ar);
  h.ent_row = (const short*)(B0 + o_er); h.dof_topmask = (const unsigned*)(B0 + o_tpm);
  h.task_childmask = (const unsigned*)(B0 + o_chm); h.tasklev_start = (const int*)(B0 + o_tls);
  for (int k = 0; k < 2; k++) {
    h.aba_nup[k] = aba_nup[k]; h.aba_ndn[k] = aba_ndn[k]; h.aba_nmerge[k] = aba_nmerge[k];
    h.aba_root[k] = aba_root[k]; h.aba_rootdof[k] = aba_rootdof[k];
    h.aba_kktmask[k] = aba_kktmask[k]; h.aba_up0[k] = aba_up0[k]; h.aba_dn0[k] = aba_dn0[k];
  }
  h.aba_doubles = aba_doubles;
  h.aba_up = (const int4*)(B0 + o_ap1);
  h.aba_ch = (const int2*)(B0 + o_ach); h.aba_dn = (const int4*)(B0 + o_adn); h.aba_merge = (const int2*)(B0 + o_amg);
  h.task_dof = (const int*)(B0 + o_atd);
  h.fk_rounds = fk_rounds; h.fk_jump = (const short*)(B0 + o_fj);
  h.n_fk = (int)fk_body.size(); h.fk_body = (const short*)(B0 + o_fb);
  h.n_cdof = (int)cdof.size(); h.cdof = (const short*)(B0 + o_cd); h.n_udof = (int)udof.size(); h.udof = (const short*)(B0 + o_ud);
  h.tasklev_task = (const int*)(B0 + o_tlt); h.n_tasklev[0] = n_tasklev[0]; h.n_tasklev[1] = n_tasklev[1];
  m->max_task = std::max(h.n_task[0], h.n_task[1]);
  char* T0 = m->tail_blob;
  m->tail.nbody = nbt; m->tail.parent = (const int*)(T0 + t_p); m->tail.trans = (const float*)(T0 + t_t);
  m->tail.rot_xyzw = (const float*)(T0 + t_r); m->tail.axis = (const float*)(T0 + t_a); m->tail.dof = (const int*)(T0 + t_d);
  m->tail.ndof = nv - root_dofs;
  *out = m;
  return 0;
}

extern "C" int gmr_model_destroy(gmr_model* m) {
  if (!m) return 0;
  if (m->device >= 0) {
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(m->device);
    if (m->blob) cudaFree(m->blob);
    if (m->tail_blob) cudaFree(m->tail_blob);
    cudaSetDevice(prev);
  }
  delete m;
  return 0;
}

// Host-side view of the static schedule of the box-QP solver (DevModel.aba_*), for tests of the flattener:
// copies table `which` of `stage` into out[0..cap) and returns the number of int32 it holds (or < 0).
extern "C" int gmr_model_table(const gmr_model* m, int which, int stage, int32_t* out, int cap) {
  if (!m || stage < 0 || stage > 1) return fail(-1, "bad argument");
  if (m->device >= 0) return fail(-6, "gmr_model_table needs a host-only model (device = GMR_HOST_ONLY)");
  const DevModel& h = m->host;
  std::vector<int32_t> v;
  auto span = [&](const void* p, size_t n) { const int32_t* q = (const int32_t*)p; v.assign(q, q + n); };
  const int up_end = h.aba_up0[stage] + h.aba_nup[stage];
  int ch_end = 0;
  for (int i = 0; i < up_end * 4; i++) ch_end = std::max(ch_end, h.aba_up[2 * i + 1].z);
  switch (which) {
    case GMR_TABLE_INFO:
      v = {h.nv, h.n_task[stage], h.aba_nup[stage], h.aba_ndn[stage], h.aba_nmerge[stage], h.aba_root[stage],
           h.aba_rootdof[stage], (int32_t)h.aba_kktmask[stage], h.aba_doubles, GMR_KS, GMR_XS, GMR_SS};
      break;
    case GMR_TABLE_UP: span(h.aba_up + 8 * (size_t)h.aba_up0[stage], 32 * (size_t)h.aba_nup[stage]); break;
    case GMR_TABLE_CHILD: span(h.aba_ch, 2 * (size_t)ch_end); break;          // child indices are global: from child 0
    case GMR_TABLE_DN: span(h.aba_dn + 4 * (size_t)h.aba_dn0[stage], 16 * (size_t)h.aba_ndn[stage]); break;
    case GMR_TABLE_MERGE: span(h.aba_merge + stage * GMR_MAX_TASKS, 2 * (size_t)h.aba_nmerge[stage]); break;
    case GMR_TABLE_TASK_DOF: span(h.task_dof + stage * GMR_MAX_TASKS, (size_t)h.n_task[stage]); break;
    case GMR_TABLE_DOF_PARENT: span(h.dof_parent, (size_t)h.nv); break;
    case GMR_TABLE_TASKMASK: span(h.dof_taskmask + stage * h.nv, (size_t)h.nv); break;
    case GMR_TABLE_FK_BODY: v.resize(h.n_fk); for (int i = 0; i < h.n_fk; i++) v[i] = h.fk_body[i]; break;
    default: return fail(-1, "unknown table");
  }
  if (out) for (size_t i = 0; i < v.size() && (int)i < cap; i++) out[i] = v[i];
  return (int)v.size();
}

// ---------------------------------------------------------------- batch
// the instantiation of the fused kernel for (stream dtype, model staged in shared memory)
// and the compile-time shared-memory layout (gmr_layouts.cuh; -1 = the run-time layout of RunArgs)
template <int L>
static const void* clip_kernel_of(int dtype, bool staged) {
  if (dtype == GMR_F32) return staged ? (const void*)clip_kernel<float, true, L> : (const void*)clip_kernel<float, false, -1>;
  return staged ? (const void*)clip_kernel<double, true, L> : (const void*)clip_kernel<double, false, -1>;
}
static const void* clip_kernel_ptr(int dtype, bool staged, int layout = -1) {
  switch (staged ? layout : -1) {
#define GMR_LAYOUT_CASE(k) case k: return clip_kernel_of<k>(dtype, staged);
    GMR_FOR_EACH_FIXED_LAYOUT(GMR_LAYOUT_CASE)
#undef GMR_LAYOUT_CASE
    default: return clip_kernel_of<-1>(dtype, staged);
  }
}

static Lay make_layout(int nq, int s, int nv, int nbody, int m, int nnz, int h, int blocks) {
  // Scratch behind the per-frame state (q, targets, x), reused along the tick:
  //   poses + spatial axes: FK .. stiffness; during the solve the link velocities v overlay the poses
  //                         (gmr_build_qp: F_l of the assembly does);
  //   Ud: per dof [U, alpha, beta] of the solver;
  //   e:  task errors (+ sin/cos of the angle), then the task forces of the KKT check;
  //   K:  `blocks` doubles - task stiffness blocks [K_t | g_t], then the solver's accumulators;
  //       before the stiffness pass the same span holds the local FK transforms / the staged human frame
  //       (gmr_build_qp: H overlays e/K once the assembly has folded them into F).
  Lay L{};
  int o = 0;
  auto take = [&](int n) { int r = o; o += (n + 1) & ~1; return r; };   // keep 16-byte alignment
  L.q = take(nq); L.scaled = take(7 * s); L.x = take(nv); L.st = take(10);
  L.xpos = take(3 * nbody); L.xquat = take(GMR_QS * nbody);
  o = std::max(o, L.xpos + ((std::max(GMR_SS, 6) * nv + 1) & ~1));   // F_l (stride GMR_SS) and v (stride 6) alias the body poses
  L.F = L.xpos; L.v = L.xpos;
  L.S = take(GMR_SS * nv);
  L.Ud = take(8 * nv);
  const int region2 = o;
  L.e = take(std::max(GMR_ES * m, 32));                      // also the scratch of the solver's root block (27) and its task forces
  L.K = take(std::max(blocks, GMR_KS * m));
  L.tl = L.K; L.ql = L.tl + ((3 * nbody + 1) & ~1);          // FK local transforms alias the blocks
  o = std::max(o, L.ql + ((GMR_QS * nbody + 1) & ~1));
  o = std::max(o, (L.K + 7 * h + 1) & ~1);                   // one frame of the human stream is staged there too
  L.H = region2;
  o = std::max(o, (L.H + nnz + 1) & ~1);
  L.total = (o + 31) & ~31;
  L.maxt = GMR_MAX_TASKS;
  return L;
}

extern "C" int gmr_model_layout(const gmr_model* m, int32_t* out, int cap) {
  if (!m) return fail(-1, "null model");
  const DevModel& h = m->host;
  const Lay L = make_layout(h.nq, h.n_scaled, h.nv, h.nbody, std::max(m->max_task, 1), h.nnz, h.n_human, std::max(h.aba_doubles, 1));
  const int n = (int)(sizeof(Lay) / sizeof(int));
  static_assert(sizeof(Lay) % sizeof(int) == 0, "Lay is a struct of ints");
  const int* src = (const int*)&L;
  if (out) for (int i = 0; i < n && i < cap; i++) out[i] = src[i];
  return n;
}

extern "C" int gmr_batch_create(gmr_model* const* models, int n_models, const int32_t* model_of_instance,
                                int B, void* stream, gmr_batch** out) {
  if (!models || n_models < 1 || B < 1 || !out) return fail(-1, "bad argument");
  if (n_models > 1 && !model_of_instance) return fail(-1, "model_of_instance required for n_models > 1");
  gmr_batch* b = new gmr_batch();
  b->device = models[0]->device;
  b->stream = (cudaStream_t)stream;
  b->B = B; b->n_models = n_models;
  std::vector<DevModel> hm;
  for (int i = 0; i < n_models; i++) {
    gmr_model* m = models[i];
    if (!m || m->device < 0 || m->device != b->device) { delete b; return fail(-1, "models must live on one device"); }
    b->models.push_back(m);
    hm.push_back(m->host);
    b->nq_max = std::max(b->nq_max, m->host.nq); b->s_max = std::max(b->s_max, m->host.n_scaled);
    b->h_max = std::max(b->h_max, m->host.n_human); b->nbody_max = std::max(b->nbody_max, m->host.nbody);
    b->nv_max = std::max(b->nv_max, m->host.nv); b->m_max = std::max(b->m_max, m->max_task);
    b->nnz_max = std::max(b->nnz_max, m->host.nnz); b->blocks_max = std::max(b->blocks_max, m->host.aba_doubles);
  }
  if (model_of_instance)
    for (int i = 0; i < B; i++)
      if (model_of_instance[i] < 0 || model_of_instance[i] >= n_models) { delete b; return fail(-1, "model_of_instance out of range"); }
  b->lay = make_layout(b->nq_max, b->s_max, b->nv_max, b->nbody_max, std::max(b->m_max, 1), b->nnz_max, b->h_max, b->blocks_max);
  int prev = 0;
  cudaGetDevice(&prev);
  cudaError_t e = cudaSetDevice(b->device);
  auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
  alloc((void**)&b->d_models, sizeof(DevModel) * n_models);
  alloc((void**)&b->d_q, sizeof(double) * (size_t)B * b->nq_max);
  alloc((void**)&b->d_scale, sizeof(double) * (size_t)B * b->s_max);
  alloc((void**)&b->d_counter, sizeof(int));
  alloc((void**)&b->d_done, sizeof(int) * (size_t)B);
  alloc((void**)&b->d_act, sizeof(unsigned long long) * GMR_ACT_WORDS * (size_t)B);
  if (model_of_instance) alloc((void**)&b->d_model_of, sizeof(int) * (size_t)B);
  if (e == cudaSuccess) e = cudaMemcpy(b->d_models, hm.data(), sizeof(DevModel) * n_models, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && model_of_instance)
    e = cudaMemcpy(b->d_model_of, model_of_instance, sizeof(int) * (size_t)B, cudaMemcpyHostToDevice);
  // launch geometry: one instance per warp, warps independent of each other (no CTA-wide barrier in the
  // tick loop).  Single-model batches stage the model tables in shared memory (one copy per CTA).
  const size_t per_warp = sizeof(double) * (size_t)b->lay.total;
  int dev_smem = 0, sms = 0;
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, b->device);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, b->device);
  if (e == cudaSuccess) {
    // the fused kernel is compiled for 128 registers (16 warps per SM): pick the CTA width that puts
    // the most warps on an SM; ties go to the wider CTA (the staged model is per CTA)
    const size_t blob = n_models == 1 ? models[0]->staged_bytes : 0;
    // static shared memory of the fused kernel (the staged DevModel) + the 1 KB the hardware reserves per CTA
    cudaFuncAttributes fa{};
    size_t stat = 0;
    for (int dt : {GMR_F32, GMR_F64})
      for (int st = 0; st < 2; st++)
        if (cudaFuncGetAttributes(&fa, clip_kernel_ptr(dt, st != 0)) == cudaSuccess) stat = std::max(stat, fa.sharedSizeBytes);
    const size_t fixed = stat + 1024;
    const size_t sm_total = 228 * 1024;
    b->smem_model = n_models == 1 && per_warp + blob + fixed <= (size_t)dev_smem && !getenv("GMR_DEBUG_GLOBAL_MODEL");
    // a single-model batch whose layout is one of the generated ones runs the kernel compiled for it: every
    // shared-memory offset is then an immediate instead of a constant-bank load rematerialised at each use
    b->layout_id = -1;
    if (b->smem_model && !getenv("GMR_DEBUG_RUNTIME_LAYOUT"))
      for (int k = 0; k < GMR_NUM_FIXED_LAYOUTS; k++) {
        const Lay f = fixed_layout(k);
        if (memcmp(&f, &b->lay, sizeof(Lay)) == 0) { b->layout_id = k; break; }
      }
    const size_t extra = fixed + (b->smem_model ? blob : 0);
    const int max_w = GMR_CTA_THREADS / 32;
    int best_w = 1, best_total = 0;
    for (int w = 1; w <= max_w; w++) {
      if ((size_t)w * per_warp + extra > (size_t)dev_smem + 1024) break;      // dynamic + static <= the opt-in limit
      int ctas = (int)std::min<size_t>(sm_total / ((size_t)w * per_warp + extra), (size_t)(max_w / w));
      if (ctas * w >= best_total) { best_total = ctas * w; best_w = w; }
    }
    // the four schedulers of an SM should hold the same number of warps: 13, 14 or 15 warps measured slower than 12
    // (profiles/r02_warps_sweep.txt; a warp on a fuller scheduler is slower and the slowest chain sets the launch)
    if (best_w >= 4) best_w -= best_w % 4;
    if (const char* ev = getenv("GMR_DEBUG_WARPS")) best_w = std::max(1, std::min(max_w, atoi(ev)));
    b->warps_per_cta = std::min(best_w, B);
    b->smem_bytes = per_warp * b->warps_per_cta;
    b->clip_smem_bytes = b->smem_bytes + (b->smem_model ? blob : 0);
    if (b->clip_smem_bytes + stat > (size_t)dev_smem) e = cudaErrorInvalidConfiguration;
  }
  // The dynamic shared-memory opt-in is a per-function, process-wide attribute: another batch (another robot, a
  // narrower CTA) would lower it under this one, so launch_clips sets it again right before every launch.
  const void* kf = clip_kernel_ptr(GMR_F32, b->smem_model, b->layout_id);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->clip_smem_bytes);
  int per_sm = 0;
  if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kf, 32 * b->warps_per_cta, b->clip_smem_bytes);
  if (e == cudaSuccess) {
    if (per_sm < 1) per_sm = 1;
    if (const char* ev = getenv("GMR_DEBUG_CTAS_PER_SM")) per_sm = std::max(1, std::min(per_sm, atoi(ev)));
    const int need = (B + b->warps_per_cta - 1) / b->warps_per_cta;
    b->grid = std::min(need, per_sm * sms);
  }
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    std::string msg = std::string("batch create: ") + cudaGetErrorString(e);
    gmr_batch_destroy(b);
    return fail(-100, msg);
  }
  *out = b;
  int rc = gmr_batch_reset(b, nullptr);
  if (rc == 0) rc = gmr_batch_set_scale(b, nullptr);
  if (rc) { std::string msg = g_err; gmr_batch_destroy(b); *out = nullptr; return fail(rc, msg); }
  return 0;
}

extern "C" int gmr_batch_destroy(gmr_batch* b) {
  if (!b) return 0;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(b->device);
  cudaFree(b->d_models); cudaFree(b->d_q); cudaFree(b->d_scale); cudaFree(b->d_counter); cudaFree(b->d_done); cudaFree(b->d_act);
  cudaFree(b->d_model_of); cudaFree(b->d_order);
  cudaFree(b->stage_in); cudaFree(b->stage_out); cudaFree(b->stage_T); cudaFree(b->stage_cs);
  if (b->s_in) { cudaStreamDestroy(b->s_in); cudaStreamDestroy(b->s_out); for (int i = 0; i < 3; i++) cudaEventDestroy(b->ev[i]); }
  cudaSetDevice(prev);
  delete b;
  return 0;
}

namespace {
struct DevGuard {
  int prev = 0;
  explicit DevGuard(int dev) { cudaGetDevice(&prev); cudaSetDevice(dev); }
  ~DevGuard() { cudaSetDevice(prev); }
};
__global__ void fill_rows_kernel(double* dst, int B, int stride, const DevModel* models, const int* model_of, int what) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * stride) return;
  const int inst = i / stride, k = i % stride;
  const DevModel& M = models[model_of ? model_of[inst] : 0];
  if (what == 0) dst[i] = k < M.nq ? M.qpos0[k] : 0.0;
  else dst[i] = k < M.n_scaled ? M.scale[k] : 0.0;
}
}  // namespace

extern "C" int gmr_batch_reset(gmr_batch* b, const double* q0_dev) {
  if (!b) return fail(-1, "null batch");
  DevGuard g(b->device);
  const int n = b->B * b->nq_max;
  if (q0_dev) CU(cudaMemcpyAsync(b->d_q, q0_dev, sizeof(double) * n, cudaMemcpyDeviceToDevice, b->stream));
  else fill_rows_kernel<<<(n + 255) / 256, 256, 0, b->stream>>>(b->d_q, b->B, b->nq_max, b->d_models, b->d_model_of, 0);
  CU(cudaGetLastError());
  CU(cudaMemsetAsync(b->d_act, 0, sizeof(unsigned long long) * GMR_ACT_WORDS * (size_t)b->B, b->stream));      // cold start: empty active set
  return 0;
}

extern "C" int gmr_batch_set_scale(gmr_batch* b, const double* scale_dev) {
  if (!b) return fail(-1, "null batch");
  DevGuard g(b->device);
  const int n = b->B * b->s_max;
  if (scale_dev) CU(cudaMemcpyAsync(b->d_scale, scale_dev, sizeof(double) * n, cudaMemcpyDeviceToDevice, b->stream));
  else fill_rows_kernel<<<(n + 255) / 256, 256, 0, b->stream>>>(b->d_scale, b->B, b->s_max, b->d_models, b->d_model_of, 1);
  CU(cudaGetLastError());
  return 0;
}

extern "C" int gmr_batch_set_order(gmr_batch* b, const int32_t* order_dev) {
  if (!b) return fail(-1, "null batch");
  DevGuard g(b->device);
  if (!order_dev) { if (b->d_order) { cudaFree(b->d_order); b->d_order = nullptr; } return 0; }
  if (!b->d_order) CU(cudaMalloc(&b->d_order, sizeof(int) * (size_t)b->B));
  CU(cudaMemcpyAsync(b->d_order, order_dev, sizeof(int) * (size_t)b->B, cudaMemcpyDeviceToDevice, b->stream));
  return 0;
}

extern "C" int gmr_batch_get_state(gmr_batch* b, double* q_dev) {
  if (!b || !q_dev) return fail(-1, "null argument");
  DevGuard g(b->device);
  CU(cudaMemcpyAsync(q_dev, b->d_q, sizeof(double) * (size_t)b->B * b->nq_max, cudaMemcpyDeviceToDevice, b->stream));
  return 0;
}

extern "C" int gmr_batch_dims(const gmr_batch* b, int32_t* B, int32_t* nq_max, int32_t* s_max, int32_t* h_max, int32_t* nbody_max) {
  if (!b) return fail(-1, "null batch");
  if (B) *B = b->B;
  if (nq_max) *nq_max = b->nq_max;
  if (s_max) *s_max = b->s_max;
  if (h_max) *h_max = b->h_max;
  if (nbody_max) *nbody_max = b->nbody_max;
  return 0;
}

extern "C" int gmr_batch_layout_id(const gmr_batch* b) { return b ? b->layout_id : -2; }

extern "C" int gmr_batch_launch_info(const gmr_batch* b, int32_t* grid, int32_t* warps_per_cta, int32_t* smem_bytes, int32_t* smem_per_instance) {
  if (!b) return fail(-1, "null batch");
  if (grid) *grid = b->grid;
  if (warps_per_cta) *warps_per_cta = b->warps_per_cta;
  if (smem_bytes) *smem_bytes = (int)b->clip_smem_bytes;
  if (smem_per_instance) *smem_per_instance = (int)(sizeof(double) * b->lay.total);
  return 0;
}

static int launch_clips(gmr_batch* b, const void* human_dev, int dtype, const int32_t* T_of_dev, int T_max, int to_ground,
                        double* qpos_out, double* scaled_out, const gmr_diag* diag, int t_begin = 0, int t_end = -1,
                        const int64_t* clip_start_dev = nullptr, int32_t* status_out = nullptr) {
  if (!b || !human_dev) return fail(-1, "null argument");
  if (dtype != GMR_F32 && dtype != GMR_F64) return fail(-1, "human_dtype must be GMR_F32 or GMR_F64");
  if (T_max < 1) return fail(-1, "T_max must be >= 1");
  DevGuard g(b->device);
  RunArgs a{};
  a.models = b->d_models; a.model_of = b->d_model_of;
  a.human = human_dev; a.human_frame_stride = (long long)b->h_max * 7;
  a.clip_start = (const long long*)clip_start_dev;
  a.T_of = T_of_dev; a.T_max = T_max; a.to_ground = to_ground;
  a.t_begin = t_begin; a.t_end = t_end < 0 ? T_max : std::min(t_end, T_max);
  a.scale = b->d_scale; a.s_max = b->s_max; a.q_state = b->d_q; a.nq_max = b->nq_max;
  a.qpos_out = qpos_out; a.scaled_out = scaled_out;
  a.err_out = diag ? diag->err_dev : nullptr; a.iters_out = diag ? diag->iters_dev : nullptr; a.nfact_out = diag ? diag->nfact_dev : nullptr;
  a.status_out = diag && diag->status_dev ? diag->status_dev : status_out;
  a.counter = b->d_counter; a.order = b->d_order; a.B = b->B; a.lay = b->lay;
  a.unit_frames = 32;
  if (const char* ev = getenv("GMR_DEBUG_UNIT_FRAMES")) a.unit_frames = std::max(1, atoi(ev));
  a.n_ranges = std::max(1, (a.t_end - a.t_begin + a.unit_frames - 1) / a.unit_frames);
  a.done = b->d_done; a.act_state = b->d_act;
  CU(cudaMemsetAsync(b->d_done, 0, sizeof(int) * (size_t)b->B, b->stream));
  CU(cudaMemsetAsync(b->d_counter, 0, sizeof(int), b->stream));
  const dim3 grid(b->grid), block(32 * b->warps_per_cta);
  void* kargs[] = {(void*)&a};
  const void* kern = clip_kernel_ptr(dtype, b->smem_model, b->layout_id);
  CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->clip_smem_bytes));
  CU(cudaLaunchKernel(kern, grid, block, kargs, b->clip_smem_bytes, b->stream));
  CU(cudaGetLastError());
  return 0;
}

extern "C" int gmr_batch_step(gmr_batch* b, const void* human_dev, int human_dtype, int offset_to_ground,
                              double* qpos_out_dev, double* scaled_out_dev, const gmr_diag* diag) {
  return launch_clips(b, human_dev, human_dtype, nullptr, 1, offset_to_ground, qpos_out_dev, scaled_out_dev, diag);
}

extern "C" int gmr_batch_run(gmr_batch* b, const void* human_dev, int human_dtype, const int32_t* T_of_instance_dev,
                             int T_max, int offset_to_ground, double* qpos_out_dev, const gmr_diag* diag) {
  return launch_clips(b, human_dev, human_dtype, T_of_instance_dev, T_max, offset_to_ground, qpos_out_dev, nullptr, diag);
}

extern "C" int gmr_batch_run_csr(gmr_batch* b, const void* human_dev, int human_dtype, const int64_t* clip_start_dev,
                                 int T_max, int offset_to_ground, double* qpos_out_dev, const gmr_diag* diag) {
  if (!clip_start_dev) return fail(-1, "null clip_start");
  return launch_clips(b, human_dev, human_dtype, nullptr, T_max, offset_to_ground, qpos_out_dev, nullptr, diag, 0, -1, clip_start_dev);
}

namespace {
// device pointer of a page-locked host buffer (null if the memory is not page-locked)
void* mapped_ptr(const void* host) {
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, host) != cudaSuccess || at.type != cudaMemoryTypeHost || !at.devicePointer) {
    cudaGetLastError();
    return nullptr;
  }
  return at.devicePointer;
}
}  // namespace

extern "C" int gmr_batch_step_host(gmr_batch* b, const void* human_host, int human_dtype, int offset_to_ground,
                                   double* qpos_out_host, double* scaled_out_host, int sync) {
  if (!b || !human_host || !qpos_out_host) return fail(-1, "null argument");
  DevGuard g(b->device);
  void* hi = mapped_ptr(human_host);
  void* qo = mapped_ptr(qpos_out_host);
  void* so = scaled_out_host ? mapped_ptr(scaled_out_host) : nullptr;
  if (!hi || !qo || (scaled_out_host && !so)) return fail(-6, "gmr_batch_step_host needs page-locked host buffers");
  int rc = launch_clips(b, hi, human_dtype, nullptr, 1, offset_to_ground, (double*)qo, (double*)so, nullptr);
  if (rc) return rc;
  if (sync) CU(cudaStreamSynchronize(b->stream));
  return 0;
}

extern "C" int gmr_batch_run_host_csr(gmr_batch* b, const void* human_host, int human_dtype, const int64_t* clip_start_host,
                                      int offset_to_ground, int cold_start, double* qpos_out_host, int32_t* status_out_host) {
  if (!b || !human_host || !qpos_out_host || !clip_start_host) return fail(-1, "null argument");
  if (human_dtype != GMR_F32 && human_dtype != GMR_F64) return fail(-1, "human_dtype must be GMR_F32 or GMR_F64");
  DevGuard g(b->device);
  int T_max = 0;
  for (int i = 0; i < b->B; i++) {
    const int64_t len = clip_start_host[i + 1] - clip_start_host[i];
    if (len < 0 || len > INT32_MAX) return fail(-1, "clip_start must be non-decreasing");
    T_max = std::max(T_max, (int)len);
  }
  const int64_t N = clip_start_host[b->B] - clip_start_host[0];
  if (clip_start_host[0] != 0) return fail(-1, "clip_start[0] must be 0");
  if (N == 0) return 0;
  if (!b->stage_cs) CU(cudaMalloc(&b->stage_cs, sizeof(int64_t) * ((size_t)b->B + 1)));
  CU(cudaMemcpyAsync(b->stage_cs, clip_start_host, sizeof(int64_t) * ((size_t)b->B + 1), cudaMemcpyHostToDevice, b->stream));
  if (cold_start) { int rc = gmr_batch_reset(b, nullptr); if (rc) return rc; }
  void* hi = mapped_ptr(human_host);
  void* qo = mapped_ptr(qpos_out_host);
  void* so = status_out_host ? mapped_ptr(status_out_host) : nullptr;
  if (hi && qo && (!status_out_host || so))       // zero-copy: the warps read / write the page-locked buffers over the bus
    return launch_clips(b, hi, human_dtype, nullptr, T_max, offset_to_ground, (double*)qo, nullptr, nullptr, 0, -1, b->stage_cs, (int32_t*)so);
  // pageable memory: one copy in, the launch, one copy out through device staging
  const size_t esz = human_dtype == GMR_F32 ? 4 : 8;
  const size_t in_bytes = (size_t)N * b->h_max * 7 * esz, out_bytes = (size_t)N * b->nq_max * sizeof(double) + (size_t)N * sizeof(int32_t);
  if (b->stage_in_bytes < in_bytes) {
    if (b->stage_in) cudaFree(b->stage_in);
    b->stage_in = nullptr; b->stage_in_bytes = 0;
    CU(cudaMalloc(&b->stage_in, in_bytes));
    b->stage_in_bytes = in_bytes;
  }
  if (b->stage_out_bytes < out_bytes) {
    if (b->stage_out) cudaFree(b->stage_out);
    b->stage_out = nullptr; b->stage_out_bytes = 0;
    CU(cudaMalloc(&b->stage_out, out_bytes));
    b->stage_out_bytes = out_bytes;
  }
  int32_t* st_dev = (int32_t*)((char*)b->stage_out + (size_t)N * b->nq_max * sizeof(double));
  CU(cudaMemcpyAsync(b->stage_in, human_host, in_bytes, cudaMemcpyHostToDevice, b->stream));
  int rc = launch_clips(b, b->stage_in, human_dtype, nullptr, T_max, offset_to_ground, (double*)b->stage_out, nullptr, nullptr, 0, -1,
                        b->stage_cs, status_out_host ? st_dev : nullptr);
  if (rc) return rc;
  CU(cudaMemcpyAsync(qpos_out_host, b->stage_out, (size_t)N * b->nq_max * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  if (status_out_host) CU(cudaMemcpyAsync(status_out_host, st_dev, (size_t)N * sizeof(int32_t), cudaMemcpyDeviceToHost, b->stream));
  return 0;
}

// Host-buffer form of gmr_batch_run: the dataset drivers' loop (scripts/bvh_to_robot_dataset.py:96-113)
// for B clips held in HOST memory.  The clips are cut into n_chunks frame ranges; range k+1 is copied
// to the device (strided 2-D copy, one row per clip) while range k is being retargeted and range k-1
// is copied back, on three streams chained by events.  The warm start crosses chunk boundaries through
// the batch state, so the result is bit-identical to one launch over the whole clips.
extern "C" int gmr_batch_run_host(gmr_batch* b, const void* human_host, int human_dtype, const int32_t* T_of_host,
                                  int T_max, int offset_to_ground, int cold_start, int n_chunks, double* qpos_out_host) {
  if (!b || !human_host || !qpos_out_host) return fail(-1, "null argument");
  if (human_dtype != GMR_F32 && human_dtype != GMR_F64) return fail(-1, "human_dtype must be GMR_F32 or GMR_F64");
  if (T_max < 1) return fail(-1, "T_max must be >= 1");
  DevGuard g(b->device);
  if (n_chunks <= 0) {
    // zero-copy: the kernel reads the stream from, and writes qpos to, pinned host memory directly
    // (one coalesced burst per frame in, one per frame out; ~5 GB/s each way at 7 M frames/s)
    cudaPointerAttributes ai{}, ao{};
    if (cudaPointerGetAttributes(&ai, human_host) != cudaSuccess || cudaPointerGetAttributes(&ao, qpos_out_host) != cudaSuccess ||
        ai.type != cudaMemoryTypeHost || ao.type != cudaMemoryTypeHost || !ai.devicePointer || !ao.devicePointer) {
      cudaGetLastError();
      return fail(-6, "zero-copy (n_chunks = 0) needs pinned (page-locked) host buffers");
    }
    const int32_t* T_dev = nullptr;
    if (T_of_host) {
      if (!b->stage_T) CU(cudaMalloc(&b->stage_T, sizeof(int) * (size_t)b->B));
      CU(cudaMemcpyAsync(b->stage_T, T_of_host, sizeof(int) * (size_t)b->B, cudaMemcpyHostToDevice, b->stream));
      T_dev = b->stage_T;
    }
    if (cold_start) { int rc = gmr_batch_reset(b, nullptr); if (rc) return rc; }
    return launch_clips(b, ai.devicePointer, human_dtype, T_dev, T_max, offset_to_ground, (double*)ao.devicePointer, nullptr, nullptr);
  }
  const size_t esz = human_dtype == GMR_F32 ? 4 : 8;
  const size_t in_row = (size_t)T_max * b->h_max * 7 * esz, out_row = (size_t)T_max * b->nq_max * sizeof(double);
  const size_t in_bytes = in_row * b->B, out_bytes = out_row * b->B;
  if (b->stage_in_bytes < in_bytes) {
    if (b->stage_in) cudaFree(b->stage_in);
    b->stage_in = nullptr; b->stage_in_bytes = 0;
    CU(cudaMalloc(&b->stage_in, in_bytes));
    b->stage_in_bytes = in_bytes;
  }
  if (b->stage_out_bytes < out_bytes) {
    if (b->stage_out) cudaFree(b->stage_out);
    b->stage_out = nullptr; b->stage_out_bytes = 0;
    CU(cudaMalloc(&b->stage_out, out_bytes));
    b->stage_out_bytes = out_bytes;
  }
  if (T_of_host) {
    if (!b->stage_T) CU(cudaMalloc(&b->stage_T, sizeof(int) * (size_t)b->B));
    CU(cudaMemcpyAsync(b->stage_T, T_of_host, sizeof(int) * (size_t)b->B, cudaMemcpyHostToDevice, b->stream));
  }
  if (!b->s_in) {
    CU(cudaStreamCreateWithFlags(&b->s_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&b->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 3; i++) CU(cudaEventCreateWithFlags(&b->ev[i], cudaEventDisableTiming));
  }
  n_chunks = std::max(1, std::min(n_chunks, T_max));
  if (cold_start) { int rc = gmr_batch_reset(b, nullptr); if (rc) return rc; }
  // the copy streams start after everything already queued on the batch stream (buffer reuse)
  CU(cudaEventRecord(b->ev[2], b->stream));
  CU(cudaStreamWaitEvent(b->s_in, b->ev[2], 0));
  CU(cudaStreamWaitEvent(b->s_out, b->ev[2], 0));
  const size_t in_frame = (size_t)b->h_max * 7 * esz, out_frame = (size_t)b->nq_max * sizeof(double);
  // frame ranges: with 3 or more ranges the first and the last are short (about T/10, a whole number of
  // work units): only the first range's upload and the last range's download are not hidden behind a
  // kernel, so they are made cheap; the middle is split evenly
  std::vector<int> cut(1, 0);
  if (n_chunks >= 3 && T_max >= 96) {
    const int edge = std::max(32, (T_max / 10 + 31) / 32 * 32);
    cut.push_back(edge);
    for (int k = 1; k < n_chunks - 2; k++) cut.push_back(edge + (int)((long long)(T_max - 2 * edge) * k / (n_chunks - 2)));
    cut.push_back(T_max - edge);
  } else {
    for (int k = 1; k < n_chunks; k++) cut.push_back((int)((long long)T_max * k / n_chunks));
  }
  cut.push_back(T_max);
  for (size_t k = 0; k + 1 < cut.size(); k++) {
    const int t0 = cut[k], t1 = cut[k + 1];
    if (t1 <= t0) continue;
    CU(cudaMemcpy2DAsync((char*)b->stage_in + t0 * in_frame, in_row, (const char*)human_host + t0 * in_frame, in_row,
                         (t1 - t0) * in_frame, b->B, cudaMemcpyHostToDevice, b->s_in));
    CU(cudaEventRecord(b->ev[0], b->s_in));
    CU(cudaStreamWaitEvent(b->stream, b->ev[0], 0));
    int rc = launch_clips(b, b->stage_in, human_dtype, T_of_host ? b->stage_T : nullptr, T_max, offset_to_ground,
                          (double*)b->stage_out, nullptr, nullptr, t0, t1);
    if (rc) return rc;
    CU(cudaEventRecord(b->ev[1], b->stream));
    CU(cudaStreamWaitEvent(b->s_out, b->ev[1], 0));
    CU(cudaMemcpy2DAsync((char*)qpos_out_host + t0 * out_frame, out_row, (const char*)b->stage_out + t0 * out_frame, out_row,
                         (t1 - t0) * out_frame, b->B, cudaMemcpyDeviceToHost, b->s_out));
  }
  // the batch stream is done when the last copy-back is
  CU(cudaEventRecord(b->ev[2], b->s_out));
  CU(cudaStreamWaitEvent(b->stream, b->ev[2], 0));
  return 0;
}

// ---------------------------------------------------------------- stage entry points
static StageArgs stage_args(gmr_batch* b) {
  StageArgs a{};
  a.models = b->d_models; a.model_of = b->d_model_of; a.B = b->B; a.lay = b->lay;
  a.nq_max = b->nq_max; a.s_max = b->s_max; a.nbody_max = b->nbody_max; a.nv_max = b->nv_max; a.m_max = std::max(b->m_max, 1);
  a.scale = b->d_scale; a.q = b->d_q;
  return a;
}
#define STAGE_LAUNCH(kernel, b, a) do { \
    const int w_ = std::min(b->warps_per_cta, 8);      /* stage kernels carry no launch bound: stay narrow */ \
    const size_t sm_ = sizeof(double) * (size_t)b->lay.total * w_; \
    CU(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); \
    kernel<<<(b->B + w_ - 1) / w_, 32 * w_, sm_, b->stream>>>(a); \
    CU(cudaGetLastError()); } while (0)

extern "C" int gmr_target_prep(gmr_batch* b, const void* human_dev, int human_dtype, int offset_to_ground, double* scaled_out_dev) {
  if (!b || !human_dev || !scaled_out_dev) return fail(-1, "null argument");
  DevGuard g(b->device);
  StageArgs a = stage_args(b);
  a.human = human_dev; a.human_inst_stride = (long long)b->h_max * 7; a.to_ground = offset_to_ground; a.scaled_out = scaled_out_dev;
  if (human_dtype == GMR_F32) STAGE_LAUNCH(prep_kernel<float>, b, a);
  else if (human_dtype == GMR_F64) STAGE_LAUNCH(prep_kernel<double>, b, a);
  else return fail(-1, "bad human_dtype");
  return 0;
}

extern "C" int gmr_fk(gmr_batch* b, const double* q_dev, double* xpos_out_dev, double* xquat_out_dev) {
  if (!b || !xpos_out_dev || !xquat_out_dev) return fail(-1, "null argument");
  DevGuard g(b->device);
  StageArgs a = stage_args(b);
  if (q_dev) a.q = q_dev;
  a.xpos_out = xpos_out_dev; a.xquat_out = xquat_out_dev;
  STAGE_LAUNCH(fk_kernel, b, a);
  return 0;
}

extern "C" int gmr_build_qp(gmr_batch* b, int stage, const double* scaled_dev, double* e_out_dev, double* H_out_dev,
                            double* c_out_dev, double* lo_out_dev, double* hi_out_dev) {
  if (!b || !scaled_dev || !H_out_dev || !c_out_dev || !lo_out_dev || !hi_out_dev) return fail(-1, "null argument");
  if (stage < 0 || stage > 1) return fail(-1, "stage must be 0 or 1");
  DevGuard g(b->device);
  StageArgs a = stage_args(b);
  a.stage = stage; a.scaled = scaled_dev; a.e_out = e_out_dev; a.H_out = H_out_dev; a.c_out = c_out_dev;
  a.lo_out = lo_out_dev; a.hi_out = hi_out_dev;
  STAGE_LAUNCH(build_qp_kernel, b, a);
  return 0;
}

extern "C" int gmr_solve_ik(gmr_batch* b, int stage, const double* scaled_dev, double* dq_out_dev, int32_t* nfact_out_dev) {
  if (!b || !scaled_dev || !dq_out_dev) return fail(-1, "null argument");
  if (stage < 0 || stage > 1) return fail(-1, "stage must be 0 or 1");
  DevGuard g(b->device);
  StageArgs a = stage_args(b);
  a.stage = stage; a.scaled = scaled_dev; a.dq_out = dq_out_dev; a.nfact_out = nfact_out_dev;
  STAGE_LAUNCH(solve_ik_kernel, b, a);
  return 0;
}

extern "C" int gmr_qp_solve(int device, void* stream, int n_qp, int n, const double* H_dev, const double* c_dev,
                            const double* lo_dev, const double* hi_dev, double* x_out_dev, int32_t* nfact_out_dev) {
  if (n_qp < 1 || n < 1 || n > GMR_MAX_NV) return fail(-1, "bad QP size (n <= 64)");
  if (!H_dev || !c_dev || !lo_dev || !hi_dev || !x_out_dev) return fail(-1, "null argument");
  DevGuard g(device);
  const int w = 4;
  const size_t smem = sizeof(double) * (size_t)w * (n * (n + 1) + 4 * n);
  CU(cudaFuncSetAttribute(qp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  qp_kernel<<<(n_qp + w - 1) / w, 32 * w, smem, (cudaStream_t)stream>>>(n_qp, n, H_dev, c_dev, lo_dev, hi_dev, x_out_dev, nfact_out_dev);
  CU(cudaGetLastError());
  return 0;
}

extern "C" int gmr_fk_bodies_f32(gmr_model* m, void* stream, int N, const float* root_pos_dev, const float* root_rot_xyzw_dev,
                                 const float* dof_pos_dev, float* body_pos_out_dev, float* body_rot_out_dev) {
  if (!m || m->device < 0 || N < 1 || !root_pos_dev || !root_rot_xyzw_dev || !dof_pos_dev || !body_pos_out_dev) return fail(-1, "bad argument");
  DevGuard g(m->device);
  float* scratch = nullptr;
  if (!body_rot_out_dev) CU(cudaMallocAsync(&scratch, sizeof(float) * (size_t)N * m->tail.nbody * 4, (cudaStream_t)stream));
  fk_tail_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(m->tail, N, root_pos_dev, root_rot_xyzw_dev, dof_pos_dev,
                                                                    body_pos_out_dev, body_rot_out_dev, scratch);
  cudaError_t e = cudaGetLastError();
  if (scratch) cudaFreeAsync(scratch, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(-100, cudaGetErrorString(e));
  return 0;
}

// ---------------------------------------------------------------- roofline probe
// Peak of the CUDA-core FMA pipe the kernels run on (MEASURED_PEAKS.json holds only HBM and the
// bf16 tensor peak): 8 independent FMA chains per thread, every SM saturated.
namespace {
template <typename T>
__global__ void fma_probe_kernel(T* sink, int iters, T a, T b) {
  T x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
    x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
  }
  T r = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (r == (T)123456789) sink[0] = r;     // never true; keeps the chains alive
}
template <typename T>
int fma_probe(int device, double* tflops_out) {
  DevGuard g(device);
  int sms = 0;
  CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  T* sink = nullptr;
  CU(cudaMalloc(&sink, sizeof(T)));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  const int blocks = sms * 8, threads = 256, iters = 1 << 15;
  double best = 0;
  for (int rep = 0; rep < 4; rep++) {
    CU(cudaEventRecord(e0, 0));
    fma_probe_kernel<T><<<blocks, threads>>>(sink, iters, (T)0.999999, (T)1e-6);
    CU(cudaEventRecord(e1, 0));
    CU(cudaEventSynchronize(e1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = 2.0 * 8.0 * (double)iters * blocks * threads;
    if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
  *tflops_out = best;
  return 0;
}
}  // namespace

// ---------------------------------------------------------------- BVH motion -> frame stream
extern "C" int gmr_bvh_to_stream(int device, void* stream, const double* lpos_dev, const double* euler_deg_dev,
                                 const int32_t* clip_start_dev, int n_clips, int n_frames, int n_bones,
                                 const int32_t* parents_dev, const int32_t* channel_axis, const int32_t* src_pos_dev,
                                 const int32_t* src_rot_dev, int n_rows, int out_dtype, void* stream_out_dev,
                                 int head_row, const int32_t* foot_rows, double* height_out_dev) {
  if (!lpos_dev || !euler_deg_dev || !clip_start_dev || !parents_dev || !channel_axis || !src_pos_dev || !src_rot_dev ||
      !stream_out_dev || n_clips < 0 || n_frames < 0 || n_bones < 1 || n_rows < 1)
    return fail(-1, "gmr_bvh_to_stream: bad argument");
  if (out_dtype != GMR_F32 && out_dtype != GMR_F64) return fail(-1, "gmr_bvh_to_stream: dtype must be GMR_F32 or GMR_F64");
  for (int k = 0; k < 3; k++)
    if (channel_axis[k] < 0 || channel_axis[k] > 2) return fail(-1, "gmr_bvh_to_stream: channel axis must be 0, 1 or 2");
  if (head_row >= n_rows) return fail(-1, "gmr_bvh_to_stream: head row out of range");
  if (n_frames == 0 || n_clips == 0) return 0;
  DevGuard g(device);
  cudaStream_t st = (cudaStream_t)stream;
  BvhArgs a{};
  a.lpos = lpos_dev; a.euler = euler_deg_dev; a.clip_start = clip_start_dev;
  a.C = n_clips; a.N = n_frames; a.J = n_bones; a.parents = parents_dev;
  a.ax0 = channel_axis[0]; a.ax1 = channel_axis[1]; a.ax2 = channel_axis[2];
  a.src_pos = src_pos_dev; a.src_rot = src_rot_dev; a.h = n_rows; a.out = stream_out_dev;
  a.head = head_row;
  for (int k = 0; k < 4; k++) {
    a.feet[k] = foot_rows ? foot_rows[k] : -1;
    if (a.feet[k] >= n_rows) return fail(-1, "gmr_bvh_to_stream: foot row out of range");
  }
  a.height = height_out_dev;
  void *sign = nullptr, *G = nullptr;
  CU(cudaMallocAsync(&sign, (size_t)n_frames * n_bones, st));
  cudaError_t e = cudaMallocAsync(&G, (size_t)n_frames * n_bones * 8 * sizeof(double), st);
  if (e != cudaSuccess) { cudaFreeAsync(sign, st); return fail(-3, cudaGetErrorString(e)); }
  a.sign = (signed char*)sign; a.G = (double*)G;
  const int tb = 128;
  bvh_sign_kernel<<<(n_clips * n_bones + tb - 1) / tb, tb, 0, st>>>(a);
  bvh_fk_kernel<<<(n_frames + tb - 1) / tb, tb, 0, st>>>(a);
  const long long ne = (long long)n_frames * n_rows;
  if (out_dtype == GMR_F32) bvh_emit_kernel<float><<<(unsigned)((ne + 255) / 256), 256, 0, st>>>(a);
  else bvh_emit_kernel<double><<<(unsigned)((ne + 255) / 256), 256, 0, st>>>(a);
  if (height_out_dev) bvh_height_kernel<<<(n_clips + tb - 1) / tb, tb, 0, st>>>(a);
  e = cudaGetLastError();
  cudaFreeAsync(sign, st);
  cudaFreeAsync(G, st);
  if (e != cudaSuccess) return fail(-3, cudaGetErrorString(e));
  return 0;
}

// ---------------------------------------------------------------- the same without any allocation
// Scratch of gmr_bvh_to_stream_static: [G: GMR_BVH_STATIC_MAX_ELEMS x 8 doubles | sign: GMR_BVH_STATIC_MAX_ELEMS bytes]
// in the module's own global memory (32.5 MiB on every device that loads the library, until exit).  The name below is
// a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress, once per device.
#define GMR_BVH_STATIC_G_BYTES ((size_t)GMR_BVH_STATIC_MAX_ELEMS * 8 * sizeof(double))
static __device__ __align__(256) unsigned char g_bvh_static_scratch[GMR_BVH_STATIC_G_BYTES + GMR_BVH_STATIC_MAX_ELEMS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_bvh_static_mu;
void* g_bvh_static_ptr[kMaxDevices] = {};
}  // namespace

extern "C" int gmr_bvh_to_stream_static(int device, void* stream, const double* lpos_dev, const double* euler_deg_dev,
                                        const int32_t* clip_start_dev, int n_clips, int n_frames, int n_bones,
                                        const int32_t* parents_dev, const int32_t* channel_axis, const int32_t* src_pos_dev,
                                        const int32_t* src_rot_dev, int n_rows, int out_dtype, void* stream_out_dev,
                                        int head_row, const int32_t* foot_rows, double* height_out_dev) {
  if (!lpos_dev || !euler_deg_dev || !clip_start_dev || !parents_dev || !channel_axis || !src_pos_dev || !src_rot_dev ||
      !stream_out_dev || n_clips < 0 || n_frames < 0 || n_bones < 1 || n_rows < 1)
    return fail(-1, "gmr_bvh_to_stream_static: bad argument");
  if (out_dtype != GMR_F32 && out_dtype != GMR_F64) return fail(-1, "gmr_bvh_to_stream_static: dtype must be GMR_F32 or GMR_F64");
  for (int k = 0; k < 3; k++)
    if (channel_axis[k] < 0 || channel_axis[k] > 2) return fail(-1, "gmr_bvh_to_stream_static: channel axis must be 0, 1 or 2");
  if (head_row >= n_rows) return fail(-1, "gmr_bvh_to_stream_static: head row out of range");
  if (device < 0 || device >= kMaxDevices) return fail(-1, "gmr_bvh_to_stream_static: device ordinal out of range");
  const long long elems = (long long)n_frames * n_bones;
  if (elems > (long long)GMR_BVH_STATIC_MAX_ELEMS)
    return fail(-4, "gmr_bvh_to_stream_static: n_frames * n_bones = " + std::to_string(elems) + " exceeds the static scratch (GMR_BVH_STATIC_MAX_ELEMS = " +
                        std::to_string((long long)GMR_BVH_STATIC_MAX_ELEMS) + "); split the call or use gmr_bvh_to_stream");
  if (n_frames == 0 || n_clips == 0) return 0;
  DevGuard g(device);
  unsigned char* base = nullptr;
  {   // set-up, once per device, with that device current
    std::lock_guard<std::mutex> lk(g_bvh_static_mu);
    if (!g_bvh_static_ptr[device]) {
      void* p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_bvh_static_scratch);
      if (e != cudaSuccess || !p) return fail(-3, std::string("gmr_bvh_to_stream_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
      g_bvh_static_ptr[device] = p;
    }
    base = (unsigned char*)g_bvh_static_ptr[device];
  }
  cudaStream_t st = (cudaStream_t)stream;
  BvhArgs a{};
  a.lpos = lpos_dev; a.euler = euler_deg_dev; a.clip_start = clip_start_dev;
  a.C = n_clips; a.N = n_frames; a.J = n_bones; a.parents = parents_dev;
  a.ax0 = channel_axis[0]; a.ax1 = channel_axis[1]; a.ax2 = channel_axis[2];
  a.src_pos = src_pos_dev; a.src_rot = src_rot_dev; a.h = n_rows; a.out = stream_out_dev;
  a.head = head_row;
  for (int k = 0; k < 4; k++) {
    a.feet[k] = foot_rows ? foot_rows[k] : -1;
    if (a.feet[k] >= n_rows) return fail(-1, "gmr_bvh_to_stream_static: foot row out of range");
  }
  a.height = height_out_dev;
  a.G = (double*)base;
  a.sign = (signed char*)(base + GMR_BVH_STATIC_G_BYTES);
  // the array keeps the last call's contents: frames no clip covers must read "no flip", not a stale sign
  CU(cudaMemsetAsync(a.sign, 0, (size_t)elems, st));
  const int tb = 128;
  bvh_sign_kernel<<<(n_clips * n_bones + tb - 1) / tb, tb, 0, st>>>(a);
  bvh_fk_kernel<<<(n_frames + tb - 1) / tb, tb, 0, st>>>(a);
  const long long ne = (long long)n_frames * n_rows;
  if (out_dtype == GMR_F32) bvh_emit_kernel<float><<<(unsigned)((ne + 255) / 256), 256, 0, st>>>(a);
  else bvh_emit_kernel<double><<<(unsigned)((ne + 255) / 256), 256, 0, st>>>(a);
  if (height_out_dev) bvh_height_kernel<<<(n_clips + tb - 1) / tb, tb, 0, st>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(-3, cudaGetErrorString(e));
  return 0;
}

// ---------------------------------------------------------------- SMPL-X model output -> frame stream
extern "C" int gmr_smplx_to_stream(int device, void* stream, const float* full_pose_dev, const float* joints_dev,
                                   int n_joints_in, const int32_t* src_start_dev, const int32_t* clip_of_out_dev,
                                   const double* time_dev, const int32_t* aligned_dev, int n_frames_in, int n_frames_out,
                                   int n_joints, const int32_t* parents_dev, int out_dtype, void* stream_out_dev) {
  if (!full_pose_dev || !joints_dev || !src_start_dev || !clip_of_out_dev || !time_dev || !aligned_dev || !parents_dev ||
      !stream_out_dev || n_frames_in < 0 || n_frames_out < 0 || n_joints < 1 || n_joints_in < n_joints)
    return fail(-1, "gmr_smplx_to_stream: bad argument");
  if (n_joints > SMPLX_MAX_JOINTS) return fail(-1, "gmr_smplx_to_stream: at most 64 joints");
  if (out_dtype != GMR_F32 && out_dtype != GMR_F64) return fail(-1, "gmr_smplx_to_stream: dtype must be GMR_F32 or GMR_F64");
  if (n_frames_out == 0) return 0;
  DevGuard g(device);
  SmplxArgs a{};
  a.pose = full_pose_dev; a.joints = joints_dev; a.Jall = n_joints_in; a.src_start = src_start_dev;
  a.clip_of = clip_of_out_dev; a.time = time_dev; a.aligned = aligned_dev;
  a.N = n_frames_in; a.N_out = n_frames_out; a.J = n_joints; a.parents = parents_dev; a.out = stream_out_dev;
  const int tb = 64;
  if (out_dtype == GMR_F32) smplx_stream_kernel<float><<<(n_frames_out + tb - 1) / tb, tb, 0, (cudaStream_t)stream>>>(a);
  else smplx_stream_kernel<double><<<(n_frames_out + tb - 1) / tb, tb, 0, (cudaStream_t)stream>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(-3, cudaGetErrorString(e));
  return 0;
}

// ---------------------------------------------------------------- short math probe (tests only)
namespace {
__global__ void math_probe_kernel(int n, const double* x, const double* y, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[3 * i] = rcp_pos(x[i]);
  out[3 * i + 1] = rsqrt_pos(x[i]);
  out[3 * i + 2] = atan_ratio(fmin(x[i], y[i]), fmax(x[i], y[i]));
}
}  // namespace

extern "C" int gmr_probe_math(int device, int n, const double* x_dev, const double* y_dev, double* out_dev) {
  if (n < 1 || !x_dev || !y_dev || !out_dev) return fail(-1, "bad argument");
  DevGuard g(device);
  math_probe_kernel<<<(n + 127) / 128, 128>>>(n, x_dev, y_dev, out_dev);
  CU(cudaGetLastError());
  return 0;
}

extern "C" int gmr_probe_fma_peak(int device, int fp64, double* tflops_out) {
  if (!tflops_out) return fail(-1, "null argument");
  return fp64 ? fma_probe<double>(device, tflops_out) : fma_probe<float>(device, tflops_out);
}
