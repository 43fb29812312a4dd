// mjb_api.cu -- extern "C" entry points of libmjb200.so (see include/mjb200.h).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "mjb200.h"
#include "mjb_internal.h"
#include "mjb_groups.h"
#include "mjb_kernels.cuh"

static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_timing{0};
struct TimedLaunch {
  cudaEvent_t first, second;
  int group;  // StageGroup, or -1 for the utility kernels
};
static std::vector<TimedLaunch> g_events;
static std::mutex g_mu;

static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CUDA_OK(expr)                                                                            \
  do {                                                                                           \
    cudaError_t e__ = (expr);                                                                    \
    if (e__ != cudaSuccess) return fail(-10, std::string(#expr) + ": " + cudaGetErrorString(e__)); \
  } while (0)

struct MjbModel {
  ModelSizes ms;
  ModelOffsets mo;
  const unsigned char* blob;
  int blob_bytes;
  int blob_hot;  // bytes of the prefix the kernels stage in shared memory (= blob_bytes unless the model carries large cold tables)
  int col_feat;  // CollisionFeature bits the model needs (pair types present in the filtered pair list)
  int has_boxbox;  // box-box pairs present: they run GJK / EPA + multi-contact (the convex variant) unless DisableBit.NATIVECCD
};

struct MjbData {
  const MjbModel* m;
  DataSizes dsz;
  DataPtrs d;
  DataStrides st;
  int scratch_floats;
  int device;
  int nsm;
  int smem_optin;
  float* hbuf;        // nv > 32: the solver's per-world Hessians (global memory), allocated at bind
  int* queue;         // G_COUNT dynamic world counters (one per stage group), zeroed at the start of every mjb_run_stages
  float* rk_scratch;  // Runge-Kutta temporaries (reference forward.py:465-475 allocates them per call): cudaMalloc'ed by mjb_prepare / on first use
  size_t rk_floats;
  // launch configuration of the pipeline kernel, resolved on the first launch
  struct GroupCfg {  // launch configuration of one stage-group kernel, resolved on first use
    int ready, wpb, grid, in_smem, scratch_floats;
    size_t smem;
  } cfg[G_COUNT + COLF_COUNT + 5];  // [G_COUNT + v]: collision variant v > 0; then elliptic make_constraint / solve, wide smooth_velocity / euler
};

// ---- field-table introspection --------------------------------------------------------------
static const char* const k_names0[] = {
#define X(n) #n,
    MJB_MODEL_SIZES(X)
#undef X
        nullptr};
static const char* const k_names1[] = {
#define X(n) #n,
    MJB_MODEL_INT_FIELDS(X)
#undef X
        nullptr};
static const char* const k_names2[] = {
#define X(n) #n,
    MJB_MODEL_REAL_FIELDS(X)
#undef X
        nullptr};
static const char* const k_names3[] = {
#define X(n) #n,
    MJB_OPT_INT_FIELDS(X)
#undef X
        nullptr};
static const char* const k_names4[] = {
#define X(n) #n,
    MJB_OPT_REAL_FIELDS(X)
#undef X
        nullptr};
static const char* const k_names5[] = {
#define X(n) #n,
    MJB_DATA_SIZES(X)
#undef X
        nullptr};
static const char* const k_names6[] = {
#define X(n, c) #n,
    MJB_DATA_INT_FIELDS(X)
#undef X
        nullptr};
static const char* const k_names7[] = {
#define X(n, c) #n,
    MJB_DATA_REAL_FIELDS(X)
#undef X
        nullptr};
static const char* const* const k_tables[8] = {k_names0, k_names1, k_names2, k_names3, k_names4, k_names5, k_names6, k_names7};

extern "C" int mjb_field_count(int kind) {
  if (kind < 0 || kind > 7) return 0;
  int n = 0;
  while (k_tables[kind][n]) n++;
  return n;
}
extern "C" const char* mjb_field_name(int kind, int index) {
  if (index < 0 || index >= mjb_field_count(kind)) return nullptr;
  return k_tables[kind][index];
}

// ---- model / data handles -------------------------------------------------------------------
extern "C" int mjb_model_create(const int* sizes, int nsizes, const void* blob, size_t nbytes, const int* int_offsets, int nint,
                                const int* real_offsets, int nreal, MjbModel** out) {
  if (nsizes != mjb_field_count(0) || nint != mjb_field_count(1) || nreal != mjb_field_count(2))
    return fail(-1, "mjb_model_create: field table size mismatch");
  if (nbytes % 16 != 0 || (reinterpret_cast<uintptr_t>(blob) & 15)) return fail(-2, "mjb_model_create: blob must be 16-byte aligned/sized");
  MjbModel* m = new MjbModel();
  int k = 0;
#define X(n) m->ms.n = sizes[k++];
  MJB_MODEL_SIZES(X)
#undef X
  k = 0;
#define X(n) m->mo.n = int_offsets[k++];
  MJB_MODEL_INT_FIELDS(X)
#undef X
  k = 0;
#define X(n) m->mo.n = real_offsets[k++];
  MJB_MODEL_REAL_FIELDS(X)
#undef X
  if (m->ms.is_sparse) {
    delete m;
    return fail(-3, "sparse models are not implemented");
  }
  if (m->ms.nv > 128) {
    delete m;
    return fail(-3, "nv > 128 is not implemented");
  }
  m->blob = static_cast<const unsigned char*>(blob);
  m->blob_bytes = (int)nbytes;
  m->blob_hot = (m->ms.blob_hot_bytes > 0 && m->ms.blob_hot_bytes < (int)nbytes) ? m->ms.blob_hot_bytes : (int)nbytes;
  m->col_feat = 0;
  m->has_boxbox = 0;
  if (m->ms.nxn > 0) {  // pair types present in the statically filtered pair list (one small read-back at creation)
    std::vector<int> pair(2 * (size_t)m->ms.nxn), gtype((size_t)m->ms.ngeom);
    const unsigned char* db = static_cast<const unsigned char*>(blob);
    cudaError_t e = cudaMemcpy(pair.data(), db + m->mo.nxn_pair, pair.size() * sizeof(int), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(gtype.data(), db + m->mo.geom_type, gtype.size() * sizeof(int), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
      delete m;
      return fail(-10, std::string("mjb_model_create: ") + cudaGetErrorString(e));
    }
    for (int p = 0; p < m->ms.nxn; p++) {
      const int ta = gtype[pair[2 * p]], tb = gtype[pair[2 * p + 1]];
      const int t1 = ta < tb ? ta : tb, t2 = ta < tb ? tb : ta;
      if (t2 == GEOM_BOX && t1 != GEOM_PLANE) m->col_feat |= COLF_BOX;  // sphere-box, capsule-box, box-box
      if (t1 == GEOM_BOX && t2 == GEOM_BOX) m->has_boxbox = 1;
      if (t1 == GEOM_ELLIPSOID || t1 == GEOM_CYLINDER || t1 == GEOM_MESH || t2 == GEOM_ELLIPSOID || t2 == GEOM_CYLINDER || t2 == GEOM_MESH) m->col_feat |= COLF_CONVEX;
      if (t1 == GEOM_HFIELD) m->col_feat |= COLF_CONVEX;  // height fields run the per-prism GJK / EPA
    }
  }
  if (m->ms.ngrouppair > 0) m->col_feat |= COLF_SAP;  // the two-level broadphase lives in the variants that carry the sweep-and-prune workspace
  *out = m;
  return 0;
}
extern "C" void mjb_model_destroy(MjbModel* m) { delete m; }

extern "C" int mjb_data_bind(const MjbModel* m, const int* sizes, int nsizes, int* const* int_ptrs, int nint, float* const* real_ptrs,
                             int nreal, MjbData** out) {
  if (!m || !sizes || !int_ptrs || !real_ptrs || !out) return fail(-1, "mjb_data_bind: null argument");
  if (nsizes != mjb_field_count(5) || nint != mjb_field_count(6) || nreal != mjb_field_count(7))
    return fail(-1, "mjb_data_bind: field table size mismatch");
  MjbData* d = new MjbData();
  d->m = m;
  int k = 0;
#define X(n) d->dsz.n = sizes[k++];
  MJB_DATA_SIZES(X)
#undef X
  k = 0;
#define X(n, c) d->d.n = int_ptrs[k++];
  MJB_DATA_INT_FIELDS(X)
#undef X
  k = 0;
#define X(n, c) d->d.n = real_ptrs[k++];
  MJB_DATA_REAL_FIELDS(X)
#undef X
  {  // per-world strides from the size expressions of the field tables
#define X(n) const int n = m->ms.n; (void)n;
    MJB_MODEL_SIZES(X)
#undef X
#define X(n) const int n = d->dsz.n; (void)n;
    MJB_DATA_SIZES(X)
#undef X
#define X(n, c) d->st.n = (c);
    MJB_DATA_INT_FIELDS(X)
    MJB_DATA_REAL_FIELDS(X)
#undef X
  }
  {  // a field with elements must come with a pointer (empty fields -- na == 0, nmocap == 0 ... -- may be null)
    const char* missing = nullptr;
#define X(n, c) if (!d->d.n && !missing && (((#c)[0] == '0' && (#c)[1] == 0) ? d->dsz.naconmax > 0 : ((c) > 0 && d->dsz.nworld > 0))) missing = #n;
    {
#define Y(n) const int n = m->ms.n; (void)n;
      MJB_MODEL_SIZES(Y)
#undef Y
#define Y(n) const int n = d->dsz.n; (void)n;
      MJB_DATA_SIZES(Y)
#undef Y
      MJB_DATA_INT_FIELDS(X)
      MJB_DATA_REAL_FIELDS(X)
    }
#undef X
    if (missing) {
      const std::string name = missing;
      delete d;
      return fail(-2, "mjb_data_bind: null pointer for field " + name);
    }
  }
  d->scratch_floats = chol_scratch_floats(m->ms.nv, m->ms.nv_pad);
  d->rk_scratch = nullptr;
  d->rk_floats = 0;
  d->queue = nullptr;
  d->hbuf = nullptr;
  memset(d->cfg, 0, sizeof(d->cfg));
  cudaError_t e = cudaGetDevice(&d->device);
  if (e != cudaSuccess) {
    delete d;
    return fail(-10, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
  }
  cudaDeviceGetAttribute(&d->nsm, cudaDevAttrMultiProcessorCount, d->device);
  cudaDeviceGetAttribute(&d->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, d->device);
  static const int dynamic = getenv("MJB_DYNAMIC_WORLDS") ? atoi(getenv("MJB_DYNAMIC_WORLDS")) : 1;
  if (dynamic) {
    e = cudaMalloc(&d->queue, G_COUNT * sizeof(int));
    if (e != cudaSuccess) {
      delete d;
      return fail(-10, std::string("cudaMalloc of the world queue: ") + cudaGetErrorString(e));
    }
  }
  if ((m->ms.nv > 32 || MJB_SOL_JSMEM) && d->dsz.nworld > 0 && m->ms.nv > 0) {  // the solver's Hessians (mjb_smooth.cuh:solver_scr)
    e = cudaMalloc(&d->hbuf, (size_t)d->dsz.nworld * m->ms.nv * chol_ld(m->ms.nv_pad) * sizeof(float));
    if (e != cudaSuccess) {
      if (d->queue) cudaFree(d->queue);
      delete d;
      return fail(-10, std::string("cudaMalloc of the solver Hessians: ") + cudaGetErrorString(e));
    }
  }
  *out = d;
  return 0;
}
extern "C" void mjb_data_destroy(MjbData* d) {
  if (d && d->hbuf) cudaFree(d->hbuf);
  if (d && d->rk_scratch) cudaFree(d->rk_scratch);
  if (d && d->queue) cudaFree(d->queue);
  delete d;
}

float* mjb_internal_rk_scratch(MjbData* d, size_t floats) {
  if (d->rk_floats >= floats && d->rk_scratch) return d->rk_scratch;
  if (d->rk_scratch) cudaFree(d->rk_scratch);
  d->rk_scratch = nullptr;
  d->rk_floats = 0;
  cudaError_t e = cudaMalloc(&d->rk_scratch, floats * sizeof(float));
  if (e != cudaSuccess) {
    d->rk_scratch = nullptr;
    fail(-10, std::string("cudaMalloc of the Runge-Kutta scratch: ") + cudaGetErrorString(e));
    return nullptr;
  }
  d->rk_floats = floats;
  return d->rk_scratch;
}

// ---- launches -------------------------------------------------------------------------------
static void fill_params(KParams& K, const MjbModel* m, const MjbData* d, const MjbOption* opt, uint32_t stages) {
  memset(&K, 0, sizeof(K));
  K.ms = m->ms;
  K.mo = m->mo;
  K.dsz = d->dsz;
  K.d = d->d;
  K.st = d->st;
  static_assert(sizeof(OptView) == sizeof(MjbOption), "option struct mismatch");
  memcpy(&K.o, opt, sizeof(OptView));
  K.blob = m->blob;
  K.blob_bytes = m->blob_hot;
  K.cold_global = m->blob_hot < m->blob_bytes;
  K.scratch_floats = d->scratch_floats;
  K.stages = stages;
  K.queue = d->queue;
  K.hbuf = d->hbuf;
}

int mjb_internal_params(const MjbModel* m, const MjbData* d, KParams* K) {
  if (!m || !d || !K) return fail(-1, "null handle");
  MjbOption o;
  memset(&o, 0, sizeof(o));
  fill_params(*K, m, d, &o, 0);
  return 0;
}
int mjb_internal_fail(int code, const char* msg) { return fail(code, msg); }
void mjb_internal_count_launch(void) { g_launches++; }
int mjb_internal_nworld(const MjbData* d) { return d->dsz.nworld; }

// warps (= worlds in flight) per CTA; MJB_WARPS_PER_BLOCK overrides the per-group default
static int warps_per_block(int dflt) {
  static int env = -1;
  if (env < 0) {
    const char* e = getenv("MJB_WARPS_PER_BLOCK");
    env = e ? atoi(e) : 0;
    if (env > 8) env = 8;
  }
  return env > 0 ? env : dflt;
}

template <typename F>
static int timed_launch(cudaStream_t stream, F&& launch, int group = -1, bool count = true) {
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const bool timing = count && g_timing.load() != 0;
  if (timing) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, stream);
  }
  launch();
  if (timing) {
    cudaEventRecord(e1, stream);
    std::lock_guard<std::mutex> lk(g_mu);
    g_events.push_back(TimedLaunch{e0, e1, group});
  }
  if (count) g_launches++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(-11, std::string("kernel launch: ") + cudaGetErrorString(e));
  return 0;
}

// the stage-group kernels are instantiated in their own translation units (mjb_g_*.cu), compiled in parallel
static GroupKernel group_kernel_of(int g, bool in_smem, int variant) {
  switch (g) {
    case G_POS: return variant ? mjb_gk_pos_narrow(in_smem) : mjb_gk_pos(in_smem);
    case G_COL: return variant == COLF_COUNT - 1 ? mjb_gk_col_convex(in_smem) : mjb_gk_col(in_smem, variant);
    case G_CON: return mjb_gk_con(in_smem, variant);
    case G_VEL: return variant ? mjb_gk_vel_wide(in_smem) : mjb_gk_vel(in_smem);
    case G_SOL: return mjb_gk_sol(in_smem, variant);
    default: return variant ? mjb_gk_eul_wide(in_smem) : mjb_gk_eul(in_smem);
  }
}

// kernel variant of a group under the given options: collision follows the model's pair types and the broadphase option of THIS
// call; make_constraint / solve are compiled once per cone type
static int group_variant(int g, const MjbModel* m, const MjbData* d, int broadphase, int cone, int disableflags) {
  if (g == G_COL) {
    const bool convex = (m->col_feat & COLF_CONVEX) || (m->has_boxbox && !(disableflags & DSBL_NATIVECCD));
    return convex ? COLF_COUNT - 1 : (m->col_feat | (broadphase != 0 ? COLF_SAP : 0));
  }
  if (g == G_POS) {
    // smooth_position with four worlds per warp: measured on B200 at 8192 worlds, panda (nv 9) 224 -> 204 us/step, humanoid (nv 27)
    // 69 -> 80 us for the group, apollo (nv 25) -2 %: only models whose trees fit a quarter warp gain.  MJB_POS_NARROW = 0 / 1 forces it.
    static const int force = getenv("MJB_POS_NARROW") ? atoi(getenv("MJB_POS_NARROW")) : -1;
    if (force >= 0) return m->ms.nv > 32 ? 0 : force;
    return (m->ms.nv <= 16 && d->dsz.nworld >= 42 * d->nsm) ? 1 : 0;
  }
  if (g == G_VEL || g == G_EUL) {
    // Several worlds per warp (variant 0) save instructions but leave a quarter of the warps to hide the latencies of these stages: it
    // pays once the GPU is full.  Measured on B200 (148 SMs), humanoid: 1024 worlds 240 -> 205 us/step with one warp per world (variant
    // 1), 2048: 266 -> 236, 4096: equal, 8192: 564 -> 541 the other way; apollo and panda cross at the same place.  nv > 32 always takes
    // the wide build (dense factorisations want all lanes).  MJB_SMOOTH_WIDE = 0 / 1 forces the choice.
    static const int force = getenv("MJB_SMOOTH_WIDE") ? atoi(getenv("MJB_SMOOTH_WIDE")) : -1;
    if (m->ms.nv > 32) return 1;
    if (force >= 0) return force;
    return d->dsz.nworld < 42 * d->nsm ? 1 : 0;
  }
  return ((g == G_CON || g == G_SOL) && cone == 1) ? 1 : 0;
}
static MjbData::GroupCfg& group_cfg(MjbData* d, int g, int variant) {
  return d->cfg[!variant ? g : g == G_COL ? G_COUNT + variant : G_COUNT + COLF_COUNT + (g == G_CON ? 0 : g == G_SOL ? 1 : g == G_VEL ? 2 : g == G_EUL ? 3 : 4)];
}
// resolves (once) the launch configuration of a group kernel: CTA width, shared memory, resident grid.  Sets function attributes and
// queries occupancy, so it must not run under stream capture: mjb_prepare() calls it ahead of time.
static int resolve_cfg(int g, int variant, const MjbModel* m, MjbData* d) {
  MjbData::GroupCfg& c = group_cfg(d, g, variant);
  if (c.ready) return 0;
  const int wpw = 32 / group_lw(g, variant);  // worlds per warp
  int wpb = warps_per_block(g == G_SOL ? group_threads(G_SOL) / 32 : wpw > 1 ? MJB_LW_WPB : 8);
  if (32 * wpb > group_threads(g)) wpb = group_threads(g) / 32;
  c.scratch_floats = group_scratch_floats(g, m->ms.nv, m->ms.nv_pad, d->dsz.njmax_pad, m->ms.ngeom, m->ms.nxn, variant, m->ms.ngroup);
  // fewer warps per CTA when the per-world scratch of a large model would not fit beside the model tables
  {
    const size_t blob = group_uses_blob(g) && m->blob_hot <= 96 * 1024 ? (size_t)m->blob_hot + 16 : 0;
    while (wpb > 1 && (size_t)wpb * wpw * c.scratch_floats * sizeof(float) + blob > (size_t)d->smem_optin) wpb--;
  }
  if ((size_t)c.scratch_floats * sizeof(float) > 16 * 1024 && wpb > 1) {
    // large per-world scratch: the widest CTA that fits may leave half an SM's shared memory unused (4 warps x 34 kB = one CTA per SM,
    // 3 warps = two): pick the width with the most resident warps
    const size_t blob = group_uses_blob(g) && m->blob_hot <= 96 * 1024 ? (size_t)m->blob_hot + 16 : 0;
    int best = wpb, best_warps = 0;
    for (int cand = wpb; cand >= 1; cand--) {
      const size_t need = (size_t)cand * wpw * c.scratch_floats * sizeof(float) + blob;
      const int per_sm = (int)(((size_t)d->smem_optin + 1024) / (need + 1024));  // 1 kB of shared memory is reserved per CTA
      if (per_sm * cand > best_warps) { best_warps = per_sm * cand; best = cand; }
    }
    wpb = best;
  }
  const size_t scr_bytes = (size_t)wpb * wpw * c.scratch_floats * sizeof(float);
  const bool in_smem =
      group_uses_blob(g) && (size_t)m->blob_hot + 16 + scr_bytes <= (size_t)d->smem_optin && m->blob_hot <= 96 * 1024;
  const size_t smem = in_smem ? (size_t)m->blob_hot + 16 + scr_bytes : scr_bytes;
  if (smem > (size_t)d->smem_optin) return fail(-5, "model needs more shared memory than the device offers");
  GroupKernel k0 = group_kernel_of(g, in_smem, variant);
  CUDA_OK(cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k0, 32 * wpb, smem));
  if (occ < 1) occ = 1;
  int grid = (d->dsz.nworld + wpb * wpw - 1) / (wpb * wpw);
  const int maxgrid = d->nsm * occ;
  if (grid > maxgrid) grid = maxgrid;
  c.wpb = wpb;
  c.grid = grid;
  c.in_smem = in_smem;
  c.smem = smem;
  c.ready = 1;
  return 0;
}

static int launch_group(int g, const MjbModel* m, MjbData* d, KParams& K, uint32_t stages, cudaStream_t stream) {
  const int variant = group_variant(g, m, d, K.o.broadphase, K.o.cone, K.o.disableflags);
  {
    const int rc = resolve_cfg(g, variant, m, d);
    if (rc) return rc;
  }
  MjbData::GroupCfg& c = group_cfg(d, g, variant);
  K.stages = stages;
  K.scratch_floats = c.scratch_floats;
  GroupKernel kern = group_kernel_of(g, c.in_smem != 0, variant);
  const int grid = c.grid, threads = 32 * c.wpb;
  const size_t smem = c.smem;
  // Programmatic dependent launch: the next group's CTAs may be scheduled (and stage the model blob)
  // while this group's last CTAs drain; they wait on griddepcontrol.wait before touching Data.
  static const int pdl = getenv("MJB_PDL") ? atoi(getenv("MJB_PDL")) : 0;  // measured slower on B200 (0.747 vs 0.717 ms/step): off by default
  return timed_launch(stream, [&] {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kern, K);
  }, g);
}

extern "C" int mjb_run_stages(const MjbModel* m, MjbData* d, const MjbOption* opt, uint32_t stages, void* cuda_stream) {
  if (!m || !d || !opt) return fail(-1, "null handle");
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  if (opt->cone != 0 && opt->cone != 1) return fail(-4, "unknown cone type");
  if (opt->solver != SOLVER_NEWTON && opt->solver != SOLVER_CG) return fail(-4, "unsupported solver");
  if ((stages & MJB_STAGE_IMPLICIT) && (stages & MJB_STAGE_EULER)) return fail(-4, "euler and implicit stages are exclusive");
  KParams K;
  fill_params(K, m, d, opt, stages);
  // One small kernel clears the contact counters (collision_driver.py:779) and the world queues.  Deliberately not memset nodes:
  // those run on a copy engine, where they queue behind the caller's host <-> device transfers and stall the step (measured).
  {
    const bool col = (stages & MJB_STAGE_COLLISION) && opt->run_collision_detection;
    static const int memsets = getenv("MJB_CLEAR_MEMSET") ? atoi(getenv("MJB_CLEAR_MEMSET")) : 0;
    if (memsets) {
      if (col) {
        CUDA_OK(cudaMemsetAsync(d->d.nacon, 0, sizeof(int), stream));
        CUDA_OK(cudaMemsetAsync(d->d.ncollision, 0, sizeof(int), stream));
      }
      if (d->queue) CUDA_OK(cudaMemsetAsync(d->queue, 0, G_COUNT * sizeof(int), stream));
    } else if (col || d->queue) {
      const int rc = timed_launch(stream, [&] { mjb_util_clear(col ? d->d.nacon : nullptr, col ? d->d.ncollision : nullptr, d->queue, G_COUNT, stream); }, -1, false);
      if (rc) return rc;
    }
  }
  for (int g = 0; g < G_COUNT; g++) {
    const uint32_t gs = stages & group_mask(g);
    if (!gs) continue;
    if (g == G_COL && !opt->run_collision_detection) continue;
    const int rc = launch_group(g, m, d, K, gs, stream);
    if (rc) return rc;
  }
  return 0;
}

static uint32_t integrator_stage(const MjbOption* opt) {
  if (opt->integrator == 0) return MJB_STAGE_EULER;
  if (opt->integrator == 3) return MJB_STAGE_IMPLICIT;
  return 0;
}

extern "C" int mjb_forward(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) { return mjb_run_stages(m, d, opt, MJB_STAGES_FORWARD, s); }
extern "C" int mjb_step(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) {
  if (!opt) return fail(-1, "null option");
  const uint32_t integ = integrator_stage(opt);
  if (opt->integrator == 1) {  // forward.py:1013-1016: forward, then rungekutta4
    const int rc = mjb_run_stages(m, d, opt, MJB_STAGES_FORWARD, s);
    return rc ? rc : mjb_rungekutta4(m, d, opt, s);
  }
  if (!integ) return fail(-4, "integrator not implemented");
  return mjb_run_stages(m, d, opt, MJB_STAGES_FORWARD | integ, s);
}

// Everything a first call would otherwise do lazily (function attributes, occupancy queries, the Runge-Kutta scratch allocation),
// for the kernel variants `opt` selects: after it, step / forward / every stage entry can be captured into a CUDA graph.
extern "C" int mjb_prepare(const MjbModel* m, MjbData* d, const MjbOption* opt) {
  if (!m || !d || !opt) return fail(-1, "null handle");
  for (int g = 0; g < G_COUNT; g++) {
    const int rc = resolve_cfg(g, group_variant(g, m, d, opt->broadphase, opt->cone, opt->disableflags), m, d);
    if (rc) return rc;
  }
  const size_t smem = (size_t)warps_per_block(4) * d->scratch_floats * sizeof(float);
  CUDA_OK(mjb_util_solve_m_smem(smem));
  if (opt->integrator == 1 && !mjb_internal_rk_scratch(d, (size_t)mjb_rk_scratch_floats(m, d))) return -10;
  return 0;
}
extern "C" int mjb_step1(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) {
  return mjb_run_stages(m, d, opt, MJB_STAGES_FWD_POSITION | MJB_STAGE_FACTOR_M | MJB_STAGE_ZERO_SENSOR | MJB_STAGE_FWD_VELOCITY, s);
}
extern "C" int mjb_step2(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) {
  if (!opt) return fail(-1, "null option");
  uint32_t integ = integrator_stage(opt);
  if (!integ) integ = MJB_STAGE_EULER;  // forward.py:1063: RK4 defaults to Euler in step2
  return mjb_run_stages(m, d, opt, MJB_STAGE_FWD_ACTUATION | MJB_STAGE_FWD_ACCELERATION | MJB_STAGE_SOLVE | integ, s);
}

// one entry point per stage function of the reference (f(m, d) in mujoco_warp/__init__.py:25-112)
#define MJB_STAGE_ENTRY(name, bits) \
  extern "C" int mjb_##name(const MjbModel* m, MjbData* d, const MjbOption* opt, void* s) { return mjb_run_stages(m, d, opt, (bits), s); }
MJB_STAGE_ENTRY(kinematics, MJB_STAGE_KINEMATICS)
MJB_STAGE_ENTRY(com_pos, MJB_STAGE_COM_POS)
MJB_STAGE_ENTRY(crb, MJB_STAGE_CRB)
MJB_STAGE_ENTRY(factor_m, MJB_STAGE_FACTOR_M)
MJB_STAGE_ENTRY(collision, MJB_STAGE_COLLISION)
MJB_STAGE_ENTRY(make_constraint, MJB_STAGE_MAKE_CONSTRAINT)
MJB_STAGE_ENTRY(transmission, MJB_STAGE_TRANSMISSION)
MJB_STAGE_ENTRY(fwd_velocity, MJB_STAGE_FWD_VELOCITY)
MJB_STAGE_ENTRY(fwd_actuation, MJB_STAGE_FWD_ACTUATION)
MJB_STAGE_ENTRY(com_vel, MJB_STAGE_COM_VEL)
MJB_STAGE_ENTRY(passive, MJB_STAGE_PASSIVE)
MJB_STAGE_ENTRY(rne, MJB_STAGE_RNE)
MJB_STAGE_ENTRY(energy_pos, MJB_STAGE_ENERGY_POS)
MJB_STAGE_ENTRY(energy_vel, MJB_STAGE_ENERGY_VEL)
MJB_STAGE_ENTRY(solve, MJB_STAGE_SOLVE)
MJB_STAGE_ENTRY(euler, MJB_STAGE_EULER)
MJB_STAGE_ENTRY(implicit, MJB_STAGE_IMPLICIT)
#undef MJB_STAGE_ENTRY
extern "C" int mjb_fwd_position(const MjbModel* m, MjbData* d, const MjbOption* opt, int factorize, void* s) {
  return mjb_run_stages(m, d, opt, MJB_STAGES_FWD_POSITION | (factorize ? MJB_STAGE_FACTOR_M : 0), s);
}
extern "C" int mjb_fwd_acceleration(const MjbModel* m, MjbData* d, const MjbOption* opt, int factorize, void* s) {
  return mjb_run_stages(m, d, opt, factorize ? MJB_STAGE_FWD_ACC_FACTOR : MJB_STAGE_FWD_ACCELERATION, s);
}
// Reads the device status word (overflow / NaN bits ORed over worlds).  Synchronises the stream.
extern "C" int mjb_status(MjbData* d, uint32_t* host_flags, void* s) {
  if (!d || !host_flags) return fail(-1, "null handle");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  const int n = d->dsz.nworld;
  std::vector<int> h(n > 0 ? n : 1, 0);
  if (n > 0) {
    CUDA_OK(cudaMemcpyAsync(h.data(), d->d.status, sizeof(int) * n, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
  }
  uint32_t f = 0;
  for (int i = 0; i < n; i++) f |= (uint32_t)h[i];
  *host_flags = f;
  return 0;
}

extern "C" int mjb_solve_m(const MjbModel* m, MjbData* d, float* x, const float* y, void* s) {
  if (!m || !d) return fail(-1, "null handle");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  KParams K;
  MjbOption o;
  memset(&o, 0, sizeof(o));
  fill_params(K, m, d, &o, 0);
  K.x = x;
  K.y = y;
  const int wpb = warps_per_block(4);
  const size_t smem = (size_t)wpb * d->scratch_floats * sizeof(float);
  CUDA_OK(mjb_util_solve_m_smem(smem));
  int grid = (d->dsz.nworld + wpb - 1) / wpb;
  if (grid > d->nsm * 8) grid = d->nsm * 8;
  return timed_launch(stream, [&] { mjb_util_solve_m(K, grid, 32 * wpb, smem, stream); });
}

extern "C" int mjb_mul_m(const MjbModel* m, MjbData* d, float* res, const float* vec, void* s) {
  if (!m || !d) return fail(-1, "null handle");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  KParams K;
  MjbOption o;
  memset(&o, 0, sizeof(o));
  fill_params(K, m, d, &o, 0);
  K.x = res;
  K.y = vec;
  const int n = d->dsz.nworld * m->ms.nv;
  if (n == 0) return 0;
  return timed_launch(stream, [&] { mjb_util_mul_m(K, (n + 255) / 256, stream); });
}

extern "C" int mjb_xfrc_accumulate(const MjbModel* m, MjbData* d, float* qfrc, void* s) {
  if (!m || !d || !qfrc) return fail(-1, "null handle");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  KParams K;
  MjbOption o;
  memset(&o, 0, sizeof(o));
  fill_params(K, m, d, &o, 0);
  K.x = qfrc;
  const int n = d->dsz.nworld * m->ms.nv;
  if (n == 0) return 0;
  return timed_launch(stream, [&] { mjb_util_xfrc_accumulate(K, (n + 255) / 256, stream); });
}

extern "C" int mjb_deriv_smooth_vel(const MjbModel* m, MjbData* d, const MjbOption* opt, float* out, int ld_out, void* s) {
  if (!m || !d || !opt || !out) return fail(-1, "null handle");
  if (ld_out < m->ms.nv) return fail(-2, "deriv_smooth_vel: out must hold nv x nv entries per world");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  KParams K;
  fill_params(K, m, d, opt, 0);
  K.x = out;
  const int n = d->dsz.nworld * m->ms.nv * m->ms.nv;
  if (n == 0) return 0;
  return timed_launch(stream, [&] { mjb_util_deriv_smooth_vel(K, (n + 255) / 256, ld_out, stream); });
}

extern "C" int mjb_ctrl_noise(const MjbModel* m, MjbData* d, const MjbOption* opt, const float* ctrl_center, int step, float ctrlnoisestd,
                              float ctrlnoiserate, void* s) {
  if (!m || !d || !opt) return fail(-1, "null handle");
  cudaStream_t stream = static_cast<cudaStream_t>(s);
  KParams K;
  fill_params(K, m, d, opt, 0);
  const int n = d->dsz.nworld * m->ms.nu;
  if (n == 0) return 0;
  // not part of the step: neither timed nor counted
  return timed_launch(stream, [&] { mjb_util_ctrl_noise(K, (n + 255) / 256, ctrl_center, step, ctrlnoisestd, ctrlnoiserate, stream); }, -1, false);
}

extern "C" long long mjb_launch_count(void) { return g_launches.load(); }
extern "C" void mjb_timing_enable(int on) { g_timing = on; }
extern "C" void mjb_timing_reset(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (auto& p : g_events) {
    cudaEventDestroy(p.first);
    cudaEventDestroy(p.second);
  }
  g_events.clear();
}
extern "C" long long mjb_timing_launches(void) { return (long long)g_events.size(); }
extern "C" double mjb_timing_ms(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  double total = 0;
  for (auto& p : g_events) {
    cudaEventSynchronize(p.second);
    float ms = 0;
    if (cudaEventElapsedTime(&ms, p.first, p.second) == cudaSuccess) total += ms;
  }
  return total;
}
extern "C" double mjb_timing_group_ms(int group) {
  std::lock_guard<std::mutex> lk(g_mu);
  double total = 0;
  for (auto& p : g_events) {
    if (p.group != group) continue;
    cudaEventSynchronize(p.second);
    float ms = 0;
    if (cudaEventElapsedTime(&ms, p.first, p.second) == cudaSuccess) total += ms;
  }
  return total;
}
extern "C" int mjb_group_count(void) { return G_COUNT; }
extern "C" const char* mjb_group_name(int group) {
  static const char* names[G_COUNT] = {"smooth_position", "collision", "make_constraint", "smooth_velocity", "solve", "euler"};
  return (group >= 0 && group < G_COUNT) ? names[group] : nullptr;
}
extern "C" const char* mjb_last_error(void) { return g_err.c_str(); }
extern "C" const char* mjb_version(void) { return "mjb200 0.1 (sm_100a)"; }
