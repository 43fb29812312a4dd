/* mjb200.h -- C ABI of libmjb200.so, the B200 (sm_100a) implementation of MuJoCo Warp's batched
 * physics step.
 *
 * The reference exposes this path only as Python functions over Warp arrays
 * (reference mujoco_warp/__init__.py:25-112); there is no FFI in the reference to bind to, so
 * each entry point below names the Python function it replaces.  All functions are
 * asynchronous on the given CUDA stream, never synchronise the device (mjb_status excepted) and,
 * once mjb_prepare() has run, allocate nothing and are CUDA-graph capturable.  Pointers are *borrowed* device pointers owned by the caller
 * (PyTorch in the bundled host layer).  No torch / C++ types cross this boundary.
 *
 * Return value: 0 on success, negative on error (see mjb_last_error()).
 */
#ifndef MJB200_H_
#define MJB200_H_

#include <stddef.h>
#include <stdint.h>

#include "mjb_fields.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct MjbModel MjbModel; /* immutable device model (reference types.Model, types.py:834) */
typedef struct MjbData MjbData;   /* bound nworld-batched data (reference types.Data, types.py:1703) */

/* Options are passed by value on every call because reference users mutate m.opt.* freely
 * between calls (reference io.py:2498 override_model, broadphase_test.py:123). */
typedef struct MjbOption {
#define X(name) int name;
  MJB_OPT_INT_FIELDS(X)
#undef X
#define X(name) float name;
  MJB_OPT_REAL_FIELDS(X)
#undef X
} MjbOption;

/* Field-table introspection, used by the host layer to build pointer tables in the right order.
 * kind: 0 model sizes, 1 model int, 2 model real, 3 opt int, 4 opt real, 5 data sizes,
 *       6 data int, 7 data real.  Returns NULL past the end. */
const char* mjb_field_name(int kind, int index);
int mjb_field_count(int kind);

/* Replaces put_model's device upload (reference io.py:641-645).  `blob` is a device buffer of
 * `nbytes` bytes holding every model array back to back; int_offsets / real_offsets give the
 * byte offset (16-byte aligned) of each array in MJB_MODEL_INT_FIELDS / MJB_MODEL_REAL_FIELDS
 * order; sizes follow MJB_MODEL_SIZES order.  The blob must outlive the model. */
int mjb_model_create(const int* sizes, int nsizes, const void* blob, size_t nbytes,
                     const int* int_offsets, int nint, const int* real_offsets, int nreal,
                     MjbModel** out);
void mjb_model_destroy(MjbModel* m);

/* Replaces make_data/put_data's device allocation (reference io.py:859, :1016): binds borrowed
 * device pointers.  int_ptrs / real_ptrs follow MJB_DATA_INT_FIELDS / MJB_DATA_REAL_FIELDS
 * order; sizes follow MJB_DATA_SIZES order. */
int mjb_data_bind(const MjbModel* m, const int* sizes, int nsizes, int* const* int_ptrs, int nint,
                  float* const* real_ptrs, int nreal, MjbData** out);
void mjb_data_destroy(MjbData* d);

/* Runs the pipeline stages selected by `stages` (MJB_STAGE_* bits) in pipeline order: one kernel
 * launch per stage group touched (smooth_position, collision, make_constraint, smooth_velocity,
 * solve, euler -- six for a whole step), one warp per world.  Every per-stage reference function f(m, d)
 * (kinematics, com_pos, crb, factor_m, collision, make_constraint, transmission,
 * fwd_velocity, fwd_actuation, fwd_acceleration, solve, euler, implicit; reference
 * __init__.py:25-112) maps to one bit. */
int mjb_run_stages(const MjbModel* m, MjbData* d, const MjbOption* opt, uint32_t stages,
                   void* cuda_stream);

/* mjw.step (reference forward.py:1004) / mjw.forward (:973) / step1 (:1022) / step2 (:1050). */
int mjb_step(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_forward(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_step1(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_step2(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);

/* forward.rungekutta4 (reference forward.py:458): the RK4 integrator, called after `forward` like the reference's; mjb_step runs
 * forward + this when opt->integrator is RK4.  The six temporaries the reference allocates on every call (forward.py:465-475) live in
 * a scratch buffer of mjb_rk_scratch_floats() floats owned by the MjbData handle: cudaMalloc'ed by mjb_prepare(), or by the first call. */
int mjb_rungekutta4(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
long long mjb_rk_scratch_floats(const MjbModel* m, const MjbData* d);
/* No reference counterpart (the reference JIT-compiles and sizes its launches lazily inside the first call, warp_util.py cache_kernel):
 * does everything a first call would do lazily for the kernel variants `opt` selects -- function attributes, occupancy queries, the
 * Runge-Kutta scratch -- so that the next step / forward / stage call can be captured into a CUDA graph. */
int mjb_prepare(const MjbModel* m, MjbData* d, const MjbOption* opt);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same integrator as mjb_rungekutta4 -- same inputs, same
 * outputs, called after `forward` -- but its scratch is a file-scope `static __device__` array of MJB_RK_STATIC_MAX_FLOATS floats
 * (4 MiB; every GPU that loads the library holds it until the process exits) instead of the cudaMalloc'ed buffer of the MjbData
 * handle.  It calls no cudaMalloc / cudaFree / allocating library routine.  Returns -6 with a clear mjb_last_error() when
 * nworld x round_up4(nq + 3 nv + 2 na) exceeds the maximum (humanoid: 112 floats per world, i.e. up to 9362 worlds).
 * Every call on a device shares the one array: it keeps the previous call's contents (stage 0 overwrites every word a call
 * reads, so nothing needs clearing), and two calls that overlap on the same GPU (different streams, or different MjbData) RACE --
 * serialise them.  One resolve of the array's device address per device (cudaGetSymbolAddress) happens on the first call with
 * that device current; run the kernels' own first-call set-up with mjb_prepare() before capturing. */
#define MJB_RK_STATIC_MAX_FLOATS (1 << 20)
int mjb_rungekutta4_static(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
long long mjb_rk_static_max_floats(void);

/* smooth.solve_m (reference smooth.py:2848): x = M^-1 y using d.qLD; x, y are (nworld, nv). */
int mjb_solve_m(const MjbModel* m, MjbData* d, float* x, const float* y, void* cuda_stream);
/* support.xfrc_accumulate (reference support.py:229): qfrc (nworld, nv) += J^T d.xfrc_applied (needs kinematics + com_pos results). */
int mjb_xfrc_accumulate(const MjbModel* m, MjbData* d, float* qfrc, void* cuda_stream);
/* derivative.deriv_smooth_vel (reference derivative.py:321), dense: out (nworld, ld_out, ld_out), ld_out >= nv, receives
 * qM - dt * qDeriv in its leading nv x nv entries (actuator velocity derivatives and joint damping, as the flags in opt allow). */
int mjb_deriv_smooth_vel(const MjbModel* m, MjbData* d, const MjbOption* opt, float* out, int ld_out, void* cuda_stream);
/* support.mul_m (reference support.py:133): res = qM @ vec; (nworld, nv). */
int mjb_mul_m(const MjbModel* m, MjbData* d, float* res, const float* vec, void* cuda_stream);

/* One entry per stage function f(m, d) of the reference's public surface (mujoco_warp/__init__.py:25-112):
 * smooth.py:358 kinematics, :602 com_pos, :889 crb, :1104 factor_m, :2605 transmission;
 * collision_driver.py:755 collision; constraint.py:2210 make_constraint; forward.py:514 fwd_position,
 * :593 fwd_velocity, :837 fwd_actuation, :950 fwd_acceleration, :327 euler; solver.py:3297 solve. */
int mjb_kinematics(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_com_pos(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_crb(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_factor_m(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_collision(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_make_constraint(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_transmission(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_fwd_position(const MjbModel* m, MjbData* d, const MjbOption* opt, int factorize, void* cuda_stream);
int mjb_fwd_velocity(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_fwd_actuation(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
/* sub-functions of fwd_velocity / forward that the reference exports on their own: smooth.py:2015 com_vel (cvel, cdof_dot),
 * passive.py:729 passive (qfrc_spring / damper / gravcomp / passive), smooth.py:1259 rne (cacc, cfrc_int, qfrc_bias),
 * sensor.py:2854 energy_pos / :2922 energy_vel (d.energy[:, 0] / [:, 1]) */
int mjb_com_vel(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_passive(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_rne(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_energy_pos(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_energy_vel(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_fwd_acceleration(const MjbModel* m, MjbData* d, const MjbOption* opt, int factorize, void* cuda_stream);
int mjb_solve(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
int mjb_euler(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
/* forward.py:495 implicit (integrator IMPLICITFAST): (qM - dt qDeriv) solve + _advance */
int mjb_implicit(const MjbModel* m, MjbData* d, const MjbOption* opt, void* cuda_stream);
/* Replaces the reference's device printf on overflow (forward.py:194-210, collision_core.py:213):
 * ORs the per-world status words (MJB_STATUS_* bits) into *host_flags.  Synchronises the stream;
 * the step entry points never do. */
int mjb_status(MjbData* d, uint32_t* host_flags, void* cuda_stream);

/* benchmark.ctrl_noise (reference benchmark.py:42): Ornstein-Uhlenbeck control noise with
 * Halton-sequence noise; ctrl_center may be NULL. */
int mjb_ctrl_noise(const MjbModel* m, MjbData* d, const MjbOption* opt, const float* ctrl_center,
                   int step, float ctrlnoisestd, float ctrlnoiserate, void* cuda_stream);

/* ---- the callers either side of `step` (RL loops): state packing, masked reset, contact forces ----
 * State signature bits = the reference's `State` enum (types.py:598-637 = MuJoCo mjtState):
 * time, qpos, qvel, act, qacc_warmstart, ctrl, qfrc_applied, xfrc_applied, eq_active, mocap_pos, mocap_quat. */
#define MJB_STATE_TIME (1 << 0)
#define MJB_STATE_QPOS (1 << 1)
#define MJB_STATE_QVEL (1 << 2)
#define MJB_STATE_ACT (1 << 3)
#define MJB_STATE_WARMSTART (1 << 4)
#define MJB_STATE_CTRL (1 << 5)
#define MJB_STATE_QFRC_APPLIED (1 << 6)
#define MJB_STATE_XFRC_APPLIED (1 << 7)
#define MJB_STATE_EQ_ACTIVE (1 << 8)
#define MJB_STATE_MOCAP_POS (1 << 9)
#define MJB_STATE_MOCAP_QUAT (1 << 10)
#define MJB_STATE_USERDATA (1 << 11) /* accepted, carries no data (unsupported in the reference too, types.py:638) */
#define MJB_STATE_PLUGIN (1 << 12)   /* accepted, carries no data */
#define MJB_NSTATE 13
/* Floats per world of the state selected by `sig` (mj_stateSize), or < 0 on an invalid signature. */
int mjb_state_size(const MjbModel* m, const MjbData* d, int sig);
/* support.get_state / set_state (reference support.py:572 / :711): packs / unpacks the selected components between Data and
 * `state` (device, nworld rows of `ld` >= state size floats).  `active` is a device array of nworld bytes (non-zero = copy
 * that world) or NULL for every world. */
int mjb_get_state(const MjbModel* m, MjbData* d, float* state, int ld, int sig, const uint8_t* active, void* cuda_stream);
int mjb_set_state(const MjbModel* m, MjbData* d, const float* state, int ld, int sig, const uint8_t* active, void* cuda_stream);
/* io.reset_data (reference io.py:1458): default state for the worlds whose byte in `reset` (device, nworld bytes; NULL = all)
 * is non-zero, their contacts cleared in place.  No host synchronisation, graph capturable. */
int mjb_reset_data(const MjbModel* m, MjbData* d, const uint8_t* reset, void* cuda_stream);
/* support.contact_force (reference support.py:354): 6D force:torque of `n` contacts (device ids) in the contact frame, or
 * rotated into the world frame; `cone` = Option.cone; `force` = n x 6 floats (device).  Rows of ids >= nacon are not written. */
int mjb_contact_force(const MjbModel* m, MjbData* d, int cone, const int* contact_ids, int n, int to_world_frame, float* force,
                      void* cuda_stream);

/* Number of kernel launches issued by this library since load (for bench.py's gpu_launches). */
long long mjb_launch_count(void);
/* Duration bookkeeping for bench.py: records CUDA events around every stage launch when enabled
 * and returns the accumulated milliseconds of the main kernel (synchronises the events). */
void mjb_timing_enable(int on);
double mjb_timing_ms(void);
long long mjb_timing_launches(void);
void mjb_timing_reset(void);
/* The step is issued as one launch per stage group (smooth_position, collision, make_constraint,
 * smooth_velocity, solve, euler); these report the time of one group's launches since the reset. */
double mjb_timing_group_ms(int group);
int mjb_group_count(void);
const char* mjb_group_name(int group);

const char* mjb_last_error(void);
const char* mjb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MJB200_H_ */
