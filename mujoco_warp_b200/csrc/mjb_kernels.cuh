// mjb_kernels.cuh -- the persistent one-warp-per-world pipeline kernel and small utility kernels.
#pragma once

// A translation unit that instantiates ONE stage group defines MJB_UNIT_GROUP (0 smooth_position, 1 collision, 2 make_constraint,
// 3 smooth_velocity, 4 solve, 5 euler) before including this header and then sees only the stage headers that group needs, so an
// edit of one stage recompiles one unit.  Host code (mjb_api.cu) leaves it undefined and sees everything.
#ifdef MJB_UNIT_GROUP
#define MJB_UNIT_IS(g) (MJB_UNIT_GROUP == (g))
#else
#define MJB_UNIT_IS(g) 1
#endif
#if MJB_UNIT_IS(1)
#include "mjb_collision.cuh"
#endif
#if MJB_UNIT_IS(2)
#include "mjb_constraint.cuh"
#endif
#if MJB_UNIT_IS(0) || MJB_UNIT_IS(3) || MJB_UNIT_IS(5) || MJB_UNIT_IS(4)
#include "mjb_smooth.cuh"
#endif
#if MJB_UNIT_IS(4)
#include "mjb_solver.cuh"
#endif
#include "mjb_types.cuh"
#ifndef MJB_LW
#define MJB_LW 32
#endif

// The step is issued as one launch per stage GROUP.  Every group is its own kernel instantiation:
// compact code (the instruction cache holds it while warps drift apart), its own register budget
// and its own shared-memory scratch.  Intermediate results travel between groups through L2/HBM.
#ifndef MJB_SOL_MIN_CTAS
#define MJB_SOL_MIN_CTAS 5
#endif
#ifndef MJB_SOL_SYNC
#define MJB_SOL_SYNC 0
#endif
#ifndef MJB_SOL_THREADS
#define MJB_SOL_THREADS 128
#endif
#ifndef MJB_CON_MIN_CTAS
#define MJB_CON_MIN_CTAS 3
#endif
#ifndef MJB_COL_MIN_CTAS
#define MJB_COL_MIN_CTAS 2
#endif
enum StageGroup { G_POS = 0, G_COL, G_CON, G_VEL, G_SOL, G_EUL, G_COUNT };

__host__ __device__ inline uint32_t group_mask(int g) {
  switch (g) {
    case G_POS: return MJB_STAGE_KINEMATICS | MJB_STAGE_COM_POS | MJB_STAGE_CRB | MJB_STAGE_FACTOR_M | MJB_STAGE_TRANSMISSION;
    case G_COL: return MJB_STAGE_COLLISION;
    case G_CON: return MJB_STAGE_MAKE_CONSTRAINT;
    case G_VEL:
      return MJB_STAGE_ZERO_SENSOR | MJB_STAGE_FWD_VELOCITY | MJB_STAGE_FWD_ACTUATION | MJB_STAGE_FWD_ACCELERATION | MJB_STAGE_FWD_ACC_FACTOR |
             MJB_STAGE_COM_VEL | MJB_STAGE_PASSIVE | MJB_STAGE_RNE | MJB_STAGE_ENERGY_POS | MJB_STAGE_ENERGY_VEL;
    case G_SOL: return MJB_STAGE_SOLVE;
    case G_EUL: return MJB_STAGE_EULER | MJB_STAGE_IMPLICIT;
  }
  return 0;
}

#ifndef MJB_UNIT_GROUP
// per-warp shared-memory scratch (floats) of a group
__host__ __device__ inline int group_scratch_floats(int g, int nv, int nv_pad, int njmax_pad, int ngeom, int nxn, int variant = 0, int ngroup = 0) {
  int n = 0;
  switch (g) {
    case G_POS:
    case G_VEL:
    case G_EUL: n = chol_scratch_floats(nv, nv_pad); break;
    case G_COL: n = (variant & COLF_SAP) ? collision_scratch_words(ngeom, nxn, ngroup, (variant & COLF_CONVEX) != 0) : 72; break;
    case G_CON: n = 0; break;
    case G_SOL: n = solver_scratch_floats(nv, nv_pad, njmax_pad); break;
  }
  return (n + 3) & ~3;
}
#endif
__host__ __device__ inline bool group_uses_blob(int g) { return g != G_SOL; }
// lanes per world of a group's kernel (worlds per warp = 32 / lanes)
// lanes per world of a group kernel; variant 1 of smooth_velocity / euler is the one-warp-per-world build (nv > 32, or few worlds),
// variant 1 of smooth_position the four-worlds-per-warp build (small models)
__host__ __device__ constexpr int group_lw(int g, int v = 0) {
  return g == G_POS ? (v ? 8 : MJB_POS_LW) : g == G_VEL ? (v ? 32 : MJB_VEL_LW) : g == G_EUL ? (v ? 32 : MJB_EUL_LW) : 32;
}

// Stages the model blob into shared memory with one TMA bulk copy (cp.async.bulk, completion on
// an mbarrier).  All threads of the CTA wait on the barrier's phase 0.
__device__ __forceinline__ void stage_model_blob(unsigned char* dst, const unsigned char* src, int nbytes, uint64_t* bar) {
  const uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(bar);
  const uint32_t dst_s = (uint32_t)__cvta_generic_to_shared(dst);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"((uint32_t)nbytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s), "l"(src),
                 "r"((uint32_t)nbytes), "r"(bar_s)
                 : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar_s), "r"(0u)
        : "memory");
  }
}

// CTA width and minimum resident CTAs per SM of a group (register budget = 65536 / (threads x CTAs))
__host__ __device__ constexpr int group_threads(int g) { return g == G_SOL ? MJB_SOL_THREADS : 256; }
__host__ __device__ constexpr int group_min_ctas(int g) {
  // caps the register allocation (65536 / (threads x CTAs)): without it ptxas spends registers on unrolling
  // and the smooth kernels lose half of their occupancy
  return g == G_SOL ? MJB_SOL_MIN_CTAS : g == G_COL ? MJB_COL_MIN_CTAS : g == G_POS ? 4 : g == G_CON ? MJB_CON_MIN_CTAS : 3;
}

// V: compile-time variant of the group (collision: CollisionFeature bits; 0 elsewhere)
template <int G, bool BLOB_IN_SMEM, int V = 0>
__global__ void __launch_bounds__(group_threads(G), group_min_ctas(G)) mjb_group_kernel(const __grid_constant__ KParams K) {
  extern __shared__ __align__(16) unsigned char smem[];
  const unsigned char* mb;
  float* scr_base;
  // let the next group's grid start launching as soon as SM resources free up (PDL)
  asm volatile("griddepcontrol.launch_dependents;");
  if (BLOB_IN_SMEM) {
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + K.blob_bytes);
    stage_model_blob(smem, K.blob, K.blob_bytes, bar);  // the model is constant: no dependency on the previous grid
    mb = smem;
    scr_base = reinterpret_cast<float*>(smem + K.blob_bytes + 16);
  } else {
    mb = K.blob;
    scr_base = reinterpret_cast<float*>(smem);
  }
  // Data written by the previous group is visible only after the previous grid has completed
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // broadcast from lane 0 so that the compiler can treat the warp index (and every loop / branch
  // that depends on it) as warp-uniform
  // A warp carries WPW = 32 / LW worlds (LW lanes each); `lane` is the lane inside the world's group.
  constexpr int LW = group_lw(G, V), WPW = 32 / LW;
  static_assert(LW == 32 || LW == MJB_LW, "a sub-warp group kernel must be compiled with MJB_LW = its lane count");
  const int lane32 = threadIdx.x & 31, lane = lane32 % LW, grp = lane32 / LW;
  const int warp = __shfl_sync(FULL, threadIdx.x >> 5, 0), wpb = blockDim.x >> 5;
  float* scr = scr_base + (size_t)(warp * WPW + grp) * K.scratch_floats;
  const uint32_t st = K.stages;
#if MJB_SOL_SYNC && MJB_UNIT_IS(4)
  if constexpr (G == G_SOL && (V & 1) == 0) {  // the pass-synchronous experiment has a pyramidal build only: elliptic models take the default path
    s_solve_cta(K, lane, warp, wpb, scr);
    return;
  }
#endif
  // Worlds are handed out dynamically: every warp starts on world (CTA, warp) and fetches its next world from a per-launch counter.
  // Per-world time varies (contacts, solver iterations), and nworld is rarely a multiple of the resident warps: with a static
  // stride the slowest warp decides (8192 worlds on 2960 resident solver warps = 2 or 3 worlds per warp).
  // (Units of WPW consecutive worlds; the groups of a last, partial unit repeat its last world -- same inputs, same stores.)
  const int nresident = gridDim.x * wpb, nunit = (K.dsz.nworld + WPW - 1) / WPW;
  for (int u = blockIdx.x * wpb + warp; u < nunit;
       u = K.queue ? uniform_i(lane32 == 0 ? nresident + atomicAdd(K.queue + G, 1) : 0) : u + nresident) {
    const int w = min(u * WPW + grp, K.dsz.nworld - 1);
#if MJB_UNIT_IS(0)
    if constexpr (G == G_POS) {
      if (st & MJB_STAGE_KINEMATICS) s_kinematics(K, mb, w, lane);
      if (st & MJB_STAGE_COM_POS) s_com_pos(K, mb, w, lane);
      if (st & MJB_STAGE_CRB) s_crb(K, mb, w, lane);
      if (st & MJB_STAGE_FACTOR_M) s_factor_m(K, mb, w, lane, scr);
      if (st & MJB_STAGE_TRANSMISSION) s_transmission(K, mb, w, lane);
    }
#endif
#if MJB_UNIT_IS(1)
    if constexpr (G == G_COL) {
      if (K.o.run_collision_detection) s_collision<V>(K, mb, w, lane, reinterpret_cast<int*>(scr));
    }
#endif
#if MJB_UNIT_IS(2)
    if constexpr (G == G_CON) {
      s_make_constraint<(V & 1) != 0>(K, mb, w, lane);
    }
#endif
#if MJB_UNIT_IS(3)
    if constexpr (G == G_VEL) {
      const bool energy = (st & MJB_STAGE_ZERO_SENSOR) && (K.o.enableflags & 2);  // EnableBit.ENERGY (forward.py:975)
      if (st & MJB_STAGE_ZERO_SENSOR) {
        if (energy) s_energy_pos(K, mb, w, lane);
        else if (lane < 2) DW(energy)[lane] = 0.f;
      }
      if (st & MJB_STAGE_ENERGY_POS) s_energy_pos(K, mb, w, lane);
      const int parts = (st & MJB_STAGE_FWD_VELOCITY) ? FV_ALL
                                                      : ((st & MJB_STAGE_COM_VEL) ? FV_COMVEL : 0) | ((st & MJB_STAGE_PASSIVE) ? FV_PASSIVE : 0) |
                                                            ((st & MJB_STAGE_RNE) ? FV_RNE : 0);
      if (parts) s_fwd_velocity(K, mb, w, lane, parts);
      if ((energy && (st & MJB_STAGE_FWD_VELOCITY)) || (st & MJB_STAGE_ENERGY_VEL)) s_energy_vel(K, w, lane);
      if (st & MJB_STAGE_FWD_ACTUATION) s_fwd_actuation(K, mb, w, lane);
      if (st & (MJB_STAGE_FWD_ACCELERATION | MJB_STAGE_FWD_ACC_FACTOR)) s_fwd_acceleration(K, mb, w, lane, scr, (st & MJB_STAGE_FWD_ACC_FACTOR) != 0);
    }
#endif
#if MJB_UNIT_IS(4)
    if constexpr (G == G_SOL) {
      s_solve<(V & 1) != 0>(K, mb, w, lane, scr);
    }
#endif
#if MJB_UNIT_IS(5)
    if constexpr (G == G_EUL) {
      s_integrate(K, mb, w, lane, scr, (st & MJB_STAGE_IMPLICIT) != 0);
    }
#endif
  }
}
