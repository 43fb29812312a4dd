// mjb_solver.cuh -- solver.solve (reference solver.py:3297): primal Newton / CG with the exact
// iterative bracketing linesearch, pyramidal cones, dense Jacobian.  One warp per world; the whole
// iteration loop runs inside the kernel with per-world convergence.
//
// The reference drives ~12-15 kernel launches per iteration from a CUDA conditional-graph while
// node and polls a global `nsolving` counter (solver.py:3188-3343).  Here:
//   * dof vectors (qacc, Ma, grad, Mgrad, search, mv) and per-row vectors (Jaref, jv, D, force) live
//     in the warp's shared-memory scratch;
//   * H = M + J^T D J (solver.py:2368) is accumulated in a 4x8 register tile per lane over 32x32
//     blocks, only for rows whose QUADRATIC state changed (incremental update, solver.py:2055),
//     then added to the shared-memory copy of H;
//   * Cholesky + triangular solves run in shared memory (lane = row, shuffles for broadcasts);
//   * the linesearch evaluates lo/hi/mid candidates with lanes over rows and 9 warp reductions.
#pragma once

#include "mjb_smooth.cuh"
#if MJB_SOL_COMPACT
#define MJB_ROLL _Pragma("unroll 1")
#else
#define MJB_ROLL
#endif
#include "mjb_types.cuh"

struct Pt3 {
  float c, g, h;
};
__device__ __forceinline__ Pt3 operator+(Pt3 a, Pt3 b) { return Pt3{a.c + b.c, a.g + b.g, a.h + b.h}; }
// one out-of-line copy of the three interleaved reductions (the line search calls it in many places)
#if MJB_SOL_COMPACT
static __device__ __noinline__
#else
__device__ __forceinline__
#endif
Pt3 warp_sum(Pt3 p) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    p.c += __shfl_xor_sync(FULL, p.c, o);
    p.g += __shfl_xor_sync(FULL, p.g, o);
    p.h += __shfl_xor_sync(FULL, p.h, o);
  }
  return p;
}

// solver.py:571-598 (_compute_efc_eval_pt_pyramidal)
__device__ __forceinline__ Pt3 efc_eval_pt(int r, float alpha, int ne, int nf, float D, float floss, float jaref, float jv) {
  const float x = jaref + alpha * jv;
  const float jvD = jv * D;
  const Pt3 quad = Pt3{0.5f * D * x * x, jvD * x, jv * jvD};
  if (r >= ne + nf) return x < 0.f ? quad : Pt3{0.f, 0.f, 0.f};
  if (r >= ne) {  // solver.py:224
    const float rf = safe_div(floss, D);
    if (-rf < x && x < rf) return quad;
    if (x <= -rf) return Pt3{floss * (-0.5f * rf - x), -floss * jv, 0.f};
    return Pt3{floss * (-0.5f * rf + x), floss * jv, 0.f};
  }
  return quad;
}
// solver.py:198
__device__ __forceinline__ Pt3 eval_pt_quad(float q0, float q1, float q2, float alpha) {
  const float aq2 = alpha * q2;
  return Pt3{alpha * aq2 + alpha * q1 + q0, 2.0f * aq2 + q1, 2.0f * q2};
}
__device__ __forceinline__ bool in_bracket(Pt3 x, Pt3 y) { return (x.g < y.g && y.g < 0.f) || (x.g > y.g && y.g > 0.f); }

// res[row] = A[row, :n] . vec for rows handled by this lane (A row-major, 16-byte aligned rows)
__device__ __forceinline__ float row_dot(const float* __restrict__ row, const float* __restrict__ vec, int n4) {
  const float4* a = reinterpret_cast<const float4*>(row);
  const float4* v = reinterpret_cast<const float4*>(vec);
  float s = 0.f;
  if (n4 <= 8) {  // nv_pad <= 32: issue every global load of the row before the first use
    float4 x[8];
#pragma unroll
    for (int k = 0; k < 8; k++) x[k] = k < n4 ? a[k] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      if (k < n4) {
        const float4 y = v[k];
        s += x[k].x * y.x + x[k].y * y.y + x[k].z * y.z + x[k].w * y.w;
      }
    }
    return s;
  }
  for (int k = 0; k < n4; k++) {
    const float4 x = a[k], y = v[k];
    s += x.x * y.x + x.y * y.y + x.z * y.z + x.w * y.w;
  }
  return s;
}

// out[r] = A[r, :] . vec - sub[r] for r < nrows (A row-major in global memory, rows of 4*n4 floats, 16-byte
// aligned; vec in shared memory).  One out-of-line copy serves J qacc, M qacc, M p and J p: the solver kernel
// is instruction-fetch bound when its hot code outgrows the 32 kB instruction cache.
static __device__ __noinline__ void mat_vec(const float* __restrict__ A, int n4, int nrows, const float* vec, float* out, const float* sub, int lane) {
#pragma unroll 1
  for (int r = lane; r < nrows; r += 32) {
    const float s = row_dot(A + (size_t)r * (4 * n4), vec, n4);
    out[r] = sub ? s - sub[r] : s;
  }
}
static __device__ __noinline__ float warp_sum_ol(float v) { return warp_sum(v); }  // out-of-line copy

// ---- elliptic friction cones (template flag ELL; compiled only into the elliptic variant of the kernel) ---------
// Rows of one contact are contiguous here (make_constraint emits them so): row r0 = normal, r0 + j = friction
// direction j.  `EllRow` describes the contact a row belongs to.
struct EllRow {
  int r0, dim;      // primary row, condim; dim = 0: not a (complete) elliptic contact row
  const float* fri;  // contact.friction[conid]
  float mu;          // friction[0] * impratio_invsqrt
};
__device__ __forceinline__ EllRow ell_row(const KParams& K, int w, int r, int nefc) {
  EllRow e{0, 0, nullptr, 0.f};
  if (DW(efc_type)[r] != CT_CONTACT_ELLIPTIC) return e;
  const int conid = DW(efc_id)[r];
  const int r0 = K.d.contact_efc_address[conid * K.ms.nmaxpyramid], dim = K.d.contact_dim[conid];
  if (r0 < 0 || r0 + dim > nefc) return e;  // rows cut off by njmax: the contact is skipped (solver.py:1897-1908)
  e.r0 = r0;
  e.dim = dim;
  e.fri = K.d.contact_friction + 5 * conid;
  e.mu = e.fri[0] * K.o.impratio_invsqrt;
  return e;
}
// solver.py:263-319 _eval_elliptic with the sums of :986-1050 folded in: cost, slope, curvature along the search
// direction of the contact whose primary row is e.r0.
__device__ __forceinline__ Pt3 ell_eval(const SolverScr& s, const EllRow& e, float alpha) {
  const float D0 = s.D[e.r0], ja0 = s.Jaref[e.r0], jv0 = s.jv[e.r0], mu = e.mu;
  float q0 = 0.5f * ja0 * ja0 * D0, q1 = jv0 * D0 * ja0, q2 = 0.5f * jv0 * jv0 * D0;
  float uu = 0.f, uv = 0.f, vv = 0.f;
  for (int j = 1; j < e.dim; j++) {
    const float ja = s.Jaref[e.r0 + j], jv = s.jv[e.r0 + j], dj = s.D[e.r0 + j], f = e.fri[j - 1];
    q0 += 0.5f * ja * dj * ja;
    q1 += jv * dj * ja;
    q2 += 0.5f * jv * dj * jv;
    const float uj = ja * f, vj = jv * f;
    uu += uj * uj;
    uv += uj * vj;
    vv += vj * vj;
  }
  const float N = ja0 * mu + alpha * jv0 * mu, Tsqr = uu + alpha * (2.0f * uv + alpha * vv);
  const Pt3 zero{0.f, 0.f, 0.f};
  if (Tsqr <= 0.f) return N < 0.f ? eval_pt_quad(q0, q1, q2, alpha) : zero;
  const float T = sqrtf(Tsqr);
  if (N >= mu * T) return zero;
  if (mu * N + T <= 0.f) return eval_pt_quad(q0, q1, q2, alpha);
  const float dm = D0 / (mu * mu * (1.0f + mu * mu));
  const float N1 = jv0 * mu, T1 = (uv + alpha * vv) / T, T2 = vv / T - (uv + alpha * vv) * T1 / (T * T);
  const float e0 = N - mu * T, e1 = N1 - mu * T1;
  return Pt3{0.5f * dm * e0 * e0, dm * e0 * e1, dm * (e1 * e1 + e0 * (-mu * T2))};
}
// one row's contribution to the line cost at alpha, any row type
template <bool ELL>
__device__ __forceinline__ Pt3 row_eval(const KParams& K, int w, const SolverScr& s, int r, int nefc, int ne, int nf, float fl, float alpha) {
  if (ELL && r >= ne + nf && DW(efc_type)[r] == CT_CONTACT_ELLIPTIC) {
    const EllRow e = ell_row(K, w, r, nefc);
    return (e.dim && e.r0 == r) ? ell_eval(s, e, alpha) : Pt3{0.f, 0.f, 0.f};  // only the primary row contributes (:625-628)
  }
  return efc_eval_pt(r, alpha, ne, nf, s.D[r], fl, s.Jaref[r], s.jv[r]);
}

// Cone Hessian (solver.py:2592 update_gradient_JTCJ_dense) added to L (the copy of H about to be factored), for
// every contact of the world in the middle zone.  With w_a = s_a J_a (s_0 = mu, s_j = friction_j), u_j = Jaref_j
// friction_j, n = Jaref_0 mu, t = |u|, the reference's condim x condim matrix is
//   dm [ p p^T + c1 q q^T + c2 sum_j w_j w_j^T ],  q = sum_j u_j w_j,  p = w_0 - (mu / t) q,
//   c1 = (mu / t^2) (n / t - mu),  c2 = mu (mu - n / t),
// i.e. rank-one terms; p and q are staged in s.mv / s.x (both free between the constraint update and the solve).
__device__ __forceinline__ void cone_hessian(const KParams& K, int w, int lane, const SolverScr& s, int nefc) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, ld = s.ld, ncon = DW(wcon_n)[0];
  const float* J = s.Jw;
  for (int k = 0; k < ncon; k++) {
    const int conid = DW(wcon)[k];
    const int dim = K.d.contact_dim[conid], r0 = K.d.contact_efc_address[conid * K.ms.nmaxpyramid];
    if (uniform_b(dim == 1 || r0 < 0 || r0 + dim > nefc)) continue;
    if (uniform_b(DW(efc_state)[r0] != ST_CONE)) continue;
    const float* fri = K.d.contact_friction + 5 * conid;
    const float mu = fri[0] * K.o.impratio_invsqrt, mu2 = mu * mu, dm = safe_div(s.D[r0], mu2 * (1.0f + mu2));
    if (uniform_b(dm == 0.f)) continue;
    const float n = s.Jaref[r0] * mu;
    float tt = 0.f;
    for (int j = 1; j < dim; j++) {
      const float uj = s.Jaref[r0 + j] * fri[j - 1];
      tt += uj * uj;
    }
    const float t = fmaxf(tt <= 0.f ? 0.f : sqrtf(tt), MJ_MINVAL), ttt = fmaxf(t * t * t, MJ_MINVAL);
    const float mu_t = safe_div(mu, t), c1 = mu * safe_div(n, ttt) - mu_t * mu_t, c2 = mu2 - mu * safe_div(n, t);
    for (int i = lane; i < nv; i += 32) {
      float q = 0.f;
      for (int j = 1; j < dim; j++) q += s.Jaref[r0 + j] * fri[j - 1] * fri[j - 1] * J[(size_t)(r0 + j) * np + i];
      s.x[i] = q;
      s.mv[i] = mu * J[(size_t)r0 * np + i] - mu_t * q;
    }
    __syncwarp();
    for (int e = lane; e < nv * nv; e += 32) {
      const int i = e / nv, j = e - i * nv;
      if (j > i) continue;
      float h = s.mv[i] * s.mv[j] + c1 * s.x[i] * s.x[j];
      for (int a = 1; a < dim; a++) h += c2 * fri[a - 1] * fri[a - 1] * J[(size_t)(r0 + a) * np + i] * J[(size_t)(r0 + a) * np + j];
      s.L[i * ld + j] += dm * h;
    }
    __syncwarp();
  }
}

// H(lower) += sum_r coef[r] * J[r]^T J[r], register tiles of 4 rows x 8 columns per lane over
// 32x32 blocks (solver.py:2368 JTDAJ and :2055 incremental update in one routine).
#ifndef MJB_SOL_HTRI
#define MJB_SOL_HTRI 1
#endif
#ifndef MJB_SOL_HROWS
#define MJB_SOL_HROWS 2  // measured on B200 (humanoid): 2 rows 196 us, 4 rows 199 us, 8 rows 218 us (registers)
#endif
#if MJB_SOL_HTRI
// Lower triangle only, in 4 x 4 tiles: tile t = bi (bi + 1) / 2 + bj (bj <= bi) belongs to lane t % 32, so a 27-dof model keeps 28
// lanes busy with 16 useful FMAs per constraint row each (the 32 x 32-block mapping below spends 32 FMAs per lane on 20 lanes, half
// of them above the diagonal).  Per element the rows are added in the same order as before: results are bit-identical.
__device__ __forceinline__ void h_accumulate(const KParams& K, int w, int lane, const SolverScr& s, int nefc) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, ld = s.ld;
  const float* J = s.Jw;
  const int nb = (nv + 3) >> 2, ntile = (nb * (nb + 1)) >> 1;
  MJB_ROLL
  for (int t0 = 0; t0 < ntile; t0 += 32) {
    const int t = t0 + lane;
    const bool need = t < ntile;
    int bi = (int)((sqrtf(8.f * (float)t + 1.f) - 1.f) * 0.5f);
    if (((bi + 1) * (bi + 2)) >> 1 <= t) bi++;
    if ((bi * (bi + 1)) >> 1 > t) bi--;
    const int r0 = 4 * bi, c0 = 4 * (t - ((bi * (bi + 1)) >> 1));
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    // MJB_SOL_HROWS rows per iteration: their 128-bit loads are issued before the FMAs (rows in order: the sums keep their order)
    MJB_ROLL
    for (int r = 0; r < nefc; r += MJB_SOL_HROWS) {
      float cf[MJB_SOL_HROWS];
      bool any = false;
#pragma unroll
      for (int k = 0; k < MJB_SOL_HROWS; k++) {
        cf[k] = (r + k < nefc) ? s.coef[r + k] : 0.f;
        any = any || cf[k] != 0.f;
      }
      if (!any) continue;  // warp-uniform
      if (need) {
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 ji[MJB_SOL_HROWS], jj[MJB_SOL_HROWS];
#pragma unroll
        for (int k = 0; k < MJB_SOL_HROWS; k++) {
          const float* Jr = J + (size_t)min(r + k, nefc - 1) * np;
          const bool u = cf[k] != 0.f;
          ji[k] = u ? *reinterpret_cast<const float4*>(Jr + r0) : z4;
          jj[k] = u ? *reinterpret_cast<const float4*>(Jr + c0) : z4;
        }
#pragma unroll
        for (int k = 0; k < MJB_SOL_HROWS; k++) {
          const float vi[4] = {ji[k].x * cf[k], ji[k].y * cf[k], ji[k].z * cf[k], ji[k].w * cf[k]};
          const float vj[4] = {jj[k].x, jj[k].y, jj[k].z, jj[k].w};
#pragma unroll
          for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] += vi[i] * vj[j];
        }
      }
    }
    if (need) {  // whole tiles with 128-bit accesses; in a diagonal tile the elements above the diagonal get their symmetric values
#pragma unroll
      for (int i = 0; i < 4; i++) {
        if (r0 + i < nv) {
          float4* h = reinterpret_cast<float4*>(s.H + (r0 + i) * ld + c0);
          float4 v = h[0];
          v.x += acc[i][0]; v.y += acc[i][1]; v.z += acc[i][2]; v.w += acc[i][3];
          h[0] = v;
        }
      }
    }
  }
  __syncwarp();
}
#else
__device__ __forceinline__ void h_accumulate(const KParams& K, int w, int lane, const SolverScr& s, int nefc) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, ld = s.ld;
  const float* J = s.Jw;
  const int a = lane >> 2, b = lane & 3;
  const int nblk = (nv + 31) >> 5;
  MJB_ROLL
  for (int bi = 0; bi < nblk; bi++) {
    MJB_ROLL
    for (int bj = 0; bj <= bi; bj++) {
      const int r0 = 32 * bi + 4 * a, c0 = 32 * bj + 8 * b;
      const bool rin = r0 < np, cin0 = c0 < np, cin1 = c0 + 4 < np;
      const bool need = rin && cin0 && (r0 + 3 >= c0);
      float acc[4][8];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
      // two rows per iteration: the six 128-bit loads of both rows are issued before the 64 FMAs
      const int nefc2 = (nefc + 1) & ~1;
      MJB_ROLL
      for (int r = 0; r < nefc2; r += 2) {
        const float cf0 = s.coef[r], cf1 = (r + 1 < nefc) ? s.coef[r + 1] : 0.f;
        if (cf0 == 0.f && cf1 == 0.f) continue;  // warp-uniform
        if (need) {
          const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
          const float* Jr0 = J + (size_t)r * np;
          const float* Jr1 = J + (size_t)min(r + 1, nefc - 1) * np;
          const bool u0 = cf0 != 0.f, u1 = cf1 != 0.f;
          const float4 ji0 = u0 ? *reinterpret_cast<const float4*>(Jr0 + r0) : z4;
          const float4 ja0 = u0 ? *reinterpret_cast<const float4*>(Jr0 + c0) : z4;
          const float4 jb0 = (u0 && cin1) ? *reinterpret_cast<const float4*>(Jr0 + c0 + 4) : z4;
          const float4 ji1 = u1 ? *reinterpret_cast<const float4*>(Jr1 + r0) : z4;
          const float4 ja1 = u1 ? *reinterpret_cast<const float4*>(Jr1 + c0) : z4;
          const float4 jb1 = (u1 && cin1) ? *reinterpret_cast<const float4*>(Jr1 + c0 + 4) : z4;
          {
            const float vi[4] = {ji0.x * cf0, ji0.y * cf0, ji0.z * cf0, ji0.w * cf0};
            const float vj[8] = {ja0.x, ja0.y, ja0.z, ja0.w, jb0.x, jb0.y, jb0.z, jb0.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
              for (int j = 0; j < 8; j++) acc[i][j] += vi[i] * vj[j];
          }
          {
            const float vi[4] = {ji1.x * cf1, ji1.y * cf1, ji1.z * cf1, ji1.w * cf1};
            const float vj[8] = {ja1.x, ja1.y, ja1.z, ja1.w, jb1.x, jb1.y, jb1.z, jb1.w};
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
              for (int j = 0; j < 8; j++) acc[i][j] += vi[i] * vj[j];
          }
        }
      }
      if (need) {
#if MJB_SOL_COMPACT
        // whole 4x8 tiles with 128-bit accesses; elements above the diagonal receive their (symmetric) values too
#pragma unroll
        for (int i = 0; i < 4; i++) {
          if (r0 + i < nv) {
            float4* h = reinterpret_cast<float4*>(s.H + (r0 + i) * ld + c0);
            float4 v = h[0];
            v.x += acc[i][0]; v.y += acc[i][1]; v.z += acc[i][2]; v.w += acc[i][3];
            h[0] = v;
            if (cin1) {
              v = h[1];
              v.x += acc[i][4]; v.y += acc[i][5]; v.z += acc[i][6]; v.w += acc[i][7];
              h[1] = v;
            }
          }
        }
#else
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const int ri = r0 + i, cj = c0 + j;
            if (ri < nv && cj <= ri) s.H[ri * ld + cj] += acc[i][j];
          }
#endif
      }
    }
  }
  __syncwarp();
}

#endif  // MJB_SOL_HTRI

// The same update on the tensor cores: H(lower) += A B with A[m][k] = coef[k] J[k][m] (16 x 8 tiles) and B[k][n] = J[k][n] (8 x 8),
// mma.sync m16n8k8 TF32 with the 3xTF32 split (a = a_hi + a_lo, products hi*hi + hi*lo + lo*hi, FP32 accumulate), which keeps the
// result at FP32 accuracy (relative error ~2^-21; plain TF32 would be 2^-11).  One 16-row band of H at a time, its <= 8 column
// tiles of the lower triangle accumulate in registers over the constraint rows, eight rows per step.  Fragment layout (PTX ISA,
// m16n8k8 .tf32): g = lane / 4, t = lane % 4; a0..3 = A[g | g+8][t | t+4], b0..1 = B[t | t+4][g], c0..3 = C[g | g+8][2t | 2t+1].
// Measured on B200 (humanoid, 8192 worlds, steady state, nefc ~ 46 x nv 27): the tensor-core version runs the solve kernel in
// 260.8 us against 232.9 us for the FP32 FMA register tiles below (profiles/r2_tensor_core_experiment.md): at these sizes the
// fragment gathers, the six conversions per operand pair of the 3xTF32 split and the dependent accumulator chain of mma.sync cost
// more than the 2 x 46 x 28 x 28 / 2 FMAs they replace.  It stays in the source behind MJB_SOL_MMA = 1; the default is FMA.
#ifndef MJB_SOL_MMA
#define MJB_SOL_MMA 0
#endif
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
  const float r = x - __uint_as_float(hi);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void h_accumulate_mma(const KParams& K, int w, int lane, const SolverScr& s, int nefc) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, ld = s.ld;
  const float* J = s.Jw;
  const int g = lane >> 2, t = lane & 3;
  const int nmt = (nv + 15) >> 4, ntile = (nv + 7) >> 3;
  MJB_ROLL
  for (int mt = 0; mt < nmt; mt++) {
    const int m0 = 16 * mt, nnt = min(2 * mt + 2, ntile);  // column tiles of this band that touch the lower triangle
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
    const bool ra = m0 + g < np, rb = m0 + g + 8 < np;
    MJB_ROLL
    for (int k0 = 0; k0 < nefc; k0 += 8) {
      const int ka = k0 + t, kb = k0 + t + 4;
      const float ca = ka < nefc ? s.coef[ka] : 0.f, cb = kb < nefc ? s.coef[kb] : 0.f;
      if (!__any_sync(FULL, ca != 0.f || cb != 0.f)) continue;  // warp-uniform: no row of this step changed state
      const float* Ja = J + (size_t)min(ka, nefc - 1) * np;
      const float* Jb = J + (size_t)min(kb, nefc - 1) * np;
      uint32_t ah[4], al[4];
      tf32_split(ra ? ca * Ja[m0 + g] : 0.f, ah[0], al[0]);
      tf32_split(rb ? ca * Ja[m0 + g + 8] : 0.f, ah[1], al[1]);
      tf32_split(ra ? cb * Jb[m0 + g] : 0.f, ah[2], al[2]);
      tf32_split(rb ? cb * Jb[m0 + g + 8] : 0.f, ah[3], al[3]);
#pragma unroll
      for (int nt = 0; nt < 8; nt++) {
        if (nt < nnt) {
          const int n = 8 * nt + g;
          uint32_t bh0, bl0, bh1, bl1;
          tf32_split((n < np && ka < nefc) ? Ja[n] : 0.f, bh0, bl0);
          tf32_split((n < np && kb < nefc) ? Jb[n] : 0.f, bh1, bl1);
          mma_tf32(acc[nt], al, bh0, bh1);
          mma_tf32(acc[nt], ah, bl0, bl1);
          mma_tf32(acc[nt], ah, bh0, bh1);
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 8; nt++) {
      if (nt < nnt) {
        const int c = 8 * nt + 2 * t;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int r = m0 + g + 8 * h;
          if (r < nv && c < np) {  // whole tiles: elements above the diagonal receive their (symmetric) values too
            float2* p = reinterpret_cast<float2*>(s.H + r * ld + c);
            float2 v = *p;
            v.x += acc[nt][2 * h];
            v.y += acc[nt][2 * h + 1];
            *p = v;
          }
        }
      }
    }
  }
  __syncwarp();
}

// solver.py:2154-2219 _update_constraint.  Returns the cost; sets force/state (global), coef.
template <bool ELL>
__device__ __forceinline__ void update_constraint(const KParams& K, int w, int lane, const SolverScr& s, int nefc, int ne, int nf, bool first,
                                                  float& cost, float& gauss, bool& any_change, bool& any_cone) {
  const int nv = K.ms.nv, np = K.ms.nv_pad;
  float c = 0.f;
  bool changed = false, cone = false;
  int* state = DW(efc_state);
  MJB_ROLL
  for (int r = lane; r < nefc; r += 32) {
    const float D = s.D[r], ja = s.Jaref[r];
    float f;
    int st;
    if (r < ne) {
      f = -D * ja; st = ST_QUADRATIC; c += 0.5f * D * ja * ja;
    } else if (r < ne + nf) {
      const float fl = DW(efc_frictionloss)[r], rf = safe_div(fl, D);
      if (ja <= -rf) { f = fl; st = ST_LINEARNEG; c += -fl * (0.5f * rf + ja); }
      else if (ja >= rf) { f = -fl; st = ST_LINEARPOS; c += -fl * (0.5f * rf - ja); }
      else { f = -D * ja; st = ST_QUADRATIC; c += 0.5f * D * ja * ja; }
    } else if (!(ELL && DW(efc_type)[r] == CT_CONTACT_ELLIPTIC)) {
      if (ja >= 0.f) { f = 0.f; st = ST_SATISFIED; }
      else { f = -D * ja; st = ST_QUADRATIC; c += 0.5f * D * ja * ja; }
    } else {  // elliptic friction cone contact, solver.py:1886-1943
      const EllRow e = ell_row(K, w, r, nefc);
      if (!e.dim) { f = 0.f; st = ST_SATISFIED; }
      else {
        const float mu = e.mu, N = s.Jaref[e.r0] * mu;
        float TT = 0.f, ufr = 0.f;
        for (int j = 1; j < e.dim; j++) {
          const float uj = s.Jaref[e.r0 + j] * e.fri[j - 1];
          TT += uj * uj;
          if (e.r0 + j == r) ufr = uj * e.fri[j - 1];
        }
        const float T = TT <= 0.f ? 0.f : sqrtf(TT);
        if (N >= mu * T || (T <= 0.f && N >= 0.f)) { f = 0.f; st = ST_SATISFIED; }
        else if (mu * N + T <= 0.f || (T <= 0.f && N < 0.f)) { f = -D * ja; st = ST_QUADRATIC; c += 0.5f * D * ja * ja; }
        else {
          const float dm = safe_div(s.D[e.r0], mu * mu * (1.0f + mu * mu)), nmt = N - mu * T, force = -dm * nmt * mu;
          if (r == e.r0) { f = force; c += 0.5f * dm * nmt * nmt; }
          else f = -safe_div(force, T) * ufr;
          st = ST_CONE;
        }
      }
    }
    if (ELL) cone |= st == ST_CONE;
    const bool newq = st == ST_QUADRATIC;
    const bool oldq = first ? false : (state[r] == ST_QUADRATIC);
    s.coef[r] = newq == oldq ? 0.f : (newq ? D : -D);
    changed |= newq != oldq;
    s.force[r] = f;
    state[r] = st;
    DW(efc_force)[r] = f;
  }
  any_change = __any_sync(FULL, changed);
  any_cone = ELL ? __any_sync(FULL, cone) : false;
  __syncwarp();
  // qfrc_constraint = J^T force (solver.py:1988).  Lane = (128-bit column group c, row group gi): the G = 4 / 2 / 1 lanes of a
  // column group sit next to each other and take every G-th row, four rows per trip with their 128-bit loads issued together; rows
  // without force multiply by zero instead of branching.  A 27-dof model: 7 column groups x 4 row groups = 28 lanes.
  const float* J = s.Jw;
  const int n4 = np >> 2;
  const int sh = n4 <= 8 ? 2 : n4 <= 16 ? 1 : 0, G = 1 << sh;
  const int rlast = nefc - 1;  // rows past nefc may hold anything: never loaded, their force is zero
  float g = 0.f;
  MJB_ROLL
  for (int cb = 0; cb < n4; cb += 32 >> sh) {
    const int c = cb + (lane >> sh), gi = lane & (G - 1);
    const bool cin = c < n4;
    const float4* Jc = reinterpret_cast<const float4*>(J) + (cin ? c : 0);
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    MJB_ROLL
    for (int r = gi; r < nefc; r += 4 * G) {
      const int r1 = r + G, r2 = r + 2 * G, r3 = r + 3 * G;
      const float f0 = s.force[r], f1 = r1 < nefc ? s.force[r1] : 0.f, f2 = r2 < nefc ? s.force[r2] : 0.f, f3 = r3 < nefc ? s.force[r3] : 0.f;
      const float4 j0 = Jc[(size_t)r * n4], j1 = Jc[(size_t)min(r1, rlast) * n4], j2 = Jc[(size_t)min(r2, rlast) * n4], j3 = Jc[(size_t)min(r3, rlast) * n4];  // clamped rows were written this step
      q.x += j0.x * f0 + j1.x * f1 + j2.x * f2 + j3.x * f3;
      q.y += j0.y * f0 + j1.y * f1 + j2.y * f2 + j3.y * f3;
      q.z += j0.z * f0 + j1.z * f1 + j2.z * f2 + j3.z * f3;
      q.w += j0.w * f0 + j1.w * f1 + j2.w * f2 + j3.w * f3;
    }
    for (int o = 1; o < G; o <<= 1) {
      q.x += __shfl_xor_sync(FULL, q.x, o);
      q.y += __shfl_xor_sync(FULL, q.y, o);
      q.z += __shfl_xor_sync(FULL, q.z, o);
      q.w += __shfl_xor_sync(FULL, q.w, o);
    }
    if (cin && gi == 0) {
      const float qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int i = 4 * c + k;
        if (i < nv) {
          DW(qfrc_constraint)[i] = qq[k];
          // grad = Ma - qfrc_smooth - qfrc_constraint (solver.py:2238)
          s.grad[i] = s.Ma[i] - s.qfs[i] - qq[k];
          g += (s.Ma[i] - s.qfs[i]) * (s.qacc[i] - DW(qacc_smooth)[i]);
        }
      }
    }
  }
  gauss = 0.5f * warp_sum(g);
  cost = warp_sum(c) + gauss;
  __syncwarp();
}

template <bool ELL>
__device__ __forceinline__ Pt3 ls_rows(const KParams& K, int w, int lane, const SolverScr& s, int nefc, int ne, int nf, float alpha) {
  Pt3 t = Pt3{0.f, 0.f, 0.f};
  MJB_ROLL
  for (int r = lane; r < nefc; r += 32) {
    const float fl = (r >= ne && r < ne + nf) ? DW(efc_frictionloss)[r] : 0.f;
    t = t + row_eval<ELL>(K, w, s, r, nefc, ne, nf, fl, alpha);
  }
  return t;
}

// solver.py:887-1341 linesearch_iterative (pyramidal); updates qacc, Ma, Jaref in scratch.
template <bool ELL>
__device__ __forceinline__ void linesearch(const KParams& K, int w, int lane, const SolverScr& s, int nefc, int ne, int nf, float gauss,
                                           float search_dot) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, n4 = np >> 2;
  // mv = M search, jv = J search
#if MJB_SOL_COMPACT
  mat_vec(DW(qM), n4, nv, s.search, s.mv, nullptr, lane);
  mat_vec(s.Jw, n4, nefc, s.search, s.jv, nullptr, lane);
#else
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) s.mv[i] = row_dot(DW(qM) + (size_t)i * np, s.search, n4);
  MJB_ROLL
  for (int r = lane; r < nefc; r += 32) s.jv[r] = row_dot(s.Jw + (size_t)r * np, s.search, n4);
#endif
  __syncwarp();
  const float scale = K.o.meaninertia * (float)nv;
  const float gtol = fmaxf(K.o.tolerance * K.o.ls_tolerance * sqrtf(search_dot) * scale, 1e-6f);
  float q1 = 0.f, q2 = 0.f;
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) {
    q1 += s.search[i] * (s.Ma[i] - s.qfs[i]);
    q2 += 0.5f * s.search[i] * s.mv[i];
  }
  q1 = warp_sum(q1);
  q2 = warp_sum(q2);
  const float q0 = gauss;
  Pt3 p0 = warp_sum(ls_rows<ELL>(K, w, lane, s, nefc, ne, nf, 0.f)) + Pt3{q0, q1, 2.0f * q2};
  const float lo_alpha_in = -safe_div(p0.g, p0.h);
  Pt3 lo_in = warp_sum(ls_rows<ELL>(K, w, lane, s, nefc, ne, nf, lo_alpha_in)) + eval_pt_quad(q0, q1, q2, lo_alpha_in);
  float alpha;
  if (uniform_b(fabsf(lo_in.g) < gtol && lo_in.c < p0.c)) {
    alpha = lo_alpha_in;
  } else {
    alpha = 0.f;
    const bool lo_less = lo_in.g < p0.g;
    Pt3 lo = lo_less ? lo_in : p0, hi = lo_less ? p0 : lo_in;
    float lo_alpha = lo_less ? lo_alpha_in : 0.f, hi_alpha = lo_less ? 0.f : lo_alpha_in;
    MJB_ROLL
    for (int it = 0; it < K.o.ls_iterations; it++) {
      const float lo_next_alpha = lo_alpha - safe_div(lo.g, lo.h);
      const float hi_next_alpha = hi_alpha - safe_div(hi.g, hi.h);
      const float mid_alpha = 0.5f * (lo_alpha + hi_alpha);
      Pt3 a = Pt3{0.f, 0.f, 0.f}, b = a, c = a;
      MJB_ROLL
      for (int r = lane; r < nefc; r += 32) {
        const float fl = (r >= ne && r < ne + nf) ? DW(efc_frictionloss)[r] : 0.f;
        if (ELL) {
          a = a + row_eval<true>(K, w, s, r, nefc, ne, nf, fl, lo_next_alpha);
          b = b + row_eval<true>(K, w, s, r, nefc, ne, nf, fl, hi_next_alpha);
          c = c + row_eval<true>(K, w, s, r, nefc, ne, nf, fl, mid_alpha);
          continue;
        }
        const float D = s.D[r], ja = s.Jaref[r], jv = s.jv[r];
        a = a + efc_eval_pt(r, lo_next_alpha, ne, nf, D, fl, ja, jv);
        b = b + efc_eval_pt(r, hi_next_alpha, ne, nf, D, fl, ja, jv);
        c = c + efc_eval_pt(r, mid_alpha, ne, nf, D, fl, ja, jv);
      }
      const Pt3 lo_next = warp_sum(a) + eval_pt_quad(q0, q1, q2, lo_next_alpha);
      const Pt3 hi_next = warp_sum(b) + eval_pt_quad(q0, q1, q2, hi_next_alpha);
      const Pt3 mid = warp_sum(c) + eval_pt_quad(q0, q1, q2, mid_alpha);
      // bracket swapping, solver.py:1293-1316
      const bool s1 = in_bracket(lo, lo_next);
      if (s1) { lo = lo_next; lo_alpha = lo_next_alpha; }
      const bool s2 = in_bracket(lo, mid);
      if (s2) { lo = mid; lo_alpha = mid_alpha; }
      const bool s3 = in_bracket(lo, hi_next);
      if (s3) { lo = hi_next; lo_alpha = hi_next_alpha; }
      const bool h1 = in_bracket(hi, hi_next);
      if (h1) { hi = hi_next; hi_alpha = hi_next_alpha; }
      const bool h2 = in_bracket(hi, mid);
      if (h2) { hi = mid; hi_alpha = mid_alpha; }
      const bool h3 = in_bracket(hi, lo_next);
      if (h3) { hi = lo_next; hi_alpha = lo_next_alpha; }
      const bool swap_lo = s1 || s2 || s3, swap_hi = h1 || h2 || h3;
      const bool ls_done = (!swap_lo && !swap_hi) || (lo.g < 0.f && lo.g > -gtol) || (hi.g > 0.f && hi.g < gtol);
      const bool improved = lo.c < p0.c || hi.c < p0.c, lo_better = lo.c < hi.c;
      if (improved) alpha = lo_better ? lo_alpha : hi_alpha;
      if (uniform_b(ls_done)) break;
    }
  }
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) {
    s.qacc[i] += alpha * s.search[i];
    s.Ma[i] += alpha * s.mv[i];
  }
  MJB_ROLL
  for (int r = lane; r < nefc; r += 32) s.Jaref[r] += alpha * s.jv[r];
  __syncwarp();
}

// Newton: factor H, Mgrad = H^-1 grad.  CG: Mgrad = M^-1 grad with the factor held in s.L.
// The factor is only recomputed when H changed since the last factorisation.
template <bool ELL>
__device__ __forceinline__ void solve_direction(const KParams& K, int w, int lane, const SolverScr& s, bool refactor, bool cones, int nefc) {
  const int nv = K.ms.nv, ld = s.ld;
  if (refactor) {
    const float4* h4 = reinterpret_cast<const float4*>(s.H);
    float4* l4 = reinterpret_cast<float4*>(s.L);
#pragma unroll 4
    for (int i = lane; i < (nv * ld) >> 2; i += 32) l4[i] = h4[i];
    __syncwarp();
    if (ELL && cones) cone_hessian(K, w, lane, s, nefc);
    warp_cholesky<!MJB_SOL_COMPACT>(s.L, ld, nv, s.rd, lane);
  }
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) s.x[i] = s.grad[i];
  __syncwarp();
  warp_chol_solve<!MJB_SOL_COMPACT>(s.L, ld, nv, s.rd, s.x, lane);
}

// Solver state of one world that lives in registers across passes.
struct SolveState {
  int it, niter, nefc, ne, nf;
  float cost, gauss, prev_cost, search_dot;
  bool had_cone;  // the current factor contains cone terms (elliptic variant)
};

// init_context part 1 (solver.py:3257): loads the world into shared memory.  Returns false when there is
// nothing to solve (solver.py:3298).
__device__ __forceinline__ bool solve_begin(const KParams& K, int w, int lane, SolverScr& s, SolveState& st) {
  const int nv = K.ms.nv, np = K.ms.nv_pad, njmax = K.dsz.njmax;
  if (njmax == 0 || nv == 0) {
    MJB_ROLL
    for (int i = lane; i < nv; i += 32) DW(qacc)[i] = DW(qacc_smooth)[i];
    if (lane == 0) DW(solver_niter)[0] = 0;
    __syncwarp();
    return false;
  }
  st.nefc = uniform_i(min(DW(nefc)[0], njmax));
  st.ne = uniform_i(DW(ne)[0]);
  st.nf = uniform_i(DW(nf)[0]);
  s.Jw = DW(efc_J);
  if (s.Jst != nullptr && st.nefc <= s.jcap) {  // stage the world's rows: coalesced 128-bit copies, all in flight
    const float4* g4 = reinterpret_cast<const float4*>(DW(efc_J));
    float4* s4 = reinterpret_cast<float4*>(s.Jst);
    const int n4tot = (st.nefc * np) >> 2;
#pragma unroll 4
    for (int i = lane; i < n4tot; i += 32) s4[i] = g4[i];
    s.Jw = s.Jst;
  }
  st.it = 0;
  st.niter = 0;
  st.had_cone = false;
  st.cost = st.gauss = st.prev_cost = st.search_dot = 0.f;
  const bool newton = K.o.solver == SOLVER_NEWTON;
  const float* src = (K.o.disableflags & DSBL_WARMSTART) ? DW(qacc_smooth) : DW(qacc_warmstart);
  MJB_ROLL
  for (int i = lane; i < np; i += 32) {
    s.qacc[i] = i < nv ? src[i] : 0.f;
    s.qfs[i] = i < nv ? DW(qfrc_smooth)[i] : 0.f;
    s.search[i] = 0.f;
  }
  MJB_ROLL
  for (int r = lane; r < st.nefc; r += 32) s.D[r] = DW(efc_D)[r];
  __syncwarp();
  // Jaref = J qacc - aref, Ma = M qacc
  const int n4 = np >> 2;
#if MJB_SOL_COMPACT
  mat_vec(s.Jw, n4, st.nefc, s.qacc, s.Jaref, DW(efc_aref), lane);
  mat_vec(DW(qM), n4, nv, s.qacc, s.Ma, nullptr, lane);
#else
  MJB_ROLL
  for (int r = lane; r < st.nefc; r += 32) s.Jaref[r] = row_dot(s.Jw + (size_t)r * np, s.qacc, n4) - DW(efc_aref)[r];
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) s.Ma[i] = row_dot(DW(qM) + (size_t)i * np, s.qacc, n4);
#endif
  __syncwarp();
  // Newton: H starts as qM.  CG: the preconditioner is the factor of qM (qLD).
  load_rows(newton ? s.H : s.L, s.ld, newton ? DW(qM) : DW(qLD), newton ? np : nv, nv, lane, nullptr, 0.f);
  if (!newton) chol_rdiag(s.L, s.ld, nv, s.rd, lane);
  return true;
}

// One pass.  Pass 0 is the rest of init_context (constraint update, gradient, first search direction);
// passes 1.. are solver.py:3188 _solver_iteration.  One body keeps a single copy of every routine in
// the instruction stream.  Returns true when the world is finished.
template <bool ELL>
__device__ __forceinline__ bool solve_pass(const KParams& K, int w, int lane, const SolverScr& s, SolveState& st) {
  const int nv = K.ms.nv, np = K.ms.nv_pad;
  const bool newton = K.o.solver == SOLVER_NEWTON;
  const bool init = st.it == 0;
  float* prev_grad = s.H;  // CG only (H is unused by CG)
  float* prev_Mgrad = s.H + np;
  if (!init) {
    linesearch<ELL>(K, w, lane, s, st.nefc, st.ne, st.nf, st.gauss, st.search_dot);
    if (!newton) {
      MJB_ROLL
      for (int i = lane; i < nv; i += 32) { prev_grad[i] = s.grad[i]; prev_Mgrad[i] = s.x[i]; }
      __syncwarp();
    }
    st.prev_cost = st.cost;
  }
  bool any_change, any_cone;
  update_constraint<ELL>(K, w, lane, s, st.nefc, st.ne, st.nf, init, st.cost, st.gauss, any_change, any_cone);
  // solver.py:3177 solve_done.  The reference evaluates it after the next search direction has been
  // computed; the direction of a finished world is never used, so it is tested first and the Hessian
  // update, factorisation and solve of the last pass are skipped (outputs are identical).
  if (!init) {
    float grad_dot = 0.f;
    MJB_ROLL
    for (int i = lane; i < nv; i += 32) grad_dot += s.grad[i] * s.grad[i];
    grad_dot = warp_sum(grad_dot);
    st.niter++;
    const float scl = 1.0f / (K.o.meaninertia * (float)nv);
    const float improvement = (st.prev_cost - st.cost) * scl, gradient = sqrtf(grad_dot) * scl;
    if (uniform_b(improvement < K.o.tolerance || gradient < K.o.tolerance) || st.niter >= K.o.iterations) return true;
  } else if (K.o.iterations <= 0) {
    return true;
  }
#if MJB_SOL_MMA
  if (newton && any_change) h_accumulate_mma(K, w, lane, s, st.nefc);
#else
  if (newton && any_change) h_accumulate(K, w, lane, s, st.nefc);
#endif
  // the cone part of the Hessian depends on Jaref: refactor whenever a cone is or was in the factor (solver.py:3205-3209)
  solve_direction<ELL>(K, w, lane, s, newton && (any_change || init || (ELL && (any_cone || st.had_cone))), any_cone, st.nefc);
  if (ELL) st.had_cone = any_cone;
  float beta = 0.f;
  if (!newton && !init) {  // Polak-Ribiere, solver.py:3081
    float num = 0.f, den = 0.f;
    MJB_ROLL
    for (int i = lane; i < nv; i += 32) { num += s.grad[i] * (s.x[i] - prev_Mgrad[i]); den += prev_grad[i] * prev_Mgrad[i]; }
    beta = fmaxf(0.f, warp_sum(num) / fmaxf(MJ_MINVAL, warp_sum(den)));
  }
  float search_dot = 0.f;
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) {
    float v = -s.x[i];
    if (!newton && !init) v += beta * s.search[i];
    s.search[i] = v;
    search_dot += v * v;
  }
  st.search_dot = warp_sum(search_dot);
  __syncwarp();
  st.it++;
  return false;
}

__device__ __forceinline__ void solve_finish(const KParams& K, int w, int lane, const SolverScr& s, const SolveState& st) {
  const int nv = K.ms.nv;
  const int niter = st.niter;
  MJB_ROLL
  for (int i = lane; i < nv; i += 32) {
    const float q = s.qacc[i];
    DW(qacc)[i] = q;
    DW(efc_Ma)[i] = s.Ma[i];
    if (!(q == q)) atomicOr(DW(status), MJB_STATUS_NAN_QACC);
  }
  if (lane == 0) DW(solver_niter)[0] = niter;
  __syncwarp();
}

// one world, start to finish (used when the solver is fused with other stages)
template <bool ELL>
__device__ __forceinline__ void s_solve(const KParams& K, const unsigned char* mb, int w, int lane, float* scr) {
  SolverScr s = solver_scr(scr, K.ms.nv, K.ms.nv_pad, K.dsz.njmax_pad, K.hbuf ? K.hbuf + (size_t)w * K.ms.nv * chol_ld(K.ms.nv_pad) : nullptr);
  SolveState st;
  if (!solve_begin(K, w, lane, s, st)) return;
  while (!solve_pass<ELL>(K, w, lane, s, st)) {
  }
  solve_finish(K, w, lane, s, st);
}

// All worlds of a CTA, pass-synchronous: the warps of the CTA take their solver passes together (a
// CTA-wide barrier per pass), each on its own world and at its own iteration number; a warp whose world
// has converged picks up its next world and rejoins at pass 0.  The barrier keeps the resident warps of
// an SM in the same phase of the (large) solver code, so they share instruction-cache lines instead of
// evicting each other's.
__device__ __forceinline__ void s_solve_cta(const KParams& K, int lane, int warp, int wpb, float* scr) {
  SolverScr s = solver_scr(scr, K.ms.nv, K.ms.nv_pad, K.dsz.njmax_pad, K.hbuf);  // experiment: see the note at MJB_SOL_SYNC
  const int stride = gridDim.x * wpb;
  int w = blockIdx.x * wpb + warp;
  SolveState st;
  bool running = false;  // a world is loaded and not yet finished
  while (true) {
    while (!running && w < K.dsz.nworld) {  // fetch worlds until one needs solving
      running = solve_begin(K, w, lane, s, st);
      if (!running) w += stride;
    }
    if (!__syncthreads_or(running ? 1 : 0)) break;
    if (running && solve_pass<false>(K, w, lane, s, st)) {
      solve_finish(K, w, lane, s, st);
      running = false;
      w += stride;
    }
  }
}
