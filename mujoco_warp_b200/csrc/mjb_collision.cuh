// mjb_collision.cuh -- NXN broadphase + primitive narrowphase, one warp per world.
//
// Reference: collision_driver.py:755 collision, :647 _nxn_broadphase, :274 _broadphase_filter,
// collision_primitive.py:1300 _primitive_narrowphase, collision_core.py:160 write_contact,
// :236 contact_params, collision_primitive_core.py:107-443 pair functions.
//
// The reference appends geom pairs and contacts to global pools with one atomic per element, so
// its ordering is nondeterministic.  Here each warp filters the static pair list of its world
// 32 pairs at a time, compacts survivors with ballots into a small shared-memory queue, runs the
// narrowphase with one lane per surviving pair, and reserves pool slots with ONE atomicAdd per
// queue flush (usually one per world): contacts of a world are emitted in pair-list order and
// the per-world list `wcon` records their pool indices for make_constraint.
#pragma once

#include "mjb_types.cuh"

// Compile-time variants of the collision kernel (template parameter FEAT): code a model / option set does not need
// is not compiled into the variant it runs, so it costs the common case neither registers nor instruction cache.
enum CollisionFeature {
  COLF_BOX = 1 /* sphere-box, capsule-box, box-box pairs present */,
  COLF_SAP = 2 /* sweep-and-prune broadphase */,
  COLF_CONVEX = 4 /* ellipsoid / cylinder geoms: their primitive pairs and the GJK / EPA pairs */,
  COLF_COUNT = 8
};

// Convex narrowphase core (GJK + EPA, reference collision_gjk.py), float32 device build of the shared C source
#define CCD_REAL float
#define CCD_FN static __device__ __forceinline__
#define CCD_FN_NOINLINE static __device__ __noinline__
#define CCD_SQRT sqrtf
#define CCD_FABS fabsf
// Warp-cooperative mesh support (product path only; the oracle keeps the reference's hill climbing, collision_gjk.py:153-176):
// when all 32 lanes run GJK / EPA on the SAME pair, the support vertex of a hull is the arg max of <vertex, dir> over ALL its
// vertices, lanes striding over the (coalesced) vertex table, instead of a serial walk over the hull's edge graph whose every
// step is a dependent global-memory round trip.  Same vertex as the hill climb on a convex hull (ties: lowest index).
__device__ __forceinline__ int ccd_coop_mesh_support(const float* vert, int nvert, const float* ld) {
  const int lane = threadIdx.x & 31;
  float best = -3.0e38f;
  int ibest = 0x7fffffff;
#pragma unroll 4
  for (int i = lane; i < nvert; i += 32) {
    const float d = vert[3 * i] * ld[0] + vert[3 * i + 1] * ld[1] + vert[3 * i + 2] * ld[2];
    if (d > best) { best = d; ibest = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, ibest, o);
    if (ob > best || (ob == best && oi < ibest)) { best = ob; ibest = oi; }
  }
  return ibest;
}
#define CCD_COOP_MESH_SUPPORT ccd_coop_mesh_support
#include "mjb_ccd_core.h"
// words of one EPA polytope, padded to 16 bytes (the warp-cooperative mesh pairs keep theirs in shared memory)
#define CCD_POLY_WORDS ((int)((sizeof(CcdPolytope) + 15) / 16 * 4))

// collision_driver.py:43 MJ_COLLISION_TABLE: pair types of CollisionType.CONVEX among the implemented geom types
// (box-box stays on the primitive, the reference's NATIVECCD-disabled route, :736)
__device__ __forceinline__ bool is_convex_pair(int t1, int t2) {
  if (t1 == GEOM_SPHERE || t1 == GEOM_CAPSULE) return t2 == GEOM_ELLIPSOID || (t1 == GEOM_CAPSULE && t2 == GEOM_CYLINDER);
  if (t1 == GEOM_ELLIPSOID) return t2 == GEOM_ELLIPSOID || t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  if (t1 == GEOM_CYLINDER) return t2 == GEOM_CYLINDER || t2 == GEOM_BOX;
  return false;
}
// nativeccd: box-box goes through GJK / EPA + multi-contact unless DisableBit.NATIVECCD (collision_driver.py:74, :736)
__device__ __forceinline__ bool is_ccd_pair(int t1, int t2, bool nativeccd = false) {
  return (t2 == GEOM_MESH && t1 >= GEOM_SPHERE) || is_convex_pair(t1, t2) || (nativeccd && t1 == GEOM_BOX && t2 == GEOM_BOX);
}

// collision_convex.py:712 eval_ccd_write_contact for one pair (single contact, no collision sensor): returns the number of
// contacts (0 / 1); dist is the surface distance, pos the midpoint of the witness points, nrm = x1 - x2 (unnormalised).
static __device__ __noinline__ int c_ccd_pair(const float* p1, const float* r1, const float* s1, int t1, const float* v1, int nv1, const int* gr1,
                                       const float* p2, const float* r2, const float* s2, int t2, const float* v2, int nv2, const int* gr2,
                                       float margin, float tolerance, int iterations, float* dist, float* pos, float* nrm, int coop,
                                       CcdPolytope* shared_pt) {
  CcdGeom g1, g2;
  g1.coop = g2.coop = coop;
  // ~5 kB of local memory per thread; the warp-cooperative path passes ONE polytope in shared memory instead (every lane computes
  // and stores the same values), which takes the EPA's face / vertex tables off the local-memory path
  CcdPolytope pt_local;
  CcdPolytope& pt = shared_pt ? *shared_pt : pt_local;
  CcdGjk res;
#pragma unroll
  for (int i = 0; i < 3; i++) { g1.pos[i] = p1[i]; g2.pos[i] = p2[i]; g1.size[i] = s1[i]; g2.size[i] = s2[i]; }
#pragma unroll
  for (int i = 0; i < 9; i++) { g1.rot[i] = r1[i]; g2.rot[i] = r2[i]; }
  g1.margin = g2.margin = margin;
  g1.type = t1;
  g2.type = t2;
  g1.vert = v1; g1.nvert = nv1; g1.graph = gr1;
  g2.vert = v2; g2.nvert = nv2; g2.graph = gr2;
  if (iterations > CCD_MAX_ITER) iterations = CCD_MAX_ITER;
  float d, x1[3], x2[3];
  int face_idx = -1;
  const int nc = ccd_run(tolerance, 0.f, iterations, iterations, g1, g2, &pt, &res, &d, x1, x2, &face_idx);
  if (nc == 0 || d >= 0.f) return 0;
#pragma unroll
  for (int i = 0; i < 3; i++) nrm[i] = x1[i] - x2[i];
  if (t1 == GEOM_BOX && t2 == GEOM_BOX && face_idx > -1) {  // collision_convex.py:809-846: up to 4 contacts by face clipping
    CcdMulti mc;
    float w1[4][3], w2[4][3];
#pragma unroll
    for (int i = 0; i < 3; i++) { w1[0][i] = x1[i]; w2[0][i] = x2[i]; }
    const int n = ccd_multicontact_box(&mc, &pt, pt.face[face_idx], &g1, &g2, w1, w2);
    for (int k = 0; k < n; k++) {
      dist[k] = d + margin;
      for (int i = 0; i < 3; i++) pos[3 * k + i] = 0.5f * (w1[k][i] + w2[k][i]);
    }
    for (int i = 0; i < 3; i++) nrm[i] = w1[0][i] - w2[0][i];
    return n;
  }
  *dist = d + margin;
#pragma unroll
  for (int i = 0; i < 3; i++) pos[i] = 0.5f * (x1[i] + x2[i]);
  return 1;
}

// collision_convex.py:56 _hfield_filter + :158 ccd_hfield_kernel: a convex geom against a height field.  The geom is brought into the
// field's frame, the grid cells under its bounding box are walked, every cell contributes two triangular prisms (bottom at
// -z_bottom, top at the elevations + margin) that run GJK / EPA against the geom, and up to four of the resulting contacts are kept:
// the deepest, the one furthest from it, the one furthest from their line, the one furthest from the other triangle edge.  Every
// contact carries its own normal.  One lane per pair (the prisms are walked serially).
#define MJB_MAXCONPAIR 50
static __device__ __noinline__ void c_hfield_pair(const KParams& K, const unsigned char* mb, int w, int g1, int g2, float margin, float* dist4,
                                                  float* pos12, float* nrm12) {
  for (int k = 0; k < 4; k++) dist4[k] = INFINITY;
  const int hid = MI(geom_dataid)[g1];
  const float* size1 = MR(hfield_size) + 4 * hid;
  const float *pos1 = DW(geom_xpos) + 3 * g1, *mat1 = DW(geom_xmat) + 9 * g1, *pos2 = DW(geom_xpos) + 3 * g2, *mat2 = DW(geom_xmat) + 9 * g2;
  const float dp[3] = {pos2[0] - pos1[0], pos2[1] - pos1[1], pos2[2] - pos1[2]};
  float pos[3];
  for (int i = 0; i < 3; i++) pos[i] = mat1[i] * dp[0] + mat1[3 + i] * dp[1] + mat1[6 + i] * dp[2];
  const float r2 = MR(geom_rbound)[g2], fm = MR(geom_margin)[g1] + MR(geom_margin)[g2];
  for (int i = 0; i < 2; i++)
    if (size1[i] < pos[i] - r2 - fm || -size1[i] > pos[i] + r2 + fm) return;
  if (size1[2] < pos[2] - r2 - fm) return;
  if (-size1[3] > pos[2] + r2 + fm) return;
  CcdGeom ga, gb;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) gb.rot[3 * i + j] = mat1[i] * mat2[j] + mat1[3 + i] * mat2[3 + j] + mat1[6 + i] * mat2[6 + j];
  for (int i = 0; i < 3; i++) { gb.pos[i] = pos[i]; gb.size[i] = MR(geom_size)[3 * g2 + i]; ga.size[i] = 0.f; }
  gb.margin = 0.f;
  gb.type = MI(geom_type)[g2];
  gb.vert = nullptr; gb.nvert = 0; gb.graph = nullptr; gb.coop = 0;
  ga.vert = nullptr; ga.nvert = 0; ga.graph = nullptr; ga.coop = 0;
  if (gb.type == GEOM_MESH) {
    const int id = MI(geom_dataid)[g2];
    gb.vert = MRC(mesh_vert) + 3 * MI(mesh_vertadr)[id];
    gb.nvert = MI(mesh_vertnum)[id];
    gb.graph = MI(mesh_graphadr)[id] >= 0 ? MIC(mesh_graph) + MI(mesh_graphadr)[id] : nullptr;
  }
  float ext[6];  // xmax xmin ymax ymin zmax zmin by support queries
  for (int a = 0; a < 3; a++)
    for (int sgn = 0; sgn < 2; sgn++) {
      float dir[3] = {0.f, 0.f, 0.f}, pnt[3];
      int cache = -1;
      dir[a] = sgn ? -1.f : 1.f;
      ccd_support(&gb, dir, pnt, &cache);
      ext[2 * a + sgn] = pnt[a];
    }
  const float xmax = ext[0], xmin = ext[1], ymax = ext[2], ymin = ext[3], zmax = ext[4], zmin = ext[5];
  if (xmin - fm > size1[0] || xmax + fm < -size1[0] || ymin - fm > size1[1] || ymax + fm < -size1[1] || zmin - fm > size1[2] || zmax + fm < -size1[3])
    return;
  ga.type = GEOM_HFIELD;
  ga.margin = 0.f;
  for (int i = 0; i < 9; i++) ga.rot[i] = (i == 0 || i == 4 || i == 8) ? 1.f : 0.f;
  const int nrow = MI(hfield_nrow)[hid], ncol = MI(hfield_ncol)[hid], adr = MI(hfield_adr)[hid];
  const float* hdata = MRC(hfield_data) + adr;
  const float x_scale = 0.5f * (float)(ncol - 1) / size1[0], y_scale = 0.5f * (float)(nrow - 1) / size1[1];
  const int cmin = max(0, (int)floorf((xmin + size1[0]) * x_scale)), cmax = min(ncol - 1, (int)ceilf((xmax + size1[0]) * x_scale));
  const int rmin = max(0, (int)floorf((ymin + size1[1]) * y_scale)), rmax = min(nrow - 1, (int)ceilf((ymax + size1[1]) * y_scale));
  const float dx = 2.0f * size1[0] / (float)(ncol - 1), dy = 2.0f * size1[1] / (float)(nrow - 1);
  float prism[6][3];
  for (int i = 0; i < 6; i++) prism[i][0] = prism[i][1] = prism[i][2] = 0.f;
  prism[0][2] = prism[1][2] = prism[2][2] = -size1[3];
  float cdist[MJB_MAXCONPAIR], cpos[MJB_MAXCONPAIR][3], cnrm[MJB_MAXCONPAIR][3];
  float min_dist = MJ_MAXVAL;
  int min_id = -1, count = 0;
  gb.margin = margin;
  CcdPolytope pt;
  CcdGjk res;
  int iterations = K.o.ccd_iterations;
  if (iterations > CCD_MAX_ITER) iterations = CCD_MAX_ITER;
  auto add_vert = [&](int cc, int rr) {
    const float x = dx * (float)cc - size1[0], y = dy * (float)rr - size1[1], z = hdata[rr * ncol + cc] * size1[2] + margin;
    for (int k = 0; k < 3; k++) { prism[0][k] = prism[1][k]; prism[1][k] = prism[2][k]; prism[3][k] = prism[4][k]; prism[4][k] = prism[5][k]; }
    prism[2][0] = x; prism[5][0] = x; prism[2][1] = y; prism[5][1] = y; prism[5][2] = z;
  };
  for (int r = rmin; r < rmax; r++) {
    add_vert(cmin, r + 1);
    add_vert(cmin, r);
    for (int c = cmin + 1; c <= cmax; c++) {
      for (int i = 0; i < 2; i++) {
        if (count >= MJB_MAXCONPAIR) continue;  // the reference reports an overflow and skips (without advancing the prism)
        add_vert(c, r + (i == 0 ? 1 : 0));
        if (prism[3][2] < zmin && prism[4][2] < zmin && prism[5][2] < zmin) continue;
        for (int k = 0; k < 6; k++) { ga.prism[k][0] = prism[k][0]; ga.prism[k][1] = prism[k][1]; ga.prism[k][2] = prism[k][2]; }
        for (int k = 0; k < 3; k++) ga.pos[k] = (prism[0][k] + prism[1][k] + prism[2][k] + prism[3][k] + prism[4][k] + prism[5][k]) * (1.0f / 6.0f);
        float d, w1[3], w2[3];
        int face_idx;
        const int nc = ccd_run(K.o.ccd_tolerance, 0.f, iterations, iterations, ga, gb, &pt, &res, &d, w1, w2, &face_idx);
        if (nc == 0) continue;
        const V3 pl{0.5f * (w1[0] + w2[0]), 0.5f * (w1[1] + w2[1]), 0.5f * (w1[2] + w2[2])};
        {
          // Float32 safeguard (not in the reference): the contact point lies between a point of the geom and a point at distance |d|
          // from it, i.e. within rbound + |d| of the geom's centre.  A degenerate prism can make the float32 EPA return a witness
          // hundreds of metres away (seen once in 4096 x 370 steps of the apollo height-field scene); such a contact row has a
          // Jacobian norm ~500 and turns the world's solve into NaN.  Dropped instead.
          const float ex = pl.x - pos[0], ey = pl.y - pos[1], ez = pl.z - pos[2];
          const float lim = r2 + fabsf(d) + fm + 1e-3f * (1.0f + r2);
          if (!(ex * ex + ey * ey + ez * ez <= lim * lim)) continue;
        }
        cdist[count] = d;
        const M3 fr = make_frame(V3{w1[0] - w2[0], w1[1] - w2[1], w1[2] - w2[2]});
        const M3 m1 = ldm3(mat1);
        const V3 pg = mat_vec(m1, pl) + ld3(pos1), ng = mat_vec(m1, V3{fr.m[0], fr.m[1], fr.m[2]});
        st3(cpos[count], pg);
        st3(cnrm[count], ng);
        if (d < min_dist) { min_dist = d; min_id = count; }
        count++;
      }
    }
  }
  if (min_id < 0) return;
  const V3 min_pos = ld3(cpos[min_id]), min_normal = ld3(cnrm[min_id]);
  auto put = [&](int k, int id) {
    dist4[k] = cdist[id];
    for (int i = 0; i < 3; i++) { pos12[3 * k + i] = cpos[id][i]; nrm12[3 * k + i] = cnrm[id][i]; }
  };
  put(0, min_id);
  const float MIN_NEXT = 1.0e-3f;
  int id1 = -1, id2 = -1, id3 = -1;
  float dist1 = -MJ_MAXVAL, dist12 = -MJ_MAXVAL, dist3 = -MJ_MAXVAL;
  for (int i = 0; i < count; i++) {
    if (i == min_id) continue;
    const float dd = norm(ld3(cpos[i]) - min_pos);
    if (dd > dist1) { id1 = i; dist1 = dd; }
  }
  if (id1 == -1 || (0.f < dist1 && dist1 < MIN_NEXT)) return;
  put(1, id1);
  const V3 p1 = ld3(cpos[id1]), d_min1 = cross(min_normal, min_pos - p1);
  for (int i = 0; i < count; i++) {
    if (i == min_id || i == id1) continue;
    const float dd = fabsf(dot(ld3(cpos[i]) - min_pos, d_min1));
    if (dd > dist12) { id2 = i; dist12 = dd; }
  }
  if (id2 == -1 || (0.f < dist12 && dist12 < MIN_NEXT)) return;
  put(2, id2);
  const V3 p2 = ld3(cpos[id2]), vec_min2 = cross(min_normal, min_pos - p2), vec_12 = cross(min_normal, p1 - p2);
  for (int i = 0; i < count; i++) {
    if (i == min_id || i == id1 || i == id2) continue;
    const V3 hp = ld3(cpos[i]);
    const float dd = fabsf(dot(hp - min_pos, vec_min2)) + fabsf(dot(p1 - hp, vec_12));
    if (dd > dist3) { id3 = i; dist3 = dd; }
  }
  if (id3 == -1 || (0.f < dist3 && dist3 < MIN_NEXT)) return;
  put(3, id3);
}

// collision_primitive.py:52-277 plane_convex (hill climbing over the mesh graph for >= 10 vertices, else exhaustive): up to 4 contacts at the
// deepest vertex a, the in-threshold vertex b furthest from a, c furthest from the line ab, d furthest from the edges ac / bc.
static __device__ __noinline__ int c_plane_mesh(const float* p1, const float* n1, const float* p2, const float* r2, const float* vert, int nvert,
                                         const int* graph, float* dist, float* pos) {
  const float HUGE_V = 1e6f;
  float pl[3], n[3];
  const float dp[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
#pragma unroll
  for (int i = 0; i < 3; i++) {
    pl[i] = r2[i] * dp[0] + r2[3 + i] * dp[1] + r2[6 + i] * dp[2];
    n[i] = r2[i] * n1[0] + r2[3 + i] * n1[1] + r2[6 + i] * n1[2];
  }
  int idx[4] = {-1, -1, -1, -1};
  float max_support = -HUGE_V, a[3] = {0.f, 0.f, 0.f}, b[3] = {0.f, 0.f, 0.f}, c[3] = {0.f, 0.f, 0.f};
  auto support = [&](int i) {
    return (pl[0] - vert[3 * i]) * n[0] + (pl[1] - vert[3 * i + 1]) * n[1] + (pl[2] - vert[3 * i + 2]) * n[2];
  };
  const V3 nn{n[0], n[1], n[2]};
  bool climbed = false;
  if (graph && nvert >= 10) {  // collision_primitive.py:140-255: the same four searches by hill climbing over the hull's edge graph
    climbed = true;
    const int numvert = graph[0];
    const int *vert_edgeadr = graph + 2, *vert_globalid = graph + 2 + numvert, *edge_localid = graph + 2 + 2 * numvert;
    int imax = 0;
    float threshold = 0.f, best = -HUGE_V;
    V3 ab{0.f, 0.f, 0.f}, ac = ab, bc = ab;
    for (int phase = 0; phase < 5; phase++) {  // 0: deepest vertex, 1: a (deepest above the threshold), 2: b, 3: c, 4: d
      if (phase > 0) best = -HUGE_V;
      for (;;) {
        const int prev = imax;
        int i = vert_edgeadr[imax];
        while (edge_localid[i] >= 0) {
          const int subidx = edge_localid[i], vi = vert_globalid[subidx];
          const float sp = support(vi), mask = sp > threshold ? 0.f : -HUGE_V;
          const V3 ap{a[0] - vert[3 * vi], a[1] - vert[3 * vi + 1], a[2] - vert[3 * vi + 2]};
          float val;
          if (phase == 0) val = sp;
          else if (phase == 1) val = sp > threshold ? sp : -HUGE_V;
          else if (phase == 2) val = dot(ap, ap) + mask;
          else if (phase == 3) val = fabsf(dot(ap, ab)) + mask;
          else {
            const V3 bp{b[0] - vert[3 * vi], b[1] - vert[3 * vi + 1], b[2] - vert[3 * vi + 2]};
            val = (fabsf(dot(ap, ac)) + mask) + (fabsf(dot(bp, bc)) + mask);
          }
          if (val > best) { best = val; imax = subidx; }
          i++;
        }
        if (imax == prev) break;
      }
      const int gi = vert_globalid[imax];
      const float* v = vert + 3 * gi;
      if (phase == 0) { max_support = best; threshold = fmaxf(0.f, max_support - 1e-3f); }
      else if (phase == 1) { idx[0] = gi; a[0] = v[0]; a[1] = v[1]; a[2] = v[2]; }
      else if (phase == 2) {
        idx[1] = gi; b[0] = v[0]; b[1] = v[1]; b[2] = v[2];
        ab = cross(nn, V3{a[0] - b[0], a[1] - b[1], a[2] - b[2]});
      } else if (phase == 3) {
        idx[2] = gi; c[0] = v[0]; c[1] = v[1]; c[2] = v[2];
        ac = cross(nn, V3{a[0] - c[0], a[1] - c[1], a[2] - c[2]});
        bc = cross(nn, V3{b[0] - c[0], b[1] - c[1], b[2] - c[2]});
      } else idx[3] = gi;
    }
  }
  if (!climbed) {
  for (int i = 0; i < nvert; i++) {
    const float sp = support(i);
    if (sp > max_support) { max_support = sp; idx[0] = i; a[0] = vert[3 * i]; a[1] = vert[3 * i + 1]; a[2] = vert[3 * i + 2]; }
  }
  if (max_support < 0.f) return 0;
  const float threshold = max_support - 1e-3f;
  float bd = -HUGE_V, cd = -HUGE_V, dd = -HUGE_V;
  for (int i = 0; i < nvert; i++) {
    const float mask = support(i) > threshold ? 0.f : -HUGE_V;
    const float t0 = a[0] - vert[3 * i], t1 = a[1] - vert[3 * i + 1], t2 = a[2] - vert[3 * i + 2];
    const float d = t0 * t0 + t1 * t1 + t2 * t2 + mask;
    if (d > bd) { idx[1] = i; bd = d; b[0] = vert[3 * i]; b[1] = vert[3 * i + 1]; b[2] = vert[3 * i + 2]; }
  }
  const V3 ab = cross(nn, V3{a[0] - b[0], a[1] - b[1], a[2] - b[2]});
  for (int i = 0; i < nvert; i++) {
    const float mask = support(i) > threshold ? 0.f : -HUGE_V;
    const V3 ap{a[0] - vert[3 * i], a[1] - vert[3 * i + 1], a[2] - vert[3 * i + 2]};
    const float d = fabsf(dot(ap, ab)) + mask;
    if (d > cd) { idx[2] = i; cd = d; c[0] = vert[3 * i]; c[1] = vert[3 * i + 1]; c[2] = vert[3 * i + 2]; }
  }
  const V3 ac = cross(nn, V3{a[0] - c[0], a[1] - c[1], a[2] - c[2]}), bc = cross(nn, V3{b[0] - c[0], b[1] - c[1], b[2] - c[2]});
  for (int i = 0; i < nvert; i++) {
    const float mask = support(i) > threshold ? 0.f : -HUGE_V;
    const V3 ap{a[0] - vert[3 * i], a[1] - vert[3 * i + 1], a[2] - vert[3 * i + 2]};
    const V3 bp{b[0] - vert[3 * i], b[1] - vert[3 * i + 1], b[2] - vert[3 * i + 2]};
    const float d = (fabsf(dot(ap, ac)) + mask) + (fabsf(dot(bp, bc)) + mask);
    if (d > dd) { idx[3] = i; dd = d; }
  }
  }  // exhaustive search
  int count = 0;
  for (int i = 3; i >= 0; i--) {  // :258-275 one contact per index that appears once among idx[0..i]
    int rep = 0;
    for (int j = 0; j <= i; j++) rep += idx[j] == idx[i];
    if (rep != 1) continue;
    const float* v = vert + 3 * idx[i];
    const float d = -support(idx[i]);
    dist[count] = d;
#pragma unroll
    for (int k = 0; k < 3; k++) pos[3 * count + k] = p2[k] + r2[3 * k] * v[0] + r2[3 * k + 1] * v[1] + r2[3 * k + 2] * v[2] - 0.5f * d * n1[k];
    count++;
  }
  return count;
}

// collision_driver.py:91-111, 218-271
__device__ __forceinline__ bool bp_filter(const KParams& K, const unsigned char* mb, int w, int g1, int g2) {
  const float rb1 = MR(geom_rbound)[g1], rb2 = MR(geom_rbound)[g2];
  const float margin = MR(geom_margin)[g1] + MR(geom_margin)[g2];
  const V3 p1 = ld3(DW(geom_xpos) + 3 * g1), p2 = ld3(DW(geom_xpos) + 3 * g2);
  const int filter = K.o.broadphase_filter;
  if (rb1 == 0.f || rb2 == 0.f) {
    if (filter & BF_PLANE) {
      if (rb1 == 0.f) {
        const float* r = DW(geom_xmat) + 9 * g1;
        return dot(p2 - p1, V3{r[2], r[5], r[8]}) <= rb2 + margin;
      } else {
        const float* r = DW(geom_xmat) + 9 * g2;
        return dot(p1 - p2, V3{r[2], r[5], r[8]}) <= rb1 + margin;
      }
    }
    return true;
  }
  if (filter & BF_SPHERE) {
    const float bound = rb1 + rb2 + margin;
    const V3 dif = p2 - p1;
    if (!(dot(dif, dif) <= bound * bound)) return false;
  }
  if (filter & (BF_AABB | BF_OBB)) {
    const M3 r1 = ldm3(DW(geom_xmat) + 9 * g1), r2 = ldm3(DW(geom_xmat) + 9 * g2);
    const float* a1 = MR(geom_aabb) + 6 * g1;
    const float* a2 = MR(geom_aabb) + 6 * g2;
    const V3 c1 = mat_vec(r1, ld3(a1)) + p1, c2 = mat_vec(r2, ld3(a2)) + p2;
    const V3 s1 = ld3(a1 + 3), s2 = ld3(a2 + 3);
    if (filter & BF_AABB) {  // collision_driver.py:116-213: world AABBs of the two boxes
      float h1[3], h2[3];
#pragma unroll
      for (int i = 0; i < 3; i++) {
        h1[i] = fabsf(r1.m[3 * i] * s1.x) + fabsf(r1.m[3 * i + 1] * s1.y) + fabsf(r1.m[3 * i + 2] * s1.z);
        h2[i] = fabsf(r2.m[3 * i] * s2.x) + fabsf(r2.m[3 * i + 1] * s2.y) + fabsf(r2.m[3 * i + 2] * s2.z);
      }
      const float cc1[3] = {c1.x, c1.y, c1.z}, cc2[3] = {c2.x, c2.y, c2.z};
#pragma unroll
      for (int i = 0; i < 3; i++) {
        if (cc1[i] + h1[i] + margin < cc2[i] - h2[i]) return false;
        if (cc2[i] + h2[i] + margin < cc1[i] - h1[i]) return false;
      }
    }
    if (filter & BF_OBB) {  // 6 face normals only
#pragma unroll
      for (int j = 0; j < 2; j++) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
          const V3 n = j == 0 ? mat_col(r1, k) : mat_col(r2, k);
          const float rad1 = fabsf(s1.x * dot(mat_col(r1, 0), n)) + fabsf(s1.y * dot(mat_col(r1, 1), n)) + fabsf(s1.z * dot(mat_col(r1, 2), n));
          const float rad2 = fabsf(s2.x * dot(mat_col(r2, 0), n)) + fabsf(s2.y * dot(mat_col(r2, 1), n)) + fabsf(s2.z * dot(mat_col(r2, 2), n));
          if (rad1 + rad2 + margin < fabsf(dot(c2, n) - dot(c1, n))) return false;
        }
      }
    }
  }
  return true;
}

struct PairGeom {
  V3 p1, p2, n1, n2, s1, s2;  // positions, z axes, sizes
  int t1, t2;
};

// collision_primitive_core.py:115-143
__device__ __forceinline__ void c_sphere_sphere(V3 p1, float r1, V3 p2, float r2, float& dist, V3& pos, V3& n) {
  const V3 dir = p2 - p1;
  const float dd = norm(dir);
  n = dd == 0.f ? V3{1.f, 0.f, 0.f} : dir * (1.0f / dd);
  dist = dd - (r1 + r2);
  pos = p1 + n * (r1 + 0.5f * dist);
}

__device__ __forceinline__ float safe_divf(float x, float y) { return x / (y != 0.f ? y : MJ_MINVAL); }

// Sphere (geom 1) vs box (geom 2), collision_primitive_core.py:1103-1154.  `br` row-major box rotation.
__device__ __forceinline__ void c_sphere_box(V3 sp, float r, V3 bp, const float* br, const float* bs, float& dist, V3& pos, V3& nrm) {
  const V3 d0 = sp - bp;
  float c[3], q[3], dl[3], nl[3] = {0.f, 0.f, 0.f}, loc[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    c[i] = br[i] * d0.x + br[3 + i] * d0.y + br[6 + i] * d0.z;
    q[i] = clampf(c[i], -bs[i], bs[i]);
    dl[i] = q[i] - c[i];
  }
  const float len = sqrtf(dl[0] * dl[0] + dl[1] * dl[1] + dl[2] * dl[2]);
  if (len <= MJ_MINVAL) {  // centre inside the box: nearest face, scanned -x,+x,-y,+y,-z,+z
    float closest = 2.f * (bs[0] + bs[1] + bs[2]);
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) {
      const float fd = fabsf(((i & 1) ? 1.f : -1.f) * bs[i / 2] - c[i / 2]);
      if (closest > fd) { closest = fd; k = i; }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
      nl[i] = (i == k / 2) ? ((k & 1) ? -1.f : 1.f) : 0.f;
      loc[i] = c[i] + nl[i] * (r - closest) * 0.5f;
    }
    dist = -closest - r;
  } else {
    const float inv = 1.f / len;
#pragma unroll
    for (int i = 0; i < 3; i++) { nl[i] = dl[i] * inv; loc[i] = 0.5f * (q[i] + c[i] + nl[i] * r); }
    dist = len - r;
  }
  nrm = V3{br[0] * nl[0] + br[1] * nl[1] + br[2] * nl[2], br[3] * nl[0] + br[4] * nl[1] + br[5] * nl[2], br[6] * nl[0] + br[7] * nl[1] + br[8] * nl[2]};
  pos = bp + V3{br[0] * loc[0] + br[1] * loc[1] + br[2] * loc[2], br[3] * loc[0] + br[4] * loc[1] + br[5] * loc[2],
                br[6] * loc[0] + br[7] * loc[1] + br[8] * loc[2]};
}

// Capsule (geom 1) vs box (geom 2), collision_primitive_core.py:1158-1491: closest feature pair between the capsule
// segment and the box (a segment end against a face, then the segment against each of the 12 edges), the discrete
// configuration decides where a second support point along the segment lies, and both points are collided as spheres.
// Returns the segment parameters t[0], t[1] (in half-lengths from the centre); n = number of points (0..2).
// Out of line: keeps its register needs out of the humanoid's capsule / sphere / plane paths.
__device__ __forceinline__ int capsule_box_points(V3 cp, V3 caxis, float hl, V3 bp, const float* br, const float* bs, float* t) {
  const V3 d0 = cp - bp;
  float pl[3], ax[3], half[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    pl[i] = br[i] * d0.x + br[3 + i] * d0.y + br[6 + i] * d0.z;
    ax[i] = br[i] * caxis.x + br[3 + i] * caxis.y + br[6 + i] * caxis.z;
    half[i] = ax[i] * hl;
  }
  const int axisdir = (half[0] > 0.f ? 1 : 0) + (half[1] > 0.f ? 2 : 0) + (half[2] > 0.f ? 4 : 0);
  float best = 1.0e32f, bestseg = -12.f, bestbox = 0.f;
  int cltype = -4, clface = -12, clcorner = -123, cledge = -123;
#pragma unroll
  for (int sg = -1; sg <= 1; sg += 2) {
    float dd = 0.f;
    int nout = 0, axout = -1;
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const float tip = pl[j] + (float)sg * half[j];
      float cj = tip;
      if (cj < -bs[j]) { nout++; axout = j; cj = -bs[j]; } else if (cj > bs[j]) { nout++; axout = j; cj = bs[j]; }
      dd += (cj - tip) * (cj - tip);
    }
    if (nout <= 1 && dd < best) { best = dd; bestseg = (float)sg; cltype = -2 + sg; clface = axout; }
  }
  const float mc = hl * hl;
  for (int i = 0; i < 8; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) {
      if (i & (1 << j)) continue;
      float dif[3];
#pragma unroll
      for (int k = 0; k < 3; k++) dif[k] = (k == j ? 0.f : ((i >> k) & 1 ? bs[k] : -bs[k])) - pl[k];
      const float u = -bs[j] * dif[j], v = half[0] * dif[0] + half[1] * dif[1] + half[2] * dif[2];
      const float ma = bs[j] * bs[j], mb = -bs[j] * half[j], det = ma * mc - mb * mb;
      if (fabsf(det) < MJ_MINVAL) continue;
      const float idet = 1.f / det;
      float x1 = (mc * u - mb * v) * idet, x2 = (ma * v - mb * u) * idet;
      int s1 = 1, s2 = 1;
      if (x1 > 1.f) { x1 = 1.f; s1 = 2; x2 = safe_divf(v - mb, mc); }
      else if (x1 < -1.f) { x1 = -1.f; s1 = 0; x2 = safe_divf(v + mb, mc); }
      if (x2 > 1.f || x2 < -1.f) {
        if (x2 > 1.f) { x2 = 1.f; s2 = 2; x1 = safe_divf(u - mb, ma); } else { x2 = -1.f; s2 = 0; x1 = safe_divf(u + mb, ma); }
        if (x1 > 1.f) { x1 = 1.f; s1 = 2; } else if (x1 < -1.f) { x1 = -1.f; s1 = 0; }
      }
      float dd = 0.f;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const float e = dif[k] - half[k] * x2 + (k == j ? bs[j] * x1 : 0.f);
        dd += e * e;
      }
      if (dd < best - MJ_MINVAL) {
        best = dd; bestseg = x2; bestbox = x1;
        cltype = s1 * 3 + s2;
        clcorner = i + (1 << j) * (cltype / 6);
        cledge = j;
      }
    }
  }
  if (cltype == -4) return 0;
  float second = -4.f;
  if (cltype >= 0 && cltype / 3 != 1) {  // a box corner is closest
    int c1 = axisdir ^ clcorner;
    if (c1 != 0 && c1 != 7) {
      float mul = 1.f;
      if (!(c1 == 1 || c1 == 2 || c1 == 4)) { mul = -1.f; c1 = 7 - c1; }
      const int a = c1 == 1 ? 0 : (c1 == 2 ? 1 : 2), a1 = (a + 1) % 3, a2 = (a + 2) % 3;
      if (ax[a] * ax[a] > 0.5f) second = fminf(1.f - mul * bestseg, 2.f * safe_divf(bs[a], fabsf(half[a])));
      else second = -fminf(1.f + mul * bestseg, 2.f * fminf(safe_divf(bs[a1], fabsf(half[a1])), safe_divf(bs[a2], fabsf(half[a2]))));
      second *= mul;
    }
  } else if (cltype >= 0) {  // the interior of a box edge is closest
    const int c1 = (axisdir ^ clcorner) & (7 - (1 << cledge));
    if (c1 == 1 || c1 == 2 || c1 == 4) {
      const int a = cledge;
      int a1 = (a + 1) % 3, a2 = (a + 2) % 3;
      if (fabsf(ax[a1]) > fabsf(ax[a2])) a1 = a2;
      a2 = 3 - a - a1;
      float mul;
      if (c1 & (1 << a2)) { mul = 1.f; second = 1.f - bestseg; } else { mul = -1.f; second = 1.f + bestseg; }
      second = fminf(2.f * safe_divf(bs[a2], fabsf(half[a2])), second);
      const float e2 = (((axisdir >> a) & 1) == ((c1 >> a2) & 1)) ? 1.f - bestbox : 1.f + bestbox;
      second = fminf(bs[a] * safe_divf(e2, fabsf(half[a])), second) * mul;
    }
  } else if (clface != -1) {  // a segment end is closest to a face, from outside
    const float mul = cltype == -3 ? 1.f : -1.f;
    second = 2.f;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      if (i == clface) continue;
      const float t1 = pl[i] - half[i] * mul, har = safe_divf(mul, half[i]);
      float e1 = (bs[i] - t1) * har;
      if (0.f < e1 && e1 < second) second = e1;
      e1 = (-bs[i] - t1) * har;
      if (0.f < e1 && e1 < second) second = e1;
    }
    second *= mul;
  }
  t[0] = bestseg;
  t[1] = second + bestseg;
  return second > -3.f ? 2 : 1;
}

// ---- box-box (collision_primitive_core.py:648-1099), float restatement of the routine in oracle/mjoracle.c -------
__device__ __forceinline__ void bb_mat_vec(float* r, const float* m, const float* v) {
#pragma unroll
  for (int i = 0; i < 3; i++) r[i] = m[3 * i] * v[0] + m[3 * i + 1] * v[1] + m[3 * i + 2] * v[2];
}
__device__ __forceinline__ float bb_dot3(const float* a, const float* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ float bb_norm3(const float* a) { return sqrtf(bb_dot3(a, a)); }
__device__ __forceinline__ void m3_mul(float* out, const float* a, const float* b) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) out[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
}
__device__ __forceinline__ void m3_tr(float* out, const float* a) {
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) out[3 * i + j] = a[3 * j + i];
}
// signed permutation that turns face f (0..2 = +x,+y,+z, 3..5 = -x,-y,-z) into the +z face (:616-644)
__device__ __forceinline__ void face_rotation(float* m, int f) {
  const int col[6][3] = {{2, 1, 0}, {0, 2, 1}, {0, 1, 2}, {2, 1, 0}, {0, 2, 1}, {0, 1, 2}};
  const float sgn[6][3] = {{-1.f, 1.f, 1.f}, {1.f, -1.f, 1.f}, {1.f, 1.f, 1.f}, {1.f, 1.f, -1.f}, {1.f, 1.f, -1.f}, {-1.f, 1.f, -1.f}};
  for (int i = 0; i < 9; i++) m[i] = 0.f;
  for (int i = 0; i < 3; i++) m[3 * i + col[f][i]] = sgn[f][i];
}
static __device__ __noinline__ int c_box_box(const float* p1, const float* R1, const float* s1, const float* p2, const float* R2, const float* s2, float margin,
                                      float* dist, float* pos, float* nrm) {  // pos: [8][3]
  float R1t[9], R2t[9], rot21[9], rot12[9], a21[9], a12[9], pos21[3], pos12[3], d21[3], d12[3], plen1[3], plen2[3];
  m3_tr(R1t, R1); m3_tr(R2t, R2);
  for (int i = 0; i < 3; i++) { d21[i] = p2[i] - p1[i]; d12[i] = p1[i] - p2[i]; }
  bb_mat_vec(pos21, R1t, d21); bb_mat_vec(pos12, R2t, d12);
  m3_mul(rot21, R1t, R2); m3_tr(rot12, rot21);
  for (int i = 0; i < 9; i++) { a21[i] = fabsf(rot21[i]); a12[i] = fabsf(rot12[i]); }
  bb_mat_vec(plen2, a21, s2); bb_mat_vec(plen1, a12, s1);
  float separation = margin + 3.0f * (s1[0] + s2[0]) + 3.0f * (s1[1] + s2[1]) + 3.0f * (s1[2] + s2[2]);
  int code = -1;
  for (int i = 0; i < 3; i++) { // face normals of both boxes
    float c1 = -fabsf(pos21[i]) + s1[i] + plen2[i], c2 = -fabsf(pos12[i]) + s2[i] + plen1[i];
    if (c1 < -margin || c2 < -margin) return 0;
    if (c1 < separation) { separation = c1; code = i + 3 * (pos21[i] < 0); }
    if (c2 < separation) { separation = c2; code = i + 3 * (pos12[i] < 0) + 6; }
  }
  float clnorm[3] = {0, 0, 0};
  int inv = 0, cle1 = 0, cle2 = 0;
  for (int i = 0; i < 3; i++) { // edge i of box 1 x edge j of box 2
    for (int j = 0; j < 3; j++) {
      const float* a = rot12 + 3 * j; // axis j of box 2 in the frame of box 1
      float cr[3];
      if (i == 0) { cr[0] = 0; cr[1] = -a[2]; cr[2] = a[1]; }
      else if (i == 1) { cr[0] = a[2]; cr[1] = 0; cr[2] = -a[0]; }
      else { cr[0] = -a[1]; cr[1] = a[0]; cr[2] = 0; }
      float len = bb_norm3(cr);
      if (len < MJ_MINVAL) continue;
      for (int k = 0; k < 3; k++) cr[k] /= len;
      float bd = bb_dot3(pos21, cr), c3 = 0;
      for (int k = 0; k < 3; k++) {
        if (k != i) c3 += s1[k] * fabsf(cr[k]);
        if (k != j) c3 += s2[k] * a21[3 * i + 3 - k - j] / len;
      }
      c3 -= fabsf(bd);
      if (c3 < -margin) return 0;
      if (c3 < separation * (1.0f - 1e-12f)) {
        separation = c3;
        cle1 = cle2 = 0;
        for (int k = 0; k < 3; k++) {
          if (k != i && ((cr[k] > 0) ^ (bd < 0))) cle1 += 1 << k;
          if (k != j && ((rot21[3 * i + 3 - k - j] > 0) ^ (bd < 0) ^ ((k - j + 3) % 3 == 1))) cle2 += 1 << k;
        }
        code = 12 + 3 * i + j;
        clnorm[0] = cr[0]; clnorm[1] = cr[1]; clnorm[2] = cr[2];
        inv = bd < 0;
      }
    }
  }
  if (code == -1) return 0;
  float pts[8][3], depth[8], rw[9], pw[3], rotmore[9], rmt[9], hz;
  int n = 0;
  for (int i = 0; i < 8; i++) pts[i][0] = pts[i][1] = pts[i][2] = 0.f;
  if (code < 12) { // a face of box `bi` is the reference face; the other box is incident
    int bi = code / 6;
    face_rotation(rotmore, code % 6);
    float r[9], rt[9], p[3], ss[3], tmp[3];
    m3_mul(r, rotmore, bi ? rot12 : rot21);
    bb_mat_vec(p, rotmore, bi ? pos12 : pos21);
    bb_mat_vec(tmp, rotmore, bi ? s2 : s1);
    for (int i = 0; i < 3; i++) ss[i] = fabsf(tmp[i]);
    const float* s = bi ? s1 : s2;
    m3_tr(rt, r);
    float lx = ss[0], ly = ss[1];
    hz = ss[2];
    p[2] -= hz;
    int clcorner = 0, dirs = 0;
    for (int i = 0; i < 3; i++) if (r[6 + i] < 0) clcorner += 1 << i;
    float lp[3] = {p[0], p[1], p[2]}, cn[2][3] = {{0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < 3; i++)
      for (int k = 0; k < 3; k++) lp[k] += rt[3 * i + k] * s[i] * ((clcorner & (1 << i)) ? 1.0f : -1.0f);
    for (int i = 0; i < 3; i++) {
      if (fabsf(r[6 + i]) < 0.5f) {
        for (int k = 0; k < 3; k++) cn[dirs ? 1 : 0][k] = rt[3 * i + k] * s[i] * ((clcorner & (1 << i)) ? -2.0f : 2.0f);
        dirs++;
      }
    }
    for (int i = 0; i < dirs * dirs; i++) { // incident edges against the sides of the reference rectangle
      for (int q = 0; q < 2; q++) {
        float lav[3], lbv[3];
        for (int k = 0; k < 3; k++) {
          lav[k] = lp[k] + (i < 2 ? 0.0f : (i == 2 ? cn[0][k] : cn[1][k]));
          lbv[k] = (i == 0 || i == 3) ? cn[0][k] : cn[1][k];
        }
        if (fabsf(lbv[q]) > MJ_MINVAL) {
          float br = 1.0f / lbv[q];
          for (int j = -1; j <= 1; j += 2) {
            float l = ss[q] * j, c1 = (l - lav[q]) * br;
            if (c1 < 0 || c1 > 1) continue;
            float c2 = lav[1 - q] + lbv[1 - q] * c1;
            if (fabsf(c2) > ss[1 - q]) continue;
            if (n < 8) { for (int k = 0; k < 3; k++) pts[n][k] = lav[k] + c1 * lbv[k]; n++; }
          }
        }
      }
    }
    if (dirs == 2) { // corners of the reference rectangle inside the incident face
      float ax = cn[0][0], bx = cn[1][0], ay = cn[0][1], by = cn[1][1], C = safe_divf(1.0f, ax * by - bx * ay);
      for (int i = 0; i < 4; i++) {
        float llx = (i / 2) ? lx : -lx, lly = (i % 2) ? ly : -ly, x = llx - lp[0], y = lly - lp[1];
        float u = (x * by - y * bx) * C, v = (y * ax - x * ay) * C;
        if (u > 0 && v > 0 && u < 1 && v < 1 && n < 8) {
          pts[n][0] = llx; pts[n][1] = lly; pts[n][2] = lp[2] + u * cn[0][2] + v * cn[1][2];
          n++;
        }
      }
    }
    for (int i = 0; i < (1 << dirs); i++) { // incident corners inside the reference rectangle
      float t[3];
      for (int k = 0; k < 3; k++) t[k] = lp[k] + (float)(i & 1) * cn[0][k] + (float)((i & 2) != 0) * cn[1][k];
      if (t[0] > -lx && t[0] < lx && t[1] > -ly && t[1] < ly && n < 8) { pts[n][0] = t[0]; pts[n][1] = t[1]; pts[n][2] = t[2]; n++; }
    }
    int m = n;
    n = 0;
    for (int i = 0; i < m; i++) {
      if (pts[i][2] > margin) continue;
      if (i != n) { pts[n][0] = pts[i][0]; pts[n][1] = pts[i][1]; pts[n][2] = pts[i][2]; }
      depth[n] = pts[n][2];
      pts[n][2] *= 0.5f;
      n++;
    }
    m3_tr(rmt, rotmore);
    m3_mul(rw, bi ? R2 : R1, rmt);
    { const float* pp = bi ? p2 : p1; pw[0] = pp[0]; pw[1] = pp[1]; pw[2] = pp[2]; }
    for (int k = 0; k < 3; k++) nrm[k] = (bi ? -1.0f : 1.0f) * rw[3 * k + 2];
  } else { // edge e1 of box 1 against edge e2 of box 2
    int e1 = (code - 12) / 3, e2 = (code - 12) % 3;
    int ax1 = 1 - (e2 & 1), ax2 = 2 - (e2 & 2), pax1 = 1 - (e1 & 1), pax2 = 2 - (e1 & 2);
    if (a21[3 * e1 + ax1] < a21[3 * e1 + ax2]) { int t = ax1; ax1 = ax2; ax2 = t; }
    if (a12[3 * e2 + pax1] < a12[3 * e2 + pax2]) { int t = pax1; pax1 = pax2; pax2 = t; }
    (void)pax1;
    face_rotation(rotmore, (cle1 & (1 << pax2)) ? pax2 : pax2 + 3);
    m3_tr(rmt, rotmore);
    float p[3], rn[3], r[9], rt[9], s[3], tmp[3];
    bb_mat_vec(p, rotmore, pos21); bb_mat_vec(rn, rotmore, clnorm); m3_mul(r, rotmore, rot21); m3_tr(rt, r);
    bb_mat_vec(tmp, rmt, s1);
    for (int i = 0; i < 3; i++) s[i] = fabsf(tmp[i]);
    float lx = s[0], ly = s[1];
    hz = s[2];
    p[2] -= hz;
    float sg1 = (cle2 & (1 << ax1)) ? 1.0f : -1.0f, sg2 = (cle2 & (1 << ax2)) ? 1.0f : -1.0f, pu[4][3];
    for (int k = 0; k < 3; k++) { // the face of box 2 nearest to box 1: two edges parallel to e2
      float e = rt[3 * e2 + k] * s2[e2];
      float q0 = p[k] + rt[3 * ax1 + k] * s2[ax1] * sg1 + rt[3 * ax2 + k] * s2[ax2] * sg2;
      float q2 = p[k] - rt[3 * ax1 + k] * s2[ax1] * sg1 + rt[3 * ax2 + k] * s2[ax2] * sg2;
      pts[0][k] = q0 + e; pts[1][k] = q0 - e; pts[2][k] = q2 + e; pts[3][k] = q2 - e;
    }
    float alp[3], acn1[3], acn2[3];
    for (int k = 0; k < 3; k++) { alp[k] = pts[0][k]; acn1[k] = pts[1][k] - pts[0][k]; acn2[k] = pts[2][k] - pts[0][k]; }
    if (fabsf(rn[2]) < MJ_MINVAL) return 0;
    float sgn = inv ? -1.0f : 1.0f, innorm = sgn / rn[2];
    for (int i = 0; i < 4; i++) { // project along the separating axis onto the reference plane
      pu[i][0] = pts[i][0]; pu[i][1] = pts[i][1]; pu[i][2] = pts[i][2];
      float sc = pts[i][2] * sgn * innorm;
      for (int k = 0; k < 3; k++) pts[i][k] -= rn[k] * sc;
    }
    float plp[3], pcn1[3], pcn2[3];
    for (int k = 0; k < 3; k++) { plp[k] = pts[0][k]; pcn1[k] = pts[1][k] - pts[0][k]; pcn2[k] = pts[2][k] - pts[0][k]; }
    for (int i = 0; i < 4; i++) {
      for (int q = 0; q < 2; q++) {
        float la = plp[q] + (i < 2 ? 0.0f : (i == 2 ? pcn1[q] : pcn2[q])), lb = (i == 0 || i == 3) ? pcn1[q] : pcn2[q];
        float lc = plp[1 - q] + (i < 2 ? 0.0f : (i == 2 ? pcn1[1 - q] : pcn2[1 - q])), ld = (i == 0 || i == 3) ? pcn1[1 - q] : pcn2[1 - q];
        float lua[3], lub[3];
        for (int k = 0; k < 3; k++) {
          lua[k] = alp[k] + (i < 2 ? 0.0f : (i == 2 ? acn1[k] : acn2[k]));
          lub[k] = (i == 0 || i == 3) ? acn1[k] : acn2[k];
        }
        if (fabsf(lb) > MJ_MINVAL) {
          float br = 1.0f / lb;
          for (int j = -1; j <= 1; j += 2) {
            if (n == 8) break;
            float l = s[q] * j, c1 = (l - la) * br;
            if (c1 < 0 || c1 > 1) continue;
            float c2 = lc + ld * c1;
            if (fabsf(c2) > s[1 - q]) continue;
            if ((lua[2] + lub[2] * c1) * innorm > margin) continue;
            for (int k = 0; k < 3; k++) pts[n][k] = lua[k] * 0.5f + c1 * lub[k] * 0.5f;
            pts[n][q] += 0.5f * l;
            pts[n][1 - q] += 0.5f * c2;
            depth[n] = pts[n][2] * innorm * 2.0f;
            n++;
          }
        }
      }
    }
    int nl = n;
    float ax = pcn1[0], bx = pcn2[0], ay = pcn1[1], by = pcn2[1], C = safe_divf(1.0f, ax * by - bx * ay);
    for (int i = 0; i < 4; i++) { // corners of the reference rectangle under the projected face
      if (n == 8) break;
      float llx = (i / 2) ? lx : -lx, lly = (i % 2) ? ly : -ly, x = llx - plp[0], y = lly - plp[1];
      float u = (x * by - y * bx) * C, v = (y * ax - x * ay) * C;
      if (nl == 0) { if ((u < 0 || u > 1) && (v < 0 || v > 1)) continue; }
      else if (u < 0 || v < 0 || u > 1 || v > 1) continue;
      u = clampf(u, 0.0f, 1.0f); v = clampf(v, 0.0f, 1.0f);
      float w = 1.0f - u - v, vt[3], tc1 = 0, corner[3] = {llx, lly, 0.0f};
      for (int k = 0; k < 3; k++) { vt[k] = pu[0][k] * w + pu[1][k] * u + pu[2][k] * v; tc1 += (corner[k] - vt[k]) * (corner[k] - vt[k]); }
      if (vt[2] > 0 && tc1 > margin * margin) continue;
      for (int k = 0; k < 3; k++) pts[n][k] = 0.5f * (corner[k] + vt[k]);
      depth[n] = sqrtf(tc1) * (vt[2] < 0 ? -1.0f : 1.0f);
      n++;
    }
    int nf = n;
    for (int i = 0; i < 4; i++) { // corners of the face of box 2 over the reference rectangle
      if (n >= 8) break;
      float x = pu[i][0], y = pu[i][1];
      if (nl == 0 && nf != 0) { if ((x < -lx || x > lx) && (y < -ly || y > ly)) continue; }
      else if (x < -lx || x > lx || y < -ly || y > ly) continue;
      float c1 = 0, tp[3] = {pu[i][0], pu[i][1], 0.0f};
      for (int j = 0; j < 2; j++) {
        if (pu[i][j] < -s[j]) { c1 += (pu[i][j] + s[j]) * (pu[i][j] + s[j]); tp[j] = -s[j] * 0.5f; }
        else if (pu[i][j] > s[j]) { c1 += (pu[i][j] - s[j]) * (pu[i][j] - s[j]); tp[j] = s[j] * 0.5f; }
      }
      c1 += pu[i][2] * innorm * pu[i][2] * innorm;
      if (pu[i][2] > 0 && c1 > margin * margin) continue;
      for (int k = 0; k < 3; k++) pts[n][k] = (tp[k] + pu[i][k]) * 0.5f;
      depth[n] = sqrtf(c1) * (pu[i][2] < 0 ? -1.0f : 1.0f);
      n++;
    }
    m3_mul(rw, R1, rmt);
    pw[0] = p1[0]; pw[1] = p1[1]; pw[2] = p1[2];
    float t[3];
    bb_mat_vec(t, rw, rn);
    for (int k = 0; k < 3; k++) nrm[k] = sgn * t[k];
  }
  for (int i = 0; i < n; i++) {
    pts[i][2] += hz;
    bb_mat_vec(pos + 3 * i, rw, pts[i]);
    for (int k = 0; k < 3; k++) pos[3 * i + k] += pw[k];
    dist[i] = depth[i];
  }
  return n;
}

// Number of candidate contacts of a pair type (collision_primitive.py wrappers).
__device__ __forceinline__ int pair_ncand(int t1, int t2) {
  if (t1 == GEOM_HFIELD) return 4;
  if (t1 == GEOM_PLANE) return t2 == GEOM_SPHERE || t2 == GEOM_ELLIPSOID ? 1 : (t2 == GEOM_CAPSULE ? 2 : (t2 == GEOM_BOX ? 8 : (t2 == GEOM_CYLINDER || t2 == GEOM_MESH ? 4 : 0)));
  if (is_ccd_pair(t1, t2)) return 1;
  if (t1 == GEOM_SPHERE) return (t2 == GEOM_SPHERE || t2 == GEOM_CAPSULE || t2 == GEOM_BOX || t2 == GEOM_CYLINDER) ? 1 : 0;
  if (t1 == GEOM_CAPSULE) return (t2 == GEOM_CAPSULE || t2 == GEOM_BOX) ? 2 : 0;
  if (t1 == GEOM_BOX && t2 == GEOM_BOX) return 8;
  return 0;
}

// k-th candidate contact of the pair: distance, position, frame.  `r2` = rotation of geom 2.
template <int FEAT>
__device__ __forceinline__ void pair_contact(const PairGeom& g, const float* r2, float margin, int k, float& dist, V3& pos, M3& frame) {
  dist = INFINITY;
  pos = V3{0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 9; i++) frame.m[i] = 0.f;
  if (g.t1 == GEOM_PLANE) {
    const V3 n = g.n1;
    if (g.t2 == GEOM_SPHERE) {  // core :107
      dist = dot(g.p2 - g.p1, n) - g.s2.x;
      pos = g.p2 - n * (g.s2.x + 0.5f * dist);
      frame = make_frame(n);
    } else if (g.t2 == GEOM_CAPSULE) {  // core :312-361
      V3 b = g.n2 - n * dot(n, g.n2);
      const float nb = norm(b);
      if (nb != 0.f) b = b * (1.0f / nb);
      if (nb < 0.5f) b = (-0.5f < n.y && n.y < 0.5f) ? V3{0.f, 1.f, 0.f} : V3{0.f, 0.f, 1.f};
      const V3 c = cross(n, b);
      frame.m[0] = n.x; frame.m[1] = n.y; frame.m[2] = n.z; frame.m[3] = b.x; frame.m[4] = b.y; frame.m[5] = b.z;
      frame.m[6] = c.x; frame.m[7] = c.y; frame.m[8] = c.z;
      const V3 e = g.p2 + g.n2 * (k == 0 ? g.s2.y : -g.s2.y);
      dist = dot(e - g.p1, n) - g.s2.x;
      pos = e - n * (g.s2.x + 0.5f * dist);
    } else if ((FEAT & COLF_CONVEX) && g.t2 == GEOM_ELLIPSOID) {  // core :365-392
      V3 ld = V3{(r2[0] * n.x + r2[3] * n.y + r2[6] * n.z) * g.s2.x, (r2[1] * n.x + r2[4] * n.y + r2[7] * n.z) * g.s2.y,
                 (r2[2] * n.x + r2[5] * n.y + r2[8] * n.z) * g.s2.z};
      const float nn = norm(ld);
      ld = nn > 0.f ? ld * (-1.0f / nn) : V3{0.f, 0.f, 0.f};
      const V3 loc = V3{ld.x * g.s2.x, ld.y * g.s2.y, ld.z * g.s2.z};
      const V3 pe = g.p2 + V3{r2[0] * loc.x + r2[1] * loc.y + r2[2] * loc.z, r2[3] * loc.x + r2[4] * loc.y + r2[5] * loc.z,
                              r2[6] * loc.x + r2[7] * loc.y + r2[8] * loc.z};
      dist = dot(n, pe - g.p1);
      pos = pe - n * (dist * 0.5f);
      frame = make_frame(n);
    } else if ((FEAT & COLF_CONVEX) && g.t2 == GEOM_CYLINDER) {  // core :519-612
      V3 axis = g.n2;
      float prjaxis = dot(n, axis);
      if (prjaxis > 0.f) { axis = axis * -1.0f; prjaxis = -prjaxis; }
      const float dist0 = dot(g.p2 - g.p1, n), radius = g.s2.x, hh = g.s2.y;
      V3 vec = axis * prjaxis - n;
      const float len_sqr = dot(vec, vec);
      vec = len_sqr >= 1e-12f ? vec * safe_divf(radius, sqrtf(len_sqr)) : V3{radius, 0.f, 0.f};
      const float prjvec = dot(vec, n);
      axis = axis * hh;
      prjaxis *= hh;
      if (k < 2) {
        dist = dist0 + (k == 0 ? prjaxis : -prjaxis) + prjvec;
        pos = g.p2 + vec + (k == 0 ? axis : axis * -1.0f) - n * (dist * 0.5f);
      } else {
        dist = dist0 + prjaxis - prjvec * 0.5f;
        V3 vec1 = cross(vec, axis);
        const float nv1 = norm(vec1);
        vec1 = nv1 > 0.f ? vec1 * (radius * sqrtf(3.0f) * 0.5f / nv1) : V3{0.f, 0.f, 0.f};
        pos = g.p2 + (k == 2 ? vec1 : vec1 * -1.0f) + axis - vec * 0.5f - n * (dist * 0.5f);
      }
      frame = make_frame(n);
    } else {  // plane-box, core :396-443
      const V3 corner = V3{(k & 1) ? g.s2.x : -g.s2.x, (k & 2) ? g.s2.y : -g.s2.y, (k & 4) ? g.s2.z : -g.s2.z};
      const V3 cw = V3{r2[0] * corner.x + r2[1] * corner.y + r2[2] * corner.z, r2[3] * corner.x + r2[4] * corner.y + r2[5] * corner.z,
                       r2[6] * corner.x + r2[7] * corner.y + r2[8] * corner.z};
      dist = dot(g.p2 - g.p1, n) + dot(n, cw);
      pos = cw + g.p2 - n * (0.5f * dist);
      frame = make_frame(n);
    }
    return;
  }
  V3 nrm = V3{1.f, 0.f, 0.f};
  if (g.t1 == GEOM_SPHERE && g.t2 == GEOM_SPHERE) {
    c_sphere_sphere(g.p1, g.s1.x, g.p2, g.s2.x, dist, pos, nrm);
  } else if ((FEAT & COLF_BOX) && g.t2 == GEOM_BOX) {  // sphere-box :1103, capsule-box :1158 (sphere tests at the support points)
    const float bs[3] = {g.s2.x, g.s2.y, g.s2.z};
    V3 centre = g.p1;
    if (g.t1 == GEOM_CAPSULE) {
      float t[2];
      const int n = capsule_box_points(g.p1, g.n1, g.s1.y, g.p2, r2, bs, t);
      if (k >= n) return;
      centre = g.p1 + g.n1 * (g.s1.y * t[k]);
    }
    c_sphere_box(centre, g.s1.x, g.p2, r2, bs, dist, pos, nrm);
  } else if ((FEAT & COLF_CONVEX) && g.t1 == GEOM_SPHERE && g.t2 == GEOM_CYLINDER) {  // core :447-515
    const V3 vec = g.p1 - g.p2;
    const float x = dot(vec, g.n2);
    const V3 aproj = g.n2 * x, pproj = vec - aproj;
    const float psq = dot(pproj, pproj);
    bool side = fabsf(x) < g.s2.y, cap = psq < g.s2.x * g.s2.x;
    if (side && cap) {
      if (g.s2.y - fabsf(x) < g.s2.x - sqrtf(psq)) side = false;
      else cap = false;
    }
    if (side) {
      c_sphere_sphere(g.p1, g.s1.x, g.p2 + aproj, g.s2.x, dist, pos, nrm);
    } else if (cap) {
      const float sg = x > 0.f ? 1.f : -1.f;
      const V3 pn = g.n2 * sg, pc = g.p2 + pn * g.s2.y;
      dist = dot(g.p1 - pc, pn) - g.s1.x;
      pos = g.p1 - pn * (g.s1.x + 0.5f * dist);
      nrm = pn * -1.0f;
    } else {
      const float inv = safe_divf(1.0f, sqrtf(psq)), sg = x < 0.f ? -1.f : 1.f;
      c_sphere_sphere(g.p1, g.s1.x, g.p2 + g.n2 * (sg * g.s2.y) + pproj * (g.s2.x * inv), 0.f, dist, pos, nrm);
    }
  } else if (g.t1 == GEOM_SPHERE) {  // sphere-capsule, core :147-178
    const V3 a = g.p2 - g.n2 * g.s2.y, ab = g.n2 * (2.0f * g.s2.y);
    const float t = clampf(dot(g.p1 - a, ab) / (dot(ab, ab) + 1e-6f), 0.f, 1.f);
    c_sphere_sphere(g.p1, g.s1.x, a + ab * t, g.s2.x, dist, pos, nrm);
  } else {  // capsule-capsule, core :182-308
    const V3 ax1 = g.n1 * g.s1.y, ax2 = g.n2 * g.s2.y, dif = g.p1 - g.p2;
    const float ma = dot(ax1, ax1), mb = -dot(ax1, ax2), mc = dot(ax2, ax2), u = -dot(ax1, dif), v = dot(ax2, dif);
    const float det = ma * mc - mb * mb;
    if (fabsf(det) >= MJ_MINVAL) {
      if (k > 0) return;
      float x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
      if (x1 > 1.f) { x1 = 1.f; x2 = (v - mb) / mc; } else if (x1 < -1.f) { x1 = -1.f; x2 = (v + mb) / mc; }
      if (x2 > 1.f) { x2 = 1.f; x1 = clampf((u - mb) / ma, -1.f, 1.f); }
      else if (x2 < -1.f) { x2 = -1.f; x1 = clampf((u + mb) / ma, -1.f, 1.f); }
      float dd;
      V3 pp;
      c_sphere_sphere(g.p1 + ax1 * x1, g.s1.x, g.p2 + ax2 * x2, g.s2.x, dd, pp, nrm);
      if (dd <= margin) { dist = dd; pos = pp; }
    } else {  // parallel axes: first two of the four end-point tests that pass the margin check
      int cnt = 0;
      for (int e = 0; e < 4; e++) {
        float x1, x2;
        if (e == 0) { x1 = 1.f; x2 = clampf((v - mb) / mc, -1.f, 1.f); }
        else if (e == 1) { x1 = -1.f; x2 = clampf((v + mb) / mc, -1.f, 1.f); }
        else if (e == 2) { x2 = 1.f; x1 = clampf((u - mb) / ma, -1.f, 1.f); }
        else { x2 = -1.f; x1 = clampf((u + mb) / ma, -1.f, 1.f); }
        float dd;
        V3 pp, nn;
        c_sphere_sphere(g.p1 + ax1 * x1, g.s1.x, g.p2 + ax2 * x2, g.s2.x, dd, pp, nn);
        if (dd <= margin) {
          if (cnt == k) { dist = dd; pos = pp; nrm = nn; }
          cnt++;
          if (cnt > k) break;
        }
      }
    }
    if (dist == INFINITY) return;
  }
  frame = make_frame(nrm);
}

struct ContactParams {
  float margin, gap, friction[5], solref[2], solreffriction[2], solimp[5];
  int condim;
};

// collision_core.py:236-341
__device__ __forceinline__ void contact_params(const KParams& K, const unsigned char* mb, int g1, int g2, int pairid, ContactParams& c) {
  if (pairid > -1) {
    c.margin = MR(pair_margin)[pairid];
    c.gap = MR(pair_gap)[pairid];
    c.condim = MI(pair_dim)[pairid];
#pragma unroll
    for (int i = 0; i < 5; i++) { c.friction[i] = MR(pair_friction)[5 * pairid + i]; c.solimp[i] = MR(pair_solimp)[5 * pairid + i]; }
#pragma unroll
    for (int i = 0; i < 2; i++) { c.solref[i] = MR(pair_solref)[2 * pairid + i]; c.solreffriction[i] = MR(pair_solreffriction)[2 * pairid + i]; }
  } else {
    const float sm1 = MR(geom_solmix)[g1], sm2 = MR(geom_solmix)[g2];
    const int p1 = MI(geom_priority)[g1], p2 = MI(geom_priority)[g2];
    const V3 f1 = ld3(MR(geom_friction) + 3 * g1), f2 = ld3(MR(geom_friction) + 3 * g2);
    float mix;
    V3 f;
    if (p1 > p2) { mix = 1.f; c.condim = MI(geom_condim)[g1]; f = f1; }
    else if (p2 > p1) { mix = 0.f; c.condim = MI(geom_condim)[g2]; f = f2; }
    else {
      mix = safe_div(sm1, sm1 + sm2);
      if (sm1 < MJ_MINVAL && sm2 < MJ_MINVAL) mix = 0.5f;
      if (sm1 < MJ_MINVAL && sm2 >= MJ_MINVAL) mix = 0.f;
      if (sm1 >= MJ_MINVAL && sm2 < MJ_MINVAL) mix = 1.f;
      c.condim = max(MI(geom_condim)[g1], MI(geom_condim)[g2]);
      f = V3{fmaxf(f1.x, f2.x), fmaxf(f1.y, f2.y), fmaxf(f1.z, f2.z)};
    }
    c.friction[0] = f.x; c.friction[1] = f.x; c.friction[2] = f.y; c.friction[3] = f.z; c.friction[4] = f.z;
    const float* sr1 = MR(geom_solref) + 2 * g1;
    const float* sr2 = MR(geom_solref) + 2 * g2;
    if (sr1[0] > 0.f && sr2[0] > 0.f) {
      c.solref[0] = mix * sr1[0] + (1.f - mix) * sr2[0];
      c.solref[1] = mix * sr1[1] + (1.f - mix) * sr2[1];
    } else {
      c.solref[0] = fminf(sr1[0], sr2[0]);
      c.solref[1] = fminf(sr1[1], sr2[1]);
    }
    c.solreffriction[0] = c.solreffriction[1] = 0.f;
#pragma unroll
    for (int i = 0; i < 5; i++) c.solimp[i] = mix * MR(geom_solimp)[5 * g1 + i] + (1.f - mix) * MR(geom_solimp)[5 * g2 + i];
    c.margin = MR(geom_margin)[g1] + MR(geom_margin)[g2];
    c.gap = MR(geom_gap)[g1] + MR(geom_gap)[g2];
  }
#pragma unroll
  for (int i = 0; i < 5; i++) c.friction[i] = fmaxf(MJ_MINMU, c.friction[i]);
}

// Narrowphase for up to 32 queued pairs (one lane each); appends contacts to the pool.
template <int FEAT>
__device__ __forceinline__ void narrowphase_flush(const KParams& K, const unsigned char* mb, int w, int lane, const int* queue, int count,
                                                  int& nw) {
  const bool has = lane < count;
  int g1 = 0, g2 = 0, pid0 = -2, pid1 = -1, ncand = 0, nact = 0;
  PairGeom g;
  ContactParams cp;
  const float* r2 = nullptr;
  float bdist[(FEAT & COLF_BOX) ? 8 : 1], bpos[(FEAT & COLF_BOX) ? 24 : 3], bnrm[3] = {1.f, 0.f, 0.f};  // box-box results
  float hnrm[(FEAT & COLF_CONVEX) ? 12 : 1];  // height-field contacts carry one normal each
  bool pernormal = false;
  int nbb = 0;
  bool boxbox = false, mesh_ccd = false;
  if (has) {
    const int p = queue[lane];
    const int ga = MIC(nxn_pair)[2 * p], gb = MIC(nxn_pair)[2 * p + 1];
    pid0 = MIC(nxn_pairid)[2 * p];
    pid1 = MIC(nxn_pairid)[2 * p + 1];
    const bool swap = MI(geom_type)[ga] > MI(geom_type)[gb];  // collision_driver.py:351
    g1 = swap ? gb : ga;
    g2 = swap ? ga : gb;
    g.t1 = MI(geom_type)[g1];
    g.t2 = MI(geom_type)[g2];
    g.p1 = ld3(DW(geom_xpos) + 3 * g1);
    g.p2 = ld3(DW(geom_xpos) + 3 * g2);
    const float* r1 = DW(geom_xmat) + 9 * g1;
    r2 = DW(geom_xmat) + 9 * g2;
    g.n1 = V3{r1[2], r1[5], r1[8]};
    g.n2 = V3{r2[2], r2[5], r2[8]};
    g.s1 = ld3(MR(geom_size) + 3 * g1);
    g.s2 = ld3(MR(geom_size) + 3 * g2);
    contact_params(K, mb, g1, g2, pid0, cp);
    ncand = pair_ncand(g.t1, g.t2);
    const bool nativeccd = (FEAT & COLF_CONVEX) && !(K.o.disableflags & DSBL_NATIVECCD);
    if ((FEAT & COLF_BOX) && g.t1 == GEOM_BOX && g.t2 == GEOM_BOX && !nativeccd) {  // all (<= 8) contacts of the pair in one evaluation
      const float pa[3] = {g.p1.x, g.p1.y, g.p1.z}, pb[3] = {g.p2.x, g.p2.y, g.p2.z};
      const float sa[3] = {g.s1.x, g.s1.y, g.s1.z}, sb[3] = {g.s2.x, g.s2.y, g.s2.z};
      nbb = c_box_box(pa, DW(geom_xmat) + 9 * g1, sa, pb, r2, sb, cp.margin, bdist, bpos, bnrm);
      boxbox = true;
    }
    if ((FEAT & COLF_CONVEX) && (FEAT & COLF_BOX) && g.t1 == GEOM_HFIELD) {
      c_hfield_pair(K, mb, w, g1, g2, cp.margin, bdist, bpos, hnrm);
      nbb = 4;
      boxbox = true;
      pernormal = true;
    } else if ((FEAT & COLF_CONVEX) && g.t2 == GEOM_MESH && is_ccd_pair(g.t1, g.t2)) {
      mesh_ccd = true;  // evaluated below by the whole warp, one pair at a time
    } else if ((FEAT & COLF_CONVEX) && is_ccd_pair(g.t1, g.t2, nativeccd)) {  // GJK / EPA once per pair; the two passes below read the result
      const float pa[3] = {g.p1.x, g.p1.y, g.p1.z}, pb[3] = {g.p2.x, g.p2.y, g.p2.z};
      const float sa[3] = {g.s1.x, g.s1.y, g.s1.z}, sb[3] = {g.s2.x, g.s2.y, g.s2.z};
      const float *v1 = nullptr, *v2 = nullptr;
      const int *gr1 = nullptr, *gr2 = nullptr;
      int nv1 = 0, nv2 = 0;
      if (g.t1 == GEOM_MESH) {
        const int id = MI(geom_dataid)[g1];
        v1 = MRC(mesh_vert) + 3 * MI(mesh_vertadr)[id];
        nv1 = MI(mesh_vertnum)[id];
        if (MI(mesh_graphadr)[id] >= 0) gr1 = MIC(mesh_graph) + MI(mesh_graphadr)[id];
      }
      if (g.t2 == GEOM_MESH) {
        const int id = MI(geom_dataid)[g2];
        v2 = MRC(mesh_vert) + 3 * MI(mesh_vertadr)[id];
        nv2 = MI(mesh_vertnum)[id];
        if (MI(mesh_graphadr)[id] >= 0) gr2 = MIC(mesh_graph) + MI(mesh_graphadr)[id];
      }
      nbb = c_ccd_pair(pa, DW(geom_xmat) + 9 * g1, sa, g.t1, v1, nv1, gr1, pb, r2, sb, g.t2, v2, nv2, gr2, cp.margin, K.o.ccd_tolerance,
                       K.o.ccd_iterations, bdist, bpos, bnrm, 0, nullptr);
      boxbox = true;
    }
    if ((FEAT & COLF_CONVEX) && g.t1 == GEOM_PLANE && g.t2 == GEOM_MESH) {
      const float pa[3] = {g.p1.x, g.p1.y, g.p1.z}, pb[3] = {g.p2.x, g.p2.y, g.p2.z};
      const int id = MI(geom_dataid)[g2];
      bnrm[0] = g.n1.x; bnrm[1] = g.n1.y; bnrm[2] = g.n1.z;
      nbb = c_plane_mesh(pa, bnrm, pb, r2, MRC(mesh_vert) + 3 * MI(mesh_vertadr)[id], MI(mesh_vertnum)[id],
                         MI(mesh_graphadr)[id] >= 0 ? MIC(mesh_graph) + MI(mesh_graphadr)[id] : nullptr, bdist, bpos);
      boxbox = true;
    }
  }
  if (FEAT & COLF_CONVEX) {
    // Pairs with a mesh: GJK / EPA by the WHOLE warp, one pair at a time.  Every lane runs the same (warp-uniform) iteration on the
    // same pair and the hull support function searches all vertices with the lanes in parallel (ccd_coop_mesh_support): a mesh
    // pair costs ~50 support calls on hulls of up to ~1200 vertices, and one lane walking the hull's edge graph through global
    // memory for each of them is what made this stage 80 % of the aloha_pot step.
    unsigned todo = __ballot_sync(FULL, has && mesh_ccd);
    while (todo) {  // warp-uniform
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      const int p = queue[src];
      const int ga = MIC(nxn_pair)[2 * p], gb = MIC(nxn_pair)[2 * p + 1];
      const bool swap = MI(geom_type)[ga] > MI(geom_type)[gb];
      const int u1 = swap ? gb : ga, u2 = swap ? ga : gb;
      const int ut1 = MI(geom_type)[u1];
      const float* v1 = nullptr;
      const int* gr1 = nullptr;
      int nv1 = 0;
      if (ut1 == GEOM_MESH) {
        const int id = MI(geom_dataid)[u1];
        v1 = MRC(mesh_vert) + 3 * MI(mesh_vertadr)[id];
        nv1 = MI(mesh_vertnum)[id];
        if (MI(mesh_graphadr)[id] >= 0) gr1 = MIC(mesh_graph) + MI(mesh_graphadr)[id];
      }
      const int id2 = MI(geom_dataid)[u2];
      const float* v2 = MRC(mesh_vert) + 3 * MI(mesh_vertadr)[id2];
      const int* gr2 = MI(mesh_graphadr)[id2] >= 0 ? MIC(mesh_graph) + MI(mesh_graphadr)[id2] : nullptr;
      ContactParams ucp;
      contact_params(K, mb, u1, u2, MIC(nxn_pairid)[2 * p], ucp);
      float tdist = 0.f, tpos[3] = {0.f, 0.f, 0.f}, tnrm[3] = {1.f, 0.f, 0.f};
      const int n = c_ccd_pair(DW(geom_xpos) + 3 * u1, DW(geom_xmat) + 9 * u1, MR(geom_size) + 3 * u1, ut1, v1, nv1, gr1, DW(geom_xpos) + 3 * u2,
                               DW(geom_xmat) + 9 * u2, MR(geom_size) + 3 * u2, GEOM_MESH, v2, MI(mesh_vertnum)[id2], gr2, ucp.margin,
                               K.o.ccd_tolerance, K.o.ccd_iterations, &tdist, tpos, tnrm, 1,
                               reinterpret_cast<CcdPolytope*>(const_cast<int*>(queue) + K.scratch_floats - 64 - CCD_POLY_WORDS));
      if (lane == src) {
        nbb = n;
        boxbox = true;
        bdist[0] = tdist;
#pragma unroll
        for (int i = 0; i < 3; i++) { bpos[i] = tpos[i]; bnrm[i] = tnrm[i]; }
      }
      __syncwarp();
    }
  }
  if (has) {
    for (int k = 0; k < ncand; k++) {  // pass 1: count (write_contact filter, collision_core.py:199-203)
      float dist;
      V3 pos;
      M3 frame;
      if ((FEAT & (COLF_BOX | COLF_CONVEX)) && boxbox) dist = k < nbb ? bdist[k] : INFINITY;
      else pair_contact<FEAT>(g, r2, cp.margin, k, dist, pos, frame);
      const bool active = dist < cp.margin;
      if (!((pid0 == -2 || !active) && pid1 == -1)) nact++;
    }
  }
  const int off = warp_excl_scan(nact, lane);
  const int total = __shfl_sync(FULL, off + nact, 31);
  if (total == 0) return;
  int base = 0;
  if (lane == 0) base = atomicAdd(K.d.nacon, total);
  base = __shfl_sync(FULL, base, 0);
  if (has && nact > 0) {
    const int nmp = K.ms.nmaxpyramid;
    int slot = off;
    for (int k = 0; k < ncand; k++) {  // pass 2: write
      float dist;
      V3 pos;
      M3 frame;
      if ((FEAT & (COLF_BOX | COLF_CONVEX)) && boxbox) {
        dist = k < nbb ? bdist[k] : INFINITY;
        pos = V3{bpos[3 * (k & 7)], bpos[3 * (k & 7) + 1], bpos[3 * (k & 7) + 2]};
        if ((FEAT & COLF_CONVEX) && pernormal) frame = make_frame(V3{hnrm[3 * (k & 3)], hnrm[3 * (k & 3) + 1], hnrm[3 * (k & 3) + 2]});
        else frame = make_frame(V3{bnrm[0], bnrm[1], bnrm[2]});
      } else {
        pair_contact<FEAT>(g, r2, cp.margin, k, dist, pos, frame);
      }
      const bool active = dist < cp.margin;
      if ((pid0 == -2 || !active) && pid1 == -1) continue;
      const int cid = base + slot;
      const int widx = nw + slot;
      slot++;
      if (cid >= K.dsz.naconmax) {
        atomicOr(DW(status), MJB_STATUS_NCON_OVERFLOW);
        continue;
      }
      int type = 0;
      if (pid0 >= -1 && active) type |= 1;
      if (pid1 >= 0) type |= 2;
      K.d.contact_dist[cid] = dist;
      st3(K.d.contact_pos + 3 * cid, pos);
      stm3(K.d.contact_frame + 9 * cid, frame);
      K.d.contact_includemargin[cid] = cp.margin - cp.gap;
#pragma unroll
      for (int i = 0; i < 5; i++) { K.d.contact_friction[5 * cid + i] = cp.friction[i]; K.d.contact_solimp[5 * cid + i] = cp.solimp[i]; }
      K.d.contact_solref[2 * cid] = cp.solref[0];
      K.d.contact_solref[2 * cid + 1] = cp.solref[1];
      K.d.contact_solreffriction[2 * cid] = cp.solreffriction[0];
      K.d.contact_solreffriction[2 * cid + 1] = cp.solreffriction[1];
      K.d.contact_dim[cid] = cp.condim;
      K.d.contact_geom[2 * cid] = g1;
      K.d.contact_geom[2 * cid + 1] = g2;
      K.d.contact_worldid[cid] = w;
      K.d.contact_type[cid] = type;
      K.d.contact_geomcollisionid[cid] = k;
      for (int i = 0; i < nmp; i++) K.d.contact_efc_address[cid * nmp + i] = -1;
      if (widx < K.dsz.nwconmax) DW(wcon)[widx] = cid;
      else atomicOr(DW(status), MJB_STATUS_WCON_OVERFLOW);
    }
  }
  // only the contacts that fitted into the pool enter the world's list (they are the first slots)
  nw += max(0, min(total, K.dsz.naconmax - base));
}

// Per-warp shared scratch of the collision stage, in 32-bit words: the pair queue (64 + 8), then the
// sweep-and-prune workspace (lower, upper, order, range prefix: 4 ngeom words; candidate bit set).
// The two-level broadphase (ngroup > 0) uses the sweep-and-prune workspace for the geoms' world spheres and 4 ngroup more words for
// the groups' (the candidate bit set, unused on that path, follows and is simply large enough).
__host__ __device__ inline int collision_scratch_words(int ngeom, int nxn, int ngroup = 0, bool convex = false) {
  // ... + alignment + [the polytope of the warp-cooperative convex pairs] + the 64-entry candidate queue at the very end
  return ((72 + 4 * ngeom + 4 * ngroup + (nxn + 31) / 32 + 1 + 3) & ~3) + (convex ? CCD_POLY_WORDS : 0) + 64;
}

// Sweep-and-prune candidate set (collision_driver.py:554 sap_broadphase).  The reference projects the bounding
// spheres on a fixed direction (:573), sorts per world by the lower end (tile sort or segmented sort, :458/:617),
// finds per sorted geom the range of followers whose lower end does not exceed its upper end -- plus one
// (_sap_range :411: limit = min(ngeom - 1, first lower > upper)) -- and tests the flattened (geom, follower) work
// items (:485-551).  One warp does the same for its world: rank sort (lane = geom), warp scan of the ranges, lanes
// over the flattened items; items that exist in the statically filtered pair table set a bit of `bits`, which the
// pair loop of s_collision then uses as a pre-filter, so contacts keep the pair-list order of the NXN path.
__device__ __forceinline__ void sap_candidates(const KParams& K, const unsigned char* mb, int w, int lane, int* ws, unsigned* bits) {
  const int ng = K.ms.ngeom, nxn = K.ms.nxn;
  float* lower = reinterpret_cast<float*>(ws);
  float* upper = lower + ng;
  int* order = ws + 2 * ng;
  int* cum = ws + 3 * ng;
  const float inv = rsqrtf(0.5935f * 0.5935f + 0.7790f * 0.7790f + 0.1235f * 0.1235f);
  const V3 dir{0.5935f * inv, 0.7790f * inv, 0.1235f * inv};
  for (int g = lane; g < ng; g += 32) {
    float rb = MR(geom_rbound)[g];
    if (rb == 0.f) rb = MJ_MAXVAL;
    const float radius = rb + MR(geom_margin)[g], c = dot(dir, ld3(DW(geom_xpos) + 3 * g));
    const bool bad = c != c;
    lower[g] = bad ? MJ_MAXVAL : c - radius;
    upper[g] = bad ? MJ_MAXVAL : c + radius;
  }
  for (int i = lane; i < (nxn + 31) / 32; i += 32) bits[i] = 0u;
  __syncwarp();
  for (int g = lane; g < ng; g += 32) {  // stable rank sort by lower end
    const float lg = lower[g];
    int rank = 0;
    for (int h = 0; h < ng; h++) {
      const float lh = lower[h];
      rank += (lh < lg || (lh == lg && h < g)) ? 1 : 0;
    }
    order[rank] = g;
  }
  __syncwarp();
  int running = 0;
  for (int i0 = 0; i0 < ng; i0 += 32) {  // followers per sorted position, inclusive prefix sums into cum[]
    const int i = i0 + lane;
    int range = 0;
    if (i < ng) {
      const float up = upper[order[i]];
      int limit = i + 1;
      while (limit < ng && !(lower[order[limit]] > up)) limit++;
      range = min(limit, ng - 1) - i;
    }
    const int ex = warp_excl_scan(range, lane);
    if (i < ng) cum[i] = running + ex + range;
    running += __shfl_sync(FULL, ex + range, 31);
  }
  __syncwarp();
  for (int k = lane; k < running; k += 32) {  // flattened work items (collision_driver.py:488-506)
    int lo = 0, hi = ng;
    while (lo < hi) {  // first position whose inclusive prefix exceeds k
      const int mid = (lo + hi) >> 1;
      if (cum[mid] > k) hi = mid; else lo = mid + 1;
    }
    const int i = hi, j = i + 1 + k - (i > 0 ? cum[i - 1] : 0);
    int a = order[i], b = order[j];
    if (b < a) { const int t = a; a = b; b = t; }
    const int p = MIC(nxn_lookup)[(a * (2 * ng - a - 3)) / 2 + b - 1];
    if (p >= 0) atomicOr(bits + (p >> 5), 1u << (p & 31));
  }
  __syncwarp();
}

// collision_driver.collision (collision_driver.py:755).  scratch: collision_scratch_words() ints per warp.
template <int FEAT>
__device__ __forceinline__ void s_collision(const KParams& K, const unsigned char* mb, int w, int lane, int* queue) {
  int nw = 0, ncol = 0, qn = 0;
  const bool enabled = K.dsz.naconmax > 0 && !(K.o.disableflags & (DSBL_CONSTRAINT | DSBL_CONTACT));
  if (enabled) {
    const int nxn = K.ms.nxn;
    // the all-features variant (convex models) carries the sweep-and-prune code but runs it only when the option asks for it
    const bool sap = (FEAT & COLF_SAP) != 0 && nxn > 0 && K.o.broadphase != 0;
    unsigned* bits = reinterpret_cast<unsigned*>(queue + 72 + 4 * K.ms.ngeom);
    if (sap) sap_candidates(K, mb, w, lane, queue + 72, bits);
    // one chunk of up to 32 pairs: lane = pair, broadphase filter, ballot-compacted append to the queue, narrowphase when 32 wait
    auto chunk = [&](int p, bool valid) {
      bool pass = false;
      if (valid && (!sap || (bits[p >> 5] >> (p & 31) & 1u))) pass = bp_filter(K, mb, w, MIC(nxn_pair)[2 * p], MIC(nxn_pair)[2 * p + 1]) || MIC(nxn_pairid)[2 * p + 1] >= 0;
      const unsigned mask = __ballot_sync(FULL, pass);
      if (pass) queue[qn + __popc(mask & ((1u << lane) - 1u))] = p;
      qn += __popc(mask);
      ncol += __popc(mask);
      __syncwarp();
      if (qn >= 32) {
        narrowphase_flush<FEAT>(K, mb, w, lane, queue, 32, nw);
        __syncwarp();
        const int rem = qn - 32;
        int v = 0;
        if (lane < rem) v = queue[32 + lane];
        __syncwarp();
        if (lane < rem) queue[lane] = v;
        qn = rem;
        __syncwarp();
      }
    };
    const int ngp = K.ms.ngrouppair;
    if ((FEAT & COLF_SAP) != 0 && ngp > 0 && !sap) {
      // Two-level broadphase (large pair lists, io.py host_model): bounding spheres of the geom groups in the world frame, lanes
      // over group pairs, then the geom pairs of the surviving group pairs only.  Conservative w.r.t. the per-pair sphere filter.
      float4* gg = reinterpret_cast<float4*>(queue + 72);  // the sweep-and-prune workspace is free on this path
      float4* gs = gg + K.ms.ngeom;
      int* cand = queue + K.scratch_floats - 64;  // pairs that passed the cheap sphere test, waiting for the full filter
      int cn = 0;
      const bool sphere = (K.o.broadphase_filter & BF_SPHERE) != 0;
      // world bounding spheres of the geoms: the quantities of bp_filter's sphere test, |p1 - p2| <= rbound1 + rbound2 + margins
      for (int g = lane; g < K.ms.ngeom; g += 32) {
        const V3 pw = ld3(DW(geom_xpos) + 3 * g);
        const float rb = MR(geom_rbound)[g];
        gg[g] = make_float4(pw.x, pw.y, pw.z, (rb == 0.f || !sphere) ? -1.f : rb + MR(geom_margin)[g]);  // < 0: no sphere test
      }
      for (int g = lane; g < K.ms.ngroup; g += 32) {
        const int b = MI(group_bodyid)[g];
        const float* c = MR(group_sphere) + 4 * g;
        const V3 pw = ld3(DW(xpos) + 3 * b) + mat_vec(ldm3(DW(xmat) + 9 * b), ld3(c));
        gs[g] = make_float4(pw.x, pw.y, pw.z, c[3]);
      }
      __syncwarp();
      // stage 1 (cheap, shared memory only): sphere test per pair; survivors are compacted and go to `chunk` 32 at a time
      auto feed = [&](int p, bool valid) {
        bool near = false;
        if (valid) {
          const float4 a = gg[MIC(nxn_pair)[2 * p]], b = gg[MIC(nxn_pair)[2 * p + 1]];
          const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, bound = a.w + b.w;
          near = a.w < 0.f || b.w < 0.f || dx * dx + dy * dy + dz * dz <= bound * bound || MIC(nxn_pairid)[2 * p + 1] >= 0;
        }
        const unsigned mask = __ballot_sync(FULL, near);
        if (near) cand[cn + __popc(mask & ((1u << lane) - 1u))] = p;
        cn += __popc(mask);
        __syncwarp();
        if (cn >= 32) {
          chunk(cand[lane], true);
          const int rem = cn - 32;
          int v = 0;
          if (lane < rem) v = cand[32 + lane];
          __syncwarp();
          if (lane < rem) cand[lane] = v;
          cn = rem;
          __syncwarp();
        }
      };
      for (int gp0 = 0; gp0 < ngp; gp0 += 32) {
        const int gp = gp0 + lane;
        bool live = false;
        int adr = 0, len = 0;
        if (gp < ngp) {
          live = !sphere || MI(grouppair_always)[gp] != 0;
          if (!live) {
            const float4 a = gs[MI(grouppair)[2 * gp]], b = gs[MI(grouppair)[2 * gp + 1]];
            const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, bound = a.w + b.w;
            live = dx * dx + dy * dy + dz * dz <= bound * bound;
          }
          if (live) {
            adr = MI(grouppair_adr)[gp];
            len = MI(grouppair_adr)[gp + 1] - adr;
          }
        }
        // the geom pairs of this batch's live group pairs, flattened: lane k takes the k-th pair of the concatenated ranges
        const int excl = warp_excl_scan(len, lane);
        const int total = __shfl_sync(FULL, excl + len, 31);
        for (int k0 = 0; k0 < total; k0 += 32) {
          const int k = k0 + lane;
          int owner = 0;  // the last lane whose range starts at or before k (ranges are contiguous in k, empty ones share a start)
#pragma unroll
          for (int step = 16; step > 0; step >>= 1) {
            const int probe = owner + step;
            const int start = __shfl_sync(FULL, excl, probe & 31);
            if (probe < 32 && start <= k) owner = probe;
          }
          const int oadr = __shfl_sync(FULL, adr, owner), oexcl = __shfl_sync(FULL, excl, owner);
          feed(oadr + k - oexcl, k < total);
        }
      }
      if (cn > 0) chunk(lane < cn ? cand[lane] : 0, lane < cn);
    } else {
      for (int p0 = 0; p0 < nxn; p0 += 32) chunk(p0 + lane, p0 + lane < nxn);
    }
    if (qn > 0) narrowphase_flush<FEAT>(K, mb, w, lane, queue, qn, nw);
  }
  if (lane == 0) {
    DW(wcon_n)[0] = min(nw, K.dsz.nwconmax);
    if (ncol) atomicAdd(K.d.ncollision, ncol);
  }
  __syncwarp();
}
