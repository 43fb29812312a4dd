/* mjb_ccd_core.h -- convex narrowphase: GJK distance + EPA penetration for sphere / capsule / ellipsoid / cylinder / box /
 * convex mesh (hull vertices, exhaustive support search).
 *
 * Restates reference mujoco_warp/_src/collision_gjk.py: support :98, _attach_face :195, _epa_support :217, _subdistance
 * :253 (_S3D :319, _S2D :395, _S1D :540), gjk :563, _polytope2 :937, _polytope3 :1027, _polytope4 :1115, _epa :1202,
 * _epa_witness :863, _inflate :2153, ccd :2200 (itself a restatement of MuJoCo C engine_collision_gjk.c).
 * The hill-climbing mesh support (mesh_graph), height fields and the multi-contact face clipping (:1930) are not included.
 *
 * Plain C (no C++), compiled TWICE from this one source: in float32 as __device__ code by mjb_collision.cuh (the product
 * path) and in float64 by oracle/mjoracle.c (the CPU checker).  Because the two builds share this text, device-vs-oracle
 * agreement alone proves nothing about it: what pins it is tests/golden/ccd_pairs.npz, produced by executing the
 * reference's own `ccd` (tests/golden/make_ccd_golden.py).
 *
 * The includer defines:  CCD_REAL (float | double),  CCD_FN (qualifiers of the small helpers),  CCD_FN_NOINLINE (qualifiers of
 * the large routines: kept out of line on the device, where inlining the S3D -> S2D -> S1D recursion into every caller multiplies
 * the code size),  CCD_SQRT / CCD_FABS.
 */
#ifndef MJB_CCD_CORE_H_
#define MJB_CCD_CORE_H_

#ifndef CCD_FN_NOINLINE
#define CCD_FN_NOINLINE CCD_FN
#endif
typedef CCD_REAL creal;
#ifndef CCD_VERT_REAL
#define CCD_VERT_REAL CCD_REAL /* element type of the mesh vertex table */
#endif
#define CR(x) ((creal)(x))

#define CCD_GEOM_HFIELD 1 /* one prism of a height field (collision_convex.py:158 ccd_hfield_kernel), in the field's frame */
#define CCD_GEOM_SPHERE 2
#define CCD_GEOM_CAPSULE 3
#define CCD_GEOM_ELLIPSOID 4
#define CCD_GEOM_CYLINDER 5
#define CCD_GEOM_BOX 6
#define CCD_GEOM_MESH 7

#define CCD_FLOAT_MAX CR(1e30)
#define CCD_MINVAL CR(1e-15)
#define CCD_MIN_DIST CR(1e-10)
#define CCD_MAX_ITER 35                        /* opt.ccd_iterations cap of this build (MuJoCo default 35, reference types.py Option) */
#define CCD_NVERT (5 + CCD_MAX_ITER)           /* collision_convex.py:1146: 10 + 2 * epa_iterations points = 5 + iterations vertex pairs */
#define CCD_NFACE (6 + 5 * CCD_MAX_ITER)       /* collision_convex.py:1150: 6 + MJ_MAX_EPAFACES * epa_iterations */
#define CCD_NHORIZON 24                        /* types.py:30 MJ_MAX_EPAHORIZON */

typedef struct {
  creal pos[3], rot[9], size[3], margin;
  int type;
  const CCD_VERT_REAL* vert; /* mesh: hull vertices in the geom frame (the model's mesh_vert + 3 * mesh_vertadr), nvert of them */
  int nvert;
  const int* graph; /* mesh: MuJoCo-layout convex graph of the hull (mesh_graph + mesh_graphadr), or NULL: exhaustive support search */
  creal prism[6][3]; /* CCD_GEOM_HFIELD: bottom triangle 0-2, top triangle 3-5 */
  int coop;         /* device build only: 1 = every lane of the warp runs THIS pair, mesh support = lane-parallel search (CCD_COOP_MESH_SUPPORT) */
} CcdGeom;

typedef struct {
  creal dist, x1[3], x2[3];
  int dim;
  creal simplex[4][3], simplex1[4][3], simplex2[4][3];
  int index1[4], index2[4];
} CcdGjk;

typedef struct {
  int status, nvert, nface, nhorizon, maxface;
  creal vert[2 * CCD_NVERT][3];
  int vert_index[2 * CCD_NVERT];
  int face[CCD_NFACE];
  creal face_pr[CCD_NFACE][3];
  creal face_norm2[CCD_NFACE];
  int horizon[CCD_NHORIZON];
} CcdPolytope;

CCD_FN creal ccd_dot(const creal* a, const creal* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
CCD_FN void ccd_cross(creal* r, const creal* a, const creal* b) {
  creal x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
CCD_FN void ccd_sub(creal* r, const creal* a, const creal* b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
CCD_FN void ccd_copy(creal* r, const creal* a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
CCD_FN creal ccd_norm(const creal* a) { return CCD_SQRT(ccd_dot(a, a)); }
CCD_FN creal ccd_sign(creal x) { return x < CR(0.0) ? CR(-1.0) : CR(1.0); } /* wp.sign: -1 for negative, +1 otherwise */
CCD_FN creal ccd_det3(const creal* v1, const creal* v2, const creal* v3) {
  creal c[3];
  ccd_cross(c, v2, v3);
  return ccd_dot(v1, c);
}
/* rot (row-major 3x3) times v, and its transpose */
CCD_FN void ccd_rot(creal* r, const creal* m, const creal* v) {
  creal x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2], z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
CCD_FN void ccd_rot_t(creal* r, const creal* m, const creal* v) {
  creal x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2], z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}

/* collision_gjk.py:98 support: furthest point of a geom along `dir`; returns the vertex index (boxes, meshes) or -1.
 * *cached: in = the mesh vertex found by the previous call of this search (geom.index, -1 = none), out = this call's. */
CCD_FN_NOINLINE int ccd_support(const CcdGeom* g, const creal* dir, creal* point, int* cached) {
  int vertex_index = -1, cache_in = *cached;
  *cached = -1;
  if (g->type == CCD_GEOM_SPHERE) {
    creal s = g->size[0] + CR(0.5) * g->margin;
    for (int i = 0; i < 3; i++) point[i] = g->pos[i] + s * dir[i];
    return -1;
  }
  if (g->type == CCD_GEOM_HFIELD) { /* :178-188: the prism vertex of largest projection; identity pose */
    creal max_dist = CR(-1e30);
    for (int i = 0; i < 6; i++) {
      creal d = ccd_dot(g->prism[i], dir);
      if (d > max_dist) { max_dist = d; ccd_copy(point, g->prism[i]); }
    }
    if (g->margin > CR(0.0))
      for (int i = 0; i < 3; i++) point[i] += dir[i] * (CR(0.5) * g->margin);
    return dir[2] < CR(0.0) ? -2 : -3;
  }
  creal ld[3], res[3] = {CR(0.0), CR(0.0), CR(0.0)};
  ccd_rot_t(ld, g->rot, dir);
  if (g->type == CCD_GEOM_BOX) {
    creal t0 = ccd_sign(ld[0]), t1 = ccd_sign(ld[1]), t2 = ccd_sign(ld[2]);
    res[0] = t0 * g->size[0]; res[1] = t1 * g->size[1]; res[2] = t2 * g->size[2];
    vertex_index = (t0 > CR(0.0) ? 1 : 0) + (t1 > CR(0.0) ? 2 : 0) + (t2 > CR(0.0) ? 4 : 0);
  } else if (g->type == CCD_GEOM_CAPSULE) {
    res[0] = ld[0] * g->size[0]; res[1] = ld[1] * g->size[0];
    res[2] = ld[2] * g->size[0] + ccd_sign(ld[2]) * g->size[1];
  } else if (g->type == CCD_GEOM_ELLIPSOID) {
    creal r[3] = {ld[0] * g->size[0], ld[1] * g->size[1], ld[2] * g->size[2]};
    creal n = ccd_norm(r);
    if (n > CR(0.0)) { r[0] /= n; r[1] /= n; r[2] /= n; } else { r[0] = r[1] = r[2] = CR(0.0); } /* wp.normalize */
    res[0] = r[0] * g->size[0]; res[1] = r[1] * g->size[1]; res[2] = r[2] * g->size[2];
  } else if (g->type == CCD_GEOM_CYLINDER) {
    creal d = CCD_SQRT(ld[0] * ld[0] + ld[1] * ld[1]);
    if (d > CCD_MINVAL) {
      creal scl = g->size[0] / d;
      res[0] = ld[0] * scl; res[1] = ld[1] * scl;
    }
    res[2] = ccd_sign(ld[2]) * g->size[1];
#ifdef CCD_COOP_MESH_SUPPORT
  } else if (g->type == CCD_GEOM_MESH && g->coop) { /* the hull vertex of largest projection, searched by all lanes together */
    const int best = CCD_COOP_MESH_SUPPORT(g->vert, g->nvert, ld);
    res[0] = (creal)g->vert[3 * best]; res[1] = (creal)g->vert[3 * best + 1]; res[2] = (creal)g->vert[3 * best + 2];
    *cached = best;
    vertex_index = best;
#endif
  } else if (g->type == CCD_GEOM_MESH && g->graph && g->nvert >= 10) { /* :153-176 hill climbing over the hull's edge graph */
    const int* graph = g->graph;
    const int numvert = graph[0];
    const int *vert_edgeadr = graph + 2, *vert_globalid = graph + 2 + numvert, *edge_localid = graph + 2 + 2 * numvert;
    creal max_dist = CR(-1e30);
    int prev = -1, imax = cache_in > -1 ? cache_in : 0;
    while (imax != prev) { /* the start vertex itself is not evaluated (max_dist starts at FLOAT_MIN), as in the reference */
      prev = imax;
      int i = vert_edgeadr[imax], subidx = edge_localid[i];
      while (subidx >= 0) {
        const int idx = vert_globalid[subidx];
        creal d = ld[0] * (creal)g->vert[3 * idx] + ld[1] * (creal)g->vert[3 * idx + 1] + ld[2] * (creal)g->vert[3 * idx + 2];
        if (d > max_dist) { imax = subidx; max_dist = d; }
        i++;
        subidx = edge_localid[i];
      }
    }
    *cached = imax;
    vertex_index = vert_globalid[imax];
    res[0] = (creal)g->vert[3 * vertex_index]; res[1] = (creal)g->vert[3 * vertex_index + 1]; res[2] = (creal)g->vert[3 * vertex_index + 2];
  } else if (g->type == CCD_GEOM_MESH) { /* :140-154 exhaustive search (graphadr == -1); ties keep the cached vertex, then the first */
    creal max_dist = CR(-1e30);
    int best = -1;
    if (cache_in > -1) {
      best = cache_in;
      max_dist = (creal)g->vert[3 * cache_in] * ld[0] + (creal)g->vert[3 * cache_in + 1] * ld[1] + (creal)g->vert[3 * cache_in + 2] * ld[2];
    }
    for (int i = 0; i < g->nvert; i++) {
      creal d = (creal)g->vert[3 * i] * ld[0] + (creal)g->vert[3 * i + 1] * ld[1] + (creal)g->vert[3 * i + 2] * ld[2];
      if (d > max_dist) { max_dist = d; best = i; }
    }
    if (best > -1) { res[0] = (creal)g->vert[3 * best]; res[1] = (creal)g->vert[3 * best + 1]; res[2] = (creal)g->vert[3 * best + 2]; }
    *cached = best;
    vertex_index = best;
  }
  ccd_rot(point, g->rot, res);
  for (int i = 0; i < 3; i++) point[i] += g->pos[i];
  if (g->margin > CR(0.0))
    for (int i = 0; i < 3; i++) point[i] += dir[i] * (CR(0.5) * g->margin);
  return vertex_index;
}

CCD_FN int ccd_same_sign(creal a, creal b) {
  if (a > CR(0.0) && b > CR(0.0)) return 1;
  if (a < CR(0.0) && b < CR(0.0)) return -1;
  return 0;
}

/* :280 */
CCD_FN void ccd_project_origin_line(creal* r, const creal* v1, const creal* v2) {
  creal diff[3];
  ccd_sub(diff, v2, v1);
  creal scl = -(ccd_dot(v2, diff) / ccd_dot(diff, diff));
  for (int i = 0; i < 3; i++) r[i] = v2[i] + scl * diff[i];
}

/* :287 returns 1 on failure */
CCD_FN_NOINLINE int ccd_project_origin_plane(creal* r, const creal* v1, const creal* v2, const creal* v3) {
  creal d21[3], d31[3], d32[3], n[3], nv, nn;
  ccd_sub(d21, v2, v1); ccd_sub(d31, v3, v1); ccd_sub(d32, v3, v2);
  r[0] = r[1] = r[2] = CR(0.0);
  ccd_cross(n, d32, d21);
  nv = ccd_dot(n, v2); nn = ccd_dot(n, n);
  if (nn == CR(0.0)) return 1;
  if (nv != CR(0.0) && nn > CCD_MINVAL) { for (int i = 0; i < 3; i++) r[i] = (nv / nn) * n[i]; return 0; }
  ccd_cross(n, d21, d31);
  nv = ccd_dot(n, v1); nn = ccd_dot(n, n);
  if (nn == CR(0.0)) return 1;
  if (nv != CR(0.0) && nn > CCD_MINVAL) { for (int i = 0; i < 3; i++) r[i] = (nv / nn) * n[i]; return 0; }
  ccd_cross(n, d31, d32);
  nv = ccd_dot(n, v3); nn = ccd_dot(n, n);
  for (int i = 0; i < 3; i++) r[i] = (nv / nn) * n[i];
  return 0;
}

/* :540 */
CCD_FN void ccd_S1D(creal* out, const creal* s1, const creal* s2) {
  creal po[3], mu_max = CR(0.0);
  int index = 0;
  ccd_project_origin_line(po, s1, s2);
  for (int i = 0; i < 3; i++) {
    creal mu = s1[i] - s2[i];
    if (CCD_FABS(mu) >= CCD_FABS(mu_max)) { mu_max = mu; index = i; }
  }
  creal C1 = po[index] - s2[index], C2 = s1[index] - po[index];
  if (ccd_same_sign(mu_max, C1) && ccd_same_sign(mu_max, C2)) { out[0] = C1 / mu_max; out[1] = C2 / mu_max; return; }
  out[0] = CR(0.0); out[1] = CR(1.0);
}

/* :395 */
CCD_FN_NOINLINE void ccd_S2D(creal* out, const creal* s1, const creal* s2, const creal* s3) {
  creal po[3];
  if (ccd_project_origin_plane(po, s1, s2, s3)) {
    creal v[2];
    ccd_S1D(v, s1, s2);
    out[0] = v[0]; out[1] = v[1]; out[2] = CR(0.0);
    return;
  }
  creal M14 = s2[1] * s3[2] - s2[2] * s3[1] - s1[1] * s3[2] + s1[2] * s3[1] + s1[1] * s2[2] - s1[2] * s2[1];
  creal M24 = s2[0] * s3[2] - s2[2] * s3[0] - s1[0] * s3[2] + s1[2] * s3[0] + s1[0] * s2[2] - s1[2] * s2[0];
  creal M34 = s2[0] * s3[1] - s2[1] * s3[0] - s1[0] * s3[1] + s1[1] * s3[0] + s1[0] * s2[1] - s1[1] * s2[0];
  creal mu1 = CCD_FABS(M14), mu2 = CCD_FABS(M24), mu3 = CCD_FABS(M34), Mmax;
  int x, y;
  if (mu1 >= mu2 && mu1 >= mu3) { Mmax = M14; x = 1; y = 2; }
  else if (mu2 >= mu3) { Mmax = M24; x = 0; y = 2; }
  else { Mmax = M34; x = 0; y = 1; }
  creal a0 = s1[x], a1 = s1[y], b0 = s2[x], b1 = s2[y], c0 = s3[x], c1 = s3[y], p0 = po[x], p1 = po[y];
  creal C31 = p0 * b1 + p1 * c0 + b0 * c1 - p0 * c1 - p1 * b0 - c0 * b1;
  creal C32 = p0 * c1 + p1 * a0 + c0 * a1 - p0 * a1 - p1 * c0 - a0 * c1;
  creal C33 = p0 * a1 + p1 * b0 + a0 * b1 - p0 * b1 - p1 * a0 - b0 * a1;
  int comp1 = ccd_same_sign(Mmax, C31), comp2 = ccd_same_sign(Mmax, C32), comp3 = ccd_same_sign(Mmax, C33);
  if (comp1 && comp2 && comp3) { out[0] = C31 / Mmax; out[1] = C32 / Mmax; out[2] = C33 / Mmax; return; }
  creal dmin = CCD_FLOAT_MAX, sub[2], xx[3], d;
  out[0] = out[1] = out[2] = CR(0.0);
  if (!comp1) {
    ccd_S1D(sub, s2, s3);
    for (int i = 0; i < 3; i++) xx[i] = sub[0] * s2[i] + sub[1] * s3[i];
    d = ccd_dot(xx, xx);
    out[0] = CR(0.0); out[1] = sub[0]; out[2] = sub[1];
    dmin = d;
  }
  if (!comp2) {
    ccd_S1D(sub, s1, s3);
    for (int i = 0; i < 3; i++) xx[i] = sub[0] * s1[i] + sub[1] * s3[i];
    d = ccd_dot(xx, xx);
    if (d < dmin) { out[0] = sub[0]; out[1] = CR(0.0); out[2] = sub[1]; dmin = d; }
  }
  if (!comp3) {
    ccd_S1D(sub, s1, s2);
    for (int i = 0; i < 3; i++) xx[i] = sub[0] * s1[i] + sub[1] * s2[i];
    d = ccd_dot(xx, xx);
    if (d < dmin) { out[0] = sub[0]; out[1] = sub[1]; out[2] = CR(0.0); }
  }
}

/* :319 */
CCD_FN_NOINLINE void ccd_S3D(creal* out, const creal* s1, const creal* s2, const creal* s3, const creal* s4) {
  creal C41 = -ccd_det3(s2, s3, s4), C42 = ccd_det3(s1, s3, s4), C43 = -ccd_det3(s1, s2, s4), C44 = ccd_det3(s1, s2, s3);
  creal m_det = C41 + C42 + C43 + C44;
  int comp1 = ccd_same_sign(m_det, C41), comp2 = ccd_same_sign(m_det, C42), comp3 = ccd_same_sign(m_det, C43), comp4 = ccd_same_sign(m_det, C44);
  if (comp1 && comp2 && comp3 && comp4) { out[0] = C41 / m_det; out[1] = C42 / m_det; out[2] = C43 / m_det; out[3] = C44 / m_det; return; }
  creal dmin = CCD_FLOAT_MAX, sub[3], x[3], d;
  out[0] = out[1] = out[2] = out[3] = CR(0.0);
  if (!comp1) {
    ccd_S2D(sub, s2, s3, s4);
    for (int i = 0; i < 3; i++) x[i] = sub[0] * s2[i] + sub[1] * s3[i] + sub[2] * s4[i];
    d = ccd_dot(x, x);
    out[0] = CR(0.0); out[1] = sub[0]; out[2] = sub[1]; out[3] = sub[2];
    dmin = d;
  }
  if (!comp2) {
    ccd_S2D(sub, s1, s3, s4);
    for (int i = 0; i < 3; i++) x[i] = sub[0] * s1[i] + sub[1] * s3[i] + sub[2] * s4[i];
    d = ccd_dot(x, x);
    if (d < dmin) { out[0] = sub[0]; out[1] = CR(0.0); out[2] = sub[1]; out[3] = sub[2]; dmin = d; }
  }
  if (!comp3) {
    ccd_S2D(sub, s1, s2, s4);
    for (int i = 0; i < 3; i++) x[i] = sub[0] * s1[i] + sub[1] * s2[i] + sub[2] * s4[i];
    d = ccd_dot(x, x);
    if (d < dmin) { out[0] = sub[0]; out[1] = sub[1]; out[2] = CR(0.0); out[3] = sub[2]; dmin = d; }
  }
  if (!comp4) {
    ccd_S2D(sub, s1, s2, s3);
    for (int i = 0; i < 3; i++) x[i] = sub[0] * s1[i] + sub[1] * s2[i] + sub[2] * s3[i];
    d = ccd_dot(x, x);
    if (d < dmin) { out[0] = sub[0]; out[1] = sub[1]; out[2] = sub[2]; out[3] = CR(0.0); }
  }
}

/* :253 */
CCD_FN void ccd_subdistance(creal* c, int n, creal simplex[4][3]) {
  c[0] = c[1] = c[2] = c[3] = CR(0.0);
  if (n == 4) { ccd_S3D(c, simplex[0], simplex[1], simplex[2], simplex[3]); return; }
  if (n == 3) { ccd_S2D(c, simplex[0], simplex[1], simplex[2]); return; }
  if (n == 2) { ccd_S1D(c, simplex[0], simplex[1]); return; }
  c[0] = CR(1.0);
}

/* :234 */
CCD_FN void ccd_linear_combine(creal* v, int n, const creal* c, creal mat[4][3]) {
  v[0] = v[1] = v[2] = CR(0.0);
  if (n < 1) n = 4; /* the reference's `else` branch */
  for (int k = 0; k < n && k < 4; k++)
    for (int i = 0; i < 3; i++) v[i] += c[k] * mat[k][i];
}

/* :563 gjk.  x1_0 / x2_0: initial guess (geom centres). */
CCD_FN_NOINLINE void ccd_gjk(CcdGjk* res, creal tolerance, int iterations, const CcdGeom* g1, const CcdGeom* g2, const creal* x1_0,
                    const creal* x2_0, creal cutoff, int is_discrete) {
  creal cutoff2 = cutoff * cutoff, tol2 = tolerance * tolerance, epsilon = is_discrete ? CR(0.0) : CR(0.5) * tol2;
  creal coord[4] = {CR(0.0), CR(0.0), CR(0.0), CR(0.0)}, xk[3], xnorm_old = CCD_FLOAT_MAX;
  int n = 0;
  for (int k = 0; k < 4; k++) {
    res->index1[k] = res->index2[k] = 0;
    for (int i = 0; i < 3; i++) res->simplex[k][i] = res->simplex1[k][i] = res->simplex2[k][i] = CR(0.0);
  }
  for (int i = 0; i < 3; i++) res->x1[i] = res->x2[i] = CR(0.0); /* a fresh GJKResult: early exits leave the witness points at zero */
  int cache1 = -1, cache2 = -1; /* geom.index: lives in gjk's by-value copies of the geoms (:602, :608) */
  ccd_sub(xk, x1_0, x2_0);
  for (int it = 0; it < iterations; it++) {
    creal xnorm = ccd_dot(xk, xk);
    if (xnorm < tol2 || CCD_FABS(xnorm_old - xnorm) < tol2) break;
    xnorm_old = xnorm;
    creal inv = CR(1.0) / CCD_SQRT(xnorm), dneg[3] = {xk[0] * inv, xk[1] * inv, xk[2] * inv}, dpos[3] = {-dneg[0], -dneg[1], -dneg[2]};
    res->index1[n] = ccd_support(g1, dpos, res->simplex1[n], &cache1);
    res->index2[n] = ccd_support(g2, dneg, res->simplex2[n], &cache2);
    ccd_sub(res->simplex[n], res->simplex1[n], res->simplex2[n]);
    if (cutoff == CR(0.0)) {
      if (ccd_dot(xk, res->simplex[n]) > CR(0.0)) { res->dim = 0; res->dist = CCD_FLOAT_MAX; return; }
    } else if (cutoff < CCD_FLOAT_MAX) {
      creal vs = ccd_dot(xk, res->simplex[n]);
      if (vs > CR(0.0) && (vs * vs / xnorm) >= cutoff2) { res->dim = 0; res->dist = CCD_FLOAT_MAX; return; }
    }
    creal dif[3];
    ccd_sub(dif, xk, res->simplex[n]);
    if (ccd_dot(xk, dif) < epsilon) break;
    ccd_subdistance(coord, n + 1, res->simplex);
    n = 0;
    for (int i = 0; i < 4; i++) {
      if (coord[i] == CR(0.0)) continue;
      ccd_copy(res->simplex[n], res->simplex[i]);
      ccd_copy(res->simplex1[n], res->simplex1[i]);
      ccd_copy(res->simplex2[n], res->simplex2[i]);
      res->index1[n] = res->index1[i];
      res->index2[n] = res->index2[i];
      coord[n] = coord[i];
      n++;
    }
    if (n < 1) break;
    creal xnext[3];
    ccd_linear_combine(xnext, n, coord, res->simplex);
    if (CCD_FABS(xnext[0] - xk[0]) < CCD_MINVAL && CCD_FABS(xnext[1] - xk[1]) < CCD_MINVAL && CCD_FABS(xnext[2] - xk[2]) < CCD_MINVAL) break;
    ccd_copy(xk, xnext);
    if (n == 4) break;
  }
  if (n == 0) { ccd_copy(res->x1, x1_0); ccd_copy(res->x2, x2_0); }
  else { ccd_linear_combine(res->x1, n, coord, res->simplex1); ccd_linear_combine(res->x2, n, coord, res->simplex2); }
  res->dist = ccd_norm(xk);
  res->dim = n;
#ifdef CCD_DEBUG_PRINT
  if (g1->type == CCD_GEOM_HFIELD && g2->type == CCD_GEOM_MESH) {
    CCD_DEBUG_PRINT("ORAGJK dim %d dist %.6f i1 [%d %d %d %d] i2 [%d %d %d %d] s2", n, (double)res->dist, res->index1[0], res->index1[1], res->index1[2], res->index1[3], res->index2[0], res->index2[1], res->index2[2], res->index2[3]);
    for (int k = 0; k < 4; k++) CCD_DEBUG_PRINT(" [%.4f %.4f %.4f]", (double)res->simplex2[k][0], (double)res->simplex2[k][1], (double)res->simplex2[k][2]);
    CCD_DEBUG_PRINT("\n");
  }
#endif
}

/* :689 */
CCD_FN int ccd_same_side(const creal* p0, const creal* p1, const creal* p2, const creal* p3) {
  creal a[3], b[3], n[3], c[3], m[3] = {-p0[0], -p0[1], -p0[2]};
  ccd_sub(a, p1, p0); ccd_sub(b, p2, p0); ccd_cross(n, a, b);
  ccd_sub(c, p3, p0);
  creal d1 = ccd_dot(n, c), d2 = ccd_dot(n, m);
  return (d1 > CR(0.0) && d2 > CR(0.0)) || (d1 < CR(0.0) && d2 < CR(0.0));
}
CCD_FN int ccd_test_tetra(const creal* p0, const creal* p1, const creal* p2, const creal* p3) {
  return ccd_same_side(p0, p1, p2, p3) && ccd_same_side(p1, p2, p3, p0) && ccd_same_side(p2, p3, p0, p1) && ccd_same_side(p3, p0, p1, p2);
}

/* :702 */
CCD_FN_NOINLINE void ccd_tri_affine_coord(creal* out, const creal* v1, const creal* v2, const creal* v3, const creal* p) {
  creal M14 = v2[1] * v3[2] - v2[2] * v3[1] - v1[1] * v3[2] + v1[2] * v3[1] + v1[1] * v2[2] - v1[2] * v2[1];
  creal M24 = v2[0] * v3[2] - v2[2] * v3[0] - v1[0] * v3[2] + v1[2] * v3[0] + v1[0] * v2[2] - v1[2] * v2[0];
  creal M34 = v2[0] * v3[1] - v2[1] * v3[0] - v1[0] * v3[1] + v1[1] * v3[0] + v1[0] * v2[1] - v1[1] * v2[0];
  creal mu1 = CCD_FABS(M14), mu2 = CCD_FABS(M24), mu3 = CCD_FABS(M34), Mmax;
  int x, y;
  if (mu1 >= mu2 && mu1 >= mu3) { Mmax = M14; x = 1; y = 2; }
  else if (mu2 >= mu3) { Mmax = M24; x = 0; y = 2; }
  else { Mmax = M34; x = 0; y = 1; }
  creal C31 = p[x] * v2[y] + p[y] * v3[x] + v2[x] * v3[y] - p[x] * v3[y] - p[y] * v2[x] - v3[x] * v2[y];
  creal C32 = p[x] * v3[y] + p[y] * v1[x] + v3[x] * v1[y] - p[x] * v1[y] - p[y] * v3[x] - v1[x] * v3[y];
  creal C33 = p[x] * v1[y] + p[y] * v2[x] + v1[x] * v2[y] - p[x] * v2[y] - p[y] * v1[x] - v2[x] * v1[y];
  out[0] = C31 / Mmax; out[1] = C32 / Mmax; out[2] = C33 / Mmax;
}

/* :745 */
CCD_FN int ccd_tri_point_intersect(const creal* v1, const creal* v2, const creal* v3, const creal* p) {
  creal l[3], pr[3], dd[3];
  ccd_tri_affine_coord(l, v1, v2, v3, p);
  if (l[0] < CR(0.0) || l[1] < CR(0.0) || l[2] < CR(0.0)) return 0;
  for (int i = 0; i < 3; i++) pr[i] = v1[i] * l[0] + v2[i] * l[1] + v3[i] * l[2];
  ccd_sub(dd, pr, p);
  return ccd_norm(dd) < CCD_MINVAL;
}

/* :195 */
CCD_FN_NOINLINE creal ccd_attach_face(CcdPolytope* pt, int idx, int v1, int v2, int v3) {
  if (pt->nface == pt->maxface) return CR(0.0);
  creal p1[3], p2[3], p3[3], r[3];
  ccd_sub(p1, pt->vert[2 * v1], pt->vert[2 * v1 + 1]);
  ccd_sub(p2, pt->vert[2 * v2], pt->vert[2 * v2 + 1]);
  ccd_sub(p3, pt->vert[2 * v3], pt->vert[2 * v3 + 1]);
  if (ccd_project_origin_plane(r, p3, p2, p1)) return CR(0.0);
  pt->face[idx] = v1 + (v2 << 10) + (v3 << 20);
  ccd_copy(pt->face_pr[idx], r);
  pt->face_norm2[idx] = ccd_dot(r, r);
  return pt->face_norm2[idx];
}

/* :217 */
CCD_FN void ccd_epa_support(CcdPolytope* pt, int idx, const CcdGeom* g1, const CcdGeom* g2, const creal* dir, int* cache1, int* cache2) {
  creal nd[3] = {-dir[0], -dir[1], -dir[2]};
  pt->vert_index[2 * idx] = ccd_support(g1, dir, pt->vert[2 * idx], cache1);
  pt->vert_index[2 * idx + 1] = ccd_support(g2, nd, pt->vert[2 * idx + 1], cache2);
}

/* :762 */
CCD_FN void ccd_replace_simplex3(CcdGjk* res, const CcdPolytope* pt, int v1, int v2, int v3) {
  int v[3] = {v1, v2, v3};
  for (int k = 0; k < 4; k++) {
    res->index1[k] = res->index2[k] = 0;
    for (int i = 0; i < 3; i++) res->simplex[k][i] = res->simplex1[k][i] = res->simplex2[k][i] = CR(0.0);
  }
  for (int k = 0; k < 3; k++) {
    ccd_copy(res->simplex1[k], pt->vert[2 * v[k]]);
    ccd_copy(res->simplex2[k], pt->vert[2 * v[k] + 1]);
    ccd_sub(res->simplex[k], res->simplex1[k], res->simplex2[k]);
    res->index1[k] = pt->vert_index[2 * v[k]];
    res->index2[k] = pt->vert_index[2 * v[k] + 1];
  }
}

/* :823 */
CCD_FN int ccd_ray_triangle(const creal* v1, const creal* v2, const creal* v3, const creal* v4, const creal* v5) {
  creal a[3], b[3], c[3], e[3];
  ccd_sub(a, v3, v1); ccd_sub(b, v4, v1); ccd_sub(c, v5, v1); ccd_sub(e, v2, v1);
  creal vol1 = ccd_det3(a, b, e), vol2 = ccd_det3(b, c, e), vol3 = ccd_det3(c, a, e);
  if (vol1 >= CR(0.0) && vol2 >= CR(0.0) && vol3 >= CR(0.0)) return 1;
  if (vol1 <= CR(0.0) && vol2 <= CR(0.0) && vol3 <= CR(0.0)) return -1;
  return 0;
}

CCD_FN void ccd_load_simplex(CcdPolytope* pt, const CcdGjk* r, int n) {
  for (int k = 0; k < n; k++) {
    ccd_copy(pt->vert[2 * k], r->simplex1[k]);
    ccd_copy(pt->vert[2 * k + 1], r->simplex2[k]);
    pt->vert_index[2 * k] = r->index1[k];
    pt->vert_index[2 * k + 1] = r->index2[k];
  }
}

/* :937 polytope from a 1-simplex.  On status -1 `r` is replaced by a 2-simplex (the caller continues with _polytope3). */
CCD_FN_NOINLINE void ccd_polytope2(CcdPolytope* pt, CcdGjk* r, const CcdGeom* g1, const CcdGeom* g2) {
  creal diff[3], value = CCD_FLOAT_MAX, e[3] = {CR(0.0), CR(0.0), CR(0.0)}, d1[3], d2[3], d3[3], R[9];
  creal s0[3], s1[3];
  int index = 0;
  ccd_copy(s0, r->simplex[0]); ccd_copy(s1, r->simplex[1]);
  ccd_sub(diff, s1, s0);
  for (int i = 0; i < 3; i++)
    if (CCD_FABS(diff[i]) < value) { value = CCD_FABS(diff[i]); index = i; }
  e[index] = CR(1.0);
  ccd_cross(d1, e, diff);
  { /* :801 _rotmat: rotation by 120 degrees about diff */
    creal n = ccd_norm(diff), u1 = diff[0] / n, u2 = diff[1] / n, u3 = diff[2] / n, sn = CR(0.86602540378), cs = CR(-0.5), oc = CR(1.0) - cs;
    R[0] = cs + u1 * u1 * oc; R[1] = u1 * u2 * oc - u3 * sn; R[2] = u1 * u3 * oc + u2 * sn;
    R[3] = u2 * u1 * oc + u3 * sn; R[4] = cs + u2 * u2 * oc; R[5] = u2 * u3 * oc - u1 * sn;
    R[6] = u1 * u3 * oc - u2 * sn; R[7] = u2 * u3 * oc + u1 * sn; R[8] = cs + u3 * u3 * oc;
  }
  ccd_rot(d2, R, d1);
  ccd_rot(d3, R, d2);
  ccd_load_simplex(pt, r, 2);
  creal n1 = ccd_norm(d1), n2 = ccd_norm(d2), n3 = ccd_norm(d3);
  for (int i = 0; i < 3; i++) { d1[i] /= n1; d2[i] /= n2; d3[i] /= n3; }
  { int c1 = -1, c2 = -1; ccd_epa_support(pt, 2, g1, g2, d1, &c1, &c2); }
  { int c1 = -1, c2 = -1; ccd_epa_support(pt, 3, g1, g2, d2, &c1, &c2); }
  { int c1 = -1, c2 = -1; ccd_epa_support(pt, 4, g1, g2, d3, &c1, &c2); }
  const int f[6][3] = {{0, 2, 3}, {0, 4, 2}, {0, 3, 4}, {1, 3, 2}, {1, 2, 4}, {1, 4, 3}};
  for (int k = 0; k < 6; k++) {
    if (ccd_attach_face(pt, k, f[k][0], f[k][1], f[k][2]) < CCD_MIN_DIST) {
      pt->status = -1;
      ccd_replace_simplex3(r, pt, f[k][0], f[k][1], f[k][2]);
      return;
    }
  }
  creal v2[3], v3[3], v4[3];
  ccd_sub(v2, pt->vert[4], pt->vert[5]); ccd_sub(v3, pt->vert[6], pt->vert[7]); ccd_sub(v4, pt->vert[8], pt->vert[9]);
  if (!ccd_ray_triangle(s0, s1, v2, v3, v4)) { pt->status = 1; return; }
  pt->nvert = 5; pt->nface = 6; pt->status = 0;
}

/* :1027 polytope from a 2-simplex */
CCD_FN_NOINLINE void ccd_polytope3(CcdPolytope* pt, const CcdGjk* r, const CcdGeom* g1, const CcdGeom* g2) {
  creal a[3], b[3], n[3], nn[3], v1[3], v2[3], v3[3], v4[3], v5[3];
  ccd_copy(v1, r->simplex[0]); ccd_copy(v2, r->simplex[1]); ccd_copy(v3, r->simplex[2]);
  ccd_sub(a, v2, v1); ccd_sub(b, v3, v1); ccd_cross(n, a, b);
  if (ccd_norm(n) < CCD_MINVAL) { pt->status = 2; return; }
  ccd_load_simplex(pt, r, 3);
  nn[0] = -n[0]; nn[1] = -n[1]; nn[2] = -n[2];
  { int c1 = -1, c2 = -1; ccd_epa_support(pt, 3, g1, g2, nn, &c1, &c2); }
  { int c1 = -1, c2 = -1; ccd_epa_support(pt, 4, g1, g2, n, &c1, &c2); }
  ccd_sub(v4, pt->vert[6], pt->vert[7]);
  ccd_sub(v5, pt->vert[8], pt->vert[9]);
  if (ccd_tri_point_intersect(v1, v2, v3, v4)) { pt->status = 3; return; }
  if (ccd_tri_point_intersect(v1, v2, v3, v5)) { pt->status = 4; return; }
  if (r->dist > CR(1e-5) && !ccd_test_tetra(v1, v2, v3, v4) && !ccd_test_tetra(v1, v2, v3, v5)) { pt->status = 5; return; }
  const int f[6][3] = {{4, 0, 1}, {4, 2, 0}, {4, 1, 2}, {3, 1, 0}, {3, 0, 2}, {3, 2, 1}};
  for (int k = 0; k < 6; k++)
    if (ccd_attach_face(pt, k, f[k][0], f[k][1], f[k][2]) < CCD_MIN_DIST) { pt->status = 6 + k; return; }
  pt->nvert = 5; pt->nface = 6; pt->status = 0;
}

/* :1115 polytope from a 3-simplex */
CCD_FN_NOINLINE void ccd_polytope4(CcdPolytope* pt, CcdGjk* r) {
  ccd_load_simplex(pt, r, 4);
  const int f[4][3] = {{0, 1, 2}, {0, 3, 1}, {0, 2, 3}, {3, 2, 1}};
  for (int k = 0; k < 4; k++) {
    if (ccd_attach_face(pt, k, f[k][0], f[k][1], f[k][2]) < CCD_MIN_DIST) {
      pt->status = -1;
      ccd_replace_simplex3(r, pt, f[k][0], f[k][1], f[k][2]);
      return;
    }
  }
  if (!ccd_test_tetra(r->simplex[0], r->simplex[1], r->simplex[2], r->simplex[3])) { pt->status = 12; return; }
  pt->nvert = 4; pt->nface = 4; pt->status = 0;
}

#define CCD_FACE_DELETED 0x80000000u
#define CCD_FACE_INVALID 0x40000000u
#define CCD_FACE_MASK 0xC0000000u

/* :841 returns the new horizon size, -1 when the horizon buffer is full */
CCD_FN int ccd_add_edge(CcdPolytope* pt, int e1, int e2) {
  int n = pt->nhorizon;
  if (n < 0) return -1;
  int lo = e1 < e2 ? e1 : e2, hi = e1 < e2 ? e2 : e1, edge = (lo << 10) | hi;
  for (int i = 0; i < n; i++)
    if (edge == pt->horizon[i]) { pt->horizon[i] = pt->horizon[n - 1]; return n - 1; }
  if (n == CCD_NHORIZON) return -1;
  pt->horizon[n] = edge;
  return n + 1;
}

/* :886-921 and :2160-2185: witness points against the TOP triangle of a height-field prism when the simplex / face touches more than
 * one of its sides.  x2 is the point found on geom 2; returns x1 on the prism and dist = -|x1 - x2|. */
CCD_FN creal ccd_hfield_witness(const CcdGeom* g1, creal* x1, const creal* x2) {
  const creal *a = g1->prism[3], *b = g1->prism[4], *c = g1->prism[5];
  creal co[3];
  ccd_tri_affine_coord(co, a, b, c, x2);
  if (co[0] > CR(0.0) && co[1] > CR(0.0) && co[2] > CR(0.0)) {
    for (int i = 0; i < 3; i++) x1[i] = co[0] * a[i] + co[1] * b[i] + co[2] * c[i];
  } else {
    const creal* p = c;
    if (co[1] > CR(0.0)) p = b;
    if (co[0] > CR(0.0)) p = a;
    const creal h = x2[2] - p[2]; /* dot(x2 - p, n) with n = (0, 0, 1) */
    x1[0] = x2[0]; x1[1] = x2[1]; x1[2] = x2[2] - h;
  }
  creal d[3];
  ccd_sub(d, x1, x2);
  return -ccd_norm(d);
}

/* :1202 _epa + :863 _epa_witness.  Returns the face index (-1: no contact) and dist / x1 / x2. */
CCD_FN_NOINLINE int ccd_epa(creal tolerance, int iterations, CcdPolytope* pt, const CcdGeom* g1, const CcdGeom* g2, int is_discrete, creal* dist,
                   creal* x1, creal* x2) {
  creal upper = CCD_FLOAT_MAX, upper2 = CCD_FLOAT_MAX, epsilon = is_discrete ? CR(1e-15) : tolerance;
  int idx = -1, pidx = -1, nvalid = pt->nface, cache1 = -1, cache2 = -1;
  if (iterations > 1000) iterations = 1000;
  for (int it = 0; it < iterations; it++) {
    pidx = idx;
    idx = -1;
    creal lower2 = CCD_FLOAT_MAX;
    for (int i = 0; i < pt->nface; i++)
      if (!((unsigned)pt->face[i] & CCD_FACE_MASK) && pt->face_norm2[i] < lower2) { idx = i; lower2 = pt->face_norm2[i]; }
    if (lower2 > upper2 || idx < 0) { idx = pidx; break; }
    if (lower2 <= CR(0.0)) break;
    creal lower = CCD_SQRT(lower2), dir[3] = {pt->face_pr[idx][0] / lower, pt->face_pr[idx][1] / lower, pt->face_pr[idx][2] / lower}, w[3];
    int wi = pt->nvert;
    ccd_epa_support(pt, wi, g1, g2, dir, &cache1, &cache2); /* :1248-1251: geom.index follows the last support */
    ccd_sub(w, pt->vert[2 * wi], pt->vert[2 * wi + 1]);
    pt->nvert++;
    creal upper_k = ccd_dot(dir, w);
    if (upper_k < upper) { upper = upper_k; upper2 = upper * upper; }
    if (upper - lower < epsilon) break;
#ifdef CCD_DEBUG_PRINT
    if (g1->type == CCD_GEOM_HFIELD && g2->type == CCD_GEOM_MESH) {
      CCD_DEBUG_PRINT("  EPA it %d idx %d wi %d disc %d idx(wi) %d %d prev:", it, idx, wi, is_discrete, pt->vert_index[2 * wi], pt->vert_index[2 * wi + 1]);
      for (int i = 0; i < wi; i++) CCD_DEBUG_PRINT(" (%d %d)", pt->vert_index[2 * i], pt->vert_index[2 * i + 1]);
      CCD_DEBUG_PRINT(" upper-lower %.3e\n", (double)(upper - lower));
    }
#endif
    if (is_discrete) {
      int found = 0;
      for (int i = 0; i < pt->nvert - 1; i++)
        if (pt->vert_index[2 * i] == pt->vert_index[2 * wi] && pt->vert_index[2 * i + 1] == pt->vert_index[2 * wi + 1]) { found = 1; break; }
      if (found) break;
    }
    nvalid--;
    pt->face[idx] = (int)((unsigned)pt->face[idx] | CCD_FACE_DELETED);
    {
      int fc = pt->face[idx], a = fc & 0x3FF, b = (fc >> 10) & 0x3FF, c = (fc >> 20) & 0x3FF;
      pt->nhorizon = ccd_add_edge(pt, a, b);
      pt->nhorizon = ccd_add_edge(pt, b, c);
      pt->nhorizon = ccd_add_edge(pt, c, a);
    }
    if (pt->nhorizon == -1) { idx = -1; break; }
    for (int i = 0; i < pt->nface; i++) {
      if ((unsigned)pt->face[i] & CCD_FACE_DELETED) continue;
      if (ccd_dot(pt->face_pr[i], w) - pt->face_norm2[i] > CR(1e-10)) {
        if (!((unsigned)pt->face[i] & CCD_FACE_MASK)) nvalid--;
        pt->face[i] = (int)((unsigned)pt->face[i] | CCD_FACE_DELETED);
        int fc = pt->face[i], a = fc & 0x3FF, b = (fc >> 10) & 0x3FF, c = (fc >> 20) & 0x3FF;
        pt->nhorizon = ccd_add_edge(pt, a, b);
        pt->nhorizon = ccd_add_edge(pt, b, c);
        pt->nhorizon = ccd_add_edge(pt, c, a);
        if (pt->nhorizon == -1) { idx = -1; break; }
      }
    }
    for (int i = 0; i < pt->nhorizon; i++) {
      int e0 = pt->horizon[i] & 0x3FF, e1 = (pt->horizon[i] >> 10) & 0x3FF;
      creal dist2 = ccd_attach_face(pt, pt->nface, wi, e0, e1);
      if (dist2 == CR(0.0)) { idx = -1; break; }
      pt->nface++;
      if (dist2 >= lower2 && dist2 <= upper2) nvalid++;
      else pt->face[pt->nface - 1] = (int)((unsigned)pt->face[pt->nface - 1] | CCD_FACE_INVALID);
    }
    if (nvalid == 0 || idx == -1) break;
    pt->nhorizon = 0;
  }
  if (idx > -1) { /* :863 witness points from the affine coordinates of the projected origin on the face */
    int fc = pt->face[idx], f0 = fc & 0x3FF, f1 = (fc >> 10) & 0x3FF, f2 = (fc >> 20) & 0x3FF;
    creal v1[3], v2[3], v3[3], l[3];
    ccd_sub(v1, pt->vert[2 * f0], pt->vert[2 * f0 + 1]);
    ccd_sub(v2, pt->vert[2 * f1], pt->vert[2 * f1 + 1]);
    ccd_sub(v3, pt->vert[2 * f2], pt->vert[2 * f2 + 1]);
    ccd_tri_affine_coord(l, v1, v2, v3, pt->face_pr[idx]);
    for (int i = 0; i < 3; i++) {
      x2[i] = pt->vert[2 * f0 + 1][i] * l[0] + pt->vert[2 * f1 + 1][i] * l[1] + pt->vert[2 * f2 + 1][i] * l[2];
      x1[i] = pt->vert[2 * f0][i] * l[0] + pt->vert[2 * f1][i] * l[1] + pt->vert[2 * f2][i] * l[2];
    }
    *dist = -CCD_SQRT(pt->face_norm2[idx]);
    {
      const int i1 = pt->vert_index[2 * f0], i2 = pt->vert_index[2 * f1], i3 = pt->vert_index[2 * f2];
#ifdef CCD_DEBUG_PRINT
      if (g1->type == CCD_GEOM_HFIELD && g2->type == CCD_GEOM_MESH) CCD_DEBUG_PRINT("ORAWIT face_idx %d face (%d %d %d) i [%d %d %d] nvert %d nface %d dist %.6f\n", idx, f0, f1, f2, i1, i2, i3, pt->nvert, pt->nface, (double)*dist);
#endif
      if (g1->type == CCD_GEOM_HFIELD && (i1 != i2 || i1 != i3)) { /* :886-921 correct the witness points for height fields */
        CcdGeom t = *g2;
        creal sp[3], dirn[3];
        int cache = -1;
        if (g2->type == CCD_GEOM_CAPSULE || g2->type == CCD_GEOM_SPHERE) {
          const creal radius = g2->size[0], margin = g2->margin;
          t.margin = CR(0.0);
          t.size[0] = CR(0.0);
          ccd_support(&t, x2, sp, &cache); /* the reference passes x2 itself as the direction here */
          x2[0] = sp[0]; x2[1] = sp[1]; x2[2] = sp[2] - (CR(0.5) * margin + radius);
        } else {
          ccd_copy(dirn, x2);
          creal n = ccd_norm(dirn);
          if (n > CR(0.0)) { dirn[0] /= n; dirn[1] /= n; dirn[2] /= n; } else { dirn[0] = dirn[1] = dirn[2] = CR(0.0); }
          ccd_support(&t, dirn, sp, &cache);
          ccd_copy(x2, sp);
        }
        *dist = ccd_hfield_witness(g1, x1, x2);
      }
    }
    return idx;
  }
  *dist = CR(0.0);
  x1[0] = x1[1] = x1[2] = x2[0] = x2[1] = x2[2] = CR(0.0);
  return -1;
}

/* :2200 ccd.  g1 / g2 are taken by value (the routine shrinks spheres / capsules to points / segments).  `pt` is scratch.
 * Returns the number of contacts (0 or 1); dist is relative to the margin-inflated shapes as in the reference. */
CCD_FN_NOINLINE int ccd_run(creal tolerance, creal cutoff, int gjk_iterations, int epa_iterations, CcdGeom g1, CcdGeom g2, CcdPolytope* pt,
                   CcdGjk* res, creal* dist, creal* x1, creal* x2, int* multiccd_idx) {
  *multiccd_idx = -1; /* the EPA face the contact was found on, when multi-contact recovery applies (:2336-2343) */
  creal full1 = CR(0.0), full2 = CR(0.0), size1 = CR(0.0), size2 = CR(0.0), x_1[3], x_2[3];
  int is_discrete = ((g1.type == CCD_GEOM_BOX || g1.type == CCD_GEOM_MESH || g1.type == CCD_GEOM_HFIELD) &&
                     (g2.type == CCD_GEOM_BOX || g2.type == CCD_GEOM_MESH || g2.type == CCD_GEOM_HFIELD)) &&
                    (g1.margin == CR(0.0) && g2.margin == CR(0.0)); /* :91 _discrete_geoms */
  ccd_copy(x_1, g1.pos); ccd_copy(x_2, g2.pos);
  if (g1.type == CCD_GEOM_SPHERE || g1.type == CCD_GEOM_CAPSULE) {
    size1 = g1.size[0]; full1 = size1 + CR(0.5) * g1.margin; g1.margin = CR(0.0); g1.size[0] = CR(0.0);
  }
  if (g2.type == CCD_GEOM_SPHERE || g2.type == CCD_GEOM_CAPSULE) {
    size2 = g2.size[0]; full2 = size2 + CR(0.5) * g2.margin; g2.margin = CR(0.0); g2.size[0] = CR(0.0);
  }
  if (size1 + size2 > CR(0.0)) {
    cutoff += full1 + full2;
    ccd_gjk(res, tolerance, gjk_iterations, &g1, &g2, x_1, x_2, cutoff, is_discrete);
    if (res->dist > tolerance) {
      ccd_copy(x1, res->x1); ccd_copy(x2, res->x2);
      if (res->dist == CCD_FLOAT_MAX) { *dist = res->dist; return 1; }
      /* :2153 _inflate */
      if (g1.type == CCD_GEOM_HFIELD) { /* :2160-2185 the simplex touches a side of the prism: witness against its top triangle */
        int is_side = 0;
        for (int i = 0; i < res->dim; i++)
          if (res->index1[i] != res->index1[0]) { is_side = 1; break; }
        if (is_side) {
          creal sp[3];
          int cache = -1;
          ccd_support(&g2, x2, sp, &cache);
          x2[0] = sp[0]; x2[1] = sp[1]; x2[2] = sp[2] - full2;
          *dist = ccd_hfield_witness(&g1, x1, x2);
          return 1;
        }
      }
      creal n[3], nn;
      ccd_sub(n, x2, x1);
      nn = ccd_norm(n);
      if (nn > CR(0.0)) { n[0] /= nn; n[1] /= nn; n[2] /= nn; } else { n[0] = n[1] = n[2] = CR(0.0); }
      if (full1 > CR(0.0)) for (int i = 0; i < 3; i++) x1[i] += full1 * n[i];
      if (full2 > CR(0.0)) for (int i = 0; i < 3; i++) x2[i] -= full2 * n[i];
      *dist = res->dist - (full1 + full2);
      return 1;
    }
    g1.margin = full1 - size1; g1.size[0] = size1;
    g2.margin = full2 - size2; g2.size[0] = size2;
    cutoff -= full1 + full2;
  }
  ccd_gjk(res, tolerance, gjk_iterations, &g1, &g2, x_1, x_2, cutoff, is_discrete);
  if (res->dist > tolerance || res->dim < 2) { *dist = res->dist; ccd_copy(x1, res->x1); ccd_copy(x2, res->x2); return 1; }
  pt->nface = pt->nvert = pt->nhorizon = 0;
  pt->status = 0;
  pt->maxface = 6 + 5 * epa_iterations;
  if (pt->maxface > CCD_NFACE) pt->maxface = CCD_NFACE;
  if (res->dim == 2) {
    ccd_polytope2(pt, res, &g1, &g2);
    if (pt->status == -1) res->dim = 3;
  } else if (res->dim == 4) {
    ccd_polytope4(pt, res);
    if (pt->status == -1) res->dim = 3;
  }
  if (res->dim == 3) ccd_polytope3(pt, res, &g1, &g2);
  if (pt->status) { *dist = res->dist; ccd_copy(x1, res->x1); ccd_copy(x2, res->x2); return 1; }
  int idx = ccd_epa(tolerance, epa_iterations, pt, &g1, &g2, is_discrete, dist, x1, x2);
  if (idx == -1) { *dist = CCD_FLOAT_MAX; return 0; }
  /* multicontact: not with a margin, boxes and meshes only */
  if (g1.margin == CR(0.0) && g2.margin == CR(0.0) && (g1.type == CCD_GEOM_BOX || g1.type == CCD_GEOM_MESH) &&
      (g2.type == CCD_GEOM_BOX || g2.type == CCD_GEOM_MESH))
    *multiccd_idx = idx;
  return 1;
}

/* ---- multi-contact recovery by face clipping (collision_gjk.py:1332-2150), boxes --------------------------------------------------
 * The reference also clips mesh polygons (_mesh_normals :1460, _mesh_edge_normals :1530, _mesh_face :1766); those need MuJoCo's
 * mesh_poly* tables, which the MJCF-lite compiler does not build: pairs with a mesh keep their single EPA contact. */
#define CCD_FACE_TOL CR(0.99999872000027307) /* cos(0.0016) */
#define CCD_EDGE_TOL CR(0.0015999993173334207) /* sin(0.0016) */
#define CCD_INTERSECT_TOL CR(0.0000003)

typedef struct {
  creal polygon[8][3], clipped[8][3], pnormal[4][3], pdist[4]; /* polygon buffers: 2 * max sides (collision_convex.py:1158) */
  creal n1[3][3], n2[3][3], endvert[3][3], face1[4][3], face2[4][3];
  int idx1[3], idx2[3];
} CcdMulti;

CCD_FN void ccd_matvec(creal* r, const creal* m, creal x, creal y, creal z) {
  r[0] = m[0] * x + m[1] * y + m[2] * z; r[1] = m[3] * x + m[4] * y + m[5] * z; r[2] = m[6] * x + m[7] * y + m[8] * z;
}
CCD_FN void ccd_normalize(creal* v) {
  creal n = ccd_norm(v);
  if (n > CR(0.0)) { v[0] /= n; v[1] /= n; v[2] /= n; } else { v[0] = v[1] = v[2] = CR(0.0); }
}
/* :1332 */
CCD_FN creal ccd_area4(const creal* a, const creal* b, const creal* c, const creal* d) {
  creal ad[3], db[3], bc[3], ca[3], c1[3], c2[3];
  ccd_sub(ad, a, d); ccd_sub(db, d, b); ccd_sub(bc, b, c); ccd_sub(ca, c, a);
  ccd_cross(c1, ad, db); ccd_cross(c2, bc, ca);
  c1[0] += c2[0]; c1[1] += c2[1]; c1[2] += c2[2];
  return CR(0.5) * ccd_norm(c1);
}
/* :1338 quadrilateral of maximum area in a convex polygon (npolygon > 4) */
CCD_FN void ccd_polygon_quad(int* res, creal (*poly)[3], int np) {
  int b = 1, c = 2, d = 3;
  res[0] = 0; res[1] = b; res[2] = c; res[3] = d;
  creal m = ccd_area4(poly[0], poly[b], poly[c], poly[d]);
  for (int a = 0; a < np; a++) {
    for (;;) {
      creal m_next = ccd_area4(poly[a], poly[b], poly[c], poly[(d + 1) % np]);
      if (m_next <= m) break;
      m = m_next; d = (d + 1) % np;
      res[0] = a; res[1] = b; res[2] = c; res[3] = d;
      for (;;) {
        m_next = ccd_area4(poly[a], poly[b], poly[(c + 1) % np], poly[d]);
        if (m_next <= m) break;
        m = m_next; c = (c + 1) % np;
        res[0] = a; res[1] = b; res[2] = c; res[3] = d;
      }
      for (;;) {
        m_next = ccd_area4(poly[a], poly[(b + 1) % np], poly[c], poly[d]);
        if (m_next <= m) break;
        m = m_next; b = (b + 1) % np;
        res[0] = a; res[1] = b; res[2] = c; res[3] = d;
      }
    }
    if (b == a) {
      b = (b + 1) % np;
      if (c == b) {
        c = (c + 1) % np;
        if (d == c) d = (d + 1) % np;
      }
    }
  }
}
/* :1378 dimension (1, 2, 3) of the feature of geom `offset` spanned by an EPA face; vertices reordered if necessary */
CCD_FN int ccd_feature_dim(const int* face, const CcdPolytope* pt, int offset, int* findex, creal (*fvert)[3]) {
  int v1i = pt->vert_index[2 * face[0] + offset], v2i = pt->vert_index[2 * face[1] + offset], v3i = pt->vert_index[2 * face[2] + offset];
  findex[0] = v1i; findex[1] = v2i; findex[2] = v3i;
  for (int k = 0; k < 3; k++) ccd_copy(fvert[k], pt->vert[2 * face[k] + offset]);
  if (v1i != v2i) return (v3i == v1i || v3i == v2i) ? 2 : 3;
  findex[1] = v3i;
  ccd_copy(fvert[1], pt->vert[2 * face[2] + offset]);
  return v1i != v3i ? 2 : 1;
}

/* :1578 the box face (if any) whose normal is within FACE_TOL of n (world frame) */
CCD_FN int ccd_box_normals2(const creal* mat, const creal* n, creal (*normal)[3], int* index) {
  const creal fn[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
  creal ln[3];
  ccd_rot_t(ln, mat, n);
  ccd_normalize(ln);
  for (int i = 0; i < 6; i++) {
    if (ln[0] * fn[i][0] + ln[1] * fn[i][1] + ln[2] * fn[i][2] > CCD_FACE_TOL) {
      ccd_matvec(normal[0], mat, fn[i][0], fn[i][1], fn[i][2]);
      index[0] = i;
      return 1;
    }
  }
  return 0;
}
/* :1611 possible face normals of a box given up to 3 of its vertices (vertex index bits: x 1, y 2, z 4) */
CCD_FN int ccd_box_normals(int fdim, const int* fi, const creal* mat, const creal* dir, creal (*normal)[3], int* index) {
  const int v1 = fi[0], v2 = fi[1], v3 = fi[2];
  if (fdim == 3) {
    int c = 0;
    creal x = (creal)(((v1 & 1) && (v2 & 1) && (v3 & 1)) ? 1 : 0) - (creal)((!(v1 & 1) && !(v2 & 1) && !(v3 & 1)) ? 1 : 0);
    creal y = (creal)(((v1 & 2) && (v2 & 2) && (v3 & 2)) ? 1 : 0) - (creal)((!(v1 & 2) && !(v2 & 2) && !(v3 & 2)) ? 1 : 0);
    creal z = (creal)(((v1 & 4) && (v2 & 4) && (v3 & 4)) ? 1 : 0) - (creal)((!(v1 & 4) && !(v2 & 4) && !(v3 & 4)) ? 1 : 0);
    ccd_matvec(normal[0], mat, x, y, z);
    creal sgn = x + y + z;
    if (x != CR(0.0)) index[c++] = 0;
    if (y != CR(0.0)) index[c++] = 2;
    if (z != CR(0.0)) index[c++] = 4;
    if (sgn == CR(-1.0)) index[0] = index[0] + 1;
    if (c == 1) return 1;
    return ccd_box_normals2(mat, dir, normal, index);
  }
  if (fdim == 2) {
    int c = 0;
    creal x = (creal)(((v1 & 1) && (v2 & 1)) ? 1 : 0) - (creal)((!(v1 & 1) && !(v2 & 1)) ? 1 : 0);
    creal y = (creal)(((v1 & 2) && (v2 & 2)) ? 1 : 0) - (creal)((!(v1 & 2) && !(v2 & 2)) ? 1 : 0);
    creal z = (creal)(((v1 & 4) && (v2 & 4)) ? 1 : 0) - (creal)((!(v1 & 4) && !(v2 & 4)) ? 1 : 0);
    if (x != CR(0.0)) { ccd_matvec(normal[c], mat, x, CR(0.0), CR(0.0)); index[c] = x > CR(0.0) ? 0 : 1; c++; }
    if (y != CR(0.0)) { ccd_matvec(normal[c], mat, CR(0.0), y, CR(0.0)); index[c] = y > CR(0.0) ? 2 : 3; c++; }
    if (z != CR(0.0)) { ccd_matvec(normal[c], mat, CR(0.0), CR(0.0), z); index[c] = z > CR(0.0) ? 4 : 5; c++; }
    if (c == 1 || c == 2) return c; /* 1: the edge is a face diagonal, 2: an external edge of the box */
    return ccd_box_normals2(mat, dir, normal, index);
  }
  if (fdim == 1) {
    creal x = (v1 & 1) ? CR(1.0) : CR(-1.0), y = (v1 & 2) ? CR(1.0) : CR(-1.0), z = (v1 & 4) ? CR(1.0) : CR(-1.0);
    ccd_matvec(normal[0], mat, x, CR(0.0), CR(0.0));
    ccd_matvec(normal[1], mat, CR(0.0), y, CR(0.0));
    ccd_matvec(normal[2], mat, CR(0.0), CR(0.0), z);
    index[0] = x > CR(0.0) ? 0 : 1; index[1] = y > CR(0.0) ? 2 : 3; index[2] = z > CR(0.0) ? 4 : 5;
    return 3;
  }
  return 0;
}
/* :1685 edge directions of a box for edge collisions */
CCD_FN int ccd_box_edge_normals(int dim, const CcdGeom* g, const creal* v1, const creal* v2, int v1i, creal (*normal)[3], creal (*endvert)[3]) {
  if (dim == 2) {
    ccd_copy(endvert[0], v2);
    ccd_sub(normal[0], v2, v1);
    ccd_normalize(normal[0]);
    return 1;
  }
  if (dim == 1) { /* the three vertices adjacent to v1 */
    creal x = (v1i & 1) ? g->size[0] : -g->size[0], y = (v1i & 2) ? g->size[1] : -g->size[1], z = (v1i & 4) ? g->size[2] : -g->size[2];
    const creal c[3][3] = {{-x, y, z}, {x, -y, z}, {x, y, -z}};
    for (int k = 0; k < 3; k++) {
      ccd_matvec(endvert[k], g->rot, c[k][0], c[k][1], c[k][2]);
      for (int i = 0; i < 3; i++) endvert[k][i] += g->pos[i];
      ccd_sub(normal[k], endvert[k], v1);
      ccd_normalize(normal[k]);
    }
    return 3;
  }
  return 0;
}
/* :1723 corners of box face idx (0 +x, 1 -x, 2 +y, 3 -y, 4 +z, 5 -z) in world coordinates */
CCD_FN int ccd_box_face(const CcdGeom* g, int idx, creal (*face)[3]) {
  static const signed char sg[6][4][3] = {
      {{1, 1, 1}, {1, 1, -1}, {1, -1, -1}, {1, -1, 1}},     {{-1, 1, -1}, {-1, 1, 1}, {-1, -1, 1}, {-1, -1, -1}},
      {{-1, 1, -1}, {1, 1, -1}, {1, 1, 1}, {-1, 1, 1}},     {{-1, -1, 1}, {1, -1, 1}, {1, -1, -1}, {-1, -1, -1}},
      {{-1, 1, 1}, {1, 1, 1}, {1, -1, 1}, {-1, -1, 1}},     {{1, 1, -1}, {-1, 1, -1}, {-1, -1, -1}, {1, -1, -1}}};
  if (idx < 0 || idx > 5) return 0;
  for (int k = 0; k < 4; k++) {
    ccd_matvec(face[k], g->rot, sg[idx][k][0] * g->size[0], sg[idx][k][1] * g->size[1], sg[idx][k][2] * g->size[2]);
    for (int i = 0; i < 3; i++) face[k][i] += g->pos[i];
  }
  return 4;
}

/* :1816 clips polygon face2 against the edges of face1 (normal n); up to 4 witness pairs (w2 on the clipped polygon, w1 = w2 - dir) */
CCD_FN int ccd_polygon_clip(CcdMulti* mc, creal (*face1)[3], int nface1, creal (*face2)[3], int nface2, const creal* n, const creal* dir,
                            creal (*w1)[3], creal (*w2)[3]) {
  if (nface1 < 3) return 0; /* the clipping face needs to be at least a triangle */
  for (int i = 0; i < nface1; i++) { /* :1791 _plane_normal of every edge */
    const creal *a = face1[i], *b = face1[(i + 1) % nface1];
    creal ba[3];
    ccd_sub(ba, b, a);
    ccd_cross(mc->pnormal[i], ba, n);
    mc->pdist[i] = ccd_dot(mc->pnormal[i], a);
  }
  creal (*poly)[3] = mc->polygon, (*clip)[3] = mc->clipped;
  int npoly = nface2, nclip = 0;
  for (int i = 0; i < nface2; i++) ccd_copy(poly[i], face2[i]);
  for (int e = 0; e < nface1; e++) {
    for (int i = 0; i < npoly; i++) {
      const creal *P = poly[i], *Q = poly[(i + 1) % npoly];
      creal pa[3], qa[3];
      ccd_sub(pa, P, face1[e]); ccd_sub(qa, Q, face1[e]);
      const int in1 = ccd_dot(pa, mc->pnormal[e]) > CR(-1e-10), in2 = ccd_dot(qa, mc->pnormal[e]) > CR(-1e-10); /* :1798 _halfspace */
      if (!in1 && !in2) continue;
      if (in1 && in2) { ccd_copy(clip[nclip++], Q); continue; }
      creal qp[3];
      ccd_sub(qp, Q, P);
      const creal dt = ccd_dot(mc->pnormal[e], qp); /* :1803 _plane_intersect */
      const creal t = CCD_FABS(dt) < CR(1e-10) ? CCD_FLOAT_MAX : (mc->pdist[e] - ccd_dot(mc->pnormal[e], P)) / dt;
      if (t > -CCD_INTERSECT_TOL && t < CR(1.0) + CCD_INTERSECT_TOL) {
        const creal tc = t < CR(0.0) ? CR(0.0) : (t > CR(1.0) ? CR(1.0) : t);
        for (int k = 0; k < 3; k++) clip[nclip][k] = P[k] + tc * qp[k];
        nclip++;
      }
      if (in2) ccd_copy(clip[nclip++], Q);
    }
    creal (*tmp)[3] = poly; poly = clip; clip = tmp;
    npoly = nclip; nclip = 0;
  }
  if (npoly < 1) return 0;
  if (npoly > 4) {
    int quad[4];
    ccd_polygon_quad(quad, poly, npoly);
    for (int i = 0; i < 4; i++) { ccd_copy(w2[i], poly[quad[i]]); ccd_sub(w1[i], w2[i], dir); }
    return 4;
  }
  for (int i = 0; i < npoly; i++) { ccd_copy(w2[i], poly[i]); ccd_sub(w1[i], w2[i], dir); }
  return npoly;
}

/* :1930 multicontact for a box-box pair: recovers up to 4 contacts from the EPA face `epa_face` by clipping the boxes' matching
 * faces (or an edge against a face).  w1[0] / w2[0] hold the EPA witness points on entry.  Returns the number of contacts. */
CCD_FN_NOINLINE int ccd_multicontact_box(CcdMulti* mc, const CcdPolytope* pt, int epa_face, const CcdGeom* g1, const CcdGeom* g2,
                                         creal (*w1)[3], creal (*w2)[3]) {
  const int face[3] = {epa_face & 0x3FF, (epa_face >> 10) & 0x3FF, (epa_face >> 20) & 0x3FF};
  int fi1[3], fi2[3];
  creal fv1[3][3], fv2[3][3], dir[3], dirn[3];
  int nface1 = ccd_feature_dim(face, pt, 0, fi1, fv1), nface2 = ccd_feature_dim(face, pt, 1, fi2, fv2);
  ccd_sub(dir, w2[0], w1[0]);
  dirn[0] = -dir[0]; dirn[1] = -dir[1]; dirn[2] = -dir[2];
  int nn1 = ccd_box_normals(nface1, fi1, g1->rot, dirn, mc->n1, mc->idx1);
  int nn2 = ccd_box_normals(nface2, fi2, g2->rot, dir, mc->n2, mc->idx2);
  int edge1 = 0, edge2 = 0, ri = 0, rj = 0, found = 0;
  for (int i = 0; i < nn1 && !found; i++) /* :1404 _aligned_faces */
    for (int j = 0; j < nn2; j++)
      if (ccd_dot(mc->n1[i], mc->n2[j]) < -CCD_FACE_TOL) { ri = i; rj = j; found = 1; break; }
  if (!found) {
    if (nface1 < 3 && nface1 <= nface2) { /* edge (or vertex) of geom 1 against a face of geom 2 */
      nn1 = ccd_box_edge_normals(nface1, g1, fv1[0], fv1[1], fi1[0], mc->n1, mc->endvert);
      for (int i = 0; i < nn2 && !found; i++) /* :1418 _aligned_face_edge(edge = n1, face = n2) */
        for (int j = 0; j < nn1; j++)
          if (CCD_FABS(ccd_dot(mc->n1[j], mc->n2[i])) < CCD_EDGE_TOL) { ri = j; rj = i; found = 1; break; }
      if (!found) return 1;
      edge1 = 1;
    } else if (nface2 < 3) { /* face of geom 1 against an edge of geom 2 */
      nn2 = ccd_box_edge_normals(nface2, g2, fv2[0], fv2[1], fi2[0], mc->n2, mc->endvert);
      for (int i = 0; i < nn1 && !found; i++)
        for (int j = 0; j < nn2; j++)
          if (CCD_FABS(ccd_dot(mc->n2[j], mc->n1[i])) < CCD_EDGE_TOL) { ri = j; rj = i; found = 1; break; }
      if (!found) return 1;
      edge2 = 1;
    } else {
      return 1;
    }
  }
  const int i = ri, j = rj;
  if (edge1) { /* :1913 _set_edge */
    ccd_copy(mc->face1[0], pt->vert[2 * face[0]]);
    ccd_copy(mc->face1[1], mc->endvert[i]);
    nface1 = 2;
  } else {
    nface1 = ccd_box_face(g1, edge2 ? mc->idx1[j] : mc->idx1[i], mc->face1);
  }
  if (edge2) {
    ccd_copy(mc->face2[0], pt->vert[2 * face[0] + 1]);
    ccd_copy(mc->face2[1], mc->endvert[i]);
    nface2 = 2;
  } else {
    nface2 = ccd_box_face(g2, mc->idx2[j], mc->face2);
  }
  const creal nd = ccd_norm(dir);
  creal approx[3];
  if (edge1) {
    for (int k = 0; k < 3; k++) approx[k] = nd * mc->n2[j][k];
    return ccd_polygon_clip(mc, mc->face2, nface2, mc->face1, nface1, mc->n2[j], approx, w1, w2);
  }
  if (edge2) {
    for (int k = 0; k < 3; k++) approx[k] = -nd * mc->n1[j][k];
    return ccd_polygon_clip(mc, mc->face1, nface1, mc->face2, nface2, mc->n1[j], approx, w1, w2);
  }
  for (int k = 0; k < 3; k++) approx[k] = nd * mc->n2[j][k];
  return ccd_polygon_clip(mc, mc->face1, nface1, mc->face2, nface2, mc->n1[i], approx, w1, w2);
}

#endif /* MJB_CCD_CORE_H_ */
