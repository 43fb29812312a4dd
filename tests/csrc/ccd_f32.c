/* Test helper: the FLOAT32 build of mujoco_warp_b200/csrc/mjb_ccd_core.h compiled for the host, i.e. the arithmetic the CUDA
 * collision kernel runs (same source, same precision), so tests/test_oracle_ccd_golden.py can quantify the float32 error of
 * GJK / EPA against the reference's outputs without a GPU.  Not part of the product. */
#include <math.h>
#include <string.h>
#define CCD_REAL float
#define CCD_FN static inline
#define CCD_FN_NOINLINE static
#define CCD_SQRT sqrtf
#define CCD_FABS fabsf
#include "../../mujoco_warp_b200/csrc/mjb_ccd_core.h"

int ccd32(int t1, int t2, const double* p1, const double* r1, const double* s1, const double* v1, int nv1, const double* p2,
          const double* r2, const double* s2, const double* v2, int nv2, double margin, double cutoff, double tolerance, int iterations,
          double* dist, double* x1, double* x2) {
  static _Thread_local CcdPolytope pt;
  static _Thread_local float fv1[3 * 64], fv2[3 * 64];
  CcdGeom g1, g2;
  CcdGjk res;
  float d, a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
  if (nv1 > 64 || nv2 > 64) return -1;
  for (int i = 0; i < 3; i++) { g1.pos[i] = (float)p1[i]; g2.pos[i] = (float)p2[i]; g1.size[i] = (float)s1[i]; g2.size[i] = (float)s2[i]; }
  for (int i = 0; i < 9; i++) { g1.rot[i] = (float)r1[i]; g2.rot[i] = (float)r2[i]; }
  for (int i = 0; i < 3 * nv1; i++) fv1[i] = (float)v1[i];
  for (int i = 0; i < 3 * nv2; i++) fv2[i] = (float)v2[i];
  g1.margin = g2.margin = (float)margin; g1.type = t1; g2.type = t2;
  g1.vert = fv1; g1.nvert = nv1; g2.vert = fv2; g2.nvert = nv2;
  g1.graph = g2.graph = 0;
  int face_idx;
  int n = ccd_run((float)tolerance, (float)cutoff, iterations, iterations, g1, g2, &pt, &res, &d, a, b, &face_idx);
  *dist = d;
  for (int i = 0; i < 3; i++) { x1[i] = a[i]; x2[i] = b[i]; }
  return n;
}
