/* hmj_sincos.h — one deterministic fp64 sin/cos shared by the CUDA kernels and the CPU oracle.
 *
 * Why: mj_kinematics turns each hinge angle into a quaternion with sin/cos (mju_axisAngle2Quat),
 * and the geom frames that come out feed libccd's MPR, whose result for flat (face-face) contacts
 * flips on ulp-level input changes (support-function ties).  MuJoCo calls the platform libm; CUDA's
 * sincos and glibc's differ in the last bit, so "same state in => bit-identical contacts out" needs
 * ONE definition evaluated as the same IEEE-754 operation sequence on both sides.  This is the
 * classic Cody-Waite reduction + fdlibm minimax kernels (error < 2 ulp for |x| < 1e4); both
 * translation units that include it are built without FMA contraction (-fmad=false /
 * -ffp-contract=off), so every operation below is a single correctly-rounded add or multiply.
 */
#ifndef HMJ_SINCOS_H_
#define HMJ_SINCOS_H_

#include <math.h>

#ifdef __CUDACC__
#define HMJ_HD __host__ __device__ __forceinline__
#else
#define HMJ_HD static inline
#endif
/* device code: explicit round-to-nearest intrinsics are never contracted into FMA, whatever the -fmad setting */
#ifdef __CUDA_ARCH__
#define HMJ_M(a, b) __dmul_rn((a), (b))
#define HMJ_A(a, b) __dadd_rn((a), (b))
#define HMJ_S(a, b) __dsub_rn((a), (b))
#else
#define HMJ_M(a, b) ((a) * (b))
#define HMJ_A(a, b) ((a) + (b))
#define HMJ_S(a, b) ((a) - (b))
#endif

HMJ_HD void hmj_sincos(double x, double* sn, double* cs) {
  /* n = nearest integer to x / (pi/2); r = x - n pi/2 with a two-part pi/2 (first part has 33 bits: n*PIO2_1 is exact) */
  const double fn = rint(HMJ_M(x, 6.36619772367581382433e-01));
  const double r = HMJ_S(HMJ_S(x, HMJ_M(fn, 1.57079632673412561417e+00)), HMJ_M(fn, 6.07710050650619224932e-11));
  const double z = HMJ_M(r, r);
  /* fdlibm __kernel_sin / __kernel_cos coefficients, Horner form */
  double ps = HMJ_A(-2.50507602534068634195e-08, HMJ_M(z, 1.58969099521155010221e-10));
  ps = HMJ_A(2.75573137070700676789e-06, HMJ_M(z, ps));
  ps = HMJ_A(-1.98412698298579493134e-04, HMJ_M(z, ps));
  ps = HMJ_A(8.33333333332248946124e-03, HMJ_M(z, ps));
  ps = HMJ_A(-1.66666666666666324348e-01, HMJ_M(z, ps));
  double pc = HMJ_A(2.08757232129817482790e-09, HMJ_M(z, -1.13596475577881948265e-11));
  pc = HMJ_A(-2.75573143513906633035e-07, HMJ_M(z, pc));
  pc = HMJ_A(2.48015872894767294178e-05, HMJ_M(z, pc));
  pc = HMJ_A(-1.38888888888741095749e-03, HMJ_M(z, pc));
  pc = HMJ_A(4.16666666666666019037e-02, HMJ_M(z, pc));
  const double s = HMJ_A(r, HMJ_M(HMJ_M(r, z), ps));
  const double c = HMJ_A(HMJ_S(1.0, HMJ_M(0.5, z)), HMJ_M(HMJ_M(z, z), pc));
  const int q = ((int)fn) & 3;
  *sn = (q == 0) ? s : (q == 1) ? c : (q == 2) ? -s : -c;
  *cs = (q == 0) ? c : (q == 1) ? -s : (q == 2) ? -c : s;
}

#endif /* HMJ_SINCOS_H_ */
