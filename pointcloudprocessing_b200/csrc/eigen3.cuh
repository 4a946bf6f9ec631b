// eigen3.cuh — in-kernel symmetric 3×3 eigen-decomposition.
//
// Replaces Eigen::SelfAdjointEigenSolver<Matrix3f> at src/pcaAnalysis.cpp:57-59,98-100
// (iterative tridiagonal QR, up to 90 data-dependent steps).  Here: cyclic Jacobi in fp64,
// every rotation is the closed-form 2×2 symmetric Schur decomposition, and the sweep count
// is bounded (PCP_JACOBI_SWEEPS; 3×3 Jacobi converges quadratically, 4-5 sweeps reach fp64
// round-off).  The fp32 moment matrix is bitwise the reference's; solving it in fp64 gives
// its exact eigenpairs to ~1e-16·λmax, so the difference to the reference is the reference
// solver's own fp32 error.  Eigenvalues ascending, eigenvectors unit-norm, sign unspecified
// (as upstream).
#pragma once
#include <cuda_runtime.h>

#define PCP_JACOBI_SWEEPS 8

struct Sym3Eig {
  float val[3];     // ascending
  float vec[3][3];  // vec[k][r] = component r of eigenvector k
};

// rotate rows/cols (p,q) of symmetric A (app, aqq, apq on the pair; apr, aqr the couplings to
// the third index r) and columns p,q of V.
#define PCP_JACOBI_ROT(app, aqq, apq, apr, aqr, vp0, vp1, vp2, vq0, vq1, vq2)                 \
  do {                                                                                        \
    if (apq != 0.0) {                                                                         \
      double theta = (aqq - app) / (2.0 * apq);                                               \
      double t = 1.0 / (fabs(theta) + sqrt(theta * theta + 1.0));                             \
      if (theta < 0.0) t = -t;                                                                \
      if (!(fabs(theta) < 1e150)) t = 0.5 / theta; /* huge ratio: avoid inf*0 */              \
      double c = rsqrt(t * t + 1.0);                                                          \
      double s = t * c;                                                                       \
      double h = t * apq;                                                                     \
      app -= h;                                                                               \
      aqq += h;                                                                               \
      apq = 0.0;                                                                              \
      double n_pr = c * apr - s * aqr;                                                        \
      double n_qr = s * apr + c * aqr;                                                        \
      apr = n_pr;                                                                             \
      aqr = n_qr;                                                                             \
      double w;                                                                               \
      w = c * vp0 - s * vq0; vq0 = s * vp0 + c * vq0; vp0 = w;                                \
      w = c * vp1 - s * vq1; vq1 = s * vp1 + c * vq1; vp1 = w;                                \
      w = c * vp2 - s * vq2; vq2 = s * vp2 + c * vq2; vp2 = w;                                \
    }                                                                                         \
  } while (0)

// C = {c00, c01, c02, c11, c12, c22} (fp32 moment); result cast to fp32 at the end.
__device__ __forceinline__ void sym3_eigen(const float C[6], Sym3Eig& out) {
  double a00 = C[0], a01 = C[1], a02 = C[2], a11 = C[3], a12 = C[4], a22 = C[5];
  double v00 = 1, v01 = 0, v02 = 0;  // v{row}{col}; column k = eigenvector k
  double v10 = 0, v11 = 1, v12 = 0;
  double v20 = 0, v21 = 0, v22 = 1;
#pragma unroll 1
  for (int sweep = 0; sweep < PCP_JACOBI_SWEEPS; ++sweep) {
    double off = a01 * a01 + a02 * a02 + a12 * a12;
    double dg = a00 * a00 + a11 * a11 + a22 * a22;
    if (!(off > dg * 1e-34)) break;  // also exits on NaN and on the zero matrix
    PCP_JACOBI_ROT(a00, a11, a01, a02, a12, v00, v10, v20, v01, v11, v21);
    PCP_JACOBI_ROT(a00, a22, a02, a01, a12, v00, v10, v20, v02, v12, v22);
    PCP_JACOBI_ROT(a11, a22, a12, a01, a02, v01, v11, v21, v02, v12, v22);
  }
  // ascending order (3-element network), carrying the columns
  double e0 = a00, e1 = a11, e2 = a22;
#define PCP_SWAPCOL(ea, eb, xa0, xa1, xa2, xb0, xb1, xb2) \
  if (eb < ea) {                                          \
    double t_;                                            \
    t_ = ea; ea = eb; eb = t_;                            \
    t_ = xa0; xa0 = xb0; xb0 = t_;                        \
    t_ = xa1; xa1 = xb1; xb1 = t_;                        \
    t_ = xa2; xa2 = xb2; xb2 = t_;                        \
  }
  PCP_SWAPCOL(e0, e1, v00, v10, v20, v01, v11, v21)
  PCP_SWAPCOL(e1, e2, v01, v11, v21, v02, v12, v22)
  PCP_SWAPCOL(e0, e1, v00, v10, v20, v01, v11, v21)
#undef PCP_SWAPCOL
  out.val[0] = (float)e0; out.val[1] = (float)e1; out.val[2] = (float)e2;
  out.vec[0][0] = (float)v00; out.vec[0][1] = (float)v10; out.vec[0][2] = (float)v20;
  out.vec[1][0] = (float)v01; out.vec[1][1] = (float)v11; out.vec[1][2] = (float)v21;
  out.vec[2][0] = (float)v02; out.vec[2][1] = (float)v12; out.vec[2][2] = (float)v22;
}

// One shared copy of the solver for the kernels that would otherwise inline it several times.
static __device__ __noinline__ void sym3_eigen_call(const float C[6], Sym3Eig& out) { sym3_eigen(C, out); }
