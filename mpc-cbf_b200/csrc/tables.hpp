// Host-side constant tables of the MPC-CBF QP (built once per handle, FP64 results computed in
// long double). This replaces what the reference rebuilds on every optimize() call:
//   PiecewiseBezierMPCQPOperations ctor      mpc/src/optimization/PiecewiseBezierMPCQPOperations.cpp:9-60
//   positionErrorPenaltyCost                 same file :64-91
//   integratedSquaredDerivativeCost          splines/src/optimization/BezierQPOperations.cpp:73-111
//   evalConstraint / evalBound / continuity  splines/.../BezierQPOperations.cpp:149-186,
//                                            mpc/src/optimization/PiecewiseBezierMPCQPGenerator.cpp:148-226
//   DoubleIntegrator get_A0 / get_lambda     model/src/DoubleIntegrator.cpp:9-51
// (paths relative to /root/reference/workspace/lib).
//
// Key structural fact used throughout: cost, equality rows and bound rows are identical and
// decoupled for the three axes (x, y, yaw); only the CBF rows couple x and y. Everything is
// therefore derived for ONE axis (m1 = pieces*ctrl_pts variables) and replicated.
#pragma once
#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/mpccbf.h"

namespace mpccbf {

typedef long double xreal;

struct XMat {
    int rows = 0, cols = 0;
    std::vector<xreal> a;
    XMat() {}
    XMat(int r, int c) : rows(r), cols(c), a(static_cast<size_t>(r) * c, 0.0L) {}
    xreal& operator()(int i, int j) { return a[static_cast<size_t>(i) * cols + j]; }
    xreal operator()(int i, int j) const { return a[static_cast<size_t>(i) * cols + j]; }
};

XMat xmul(const XMat& A, const XMat& B);
XMat xtrans(const XMat& A);

// d-th derivative of the degree-(nctrl-1) Bernstein basis on [0,T] at local parameter t.
void bernstein_deriv(int nctrl, xreal T, xreal t, int d, xreal* out);

struct HostTables {
    mpccbf_config cfg;
    int K = 0, p = 0, c = 0, deg = 0;
    int m1 = 0;        // per-axis decision variables  (p*c)
    int n = 0;         // full decision vector         (3*m1), layout x[piece*3c + axis*c + cp]
    int me1 = 0;       // per-axis equality rows: 2 + (deg+1)(p-1)
    int r1 = 0;        // per-axis free dof: m1 - me1
    bool reducible = false;  // per-axis equality block has full row rank
    std::vector<double> h_samples;   // [K]   Eigen LinSpaced(K, 0, (K-1)h)
    std::vector<int> piece_of_k;     // [K]   lower_bound piece lookup
    std::vector<double> local_of_k;  // [K]   local parameter inside that piece

    // ---- per-axis tables (long double) -----------------------------------------------------
    XMat B[3];     // [d] (K x m1): row k = d-th derivative basis at t_k   (d = 0 pos, 1 vel, 2 acc)
    XMat E1;       // (me1 x m1) equality rows: pos(0), vel(0), then continuity per knot, d = 0..deg
    XMat P1;       // (m1 x m1)  objective x'P1x + q'x (reference convention)
    XMat Phi1;     // (K x m1)   Lambda_pos * U
    std::vector<xreal> Qd;   // [K] diagonal of the tracking weight
    std::vector<xreal> tau;  // [K] (k+1)h : free-motion position = pos + tau_k vel (A0_pos)
    XMat Z1, Y1;   // (m1 x r1), (m1 x 2): x_axis = Y1 [pos;vel] + Z1 w
    XMat L1;       // (r1 x r1) Cholesky factor of H1 = 2 Z1' P1 Z1
    XMat W1;       // (m1 x r1) = Z1 L1^-T : x_axis = Y1 s + W1 v, objective Hessian = I in v
    XMat Vs, Vr;   // (r1 x 2), (r1 x K): unconstrained minimiser v0 = Vs s + Vr ref_axis
    XMat Ct[3], Ot[3];  // [d] (K x r1), (K x 2): row value = Ct[d][k].v + Ot[d][k].s
    XMat Kc, Kr;   // (2 x 2), (2 x K): constant objective part c0 = s'Kc s + (Kr ref).s

    std::string error;

    bool build(const mpccbf_config& cfg);

    // ---- full-space views (double, row-major) for the dense path, tests and integrators -------
    int var_index(int piece, int axis, int cp) const { return piece * 3 * c + axis * c + cp; }
    std::vector<double> full_U() const;       // [3K][n]  acceleration sampling basis (U_basis_)
    std::vector<double> full_basis(int d) const;  // [3K][n]
    std::vector<double> full_P() const;       // [n][n]
    std::vector<double> full_E() const;       // [3*me1][n]  reference insertion order
    std::vector<double> full_Afix() const;    // [6K][n]     acc rows then vel rows
    std::vector<double> full_Z() const;       // [n][3*r1]   column = axis*r1 + j
    std::vector<double> full_Y() const;       // [n][6]      x = Y [pos;vel] + Z w
    std::vector<double> full_Hred() const;    // [3r1][3r1]
    // linear cost q (n) for one agent: q = Phi' Q (2 A0 x0 - 2 ref)
    void linear_cost(const double* x0, const double* ref, double* q) const;
    int num_eq_rows() const { return 3 * me1; }
};

}  // namespace mpccbf
