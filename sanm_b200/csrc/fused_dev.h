/* sanm_b200/csrc/fused_dev.h -- register-resident Taylor propagation of the
 * three stress graphs `fea` runs (fea/material.cpp:55-99 behind
 * fea/mesh_template.h:191-219):  F = (x + bias) DmInv,  P = pk1(F).
 *
 * The generic path (taylor_dev.h) keeps every VarNode of the graph as a stack
 * in HBM, as the reference does (libsanm/symbolic.h:48-50).  Here one thread
 * owns one tetrahedron, every intermediate lives in registers and only the
 * minimal Taylor state that later orders must re-read goes to HBM:
 *
 *   neo-Hookean compressible   F, G = F^-1, L = log det F              S = 19
 *   neo-Hookean incompressible F, G, L, J, Q = J^p, Ic = |F|^2, T       S = 23
 *   ARAP (polar mode)          F, Ps (symmetric factor), W (rotation)  S = 27
 *
 * Recurrences (restating the reference operators, see taylor_dev.h for the
 * per-operator citations), with x_k unknown during the bias phase:
 *   G_k  = (-sum_{i=1}^{k-1} G_i F_{k-i} - G_0 F_k) G_0     BatchMatInvMul
 *   L_k  = <G_0^T, F_k> + (1/k) sum_{i=1}^{k-1} (k-i) <G_i^T, F_{k-i}>
 *          -- (log det F)' = tr(F^-1 F'); equals BatchDeterminant followed by
 *          the log recurrence of libsanm/analytic_unary.cpp:25-33 but needs
 *          neither the determinant stack nor the O(k^2) polynomial expansion
 *          of libsanm/tensor_polymat.cpp:138-341
 *   J_k  = J_0 L_k + (1/k) sum m L_m J_{k-m}                 J' = J L'
 *   Q_k  = p (Q_0 L_k + (1/k) sum m L_m Q_{k-m})             Q' = p Q L'
 *   Ic_k = sum <F_i, F_{k-i}>;  products by Cauchy convolution (Multiply)
 *   ARAP: D = sum M_i M_{k-i}^T - Ps_i^T Ps_{k-i}, Bpw = sum Ps_i W_{k-i},
 *         then svd_w_taylor_fwd_p (libsanm/tensor_svd.cpp:389-475)
 * Every order-k quantity is affine in F_k, so the bias phase stores the value
 * for F_k = 0 ("bias part") and the coefficient phase adds the tangent.
 *
 * HBM layout: tiles of TILE tets; element c of order k of tile T, lane l at
 *     arena[((T * (K+1) + k) * S + c) * TILE + l]
 * so a warp reads 32 consecutive doubles and one (tile, order) slot is one
 * contiguous S*TILE*8-byte block.
 */
#pragma once
#ifndef SANM_STREAMING_SLOTS
#define SANM_STREAMING_SLOTS 1
#endif

#include "taylor_dev.h"

namespace sanm_b200 {
namespace fused {

constexpr int TILE = 128;

enum Material : int { MAT_NONE = -1, MAT_NHC = 0, MAT_NHI = 1, MAT_ARAP = 2 };

struct Params {
    double c[5];  //!< material-specific linear coefficients, see each struct
    double p;     //!< NHI: exponent of J
    int require_rotation;  //!< ARAP
};

using taylor::clip_div;
using taylor::det3;
using taylor::inv3;

template <int N>
SANM_HD inline void ld_strided(double* dst, const double* p) {
#pragma unroll
    for (int c = 0; c < N; ++c) {
#if defined(__CUDA_ARCH__) && SANM_STREAMING_SLOTS
        dst[c] = __ldcs(p + c * TILE);  // read once per launch: evict first
#else
        dst[c] = p[c * TILE];
#endif
    }
}
template <int N>
SANM_HD inline void st_strided(double* p, const double* src) {
#pragma unroll
    for (int c = 0; c < N; ++c) {
#if defined(__CUDA_ARCH__) && SANM_STREAMING_SLOTS
        __stcs(p + c * TILE, src[c]);
#else
        p[c * TILE] = src[c];
#endif
    }
}

//! C += s * A B (3x3 row-major)
SANM_HD inline void mm3_acc(double* C, const double* A, const double* B,
                            double s) {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            C[r * 3 + c] += s * (A[r * 3] * B[c] + A[r * 3 + 1] * B[3 + c] +
                                 A[r * 3 + 2] * B[6 + c]);
        }
    }
}

//! <A^T, B> = sum_{r,c} A[c][r] B[r][c] = tr(A B)
SANM_HD inline double tr_mm3(const double* A, const double* B) {
    double s = 0;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) s += A[c * 3 + r] * B[r * 3 + c];
    }
    return s;
}

SANM_HD inline double dot9(const double* A, const double* B) {
    double s = 0;
#pragma unroll
    for (int q = 0; q < 9; ++q) s += A[q] * B[q];
    return s;
}

//! R = -G0 Fk G0
SANM_HD inline void neg_gfg(double* R, const double* G0, const double* Fk) {
    double t[9];
    taylor::mm3(t, G0, Fk);
    double u[9];
    taylor::mm3(u, t, G0);
#pragma unroll
    for (int q = 0; q < 9; ++q) R[q] = -u[q];
}

/* ------------------------------------------------ neo-Hookean compressible
 * P = c0 F + c1 G^T + c2 L G^T      (fea/material.cpp:82-87: mu, -mu, lambda)
 */
struct NHC {
    static constexpr int S = 19;
    //! resident 128-thread blocks per SM the order-step kernel is compiled for
    //! (__launch_bounds__(128, MINB)); measured on B200, 2 M tets, K = 20:
    //! unbounded (ptxas picks 128 regs = 4 blocks) 2.43 G tet*order/s, 4: 2.43,
    //! 5: 2.49, 6: 2.33 (spills); with slot 0 reloaded per phase (fused.cu
    //! load_base) 4: 2.37, 5: 2.56 (96 registers, no spill), 6: 2.52
    static constexpr int MINB = 5;
    static constexpr int AUX = 0;
    enum { oF = 0, oG = 9, oL = 18 };
    struct Acc {
        double G[9], M[9], L;
    };
    SANM_HD static void zero(Acc& a) {
#pragma unroll
        for (int q = 0; q < 9; ++q) a.G[q] = a.M[q] = 0;
        a.L = 0;
    }
    SANM_HD static void eval0(const double* F0, double* s0, double* /*aux*/,
                              double* P0, const Params& pr, int* /*err*/) {
        double* G0 = s0 + oG;
#pragma unroll
        for (int q = 0; q < 9; ++q) s0[oF + q] = F0[q];
        inv3(G0, F0);
        const double L0 = log(det3(F0));
        s0[oL] = L0;
        const double w = pr.c[1] + pr.c[2] * L0;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                P0[r * 3 + c] = pr.c[0] * F0[r * 3 + c] + w * G0[c * 3 + r];
    }
    //! contribution of the ordered pair (i, j), i + j = kb
    SANM_HD static void accum(Acc& a, const double* Si, const double* Sj,
                              int /*i*/, int j, const Params& /*pr*/) {
        const double* Gi = Si + oG;
        const double* Fj = Sj + oF;
        mm3_acc(a.G, Gi, Fj, -1.0);
        a.L += double(j) * tr_mm3(Gi, Fj);
        const double Lj = Sj[oL];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) a.M[r * 3 + c] += Lj * Gi[c * 3 + r];
    }
    //! bias parts of state[9..S) and of the output for order kb
    SANM_HD static void bias(const Acc& a, const double* s0,
                             const double* /*aux*/, int kb, const Params& pr,
                             double* bs, double* Pb) {
        const double* G0 = s0 + oG;
        double* Gb = bs + oG;
        taylor::mm3(Gb, a.G, G0);
        const double Lb = a.L / double(kb);
        bs[oL] = Lb;
        const double w = pr.c[1] + pr.c[2] * s0[oL];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                Pb[r * 3 + c] = w * Gb[c * 3 + r] +
                                pr.c[2] * (Lb * G0[c * 3 + r] + a.M[r * 3 + c]);
    }
    //! tangent parts for a given F_k
    SANM_HD static void tangent(const double* Fk, const double* s0,
                                const double* /*aux*/, const Params& pr,
                                double* ts, double* Pt) {
        const double* G0 = s0 + oG;
        double* Gt = ts + oG;
        neg_gfg(Gt, G0, Fk);
        const double Lt = tr_mm3(G0, Fk);
        ts[oL] = Lt;
        const double w = pr.c[1] + pr.c[2] * s0[oL];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                Pt[r * 3 + c] = pr.c[0] * Fk[r * 3 + c] + w * Gt[c * 3 + r] +
                                pr.c[2] * Lt * G0[c * 3 + r];
    }
};

/* ---------------------------------------------- neo-Hookean incompressible
 * Q = J^p, Ic = sum F^2, T = c2 Q Ic + c3 J J + c4 J, P = c0 Q F + c1 T G^T
 * (fea/material.cpp:70-81: mu, 1, -mu/3, k, -k, p = -2/3)
 */
struct NHI {
    static constexpr int S = 23;
    //! unbounded (236 regs = 2 blocks) 1.61 G, 4: 1.85, 5: 1.61; with slot 0
    //! reloaded per phase 4: 1.99 (128 registers, no spill), 5: 1.75, 6: 1.47
    static constexpr int MINB = 4;
    static constexpr int AUX = 0;
    enum { oF = 0, oG = 9, oL = 18, oJ = 19, oQ = 20, oI = 21, oT = 22 };
    //! P = c0 sum Q_i F_j + c1 sum T_j G_i^T: the two products only ever
    //! enter the stress bias in this combination, so they share one
    //! accumulator (24 doubles of accumulators instead of 33: registers that
    //! hold loads in flight instead)
    struct Acc {
        double G[9], P[9], L, J, Q, Ic, A, JJ;
    };
    SANM_HD static void zero(Acc& a) {
#pragma unroll
        for (int q = 0; q < 9; ++q) a.G[q] = a.P[q] = 0;
        a.L = a.J = a.Q = a.Ic = a.A = a.JJ = 0;
    }
    SANM_HD static void eval0(const double* F0, double* s0, double* /*aux*/,
                              double* P0, const Params& pr, int* err) {
        double* G0 = s0 + oG;
#pragma unroll
        for (int q = 0; q < 9; ++q) s0[oF + q] = F0[q];
        inv3(G0, F0);
        const double J0 = det3(F0);
        if (fabs(J0) < 1e-3) {
            // libsanm/analytic_unary.cpp:116-120: 0^p with non-integer p
            taylor::rt_flag_or(err, 1);
        }
        const double Q0 = taylor::pow_ref(J0, pr.p);
        const double Ic0 = dot9(F0, F0);
        const double T0 = pr.c[2] * Q0 * Ic0 + pr.c[3] * J0 * J0 + pr.c[4] * J0;
        s0[oL] = 0;  // log J itself is never needed, only its derivatives
        s0[oJ] = J0;
        s0[oQ] = Q0;
        s0[oI] = Ic0;
        s0[oT] = T0;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                P0[r * 3 + c] = pr.c[0] * Q0 * F0[r * 3 + c] +
                                pr.c[1] * T0 * G0[c * 3 + r];
    }
    SANM_HD static void accum(Acc& a, const double* Si, const double* Sj,
                              int /*i*/, int j, const Params& pr) {
        const double* Gi = Si + oG;
        const double* Fi = Si + oF;
        const double* Fj = Sj + oF;
        const double Ji = Si[oJ], Qi = Si[oQ];
        const double Lj = Sj[oL], Jj = Sj[oJ], Ij = Sj[oI], Tj = Sj[oT];
        mm3_acc(a.G, Gi, Fj, -1.0);
        a.L += double(j) * tr_mm3(Gi, Fj);
        const double jl = double(j) * Lj;
        a.J += jl * Ji;
        a.Q += jl * Qi;
        a.Ic += dot9(Fi, Fj);
        a.A += Qi * Ij;
        a.JJ += Ji * Jj;
        const double cq = pr.c[0] * Qi, ct = pr.c[1] * Tj;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                a.P[r * 3 + c] += cq * Fj[r * 3 + c] + ct * Gi[c * 3 + r];
    }
    SANM_HD static void scalars(double L, double sJ, double sQ, double Ic,
                                double sA, double sJJ, const double* s0,
                                const Params& pr, double* out) {
        const double J0 = s0[oJ], Q0 = s0[oQ], I0 = s0[oI];
        const double J = J0 * L + sJ;
        const double Q = pr.p * (Q0 * L + sQ);
        const double A = Q0 * Ic + Q * I0 + sA;
        const double JJ = 2 * J0 * J + sJJ;
        out[oL] = L;
        out[oJ] = J;
        out[oQ] = Q;
        out[oI] = Ic;
        out[oT] = pr.c[2] * A + pr.c[3] * JJ + pr.c[4] * J;
    }
    SANM_HD static void bias(const Acc& a, const double* s0,
                             const double* /*aux*/, int kb, const Params& pr,
                             double* bs, double* Pb) {
        const double* G0 = s0 + oG;
        const double* F0 = s0 + oF;
        double* Gb = bs + oG;
        taylor::mm3(Gb, a.G, G0);
        const double ik = 1.0 / double(kb);
        scalars(a.L * ik, a.J * ik, a.Q * ik, a.Ic, a.A, a.JJ, s0, pr, bs);
        const double Qb = bs[oQ], Tb = bs[oT], T0 = s0[oT];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                Pb[r * 3 + c] =
                        pr.c[0] * Qb * F0[r * 3 + c] +
                        pr.c[1] * (T0 * Gb[c * 3 + r] + Tb * G0[c * 3 + r]) +
                        a.P[r * 3 + c];
    }
    SANM_HD static void tangent(const double* Fk, const double* s0,
                                const double* /*aux*/, const Params& pr,
                                double* ts, double* Pt) {
        const double* G0 = s0 + oG;
        const double* F0 = s0 + oF;
        double* Gt = ts + oG;
        neg_gfg(Gt, G0, Fk);
        scalars(tr_mm3(G0, Fk), 0, 0, 2 * dot9(F0, Fk), 0, 0, s0, pr, ts);
        const double Qt = ts[oQ], Tt = ts[oT], Q0 = s0[oQ], T0 = s0[oT];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                Pt[r * 3 + c] =
                        pr.c[0] * (Q0 * Fk[r * 3 + c] + Qt * F0[r * 3 + c]) +
                        pr.c[1] * (T0 * Gt[c * 3 + r] + Tt * G0[c * 3 + r]);
    }
};

/* --------------------------------------------------------- ARAP, polar mode
 * F = Ps W (Ps symmetric, W rotation), P = c0 F + c1 W
 * (fea/material.cpp:88-90: mu, -mu; BatchSVDW with require_rotation)
 * aux = [U0 (9), S0 (3)]
 */
struct ARAP {
    static constexpr int S = 27;
    //! unbounded (240 regs = 2 blocks) 1.43 G, 4: 1.56, 5: 1.55; with slot 0
    //! reloaded per phase 4: 1.67, 5: 1.68, 6: 1.63
    static constexpr int MINB = 5;
    static constexpr int AUX = 12;
    enum { oF = 0, oP = 9, oW = 18 };
    struct Acc {
        double D[9], Bpw[9];
    };
    SANM_HD static void zero(Acc& a) {
#pragma unroll
        for (int q = 0; q < 9; ++q) a.D[q] = a.Bpw[q] = 0;
    }
    SANM_HD static void eval0(const double* F0, double* s0, double* aux,
                              double* P0, const Params& pr, int* /*err*/) {
        double* U = aux;
        double* Sv = aux + 9;
        double* W = s0 + oW;
        taylor::svdw3(F0, pr.require_rotation != 0, U, Sv, W);
#pragma unroll
        for (int q = 0; q < 9; ++q) s0[oF + q] = F0[q];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                s0[oP + r * 3 + c] = U[r * 3] * Sv[0] * U[c * 3] +
                                     U[r * 3 + 1] * Sv[1] * U[c * 3 + 1] +
                                     U[r * 3 + 2] * Sv[2] * U[c * 3 + 2];
#pragma unroll
        for (int q = 0; q < 9; ++q) P0[q] = pr.c[0] * F0[q] + pr.c[1] * W[q];
    }
    SANM_HD static void accum(Acc& a, const double* Si, const double* Sj,
                              int /*i*/, int /*j*/, const Params& /*pr*/) {
        const double *Mi = Si + oF, *Pi = Si + oP;
        const double *Mj = Sj + oF, *Pj = Sj + oP, *Wj = Sj + oW;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c)
                a.D[r * 3 + c] +=
                        (Mi[r * 3] * Mj[c * 3] + Mi[r * 3 + 1] * Mj[c * 3 + 1] +
                         Mi[r * 3 + 2] * Mj[c * 3 + 2]) -
                        (Pi[r] * Pj[c] + Pi[3 + r] * Pj[3 + c] +
                         Pi[6 + r] * Pj[6 + c]);
        mm3_acc(a.Bpw, Pi, Wj, 1.0);
    }
    SANM_HD static void bias(const Acc& a, const double* s0, const double* aux,
                             int /*kb*/, const Params& pr, double* bs,
                             double* Pb) {
        const double Z[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        taylor::svdw3_fwd_p(bs + oP, bs + oW, Z, aux, aux + 9, s0 + oW, a.D,
                            a.Bpw);
#pragma unroll
        for (int q = 0; q < 9; ++q) Pb[q] = pr.c[1] * bs[oW + q];
    }
    SANM_HD static void tangent(const double* Fk, const double* s0,
                                const double* aux, const Params& pr, double* ts,
                                double* Pt) {
        const double Z[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        taylor::svdw3_fwd_p(ts + oP, ts + oW, Fk, aux, aux + 9, s0 + oW, Z, Z);
#pragma unroll
        for (int q = 0; q < 9; ++q)
            Pt[q] = pr.c[0] * Fk[q] + pr.c[1] * ts[oW + q];
    }
};

}  // namespace fused
}  // namespace sanm_b200
