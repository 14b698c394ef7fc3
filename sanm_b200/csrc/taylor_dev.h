/* sanm_b200/csrc/taylor_dev.h -- device-side Taylor-coefficient propagation.
 *
 * One thread owns one batch item (one tetrahedron in `fea`) and walks the
 * lowered operator list.  All per-item data is structure-of-arrays:
 * element c of order k of variable v for item t lives at
 *     base(v, k)[c * npad + t]
 * so that the 32 lanes of a warp touch 32 consecutive doubles (256 B) on every
 * load/store.
 *
 * The math of each operator restates the reference's OperatorMeta
 * implementations (eval_bias / compute_order_bias / compute_coeff):
 *   LinearCombination  libsanm/oprs/elem_arith.cpp:45-106
 *   Multiply           libsanm/oprs/elem_arith.cpp:133-209
 *   AnalyticUnary      libsanm/oprs/analytic_unary.cpp:116-158,
 *                      libsanm/analytic_unary.cpp:13-139 (log / pow recurrences)
 *   Reduce             libsanm/oprs/reduce.cpp:30-90
 *   BatchMatMul        libsanm/oprs/linalg.cpp:15-40,330-409
 *   BatchMatInvMul     libsanm/oprs/linalg.cpp:66-197
 *   BatchDeterminant   libsanm/oprs/linalg.cpp:222-273 (value of
 *                      libsanm/tensor_polymat.cpp:344-379, computed here with an
 *                      O(k) cross-product stack instead of the O(k^2) expansion)
 *   BatchMatTranspose  libsanm/oprs/linalg.cpp:286-326
 *   BatchMulEye        libsanm/oprs/linalg.cpp:411-466
 *   BatchSVDW          libsanm/oprs/linalg.cpp:468-600,
 *                      libsanm/tensor_svd.cpp:48-145,389-475 (polar mode)
 *   Slice / Concat     libsanm/oprs/misc.cpp:93-331
 *
 * Propagation protocol per order k >= 1 (libsanm/symbolic.cpp:162-289):
 *   phase A ("bias", x_k unknown): every op computes its self-bias from the
 *     stacks of orders < k (the convolution part), caches it, and propagates
 *     out_k = Lin(in_k) + self_bias with in_k := bias of the inputs;
 *   phase B ("coeff", x_k known): out_k = Lin(in_k) + cached self_bias.
 * Both phases write slot k of the output variable; phase B overwrites A.
 */
#pragma once

#include "rt.h"

namespace sanm_b200 {
namespace taylor {

constexpr int MAX_IN = 8;      //!< max inputs of one op (LinearCombination / Concat)
constexpr int MAXD = 7;        //!< max matrix dim handled by the generic paths
//! max elements per item of any variable (matrices up to MAXD x MAXD; plain
//! element-wise / reduction operands up to 64: the reference's
//! tests/symbolic.cpp reduces 9 x 7 items)
constexpr int MAXE = 64;
static_assert(MAXE >= MAXD * MAXD, "MAXE");
constexpr int MAX_ORDER = 48;  //!< bound for per-thread polynomial scratch

enum OpType : int {
    OP_PLACEHOLDER = 0,
    OP_CONSTANT,
    OP_LINCOMB,
    OP_MULTIPLY,
    OP_POW,
    OP_LOG,
    OP_REDUCE_SUM,
    OP_MATMUL,
    OP_TRANSPOSE,
    OP_DET,
    OP_MULEYE,
    OP_MATINVMUL,
    OP_SVDW,
    OP_SLICE,
    OP_CONCAT,
};

//! device descriptor of one variable
struct DVar {
    double* p0;       //!< order 0, [e][npad]
    double* pk;       //!< orders >= 1: pk + (k-1)*kstride
    int64_t kstride;  //!< 0 when the variable keeps no stack (orders alias)
    int e;            //!< elements per item
    int d1, d2;       //!< per-item matrix dims (vector: d2 == 1)
    int is_const;     //!< does not depend on the placeholder
};

//! device descriptor of one operator
struct DOp {
    int type;
    int nin, nout;
    int in[MAX_IN];
    int out[3];
    double f[MAX_IN + 1];  //!< LINCOMB: coeffs + bias; POW: exponent
    int i[6];              //!< op-specific integer params
    int const_op;          //!< every input is constant: evaluated at order 0 only
    double* sb;            //!< cached self-bias, [esb][npad]
    double* aux0;          //!< order-0 cache (derivative, inverse, cofactor, U0/S0)
    double* auxk;          //!< per-order hidden stack (det cross products, polar P)
    int64_t auxk_stride;
    int* flag;             //!< POW: has_zero flag
};

struct DProgram {
    const DVar* vars;
    const DOp* ops;
    int nops;
    int64_t npad;
    int* err;  //!< device error word: bit0 = 0^p with non-integer p
};

SANM_HD inline double* vbase(const DVar& v, int k) {
    return k == 0 ? v.p0 : v.pk + (int64_t)(k - 1) * v.kstride;
}

/* ------------------------------------------------------------------ helpers */

SANM_HD inline void rt_flag_or(int* p, int bits) {
#if defined(__CUDA_ARCH__)
    atomicOr(p, bits);
#else
    *p |= bits;
#endif
}

//! load a whole per-item matrix (E elements) of variable v, order k
template <int E>
SANM_HD inline void ld_mat(double* dst, const DVar& v, int k, int64_t np,
                           size_t t, int e_rt) {
    const double* p = vbase(v, k) + t;
    const int e = E ? E : e_rt;
#pragma unroll
    for (int c = 0; c < e; ++c) dst[c] = p[(int64_t)c * np];
}

template <int E>
SANM_HD inline void ld_buf(double* dst, const double* base, int64_t np,
                           size_t t, int e_rt) {
    const double* p = base + t;
    const int e = E ? E : e_rt;
#pragma unroll
    for (int c = 0; c < e; ++c) dst[c] = p[(int64_t)c * np];
}

template <int E>
SANM_HD inline void st_buf(double* base, const double* src, int64_t np,
                           size_t t, int e_rt) {
    double* p = base + t;
    const int e = E ? E : e_rt;
#pragma unroll
    for (int c = 0; c < e; ++c) p[(int64_t)c * np] = src[c];
}

//! C (+)= A[m,k] * B[k,n], row-major
template <int M, int K, int N>
SANM_HD inline void mm_acc(double* C, const double* A, const double* B, int m,
                           int kk, int n) {
    const int mm = M ? M : m, kd = K ? K : kk, nn = N ? N : n;
#pragma unroll
    for (int r = 0; r < mm; ++r) {
#pragma unroll
        for (int c = 0; c < nn; ++c) {
            double s = C[r * nn + c];
#pragma unroll
            for (int p = 0; p < kd; ++p) s += A[r * kd + p] * B[p * nn + c];
            C[r * nn + c] = s;
        }
    }
}

template <int E>
SANM_HD inline void zero_arr(double* a, int e_rt) {
    const int e = E ? E : e_rt;
#pragma unroll
    for (int c = 0; c < e; ++c) a[c] = 0;
}

SANM_HD inline double det3(const double* a) {
    return a[0] * (a[4] * a[8] - a[5] * a[7]) -
           a[1] * (a[3] * a[8] - a[5] * a[6]) +
           a[2] * (a[3] * a[7] - a[4] * a[6]);
}

//! cofactor matrix of a 3x3 (robust for singular input: pure minors)
SANM_HD inline void cof3(double* c, const double* a) {
    c[0] = a[4] * a[8] - a[5] * a[7];
    c[1] = a[5] * a[6] - a[3] * a[8];
    c[2] = a[3] * a[7] - a[4] * a[6];
    c[3] = a[2] * a[7] - a[1] * a[8];
    c[4] = a[0] * a[8] - a[2] * a[6];
    c[5] = a[1] * a[6] - a[0] * a[7];
    c[6] = a[1] * a[5] - a[2] * a[4];
    c[7] = a[2] * a[3] - a[0] * a[5];
    c[8] = a[0] * a[4] - a[1] * a[3];
}

SANM_HD inline void inv3(double* r, const double* a) {
    double c[9];
    cof3(c, a);
    double d = a[0] * c[0] + a[1] * c[1] + a[2] * c[2];
    double id = 1.0 / d;
    // inverse = cofactor^T / det
    r[0] = c[0] * id;
    r[1] = c[3] * id;
    r[2] = c[6] * id;
    r[3] = c[1] * id;
    r[4] = c[4] * id;
    r[5] = c[7] * id;
    r[6] = c[2] * id;
    r[7] = c[5] * id;
    r[8] = c[8] * id;
}

//! determinant of an n x n matrix by LU with partial pivoting (n <= MAXD)
SANM_HD inline double det_n(const double* a_in, int n) {
    if (n == 1) return a_in[0];
    if (n == 2) return a_in[0] * a_in[3] - a_in[1] * a_in[2];
    if (n == 3) return det3(a_in);
    double a[MAXE];
    for (int i = 0; i < n * n; ++i) a[i] = a_in[i];
    double det = 1;
    for (int c = 0; c < n; ++c) {
        int piv = c;
        double best = fabs(a[c * n + c]);
        for (int r = c + 1; r < n; ++r) {
            double v = fabs(a[r * n + c]);
            if (v > best) {
                best = v;
                piv = r;
            }
        }
        if (best == 0) return 0;
        if (piv != c) {
            for (int j = 0; j < n; ++j) {
                double tmp = a[c * n + j];
                a[c * n + j] = a[piv * n + j];
                a[piv * n + j] = tmp;
            }
            det = -det;
        }
        double p = a[c * n + c];
        det *= p;
        double ip = 1.0 / p;
        for (int r = c + 1; r < n; ++r) {
            double f = a[r * n + c] * ip;
            for (int j = c + 1; j < n; ++j) a[r * n + j] -= f * a[c * n + j];
        }
    }
    return det;
}

//! inverse of an n x n matrix by Gauss-Jordan with partial pivoting
SANM_HD inline void inv_n(double* r, const double* a_in, int n) {
    if (n == 3) {
        inv3(r, a_in);
        return;
    }
    if (n == 1) {
        r[0] = 1.0 / a_in[0];
        return;
    }
    if (n == 2) {
        double id = 1.0 / (a_in[0] * a_in[3] - a_in[1] * a_in[2]);
        r[0] = a_in[3] * id;
        r[1] = -a_in[1] * id;
        r[2] = -a_in[2] * id;
        r[3] = a_in[0] * id;
        return;
    }
    double a[MAXE];
    for (int i = 0; i < n * n; ++i) {
        a[i] = a_in[i];
        r[i] = 0;
    }
    for (int i = 0; i < n; ++i) r[i * n + i] = 1;
    for (int c = 0; c < n; ++c) {
        int piv = c;
        double best = fabs(a[c * n + c]);
        for (int q = c + 1; q < n; ++q) {
            double v = fabs(a[q * n + c]);
            if (v > best) {
                best = v;
                piv = q;
            }
        }
        if (piv != c) {
            for (int j = 0; j < n; ++j) {
                double tmp = a[c * n + j];
                a[c * n + j] = a[piv * n + j];
                a[piv * n + j] = tmp;
                tmp = r[c * n + j];
                r[c * n + j] = r[piv * n + j];
                r[piv * n + j] = tmp;
            }
        }
        double ip = 1.0 / a[c * n + c];
        for (int j = 0; j < n; ++j) {
            a[c * n + j] *= ip;
            r[c * n + j] *= ip;
        }
        for (int q = 0; q < n; ++q) {
            if (q == c) continue;
            double f = a[q * n + c];
            if (f == 0) continue;
            for (int j = 0; j < n; ++j) {
                a[q * n + j] -= f * a[c * n + j];
                r[q * n + j] -= f * r[c * n + j];
            }
        }
    }
}

//! cofactor matrix of an n x n matrix via minors (n <= MAXD; O(n^5), n small)
SANM_HD inline void cof_n(double* c, const double* a, int n) {
    if (n == 3) {
        cof3(c, a);
        return;
    }
    if (n == 2) {
        c[0] = a[3];
        c[1] = -a[2];
        c[2] = -a[1];
        c[3] = a[0];
        return;
    }
    double sub[MAXE];
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) {
            int q = 0;
            for (int r = 0; r < n; ++r) {
                if (r == i) continue;
                for (int s = 0; s < n; ++s) {
                    if (s == j) continue;
                    sub[q++] = a[r * n + s];
                }
            }
            double d = det_n(sub, n - 1);
            c[i * n + j] = ((i + j) & 1) ? -d : d;
        }
    }
}

//! value of pow(x, p) with the reference's special cases
//! (libsanm/tensor.cpp:405-453)
SANM_HD inline double pow_ref(double x, double p) {
    if (p == -3.) return 1.0 / (x * x * x);
    if (p == -2.) return 1.0 / (x * x);
    if (p == -1.) return 1.0 / x;
    if (p == -.5) return 1.0 / sqrt(x);
    if (p == 0.) return 1.0;
    if (p == .5) return sqrt(x);
    if (p == 1.) return x;
    if (p == 2.) return x * x;
    if (p == 3.) return x * x * x;
    return pow(x, p);
}

SANM_HD inline double clip_div(double x, double y) {
    // libsanm/tensor_svd.cpp:28-31
    return x * y / (y * y + 1e-12);
}

/* -------------------------------------------- 3x3 SVD (one-sided Jacobi) */

/*!
 * M = L diag(s) R^T with s sorted decreasing and non-negative; L, R orthogonal.
 * One-sided (Hestenes) Jacobi on the columns of M: high relative accuracy and
 * no squaring of the condition number.  Row-major 3x3 arrays.
 */
SANM_HD inline void svd3(const double* M, double* L, double* s, double* R) {
    double A[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int i = 0; i < 9; ++i) A[i] = M[i];
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0;
        for (int p = 0; p < 2; ++p) {
            for (int q = p + 1; q < 3; ++q) {
                double app = 0, aqq = 0, apq = 0;
                for (int r = 0; r < 3; ++r) {
                    app += A[r * 3 + p] * A[r * 3 + p];
                    aqq += A[r * 3 + q] * A[r * 3 + q];
                    apq += A[r * 3 + p] * A[r * 3 + q];
                }
                double lim = 1e-32 * app * aqq;  // (1e-16)^2 relative
                if (apq * apq <= lim || apq == 0) continue;
                off += fabs(apq);
                double zeta = (aqq - app) / (2 * apq);
                double tt = (zeta >= 0 ? 1.0 : -1.0) /
                            (fabs(zeta) + sqrt(1 + zeta * zeta));
                double cs = 1.0 / sqrt(1 + tt * tt), sn = cs * tt;
                for (int r = 0; r < 3; ++r) {
                    double x = A[r * 3 + p], y = A[r * 3 + q];
                    A[r * 3 + p] = cs * x - sn * y;
                    A[r * 3 + q] = sn * x + cs * y;
                    x = V[r * 3 + p];
                    y = V[r * 3 + q];
                    V[r * 3 + p] = cs * x - sn * y;
                    V[r * 3 + q] = sn * x + cs * y;
                }
            }
        }
        if (off == 0) break;
    }
    double nrm[3];
    for (int c = 0; c < 3; ++c) {
        nrm[c] = sqrt(A[c] * A[c] + A[3 + c] * A[3 + c] + A[6 + c] * A[6 + c]);
    }
    // sort columns by decreasing norm
    int idx[3] = {0, 1, 2};
    for (int i = 0; i < 2; ++i) {
        for (int j = 0; j < 2 - i; ++j) {
            if (nrm[idx[j]] < nrm[idx[j + 1]]) {
                int tmp = idx[j];
                idx[j] = idx[j + 1];
                idx[j + 1] = tmp;
            }
        }
    }
    for (int c = 0; c < 3; ++c) {
        int sc = idx[c];
        s[c] = nrm[sc];
        for (int r = 0; r < 3; ++r) R[r * 3 + c] = V[r * 3 + sc];
    }
    // left vectors: normalised columns; rank-deficient columns are completed
    // with cross products so that L stays orthogonal
    double smax = s[0] > 0 ? s[0] : 1.0;
    int good[3];
    for (int c = 0; c < 3; ++c) {
        good[c] = s[c] > 1e-300 && s[c] > 1e-14 * smax;
        if (good[c]) {
            double inv = 1.0 / s[c];
            for (int r = 0; r < 3; ++r) L[r * 3 + c] = A[r * 3 + idx[c]] * inv;
        }
    }
    if (!good[0]) {
        // zero matrix
        for (int i = 0; i < 9; ++i) L[i] = (i % 4 == 0) ? 1.0 : 0.0;
        return;
    }
    if (!good[1]) {
        // pick any unit vector orthogonal to column 0
        double x = L[0], y = L[3], z = L[6];
        double ax = fabs(x), ay = fabs(y), az = fabs(z);
        double u0, u1, u2;
        if (ax <= ay && ax <= az) {
            u0 = 0;
            u1 = -z;
            u2 = y;
        } else if (ay <= az) {
            u0 = -z;
            u1 = 0;
            u2 = x;
        } else {
            u0 = -y;
            u1 = x;
            u2 = 0;
        }
        double inv = 1.0 / sqrt(u0 * u0 + u1 * u1 + u2 * u2);
        L[1] = u0 * inv;
        L[4] = u1 * inv;
        L[7] = u2 * inv;
    }
    if (!good[2] || !good[1]) {
        // column 2 = column 0 x column 1
        L[2] = L[3] * L[7] - L[6] * L[4];
        L[5] = L[6] * L[1] - L[0] * L[7];
        L[8] = L[0] * L[4] - L[3] * L[1];
    }
}

/*!
 * SVD-W of a row-major 3x3 M: M = U diag(S) U^T W  (libsanm/tensor_svd.cpp:48-145).
 * U = left singular vectors of M, W = U Rsv^T with Rsv the right singular
 * vectors.  With require_rotation and det(L) det(R) < 0 a singular value (and
 * its U column) is negated, chosen exactly as the reference does (:93-121).
 */
SANM_HD inline void svdw3(const double* M, bool require_rotation, double* U,
                          double* S, double* W) {
    double R[9];
    svd3(M, U, S, R);
    if (require_rotation && ((det3(U) < 0) != (det3(R) < 0))) {
        const double EPS = 1e-3;
        int best_idx = -1, best_nr = 4;
        for (int i = 0; i < 3; ++i) {
            int j = i + 1;
            while (j < 3 && fabs(S[i] - S[j]) < EPS) ++j;
            int nr = j - i;
            if (nr <= best_nr || (nr == best_nr + 1 && nr % 2 == 1)) {
                best_idx = i;
                best_nr = nr;
                if (nr == 1) break;
            }
            i = j;  // note: the reference's loop also does ++i afterwards
        }
        int lo = best_idx, hi = best_idx + 1;
        if (!(best_nr == 1 || best_nr % 2 == 0)) hi = best_idx + best_nr;
        for (int c = lo; c < hi; ++c) {
            S[c] = -S[c];
            for (int r = 0; r < 3; ++r) U[r * 3 + c] = -U[r * 3 + c];
        }
    }
    // W = U R^T
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            W[r * 3 + c] = U[r * 3 + 0] * R[c * 3 + 0] +
                           U[r * 3 + 1] * R[c * 3 + 1] +
                           U[r * 3 + 2] * R[c * 3 + 2];
        }
    }
}

SANM_HD inline void mm3(double* C, const double* A, const double* B) {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            C[r * 3 + c] = A[r * 3] * B[c] + A[r * 3 + 1] * B[3 + c] +
                           A[r * 3 + 2] * B[6 + c];
        }
    }
}
SANM_HD inline void mm3_tn(double* C, const double* A, const double* B) {
    // C = A^T B
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            C[r * 3 + c] = A[r] * B[c] + A[3 + r] * B[3 + c] + A[6 + r] * B[6 + c];
        }
    }
}
SANM_HD inline void mm3_nt(double* C, const double* A, const double* B) {
    // C = A B^T
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            C[r * 3 + c] = A[r * 3] * B[c * 3] + A[r * 3 + 1] * B[c * 3 + 1] +
                           A[r * 3 + 2] * B[c * 3 + 2];
        }
    }
}

/*!
 * Order-k polar-mode forward solve, restating svd_w_taylor_fwd_p
 * (libsanm/tensor_svd.cpp:389-475) on row-major 3x3 arrays.  With
 * P(a) = U S U^T the symmetric factor and M = P W:
 *   D   = Bm - Bp           (Bm = sum M_i M_{k-i}^T, Bp = sum P_i^T P_{k-i})
 *   G   = S (U^T W) Mk^T U;   E = U^T D^T U + G + G^T
 *   x_ij = clip_div(E_ij, S_i + S_j);   Pk^T = U x U^T
 *   Wk^T = (Mk^T - Bpw^T - W^T Pk^T) U S^-1 U^T
 */
SANM_HD inline void svdw3_fwd_p(double* Pk, double* Wk, const double* Mk,
                                const double* U, const double* S,
                                const double* W, const double* D,
                                const double* Bpw) {
    double t0[9], t1[9], E[9], G[9];
    // E = U^T D^T U
    mm3_tn(t0, U, D);  // hmm: U^T D ; we need U^T D^T U = (U^T D U)^T
    mm3(t1, t0, U);    // t1 = U^T D U
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) E[r * 3 + c] = t1[c * 3 + r];
    // G = diag(S) (U^T W) Mk^T U
    mm3_tn(t0, U, W);   // U^T W
    mm3_nt(t1, t0, Mk); // (U^T W) Mk^T
    mm3(G, t1, U);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) G[r * 3 + c] *= S[r];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) t0[r * 3 + c] = E[r * 3 + c] + G[r * 3 + c] + G[c * 3 + r];
    double x[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) x[r * 3 + c] = clip_div(t0[r * 3 + c], S[r] + S[c]);
    // PkT = U x U^T  ->  Pk = U x^T U^T
    double PkT[9];
    mm3(t0, U, x);
    mm3_nt(PkT, t0, U);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Pk[r * 3 + c] = PkT[c * 3 + r];
    // Q = Mk^T - Bpw^T - W^T PkT
    mm3_tn(t0, W, PkT);
    double Q[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            Q[r * 3 + c] = Mk[c * 3 + r] - Bpw[c * 3 + r] - t0[r * 3 + c];
    // WkT = Q U S^-1 U^T
    mm3(t0, Q, U);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        double si = clip_div(1.0, S[c]);
#pragma unroll
        for (int r = 0; r < 3; ++r) t0[r * 3 + c] *= si;
    }
    double WkT[9];
    mm3_nt(WkT, t0, U);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Wk[r * 3 + c] = WkT[c * 3 + r];
}

/* ---------------- general n (2 .. MAXD): runtime-size versions of the above.
 * The reference runs SVD-W on any square size (libsanm/tensor_svd.cpp:48-145,
 * Eigen::Dynamic path; tests/symbolic.cpp:658-716 use 4x4); `fea` only ever
 * uses 3x3, which keeps the unrolled code above.  Row-major n x n arrays. */
constexpr int MAXM = MAXD * MAXD;  //!< elements of the largest matrix

//! C = op(A) op(B), n x n, ta / tb = transpose flags
SANM_HD inline void mmn(double* C, const double* A, const double* B, int n,
                        bool ta, bool tb) {
    for (int r = 0; r < n; ++r) {
        for (int c = 0; c < n; ++c) {
            double s = 0;
            for (int q = 0; q < n; ++q) {
                s += (ta ? A[q * n + r] : A[r * n + q]) *
                     (tb ? B[c * n + q] : B[q * n + c]);
            }
            C[r * n + c] = s;
        }
    }
}

//! sign of det(A) by elimination with partial pivoting (A orthogonal here)
SANM_HD inline double det_sign_n(const double* A, int n) {
    double T[MAXM];
    for (int i = 0; i < n * n; ++i) T[i] = A[i];
    double sg = 1;
    for (int k = 0; k < n; ++k) {
        int piv = k;
        for (int i = k + 1; i < n; ++i) {
            if (fabs(T[i * n + k]) > fabs(T[piv * n + k])) piv = i;
        }
        if (T[piv * n + k] == 0) return 0;
        if (piv != k) {
            sg = -sg;
            for (int c = 0; c < n; ++c) {
                double x = T[k * n + c];
                T[k * n + c] = T[piv * n + c];
                T[piv * n + c] = x;
            }
        }
        if (T[k * n + k] < 0) sg = -sg;
        for (int i = k + 1; i < n; ++i) {
            double f = T[i * n + k] / T[k * n + k];
            for (int c = k + 1; c < n; ++c) T[i * n + c] -= f * T[k * n + c];
        }
    }
    return sg;
}

//! svd3 for n x n: one-sided Jacobi, s sorted decreasing
SANM_HD inline void svdn(const double* M, int n, double* L, double* s, double* R) {
    double A[MAXM], V[MAXM], nrm[MAXD];
    int idx[MAXD], good[MAXD];
    for (int i = 0; i < n * n; ++i) {
        A[i] = M[i];
        V[i] = (i % (n + 1) == 0) ? 1.0 : 0.0;
    }
    for (int sweep = 0; sweep < 40; ++sweep) {
        double off = 0;
        for (int p = 0; p < n - 1; ++p) {
            for (int q = p + 1; q < n; ++q) {
                double app = 0, aqq = 0, apq = 0;
                for (int r = 0; r < n; ++r) {
                    app += A[r * n + p] * A[r * n + p];
                    aqq += A[r * n + q] * A[r * n + q];
                    apq += A[r * n + p] * A[r * n + q];
                }
                double lim = 1e-32 * app * aqq;
                if (apq * apq <= lim || apq == 0) continue;
                off += fabs(apq);
                double zeta = (aqq - app) / (2 * apq);
                double tt = (zeta >= 0 ? 1.0 : -1.0) /
                            (fabs(zeta) + sqrt(1 + zeta * zeta));
                double cs = 1.0 / sqrt(1 + tt * tt), sn = cs * tt;
                for (int r = 0; r < n; ++r) {
                    double x = A[r * n + p], y = A[r * n + q];
                    A[r * n + p] = cs * x - sn * y;
                    A[r * n + q] = sn * x + cs * y;
                    x = V[r * n + p];
                    y = V[r * n + q];
                    V[r * n + p] = cs * x - sn * y;
                    V[r * n + q] = sn * x + cs * y;
                }
            }
        }
        if (off == 0) break;
    }
    for (int c = 0; c < n; ++c) {
        double q = 0;
        for (int r = 0; r < n; ++r) q += A[r * n + c] * A[r * n + c];
        nrm[c] = sqrt(q);
        idx[c] = c;
    }
    for (int i = 0; i < n - 1; ++i) {
        for (int j = 0; j < n - 1 - i; ++j) {
            if (nrm[idx[j]] < nrm[idx[j + 1]]) {
                int tmp = idx[j];
                idx[j] = idx[j + 1];
                idx[j + 1] = tmp;
            }
        }
    }
    for (int c = 0; c < n; ++c) {
        s[c] = nrm[idx[c]];
        for (int r = 0; r < n; ++r) R[r * n + c] = V[r * n + idx[c]];
    }
    const double smax = s[0] > 0 ? s[0] : 1.0;
    for (int c = 0; c < n; ++c) {
        good[c] = s[c] > 1e-300 && s[c] > 1e-14 * smax;
        if (good[c]) {
            double inv = 1.0 / s[c];
            for (int r = 0; r < n; ++r) L[r * n + c] = A[r * n + idx[c]] * inv;
        }
    }
    // rank-deficient columns: complete L to an orthogonal matrix (the unit
    // vector with the largest remainder after projecting out the columns
    // already there; the singular values are sorted, so those come first)
    for (int c = 0; c < n; ++c) {
        if (good[c]) continue;
        double best = -1, v[MAXD], bv[MAXD];
        for (int j = 0; j < n; ++j) {
            for (int r = 0; r < n; ++r) v[r] = r == j ? 1.0 : 0.0;
            for (int pass = 0; pass < 2; ++pass) {
                for (int q = 0; q < c; ++q) {
                    double d = 0;
                    for (int r = 0; r < n; ++r) d += L[r * n + q] * v[r];
                    for (int r = 0; r < n; ++r) v[r] -= d * L[r * n + q];
                }
            }
            double q2 = 0;
            for (int r = 0; r < n; ++r) q2 += v[r] * v[r];
            if (q2 > best) {
                best = q2;
                for (int r = 0; r < n; ++r) bv[r] = v[r];
            }
        }
        double inv = 1.0 / sqrt(best);
        for (int r = 0; r < n; ++r) L[r * n + c] = bv[r] * inv;
    }
}

//! svdw3 for n x n (libsanm/tensor_svd.cpp:48-145)
SANM_HD inline void svdwn(const double* M, int n, bool require_rotation,
                          double* U, double* S, double* W) {
    double R[MAXM];
    svdn(M, n, U, S, R);
    if (require_rotation && ((det_sign_n(U, n) < 0) != (det_sign_n(R, n) < 0))) {
        const double EPS = 1e-3;
        int best_idx = -1, best_nr = n + 1;
        for (int i = 0; i < n; ++i) {
            int j = i + 1;
            while (j < n && fabs(S[i] - S[j]) < EPS) ++j;
            int nr = j - i;
            if (nr <= best_nr || (nr == best_nr + 1 && nr % 2 == 1)) {
                best_idx = i;
                best_nr = nr;
                if (nr == 1) break;
            }
            i = j;  // the reference's loop also does ++i afterwards
        }
        int lo = best_idx, hi = best_idx + 1;
        if (!(best_nr == 1 || best_nr % 2 == 0)) hi = best_idx + best_nr;
        for (int c = lo; c < hi; ++c) {
            S[c] = -S[c];
            for (int r = 0; r < n; ++r) U[r * n + c] = -U[r * n + c];
        }
    }
    mmn(W, U, R, n, false, true);  // W = U R^T
}

//! svdw3_fwd_p for n x n (libsanm/tensor_svd.cpp:389-475), same formulas
SANM_HD inline void svdwn_fwd_p(double* Pk, double* Wk, const double* Mk,
                                const double* U, const double* S,
                                const double* W, const double* D,
                                const double* Bpw, int n) {
    double t0[MAXM], t1[MAXM], G[MAXM], x[MAXM];
    mmn(t0, U, D, n, true, false);   // U^T D
    mmn(t1, t0, U, n, false, false); // U^T D U; E = its transpose
    mmn(t0, U, W, n, true, false);   // U^T W
    mmn(x, t0, Mk, n, false, true);  // (U^T W) Mk^T
    mmn(G, x, U, n, false, false);
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c) G[r * n + c] *= S[r];
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c)
            x[r * n + c] = clip_div(t1[c * n + r] + G[r * n + c] + G[c * n + r],
                                    S[r] + S[c]);
    // PkT = U x U^T
    mmn(t0, U, x, n, false, false);
    mmn(t1, t0, U, n, false, true);  // t1 = PkT
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c) Pk[r * n + c] = t1[c * n + r];
    // Q = Mk^T - Bpw^T - W^T PkT
    mmn(t0, W, t1, n, true, false);
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c)
            G[r * n + c] = Mk[c * n + r] - Bpw[c * n + r] - t0[r * n + c];
    // WkT = Q U S^-1 U^T
    mmn(t0, G, U, n, false, false);
    for (int c = 0; c < n; ++c) {
        const double si = clip_div(1.0, S[c]);
        for (int r = 0; r < n; ++r) t0[r * n + c] *= si;
    }
    mmn(t1, t0, U, n, false, true);
    for (int r = 0; r < n; ++r)
        for (int c = 0; c < n; ++c) Wk[r * n + c] = t1[c * n + r];
}

/*!
 * Order-k FULL-mode forward solve (U_k, S_k, W_k all wanted), restating
 * svd_w_taylor_fwd (libsanm/tensor_svd.cpp:275-387) on row-major 3x3 arrays.
 * M = U S U^T W;  Mb = the part of [U S U^T W]_k without order-k unknowns,
 * Bu = sum U_i^T U_{k-i},  Bw = sum W_i^T W_{k-i}  (i = 1 .. k-1).  With
 * V0 = W0^T U0:
 *   E  = V0^T (Mk - Mb)^T U0
 *   R  = E^T - E - V0^T Bw V0 diag(S0);   X_ij = clip_div(R_ij, S0_i + S0_j)
 *   Wk = U0 X V0^T
 *   E' = E - X^T diag(S0) + Bu^T diag(S0);   Sk_i = E'_ii
 *   Q_ij = clip_div(E'_ij, S0_i - S0_j), Q_ji = -Bu_ij - Q_ij (i < j),
 *   Q_jj = -Bu_jj / 2;   Uk = U0 Q^T
 */
SANM_HD inline void svdw3_fwd_full(double* Uk, double* Sk, double* Wk,
                                   const double* Mk, const double* Mb,
                                   const double* U0, const double* S0,
                                   const double* W0, const double* Bu,
                                   const double* Bw) {
    double V0[9], D[9], t0[9], t1[9], E[9], X[9];
    mm3_tn(V0, W0, U0);  // W0^T U0
#pragma unroll
    for (int q = 0; q < 9; ++q) D[q] = Mk[q] - Mb[q];
    // E = V0^T D^T U0, with V0^T D^T = (D V0)^T
    mm3(t1, D, V0);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) t0[r * 3 + c] = t1[c * 3 + r];
    mm3(E, t0, U0);
    // R = E^T - E - V0^T Bw V0 diag(S0)
    mm3_tn(t0, V0, Bw);
    mm3(t1, t0, V0);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double rv = E[c * 3 + r] - E[r * 3 + c] - t1[r * 3 + c] * S0[c];
            X[r * 3 + c] = clip_div(rv, S0[r] + S0[c]);
        }
    // Wk = U0 X V0^T
    mm3(t0, U0, X);
    mm3_nt(Wk, t0, V0);
    // E' = E - X^T diag(S0) + Bu^T diag(S0)
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            E[r * 3 + c] += (Bu[c * 3 + r] - X[c * 3 + r]) * S0[c];
        }
    double Q[9];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        Sk[j] = E[j * 3 + j];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (i < j) {
                const double v = clip_div(E[i * 3 + j], S0[i] - S0[j]);
                Q[i * 3 + j] = v;
                Q[j * 3 + i] = -Bu[i * 3 + j] - v;
            }
        }
        Q[j * 3 + j] = -Bu[j * 3 + j] / 2;
    }
    mm3_nt(Uk, U0, Q);  // U0 Q^T
}

/* -------------------------------------------------------- op evaluation */

enum Phase : int {
    PH_EVAL = 0,  //!< order 0
    PH_BIAS = 1,  //!< phase A: compute + cache self-bias, propagate bias
    PH_COEF = 2,  //!< phase B: propagate coefficient with cached self-bias
};

struct ItemCtx {
    const DProgram* prog;
    size_t t;
    int k;         //!< order
    int phase;
    int sb_zero;   //!< PH_COEF: treat cached self-bias (and hidden partials) as 0
    int64_t np;
    SANM_HD const DVar& var(int v) const { return prog->vars[v]; }
    SANM_HD double ld(int v, int kk, int c) const {
        return vbase(var(v), kk)[(int64_t)c * np + t];
    }
    SANM_HD void st(int v, int kk, int c, double x) const {
        vbase(var(v), kk)[(int64_t)c * np + t] = x;
    }
};

SANM_HD inline void op_lincomb(const DOp& op, const ItemCtx& c) {
    const DVar& o = c.var(op.out[0]);
    for (int e = 0; e < o.e; ++e) {
        double s = c.phase == PH_EVAL ? op.f[op.nin] : 0.0;
        for (int j = 0; j < op.nin; ++j) {
            const DVar& iv = c.var(op.in[j]);
            if (c.phase != PH_EVAL && iv.is_const) continue;
            s += c.ld(op.in[j], c.k, iv.e == 1 ? 0 : e) * op.f[j];
        }
        c.st(op.out[0], c.k, e, s);
    }
}

SANM_HD inline void op_multiply(const DOp& op, const ItemCtx& c) {
    const int a = op.in[0], b = op.in[1], o = op.out[0];
    const DVar &av = c.var(a), &bv = c.var(b), &ov = c.var(o);
    const bool conv = !av.is_const && !bv.is_const;
    for (int e = 0; e < ov.e; ++e) {
        const int ea = av.e == 1 ? 0 : e, eb = bv.e == 1 ? 0 : e;
        if (c.phase == PH_EVAL) {
            c.st(o, 0, e, c.ld(a, 0, ea) * c.ld(b, 0, eb));
            continue;
        }
        double sb = 0;
        if (conv) {
            if (c.phase == PH_BIAS) {
                for (int i = 1; i < c.k; ++i) {
                    sb += c.ld(a, i, ea) * c.ld(b, c.k - i, eb);
                }
                op.sb[(int64_t)e * c.np + c.t] = sb;
            } else if (!c.sb_zero) {
                sb = op.sb[(int64_t)e * c.np + c.t];
            }
        }
        double r = sb;
        if (!bv.is_const) r += c.ld(a, 0, ea) * c.ld(b, c.k, eb);
        if (!av.is_const) r += c.ld(a, c.k, ea) * c.ld(b, 0, eb);
        c.st(o, c.k, e, r);
    }
}

//! coefficient k of (sum_{i<k} x_i a^i)^p for integer p >= 2: the self-bias of
//! pow when x_0 ~ 0 (libsanm/analytic_unary.cpp:46-92)
SANM_HD inline double pow_int_selfbias(const double* x, int k, int p) {
    // x[0..k-1] valid, x[k] := 0;  result = [a^k] x(a)^p by repeated squaring
    double base[MAX_ORDER + 1], prod[MAX_ORDER + 1], tmp[MAX_ORDER + 1];
    for (int i = 0; i <= k; ++i) {
        base[i] = i < k ? x[i] : 0.0;
        prod[i] = i == 0 ? 1.0 : 0.0;
    }
    bool prod_is_one = true;
    while (p > 0) {
        if (p & 1) {
            if (prod_is_one) {
                for (int i = 0; i <= k; ++i) prod[i] = base[i];
                prod_is_one = false;
            } else {
                for (int i = 0; i <= k; ++i) {
                    double s = 0;
                    for (int j = 0; j <= i; ++j) s += prod[j] * base[i - j];
                    tmp[i] = s;
                }
                for (int i = 0; i <= k; ++i) prod[i] = tmp[i];
            }
        }
        p >>= 1;
        if (p) {
            for (int i = 0; i <= k; ++i) {
                double s = 0;
                for (int j = 0; j <= i; ++j) s += base[j] * base[i - j];
                tmp[i] = s;
            }
            for (int i = 0; i <= k; ++i) base[i] = tmp[i];
        }
    }
    return prod[k];
}

SANM_HD inline void op_unary(const DOp& op, const ItemCtx& c) {
    const int x = op.in[0], o = op.out[0];
    const DVar& ov = c.var(o);
    const bool is_log = op.type == OP_LOG;
    const double p = op.f[0];
    for (int e = 0; e < ov.e; ++e) {
        const int64_t off = (int64_t)e * c.np + c.t;
        if (c.phase == PH_EVAL) {
            double x0 = c.ld(x, 0, e);
            double f0, d;
            if (is_log) {
                f0 = log(x0);
                d = 1.0 / x0;
            } else {
                f0 = pow_ref(x0, p);
                d = pow_ref(x0, p - 1) * p;
                if (fabs(x0) < 1e-3) {
                    // zero base: only integer exponents have a Taylor series
                    if (p <= 0.5 || floor(p) != p) {
                        rt_flag_or(c.prog->err, 1);
                    }
                    rt_flag_or(op.flag, 1);
                }
            }
            c.st(o, 0, e, f0);
            op.aux0[off] = d;
            continue;
        }
        double sb = 0;
        if (c.phase == PH_BIAS) {
            const int k = c.k;
            if (is_log) {
                for (int i = 1; i < k; ++i) {
                    sb += c.ld(x, k - i, e) * c.ld(o, i, e) * (-double(i) / double(k));
                }
                sb /= c.ld(x, 0, e);
            } else if (*op.flag) {
                double xs[MAX_ORDER + 1];
                for (int i = 0; i < k; ++i) xs[i] = c.ld(x, i, e);
                sb = pow_int_selfbias(xs, k, (int)p);
            } else {
                for (int i = 1; i < k; ++i) {
                    sb += c.ld(o, k - i, e) * c.ld(x, i, e) *
                          (double(i) / double(k) * (p + 1) - 1);
                }
                sb /= c.ld(x, 0, e);
            }
            op.sb[off] = sb;
        } else if (!c.sb_zero) {
            sb = op.sb[off];
        }
        c.st(o, c.k, e, c.ld(x, c.k, e) * op.aux0[off] + sb);
    }
}

SANM_HD inline void op_reduce_sum(const DOp& op, const ItemCtx& c) {
    // i[0] = axis (-1: everything per item; 1 or 2: per-item axis)
    const int x = op.in[0], o = op.out[0];
    const DVar& xv = c.var(x);
    const int axis = op.i[0];
    if (axis == -1) {
        double s = 0;
        for (int e = 0; e < xv.e; ++e) s += c.ld(x, c.k, e);
        c.st(o, c.k, 0, s);
    } else if (axis == 1) {
        for (int j = 0; j < xv.d2; ++j) {
            double s = 0;
            for (int i = 0; i < xv.d1; ++i) s += c.ld(x, c.k, i * xv.d2 + j);
            c.st(o, c.k, j, s);
        }
    } else {
        for (int i = 0; i < xv.d1; ++i) {
            double s = 0;
            for (int j = 0; j < xv.d2; ++j) s += c.ld(x, c.k, i * xv.d2 + j);
            c.st(o, c.k, i, s);
        }
    }
}

SANM_HD inline void op_transpose(const DOp& op, const ItemCtx& c) {
    const int x = op.in[0], o = op.out[0];
    const DVar& xv = c.var(x);
    for (int i = 0; i < xv.d1; ++i) {
        for (int j = 0; j < xv.d2; ++j) {
            c.st(o, c.k, j * xv.d1 + i, c.ld(x, c.k, i * xv.d2 + j));
        }
    }
}

SANM_HD inline void op_muleye(const DOp& op, const ItemCtx& c) {
    const int x = op.in[0], o = op.out[0];
    const int d = op.i[0];
    double v = c.ld(x, c.k, 0);
    for (int i = 0; i < d; ++i) {
        for (int j = 0; j < d; ++j) c.st(o, c.k, i * d + j, i == j ? v : 0.0);
    }
}

SANM_HD inline void op_slice(const DOp& op, const ItemCtx& c) {
    // i[0] = per-item axis (1 or 2), i[1] = begin, i[2] = stride
    const int x = op.in[0], o = op.out[0];
    const DVar &xv = c.var(x), &ov = c.var(o);
    const int axis = op.i[0], begin = op.i[1], stride = op.i[2];
    for (int i = 0; i < ov.d1; ++i) {
        for (int j = 0; j < ov.d2; ++j) {
            int si = axis == 1 ? begin + i * stride : i;
            int sj = axis == 2 ? begin + j * stride : j;
            c.st(o, c.k, i * ov.d2 + j, c.ld(x, c.k, si * xv.d2 + sj));
        }
    }
}

SANM_HD inline void op_concat(const DOp& op, const ItemCtx& c) {
    // i[0] = per-item axis (1 or 2)
    const int o = op.out[0];
    const DVar& ov = c.var(o);
    const int axis = op.i[0];
    int off = 0;
    for (int q = 0; q < op.nin; ++q) {
        const DVar& xv = c.var(op.in[q]);
        const bool skip = c.phase != PH_EVAL && xv.is_const;
        for (int i = 0; i < xv.d1; ++i) {
            for (int j = 0; j < xv.d2; ++j) {
                int oi = axis == 1 ? off + i : i, oj = axis == 2 ? off + j : j;
                c.st(o, c.k, oi * ov.d2 + oj,
                     skip ? 0.0 : c.ld(op.in[q], c.k, i * xv.d2 + j));
            }
        }
        off += axis == 1 ? xv.d1 : xv.d2;
    }
}

template <int M, int K, int N>
SANM_HD inline void op_matmul(const DOp& op, const ItemCtx& c) {
    const int a = op.in[0], b = op.in[1], o = op.out[0];
    const DVar &av = c.var(a), &bv = c.var(b);
    const int m = M ? M : av.d1, kd = K ? K : av.d2, n = N ? N : bv.d2;
    constexpr int EA = M ? M * K : MAXE, EB = K ? K * N : MAXE,
                  EO = M ? M * N : MAXE;
    const int ea = m * kd, eb = kd * n, eo = m * n;
    double A[EA], B[EB], C[EO];
    if (c.phase == PH_EVAL) {
        ld_mat<M * K>(A, av, 0, c.np, c.t, ea);
        ld_mat<K * N>(B, bv, 0, c.np, c.t, eb);
        zero_arr<M * N>(C, eo);
        mm_acc<M, K, N>(C, A, B, m, kd, n);
        st_buf<M * N>(c.var(o).p0, C, c.np, c.t, eo);
        return;
    }
    const bool conv = !av.is_const && !bv.is_const;
    zero_arr<M * N>(C, eo);
    if (conv) {
        if (c.phase == PH_BIAS) {
            for (int i = 1; i < c.k; ++i) {
                ld_mat<M * K>(A, av, i, c.np, c.t, ea);
                ld_mat<K * N>(B, bv, c.k - i, c.np, c.t, eb);
                mm_acc<M, K, N>(C, A, B, m, kd, n);
            }
            st_buf<M * N>(op.sb, C, c.np, c.t, eo);
        } else if (!c.sb_zero) {
            ld_buf<M * N>(C, op.sb, c.np, c.t, eo);
        }
    }
    if (!av.is_const) {
        ld_mat<M * K>(A, av, c.k, c.np, c.t, ea);
        ld_mat<K * N>(B, bv, 0, c.np, c.t, eb);
        mm_acc<M, K, N>(C, A, B, m, kd, n);
    }
    if (!bv.is_const) {
        ld_mat<M * K>(A, av, 0, c.np, c.t, ea);
        ld_mat<K * N>(B, bv, c.k, c.np, c.t, eb);
        mm_acc<M, K, N>(C, A, B, m, kd, n);
    }
    st_buf<M * N>(vbase(c.var(o), c.k), C, c.np, c.t, eo);
}

template <int M>
SANM_HD inline void op_matinvmul(const DOp& op, const ItemCtx& c) {
    // i[0] = is_left, i[1] = use_identity
    const int x = op.in[0], o = op.out[0];
    const bool is_left = op.i[0], ident = op.i[1];
    const int a = ident ? -1 : op.in[1];
    const DVar &xv = c.var(x), &ov = c.var(o);
    const int m = M ? M : xv.d1;
    constexpr int E = M ? M * M : MAXE;
    const int e = m * m;
    double X[E], Y[E], T[E], R[E];
    if (c.phase == PH_EVAL) {
        ld_mat<M * M>(X, xv, 0, c.np, c.t, e);
        inv_n(R, X, m);
        if (ident) {
            st_buf<M * M>(ov.p0, R, c.np, c.t, e);
        } else {
            st_buf<M * M>(op.aux0, R, c.np, c.t, e);
            ld_mat<M * M>(Y, c.var(a), 0, c.np, c.t, e);
            zero_arr<M * M>(T, e);
            if (is_left) {
                mm_acc<M, M, M>(T, Y, R, m, m, m);
            } else {
                mm_acc<M, M, M>(T, R, Y, m, m, m);
            }
            st_buf<M * M>(ov.p0, T, c.np, c.t, e);
        }
        return;
    }
    // self bias: -sum_{i=1}^{k-1} Y_i X_{k-i}  (left)  /  X_i Y_{k-i} (right)
    zero_arr<M * M>(T, e);
    if (!xv.is_const) {
        if (c.phase == PH_BIAS) {
            for (int i = 1; i < c.k; ++i) {
                if (is_left) {
                    ld_mat<M * M>(Y, ov, i, c.np, c.t, e);
                    ld_mat<M * M>(X, xv, c.k - i, c.np, c.t, e);
                    mm_acc<M, M, M>(T, Y, X, m, m, m);
                } else {
                    ld_mat<M * M>(X, xv, i, c.np, c.t, e);
                    ld_mat<M * M>(Y, ov, c.k - i, c.np, c.t, e);
                    mm_acc<M, M, M>(T, X, Y, m, m, m);
                }
            }
#pragma unroll
            for (int q = 0; q < e; ++q) T[q] = -T[q];
            st_buf<M * M>(op.sb, T, c.np, c.t, e);
        } else if (!c.sb_zero) {
            ld_buf<M * M>(T, op.sb, c.np, c.t, e);
        }
    }
    // tmp0 = a_k + self_bias
    if (!ident && !c.var(a).is_const) {
        ld_mat<M * M>(Y, c.var(a), c.k, c.np, c.t, e);
#pragma unroll
        for (int q = 0; q < e; ++q) T[q] += Y[q];
    }
    // tmp0 - y_0 x_k   (left)   /   tmp0 - x_k y_0   (right)
    if (!xv.is_const) {
        ld_mat<M * M>(Y, ov, 0, c.np, c.t, e);
        ld_mat<M * M>(X, xv, c.k, c.np, c.t, e);
#pragma unroll
        for (int q = 0; q < e; ++q) {
            if (is_left) Y[q] = -Y[q]; else X[q] = -X[q];
        }
        if (is_left) {
            mm_acc<M, M, M>(T, Y, X, m, m, m);
        } else {
            mm_acc<M, M, M>(T, X, Y, m, m, m);
        }
    }
    // times xinv
    if (ident) {
        ld_mat<M * M>(R, ov, 0, c.np, c.t, e);
    } else {
        ld_buf<M * M>(R, op.aux0, c.np, c.t, e);
    }
    zero_arr<M * M>(X, e);
    if (is_left) {
        mm_acc<M, M, M>(X, T, R, m, m, m);
    } else {
        mm_acc<M, M, M>(X, R, T, m, m, m);
    }
    st_buf<M * M>(vbase(ov, c.k), X, c.np, c.t, e);
}

SANM_HD inline void cross3(double* r, const double* a, const double* b) {
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}

/*!
 * Determinant.  out_k = <cof(X_0), X_k> + self_bias_k where self_bias_k is the
 * a^k coefficient of det(sum_{i<k} X_i a^i) (libsanm/oprs/linalg.cpp:250-273).
 * 3x3: det = r0 . (r1 x r2).  The hidden stack c_j = sum_{i<=j} r1_i x r2_{j-i}
 * (3 doubles per order) makes the self-bias an O(k) sum:
 *   self_bias_k = sum_{i=1}^{k-1} r0_i . c_{k-i}  +  r0_0 . cp_k,
 *   cp_k = sum_{i=1}^{k-1} r1_i x r2_{k-i},   c_k = cp_k + r1_0 x r2_k + r1_k x r2_0.
 * 2x2: direct convolution.  m >= 4 (the reference expands the polynomial
 * determinant, libsanm/tensor_polymat.cpp:30-136,138-341, O(k^2) per order):
 * here the series of det X follows from (det X)' = det X tr(X^-1 X'):
 *     G = X^-1 :   G_k = -G_0 sum_{i=1}^{k} X_i G_{k-i}
 *     T = tr(G X'): T_n = sum_{p+q=n} (q+1) tr(G_p X_{q+1})
 *     k D_k = sum_{i=0}^{k-1} D_i T_{k-1-i}
 * The only term of D_k with the unknown X_k is D_0 k tr(G_0 X_k) / k =
 * <cof X_0, X_k>, so the self-bias is everything else.  Hidden stack per
 * order j: [G_j (m^2), T_j, D_j]; O(k m^3) per order.  Needs det X_0 != 0
 * (as every use of the determinant in `fea` and in the reference's tests).
 */
SANM_HD inline void op_det(const DOp& op, const ItemCtx& c) {
    const int x = op.in[0], o = op.out[0];
    const DVar& xv = c.var(x);
    const int m = xv.d1;
    const int64_t np = c.np;
    const size_t t = c.t;
    if (c.phase == PH_EVAL) {
        double X[MAXE], C[MAXE];
        ld_mat<0>(X, xv, 0, np, t, xv.e);
        c.st(o, 0, 0, det_n(X, m));
        cof_n(C, X, m);
        st_buf<0>(op.aux0, C, np, t, xv.e);
        if (m == 3) {
            double cr[3];
            cross3(cr, X + 3, X + 6);
            st_buf<3>(op.auxk, cr, np, t, 3);  // slot 0
        } else if (m >= 4 && op.auxk) {
            double G[MAXE];
            inv_n(G, X, m);
            st_buf<0>(op.auxk, G, np, t, xv.e);
            op.auxk[(int64_t)(xv.e + 1) * np + t] = det_n(X, m);  // D_0
        }
        return;
    }
    if (xv.is_const) {
        c.st(o, c.k, 0, 0.0);
        return;
    }
    const int k = c.k;
    double sb = 0;
    if (m == 3) {
        double* ck = op.auxk + (int64_t)k * op.auxk_stride;
        double cp[3] = {0, 0, 0};
        if (c.phase == PH_BIAS) {
            double r0[3], r1[3], r2[3], cj[3], tmp[3];
            for (int i = 1; i < k; ++i) {
                const double* pi = vbase(xv, i) + t;
                const double* pj = vbase(xv, k - i) + t;
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    r0[q] = pi[(int64_t)q * np];
                    r1[q] = pi[(int64_t)(3 + q) * np];
                    r2[q] = pj[(int64_t)(6 + q) * np];
                }
                ld_buf<3>(cj, op.auxk + (int64_t)(k - i) * op.auxk_stride, np, t, 3);
                sb += r0[0] * cj[0] + r0[1] * cj[1] + r0[2] * cj[2];
                cross3(tmp, r1, r2);
                cp[0] += tmp[0];
                cp[1] += tmp[1];
                cp[2] += tmp[2];
            }
            ld_mat<3>(r0, xv, 0, np, t, 3);  // row 0 of X_0
            sb += r0[0] * cp[0] + r0[1] * cp[1] + r0[2] * cp[2];
            op.sb[t] = sb;
            st_buf<3>(ck, cp, np, t, 3);
        } else {
            if (!c.sb_zero) {
                sb = op.sb[t];
                ld_buf<3>(cp, ck, np, t, 3);
            }
            // finish c_k with the now-known X_k
            double X0[9], Xk[9], t1[3], t2[3];
            ld_mat<9>(X0, xv, 0, np, t, 9);
            ld_mat<9>(Xk, xv, k, np, t, 9);
            cross3(t1, X0 + 3, Xk + 6);
            cross3(t2, Xk + 3, X0 + 6);
            cp[0] += t1[0] + t2[0];
            cp[1] += t1[1] + t2[1];
            cp[2] += t1[2] + t2[2];
            st_buf<3>(ck, cp, np, t, 3);
        }
    } else if (m == 2) {
        if (c.phase == PH_BIAS) {
            for (int i = 1; i < k; ++i) {
                sb += c.ld(x, i, 0) * c.ld(x, k - i, 3) -
                      c.ld(x, i, 1) * c.ld(x, k - i, 2);
            }
            op.sb[t] = sb;
        } else if (!c.sb_zero) {
            sb = op.sb[t];
        }
    } else {
        // m >= 4: hidden slot j = [G_j (E), T_j, D_j]
        const int E = xv.e;
        auto slot = [&](int j) { return op.auxk + (int64_t)j * op.auxk_stride + t; };
        auto tr_gx = [&](int p, int q) {  // tr(G_p X_q)
            const double* g = slot(p);
            const double* xq = vbase(xv, q) + t;
            double s = 0;
            for (int r = 0; r < m; ++r) {
                for (int cc = 0; cc < m; ++cc) {
                    s += g[(int64_t)(r * m + cc) * np] * xq[(int64_t)(cc * m + r) * np];
                }
            }
            return s;
        };
        if (c.phase == PH_BIAS) {
            double s1 = 0, s2 = 0;
            for (int i = 1; i < k; ++i) {
                s1 += slot(i)[(int64_t)(E + 1) * np] * slot(k - 1 - i)[(int64_t)E * np];
            }
            for (int p = 1; p < k; ++p) s2 += (double)(k - p) * tr_gx(p, k - p);
            const double d0 = slot(0)[(int64_t)(E + 1) * np];
            sb = (s1 + d0 * s2) / (double)k;
            op.sb[t] = sb;
            slot(k - 1)[(int64_t)E * np] = s2;  // known part of T_{k-1}
        } else {
            double tpart = 0;
            if (!c.sb_zero) {
                sb = op.sb[t];
                tpart = slot(k - 1)[(int64_t)E * np];
            }
            // T_{k-1} complete, G_k = -G_0 sum_{i=1}^{k} X_i G_{k-i}
            slot(k - 1)[(int64_t)E * np] = tpart + (double)k * tr_gx(0, k);
            double acc[MAXE], A[MAXE], B[MAXE];
            for (int e = 0; e < E; ++e) acc[e] = 0;
            for (int i = 1; i <= k; ++i) {
                ld_mat<0>(A, xv, i, np, t, E);
                ld_buf<0>(B, op.auxk + (int64_t)(k - i) * op.auxk_stride, np, t, E);
                mm_acc<0, 0, 0>(acc, A, B, m, m, m);
            }
            ld_buf<0>(A, op.auxk, np, t, E);  // G_0
            for (int e = 0; e < E; ++e) B[e] = 0;
            mm_acc<0, 0, 0>(B, A, acc, m, m, m);
            for (int e = 0; e < E; ++e) B[e] = -B[e];
            st_buf<0>(op.auxk + (int64_t)k * op.auxk_stride, B, np, t, E);
        }
    }
    double r = sb;
    for (int e = 0; e < xv.e; ++e) {
        r += op.aux0[(int64_t)e * np + t] * c.ld(x, k, e);
    }
    c.st(o, k, 0, r);
    if (m >= 4 && c.phase != PH_BIAS) {
        op.auxk[(int64_t)k * op.auxk_stride + (int64_t)(xv.e + 1) * np + t] = r;  // D_k
    }
}

//! orders >= 1 of the full mode of SVD-W, see op_svdw
SANM_HD inline void op_svdw_full(const DOp& op, const ItemCtx& c) {
    const DVar &xv = c.var(op.in[0]), &uv = c.var(op.out[0]), &sv = c.var(op.out[1]),
               &wv = c.var(op.out[2]);
    const int64_t np = c.np;
    const size_t t = c.t;
    const int k = c.k;
    auto p0_slot = [&](int j) { return op.auxk + (int64_t)j * op.auxk_stride; };
    auto p1_slot = [&](int j) { return op.auxk + (int64_t)j * op.auxk_stride + 9 * np; };
    double Bu[9], Bw[9], Mb[9], U0[9], S0[3], W0[9], A[9], B[9], Sv[3];
    ld_buf<9>(U0, op.aux0, np, t, 9);
    ld_buf<3>(S0, op.aux0 + 9 * np, np, t, 3);
    ld_mat<9>(W0, wv, 0, np, t, 9);
    zero_arr<9>(Bu, 9);
    zero_arr<9>(Bw, 9);
    zero_arr<9>(Mb, 9);
    if (c.phase == PH_BIAS) {
        double t0[9], t1[9];
        zero_arr<9>(t0, 9);
        zero_arr<9>(t1, 9);
        for (int i = 1; i < k; ++i) {
            // Bu += U_i^T U_{k-i} ; Bw += W_i^T W_{k-i}
            ld_mat<9>(A, uv, i, np, t, 9);
            ld_mat<9>(B, uv, k - i, np, t, 9);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    Bu[r * 3 + cc] += A[r] * B[cc] + A[3 + r] * B[3 + cc] + A[6 + r] * B[6 + cc];
            // t0 += U_i diag(S_{k-i})   (partial (U S)_k)
            ld_mat<3>(Sv, sv, k - i, np, t, 3);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) t0[r * 3 + cc] += A[r * 3 + cc] * Sv[cc];
            // t1 += P0_i U_{k-i}^T       (partial (U S U^T)_k)
            ld_buf<9>(A, p0_slot(i), np, t, 9);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    t1[r * 3 + cc] += A[r * 3] * B[cc * 3] + A[r * 3 + 1] * B[cc * 3 + 1] +
                                      A[r * 3 + 2] * B[cc * 3 + 2];
            // Mb += P1_i W_{k-i}
            ld_buf<9>(A, p1_slot(i), np, t, 9);
            ld_mat<9>(B, wv, k - i, np, t, 9);
            mm_acc<3, 3, 3>(Mb, A, B, 3, 3, 3);
            ld_mat<9>(A, wv, i, np, t, 9);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    Bw[r * 3 + cc] += A[r] * B[cc] + A[3 + r] * B[3 + cc] + A[6 + r] * B[6 + cc];
        }
        // t1 += t0 U_0^T ;  Mb += t1 W_0
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int cc = 0; cc < 3; ++cc)
                t1[r * 3 + cc] += t0[r * 3] * U0[cc * 3] + t0[r * 3 + 1] * U0[cc * 3 + 1] +
                                  t0[r * 3 + 2] * U0[cc * 3 + 2];
        mm_acc<3, 3, 3>(Mb, t1, W0, 3, 3, 3);
        st_buf<9>(op.sb, Bu, np, t, 9);
        st_buf<9>(op.sb + 9 * np, Bw, np, t, 9);
        st_buf<9>(op.sb + 18 * np, Mb, np, t, 9);
    } else if (!c.sb_zero) {
        ld_buf<9>(Bu, op.sb, np, t, 9);
        ld_buf<9>(Bw, op.sb + 9 * np, np, t, 9);
        ld_buf<9>(Mb, op.sb + 18 * np, np, t, 9);
    }
    double Mk[9], Uk[9], Sk[3], Wk[9];
    ld_mat<9>(Mk, xv, k, np, t, 9);
    svdw3_fwd_full(Uk, Sk, Wk, Mk, Mb, U0, S0, W0, Bu, Bw);
    st_buf<9>(vbase(uv, k), Uk, np, t, 9);
    st_buf<3>(vbase(sv, k), Sk, np, t, 3);
    st_buf<9>(vbase(wv, k), Wk, np, t, 9);
    if (c.phase == PH_COEF) {
        // P0_k = sum_{j=0}^{k} U_j diag(S_{k-j}),  P1_k = sum_{j=0}^{k} P0_j U_{k-j}^T
        double P0[9], P1[9];
        zero_arr<9>(P0, 9);
        zero_arr<9>(P1, 9);
        for (int j = 0; j <= k; ++j) {
            if (j == k) {
#pragma unroll
                for (int q = 0; q < 9; ++q) A[q] = Uk[q];
            } else {
                ld_mat<9>(A, uv, j, np, t, 9);
            }
            if (j == 0) {
#pragma unroll
                for (int q = 0; q < 3; ++q) Sv[q] = Sk[q];
            } else {
                ld_mat<3>(Sv, sv, k - j, np, t, 3);
            }
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) P0[r * 3 + cc] += A[r * 3 + cc] * Sv[cc];
        }
        for (int j = 0; j <= k; ++j) {
            if (j == k) {
#pragma unroll
                for (int q = 0; q < 9; ++q) A[q] = P0[q];
            } else {
                ld_buf<9>(A, p0_slot(j), np, t, 9);
            }
            if (j == 0) {
#pragma unroll
                for (int q = 0; q < 9; ++q) B[q] = Uk[q];
            } else {
                ld_mat<9>(B, uv, k - j, np, t, 9);
            }
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    P1[r * 3 + cc] += A[r * 3] * B[cc * 3] + A[r * 3 + 1] * B[cc * 3 + 1] +
                                      A[r * 3 + 2] * B[cc * 3 + 2];
        }
        st_buf<9>(p0_slot(k), P0, np, t, 9);
        st_buf<9>(p1_slot(k), P1, np, t, 9);
    }
}

//! polar mode of SVD-W for n x n items, n != 3 (see op_svdw for the caches:
//! aux0 = [U0 (n^2), S0 (n)], self-bias = [D, Bpw] (2 n^2), hidden stack P_k)
SANM_HD inline void op_svdw_n(const DOp& op, const ItemCtx& c, int n) {
    const DVar &xv = c.var(op.in[0]), &wv = c.var(op.out[2]);
    const int64_t np = c.np;
    const size_t t = c.t;
    const int e = n * n;
    double U[MAXM], S[MAXD], W[MAXM];
    if (c.phase == PH_EVAL) {
        double M[MAXM];
        ld_mat<0>(M, xv, 0, np, t, e);
        svdwn(M, n, op.i[0] != 0, U, S, W);
        st_buf<0>(op.aux0, U, np, t, e);
        st_buf<0>(op.aux0 + (int64_t)e * np, S, np, t, n);
        st_buf<0>(wv.p0, W, np, t, e);
        st_buf<0>(c.var(op.out[0]).p0, U, np, t, e);
        st_buf<0>(c.var(op.out[1]).p0, S, np, t, n);
        return;
    }
    const int k = c.k;
    double D[MAXM], Bpw[MAXM];
    zero_arr<0>(D, e);
    zero_arr<0>(Bpw, e);
    if (c.phase == PH_BIAS) {
        double A[MAXM], B[MAXM];
        for (int i = 1; i < k; ++i) {
            // Bm += M_i M_{k-i}^T
            ld_mat<0>(A, xv, i, np, t, e);
            ld_mat<0>(B, xv, k - i, np, t, e);
            for (int r = 0; r < n; ++r)
                for (int cc = 0; cc < n; ++cc) {
                    double q = 0;
                    for (int m = 0; m < n; ++m) q += A[r * n + m] * B[cc * n + m];
                    D[r * n + cc] += q;
                }
            // Bp += P_i^T P_{k-i} ;  Bpw += P_i W_{k-i}
            ld_buf<0>(A, op.auxk + (int64_t)i * op.auxk_stride, np, t, e);
            ld_buf<0>(B, op.auxk + (int64_t)(k - i) * op.auxk_stride, np, t, e);
            for (int r = 0; r < n; ++r)
                for (int cc = 0; cc < n; ++cc) {
                    double q = 0;
                    for (int m = 0; m < n; ++m) q += A[m * n + r] * B[m * n + cc];
                    D[r * n + cc] -= q;
                }
            ld_mat<0>(B, wv, k - i, np, t, e);
            mm_acc<0, 0, 0>(Bpw, A, B, n, n, n);
        }
        st_buf<0>(op.sb, D, np, t, e);
        st_buf<0>(op.sb + (int64_t)e * np, Bpw, np, t, e);
    } else if (!c.sb_zero) {
        ld_buf<0>(D, op.sb, np, t, e);
        ld_buf<0>(Bpw, op.sb + (int64_t)e * np, np, t, e);
    }
    double Mk[MAXM], Pk[MAXM], Wk[MAXM];
    ld_mat<0>(Mk, xv, k, np, t, e);
    ld_buf<0>(U, op.aux0, np, t, e);
    ld_buf<0>(S, op.aux0 + (int64_t)e * np, np, t, n);
    ld_mat<0>(W, wv, 0, np, t, e);
    svdwn_fwd_p(Pk, Wk, Mk, U, S, W, D, Bpw, n);
    st_buf<0>(vbase(wv, k), Wk, np, t, e);
    if (c.phase == PH_COEF) {
        st_buf<0>(op.auxk + (int64_t)k * op.auxk_stride, Pk, np, t, e);
    }
}

/*!
 * SVD-W.  Outputs (U, S, W).  Polar mode (U and S have no readers;
 * libsanm/oprs/linalg.cpp:529-550): hidden stack P_k, self-bias cache =
 * [D = Bm - Bp, Bpw] (18 doubles).  Full mode (op.i[1]; U or S is read,
 * linalg.cpp:551-563): the three outputs keep their stacks, hidden stack
 * [P0_k = (U S)_k, P1_k = (U S U^T)_k] (18 doubles) makes the triple
 * convolution of Mbias O(k), self-bias cache = [Bu, Bw, Mbias] (27 doubles).
 * aux0 = [U0 (9), S0 (3)].
 */
SANM_HD inline void op_svdw(const DOp& op, const ItemCtx& c) {
    const int x = op.in[0], ow = op.out[2];
    const DVar &xv = c.var(x), &wv = c.var(ow);
    const int64_t np = c.np;
    const size_t t = c.t;
    if (xv.d1 != 3) {
        // general n: polar mode only (graph.cpp rejects the full mode)
        op_svdw_n(op, c, xv.d1);
        return;
    }
    double U[9], S[3], W[9];
    if (c.phase == PH_EVAL) {
        double M[9];
        ld_mat<9>(M, xv, 0, np, t, 9);
        svdw3(M, op.i[0] != 0, U, S, W);
        st_buf<9>(op.aux0, U, np, t, 9);
        st_buf<3>(op.aux0 + 9 * np, S, np, t, 3);
        st_buf<9>(wv.p0, W, np, t, 9);
        // outputs U and S are also materialised at order 0 (readable by
        // eval_unary_func-style evaluation)
        st_buf<9>(c.var(op.out[0]).p0, U, np, t, 9);
        st_buf<3>(c.var(op.out[1]).p0, S, np, t, 3);
        if (op.i[1] && op.auxk) {
            double P0[9], P1[9];
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) P0[r * 3 + cc] = U[r * 3 + cc] * S[cc];
            mm3_nt(P1, P0, U);
            st_buf<9>(op.auxk, P0, np, t, 9);
            st_buf<9>(op.auxk + 9 * np, P1, np, t, 9);
        }
        return;
    }
    if (op.i[1]) {
        op_svdw_full(op, c);
        return;
    }
    const int k = c.k;
    double D[9], Bpw[9];
    zero_arr<9>(D, 9);
    zero_arr<9>(Bpw, 9);
    if (c.phase == PH_BIAS) {
        double A[9], B[9];
        for (int i = 1; i < k; ++i) {
            // Bm += M_i M_{k-i}^T
            ld_mat<9>(A, xv, i, np, t, 9);
            ld_mat<9>(B, xv, k - i, np, t, 9);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    D[r * 3 + cc] += A[r * 3] * B[cc * 3] + A[r * 3 + 1] * B[cc * 3 + 1] +
                                     A[r * 3 + 2] * B[cc * 3 + 2];
            // Bp += P_i^T P_{k-i} ;  Bpw += P_i W_{k-i}
            ld_buf<9>(A, op.auxk + (int64_t)i * op.auxk_stride, np, t, 9);
            ld_buf<9>(B, op.auxk + (int64_t)(k - i) * op.auxk_stride, np, t, 9);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int cc = 0; cc < 3; ++cc)
                    D[r * 3 + cc] -= A[r] * B[cc] + A[3 + r] * B[3 + cc] + A[6 + r] * B[6 + cc];
            ld_mat<9>(B, wv, k - i, np, t, 9);
            mm_acc<3, 3, 3>(Bpw, A, B, 3, 3, 3);
        }
        st_buf<9>(op.sb, D, np, t, 9);
        st_buf<9>(op.sb + 9 * np, Bpw, np, t, 9);
    } else if (!c.sb_zero) {
        ld_buf<9>(D, op.sb, np, t, 9);
        ld_buf<9>(Bpw, op.sb + 9 * np, np, t, 9);
    }
    double Mk[9], Pk[9], Wk[9];
    ld_mat<9>(Mk, xv, k, np, t, 9);
    ld_buf<9>(U, op.aux0, np, t, 9);
    ld_buf<3>(S, op.aux0 + 9 * np, np, t, 3);
    ld_mat<9>(W, wv, 0, np, t, 9);
    svdw3_fwd_p(Pk, Wk, Mk, U, S, W, D, Bpw);
    st_buf<9>(vbase(wv, k), Wk, np, t, 9);
    if (c.phase == PH_COEF) {
        st_buf<9>(op.auxk + (int64_t)k * op.auxk_stride, Pk, np, t, 9);
    }
}

//! run the whole operator list for one item
SANM_HD inline void run_item(const DProgram& prog, size_t t, int k, int phase,
                             int sb_zero) {
    ItemCtx c{&prog, t, k, phase, sb_zero, prog.npad};
    for (int oi = 0; oi < prog.nops; ++oi) {
        const DOp& op = prog.ops[oi];
        if (op.const_op && phase != PH_EVAL) continue;
        switch (op.type) {
            case OP_PLACEHOLDER:
                if (phase == PH_BIAS) {
                    const DVar& v = c.var(op.out[0]);
                    for (int e = 0; e < v.e; ++e) c.st(op.out[0], k, e, 0.0);
                }
                break;
            case OP_CONSTANT:
                break;
            case OP_LINCOMB:
                op_lincomb(op, c);
                break;
            case OP_MULTIPLY:
                op_multiply(op, c);
                break;
            case OP_POW:
            case OP_LOG:
                op_unary(op, c);
                break;
            case OP_REDUCE_SUM:
                op_reduce_sum(op, c);
                break;
            case OP_MATMUL:
                if (op.i[0]) {
                    op_matmul<3, 3, 3>(op, c);
                } else {
                    op_matmul<0, 0, 0>(op, c);
                }
                break;
            case OP_TRANSPOSE:
                op_transpose(op, c);
                break;
            case OP_DET:
                op_det(op, c);
                break;
            case OP_MULEYE:
                op_muleye(op, c);
                break;
            case OP_MATINVMUL:
                if (op.i[2]) {
                    op_matinvmul<3>(op, c);
                } else {
                    op_matinvmul<0>(op, c);
                }
                break;
            case OP_SVDW:
                op_svdw(op, c);
                break;
            case OP_SLICE:
                op_slice(op, c);
                break;
            case OP_CONCAT:
                op_concat(op, c);
                break;
        }
    }
}

}  // namespace taylor
}  // namespace sanm_b200
