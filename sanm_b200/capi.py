"""ctypes binding of the C ABI in ``include/sanm_b200.h``.

The classes mirror the reference's host-visible names (``ComputingGraph`` /
``SymbolVar`` sugar of ``libsanm/oprs.h:14-105``, ``ParallelTaylorCoeffProp`` of
``libsanm/symbolic.h:391-423``, ``SparseLinearDesc`` of ``libsanm/anm.h:24-85``,
``SparseSolver`` of ``libsanm/sparse_solver.h:17-87`` and the ANM drivers of
``libsanm/anm.h:98-305``) so that tests read like the reference's own.

:class:`Binding` is generic over a loaded shared library and a symbol prefix.
The product uses it with ``libsanm_b200.so`` / ``sanm_`` (see
:func:`sanm_b200.load`); the test-suite additionally instantiates it over the
oracle's ``libref_capi.so`` / ``ref_`` to drive the reference with the same
inputs.  Nothing in this package loads the oracle.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

_u64 = C.c_uint64
_dp = C.POINTER(C.c_double)
_u64p = C.POINTER(C.c_uint64)
_ip = C.POINTER(C.c_int)


class SANMError(RuntimeError):
    """libsanm/utils.h:34-50"""
    code = 1


class SANMAssertionError(SANMError):
    code = 2


class SANMNumericalError(SANMError):
    code = 3


class SANMCudaError(SANMError):
    code = 4


_ERRS = {1: SANMError, 2: SANMAssertionError, 3: SANMNumericalError,
         4: SANMCudaError}


class Hyper(C.Structure):
    """``sanm_anm_hyper_t``; defaults of libsanm/anm.h:100-114,244-248."""
    _fields_ = [("use_pade", C.c_int), ("sanity_check", C.c_int),
                ("order", C.c_int), ("maxr", C.c_double),
                ("solution_check_tol", C.c_double),
                ("xcoeff_l2_penalty", C.c_double),
                ("converge_rms", C.c_double)]

    def __init__(self, **kw):
        super().__init__(use_pade=0, sanity_check=1, order=8, maxr=1e-6,
                         solution_check_tol=1e-4, xcoeff_l2_penalty=0.0,
                         converge_rms=1e-5)
        for k, v in kw.items():
            if k not in dict(self._fields_):
                raise KeyError(k)
            setattr(self, k, v)


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def _dims(shape) -> "C.Array":
    return (C.c_uint64 * max(len(shape), 1))(*[int(i) for i in shape])


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_dp)


import os as _os
import time as _time
_TRACE_CALLS = bool(_os.environ.get("SANM_TRACE_CALLS"))


class Binding:
    """All entry points of one implementation (``sanm_`` or ``ref_``)."""

    def __init__(self, lib: C.CDLL, prefix: str):
        self.lib = lib
        self.prefix = prefix
        fn = getattr(lib, prefix + "last_error")
        fn.restype = C.c_char_p
        self._last_error = fn

    def fn(self, name):
        return getattr(self.lib, self.prefix + name)

    def has(self, name) -> bool:
        return hasattr(self.lib, self.prefix + name)

    #: host wall time per entry point when SANM_TRACE_CALLS is set (a
    #: measurement aid for tools/solve_profile.py; off by default)
    call_times: dict = {}

    def call(self, name, *args):
        if _TRACE_CALLS:
            t0 = _time.perf_counter()
            st = self.fn(name)(*args)
            rec = Binding.call_times.setdefault(self.prefix + name, [0.0, 0])
            rec[0] += _time.perf_counter() - t0
            rec[1] += 1
        else:
            st = self.fn(name)(*args)
        if st != 0:
            msg = self._last_error().decode(errors="replace")
            raise _ERRS.get(st, SANMError)(msg)

    # convenience constructors -------------------------------------------
    def ComputingGraph(self) -> "ComputingGraph":
        return ComputingGraph(self)

    def launch_count(self) -> int:
        f = self.fn("launch_count")
        f.restype = C.c_uint64
        return int(f())

    def prof_get(self, tag: str):
        ms = C.c_double()
        n = C.c_uint64()
        self.call("prof_get", tag.encode(), C.byref(ms), C.byref(n))
        return ms.value, n.value

    def prof_reset(self):
        self.call("prof_reset")

    def prof_dump(self) -> str:
        buf = C.create_string_buffer(1 << 16)
        self.call("prof_dump", buf, _u64(len(buf)))
        return buf.value.decode()


class SymbolVar:
    """libsanm/oprs.h:14-67"""

    def __init__(self, graph: "ComputingGraph", vid: int):
        self.graph = graph
        self.id = vid

    def _mk(self, name, *args) -> "SymbolVar":
        out = C.c_int()
        self.graph.b.call("graph_" + name, self.graph.h, *args, C.byref(out))
        return SymbolVar(self.graph, out.value)

    def __add__(self, rhs):
        if isinstance(rhs, SymbolVar):
            return self.graph.linear_combine([(1.0, self), (1.0, rhs)])
        return self.graph.linear_combine([(1.0, self)], float(rhs))

    __radd__ = __add__

    def __sub__(self, rhs):
        if isinstance(rhs, SymbolVar):
            return self.graph.linear_combine([(1.0, self), (-1.0, rhs)])
        return self + (-float(rhs))

    def __rsub__(self, lhs):
        return self.graph.linear_combine([(-1.0, self)], float(lhs))

    def __mul__(self, rhs):
        if isinstance(rhs, SymbolVar):
            return self._mk("multiply", C.c_int(self.id), C.c_int(rhs.id))
        return self.graph.linear_combine([(float(rhs), self)], 0.0)

    __rmul__ = __mul__

    def pow(self, e: float) -> "SymbolVar":
        return self._mk("pow", C.c_int(self.id), C.c_double(e))

    def log(self) -> "SymbolVar":
        return self._mk("log", C.c_int(self.id))

    def reduce_sum(self, axis: int, keepdim: bool = True) -> "SymbolVar":
        return self._mk("reduce_sum", C.c_int(self.id), C.c_int(axis),
                        C.c_int(int(keepdim)))

    def batched_transpose(self) -> "SymbolVar":
        return self._mk("batched_transpose", C.c_int(self.id))

    def batched_matmul(self, rhs: "SymbolVar") -> "SymbolVar":
        return self._mk("batched_matmul", C.c_int(self.id), C.c_int(rhs.id))

    def batched_det(self) -> "SymbolVar":
        return self._mk("batched_det", C.c_int(self.id))

    def batched_mul_eye(self, dim: int) -> "SymbolVar":
        return self._mk("batched_mul_eye", C.c_int(self.id), _u64(dim))

    def batched_matinv(self) -> "SymbolVar":
        return self.graph.batched_mat_inv_mul(self, None, True)

    def batched_svd_w(self, require_rotation: bool = False):
        out = (C.c_int * 3)()
        self.graph.b.call("graph_batched_svd_w", self.graph.h, C.c_int(self.id),
                          C.c_int(int(require_rotation)), out)
        return [SymbolVar(self.graph, out[i]) for i in range(3)]

    def slice(self, axis, begin=None, end=None, stride=1) -> "SymbolVar":
        return self._mk("slice", C.c_int(self.id), C.c_int(axis),
                        C.c_int(begin is not None), C.c_int(begin or 0),
                        C.c_int(end is not None), C.c_int(end or 0),
                        C.c_int(stride))


class ComputingGraph:
    """libsanm/symbolic.h:283-293 plus the free functions of oprs.h:69-105"""

    def __init__(self, b: Binding):
        self.b = b
        self.h = C.c_void_p()
        b.call("graph_create", C.byref(self.h))

    def __del__(self):
        try:
            if self.h:
                self.b.fn("graph_destroy")(self.h)
                self.h = None
        except Exception:
            pass

    def placeholder(self) -> SymbolVar:
        out = C.c_int()
        self.b.call("graph_placeholder", self.h, C.byref(out))
        return SymbolVar(self, out.value)

    def constant(self, val) -> SymbolVar:
        val = _f64(val)
        out = C.c_int()
        self.b.call("graph_constant", self.h, _ptr(val), C.c_int(val.ndim),
                    _dims(val.shape), C.byref(out))
        return SymbolVar(self, out.value)

    def linear_combine(self, terms, bias: float = 0.0) -> SymbolVar:
        n = len(terms)
        coeffs = (C.c_double * n)(*[float(t[0]) for t in terms])
        vars_ = (C.c_int * n)(*[t[1].id for t in terms])
        out = C.c_int()
        self.b.call("graph_linear_combine", self.h, C.c_int(n), coeffs, vars_,
                    C.c_double(bias), C.byref(out))
        return SymbolVar(self, out.value)

    def batched_mat_inv_mul(self, x: SymbolVar, a: Optional[SymbolVar],
                            is_left: bool) -> SymbolVar:
        out = C.c_int()
        self.b.call("graph_batched_mat_inv_mul", self.h, C.c_int(x.id),
                    C.c_int(a.id if a is not None else -1),
                    C.c_int(int(is_left)), C.byref(out))
        return SymbolVar(self, out.value)

    def concat(self, xs: Sequence[SymbolVar], axis: int) -> SymbolVar:
        n = len(xs)
        vars_ = (C.c_int * n)(*[v.id for v in xs])
        out = C.c_int()
        self.b.call("graph_concat", self.h, C.c_int(n), vars_, C.c_int(axis),
                    C.byref(out))
        return SymbolVar(self, out.value)

    def eval_unary_func(self, y: SymbolVar, xval) -> np.ndarray:
        """libsanm/symbolic.h:426"""
        xval = _f64(xval)
        orank = C.c_int()
        odims = (C.c_uint64 * 3)()
        self.b.call("eval_unary_func", self.h, C.c_int(y.id), _ptr(xval),
                    C.c_int(xval.ndim), _dims(xval.shape), C.byref(orank),
                    odims, None)
        out = np.empty([odims[i] for i in range(orank.value)])
        self.b.call("eval_unary_func", self.h, C.c_int(y.id), _ptr(xval),
                    C.c_int(xval.ndim), _dims(xval.shape), C.byref(orank),
                    odims, _ptr(out))
        return out


class TaylorCoeffProp:
    """ParallelTaylorCoeffProp, libsanm/symbolic.h:391-423."""

    def __init__(self, y: SymbolVar, max_order: int = 24):
        self.b = y.graph.b
        self.graph = y.graph
        self.h = C.c_void_p()
        if self.b.prefix == "ref_":
            self.b.call("prop_create", y.graph.h, C.c_int(y.id),
                        C.byref(self.h))
        else:
            self.b.call("prop_create", y.graph.h, C.c_int(y.id),
                        C.c_int(max_order), C.byref(self.h))

    def __del__(self):
        try:
            if self.h:
                self.b.fn("prop_destroy")(self.h)
                self.h = None
        except Exception:
            pass

    def out_shape(self):
        rank = C.c_int()
        dims = (C.c_uint64 * 3)()
        self.b.call("prop_out_shape", self.h, C.byref(rank), dims)
        return tuple(int(dims[i]) for i in range(rank.value))

    def push_xi(self, x) -> "TaylorCoeffProp":
        x = _f64(x)
        self.b.call("prop_push_xi", self.h, _ptr(x), C.c_int(x.ndim),
                    _dims(x.shape))
        return self

    def gather_yi(self) -> np.ndarray:
        out = np.empty(self.out_shape())
        self.b.call("prop_gather_yi", self.h, _ptr(out))
        return out

    def compute_next_order_bias(self) -> np.ndarray:
        out = np.empty(self.out_shape())
        self.b.call("prop_compute_next_order_bias", self.h, _ptr(out))
        return out

    def get_jacobian(self):
        """returns (type, coeff): type 1 ELEMWISE [batch, dim], 2 FULL
        [batch, odim, idim]"""
        t = C.c_int()
        batch, odim, idim = _u64(), _u64(), _u64()
        self.b.call("prop_get_jacobian", self.h, C.byref(t), C.byref(batch),
                    C.byref(odim), C.byref(idim), None)
        shape = ((batch.value, odim.value) if t.value == 1 else
                 (batch.value, odim.value, idim.value))
        out = np.empty(shape)
        self.b.call("prop_get_jacobian", self.h, C.byref(t), C.byref(batch),
                    C.byref(odim), C.byref(idim), _ptr(out))
        return t.value, out

    def jacobian_full(self) -> np.ndarray:
        t, c = self.get_jacobian()
        if t == 2:
            return c
        out = np.zeros(c.shape + (c.shape[1],))
        idx = np.arange(c.shape[1])
        out[:, idx, idx] = c
        return out


class SparseLinearDesc:
    """libsanm/anm.h:24-85, given as CSR over the flattened tensors."""

    def __init__(self, b: Binding, out_shape, inp_shape, rowptr=None, idx=None,
                 coeff=None, identity=False):
        self.b = b
        self.out_shape = tuple(int(i) for i in out_shape)
        self.inp_shape = tuple(int(i) for i in inp_shape)
        self.h = C.c_void_p()
        if identity:
            b.call("remap_identity", C.c_int(len(self.inp_shape)),
                   _dims(self.inp_shape), C.c_int(len(self.out_shape)),
                   _dims(self.out_shape), C.byref(self.h))
            return
        self.rowptr = np.ascontiguousarray(rowptr, dtype=np.uint64)
        self.idx = np.ascontiguousarray(idx, dtype=np.uint64)
        self.coeff = _f64(coeff)
        nout = int(np.prod(self.out_shape))
        assert self.rowptr.shape == (nout + 1,)
        assert self.idx.shape == self.coeff.shape == (int(self.rowptr[-1]),)
        b.call("remap_create", C.c_int(len(self.out_shape)),
               _dims(self.out_shape), C.c_int(len(self.inp_shape)),
               _dims(self.inp_shape), self.rowptr.ctypes.data_as(_u64p),
               self.idx.ctypes.data_as(_u64p), _ptr(self.coeff),
               C.byref(self.h))

    @classmethod
    def make_identity(cls, b: Binding, inp_shape, out_shape=None):
        return cls(b, out_shape if out_shape is not None else inp_shape,
                   inp_shape, identity=True)

    def __del__(self):
        try:
            if self.h:
                self.b.fn("remap_destroy")(self.h)
                self.h = None
        except Exception:
            pass

    def apply(self, x) -> np.ndarray:
        x = _f64(x)
        assert x.size == int(np.prod(self.inp_shape))
        y = np.empty(self.out_shape)
        self.b.call("remap_apply", self.h, _ptr(x), _ptr(y))
        return y

    def apply_static(self, x) -> np.ndarray:
        """sanm_remap_apply_static: the same gather staged in the library's
        static device scratch (no device allocation)"""
        x = _f64(x)
        assert x.size == int(np.prod(self.inp_shape))
        y = np.empty(self.out_shape)
        self.b.call("remap_apply_static", self.h, _ptr(x), _ptr(y))
        return y


class SparseSolver:
    """libsanm/sparse_solver.h:17-87"""

    def __init__(self, b: Binding, n: int):
        self.b = b
        self.n = int(n)
        self.h = C.c_void_p()
        b.call("sparse_create", _u64(n), C.byref(self.h))

    def __del__(self):
        try:
            if self.h:
                self.b.fn("sparse_destroy")(self.h)
                self.h = None
        except Exception:
            pass

    def add_constraints(self, cidx, xidx, coeff):
        cidx = np.ascontiguousarray(cidx, dtype=np.uint64)
        xidx = np.ascontiguousarray(xidx, dtype=np.uint64)
        coeff = _f64(coeff)
        self.b.call("sparse_add_constraints", self.h, _u64(len(coeff)),
                    cidx.ctypes.data_as(_u64p), xidx.ctypes.data_as(_u64p),
                    _ptr(coeff))
        return self

    def prepare(self, l2_penalty: float = 0.0):
        self.b.call("sparse_prepare", self.h, C.c_double(l2_penalty))
        return self

    def solve(self, b) -> np.ndarray:
        b = _f64(b)
        x = np.empty(self.n)
        self.b.call("sparse_solve", self.h, _ptr(b), _ptr(x))
        return x

    def apply(self, x) -> np.ndarray:
        x = _f64(x)
        y = np.empty(self.n)
        self.b.call("sparse_apply", self.h, _ptr(x), _ptr(y))
        return y

    def coeff_l2(self) -> float:
        out = C.c_double()
        self.b.call("sparse_coeff_l2", self.h, C.byref(out))
        return out.value

    def stats(self):
        nnz, flops, nsn, nlev = _u64(), C.c_double(), _u64(), _u64()
        self.b.call("sparse_stats", self.h, C.byref(nnz), C.byref(flops),
                    C.byref(nsn), C.byref(nlev))
        return dict(nnz_lu=nnz.value, flops=flops.value,
                    supernodes=nsn.value, levels=nlev.value)


class _ANMBase:
    def __init__(self, f: SymbolVar, remap_inp: SparseLinearDesc,
                 remap_out: SparseLinearDesc):
        self.b = f.graph.b
        self.graph = f.graph
        self.remap_inp = remap_inp
        self.remap_out = remap_out
        self.h = C.c_void_p()

    def __del__(self):
        try:
            if self.h:
                self.b.fn("anm_destroy")(self.h)
                self.h = None
        except Exception:
            pass

    def update_approx(self):
        self.b.call("anm_update_approx", self.h)
        return self

    def get_t_upper(self) -> float:
        out = C.c_double()
        self.b.call("anm_get_t_upper", self.h, C.byref(out))
        return out.value

    def solve_a(self, t: float) -> float:
        out = C.c_double()
        self.b.call("anm_solve_a", self.h, C.c_double(t), C.byref(out))
        return out.value

    def eval(self, a: float):
        x = np.empty(self._x_shape)
        t = C.c_double()
        self.b.call("anm_eval", self.h, C.c_double(a), _ptr(x), C.byref(t))
        return x, t.value

    def get_nr_iter(self) -> int:
        out = _u64()
        self.b.call("anm_nr_iter", self.h, C.byref(out))
        return out.value

    def xt_coeffs(self):
        n = _u64()
        self.b.call("anm_nr_xt_coeffs", self.h, C.byref(n))
        ret = []
        for i in range(n.value):
            v = np.empty(int(np.prod(self._x_shape)) + 1)
            self.b.call("anm_xt_coeff", self.h, _u64(i), _ptr(v))
            ret.append(v)
        return ret

    def k_csr(self):
        """the tangent matrix of the last expansion as a scipy CSR matrix:
        ``sanm_anm_k_csr`` here, the matrix last handed to SparseSolver::
        prepare for the oracle (``ref_last_prepared_csr``)"""
        import scipy.sparse as sp
        n, nnz = _u64(), _u64()
        if self.b.prefix == "ref_":
            self.b.call("last_prepared_csr", C.byref(n), C.byref(nnz), None,
                        None, None)
            ia = np.empty(n.value + 1, dtype=np.int32)
            ja = np.empty(nnz.value, dtype=np.int32)
            a = np.empty(nnz.value)
            self.b.call("last_prepared_csr", C.byref(n), C.byref(nnz),
                        ia.ctypes.data_as(C.c_void_p),
                        ja.ctypes.data_as(C.c_void_p), _ptr(a))
        else:
            self.b.call("anm_k_csr", self.h, C.byref(n), C.byref(nnz), None,
                        None, None)
            ia = np.empty(n.value + 1, dtype=np.uint32)
            ja = np.empty(nnz.value, dtype=np.uint32)
            a = np.empty(nnz.value)
            self.b.call("anm_k_csr", self.h, C.byref(n), C.byref(nnz),
                        ia.ctypes.data_as(C.c_void_p),
                        ja.ctypes.data_as(C.c_void_p), _ptr(a))
        return sp.csr_matrix((a, ja.astype(np.int64), ia.astype(np.int64)),
                             shape=(n.value, n.value))

    def element_stats(self):
        vals = [C.c_double() for _ in range(5)]
        self.b.call("anm_element_stats", self.h, *[C.byref(v) for v in vals])
        keys = ["tet_order_steps", "element_ms", "solve_ms", "factor_ms",
                "assemble_ms"]
        out = {k: v.value for k, v in zip(keys, vals)}
        if self.b.has("anm_lu_stats"):
            u = [_u64() for _ in range(3)]
            fl = C.c_double()
            w = [_u64() for _ in range(3)]
            self.b.call("anm_lu_stats", self.h, C.byref(u[0]), C.byref(u[1]),
                        C.byref(u[2]), C.byref(fl), C.byref(w[0]),
                        C.byref(w[1]), C.byref(w[2]))
            out.update(n=u[0].value, nnz_a=u[1].value,
                       factor_entries=u[2].value, factor_flops=fl.value,
                       nr_supernodes=w[0].value, nr_levels=w[1].value,
                       max_front=w[2].value)
            if self.b.has("anm_robustness_stats"):
                r = [_u64() for _ in range(3)]
                self.b.call("anm_robustness_stats", self.h,
                            *[C.byref(v) for v in r])
                out.update(pade_recovered=r[0].value,
                           lu_perturbed_pivots=r[1].value,
                           lu_refine_steps=r[2].value)
            out["nr_solves"] = self.b.prof_get("sparse_solve")[1]
            out["nr_factor"] = self.b.prof_get("sparse_prep")[1]
        return out


class ANMSolverVecScale(_ANMBase):
    """libsanm/anm.h:213-239: f(x) + t*v = 0"""

    def __init__(self, f, remap_inp, remap_out, x0, t0, v,
                 hyper: Optional[Hyper] = None):
        super().__init__(f, remap_inp, remap_out)
        x0 = _f64(x0)
        v = _f64(v)
        self._x_shape = x0.shape
        hyper = hyper or Hyper()
        self.b.call("anm_vecscale_create", f.graph.h, C.c_int(f.id),
                    remap_inp.h, remap_out.h, _ptr(x0), C.c_int(x0.ndim),
                    _dims(x0.shape), C.c_double(t0), _ptr(v), C.c_int(v.ndim),
                    _dims(v.shape), C.byref(hyper), C.byref(self.h))


class ANMEqnSolver(_ANMBase):
    """libsanm/anm.h:242-275: f(x) + y = 0"""

    def __init__(self, f, remap_inp, remap_out, x0, y,
                 hyper: Optional[Hyper] = None):
        super().__init__(f, remap_inp, remap_out)
        x0 = _f64(x0)
        y = _f64(y)
        self._x_shape = x0.shape
        hyper = hyper or Hyper()
        self.b.call("anm_eqn_create", f.graph.h, C.c_int(f.id), remap_inp.h,
                    remap_out.h, _ptr(x0), C.c_int(x0.ndim), _dims(x0.shape),
                    _ptr(y), C.c_int(y.ndim), _dims(y.shape), C.byref(hyper),
                    C.byref(self.h))

    def next_iter(self):
        self.b.call("anm_eqn_next_iter", self.h)
        return self

    def converged(self) -> bool:
        out = C.c_int()
        self.b.call("anm_eqn_converged", self.h, C.byref(out))
        return bool(out.value)

    def residual_rms(self) -> float:
        out = C.c_double()
        self.b.call("anm_eqn_residual_rms", self.h, C.byref(out))
        return out.value

    def get_x(self) -> np.ndarray:
        x = np.empty(self._x_shape)
        self.b.call("anm_eqn_get_x", self.h, _ptr(x))
        return x


class ANMImplicitSolver(_ANMBase):
    """libsanm/anm.h:277-305: F(x, t) = F(x0, t0)"""

    def __init__(self, f, remap_inp, remap_out, x0, t0,
                 hyper: Optional[Hyper] = None):
        super().__init__(f, remap_inp, remap_out)
        x0 = _f64(x0)
        self._x_shape = x0.shape
        hyper = hyper or Hyper()
        self.b.call("anm_implicit_create", f.graph.h, C.c_int(f.id),
                    remap_inp.h, remap_out.h, _ptr(x0), _u64(x0.size),
                    C.c_double(t0), C.byref(hyper), C.byref(self.h))

    def fx0(self) -> np.ndarray:
        out = np.empty(self._x_shape)
        self.b.call("anm_implicit_fx0", self.h, _ptr(out))
        return out


class ElementSweep:
    """``sanm_sweep_*``: the element phase of one series expansion
    (libsanm/anm.cpp:205-290 without the solves) resident on the device."""

    def __init__(self, f: SymbolVar, remap_inp: SparseLinearDesc,
                 remap_out: SparseLinearDesc, order: int):
        self.b = f.graph.b
        self.graph = f.graph
        self.remap_inp, self.remap_out = remap_inp, remap_out
        self.order = int(order)
        self.n_in = int(np.prod(remap_inp.inp_shape))
        self.n_out = int(np.prod(remap_out.out_shape))
        self.h = C.c_void_p()
        self.b.call("sweep_create", f.graph.h, C.c_int(f.id), remap_inp.h,
                    remap_out.h, C.c_int(self.order), C.byref(self.h))

    def __del__(self):
        try:
            if self.h:
                self.b.fn("sweep_destroy")(self.h)
                self.h = None
        except Exception:
            pass

    def set_x(self, x):
        x = _f64(x)
        assert x.shape == (self.order + 1, self.n_in)
        self.b.call("sweep_set_x", self.h, _ptr(x))

    def run(self):
        self.b.call("sweep_run", self.h)

    def get_b(self, out=None) -> np.ndarray:
        if out is None:
            out = np.empty((self.order + 1, self.n_out))
        self.b.call("sweep_get_b", self.h, _ptr(out))
        return out

    def submit_host(self, x, out):
        """asynchronous run_host; keep x and out alive until wait_host()"""
        assert x.dtype == np.float64 and x.flags.c_contiguous
        assert x.shape == (self.order + 1, self.n_in)
        assert out.dtype == np.float64 and out.flags.c_contiguous
        assert out.shape == (self.order + 1, self.n_out)
        self.b.call("sweep_submit_host", self.h, _ptr(x), _ptr(out))

    def wait_host(self):
        self.b.call("sweep_wait_host", self.h)

    def set_exchange(self, mode: int):
        """0: all-reduce of full vectors; 1: distributed vectors, halo sums"""
        self.b.call("sweep_set_exchange", self.h, C.c_int(mode))

    def local_ranges(self):
        r = (C.c_uint64 * 4)()
        self.b.call("sweep_local_ranges", self.h, r)
        return tuple(int(v) for v in r)

    def run_host(self, x, out=None) -> np.ndarray:
        """set_x + run + get_b with the copies overlapped (host buffers)"""
        x = _f64(x)
        assert x.shape == (self.order + 1, self.n_in)
        if out is None:
            out = np.empty((self.order + 1, self.n_out))
        assert out.dtype == np.float64 and out.flags.c_contiguous
        self.b.call("sweep_run_host", self.h, _ptr(x), _ptr(out))
        return out

    def stats(self):
        steps, ab = C.c_double(), C.c_double()
        sd, arena = C.c_int(), _u64()
        self.b.call("sweep_stats", self.h, C.byref(steps), C.byref(ab),
                    C.byref(sd), C.byref(arena))
        return {"steps": steps.value, "algo_bytes": ab.value,
                "state_doubles": sd.value, "arena_bytes": arena.value}


def reference_sweep(f: SymbolVar, remap_inp: SparseLinearDesc,
                    remap_out: SparseLinearDesc, x: np.ndarray) -> np.ndarray:
    """The same sweep written with the reference's public classes
    (TaylorCoeffProp + SparseLinearDesc::apply), i.e. the loop of
    libsanm/anm.cpp:205-290 without the solves.  Works on any binding; the
    tests and the CPU baseline run it on the oracle."""
    K = x.shape[0] - 1
    prop = TaylorCoeffProp(f, max_order=max(K, 1))
    n_out = int(np.prod(remap_out.out_shape))
    b = np.zeros((K + 1, n_out))
    prop.push_xi(remap_inp.apply(x[0]))
    b[0] = remap_out.apply(prop.gather_yi()).reshape(-1)
    for k in range(1, K + 1):
        b[k] = remap_out.apply(prop.compute_next_order_bias()).reshape(-1)
        if k < K:
            prop.push_xi(remap_inp.apply(x[k]))
    return b
