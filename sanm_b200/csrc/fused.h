/* sanm_b200/csrc/fused.h -- fused element engine for the stress graphs of
 * `fea` (see fused_dev.h).  TaylorProp (taylor.h) hands the whole propagation
 * to this class when match_fea_graph() recognises the graph; every other
 * graph runs through the generic operator-list kernels of taylor_dev.h.
 */
#pragma once

#include "fused_dev.h"
#include "fused_io.h"
#include "graph.h"

namespace sanm_b200 {

struct FusedSpec {
    fused::Material material = fused::MAT_NONE;
    fused::Params params{};
    int xbias_gop = -1;  //!< Graph::ops index of the constant added to x (or -1)
    int dminv_gop = -1;  //!< Graph::ops index of the constant DmInv
};

/*!
 * Recognise  out = pk1((placeholder + const) @ const)  for the three energy
 * models (graphs built by fea/material.cpp:55-99).  Returns MAT_NONE for
 * anything else.
 */
FusedSpec match_fea_graph(const Graph& g, int out_var, const Shape& x_shape);

class FusedProp {
public:
    FusedProp(const Graph& g, const FusedSpec& spec, uint64_t batch,
              int max_order);

    uint64_t batch() const { return m_n; }
    int64_t npad() const { return m_npad; }
    int state_doubles() const { return m_S; }
    uint64_t arena_bytes() const { return m_arena.size() * sizeof(double); }

    void reset();
    //! staging buffers, structure-of-arrays [9][npad]
    double* ds_slot() const { return m_ds.ptr(); }
    const double* out_slot();  //!< flushes a pending push first

    //! ds_slot() holds Ds_k: order 0 evaluates, k >= 1 is deferred so that it
    //! fuses with the following next_bias()
    void push(int k);
    //! bias of order kb = (last pushed order) + 1 into out_slot()
    void next_bias(int kb);
    /* fused I/O (fused_io.h): the kernels gather from / assemble into nodal
     * vectors themselves.  push_from(0, x0, fx) evaluates order 0 and
     * assembles f(x0) into fx (may be null); push_from(k >= 1, xk) is deferred
     * and next_bias_to(k + 1, b) runs coefficient phase k + bias phase k + 1 +
     * assembly in one kernel (plus the small second-level reduction). */
    void attach_io(std::shared_ptr<const FusedIO> io);
    bool has_io() const { return (bool)m_io; }
    //! assembled vectors are written on the rows this rank contributes to
    //! only (FusedIO::row_lo/row_hi); see ElementSweep::set_exchange
    void set_local_rows(bool on) { m_local_rows = on; }
    const FusedIO* io() const { return m_io.get(); }
    void push_from(int k, const double* x, double* fx = nullptr);
    void next_bias_to(int kb, double* b);
    //! d out / d in at the order-0 point, [81][npad] (entry o * 9 + i)
    void jacobian(double* jac);
    int* err_flag() const { return m_err.ptr(); }

private:
    fused::Material m_mat;
    fused::Params m_params;
    uint64_t m_n;
    int64_t m_npad;
    int m_K, m_S;
    int m_pending = -1;  //!< order whose push has not been launched yet
    const double* m_pending_x = nullptr;  //!< its nodal vector (fused I/O)
    std::shared_ptr<const FusedIO> m_io;
    bool m_out_is_bias = false;
    bool m_local_rows = false;
    rt::DBuf<double> m_arena, m_aux, m_ds, m_out, m_dminv, m_xbias;
    rt::DBuf<int> m_err;

    template <class M>
    void launch_step(int k_push, int kb, bool want_out, const double* x);
    void step(int k_push, int kb, bool want_out, const double* x);
    void flush();
};

}  // namespace sanm_b200
