/* sanm_b200/csrc/taylor.h -- host driver of the device Taylor propagation:
 * the B200 counterpart of sanm::symbolic::ParallelTaylorCoeffProp
 * (libsanm/symbolic.h:391-423; protocol libsanm/symbolic.cpp:162-289).
 */
#pragma once

#include "fused.h"
#include "graph.h"

namespace sanm_b200 {

class TaylorProp {
public:
    /*!
     * \param max_order highest order that will be requested; stacks are sized
     *      for it (the reference grows its TensorArrays on demand,
     *      libsanm/symbolic.h:48-50)
     */
    TaylorProp(std::shared_ptr<const Graph> g, int out_var, int max_order);
    ~TaylorProp();

    //! lower the graph for this input shape and allocate (idempotent for an
    //! unchanged shape); resets the order to "nothing pushed"
    void prepare(const Shape& x_shape);
    //! forget all pushed coefficients, keep program and memory
    void reset();

    bool prepared() const { return m_prepared; }
    const Lowered& lowered() const { return m_low; }
    const Shape& in_shape() const { return m_low.vars[m_low.input].shape; }
    const Shape& out_shape() const { return m_low.vars[m_low.output].shape; }
    uint64_t batch() const { return m_low.batch; }
    int64_t npad() const { return m_npad; }
    int in_elems() const { return m_low.vars[m_low.input].e; }
    int out_elems() const { return m_low.vars[m_low.output].e; }
    int order() const { return m_order; }
    bool xi_known() const { return m_xi_known; }

    //! device slot ([e_in][npad], structure-of-arrays) the caller fills with
    //! x_k before push_xi_filled(); k must be order() (or 0 for the first push)
    double* input_slot(int k) const;
    //! device slot of the output variable for order k
    const double* output_slot(int k) const;

    //! x_k has been written into input_slot(): propagate it (order 0: evaluate)
    void push_xi_filled();
    //! increase the order and compute the bias of the output (phase A)
    void compute_next_order_bias();

    /* fused I/O: when the graph runs on the fused element kernels and both
     * remaps have the mesh structure of fused_io.h, the kernels gather from /
     * assemble into nodal vectors themselves and the staging slots above are
     * not used.  attach_io() returns false otherwise (callers keep the staged
     * path: Remap::apply around push_xi_filled / compute_next_order_bias). */
    bool attach_io(const class Remap& rin, const class Remap& rout);
    const FusedIO* fused_io() const { return m_fused ? m_fused->io() : nullptr; }
    //! fused I/O only: assemble on the locally touched rows alone
    void set_local_rows(bool on) {
        m_local_rows = on;
        if (m_fused) m_fused->set_local_rows(on);
    }
    //! push_xi(remap_in(x)); at order 0 also fx = remap_out(f(..)) if non-null.
    //! x must stay valid until the next compute_next_order_bias_to()
    void push_xi_from(const double* x, double* fx = nullptr);
    //! b = remap_out(compute_next_order_bias())
    void compute_next_order_bias_to(double* b);

    //! host-buffer conveniences (row-major [N, ...] as in the reference)
    void push_xi_host(const double* x, const Shape& shape);
    void output_to_host(int k, double* out) const;

    /*!
     * Jacobian d out / d in at the order-0 point by forward mode.
     * FULL: device array [N][odim][idim] (the reference's batched FULL layout,
     * libsanm/tensor.h:509-601); ELEMWISE: [N][dim].
     * Must be called when order() <= 1 and before the order-1 coefficients are
     * pushed (the reference computes it inside the first
     * compute_next_order_bias, libsanm/symbolic.cpp:249-252).
     */
    enum JacType { JAC_ELEMWISE = 1, JAC_FULL = 2 };
    JacType jacobian_type() const {
        return m_low.elemwise ? JAC_ELEMWISE : JAC_FULL;
    }
    const double* ensure_jacobian();
    //! 0: the FULL Jacobian is [N][odim][idim]; otherwise it is stored
    //! structure-of-arrays [odim*idim][pitch] (fused element kernels)
    int64_t jacobian_pitch() const { return m_fused ? m_npad : 0; }
    //! copy the Jacobian to a host buffer in the reference's layout
    void jacobian_to_host(double* out);

    //! throw SANM_ERR_NUMERICAL if a kernel flagged 0^p with non-integer p;
    //! synchronises the stream
    void check_device_error();

    uint64_t arena_bytes() const {
        return m_fused ? m_fused->arena_bytes() : m_arena.size() * sizeof(double);
    }
    //! true when the graph runs on the fused element kernels (fused.h)
    bool is_fused() const { return (bool)m_fused; }
    //! doubles of Taylor state per item and order that later orders re-read
    int state_doubles() const;

private:
    std::shared_ptr<const Graph> m_graph;
    int m_out_var, m_max_order;
    bool m_prepared = false, m_xi_known = false, m_jac_done = false;
    bool m_in_jacobian = false;
    int m_order = -1;  //!< -1: nothing pushed
    int64_t m_npad = 0;
    Lowered m_low;
    rt::DBuf<double> m_arena, m_jac, m_stage;
    rt::DBuf<taylor::DVar> m_dvars;
    rt::DBuf<taylor::DOp> m_dops;
    rt::DBuf<int> m_flags;
    std::vector<taylor::DVar> m_hvars;
    std::vector<taylor::DOp> m_hops;
    taylor::DProgram m_prog{};
    std::unique_ptr<FusedProp> m_fused;
    bool m_local_rows = false;

    void launch(int k, int phase, int sb_zero);
};

//! row-major [n][e] (array of structures) <-> [e][npad] (structure of arrays)
void aos_to_soa(double* soa, const double* aos, uint64_t n, int e, int64_t npad);
void soa_to_aos(double* aos, const double* soa, uint64_t n, int e, int64_t npad);

}  // namespace sanm_b200
