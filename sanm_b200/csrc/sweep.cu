/* sanm_b200/csrc/sweep.cu -- the element phase of one series expansion as a
 * device-resident sweep over the orders: gather (remap_in) -> Taylor
 * propagation -> element-to-node assembly (remap_out), i.e. the element side of
 * ANMDriverHelper::solve_expansion_coeffs (libsanm/anm.cpp:205-290) with the
 * nodal series x_0..x_K supplied by the caller.  C ABI: sanm_sweep_*. */
#include "capi_util.h"
#include "comm.h"
#include "prof.h"

namespace sanm_b200 {

class ElementSweep {
public:
    ElementSweep(std::shared_ptr<const Graph> g, int f,
                 std::shared_ptr<const Remap> rin,
                 std::shared_ptr<const Remap> rout, int order)
            : m_rin{std::move(rin)}, m_rout{std::move(rout)}, m_order{order} {
        rt::check_device();
        sanm_check(order >= 1 && order <= taylor::MAX_ORDER, "order=%d", order);
        m_prop.reset(new TaylorProp{std::move(g), f, order});
        m_nin = m_rin->nr_in();
        m_nout = m_rout->nr_out();
        m_xbuf[0].alloc((size_t)(order + 1) * m_nin);
        m_bbuf[0].alloc((size_t)(order + 1) * m_nout);
        use_buffer(0);
        m_prop->prepare(m_rin->out_shape());
        sanm_check(m_rout->nr_in() == m_prop->batch() * m_prop->out_elems(),
                   "remap_out input size does not match the function output");
        // nodal index ranges the local tets read (x) and contribute to (b)
        m_x_lo = m_nin;
        m_x_hi = 0;
        for (uint32_t c : m_rin->h_idx()) {
            m_x_lo = std::min<uint64_t>(m_x_lo, c);
            m_x_hi = std::max<uint64_t>(m_x_hi, (uint64_t)c + 1);
        }
        if (m_x_lo >= m_x_hi) m_x_lo = m_x_hi = 0;
        const auto& rp = m_rout->h_rowptr();
        m_b_lo = m_nout;
        m_b_hi = 0;
        for (uint64_t i = 0; i + 1 < rp.size(); ++i) {
            if (rp[i + 1] > rp[i]) {
                m_b_lo = std::min(m_b_lo, i);
                m_b_hi = i + 1;
            }
        }
        if (m_b_lo >= m_b_hi) m_b_lo = m_b_hi = 0;
    }

    /*! 0: all-reduce of the full right-hand sides (what the replicated solver
     * needs; default).  1: nodal vectors stay distributed -- only the interface
     * rows shared with neighbouring ranks are exchanged (comm::halo_sum), rows
     * outside the local range [b_lo, b_hi) are left untouched, and
     * run_host() transfers only the local ranges of x and b. */
    void set_exchange(int mode) {
        sanm_check(mode >= 0 && mode <= 2, "exchange mode %d", mode);
        // 2: like 1 but WITHOUT the exchange step: every rank only assembles
        // its own tets (per-GPU reference point of a scaling run; the
        // interface rows are then partial sums)
        m_no_exchange = mode == 2;
        if (mode == 2) mode = 1;
        m_halo_mode = mode == 1;
        // distributed nodal vectors: the second assembly level and the zero
        // fills run over this rank's rows only (rows outside [b_lo, b_hi) are
        // never read or transferred in this mode)
        m_prop->set_local_rows(m_halo_mode);
        if (m_halo_mode && !m_halo.ready) {
            comm::halo_setup(m_halo, m_b_lo, m_b_hi, m_nout);
        }
    }
    void local_ranges(uint64_t* r) const {
        r[0] = m_x_lo;
        r[1] = m_x_hi;
        r[2] = m_b_lo;
        r[3] = m_b_hi;
    }

    void set_x(const double* x) {
        use_buffer(0);
        rt::h2d(m_xp, x, m_xbuf[0].size() * sizeof(double));
    }

    void run() {
        run_impl([](int) {}, [](int) {});
        join_exchange();
    }

    /*! One sweep with HOST buffers, transfers overlapped with the element
     * kernels: x_k travels on the host->device pipe while earlier orders are
     * being propagated (order k only needs x_0..x_k), b_k leaves on the
     * device->host pipe as soon as order k has been assembled.  Returns when
     * b has arrived.  x and b should be pinned for the copies to overlap. */
    void run_host(const double* x, double* b) {
        submit_host(x, b);
        wait_host();
    }

    /*! Asynchronous form: enqueue one sweep (uploads, kernels, downloads) and
     * return.  Consecutive submissions alternate between two sets of device
     * buffers, so the uploads of sweep s+1 and the downloads of sweep s overlap
     * the kernels of the other (the kernels themselves run one sweep after the
     * other: there is one Taylor state).  x must stay valid until its sweep has
     * started, b until wait_host(). */
    void submit_host(const double* x, double* b) {
        const int K = m_order;
        const int idx = (int)(m_submitted++ & 1);
        if (!m_xbuf[idx].size()) {
            m_xbuf[idx].alloc((size_t)(K + 1) * m_nin);
            m_bbuf[idx].alloc((size_t)(K + 1) * m_nout);
        }
        use_buffer(idx);
        // the sweep that used these buffers before must have read its x ...
        if (m_kernels_done[idx]) {
            rt::pipe_wait(rt::PIPE_H2D, m_kernels_done[idx]);
            m_kernels_done[idx] = nullptr;
        } else {
            rt::pipe_wait(rt::PIPE_H2D, rt::fence_main());
        }
        // ... and its b must have left before the new kernels overwrite it
        if (m_d2h_done[idx]) {
            rt::main_wait(m_d2h_done[idx]);
            m_d2h_done[idx] = nullptr;
        }
        std::vector<void*> arrived(K + 1, nullptr);
        // distributed mode: only what the local tets read / contribute to
        const uint64_t xlo = m_halo_mode ? m_x_lo : 0, xhi = m_halo_mode ? m_x_hi : m_nin;
        const uint64_t blo = m_halo_mode ? m_b_lo : 0, bhi = m_halo_mode ? m_b_hi : m_nout;
        for (int k = 0; k <= K; ++k) {
            rt::pipe_h2d(m_xp + (size_t)k * m_nin + xlo,
                         x + (size_t)k * m_nin + xlo, (xhi - xlo) * sizeof(double));
            // x_K is never pushed (libsanm/anm.cpp:205-290: the last order
            // only produces its bias), but it is part of the caller's series
            arrived[k] = rt::fence_pipe(rt::PIPE_H2D);
        }
        run_impl(
                [&](int k) {
                    rt::main_wait(arrived[k]);
                    arrived[k] = nullptr;
                },
                [&](int k) {
                    rt::pipe_wait(rt::PIPE_D2H, b_ready_fence());
                    rt::pipe_d2h(b + (size_t)k * m_nout + blo,
                                 m_bp + (size_t)k * m_nout + blo,
                                 (bhi - blo) * sizeof(double));
                });
        join_exchange();
        // x_K: nothing on the main stream waits for it; keep stream order sane
        for (void*& f : arrived) {
            if (f) {
                rt::main_wait(f);
                f = nullptr;
            }
        }
        m_kernels_done[idx] = rt::fence_main();
        m_d2h_done[idx] = rt::fence_pipe(rt::PIPE_D2H);
    }

    //! all submitted sweeps have finished and their b have arrived
    void wait_host() {
        rt::pipe_sync(rt::PIPE_D2H);
        rt::sync();
        m_prop->check_device_error();
    }

    ~ElementSweep() {
        for (void*& f : m_kernels_done) rt::fence_drop(f);
        for (void*& f : m_d2h_done) rt::fence_drop(f);
    }

private:
    /*! b_k is complete on the main stream.  Halo mode: nothing on the main
     * stream reads b_k again during this sweep (there is no solve between the
     * orders of a sweep), so the interface exchange runs on the side stream
     * beside the element kernels of the next orders; consumers of b_k (the
     * device->host pipe, the end of the sweep) wait for the side stream. */
    void exchange(double* b) {
        if (m_no_exchange) return;
        if (m_halo_mode) {
            if (comm::world() > 1 && m_async_halo) {
                rt::pipe_wait(rt::PIPE_COMM, rt::fence_main());
                rt::ScopedPipe side{rt::PIPE_COMM};
                comm::halo_sum(b, m_halo);
                m_halo_inflight = true;
            } else {
                comm::halo_sum(b, m_halo);
            }
        } else {
            comm::allreduce_sum(b, m_nout);
        }
    }
    //! fence after which b_k (exchanged) may be read
    void* b_ready_fence() {
        return m_halo_inflight ? rt::fence_pipe(rt::PIPE_COMM) : rt::fence_main();
    }
    //! main stream waits for the exchanges still running on the side stream
    void join_exchange() {
        if (m_halo_inflight) {
            rt::main_wait(rt::fence_pipe(rt::PIPE_COMM));
            m_halo_inflight = false;
        }
    }

    //! before_x(k): x_k is about to be read; after_b(k): b_k is complete
    template <typename BeforeX, typename AfterB>
    void run_impl(BeforeX&& before_x, AfterB&& after_b) {
        const int K = m_order;
        TaylorProp& p = *m_prop;
        p.reset();
        if (p.attach_io(*m_rin, *m_rout)) {
            // gather and assembly inside the element kernels
            {
                SANM_PROF("push_xi");
                before_x(0);
                p.push_xi_from(m_xp, m_bp);
                exchange(m_bp);
                after_b(0);
            }
            for (int k = 1; k <= K; ++k) {
                {
                    SANM_PROF("taylor_next_order");
                    p.compute_next_order_bias_to(m_bp + (size_t)k * m_nout);
                    // the reference computes the per-tet Jacobian inside its
                    // first compute_next_order_bias (libsanm/symbolic.cpp:249-252)
                    if (k == 1) p.ensure_jacobian();
                    exchange(m_bp + (size_t)k * m_nout);
                    after_b(k);
                }
                if (k < K) {
                    SANM_PROF("push_xi");
                    before_x(k);
                    p.push_xi_from(m_xp + (size_t)k * m_nin);
                }
            }
            return;
        }
        const VecLayout soa_in{p.in_elems(), p.npad()};
        const VecLayout soa_out{p.out_elems(), p.npad()};
        {
            SANM_PROF("push_xi");
            before_x(0);
            m_rin->apply(p.input_slot(0), soa_in, m_xp, VecLayout{});
            p.push_xi_filled();
        }
        {
            SANM_PROF("remap_out");
            m_rout->apply(m_bp, VecLayout{}, p.output_slot(0), soa_out);
            exchange(m_bp);
            after_b(0);
        }
        for (int k = 1; k <= K; ++k) {
            {
                SANM_PROF("taylor_next_order");
                p.compute_next_order_bias();
            }
            {
                SANM_PROF("remap_out");
                m_rout->apply(m_bp + (size_t)k * m_nout, VecLayout{},
                              p.output_slot(k), soa_out);
                exchange(m_bp + (size_t)k * m_nout);
                after_b(k);
            }
            if (k < K) {
                SANM_PROF("push_xi");
                before_x(k);
                m_rin->apply(p.input_slot(k), soa_in,
                             m_xp + (size_t)k * m_nin, VecLayout{});
                p.push_xi_filled();
            }
        }
    }

public:
    void get_b(double* b) {
        rt::d2h(b, m_bp, m_bbuf[0].size() * sizeof(double));
        m_prop->check_device_error();
    }

    //! doubles of Taylor state per item and order that later orders re-read
    int state_doubles() const { return m_prop->state_doubles(); }

    double steps() const { return (double)m_prop->batch() * m_order; }

    double algo_bytes() const {
        // B_elem(k) = 8 * (S * (k + 1) + 24) per item
        const double S = state_doubles();
        double per_item = 0;
        for (int k = 1; k <= m_order; ++k) per_item += 8.0 * (S * (k + 1) + 24);
        return per_item * (double)m_prop->batch();
    }

    uint64_t arena_bytes() const { return m_prop->arena_bytes(); }

private:
    std::shared_ptr<const Remap> m_rin, m_rout;
    int m_order;
    uint64_t m_nin = 0, m_nout = 0;
    uint64_t m_x_lo = 0, m_x_hi = 0, m_b_lo = 0, m_b_hi = 0;
    bool m_halo_inflight = false, m_no_exchange = false;
    //! SANM_SYNC_HALO=1: exchange on the main stream (measurement aid)
    const bool m_async_halo = getenv("SANM_SYNC_HALO") == nullptr;
    bool m_halo_mode = false;
    comm::Halo m_halo;
    std::unique_ptr<TaylorProp> m_prop;
    //! two sets of device series: consecutive submit_host() calls alternate, so
    //! the transfers of one sweep overlap the kernels of the other
    rt::DBuf<double> m_xbuf[2], m_bbuf[2];
    double* m_xp = nullptr;  //!< buffers of the sweep being enqueued
    double* m_bp = nullptr;
    uint64_t m_submitted = 0;
    void* m_kernels_done[2] = {nullptr, nullptr};  //!< fence: sweep read x / wrote b
    void* m_d2h_done[2] = {nullptr, nullptr};      //!< fence: b left the device
    void use_buffer(int i) {
        m_xp = m_xbuf[i].ptr();
        m_bp = m_bbuf[i].ptr();
    }
};

}  // namespace sanm_b200

using namespace sanm_b200;

struct sanm_sweep {
    std::unique_ptr<ElementSweep> s;
};

extern "C" {

int sanm_sweep_create(sanm_graph_t* g, int f, sanm_remap_t* remap_inp,
                      sanm_remap_t* remap_out, int order, sanm_sweep_t** out) {
    return capi::guard([&] {
        *out = new sanm_sweep{std::make_unique<ElementSweep>(
                g->g, f, remap_inp->r, remap_out->r, order)};
    });
}
int sanm_sweep_destroy(sanm_sweep_t* s) {
    delete s;
    return 0;
}
int sanm_sweep_set_x(sanm_sweep_t* s, const double* x) {
    return capi::guard([&] { s->s->set_x(x); });
}
int sanm_sweep_run(sanm_sweep_t* s) {
    return capi::guard([&] { s->s->run(); });
}
int sanm_sweep_run_host(sanm_sweep_t* s, const double* x, double* b) {
    return capi::guard([&] { s->s->run_host(x, b); });
}
int sanm_sweep_submit_host(sanm_sweep_t* s, const double* x, double* b) {
    return capi::guard([&] { s->s->submit_host(x, b); });
}
int sanm_sweep_wait_host(sanm_sweep_t* s) {
    return capi::guard([&] { s->s->wait_host(); });
}
int sanm_sweep_set_exchange(sanm_sweep_t* s, int mode) {
    return capi::guard([&] { s->s->set_exchange(mode); });
}
int sanm_sweep_local_ranges(sanm_sweep_t* s, uint64_t* ranges4) {
    return capi::guard([&] { s->s->local_ranges(ranges4); });
}
int sanm_sweep_get_b(sanm_sweep_t* s, double* b) {
    return capi::guard([&] { s->s->get_b(b); });
}
int sanm_sweep_stats(sanm_sweep_t* s, double* steps, double* algo_bytes,
                     int* state_doubles, uint64_t* arena_bytes) {
    return capi::guard([&] {
        *steps = s->s->steps();
        *algo_bytes = s->s->algo_bytes();
        *state_doubles = s->s->state_doubles();
        *arena_bytes = s->s->arena_bytes();
    });
}
}
