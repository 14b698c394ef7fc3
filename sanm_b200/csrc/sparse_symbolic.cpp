/* sanm_b200/csrc/sparse_symbolic.cpp -- ordering + symbolic factorisation.
 *
 * Everything here runs once per sparsity pattern on the host.  The reference
 * delegates this step to PARDISO phase 11 with METIS nested dissection
 * (libsanm/sparse_solver.cpp:88-127, iparm[1]); neither is available, so the
 * ordering is a plain nested dissection on the supervariable-compressed graph
 * with separators taken from BFS level structures rooted at pseudo-peripheral
 * nodes (George's automatic nested dissection).
 */
#include "sparse_symbolic.h"
#include "rt.h"

#include <algorithm>
#include <cmath>
#include <limits>
#include <atomic>
#include <numeric>

#include <chrono>
#include <cstdio>
#include <cstdlib>
namespace {
struct StageTimer {
    bool on = getenv("SANM_LU_DEBUG") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    const char* cur = "start";
    void next(const char* name) {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        printf("[lu-analyze] %-42s %8.2f ms\n", cur,
               std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
        cur = name;
    }
    ~StageTimer() { next("end"); }
};
}  // namespace
#define SANM_STAGE(name) stage_timer.next(name)

namespace sanm_b200 {

namespace {

using u32 = uint32_t;
using u64 = uint64_t;

struct Adj {
    std::vector<u64> ptr;
    std::vector<u32> idx;
    u32 n() const { return (u32)ptr.size() - 1; }
    const u32* begin(u32 i) const { return idx.data() + ptr[i]; }
    const u32* end(u32 i) const { return idx.data() + ptr[i + 1]; }
    u32 deg(u32 i) const { return (u32)(ptr[i + 1] - ptr[i]); }
};

//! symmetric adjacency without the diagonal, sorted and unique
/*! Fast path of build_sym_adj: a structurally symmetric pattern with sorted
 * rows (what the ANM drivers hand over: K = R_out blockdiag(J) R_in of a
 * Galerkin assembly, explicit zeros kept) is its own adjacency once the
 * diagonal is dropped.  Symmetry is verified entry by entry (parallel binary
 * searches); the general transposing build below costs two cache-hostile
 * scatter passes over 2 nnz entries (0.8 s at 36 M nonzeros). */
bool sym_adj_fast(u32 n, const u32* rowptr, const u32* colidx, Adj& a) {
    int bad = 0;
    std::vector<u32> len(n);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(dynamic, 512) reduction(| : bad)
    for (int64_t i64 = 0; i64 < (int64_t)n; ++i64) {
        const u32 i = (u32)i64;
        u32 m = 0;
        for (u32 p = rowptr[i]; p < rowptr[i + 1] && !bad; ++p) {
            const u32 j = colidx[p];
            if (p > rowptr[i] && colidx[p - 1] >= j) bad = 1;  // not sorted / duplicate
            if (j == i) continue;
            if (j >= n) {
                bad = 1;
                break;
            }
            ++m;
            const u32* b = colidx + rowptr[j];
            const u32* e = colidx + rowptr[j + 1];
            const u32* it = std::lower_bound(b, e, i);
            if (it == e || *it != i) bad = 1;
        }
        len[i] = m;
    }
    if (bad) return false;
    a.ptr.assign(n + 1, 0);
    for (u32 i = 0; i < n; ++i) a.ptr[i + 1] = a.ptr[i] + len[i];
    a.idx.resize(a.ptr[n]);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static)
    for (int64_t i64 = 0; i64 < (int64_t)n; ++i64) {
        const u32 i = (u32)i64;
        auto out = a.idx.begin() + a.ptr[i];
        for (u32 p = rowptr[i]; p < rowptr[i + 1]; ++p) {
            if (colidx[p] != i) *out++ = colidx[p];
        }
    }
    return true;
}

Adj build_sym_adj(u32 n, const u32* rowptr, const u32* colidx) {
    {
        Adj fast;
        if (getenv("SANM_ND_NO_FAST_ADJ") == nullptr && sym_adj_fast(n, rowptr, colidx, fast)) {
            return fast;
        }
    }
    std::vector<u64> cnt(n + 1, 0);
    for (u32 i = 0; i < n; ++i) {
        for (u32 p = rowptr[i]; p < rowptr[i + 1]; ++p) {
            u32 j = colidx[p];
            if (j == i) continue;
            ++cnt[i + 1];
            ++cnt[j + 1];
        }
    }
    for (u32 i = 0; i < n; ++i) cnt[i + 1] += cnt[i];
    std::vector<u32> raw(cnt[n]);
    std::vector<u64> pos(cnt.begin(), cnt.end() - 1);
    for (u32 i = 0; i < n; ++i) {
        for (u32 p = rowptr[i]; p < rowptr[i + 1]; ++p) {
            u32 j = colidx[p];
            if (j == i) continue;
            raw[pos[i]++] = j;
            raw[pos[j]++] = i;
        }
    }
    Adj a;
    a.ptr.assign(n + 1, 0);
    std::vector<u32> ulen(n);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(dynamic, 256)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        auto b = raw.begin() + cnt[i], e = raw.begin() + cnt[i + 1];
        std::sort(b, e);
        ulen[i] = (u32)(std::unique(b, e) - b);
    }
    for (u32 i = 0; i < n; ++i) a.ptr[i + 1] = a.ptr[i] + ulen[i];
    a.idx.resize(a.ptr[n]);
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        std::copy(raw.begin() + cnt[i], raw.begin() + cnt[i] + ulen[i],
                  a.idx.begin() + a.ptr[i]);
    }
    return a;
}

/*!
 * indistinguishable-node compression: i ~ j iff adj(i) U {i} == adj(j) U {j}
 * (e.g. the three displacement unknowns of one mesh vertex).
 * \return sv[i] = compressed node of i; nodes numbered by first member
 */
std::vector<u32> supervariables(const Adj& a, u32& nr_sv) {
    const u32 n = a.n();
    std::vector<u64> h(n);
    auto mix = [](u64 x) {
        x ^= x >> 33;
        x *= 0xff51afd7ed558ccdULL;
        x ^= x >> 33;
        x *= 0xc4ceb9fe1a85ec53ULL;
        x ^= x >> 33;
        return x;
    };
    for (u32 i = 0; i < n; ++i) {
        u64 s = mix(i);
        for (auto p = a.begin(i); p != a.end(i); ++p) s += mix(*p);
        h[i] = s;
    }
    std::vector<u32> sv(n, ~0u);
    nr_sv = 0;
    for (u32 i = 0; i < n; ++i) {
        // a representative, if any, is a smaller-numbered neighbour
        u32 rep = ~0u;
        for (auto p = a.begin(i); p != a.end(i) && *p < i; ++p) {
            u32 j = *p;
            if (h[j] != h[i] || a.deg(j) != a.deg(i)) continue;
            // compare closed neighbourhoods: adj(i) U {i} vs adj(j) U {j}
            auto pi = a.begin(i), ei = a.end(i), pj = a.begin(j), ej = a.end(j);
            bool same = true;
            while (same) {
                // skip j in adj(i) and i in adj(j): both closed sets hold both
                if (pi != ei && *pi == j) ++pi;
                if (pj != ej && *pj == i) ++pj;
                if (pi == ei || pj == ej) break;
                same = *pi == *pj;
                ++pi;
                ++pj;
            }
            if (pi != ei && *pi == j) ++pi;
            if (pj != ej && *pj == i) ++pj;
            if (same && pi == ei && pj == ej) {
                rep = j;
                break;
            }
        }
        sv[i] = rep == ~0u ? nr_sv++ : sv[rep];
    }
    return sv;
}

Adj compress_graph(const Adj& a, const std::vector<u32>& sv, u32 nsv,
                   std::vector<u32>& weight) {
    const u32 n = a.n();
    weight.assign(nsv, 0);
    std::vector<u32> rep(nsv, ~0u);
    for (u32 i = 0; i < n; ++i) {
        ++weight[sv[i]];
        if (rep[sv[i]] == ~0u) rep[sv[i]] = i;
    }
    Adj c;
    c.ptr.assign(nsv + 1, 0);
    std::vector<u32> tmp;
    for (u32 s = 0; s < nsv; ++s) {
        u32 i = rep[s];
        tmp.clear();
        for (auto p = a.begin(i); p != a.end(i); ++p) {
            if (sv[*p] != s) tmp.push_back(sv[*p]);
        }
        std::sort(tmp.begin(), tmp.end());
        tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
        c.idx.insert(c.idx.end(), tmp.begin(), tmp.end());
        c.ptr[s + 1] = c.idx.size();
    }
    return c;
}

/* ---------------------------------------------------- nested dissection */

class NestedDissection {
    const Adj& g;
    const std::vector<u32>& w;
    //! optional node coordinates [ngeo][n] (hint: further bisection keys)
    const double* geo = nullptr;
    int ngeo = 0;
    std::vector<u32> region;  //!< region id of each node (current partition)
    std::vector<u32> level;   //!< scratch: BFS level
    std::vector<u32>& order;
    std::atomic<u32> next_region{1};
    //! per-task state: sub-problems on disjoint node sets run as OpenMP tasks
    struct Ctx {
        std::vector<u32> queue, order;
        u32 stamp = 1;  //!< levels are stored as stamp + l: scratch needs no reset
    };
    static constexpr u32 LEAF_WEIGHT = 36;

    //! BFS inside region r from root; fills queue (visit order) and level;
    //! returns number of levels
    u32 bfs(Ctx& ctx, u32 root, u32 r, u32 stamp_base) {
        auto& queue = ctx.queue;
        queue.clear();
        queue.push_back(root);
        level[root] = stamp_base;
        u32 nlev = 1;
        for (size_t qh = 0; qh < queue.size(); ++qh) {
            u32 v = queue[qh];
            for (auto p = g.begin(v); p != g.end(v); ++p) {
                u32 u = *p;
                if (region[u] != r || level[u] >= stamp_base) continue;
                level[u] = level[v] + 1;
                nlev = std::max(nlev, level[u] - stamp_base + 1);
                queue.push_back(u);
            }
        }
        return nlev;
    }


    std::vector<int32_t> key;  //!< scratch: bisection key of each node
    std::vector<uint8_t> mark;  //!< scratch: 1 = boundary node, 2 = in the cover
    std::vector<u32> mate, bidx;

    /*! Marks (mark[v] = 2) a minimum vertex cover of the cut edges between
     * {key <= c} and {key > c} inside region r. */
    void min_cover_separator(const std::vector<u32>& nodes, u32 r, int32_t c) {
        std::vector<u32> lb, rb;  // boundary nodes of both sides
        for (u32 v : nodes) {
            const bool in_left = key[v] <= c;
            for (auto p = g.begin(v); p != g.end(v); ++p) {
                if (region[*p] == r && (key[*p] <= c) != in_left) {
                    mark[v] = 1;
                    (in_left ? lb : rb).push_back(v);
                    break;
                }
            }
        }
        constexpr u32 NONE = ~0u;
        for (u32 v : lb) mate[v] = NONE;
        for (u32 v : rb) mate[v] = NONE;
        auto cross = [&](u32 v, u32 u) {
            return region[u] == r && mark[u] && (key[u] <= c) != (key[v] <= c);
        };
        // maximum matching by augmenting paths (iterative DFS from each free
        // left node); the boundaries have at most a few thousand nodes
        std::vector<u32> stack, visited_r;
        std::vector<std::pair<u32, const u32*>> st;
        for (u32 root : lb) {
            // greedy first
            for (auto p = g.begin(root); p != g.end(root); ++p) {
                if (cross(root, *p) && mate[*p] == NONE) {
                    mate[*p] = root;
                    mate[root] = *p;
                    break;
                }
            }
        }
        for (u32 root : lb) {
            if (mate[root] != NONE) continue;
            // DFS over alternating paths; bidx[u] = left predecessor of right u
            visited_r.clear();
            st.clear();
            st.push_back({root, g.begin(root)});
            u32 found = NONE;
            while (!st.empty() && found == NONE) {
                auto& [v, it] = st.back();
                if (it == g.end(v)) {
                    st.pop_back();
                    continue;
                }
                u32 u = *it++;
                if (!cross(v, u) || bidx[u] != NONE) continue;
                bidx[u] = v;
                visited_r.push_back(u);
                if (mate[u] == NONE) {
                    found = u;
                } else {
                    st.push_back({mate[u], g.begin(mate[u])});
                }
            }
            if (found != NONE) {
                u32 u = found;
                while (true) {
                    u32 v = bidx[u], prev = mate[v];
                    mate[u] = v;
                    mate[v] = u;
                    if (v == root) break;
                    u = prev;
                }
            }
            for (u32 u : visited_r) bidx[u] = NONE;
        }
        // Koenig: Z = nodes reachable from free left nodes by alternating
        // paths; cover = (L \ Z) + (R & Z)
        std::vector<u32> zl, zr;
        stack.clear();
        for (u32 v : lb) {
            if (mate[v] == NONE) {
                stack.push_back(v);
                mark[v] = 3;  // left node in Z
            }
        }
        while (!stack.empty()) {
            u32 v = stack.back();
            stack.pop_back();
            for (auto p = g.begin(v); p != g.end(v); ++p) {
                u32 u = *p;
                if (!cross(v, u) || mark[u] == 2) continue;
                mark[u] = 2;  // right node in Z: in the cover
                u32 m = mate[u];
                if (m != NONE && mark[m] != 3) {
                    mark[m] = 3;
                    stack.push_back(m);
                }
            }
        }
        for (u32 v : lb) mark[v] = mark[v] == 3 ? 0 : 2;  // L \ Z is in the cover
        for (u32 v : rb) {
            if (mark[v] != 2) mark[v] = 0;
        }
    }
    struct Cut {
        bool valid = false, sep_from_left = true;
        int32_t c = 0;
        u64 sepw = 0;
    };

    /*! best threshold of key[] over the nodes of region r: minimum separator
     * weight among the thresholds that leave both sides >= 30 % (then 20 %) of
     * the weight.  O(edges). */
    Cut best_cut(const std::vector<u32>& nodes, u32 r, u64 wt) {
        int32_t kmin = INT32_MAX, kmax = INT32_MIN;
        for (u32 v : nodes) {
            kmin = std::min(kmin, key[v]);
            kmax = std::max(kmax, key[v]);
        }
        Cut best;
        if (kmax - kmin < 2) return best;
        const size_t nb = (size_t)(kmax - kmin) + 2;
        std::vector<int64_t> wk(nb, 0), dl(nb, 0), dr(nb, 0);
        for (u32 v : nodes) {
            int32_t kv = key[v], mx = kv, mn = kv;
            for (auto p = g.begin(v); p != g.end(v); ++p) {
                if (region[*p] != r) continue;
                mx = std::max(mx, key[*p]);
                mn = std::min(mn, key[*p]);
            }
            wk[kv - kmin] += w[v];
            if (mx > kv) {  // v is on the left boundary for c in [kv, mx - 1]
                dl[kv - kmin] += w[v];
                dl[mx - kmin] -= w[v];
            }
            if (mn < kv) {  // v is on the right boundary for c in [mn, kv - 1]
                dr[mn - kmin] += w[v];
                dr[kv - kmin] -= w[v];
            }
        }
        for (int pass = 0; pass < 2 && !best.valid; ++pass) {
            const u64 share = pass == 0 ? 3 : 2;
            int64_t wl = 0, bl = 0, br = 0;
            for (int32_t c = kmin; c < kmax; ++c) {
                wl += wk[c - kmin];
                bl += dl[c - kmin];
                br += dr[c - kmin];
                const bool from_left = bl <= br;
                const u64 sw = (u64)(from_left ? bl : br);
                const u64 lw = (u64)(from_left ? wl - bl : wl);
                const u64 rw = (u64)(from_left ? (int64_t)wt - wl : (int64_t)wt - wl - br);
                if (std::min(lw, rw) * 10 < wt * share) continue;
                if (!best.valid || sw < best.sepw) {
                    best.valid = true;
                    best.sepw = sw;
                    best.c = c;
                    best.sep_from_left = from_left;
                }
            }
        }
        return best;
    }

    void dissect(Ctx& ctx, std::vector<u32> nodes) {
        auto& queue = ctx.queue;
        auto& order = ctx.order;
        while (true) {
            if (nodes.empty()) return;
            const u32 r = region[nodes[0]];
            u64 wt = 0;
            for (u32 v : nodes) wt += w[v];
            if (wt <= LEAF_WEIGHT || nodes.size() <= 2) {
                emit_leaf(ctx, nodes, r);
                return;
            }
            // connected component of nodes[0]
            u32 base = new_stamp(ctx, nodes.size());
            bfs(ctx, nodes[0], r, base);
            if (queue.size() < nodes.size()) {
                // disconnected: peel this component off and handle it on its own
                u32 rc = next_region++;
                std::vector<u32> comp(queue.begin(), queue.end());
                for (u32 v : comp) region[v] = rc;
                std::vector<u32> rest;
                rest.reserve(nodes.size() - comp.size());
                for (u32 v : nodes) {
                    if (region[v] == r) rest.push_back(v);
                }
                dissect(ctx, std::move(comp));
                nodes.swap(rest);
                continue;
            }
            // pseudo-peripheral root: repeat BFS from the farthest node
            u32 root = queue.back(), nlev = 0;
            for (int it = 0; it < 3; ++it) {
                base = new_stamp(ctx, nodes.size());
                u32 nl = bfs(ctx, root, r, base);
                if (nl <= nlev) break;
                nlev = nl;
                root = queue.back();
            }
            base = new_stamp(ctx, nodes.size());
            nlev = bfs(ctx, root, r, base);
            if (nlev < 3) {
                emit_leaf(ctx, nodes, r);
                return;
            }
            // Candidate bisections: thresholds of a node key.  Key 1 is the BFS
            // level from the pseudo-peripheral root s (spherical shells); key 2
            // is d_s - d_t with t the node farthest from s, whose level sets are
            // the much flatter "perpendicular bisectors" of the body.  For a
            // threshold c the sides are {key <= c} and {key > c}; the vertex
            // separator is the lighter of the two boundaries.
            const u32 t_far = queue.back();
            std::vector<u32> nodes_bfs(queue.begin(), queue.end());
            for (u32 v : nodes_bfs) key[v] = (int32_t)(level[v] - base);
            Cut cut1 = best_cut(nodes_bfs, r, wt);
            std::vector<int32_t> key1;
            if (nodes_bfs.size() > 64) {
                key1.reserve(nodes_bfs.size());
                for (u32 v : nodes_bfs) key1.push_back(key[v]);
                u32 base2 = new_stamp(ctx, nodes.size());
                bfs(ctx, t_far, r, base2);
                for (size_t q = 0; q < nodes_bfs.size(); ++q) {
                    u32 v = nodes_bfs[q];
                    key[v] = key1[q] - (int32_t)(level[v] - base2);
                }
                Cut cut2 = best_cut(nodes_bfs, r, wt);
                std::vector<int32_t> keep;  // key of the best cut so far
                auto take = [&](const Cut& cnew) {
                    if (cnew.valid && (!cut1.valid || cnew.sepw < cut1.sepw)) {
                        cut1 = cnew;
                        keep.resize(nodes_bfs.size());
                        for (size_t q = 0; q < nodes_bfs.size(); ++q) keep[q] = key[nodes_bfs[q]];
                    }
                };
                keep = key1;
                take(cut2);
                // a second axis: u = the node farthest off the s-t axis (largest
                // d_s + d_t), u' = the node farthest from u
                if (nodes_bfs.size() > 512) {
                    u32 u_off = nodes_bfs[0];
                    int64_t best_sum = -1;
                    for (size_t q = 0; q < nodes_bfs.size(); ++q) {
                        u32 v = nodes_bfs[q];
                        int64_t sum = (int64_t)key1[q] + (int64_t)(level[v] - base2);
                        if (sum > best_sum) {
                            best_sum = sum;
                            u_off = v;
                        }
                    }
                    u32 base3 = new_stamp(ctx, nodes.size());
                    bfs(ctx, u_off, r, base3);
                    const u32 u_far = queue.back();
                    std::vector<int32_t> du(nodes_bfs.size());
                    for (size_t q = 0; q < nodes_bfs.size(); ++q) {
                        du[q] = (int32_t)(level[nodes_bfs[q]] - base3);
                    }
                    u32 base4 = new_stamp(ctx, nodes.size());
                    bfs(ctx, u_far, r, base4);
                    for (size_t q = 0; q < nodes_bfs.size(); ++q) {
                        u32 v = nodes_bfs[q];
                        key[v] = du[q] - (int32_t)(level[v] - base4);
                    }
                    take(best_cut(nodes_bfs, r, wt));
                }
                // Geometric candidates (only with a coordinate hint): thresholds
                // of each coordinate, i.e. axis-aligned planes.  On lattice-like
                // bodies the distance-based keys above give diagonal cuts, whose
                // vertex separators are about twice as heavy (a cube of 41^3
                // nodes: 10.9 k rows in the top separator instead of 5.0 k, an
                // order of magnitude in factorisation flops).
                for (int ax = 0; ax < ngeo; ++ax) {
                    const double* cx = geo + (size_t)ax * g.n();
                    double lo = cx[nodes_bfs[0]], hi = lo;
                    for (u32 v : nodes_bfs) {
                        lo = std::min(lo, cx[v]);
                        hi = std::max(hi, cx[v]);
                    }
                    if (!(hi > lo)) continue;
                    // resolution: a few keys per node along the axis
                    const double q_steps = std::min(
                            8192.0, 4.0 * std::cbrt((double)nodes_bfs.size()) + 8.0);
                    const double scale = q_steps / (hi - lo);
                    for (u32 v : nodes_bfs) {
                        key[v] = (int32_t)std::floor((cx[v] - lo) * scale + 0.5);
                    }
                    take(best_cut(nodes_bfs, r, wt));
                }
                for (size_t q = 0; q < nodes_bfs.size(); ++q) key[nodes_bfs[q]] = keep[q];
            }
            if (!cut1.valid) {
                emit_leaf(ctx, nodes, r);
                return;
            }
            std::vector<u32> left, right, sep;
            u32 rl = next_region++, rr = next_region++, rs = next_region++;
            // smallest vertex separator of this edge cut: a minimum vertex cover
            // of the bipartite graph of cut edges (Koenig: from a maximum
            // matching); it may take nodes from both boundaries
            min_cover_separator(nodes_bfs, r, cut1.c);
            for (u32 v : nodes_bfs) {
                if (mark[v] == 2) {
                    sep.push_back(v);
                } else {
                    (key[v] <= cut1.c ? left : right).push_back(v);
                }
                mark[v] = 0;
            }
            for (u32 v : left) region[v] = rl;
            for (u32 v : right) region[v] = rr;
            for (u32 v : sep) region[v] = rs;
            if (left.empty() || right.empty()) {
                for (u32 v : nodes) region[v] = r;
                emit_leaf(ctx, nodes, r);
                return;
            }
            if (left.size() + right.size() > 4096) {
                // independent sub-problems (disjoint nodes, own stamps/queues)
                Ctx cl, cr;
                cl.stamp = cr.stamp = ctx.stamp;
#pragma omp task shared(cl, left) default(shared)
                dissect(cl, std::move(left));
                dissect(cr, std::move(right));
#pragma omp taskwait
                order.insert(order.end(), cl.order.begin(), cl.order.end());
                order.insert(order.end(), cr.order.begin(), cr.order.end());
            } else {
                dissect(ctx, std::move(left));
                dissect(ctx, std::move(right));
            }
            // the separator goes last
            for (u32 v : sep) order.push_back(v);
            return;
        }
    }

    u32 new_stamp(Ctx& ctx, size_t span) {
        u32 s = ctx.stamp;
        ctx.stamp += (u32)span + 2;
        return s;
    }

    //! leaves: BFS (Cuthill-McKee-like) order, reversed
    void emit_leaf(Ctx& ctx, const std::vector<u32>& nodes, u32 r) {
        auto& queue = ctx.queue;
        auto& order = ctx.order;
        size_t start = order.size();
        std::vector<u32> rest(nodes);
        while (!rest.empty()) {
            u32 base = new_stamp(ctx, nodes.size());
            // start from a minimum-degree node of what is left
            u32 root = rest[0];
            for (u32 v : rest) {
                if (g.deg(v) < g.deg(root)) root = v;
            }
            bfs(ctx, root, r, base);
            for (u32 v : queue) {
                order.push_back(v);
                region[v] = 0;  // done
            }
            std::vector<u32> nxt;
            for (u32 v : rest) {
                if (region[v] == r) nxt.push_back(v);
            }
            rest.swap(nxt);
        }
        std::reverse(order.begin() + start, order.end());
    }

public:
    NestedDissection(const Adj& g_, const std::vector<u32>& w_,
                     std::vector<u32>& order_, const double* geo_ = nullptr,
                     int ngeo_ = 0)
            : g{g_}, w{w_}, geo{geo_}, ngeo{ngeo_}, order{order_} {
        const u32 n = g.n();
        region.assign(n, 1);
        next_region = 2;
        level.assign(n, 0);
        key.assign(n, 0);
        mark.assign(n, 0);
        mate.assign(n, ~0u);
        bidx.assign(n, ~0u);
        std::vector<u32> all(n);
        std::iota(all.begin(), all.end(), 0u);
        // stamps grow by O(|region|) per BFS: total O(n log n) << 2^32 for the
        // sizes this solver targets
        Ctx root;
        root.order.reserve(n);
#pragma omp parallel num_threads(::sanm_b200::rt::host_threads())
#pragma omp single
        dissect(root, std::move(all));
        order = std::move(root.order);
        sanm_check(order.size() == n, "nested dissection lost nodes: %zu vs %u",
                   order.size(), n);
    }
};

}  // namespace

Symbolic symbolic_analyze(uint32_t n, const uint32_t* rowptr,
                          const uint32_t* colidx, const double* node_keys) {
    StageTimer stage_timer;
    Symbolic S;
    S.n = n;
    sanm_check(n > 0);
    Adj adj = build_sym_adj(n, rowptr, colidx);

    SANM_STAGE("ordering");
    // ---- ordering
    u32 nsv = 0;
    std::vector<u32> sv = supervariables(adj, nsv);
    std::vector<u32> weight;
    Adj cg = compress_graph(adj, sv, nsv, weight);
    std::vector<u32> corder;
    // Coordinate hint: the keys of the members of a 3-unknown supervariable,
    // in index order, are taken as the coordinates of that node (the unknowns
    // of `fea` are the free vertex coordinates, three per vertex, and the
    // three share one adjacency).  Nodes of any other size get the average of
    // their neighbours.  Without enough such nodes the hint is ignored.
    std::vector<double> geo;
    int ngeo = 0;
    if (node_keys && getenv("SANM_ND_NO_GEOMETRY") == nullptr) {
        constexpr double MISSING = std::numeric_limits<double>::quiet_NaN();
        geo.assign((size_t)3 * nsv, MISSING);
        std::vector<uint8_t> cnt(nsv, 0);
        u32 full = 0;
        for (u32 i = 0; i < n; ++i) {  // members in increasing index order
            const u32 s = sv[i];
            if (weight[s] == 3 && cnt[s] < 3 && std::isfinite(node_keys[i])) {
                geo[(size_t)cnt[s] * nsv + s] = node_keys[i];
                if (++cnt[s] == 3) ++full;
            }
        }
        if ((u64)full * 2 >= nsv) {
            ngeo = 3;
            std::vector<u32> todo;
            for (u32 s = 0; s < nsv; ++s) {
                if (cnt[s] != 3) {
                    for (int k = 0; k < 3; ++k) geo[(size_t)k * nsv + s] = MISSING;
                    todo.push_back(s);
                }
            }
            for (int sweep = 0; sweep < 8 && !todo.empty(); ++sweep) {
                std::vector<u32> rest;
                std::vector<double> vals;
                for (u32 s : todo) {
                    double acc[3] = {0, 0, 0};
                    u32 m = 0;
                    for (auto p = cg.begin(s); p != cg.end(s); ++p) {
                        if (std::isnan(geo[*p])) continue;
                        for (int k = 0; k < 3; ++k) acc[k] += geo[(size_t)k * nsv + *p];
                        ++m;
                    }
                    if (!m) {
                        rest.push_back(s);
                    } else {
                        vals.push_back((double)s);
                        for (int k = 0; k < 3; ++k) vals.push_back(acc[k] / m);
                    }
                }
                for (size_t q = 0; q < vals.size(); q += 4) {
                    const u32 s = (u32)vals[q];
                    for (int k = 0; k < 3; ++k) geo[(size_t)k * nsv + s] = vals[q + 1 + k];
                }
                todo.swap(rest);
            }
            for (u32 s : todo) {
                for (int k = 0; k < 3; ++k) geo[(size_t)k * nsv + s] = 0.0;
            }
        }
    }
    { NestedDissection nd{cg, weight, corder, ngeo ? geo.data() : nullptr, ngeo}; }
    std::vector<u32> cpos(nsv);
    for (u32 i = 0; i < nsv; ++i) cpos[corder[i]] = i;
    // expand: members of a supervariable stay adjacent
    std::vector<u32> perm(n);
    {
        std::vector<u32> cnt(nsv + 1, 0);
        for (u32 i = 0; i < n; ++i) ++cnt[cpos[sv[i]] + 1];
        for (u32 i = 0; i < nsv; ++i) cnt[i + 1] += cnt[i];
        for (u32 i = 0; i < n; ++i) perm[cnt[cpos[sv[i]]]++] = i;
    }
    std::vector<u32> iperm(n);
    for (u32 i = 0; i < n; ++i) iperm[perm[i]] = i;

    SANM_STAGE("elimination tree (Liu) on the permuted s");
    // ---- elimination tree (Liu) on the permuted symmetric pattern
    auto etree = [&](const std::vector<u32>& pm, const std::vector<u32>& ipm,
                     std::vector<int32_t>& parent) {
        parent.assign(n, -1);
        std::vector<int32_t> anc(n, -1);
        for (u32 i = 0; i < n; ++i) {
            u32 oi = pm[i];
            for (auto p = adj.begin(oi); p != adj.end(oi); ++p) {
                u32 k = ipm[*p];
                if (k >= i) continue;
                // walk up from k with path compression
                int32_t j = (int32_t)k;
                while (anc[j] != -1 && anc[j] != (int32_t)i) {
                    int32_t nx = anc[j];
                    anc[j] = (int32_t)i;
                    j = nx;
                }
                if (anc[j] == -1) {
                    anc[j] = (int32_t)i;
                    parent[j] = (int32_t)i;
                }
            }
        }
    };
    std::vector<int32_t> parent;
    etree(perm, iperm, parent);

    SANM_STAGE("postorder (children before parents, subt");
    // ---- postorder (children before parents, subtrees contiguous)
    {
        std::vector<u32> head(n, ~0u), next(n, ~0u);
        for (u32 i = n; i-- > 0;) {
            if (parent[i] >= 0) {
                next[i] = head[parent[i]];
                head[parent[i]] = i;
            }
        }
        std::vector<u32> post;
        post.reserve(n);
        std::vector<u32> stack;
        for (u32 r = 0; r < n; ++r) {
            if (parent[r] >= 0) continue;
            stack.push_back(r);
            while (!stack.empty()) {
                u32 v = stack.back();
                u32 c = head[v];
                if (c != ~0u) {
                    head[v] = next[c];
                    stack.push_back(c);
                } else {
                    post.push_back(v);
                    stack.pop_back();
                }
            }
        }
        sanm_check(post.size() == n);
        std::vector<u32> perm2(n);
        for (u32 i = 0; i < n; ++i) perm2[i] = perm[post[i]];
        perm.swap(perm2);
        for (u32 i = 0; i < n; ++i) iperm[perm[i]] = i;
        // a postorder only renames the nodes of the elimination tree: relabel it
        // instead of running Liu's algorithm a second time
        std::vector<u32> where(n);
        for (u32 i = 0; i < n; ++i) where[post[i]] = i;
        std::vector<int32_t> parent2(n);
        for (u32 i = 0; i < n; ++i) {
            const int32_t p = parent[post[i]];
            parent2[i] = p >= 0 ? (int32_t)where[p] : -1;
        }
        parent.swap(parent2);
    }

    SANM_STAGE("column counts of L (off-diagonal) by row");
    // ---- column counts of L (off-diagonal) by row-subtree traversal
    std::vector<u32> cc(n, 0);
    {
        // On the supervariable graph: the members of a supervariable are
        // adjacent in the (post)order and form a chain of the elimination
        // tree, so column t of a supervariable s of weight w has
        //   cc = (w - 1 - t) + sum of the weights of the later supervariables
        //        in the column structure of s,
        // and the row-subtree traversal runs over |L| / w^2 entries (3 unknowns
        // per mesh vertex: 9 times fewer; 2.2 s -> 0.25 s for a 1 M-tet cube).
        std::vector<u32> cidx(nsv, ~0u), cfirst, clast;  // order of the supervariables
        for (u32 i = 0; i < n; ++i) {
            const u32 s = sv[perm[i]];
            if (cidx[s] == ~0u) {
                cidx[s] = (u32)cfirst.size();
                cfirst.push_back(i);
                clast.push_back(i);
            } else {
                // members must be contiguous
                sanm_check(clast[cidx[s]] + 1 == i, "supervariable split by the ordering");
                clast[cidx[s]] = i;
            }
        }
        const u32 m = (u32)cfirst.size();
        std::vector<u32> sv_at(m);  // supervariable at compressed position
        for (u32 s = 0; s < nsv; ++s) sv_at[cidx[s]] = s;
        std::vector<int32_t> cparent(m, -1);
        for (u32 c = 0; c < m; ++c) {
            const int32_t pl = parent[clast[c]];
            if (pl >= 0) cparent[c] = (int32_t)cidx[sv[perm[pl]]];
        }
        std::vector<u64> ccs(m, 0);
        std::vector<u32> mark(m, ~0u);
        for (u32 i = 0; i < m; ++i) {
            mark[i] = i;
            const u32 si = sv_at[i];
            const u64 wi = weight[si];
            for (auto p = cg.begin(si); p != cg.end(si); ++p) {
                const u32 k = cidx[*p];
                if (k >= i) continue;
                u32 j = k;
                while (mark[j] != i) {
                    ccs[j] += wi;
                    mark[j] = i;
                    j = (u32)cparent[j];
                }
            }
        }
        for (u32 c = 0; c < m; ++c) {
            const u32 w = clast[c] - cfirst[c] + 1;
            for (u32 t = 0; t < w; ++t) cc[cfirst[c] + t] = (u32)((w - 1 - t) + ccs[c]);
        }
    }

    SANM_STAGE("supernodes: fundamental, then relaxed am");
    // ---- supernodes: fundamental, then relaxed amalgamation with the parent
    struct SN {
        u32 c0, nc, nr;
    };
    std::vector<SN> sns;
    // Width cap: the block Gauss-Jordan of a front costs 2 nr^2 nc flops (it
    // sweeps the whole nr x nr block per pivot), so a long fundamental chain
    // is cheaper as several fronts of at most `max_nc` pivots each --
    // 2 sum_i nr_i^2 nc_i approaches the 2/3 nr^3 of an LU as the pieces
    // shrink -- at the price of one square front per piece.  Only the top
    // separators of large 3-D problems are that wide (a 61^3-node cube: root
    // chain of 18.7 k pivots, 1.3e13 -> 5.8e12 flops).
    static const u32 max_nc = [] {
        const char* e = getenv("SANM_LU_MAX_SN");
        return e ? (u32)std::max(64, atoi(e)) : 4096u;
    }();
    for (u32 j = 0; j < n; ++j) {
        bool join = j > 0 && parent[j - 1] == (int32_t)j && cc[j - 1] == cc[j] + 1 &&
                    sns.back().nc < max_nc;
        if (join) {
            ++sns.back().nc;
        } else {
            sns.push_back({j, 1, cc[j] + 1});
        }
    }
    {
        // merge s into its parent p when s is p's last child (columns adjacent)
        // and few explicit zeros appear.  Single ascending pass; p's record is
        // updated in place so chains of merges compose.
        const u32 m = (u32)sns.size();
        std::vector<u32> col2sn(n);
        for (u32 s = 0; s < m; ++s) {
            for (u32 c = 0; c < sns[s].nc; ++c) col2sn[sns[s].c0 + c] = s;
        }
        std::vector<char> dead(m, 0);
        for (u32 s = 0; s < m; ++s) {
            SN& a = sns[s];
            int32_t pc = parent[a.c0 + a.nc - 1];
            if (pc < 0) continue;
            u32 p = col2sn[pc];
            SN& b = sns[p];
            if (a.c0 + a.nc != b.c0) continue;
            u64 nc2 = (u64)a.nc + b.nc, nr2 = (u64)a.nc + b.nr;
            if (nc2 > max_nc) continue;
            u64 before = (u64)a.nr * a.nc + (u64)b.nr * b.nc;
            u64 after = nr2 * nc2;
            u64 zeros = after - before;
            bool ok = zeros == 0 || nc2 <= 12 ||
                      (zeros * 100 <= after * 15 && nc2 <= 96) ||
                      zeros * 100 <= after * 4;
            if (!ok) continue;
            b.c0 = a.c0;
            b.nc = (u32)nc2;
            b.nr = (u32)nr2;
            dead[s] = 1;
        }
        std::vector<SN> kept;
        for (u32 s = 0; s < m; ++s) {
            if (!dead[s]) kept.push_back(sns[s]);
        }
        sns.swap(kept);
    }
    {
        // Subtree collapse: a whole subtree of the supernodal tree whose
        // columns (contiguous in the postorder) and the update rows of its root
        // fit one small dense front becomes ONE leaf supernode.  The bottom of
        // the tree (inside the nested-dissection leaves) is a forest of 1-20
        // column supernodes several levels deep; every level costs the
        // factorisation an assembly + a front launch and every solve two
        // dependent launches, while the arithmetic of those fronts is
        // negligible.  A collapsed front stays on the shared-memory path
        // (sparse_lu.cu SMALL_NR).  Explicit zeros are bounded by the size cap.
        static const u32 cap_nr = [] {
            const char* e = getenv("SANM_LU_COLLAPSE_NR");
            return e ? (u32)atoi(e) : 104u;
        }();
        static const u32 cap_nc = [] {
            const char* e = getenv("SANM_LU_COLLAPSE_NC");
            return e ? (u32)atoi(e) : 64u;
        }();
        const u32 m = (u32)sns.size();
        if (cap_nr > 0 && m > 1) {
            std::vector<u32> col2sn(n);
            for (u32 s = 0; s < m; ++s) {
                for (u32 c = 0; c < sns[s].nc; ++c) col2sn[sns[s].c0 + c] = s;
            }
            std::vector<int32_t> par(m, -1);
            std::vector<u32> first(m);
            for (u32 s = 0; s < m; ++s) {
                first[s] = sns[s].c0;
                const int32_t pc = parent[sns[s].c0 + sns[s].nc - 1];
                if (pc >= 0) par[s] = (int32_t)col2sn[pc];
            }
            // children precede parents in the postorder
            for (u32 s = 0; s < m; ++s) {
                if (par[s] >= 0) first[par[s]] = std::min(first[par[s]], first[s]);
            }
            auto fits = [&](u32 s) {
                const u64 sub_nc = (u64)sns[s].c0 + sns[s].nc - first[s];
                const u64 nr2 = sub_nc + (sns[s].nr - sns[s].nc);
                return sub_nc <= cap_nc && nr2 <= cap_nr;
            };
            // descending: a root is met before the members of its subtree,
            // which are exactly the records starting at or after its first column
            std::vector<SN> kept;
            u32 cur_first = n;
            for (u32 s = m; s-- > 0;) {
                if (sns[s].c0 >= cur_first) continue;  // inside a collapsed subtree
                if (fits(s) && first[s] < sns[s].c0) {
                    const u32 sub_nc = sns[s].c0 + sns[s].nc - first[s];
                    kept.push_back({first[s], sub_nc, sub_nc + (sns[s].nr - sns[s].nc)});
                    cur_first = first[s];
                } else {
                    kept.push_back(sns[s]);
                }
            }
            std::reverse(kept.begin(), kept.end());
            // a subtree is a contiguous column range ending at its root, so the
            // records that survive still tile [0, n) in order
            u32 at = 0;
            for (const SN& k : kept) {
                sanm_check(k.c0 == at, "subtree collapse broke the column tiling");
                at += k.nc;
            }
            sanm_check(at == n);
            sns.swap(kept);
        }
    }
    const u32 nsn = (u32)sns.size();
    S.nsn = nsn;
    S.sn_c0.resize(nsn);
    S.sn_nc.resize(nsn);
    S.sn_nr.resize(nsn);
    S.sn_parent.assign(nsn, -1);
    std::vector<u32> col2sn(n);
    for (u32 s = 0; s < nsn; ++s) {
        S.sn_c0[s] = sns[s].c0;
        S.sn_nc[s] = sns[s].nc;
        for (u32 c = 0; c < sns[s].nc; ++c) col2sn[sns[s].c0 + c] = s;
    }

    SANM_STAGE("front structures (exact, by merging chil");
    // ---- front structures (exact, by merging children)
    S.soff.assign(nsn + 1, 0);
    {
        std::vector<std::vector<u32>> st(nsn);
        std::vector<u32> mark(n, ~0u);
        std::vector<std::vector<u32>> kids(nsn);
        for (u32 s = 0; s < nsn; ++s) {
            const u32 c0 = S.sn_c0[s], nc = S.sn_nc[s];
            auto& v = st[s];
            for (u32 c = c0; c < c0 + nc; ++c) {
                mark[c] = s;
                v.push_back(c);
            }
            for (u32 c = c0; c < c0 + nc; ++c) {
                u32 oc = perm[c];
                for (auto p = adj.begin(oc); p != adj.end(oc); ++p) {
                    u32 r = iperm[*p];
                    if (r >= c0 + nc && mark[r] != s) {
                        mark[r] = s;
                        v.push_back(r);
                    }
                }
            }
            for (u32 ch : kids[s]) {
                const auto& cv = st[ch];
                for (size_t q = S.sn_nc[ch]; q < cv.size(); ++q) {
                    u32 r = cv[q];
                    if (mark[r] != s) {
                        sanm_check(r >= c0 + nc);
                        mark[r] = s;
                        v.push_back(r);
                    }
                }
            }
            std::sort(v.begin() + nc, v.end());
            S.sn_nr[s] = (u32)v.size();
            if (v.size() > nc) {
                u32 p = col2sn[v[nc]];
                S.sn_parent[s] = (int32_t)p;
                kids[p].push_back(s);
            }
        }
        for (u32 s = 0; s < nsn; ++s) S.soff[s + 1] = S.soff[s] + st[s].size();
        S.structs.resize(S.soff[nsn]);
        for (u32 s = 0; s < nsn; ++s) {
            std::copy(st[s].begin(), st[s].end(), S.structs.begin() + S.soff[s]);
        }
        S.child_ptr.assign(nsn + 1, 0);
        for (u32 s = 0; s < nsn; ++s) S.child_ptr[s + 1] = S.child_ptr[s] + (u32)kids[s].size();
        S.child_list.resize(S.child_ptr[nsn]);
        for (u32 s = 0; s < nsn; ++s) {
            std::copy(kids[s].begin(), kids[s].end(),
                      S.child_list.begin() + S.child_ptr[s]);
        }
    }

    SANM_STAGE("front storage, child -> parent inverse m");
    // ---- front storage, child -> parent inverse maps, levels
    S.foff.assign(nsn + 1, 0);
    for (u32 s = 0; s < nsn; ++s) {
        u64 nr = S.sn_nr[s], nc = S.sn_nc[s];
        // keep every front 16-byte... 8 doubles aligned
        S.foff[s + 1] = S.foff[s] + ((nr * nr + 7) / 8) * 8;
        S.flops += 2.0 * (double)nr * (double)nr * (double)nc;
        S.factor_entries += nc * nc + 2 * nc * (nr - nc);
    }
    S.front_total = S.foff[nsn];
    S.inv_off.assign(nsn + 1, 0);
    for (u32 s = 0; s < nsn; ++s) {
        u64 sz = S.sn_parent[s] >= 0 ? S.sn_nr[S.sn_parent[s]] : 0;
        S.inv_off[s + 1] = S.inv_off[s] + sz;
    }
    S.inv.assign(S.inv_off[nsn], -1);
    for (u32 s = 0; s < nsn; ++s) {
        if (S.sn_parent[s] < 0) continue;
        u32 p = (u32)S.sn_parent[s];
        const u32* ps = S.structs.data() + S.soff[p];
        const u32* cs = S.structs.data() + S.soff[s] + S.sn_nc[s];
        u32 pn = S.sn_nr[p], cn = S.sn_nr[s] - S.sn_nc[s];
        int32_t* iv = S.inv.data() + S.inv_off[s];
        u32 q = 0;
        for (u32 i = 0; i < cn; ++i) {
            while (q < pn && ps[q] < cs[i]) ++q;
            // the pivot columns of p come first and are sorted, as are its
            // update rows, and every pivot column < every update row
            sanm_check(q < pn && ps[q] == cs[i],
                       "child structure not contained in parent");
            iv[q] = (int32_t)i;
        }
    }
    {
        std::vector<u32> lev(nsn, 0);
        u32 nlev = 0;
        for (u32 s = 0; s < nsn; ++s) {
            // children precede parents in numbering
            if (S.sn_parent[s] >= 0) {
                u32 p = (u32)S.sn_parent[s];
                lev[p] = std::max(lev[p], lev[s] + 1);
            }
            nlev = std::max(nlev, lev[s] + 1);
        }
        S.level_ptr.assign(nlev + 1, 0);
        for (u32 s = 0; s < nsn; ++s) ++S.level_ptr[lev[s] + 1];
        for (u32 l = 0; l < nlev; ++l) S.level_ptr[l + 1] += S.level_ptr[l];
        S.level_list.resize(nsn);
        std::vector<u32> pos(S.level_ptr.begin(), S.level_ptr.end() - 1);
        for (u32 s = 0; s < nsn; ++s) S.level_list[pos[lev[s]]++] = s;
    }

    if (getenv("SANM_LU_DEBUG")) {
        for (u32 l = 0; l + 1 < S.level_ptr.size(); ++l) {
            u32 mx_nr = 0, mx_nc = 0, nchain = 0;
            u64 ent = 0;
            for (u32 q = S.level_ptr[l]; q < S.level_ptr[l + 1]; ++q) {
                const u32 s = S.level_list[q];
                mx_nr = std::max(mx_nr, S.sn_nr[s]);
                mx_nc = std::max(mx_nc, S.sn_nc[s]);
                ent += (u64)S.sn_nc[s] * (2 * S.sn_nr[s] - S.sn_nc[s]);
                if (S.child_ptr[s + 1] - S.child_ptr[s] == 1) ++nchain;
            }
            fprintf(stderr, "[sanm_lu] level %2u: %5u fronts (%u with one child), max nr %5u, max nc %5u, %9llu entries\n",
                    l, S.level_ptr[l + 1] - S.level_ptr[l], nchain, mx_nr, mx_nc,
                    (unsigned long long)ent);
        }
    }

    SANM_STAGE("destinations of the original entries");
    // ---- destinations of the original entries
    S.a_dest.resize(rowptr[n]);
    int bad_entry = 0;
#pragma omp parallel for num_threads(::sanm_b200::rt::host_threads()) schedule(static) reduction(| : bad_entry)
    for (int64_t i64 = 0; i64 < (int64_t)n; ++i64) {
        const u32 i = (u32)i64;
        u32 pi = iperm[i];
        for (u32 p = rowptr[i]; p < rowptr[i + 1]; ++p) {
            u32 pj = iperm[colidx[p]];
            u32 s = col2sn[std::min(pi, pj)];
            const u32* st = S.structs.data() + S.soff[s];
            const u32 nr = S.sn_nr[s], c0 = S.sn_c0[s], nc = S.sn_nc[s];
            auto local = [&](u32 g) -> u32 {
                if (g < c0 + nc) return g - c0;
                const u32* it = std::lower_bound(st + nc, st + nr, g);
                if (it == st + nr || *it != g) {
                    bad_entry = 1;
                    return 0;
                }
                return (u32)(it - st);
            };
            S.a_dest[p] = S.foff[s] + (u64)local(pj) * nr + local(pi);
        }
    }
    sanm_check(!bad_entry, "entry outside the frontal structure");
    S.perm = std::move(perm);
    S.iperm = std::move(iperm);
    return S;
}

}  // namespace sanm_b200
