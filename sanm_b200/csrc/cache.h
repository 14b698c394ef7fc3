/* sanm_b200/csrc/cache.h -- small LRU caches for the topology-dependent host
 * analyses: the reference rebuilds ParallelTaylorCoeffProp and
 * SparseSolver for every ANM iteration and every chained stage of a run
 * (libsanm/anm.cpp:205,217; fea/main.cpp:712-768) although the mesh and the
 * sparsity pattern never change.  Keys are 64-bit content hashes (rt.h
 * hash_bytes) of the descriptor arrays; values are immutable after creation
 * (scratch buffers inside them are only touched by kernels on the library's
 * single stream).  SANM_NO_TOPO_CACHE=1 disables all of them. */
#pragma once

#include <cstdlib>
#include <list>
#include <memory>
#include <mutex>

namespace sanm_b200 {

inline bool topo_cache_enabled() {
    const char* off = getenv("SANM_NO_TOPO_CACHE");
    return !(off && off[0] == '1');
}

template <class V>
class LruCache {
public:
    explicit LruCache(size_t cap) : m_cap{cap} {}

    std::shared_ptr<V> get(uint64_t key) {
        if (!topo_cache_enabled()) return nullptr;
        std::lock_guard<std::mutex> lk{m_mtx};
        for (auto it = m_items.begin(); it != m_items.end(); ++it) {
            if (it->first == key) {
                m_items.splice(m_items.begin(), m_items, it);
                ++m_hits;
                return m_items.front().second;
            }
        }
        return nullptr;
    }

    void put(uint64_t key, std::shared_ptr<V> v) {
        if (!topo_cache_enabled()) return;
        std::lock_guard<std::mutex> lk{m_mtx};
        m_items.emplace_front(key, std::move(v));
        while (m_items.size() > m_cap) m_items.pop_back();
    }

    size_t hits() const { return m_hits; }

private:
    size_t m_cap, m_hits = 0;
    std::mutex m_mtx;
    std::list<std::pair<uint64_t, std::shared_ptr<V>>> m_items;
};

}  // namespace sanm_b200
