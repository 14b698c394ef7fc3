/* sanm_b200/csrc/prof.cu -- see prof.h. */
#include "prof.h"

#include <chrono>
#include <map>
#include <vector>

namespace sanm_b200 {
namespace prof {

namespace {
struct Pending {
#ifdef SANM_HOSTEMU
    std::chrono::steady_clock::time_point t0, t1;
#else
    cudaEvent_t e0 = nullptr, e1 = nullptr;
#endif
    int tag = -1;
    bool open = false;
};
struct Tag {
    std::string name;
    double ms = 0;
    uint64_t calls = 0;
};
std::vector<Tag> g_tags;
std::map<std::string, int> g_tag_idx;
std::vector<Pending> g_pending;
std::vector<int> g_free;
bool g_enabled = true;

int tag_id(const char* name) {
    auto it = g_tag_idx.find(name);
    if (it != g_tag_idx.end()) return it->second;
    int id = (int)g_tags.size();
    g_tags.push_back({name, 0, 0});
    g_tag_idx[name] = id;
    return id;
}

void resolve() {
#ifndef SANM_HOSTEMU
    bool any = false;
    for (auto& p : g_pending) any = any || (p.tag >= 0 && !p.open);
    if (any) rt::sync();
#endif
    for (size_t i = 0; i < g_pending.size(); ++i) {
        auto& p = g_pending[i];
        if (p.tag < 0 || p.open) continue;
        double ms;
#ifdef SANM_HOSTEMU
        ms = std::chrono::duration<double, std::milli>(p.t1 - p.t0).count();
#else
        float f = 0;
        cudaEventElapsedTime(&f, p.e0, p.e1);
        ms = f;
#endif
        g_tags[p.tag].ms += ms;
        g_tags[p.tag].calls += 1;
        p.tag = -1;
        g_free.push_back((int)i);
    }
}
}  // namespace

Scope::Scope(const char* tag) : m_slot{-1} {
    if (!g_enabled) return;
    if (g_free.empty()) {
        if (g_pending.size() >= 4096) resolve();
        if (g_free.empty()) {
            g_pending.emplace_back();
#ifndef SANM_HOSTEMU
            cudaEventCreate(&g_pending.back().e0);
            cudaEventCreate(&g_pending.back().e1);
#endif
            g_free.push_back((int)g_pending.size() - 1);
        }
    }
    m_slot = g_free.back();
    g_free.pop_back();
    auto& p = g_pending[m_slot];
    p.tag = tag_id(tag);
    p.open = true;
#ifdef SANM_HOSTEMU
    p.t0 = std::chrono::steady_clock::now();
#else
    cudaEventRecord(p.e0, rt::stream());
#endif
}

Scope::~Scope() {
    if (m_slot < 0) return;
    auto& p = g_pending[m_slot];
#ifdef SANM_HOSTEMU
    p.t1 = std::chrono::steady_clock::now();
#else
    cudaEventRecord(p.e1, rt::stream());
#endif
    p.open = false;
}

void get(const char* tag, double* ms, uint64_t* calls) {
    resolve();
    auto it = g_tag_idx.find(tag);
    if (it == g_tag_idx.end()) {
        *ms = 0;
        *calls = 0;
        return;
    }
    *ms = g_tags[it->second].ms;
    *calls = g_tags[it->second].calls;
}

void reset() {
    resolve();
    for (auto& t : g_tags) {
        t.ms = 0;
        t.calls = 0;
    }
}

std::string dump() {
    resolve();
    std::string s = "=========== sanm_b200 device-time profile ===========\n";
    for (auto& t : g_tags) {
        s += ssprintf("%-24s: tot=%.3fms nr=%zu avg=%.4fms\n", t.name.c_str(),
                      t.ms, (size_t)t.calls, t.calls ? t.ms / t.calls : 0.0);
    }
    return s;
}

void set_enabled(bool on) { g_enabled = on; }

}  // namespace prof
}  // namespace sanm_b200
