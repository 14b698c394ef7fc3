// cxx/sanm_host/base.hpp -- scalar types, errors and small utilities of the
// libsanm-compatible host layer.
//
// This directory is the C++ host side ABOVE the C ABI (include/sanm_b200.h):
// classes with the reference's names and call signatures whose bodies call
// libsanm_b200.so, so that the reference application (fea/*.cpp) and its
// tests compile unmodified and run on the B200 path.  Interfaces mirrored
// here: libsanm/typedefs.h:12-21, libsanm/utils.h:19-276, libsanm/stl.h:17-79.
// The layout, storage scheme and every body are this repository's own.
#pragma once

#include <chrono>
#include <cstdarg>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <exception>
#include <functional>
#include <limits>
#include <new>
#include <optional>
#include <span>
#include <string>
#include <type_traits>
#include <utility>

namespace sanm {

using fp_t = double;          // libsanm/typedefs.h:12
using leastsize_t = uint32_t; // libsanm/typedefs.h:17

static inline constexpr fp_t operator"" _fp(long double v) {
    return static_cast<fp_t>(v);
}

// ---- errors (libsanm/utils.h:34-50); the C ABI status codes map onto these
class SANMError : public std::exception {
public:
    explicit SANMError(std::string what) : m_what(std::move(what)) {}
    const char* what() const noexcept override { return m_what.c_str(); }

private:
    std::string m_what;
};
class SANMAssertionError : public SANMError {
public:
    using SANMError::SANMError;
};
class SANMNumericalError : public SANMError {
public:
    using SANMError::SANMError;
};

std::string svsprintf(const char* fmt, va_list ap);
std::string ssprintf(const char* fmt, ...)
        __attribute__((format(printf, 1, 2)));

[[noreturn]] void __assertion_failed(const char* expr, const char* file,
                                     const char* func, int line,
                                     const char* fmt = nullptr, ...)
        __attribute__((format(printf, 5, 6)));

#define sanm_assert(cond_, ...)                                          \
    do {                                                                 \
        if (__builtin_expect(!(cond_), 0))                               \
            ::sanm::__assertion_failed(#cond_, __FILE__, __func__,       \
                                       __LINE__, ##__VA_ARGS__);         \
    } while (0)

#define SANM_PASTE2(a, b) a##b
#define SANM_PASTE(a, b) SANM_PASTE2(a, b)

class NonCopyable {
protected:
    NonCopyable() = default;
    ~NonCopyable() = default;

public:
    NonCopyable(const NonCopyable&) = delete;
    NonCopyable& operator=(const NonCopyable&) = delete;
};

// scope guard (libsanm/utils.h:28-30, 60-66)
class DeferExec : public NonCopyable {
public:
    explicit DeferExec(std::function<void()> f) : m_f(std::move(f)) {}
    ~DeferExec() {
        if (m_f)
            m_f();
    }

private:
    std::function<void()> m_f;
};
#define SANM_DEFER(func) \
    ::sanm::DeferExec SANM_PASTE(__defer, __LINE__) { func }

// ---- optional value (libsanm/utils.h:78-160), on top of std::optional
class NoneType {};
static constexpr NoneType None{};

template <typename T>
class Maybe {
public:
    Maybe() = default;
    Maybe(NoneType) {}
    Maybe(const Maybe&) = default;
    Maybe(Maybe&&) = default;
    Maybe& operator=(const Maybe&) = default;
    Maybe& operator=(Maybe&&) = default;

    template <typename A0, typename... A,
              typename = std::enable_if_t<
                      std::is_constructible_v<T, A0, A...> &&
                      !std::is_same_v<std::decay_t<A0>, Maybe> &&
                      !std::is_same_v<std::decay_t<A0>, NoneType>>>
    Maybe(A0&& a0, A&&... a) {
        m_v.emplace(std::forward<A0>(a0), std::forward<A>(a)...);
    }

    void reset() { m_v.reset(); }

    template <typename... A>
    std::enable_if_t<std::is_constructible_v<T, A...>, T&> init(A&&... a) {
        return m_v.emplace(std::forward<A>(a)...);
    }

    bool valid() const { return m_v.has_value(); }
    T& val() {
        sanm_assert(m_v.has_value());
        return *m_v;
    }
    const T& val() const {
        sanm_assert(m_v.has_value());
        return *m_v;
    }
    T* operator->() { return &val(); }
    const T* operator->() const { return &val(); }

private:
    std::optional<T> m_v;
};

template <typename T>
static inline size_t hash_combine(size_t seed, const T& v) {
    return seed ^ (std::hash<T>{}(v) + 0x9e3779b9 + (seed << 6) + (seed >> 2));
}

class HashChain {
public:
    explicit HashChain(size_t seed = 0) : m_h(seed) {}
    template <typename T>
    HashChain& feed(T&& v) {
        m_h = hash_combine(m_h, std::forward<T>(v));
        return *this;
    }
    size_t v() const { return m_h; }

private:
    size_t m_h;
};

// accumulating wall-clock timer (libsanm/utils.h:185-217)
class Timer {
    using clock = std::chrono::steady_clock;

public:
    Timer& start() {
        sanm_assert(!m_running);
        m_running = true;
        m_t0 = clock::now();
        return *this;
    }
    Timer& stop() {
        sanm_assert(m_running);
        m_running = false;
        m_acc += std::chrono::duration<double>(clock::now() - m_t0).count();
        return *this;
    }
    double time() const { return m_acc; }
    Timer& reset() {
        sanm_assert(!m_running);
        m_acc = 0;
        return *this;
    }

private:
    bool m_running = false;
    clock::time_point m_t0{};
    double m_acc = 0;
};

// Tagged host timers (libsanm/utils.h:225-249).  report() also prints the
// DEVICE time per tag that libsanm_b200.so accumulated under the reference's
// tag names (sanm_prof_dump), which is what render/gen_table_figs.py greps.
class ScopedProfiler : public NonCopyable {
public:
    explicit ScopedProfiler(const char* tag);
    ~ScopedProfiler();

    struct Stat;
    static Stat* get_node();
    static void set_thread_root_node(Stat* node);
    static void clear_thread_root_node();
    static void report();

private:
    const char* m_tag;
    std::chrono::steady_clock::time_point m_t0;
};
#define SANM_SCOPED_PROFILER(tag) \
    ::sanm::ScopedProfiler SANM_PASTE(__profiler, __LINE__) { tag }

// xorshift128+ seeded from mt19937_64 (libsanm/utils.h:253-276,
// utils.cpp:71-80); the reference's tests draw all their inputs from it
class Xorshift128pRng {
public:
    using result_type = uint64_t;
    Xorshift128pRng(uint64_t seed = 0);
    static constexpr uint64_t min() { return 0; }
    static constexpr uint64_t max() {
        return std::numeric_limits<uint64_t>::max();
    }
    uint64_t operator()() {
        uint64_t x = m_s0;
        const uint64_t y = m_s1;
        m_s0 = y;
        x ^= x << 23;
        x ^= x >> 17;
        x ^= y ^ (y >> 26);
        m_s1 = x;
        return x + y;
    }

private:
    uint64_t m_s0, m_s1;
};

// ---- libsanm/stl.h:17-79
template <typename Obj, bool ctor_append_id = false>
class ObjArray : public NonCopyable {
public:
    template <typename... Args>
    explicit ObjArray(size_t n, Args&&... args)
            : m_p(static_cast<Obj*>(::operator new[](
                      n * sizeof(Obj), std::align_val_t{alignof(Obj)}))),
              m_n(n) {
        for (; m_live < n; ++m_live) {
            if constexpr (ctor_append_id)
                new (m_p + m_live) Obj(std::forward<Args>(args)..., m_live);
            else
                new (m_p + m_live) Obj(std::forward<Args>(args)...);
        }
    }
    ~ObjArray() {
        while (m_live)
            m_p[--m_live].~Obj();
        ::operator delete[](m_p, std::align_val_t{alignof(Obj)});
    }
    size_t size() const { return m_n; }
    operator std::span<Obj>() { return {m_p, m_n}; }
    operator std::span<const Obj>() const { return {m_p, m_n}; }
    Obj& operator[](size_t i) { return m_p[i]; }
    const Obj& operator[](size_t i) const { return m_p[i]; }
    Obj* begin() { return m_p; }
    Obj* end() { return m_p + m_n; }
    const Obj* begin() const { return m_p; }
    const Obj* end() const { return m_p + m_n; }

private:
    Obj* m_p;
    size_t m_n, m_live = 0;
};

namespace details {
template <typename T>
struct ReverseAdaptor {
    T* c;
    auto begin() const { return c->rbegin(); }
    auto end() const { return c->rend(); }
};
}  // namespace details
template <typename T>
details::ReverseAdaptor<T> reverse(T& c) {
    return {&c};
}

// ---- binding helpers (not part of the reference interface)
namespace capi {
//! turn a status code of include/sanm_b200.h into the matching exception
void check(int status);
}  // namespace capi

}  // namespace sanm
