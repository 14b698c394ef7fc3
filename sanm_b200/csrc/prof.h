/* sanm_b200/csrc/prof.h -- device-time profiler keyed by tag, the counterpart
 * of the reference's ScopedProfiler (libsanm/utils.h:225-249, tags listed in
 * SURVEY.md section 5).  CUDA build: a pair of events on the library stream
 * per scope, resolved lazily (no synchronisation inside the timed code).  Host
 * build: wall clock.
 */
#pragma once

#include "rt.h"

#include <string>

namespace sanm_b200 {
namespace prof {

class Scope {
    int m_slot;

public:
    explicit Scope(const char* tag);
    ~Scope();
    Scope(const Scope&) = delete;
    Scope& operator=(const Scope&) = delete;
};

//! accumulated milliseconds and number of scopes of a tag (synchronises)
void get(const char* tag, double* ms, uint64_t* calls);
void reset();
std::string dump();
//! enable / disable (disabled scopes cost nothing); default enabled
void set_enabled(bool on);

}  // namespace prof
}  // namespace sanm_b200

#define SANM_PROF_CAT2(a, b) a##b
#define SANM_PROF_CAT(a, b) SANM_PROF_CAT2(a, b)
#define SANM_PROF(tag) \
    ::sanm_b200::prof::Scope SANM_PROF_CAT(prof_scope_, __LINE__) { tag }
