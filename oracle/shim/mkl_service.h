/* oracle/shim/mkl_service.h -- TEST INFRASTRUCTURE (oracle build only).
 * Intel MKL is not installed in this image; the reference only needs two
 * thread-count setters from this header (libsanm/tensor.cpp:29,
 * libsanm/sparse_solver.cpp:33,36).  Both are no-ops here. */
#pragma once
static inline void mkl_set_num_threads(int) {}
static inline int mkl_set_num_threads_local(int) { return 0; }
