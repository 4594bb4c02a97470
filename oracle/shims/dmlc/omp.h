// Stand-in for <dmlc/omp.h> (dmlc-core not vendored). Test infrastructure only.
#pragma once
#if defined(_OPENMP)
#include <omp.h>
#else
inline int omp_get_thread_num() { return 0; }
inline int omp_get_num_threads() { return 1; }
#endif
