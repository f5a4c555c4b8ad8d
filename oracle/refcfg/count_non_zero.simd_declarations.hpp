#define CV_CPU_SIMD_FILENAME "count_non_zero.simd.hpp"
#define CV_CPU_DISPATCH_MODES_ALL BASELINE

#undef CV_CPU_SIMD_FILENAME
