/* OpenCL is disabled in the oracle build: no embedded kernels, only the headers the
 * reference's sources expect this (normally generated) file to pull in. */
#include "opencv2/core/ocl.hpp"
#include "opencv2/core/ocl_genbase.hpp"
#include "opencv2/core/opencl/ocl_defs.hpp"
