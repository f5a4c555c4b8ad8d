/* Hand-written build configuration for compiling the reference's CPU sources
 * (core, imgproc, objdetect) directly with g++ -- see oracle/Makefile.
 * This replaces the header the reference's own build system would generate.
 * Everything optional (CUDA, OpenCL, IPP, TBB, image codecs, GUI) is OFF;
 * the parallel backend is selected with -DHAVE_OPENMP / -DHAVE_PTHREADS_PF
 * on the compiler command line. */
#ifndef OPENCV_CVCONFIG_H_INCLUDED
#define OPENCV_CVCONFIG_H_INCLUDED
#define BUILD_SHARED_LIBS
/* the reference's default: CV_ENABLE_INTRINSICS=ON (CMakeLists.txt:501) -> its SSE2 code paths */
#define CV_ENABLE_INTRINSICS
#define HAVE_INTTYPES_H 1
#define HAVE_PTHREAD
#define HAVE_MEMALIGN 1
#define HAVE_POSIX_MEMALIGN 1
#define HAVE_MALLOC_H 1
#endif
