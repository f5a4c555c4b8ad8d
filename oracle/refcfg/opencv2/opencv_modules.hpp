/* Hand-written module list for the oracle build (oracle/Makefile). */
#define HAVE_OPENCV_CORE
#define HAVE_OPENCV_IMGPROC
#define HAVE_OPENCV_OBJDETECT
