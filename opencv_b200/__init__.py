"""opencv_b200 -- B200-native HOG multi-scale detector (the hot path behind tkortz/opencv's example_gpu_tbd).

The product is the C-ABI library ``libhogb200.so`` (include/hogb200.h, sources in opencv_b200/csrc/).  This
package holds the build script and a ctypes mirror of the ``cv::cuda::HOG`` interface used by tests and bench.
"""
from .hog import HOG, HogError, lib_path, load_library, read_model, write_model  # noqa: F401
from . import tbd  # noqa: F401
