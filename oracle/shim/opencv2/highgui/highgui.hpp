// TEST INFRASTRUCTURE — not part of the product.  The hot-path TUs include this header
// (NormalEstimator.h:41) but use nothing from it; image file I/O is bypassed with raw float frames.
#ifndef GSDF_ORACLE_SHIM_OPENCV_HIGHGUI_HPP_
#define GSDF_ORACLE_SHIM_OPENCV_HIGHGUI_HPP_
#include "../core/core.hpp"
#endif
