// ORACLE shim: <opencv2/imgproc/imgproc.hpp> as named by reference src/ORBextractor.cc:60.
#ifndef ORACLE_SHIM_IMGPROC_HPP
#define ORACLE_SHIM_IMGPROC_HPP
#include "../../cvprim.h"
#endif
