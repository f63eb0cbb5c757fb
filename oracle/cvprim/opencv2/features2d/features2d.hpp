// ORACLE shim: <opencv2/features2d/features2d.hpp> as named by reference src/ORBextractor.cc:59.
#ifndef ORACLE_SHIM_FEATURES2D_HPP
#define ORACLE_SHIM_FEATURES2D_HPP
#include "../../cvprim.h"
#endif
