// minicv: legacy umbrella header named by the reference (include/ORBextractor.h:26).
#ifndef MINICV_LEGACY_CV_H
#define MINICV_LEGACY_CV_H
#include <opencv2/core/core.hpp>
#endif
