// minicv: umbrella header (reference include/Frame.h:35, include/ORBmatcher.h:26 pull this).
#ifndef MINICV_OPENCV_HPP
#define MINICV_OPENCV_HPP
#include <opencv2/core/core.hpp>
#endif
