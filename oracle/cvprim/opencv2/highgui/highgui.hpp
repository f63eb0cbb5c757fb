// ORACLE shim: <opencv2/highgui/highgui.hpp> as named by reference src/ORBextractor.cc:58 (nothing used).
#ifndef ORACLE_SHIM_HIGHGUI_HPP
#define ORACLE_SHIM_HIGHGUI_HPP
#include <opencv2/core/core.hpp>
#endif
