// ORACLE reference build only: stands in for include/Converter.h (which pulls Eigen and g2o); the matcher path uses
// toDescriptorVector alone (src/Converter.cc:27-35, defined in ref_matcher_wrap.cc).
#ifndef CONVERTER_H
#define CONVERTER_H
#include <opencv2/core/core.hpp>
#include <vector>
namespace ORB_SLAM2 { class Converter { public: static std::vector<cv::Mat> toDescriptorVector(const cv::Mat &Descriptors); }; }
#endif
