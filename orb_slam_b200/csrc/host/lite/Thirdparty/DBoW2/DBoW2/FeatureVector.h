// Stand-in for DBoW2::FeatureVector (reference Thirdparty/DBoW2/DBoW2/FeatureVector.h:21-48): a std::map from
// vocabulary node id to the indices of the local features that descend from it.  Only used to build and test
// ORBmatcher_b200.cc in this repository (DBoW2 is an un-vendored dependency here); inside the ORB-SLAM2 tree the real
// header is found first.
#ifndef __D_T_FEATURE_VECTOR__
#define __D_T_FEATURE_VECTOR__
#include <map>
#include "BowVector.h"
#include <vector>

namespace DBoW2 {


class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {
public:
    void addFeature(NodeId id, unsigned int i_feature) { (*this)[id].push_back(i_feature); }
};

}  // namespace DBoW2
#endif
