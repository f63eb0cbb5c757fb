// Stand-in for DBoW2::BowVector (reference Thirdparty/DBoW2/DBoW2/BowVector.h:20-107): a std::map from word id to
// its (normalised) weight.  Only used to build and test ORBVocabulary_b200.cc in this repository; inside the
// ORB-SLAM2 tree the real header is found first.
#ifndef __D_T_BOW_VECTOR__
#define __D_T_BOW_VECTOR__
#include <map>

namespace DBoW2 {

typedef unsigned int WordId;
typedef double WordValue;
typedef unsigned int NodeId;

class BowVector : public std::map<WordId, WordValue> {};

}  // namespace DBoW2
#endif
