// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Stand-in for the vendored DBoW2 FeatureVector (the
// real header pulls in boost::serialization, absent here): the same public shape,
// std::map<NodeId, std::vector<unsigned int>>, which is all src/ORBmatcher.cc iterates.
#pragma once
#include <map>
#include <vector>

namespace DBoW2 {
typedef unsigned int NodeId;
class FeatureVector : public std::map<NodeId, std::vector<unsigned int>> {
public:
  void addFeature(NodeId id, unsigned int i_feature) { (*this)[id].push_back(i_feature); }
};
}  // namespace DBoW2
