// SearchSpace — the node store (reference include/sspp/searchspace.h, src/searchspace.cpp): a singly
// linked list with HEAD insertion (so list order is reverse insertion order) that refuses a pose
// already present within 1e-12 in all seven components.  The reference finds duplicates by scanning
// the list (O(N) per insert, O(N^2) per grid); here a hash of the quantised pose answers in O(1)
// while giving the same verdict: buckets are 1e-9 wide, so two poses within 1e-12 share a bucket or
// sit in neighbouring ones, and the exact test runs on those candidates only.
#pragma once
#include <cstdint>
#include <unordered_map>
#include <vector>
#include "searchspacenode.h"
#include "utils.h"

namespace SSPP {
class SearchSpace {
 public:
  SearchSpaceNode* searchspace = nullptr;  // list head
  int idCount = 0;

  SearchSpace() {}
  virtual ~SearchSpace();
  SearchSpaceNode* insertNode(const geometry_msgs::Pose& nodePose);
  SearchSpaceNode* insertNode(const geometry_msgs::Pose& nodePose, int id);
  SearchSpaceNode* insertNode(const geometry_msgs::Pose& nodePose, const geometry_msgs::PoseArray& sensorPoses);
  SearchSpaceNode* insertNode(const geometry_msgs::Pose& nodePose, const geometry_msgs::PoseArray& sensorPoses, int id);
  SearchSpaceNode* insertTempSearchSpace(const geometry_msgs::PoseArray& robotPoses,
                                         const std::vector<geometry_msgs::PoseArray>& sensorPoses);
  SearchSpaceNode* nodeExists(const geometry_msgs::Pose& nodePose);
  SearchSpaceNode* nodeExists(SearchSpaceNode* space, const geometry_msgs::Pose& nodePose);
  bool removeNode(const geometry_msgs::Pose& nodePose);
  void freeSearchSpace();
  void freeTempSearchSpace(SearchSpaceNode* space);

  // list order as a vector (head first), with SearchSpaceNode::index set; rebuilt lazily
  const std::vector<SearchSpaceNode*>& flatten();
  size_t size() const { return count_; }

 private:
  typedef std::unordered_multimap<uint64_t, SearchSpaceNode*> Index;
  Index index_;
  std::vector<SearchSpaceNode*> flat_;
  bool flat_valid_ = false;
  size_t count_ = 0;
  long order_counter_ = 0;
  static uint64_t bucket(const geometry_msgs::Pose& p, const int64_t shift[7]);
  SearchSpaceNode* lookup(const geometry_msgs::Pose& p) const;
  void index_add(SearchSpaceNode* n);
  void index_remove(SearchSpaceNode* n);
};
}  // namespace SSPP
