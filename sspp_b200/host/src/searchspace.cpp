// SearchSpace (reference src/searchspace.cpp): head-inserted singly linked list of way-points,
// duplicate poses refused.  See searchspace.h for how the duplicate test is indexed.
#include "sspp/searchspace.h"

#include <cmath>

namespace SSPP {
namespace {
const double kBucket = 1e-9;  // far wider than the 1e-12 tolerance, far narrower than any real spacing
inline int64_t quant(double v) { return (int64_t)std::floor(v / kBucket); }
inline uint64_t mix(uint64_t h, uint64_t v) {
  h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
  return h;
}
}  // namespace

SearchSpace::~SearchSpace() { freeSearchSpace(); }

uint64_t SearchSpace::bucket(const geometry_msgs::Pose& p, const int64_t shift[7]) {
  const double c[7] = {p.position.x, p.position.y, p.position.z, p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w};
  uint64_t h = 0x243F6A8885A308D3ull;
  for (int i = 0; i < 7; ++i) h = mix(h, (uint64_t)(quant(c[i]) + shift[i]));
  return h;
}

// Two values within 1e-12 fall in the same 1e-9 bucket or in adjacent ones; only components that
// sit within 1e-12 of a bucket edge need the neighbour probed, so the 3^7 worst case never happens
// in practice (typically one probe).
SearchSpaceNode* SearchSpace::lookup(const geometry_msgs::Pose& p) const {
  if (index_.empty()) return nullptr;
  const double c[7] = {p.position.x, p.position.y, p.position.z, p.orientation.x, p.orientation.y, p.orientation.z, p.orientation.w};
  int lo[7], hi[7];
  for (int i = 0; i < 7; ++i) {
    int64_t q = quant(c[i]);
    lo[i] = (quant(c[i] - 1E-12) < q) ? -1 : 0;
    hi[i] = (quant(c[i] + 1E-12) > q) ? 1 : 0;
  }
  int64_t shift[7];
  SearchSpaceNode* best = nullptr;
  // odometer over the (usually single) combination of neighbour buckets
  for (int i = 0; i < 7; ++i) shift[i] = lo[i];
  for (;;) {
    auto range = index_.equal_range(bucket(p, shift));
    for (auto it = range.first; it != range.second; ++it)
      if (samePosition(p, it->second->location)) {
        // the reference returns the first match in list order = the most recently inserted one
        if (!best || it->second->id_order_ > best->id_order_) best = it->second;
      }
    int k = 0;
    while (k < 7 && shift[k] == hi[k]) { shift[k] = lo[k]; ++k; }
    if (k == 7) break;
    ++shift[k];
  }
  return best;
}

void SearchSpace::index_add(SearchSpaceNode* n) {
  const int64_t zero[7] = {0, 0, 0, 0, 0, 0, 0};
  index_.emplace(bucket(n->location, zero), n);
}

void SearchSpace::index_remove(SearchSpaceNode* n) {
  const int64_t zero[7] = {0, 0, 0, 0, 0, 0, 0};
  auto range = index_.equal_range(bucket(n->location, zero));
  for (auto it = range.first; it != range.second; ++it)
    if (it->second == n) { index_.erase(it); return; }
}

SearchSpaceNode* SearchSpace::insertNode(const geometry_msgs::Pose& nodePose) {
  geometry_msgs::PoseArray sensors;
  sensors.poses.push_back(geometry_msgs::Pose());  // one all-zero sensor pose, as the reference adds
  return insertNode(nodePose, sensors, idCount++);
}
SearchSpaceNode* SearchSpace::insertNode(const geometry_msgs::Pose& nodePose, int id) {
  geometry_msgs::PoseArray sensors;
  sensors.poses.push_back(geometry_msgs::Pose());
  return insertNode(nodePose, sensors, id);
}
SearchSpaceNode* SearchSpace::insertNode(const geometry_msgs::Pose& nodePose, const geometry_msgs::PoseArray& sensorPoses) {
  return insertNode(nodePose, sensorPoses, idCount++);  // the id advances even when the pose is a duplicate
}
SearchSpaceNode* SearchSpace::insertNode(const geometry_msgs::Pose& nodePose, const geometry_msgs::PoseArray& sensorPoses, int id) {
  SearchSpaceNode* found = lookup(nodePose);
  if (found) return found;
  SearchSpaceNode* n = new SearchSpaceNode;
  n->location = nodePose;
  n->sensorLocation.poses = sensorPoses.poses;
  n->type = RegGridNode;
  n->id = id;
  n->id_order_ = ++order_counter_;
  n->next = searchspace;  // head insertion: the list runs from the newest node to the oldest
  searchspace = n;
  index_add(n);
  ++count_;
  flat_valid_ = false;
  return n;
}

SearchSpaceNode* SearchSpace::insertTempSearchSpace(const geometry_msgs::PoseArray& robotPoses,
                                                    const std::vector<geometry_msgs::PoseArray>& sensorPoses) {
  SearchSpaceNode* space = nullptr;
  for (size_t i = 0; i < robotPoses.poses.size(); ++i) {
    if (nodeExists(space, robotPoses.poses[i])) continue;
    SearchSpaceNode* n = new SearchSpaceNode;
    n->location = robotPoses.poses[i];
    for (size_t j = 0; j < sensorPoses.size(); ++j) n->sensorLocation.poses.push_back(sensorPoses[j].poses[i]);
    n->type = RegGridNode;
    n->next = space;
    space = n;
  }
  return space;
}

SearchSpaceNode* SearchSpace::nodeExists(SearchSpaceNode* space, const geometry_msgs::Pose& nodePose) {
  for (SearchSpaceNode* t = space; t; t = t->next)
    if (samePosition(nodePose, t->location)) return t;
  return nullptr;
}
SearchSpaceNode* SearchSpace::nodeExists(const geometry_msgs::Pose& nodePose) { return lookup(nodePose); }

bool SearchSpace::removeNode(const geometry_msgs::Pose& nodePose) {
  SearchSpaceNode* victim = lookup(nodePose);
  if (!victim) return false;
  if (searchspace == victim) {
    searchspace = victim->next;
  } else {
    SearchSpaceNode* t = searchspace;
    while (t && t->next != victim) t = t->next;
    if (!t) return false;
    t->next = victim->next;
  }
  index_remove(victim);
  delete victim;
  --count_;
  flat_valid_ = false;
  return true;
}

void SearchSpace::freeSearchSpace() {
  while (searchspace) {
    SearchSpaceNode* t = searchspace;
    searchspace = searchspace->next;
    delete t;
  }
  index_.clear();
  flat_.clear();
  flat_valid_ = false;
  count_ = 0;
}

void SearchSpace::freeTempSearchSpace(SearchSpaceNode* space) {
  while (space) {
    SearchSpaceNode* t = space;
    space = space->next;
    delete t;
  }
}

const std::vector<SearchSpaceNode*>& SearchSpace::flatten() {
  if (!flat_valid_) {
    flat_.clear();
    flat_.reserve(count_);
    int i = 0;
    for (SearchSpaceNode* t = searchspace; t; t = t->next) {
      t->index = i++;
      flat_.push_back(t);
    }
    flat_valid_ = true;
  }
  return flat_;
}
}  // namespace SSPP
