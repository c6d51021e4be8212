// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing under sspp_b200/ may include, link or call this.
//
// CPU restatement of the third-party arithmetic the sspp hot path executes but does not contain:
//   * OctoMap library (octomap::OcTree):   coordToKey / keyToCoord / computeRayKeys / search /
//                                           isNodeOccupied / castRay / setNodeValue / prune
//   * ethz-asl/volumetric_mapping octomap_world (volumetric_mapping::OctomapWorld):
//                                           getLineStatus / getLineStatusBoundingBox /
//                                           getCellStatusPoint / getCellStatusBoundingBox /
//                                           setFree / setOccupied / getAllOccupiedBoxes / getMapSize
// Neither library is vendored in /root/reference nor pinned to a version there
// (package.xml:26-27,38; CMakeLists.txt:30,36).  The only call sites are
//   src/distance_heuristic.cpp:108-111  (getLineStatusBoundingBox, CellStatus::kFree)
//   src/reactive_planner_server.cpp:42  (setFree), :60 (getAllOccupiedBoxes), :61,63 (getMapSize)
// PARITY UNPINNED at this boundary: the reference holds no golden vector, fixture or test for
// any map query.  What follows is the published algorithm of those
// libraries written out again (Amanatides & Woo DDA with octomap's float/double operand mix),
// with each function naming the upstream routine it follows.  It keeps the upstream data
// structure too (a 16-level pointer octree with pruned leaves) so that timing it is timing the
// reference's algorithm, not a flattened shortcut.
//
// Build: g++ -std=c++11 -O3 -frounding-math -ffp-contract=off   (flags of CMakeLists.txt:6-8;
// contraction must be off so x86 and the sm_100a kernels round identically).
#ifndef SSPP_ORACLE_OCTOMAP_RESTATED_HPP
#define SSPP_ORACLE_OCTOMAP_RESTATED_HPP

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <set>
#include <utility>
#include <vector>

namespace orc {

typedef uint16_t key_type;

struct OcTreeKey {
  key_type k[3];
  OcTreeKey() { k[0] = k[1] = k[2] = 0; }
  OcTreeKey(key_type a, key_type b, key_type c) { k[0] = a; k[1] = b; k[2] = c; }
  bool operator==(const OcTreeKey& o) const { return k[0] == o.k[0] && k[1] == o.k[1] && k[2] == o.k[2]; }
  bool operator!=(const OcTreeKey& o) const { return !(*this == o); }
  key_type& operator[](unsigned i) { return k[i]; }
  const key_type& operator[](unsigned i) const { return k[i]; }
};

// octomath::Vector3 — three FLOATS (this is where the double poses of the planner get narrowed).
struct point3d {
  float d[3];
  point3d() { d[0] = d[1] = d[2] = 0.f; }
  point3d(float x, float y, float z) { d[0] = x; d[1] = y; d[2] = z; }
  float& operator()(unsigned i) { return d[i]; }
  const float& operator()(unsigned i) const { return d[i]; }
  float x() const { return d[0]; }
  float y() const { return d[1]; }
  float z() const { return d[2]; }
  point3d operator-(const point3d& o) const { return point3d(d[0] - o.d[0], d[1] - o.d[1], d[2] - o.d[2]); }
  point3d operator+(const point3d& o) const { return point3d(d[0] + o.d[0], d[1] + o.d[1], d[2] + o.d[2]); }
  // Vector3::norm_sq(): float products and float sums, widened on return; norm(): double sqrt.
  double norm_sq() const { return (d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); }
  double norm() const { return std::sqrt(norm_sq()); }
  void operator/=(float x) { d[0] /= x; d[1] /= x; d[2] /= x; }
  // Vector3::normalized(): divide by (float)norm() when the norm is positive.
  point3d normalized() const {
    point3d r(*this);
    double len = norm();
    if (len > 0) r /= (float)len;
    return r;
  }
};

struct OcTreeNode {
  float value;            // log-odds
  OcTreeNode** children;  // NULL, or an array of 8 (entries may be NULL)
  OcTreeNode() : value(0.f), children(NULL) {}
};

inline float logodds(double p) { return (float)std::log(p / (1 - p)); }

// octomap::OcTree restated (OcTreeBaseImpl + OccupancyOcTreeBase).
class OcTree {
 public:
  explicit OcTree(double res)
      : resolution(res), resolution_factor(1.0 / res), tree_depth(16), tree_max_val(32768), root(NULL),
        n_nodes(0) {
    // octomap defaults (AbstractOccupancyOcTree ctor); octomap_world overrides them, see OctomapWorld.
    occ_prob_thres_log = logodds(0.5);
    clamping_thres_min = logodds(0.1192);
    clamping_thres_max = logodds(0.971);
    prob_hit_log = logodds(0.7);
    prob_miss_log = logodds(0.4);
  }
  ~OcTree() { clear(); }

  void clear() {
    if (root) { deleteRecurs(root); root = NULL; }
    n_nodes = 0;
  }

  // OcTreeBaseImpl::coordToKey(double): ((int) floor(resolution_factor * coordinate)) + tree_max_val
  inline key_type coordToKey(double coordinate) const {
    return (key_type)(((int)std::floor(resolution_factor * coordinate)) + (int)tree_max_val);
  }
  // OcTreeBaseImpl::coordToKeyChecked(double, key_type&)
  inline bool coordToKeyChecked(double coordinate, key_type& keyval) const {
    int scaled = ((int)std::floor(resolution_factor * coordinate)) + (int)tree_max_val;
    if (scaled >= 0 && ((unsigned)scaled) < (2 * tree_max_val)) {
      keyval = (key_type)scaled;
      return true;
    }
    return false;
  }
  inline bool coordToKeyChecked(const point3d& p, OcTreeKey& key) const {
    for (unsigned i = 0; i < 3; ++i)
      if (!coordToKeyChecked((double)p(i), key[i])) return false;
    return true;
  }
  // OcTreeBaseImpl::keyToCoord(key_type): (double((int)key - (int)tree_max_val) + 0.5) * resolution
  inline double keyToCoord(key_type key) const {
    return (double((int)key - (int)tree_max_val) + 0.5) * resolution;
  }
  inline point3d keyToCoord(const OcTreeKey& key) const {
    return point3d(float(keyToCoord(key[0])), float(keyToCoord(key[1])), float(keyToCoord(key[2])));
  }

  static inline unsigned computeChildIdx(const OcTreeKey& key, int depth) {
    unsigned pos = 0;
    if (key.k[0] & (1 << depth)) pos += 1;
    if (key.k[1] & (1 << depth)) pos += 2;
    if (key.k[2] & (1 << depth)) pos += 4;
    return pos;
  }

  // OcTreeBaseImpl::search(key, depth = 0): walk down 16 levels; a childless node met on the way
  // is a pruned leaf standing for its whole cube; a missing child under a node that has other
  // children means "never observed" (NULL).
  const OcTreeNode* search(const OcTreeKey& key) const {
    if (root == NULL) return NULL;
    const OcTreeNode* cur = root;
    for (int i = (int)tree_depth - 1; i >= 0; --i) {
      unsigned pos = computeChildIdx(key, i);
      if (cur->children != NULL && cur->children[pos] != NULL) {
        cur = cur->children[pos];
      } else {
        if (cur->children == NULL || !hasChildren(cur)) return cur;
        return NULL;
      }
    }
    return cur;
  }
  // search(x,y,z): key from DOUBLE coordinates (OcTreeBaseImpl::search(double,double,double)).
  const OcTreeNode* search(double x, double y, double z) const {
    OcTreeKey key;
    if (!coordToKeyChecked(x, key[0]) || !coordToKeyChecked(y, key[1]) || !coordToKeyChecked(z, key[2]))
      return NULL;
    return search(key);
  }

  inline bool isNodeOccupied(const OcTreeNode* n) const { return n->value >= occ_prob_thres_log; }

  // OccupancyOcTreeBase::setNodeValue(key, log_odds, lazy_eval) with clamping.
  OcTreeNode* setNodeValue(const OcTreeKey& key, float log_odds_value, bool lazy_eval) {
    log_odds_value = std::min(std::max(log_odds_value, clamping_thres_min), clamping_thres_max);
    bool createdRoot = false;
    if (root == NULL) { root = newNode(); createdRoot = true; }
    return setNodeValueRecurs(root, createdRoot, key, 0, log_odds_value, lazy_eval);
  }
  // setNodeValue(point3d): coordToKeyChecked on the FLOAT point, silently ignored out of range.
  OcTreeNode* setNodeValue(const point3d& p, float log_odds_value, bool lazy_eval) {
    OcTreeKey key;
    if (!coordToKeyChecked(p, key)) return NULL;
    return setNodeValue(key, log_odds_value, lazy_eval);
  }

  // OccupancyOcTreeBase::updateNode(key, occupied, lazy_eval = false): add the hit / miss log-odds
  // to the leaf (created at 0 when missing; a pruned ancestor is expanded first) and clamp.  The
  // early-out of upstream (leaf already saturated in the direction of the update) changes nothing.
  OcTreeNode* updateNode(const OcTreeKey& key, bool occupied) {
    const float update = occupied ? prob_hit_log : prob_miss_log;
    const OcTreeNode* leaf = search(key);
    if (leaf && ((update >= 0 && leaf->value >= clamping_thres_max) || (update <= 0 && leaf->value <= clamping_thres_min)))
      return const_cast<OcTreeNode*>(leaf);
    bool createdRoot = false;
    if (root == NULL) { root = newNode(); createdRoot = true; }
    return updateNodeRecurs(root, createdRoot, key, 0, update);
  }

  // OccupancyOcTreeBase::updateInnerOccupancy(): inner value = max over children.
  void updateInnerOccupancy() { if (root) updateInnerRecurs(root, 0); }
  // OcTreeBaseImpl::prune(): collapse nodes whose 8 children are childless and equal-valued.
  void prune() {
    if (!root) return;
    for (unsigned depth = tree_depth - 1; depth > 0; --depth) {
      unsigned pruned = 0;
      pruneRecurs(root, 0, depth, pruned);
      if (pruned == 0) break;
    }
  }

  // OcTreeBaseImpl::computeRayKeys — Amanatides & Woo 3-D DDA, start voxel included, end voxel
  // excluded.  Operand widths are the upstream ones: direction/length FLOAT, tMax/tDelta DOUBLE,
  // the half-voxel border shift narrowed through FLOAT.  Returns false (and an empty ray) when an
  // endpoint is outside the 16-bit key range.
  bool computeRayKeys(const point3d& origin, const point3d& end, std::vector<OcTreeKey>& ray) const {
    ray.clear();
    OcTreeKey key_origin, key_end;
    if (!coordToKeyChecked(origin, key_origin) || !coordToKeyChecked(end, key_end)) return false;
    if (key_origin == key_end) return true;  // same cell: nothing is tested
    ray.push_back(key_origin);

    point3d direction = (end - origin);
    float length = (float)direction.norm();
    direction /= length;

    int step[3];
    double tMax[3];
    double tDelta[3];
    OcTreeKey current_key = key_origin;
    for (unsigned i = 0; i < 3; ++i) {
      if (direction(i) > 0.0) step[i] = 1;
      else if (direction(i) < 0.0) step[i] = -1;
      else step[i] = 0;
      if (step[i] != 0) {
        double voxelBorder = keyToCoord(current_key[i]);
        voxelBorder += (float)(step[i] * resolution * 0.5);
        tMax[i] = (voxelBorder - origin(i)) / direction(i);
        tDelta[i] = resolution / std::fabs(direction(i));  // fabs(float) -> float, then double divide
      } else {
        tMax[i] = std::numeric_limits<double>::max();
        tDelta[i] = std::numeric_limits<double>::max();
      }
    }
    bool done = false;
    while (!done) {
      unsigned dim;
      if (tMax[0] < tMax[1]) {
        if (tMax[0] < tMax[2]) dim = 0; else dim = 2;
      } else {
        if (tMax[1] < tMax[2]) dim = 1; else dim = 2;
      }
      current_key[dim] = (key_type)(current_key[dim] + step[dim]);
      tMax[dim] += tDelta[dim];
      if (current_key == key_end) {
        done = true;
        break;
      } else {
        double dist_from_origin = std::min(std::min(tMax[0], tMax[1]), tMax[2]);
        if (dist_from_origin > length) {
          done = true;
          break;
        } else {
          ray.push_back(current_key);
        }
      }
      if (ray.size() >= 100000 - 1) break;  // KeyRay capacity (upstream asserts here)
    }
    return true;
  }

  // OccupancyOcTreeBase::castRay(origin, direction, end, ignoreUnknownCells, maxRange).
  // Differences from computeRayKeys that matter bit-for-bit: direction = normalized() of the
  // caller's direction; the half-voxel shift is added in DOUBLE; the start voxel is tested first;
  // range test on float voxel centres accumulated in double.
  bool castRay(const point3d& origin, const point3d& directionP, point3d& end, bool ignoreUnknown,
               double maxRange, OcTreeKey* hit_key = NULL, long* steps = NULL) const {
    OcTreeKey current_key;
    if (!coordToKeyChecked(origin, current_key)) return false;
    const OcTreeNode* startingNode = search(current_key);
    if (steps) ++*steps;
    if (startingNode) {
      if (isNodeOccupied(startingNode)) {
        end = keyToCoord(current_key);
        if (hit_key) *hit_key = current_key;
        return true;
      }
    } else if (!ignoreUnknown) {
      end = keyToCoord(current_key);
      return false;
    }
    point3d direction = directionP.normalized();
    bool max_range_set = (maxRange > 0.0);
    int step[3];
    double tMax[3];
    double tDelta[3];
    for (unsigned i = 0; i < 3; ++i) {
      if (direction(i) > 0.0) step[i] = 1;
      else if (direction(i) < 0.0) step[i] = -1;
      else step[i] = 0;
      if (step[i] != 0) {
        double voxelBorder = keyToCoord(current_key[i]);
        voxelBorder += double(step[i] * resolution * 0.5);
        tMax[i] = (voxelBorder - origin(i)) / direction(i);
        tDelta[i] = resolution / std::fabs(direction(i));
      } else {
        tMax[i] = std::numeric_limits<double>::max();
        tDelta[i] = std::numeric_limits<double>::max();
      }
    }
    if (step[0] == 0 && step[1] == 0 && step[2] == 0) return false;
    double maxrange_sq = maxRange * maxRange;
    bool done = false;
    while (!done) {
      unsigned dim;
      if (tMax[0] < tMax[1]) {
        if (tMax[0] < tMax[2]) dim = 0; else dim = 2;
      } else {
        if (tMax[1] < tMax[2]) dim = 1; else dim = 2;
      }
      if ((step[dim] < 0 && current_key[dim] == 0) ||
          (step[dim] > 0 && current_key[dim] == 2 * tree_max_val - 1)) {
        end = keyToCoord(current_key);
        return false;
      }
      current_key[dim] = (key_type)(current_key[dim] + step[dim]);
      tMax[dim] += tDelta[dim];
      end = keyToCoord(current_key);
      if (max_range_set) {
        double dist_from_origin_sq(0.0);
        for (unsigned j = 0; j < 3; j++)
          dist_from_origin_sq += ((end(j) - origin(j)) * (end(j) - origin(j)));  // float product
        if (dist_from_origin_sq > maxrange_sq) return false;
      }
      const OcTreeNode* currentNode = search(current_key);
      if (steps) ++*steps;
      if (currentNode) {
        if (isNodeOccupied(currentNode)) {
          done = true;
          break;
        }
      } else if (!ignoreUnknown) {
        return false;
      }
    }
    if (hit_key) *hit_key = current_key;
    return true;
  }

  // Leaf iteration (octomap leaf_iterator): visits every childless node with its key-space
  // origin and edge length in voxels.
  template <class F>
  void forEachLeaf(F f) const {
    if (root) leafRecurs(root, 0, 0, 0, 0, f);
  }

  // Bulk construction used by the synthetic scenes: builds, bottom-up, exactly the tree that
  // inserting every known voxel with lazy_eval followed by prune() yields.  `known`/`occ` are
  // x-fastest dense bit arrays over keys [kmin, kmin+dims).
  void buildFromDense(const int kmin[3], const int dims[3], const uint32_t* occ_bits, const uint32_t* known_bits,
                      float occ_value, float free_value, bool do_prune) {
    clear();
    Dense dn; dn.kmin = kmin; dn.dims = dims; dn.occ = occ_bits; dn.known = known_bits;
    dn.wpr = (dims[0] + 31) / 32;
    dn.occ_value = std::min(std::max(occ_value, clamping_thres_min), clamping_thres_max);
    dn.free_value = std::min(std::max(free_value, clamping_thres_min), clamping_thres_max);
    dn.prune = do_prune;
    root = buildRecurs(dn, 0, 0, 0, 0);
    if (root && do_prune == false) updateInnerOccupancy();
  }

  size_t size() const { return n_nodes; }

  double resolution;
  double resolution_factor;
  unsigned tree_depth;
  unsigned tree_max_val;
  float occ_prob_thres_log, clamping_thres_min, clamping_thres_max, prob_hit_log, prob_miss_log;

 private:
  OcTreeNode* root;
  size_t n_nodes;

  struct Dense {
    const int* kmin; const int* dims; const uint32_t* occ; const uint32_t* known; int wpr;
    float occ_value, free_value; bool prune;
  };

  OcTreeNode* newNode() { ++n_nodes; return new OcTreeNode(); }
  static bool hasChildren(const OcTreeNode* n) {
    if (!n->children) return false;
    for (int i = 0; i < 8; ++i) if (n->children[i]) return true;
    return false;
  }
  void deleteRecurs(OcTreeNode* n) {
    if (n->children) {
      for (int i = 0; i < 8; ++i) if (n->children[i]) deleteRecurs(n->children[i]);
      delete[] n->children;
    }
    delete n;
  }
  void allocChildren(OcTreeNode* n) {
    n->children = new OcTreeNode*[8];
    for (int i = 0; i < 8; ++i) n->children[i] = NULL;
  }
  void expandNode(OcTreeNode* n) {
    if (!n->children) allocChildren(n);
    for (int i = 0; i < 8; ++i) { n->children[i] = newNode(); n->children[i]->value = n->value; }
  }
  bool pruneNode(OcTreeNode* n) {
    // isNodeCollapsible: all 8 children exist, are childless and carry the same value.
    if (!n->children) return false;
    for (int i = 0; i < 8; ++i) if (!n->children[i] || hasChildren(n->children[i])) return false;
    for (int i = 1; i < 8; ++i) if (!(n->children[i]->value == n->children[0]->value)) return false;
    n->value = n->children[0]->value;
    for (int i = 0; i < 8; ++i) { delete n->children[i]; --n_nodes; }
    delete[] n->children;
    n->children = NULL;
    return true;
  }
  static void updateOccupancyChildren(OcTreeNode* n) {
    float mx = -std::numeric_limits<float>::max();
    for (int i = 0; i < 8; ++i) if (n->children && n->children[i]) mx = std::max(mx, n->children[i]->value);
    n->value = mx;
  }
  OcTreeNode* setNodeValueRecurs(OcTreeNode* node, bool node_just_created, const OcTreeKey& key, unsigned depth,
                                 float v, bool lazy_eval) {
    bool created_node = false;
    if (depth < tree_depth) {
      unsigned pos = computeChildIdx(key, (int)tree_depth - 1 - (int)depth);
      if (!(node->children && node->children[pos])) {
        if (!hasChildren(node) && !node_just_created) {
          expandNode(node);
        } else {
          if (!node->children) allocChildren(node);
          node->children[pos] = newNode();
          created_node = true;
        }
      }
      if (lazy_eval) return setNodeValueRecurs(node->children[pos], created_node, key, depth + 1, v, lazy_eval);
      OcTreeNode* ret = setNodeValueRecurs(node->children[pos], created_node, key, depth + 1, v, lazy_eval);
      if (pruneNode(node)) ret = node; else updateOccupancyChildren(node);
      return ret;
    }
    node->value = v;
    return node;
  }
  OcTreeNode* updateNodeRecurs(OcTreeNode* node, bool node_just_created, const OcTreeKey& key, unsigned depth, float update) {
    bool created_node = false;
    if (depth < tree_depth) {
      unsigned pos = computeChildIdx(key, (int)tree_depth - 1 - (int)depth);
      if (!(node->children && node->children[pos])) {
        if (!hasChildren(node) && !node_just_created) {
          expandNode(node);
        } else {
          if (!node->children) allocChildren(node);
          node->children[pos] = newNode();
          created_node = true;
        }
      }
      OcTreeNode* ret = updateNodeRecurs(node->children[pos], created_node, key, depth + 1, update);
      if (pruneNode(node)) ret = node; else updateOccupancyChildren(node);
      return ret;
    }
    // updateNodeLogOdds: float add, then clamp
    node->value = node->value + update;
    if (node->value < clamping_thres_min) node->value = clamping_thres_min;
    else if (node->value > clamping_thres_max) node->value = clamping_thres_max;
    return node;
  }
  void updateInnerRecurs(OcTreeNode* n, unsigned depth) {
    if (hasChildren(n)) {
      if (depth < tree_depth)
        for (int i = 0; i < 8; ++i) if (n->children[i]) updateInnerRecurs(n->children[i], depth + 1);
      updateOccupancyChildren(n);
    }
  }
  void pruneRecurs(OcTreeNode* n, unsigned depth, unsigned max_depth, unsigned& num_pruned) {
    if (depth < max_depth) {
      if (n->children)
        for (int i = 0; i < 8; ++i) if (n->children[i]) pruneRecurs(n->children[i], depth + 1, max_depth, num_pruned);
    } else {
      if (pruneNode(n)) num_pruned++;
    }
  }
  template <class F>
  void leafRecurs(const OcTreeNode* n, unsigned depth, unsigned ox, unsigned oy, unsigned oz, F& f) const {
    if (!hasChildren(n)) { f(n, ox, oy, oz, 1u << (tree_depth - depth)); return; }
    unsigned half = 1u << (tree_depth - depth - 1);
    for (int i = 0; i < 8; ++i)
      if (n->children[i])
        leafRecurs(n->children[i], depth + 1, ox + ((i & 1) ? half : 0), oy + ((i & 2) ? half : 0),
                   oz + ((i & 4) ? half : 0), f);
  }
  static inline bool denseBit(const uint32_t* bits, const Dense& dn, int kx, int ky, int kz) {
    int x = kx - dn.kmin[0], y = ky - dn.kmin[1], z = kz - dn.kmin[2];
    if (x < 0 || y < 0 || z < 0 || x >= dn.dims[0] || y >= dn.dims[1] || z >= dn.dims[2]) return false;
    size_t w = ((size_t)z * dn.dims[1] + y) * dn.wpr + (x >> 5);
    return (bits[w] >> (x & 31)) & 1u;
  }
  OcTreeNode* buildRecurs(const Dense& dn, unsigned depth, unsigned ox, unsigned oy, unsigned oz) {
    unsigned edge = 1u << (tree_depth - depth);
    // cube entirely outside the dense window: nothing known there
    if ((int)(ox + edge) <= dn.kmin[0] || (int)ox >= dn.kmin[0] + dn.dims[0] ||
        (int)(oy + edge) <= dn.kmin[1] || (int)oy >= dn.kmin[1] + dn.dims[1] ||
        (int)(oz + edge) <= dn.kmin[2] || (int)oz >= dn.kmin[2] + dn.dims[2])
      return NULL;
    if (depth == tree_depth) {
      if (!denseBit(dn.known, dn, ox, oy, oz)) return NULL;
      OcTreeNode* n = newNode();
      n->value = denseBit(dn.occ, dn, ox, oy, oz) ? dn.occ_value : dn.free_value;
      return n;
    }
    unsigned half = edge >> 1;
    OcTreeNode* kids[8];
    bool any = false;
    for (int i = 0; i < 8; ++i) {
      kids[i] = buildRecurs(dn, depth + 1, ox + ((i & 1) ? half : 0), oy + ((i & 2) ? half : 0),
                            oz + ((i & 4) ? half : 0));
      any = any || kids[i];
    }
    if (!any) return NULL;
    OcTreeNode* n = newNode();
    allocChildren(n);
    for (int i = 0; i < 8; ++i) n->children[i] = kids[i];
    if (dn.prune) { if (!pruneNode(n)) updateOccupancyChildren(n); }
    return n;
  }
};

// volumetric_mapping::OctomapWorld restated (octomap_world.cc).
class OctomapWorld {
 public:
  enum CellStatus { kFree = 0, kOccupied = 1, kUnknown = 2 };

  // Parameter defaults of OctomapParameters; the reactive launch file overrides resolution 0.3 and
  // treat_unknown_as_occupied false (launch/reactive_planner_test.launch:14,29).
  explicit OctomapWorld(double resolution, bool treat_unknown_as_occupied_ = true)
      : tree(resolution), treat_unknown_as_occupied(treat_unknown_as_occupied_) {
    tree.occ_prob_thres_log = logodds(0.7);    // threshold_occupancy
    tree.clamping_thres_min = logodds(0.12);   // threshold_min
    tree.clamping_thres_max = logodds(0.97);   // threshold_max
    tree.prob_hit_log = logodds(0.65);         // probability_hit
    tree.prob_miss_log = logodds(0.4);         // probability_miss
  }

  // per-thread counters of voxel lookups / rays issued by getLineStatus (for the byte model)
  static long& tl_steps() { static thread_local long v = 0; return v; }
  static long& tl_rays() { static thread_local long v = 0; return v; }

  double getResolution() const { return tree.resolution; }

  // OctomapWorld::getLineStatus: unknown cells only matter when treat_unknown_as_occupied.
  CellStatus getLineStatus(const double start[3], const double end[3]) const {
    point3d s((float)start[0], (float)start[1], (float)start[2]);  // pointEigenToOctomap
    point3d e((float)end[0], (float)end[1], (float)end[2]);
    static thread_local std::vector<OcTreeKey> key_ray;  // upstream: a local KeyRay per call
    key_ray.clear();
    tree.computeRayKeys(s, e, key_ray);  // return value ignored upstream: out of range == empty ray
    ++tl_rays();
    for (size_t i = 0; i < key_ray.size(); ++i) {
      ++tl_steps();
      const OcTreeNode* node = tree.search(key_ray[i]);
      if (node == NULL) {
        if (treat_unknown_as_occupied) return kUnknown;
      } else if (tree.isNodeOccupied(node)) {
        return kOccupied;
      }
    }
    return kFree;
  }

  // The lattice of parallel rays swept over the robot box; returned as the three offset tables so
  // the same host loop (double +=, inclusive upper bound) can be reproduced by any caller.
  static void rayLattice(const double box[3], double resolution, std::vector<double> off[3]) {
    const double epsilon = 0.001;
    for (int a = 0; a < 3; ++a) {
      double disc = box[a] / std::ceil((box[a] + epsilon) / resolution);
      if (disc <= 0.0) disc = 1.0;
      double half = box[a] * 0.5;
      off[a].clear();
      for (double v = -half; v <= half; v += disc) off[a].push_back(v);
    }
  }

  // OctomapWorld::getLineStatusBoundingBox: x outer, y middle, z inner; first non-free ray wins.
  CellStatus getLineStatusBoundingBox(const double start[3], const double end[3], const double box[3]) const {
    std::vector<double> off[3];
    rayLattice(box, getResolution(), off);
    for (size_t ix = 0; ix < off[0].size(); ++ix)
      for (size_t iy = 0; iy < off[1].size(); ++iy)
        for (size_t iz = 0; iz < off[2].size(); ++iz) {
          double s[3] = {start[0] + off[0][ix], start[1] + off[1][iy], start[2] + off[2][iz]};
          double e[3] = {end[0] + off[0][ix], end[1] + off[1][iy], end[2] + off[2][iz]};
          CellStatus ret = getLineStatus(s, e);
          if (ret != kFree) return ret;
        }
    return kFree;
  }

  // OctomapWorld::getCellStatusPoint
  CellStatus getCellStatusPoint(const double p[3]) const {
    const OcTreeNode* node = tree.search(p[0], p[1], p[2]);
    if (node == NULL) return kUnknown;
    if (tree.isNodeOccupied(node)) return kOccupied;
    return kFree;
  }

  // Node validity.  The reference has NO implementation of this: the hook
  // Heuristic::isFilteringConditionSatisfied (include/sspp/heuristic_interface.h:44) returns true
  // in the only heuristic present (src/distance_heuristic.cpp:134-138).  AUTHORED SPEC, modelled on
  // OctomapWorld::getCellStatusBoundingBox and stated in key space so it is exactly computable:
  //   1. the voxel holding p (key from the DOUBLE coordinates) must be free, else its status;
  //   2. key box = [coordToKey((float)(p-box/2)), coordToKey((float)(p+box/2))] per axis
  //      (pointEigenToOctomap narrowing, as leaf_bbx_iterator receives it);
  //      any occupied voxel in it -> kOccupied;
  //   3. else any never-observed voxel in it -> kUnknown (upstream checks this whatever
  //      treat_unknown_as_occupied says); else kFree.
  //   A corner outside the 16-bit key range -> kUnknown.
  CellStatus getCellStatusBoundingBox(const double p[3], const double box[3]) const {
    CellStatus center = getCellStatusPoint(p);
    if (center != kFree) return center;
    OcTreeKey kmin, kmax;
    for (int a = 0; a < 3; ++a) {
      float lo = (float)(p[a] - box[a] / 2);
      float hi = (float)(p[a] + box[a] / 2);
      if (!tree.coordToKeyChecked((double)lo, kmin[a]) || !tree.coordToKeyChecked((double)hi, kmax[a]))
        return kUnknown;
    }
    bool unknown = false;
    for (unsigned z = kmin[2]; z <= kmax[2]; ++z)
      for (unsigned y = kmin[1]; y <= kmax[1]; ++y)
        for (unsigned x = kmin[0]; x <= kmax[0]; ++x) {
          const OcTreeNode* n = tree.search(OcTreeKey((key_type)x, (key_type)y, (key_type)z));
          if (n == NULL) unknown = true;
          else if (tree.isNodeOccupied(n)) return kOccupied;
        }
    return unknown ? kUnknown : kFree;
  }

  // OctomapWorld::setLogOddsBoundingBox (setFree / setOccupied): res-spaced lattice over the box
  // grown by 0.001, lazy insertion, then updateInnerOccupancy.
  void setLogOddsBoundingBox(const double pos[3], const double box[3], float log_odds) {
    const double epsilon = 0.001;
    double res = getResolution();
    double bmin[3], bmax[3];
    for (int a = 0; a < 3; ++a) {
      bmin[a] = pos[a] - box[a] / 2 - epsilon;
      bmax[a] = pos[a] + box[a] / 2 + epsilon;
    }
    for (double x = bmin[0]; x <= bmax[0]; x += res)
      for (double y = bmin[1]; y <= bmax[1]; y += res)
        for (double z = bmin[2]; z <= bmax[2]; z += res)
          tree.setNodeValue(point3d((float)x, (float)y, (float)z), log_odds, true);
    tree.updateInnerOccupancy();
  }
  void setFree(const double pos[3], const double box[3]) { setLogOddsBoundingBox(pos, box, tree.clamping_thres_min); }
  void setOccupied(const double pos[3], const double box[3]) { setLogOddsBoundingBox(pos, box, tree.clamping_thres_max); }

  // OctomapWorld::insertPointcloudIntoMapImpl + castRay + updateOccupancy (max_free_space == 0, the
  // default), for points ALREADY in the world frame (the TF / pcl::transformPointCloud step of
  // insertPointcloudWithTf, src/reactive_planner_server.cpp:238, is the caller's).  Batch semantics:
  // a point whose voxel an EARLIER point of the same cloud already marked occupied casts no ray;
  // every voxel is updated at most once per cloud, occupied winning over free.
  void insertPointcloud(const float* pts, long n, const double origin_d[3], double sensor_max_range) {
    const point3d origin((float)origin_d[0], (float)origin_d[1], (float)origin_d[2]);
    std::set<uint64_t> free_cells, occupied_cells;
    std::vector<OcTreeKey> ray;
    for (long i = 0; i < n; ++i) {
      const point3d p(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
      OcTreeKey key(tree.coordToKey((double)p(0)), tree.coordToKey((double)p(1)), tree.coordToKey((double)p(2)));
      if (occupied_cells.find(pack(key)) != occupied_cells.end()) continue;
      if (sensor_max_range < 0.0 || (p - origin).norm() <= sensor_max_range) {
        if (tree.computeRayKeys(origin, p, ray))
          for (size_t k = 0; k < ray.size(); ++k) free_cells.insert(pack(ray[k]));
        OcTreeKey ek;
        if (tree.coordToKeyChecked(p, ek)) occupied_cells.insert(pack(ek));
      } else {
        point3d dir = (p - origin).normalized();
        const float r = (float)sensor_max_range;  // Vector3::operator*(float)
        point3d new_end(origin(0) + dir(0) * r, origin(1) + dir(1) * r, origin(2) + dir(2) * r);
        if (tree.computeRayKeys(origin, new_end, ray))
          for (size_t k = 0; k < ray.size(); ++k) free_cells.insert(pack(ray[k]));
      }
    }
    for (std::set<uint64_t>::const_iterator it = occupied_cells.begin(); it != occupied_cells.end(); ++it) {
      tree.updateNode(unpack(*it), true);
      free_cells.erase(*it);
    }
    for (std::set<uint64_t>::const_iterator it = free_cells.begin(); it != free_cells.end(); ++it) tree.updateNode(unpack(*it), false);
    tree.updateInnerOccupancy();
  }
  static uint64_t pack(const OcTreeKey& k) { return ((uint64_t)k[0] << 32) | ((uint64_t)k[1] << 16) | (uint64_t)k[2]; }
  static OcTreeKey unpack(uint64_t v) { return OcTreeKey((key_type)(v >> 32), (key_type)((v >> 16) & 0xffff), (key_type)(v & 0xffff)); }

  // OctomapWorld::getAllOccupiedBoxes / getAllFreeBoxes: (centre, edge length) per leaf.
  void getAllBoxes(bool occupied_boxes, std::vector<double>& centres, std::vector<double>& sizes) const {
    struct Visit {
      const OcTree* t; bool want_occ; std::vector<double>* c; std::vector<double>* s;
      void operator()(const OcTreeNode* n, unsigned ox, unsigned oy, unsigned oz, unsigned edge) {
        if (t->isNodeOccupied(n) != want_occ) return;
        unsigned o[3] = {ox, oy, oz};
        for (int a = 0; a < 3; ++a)
          c->push_back((double((int)o[a] - (int)t->tree_max_val) + 0.5 * edge) * t->resolution);
        s->push_back(edge * t->resolution);
      }
    } v = {&tree, occupied_boxes, &centres, &sizes};
    tree.forEachLeaf(v);
  }

  OcTree tree;
  bool treat_unknown_as_occupied;
};

}  // namespace orc

#endif
