// Astar — the search engine (reference include/sspp/astar.h:49-89, src/astar.cpp).  The observable
// behaviour — which node is expanded when, which parent wins, the ids, the abort rule — follows the
// reference line by line (the cited lines are next to each step in astar.cpp); the linear scans of
// the reference (find the way-point of a pose, find a pose in OPEN/CLOSED) are table look-ups.
#pragma once
#include <vector>
#include "heuristic_interface.h"
#include "llist.h"
#include "node.h"
#include "robot.h"
#include "searchspace.h"

namespace SSPP {
class Astar : public SearchSpace {
 public:
  explicit Astar(Robot* rob = nullptr, int progressDisplayFrequency = -1);
  virtual ~Astar();
  void setRobot(Robot* r) { robot = r; }
  void setHeuristicFucntion(Heuristic* h) { heuristic = h; }  // spelling of the reference kept
  void setProgressDisplayFrequency(int f) { progressDisplayFrequency = f; }
  void setDebugDelay(double d) { debugDelay = d; }
  Node* astarSearch(Pose start);
  void freeNode(Node* n) { delete n; }

  long int MAXNODES = 0;
  Heuristic* heuristic = nullptr;
  bool debug = false;
  Pose start;
  Robot* robot = nullptr;
  Node *root = nullptr, *current = nullptr;
  LList *openList = nullptr, *closedList = nullptr;
  std::vector<Tree> tree;
  int progressDisplayFrequency = -1;
  double debugDelay = 0;
  int globalID = 0, globalcount = 0;
  int nodesExpanded = 0;                 // of the last search
  std::vector<int> expansionOrder;       // list index of every expanded way-point, in order (for parity tests)

 protected:
  void findRoot();
  Node* makeChildrenNodes(Node* parent);
  std::vector<Node*> allocated_;         // every tree node of the running search, freed at the next one
  void freeSearchTree();
};
}  // namespace SSPP
