// Astar (reference src/astar.cpp).  Cited lines are the reference's.
#include "sspp/astar.h"

#include <limits>

namespace SSPP {
Astar::Astar(Robot* rob, int freq) : robot(rob), progressDisplayFrequency(freq) {}

Astar::~Astar() {
  if (openList) { openList->free(); delete openList; }
  if (closedList) { closedList->free(); delete closedList; }
}

// nearest way-point to the start; ties go to the first one in list order (strict '<', :76-135)
void Astar::findRoot() {
  if (!searchspace) throw SSPPException("No SearchSpace Defined");
  double shortest = std::numeric_limits<double>::max();
  root = new Node;
  for (SearchSpaceNode* t = searchspace; t; t = t->next) {
    double d = Dist(t->location, start.p);
    if (d < shortest) {
      shortest = d;
      root->pose.p = t->location;
      root->senPoses.clear();
      for (size_t i = 0; i < t->sensorLocation.poses.size(); ++i) {
        Pose sp;
        sp.p = t->sensorLocation.poses[i];
        root->senPoses.push_back(sp);
      }
      root->ssnode = t;
    }
  }
  root->id = 0;
  root->parent = root->next = root->prev = nullptr;
  root->g_value = root->h_value = root->distance = root->coverage = 0;
  heuristic->calculateHeuristic(root);
  root->depth = 0;
}

// one tree node per child of the way-point, chained by PREPENDING, so the caller meets them in
// reverse children order; ids follow children order (:375-459)
Node* Astar::makeChildrenNodes(Node* parent) {
  const SearchSpaceNode* here = parent->ssnode;
  if (!searchspace || !here) return nullptr;
  Tree t;
  t.location = parent->pose.p;
  Node* q = nullptr;
  for (size_t i = 0; i < here->children.size(); ++i) {
    const SearchSpaceNode* c = here->children[i];
    Node* p = new Node;
    p->pose.p = c->location;
    for (size_t j = 0; j < c->sensorLocation.poses.size(); ++j) {
      Pose sp;
      sp.p = c->sensorLocation.poses[j];
      p->senPoses.push_back(sp);
    }
    p->ssnode = c;
    p->parent = parent;
    p->next = q;
    p->prev = nullptr;
    p->depth = parent->depth + 1;
    p->id = globalID++;
    q = p;
    t.children.push_back(p->pose.p);
  }
  if (!t.children.empty()) tree.push_back(t);
  return q;
}

Node* Astar::astarSearch(Pose startPose) {
  nodesExpanded = 0;
  globalID = 0;
  expansionOrder.clear();
  tree.clear();
  if (!openList) openList = new LList;
  if (!closedList) closedList = new LList;
  openList->free();
  closedList->free();
  if (!searchspace) return nullptr;
  flatten();  // gives every way-point its list index (expansionOrder reports those)
  start.p = startPose.p;
  findRoot();
  const bool cost = heuristic->isCost();
  openList->add(root, cost);
  int count = 0;
  while (!openList->empty()) {
    if (progressDisplayFrequency > 0 && (count++ % progressDisplayFrequency) == 0) heuristic->displayProgress(tree);
    // the head leaves OPEN, a copy of it enters CLOSED (:268-272)
    Node* head = openList->popHead();
    current = new Node(head);
    delete head;
    closedList->add(current, cost);
    ++nodesExpanded;
    expansionOrder.push_back(current->ssnode ? current->ssnode->index : -1);
    // `current` is always a fresh copy, so the reference's `current != root` never fails: the root
    // itself may end the search (:275)
    if (heuristic->terminateConditionReached(current)) {
      Node* path = nullptr;
      for (Node* p = current; p; p = p->parent) {
        closedList->release(p);  // path nodes leave CLOSED and are handed to the caller
        p->next = path;
        p->prev = nullptr;
        if (path) path->prev = p;
        path = p;
      }
      openList->free();
      closedList->free();
      return path;
    }
    Node* childList = makeChildrenNodes(current);
    // score all children of this expansion first (one batched call for heuristics that offer it; the
    // default loops over calculateHeuristic in the same order), then file them as the reference does
    std::vector<Node*> batch;
    for (Node* c = childList; c; c = c->next) batch.push_back(c);
    heuristic->calculateHeuristicAll(batch);
    while (childList) {
      Node* child = childList;
      childList = childList->next;
      child->next = child->prev = nullptr;
      ++globalcount;
      // note: the child's f is compared with the CURRENT node's f, not with the stored duplicate's (:330-351)
      const bool condition = cost ? (current->f_value <= child->f_value) : (current->f_value >= child->f_value);
      if (openList->find(child) && condition) { delete child; continue; }
      if (closedList->find(child) && condition) { delete child; continue; }
      openList->remove(child);
      openList->add(child, cost);
    }
    if (current->id > MAXNODES) {  // expanded too many nodes without a solution (:359-368)
      closedList->free();
      openList->free();
      return nullptr;
    }
  }
  closedList->free();
  return nullptr;
}

void Astar::freeSearchTree() {}
}  // namespace SSPP
