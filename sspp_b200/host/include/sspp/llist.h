// LList — the open / closed list of the search (reference include/sspp/llist.h, src/llist.cpp): a
// doubly linked list kept ordered by f.  Parity-critical rules kept verbatim in behaviour:
//   add():    the new node goes BEFORE the first element whose f is >= its own (cost mode; <= in
//             reward mode), i.e. last-in-first-out among equal f (src/llist.cpp:99-118);
//   find()/remove(): match by pose within 1e-12; remove() deletes the first match from the head,
//             whatever its cost (src/llist.cpp:44-56,130-150).
// The reference walks the list for all three (O(n) each).  Here the order lives in a multimap keyed
// by f (insertion before equal keys = emplace_hint at lower_bound) and membership in a per-way-point
// table, so each operation is O(log n) with the same observable sequence.
#pragma once
#include <map>
#include <unordered_map>
#include "node.h"

namespace SSPP {
class LList {
 public:
  explicit LList(bool ascending = true) : ascending_(ascending) {}
  ~LList() { free(); }
  bool empty() const { return order_.empty(); }
  Node* getHead() const { return order_.empty() ? nullptr : order_.begin()->second; }
  void add(Node* n, bool ascending);
  Node* find(const Node* q) const;   // first list element with q's pose, or nullptr
  bool remove(const Node* q);        // deletes the first list element with q's pose
  Node* popHead();                   // unlinks and returns the head (caller owns it)
  void free();                       // deletes every node still in the list
  void release(Node* n);             // unlinks n without deleting it
  size_t size() const { return order_.size(); }

 private:
  typedef std::multimap<double, Node*> Order;   // key = f (cost mode) or -f (reward mode)
  Order order_;
  std::unordered_multimap<const void*, Order::iterator> where_;  // way-point -> entries
  bool ascending_;
  double key(const Node* n) const { return ascending_ ? n->f_value : -n->f_value; }
  static const void* ident(const Node* n);
};
}  // namespace SSPP
