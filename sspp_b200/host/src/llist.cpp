// LList (reference src/llist.cpp) — see llist.h for the ordering rules kept.
#include "sspp/llist.h"

namespace SSPP {
const void* LList::ident(const Node* n) { return n->ssnode ? (const void*)n->ssnode : (const void*)n; }

void LList::add(Node* n, bool ascending) {
  if (order_.empty()) ascending_ = ascending;
  const double k = key(n);
  // before the first element whose key is >= k: newest first among equal f
  Order::iterator it = order_.emplace_hint(order_.lower_bound(k), k, n);
  where_.emplace(ident(n), it);
}

Node* LList::find(const Node* q) const {
  auto range = where_.equal_range(ident(q));
  Node* best = nullptr;
  double bk = 0;
  for (auto it = range.first; it != range.second; ++it) {
    Node* c = it->second->second;
    if (!(*c == *q)) continue;
    double k = it->second->first;
    if (!best || k < bk) { best = c; bk = k; }
    else if (k == bk) {
      // same key: whichever comes first in list order
      for (auto w = order_.lower_bound(k); w != order_.end() && w->first == k; ++w) {
        if (w->second == c) { best = c; break; }
        if (w->second == best) break;
      }
    }
  }
  return best;
}

void LList::release(Node* n) {
  auto range = where_.equal_range(ident(n));
  for (auto it = range.first; it != range.second; ++it)
    if (it->second->second == n) {
      order_.erase(it->second);
      where_.erase(it);
      return;
    }
}

bool LList::remove(const Node* q) {
  Node* victim = find(q);
  if (!victim) return false;
  release(victim);
  delete victim;
  return true;
}

Node* LList::popHead() {
  if (order_.empty()) return nullptr;
  Node* n = order_.begin()->second;
  release(n);
  return n;
}

void LList::free() {
  for (auto& kv : order_) delete kv.second;
  order_.clear();
  where_.clear();
}
}  // namespace SSPP
