// bvh2_builder.cpp — see bvh2_builder.h.  Compiled with -O2 -ffp-contract=off and no -march so that every FP32 operation
// rounds like the reference's own host build; results are compared byte for byte with the oracle in tests/test_host_cpu.py.
#include "bvh2_builder.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <deque>
#include <limits>
#include <memory>
#include <mutex>
#include <thread>

namespace vxh {
namespace {

constexpr int kBins = 8;             // bvh.cpp:6
constexpr float kLarge = 1e30f;      // geometry.h:15

inline float lo_of(float a, float b) { return a < b ? a : b; }   // the reference's own fminf/fmaxf (geometry.h:178-179)
inline float hi_of(float a, float b) { return a > b ? a : b; }
inline float axis_of(const Vec3 &v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

struct Box {
  Vec3 lo{kLarge, kLarge, kLarge}, hi{-kLarge, -kLarge, -kLarge};
  void add(const Vec3 &p) {
    lo = {lo_of(lo.x, p.x), lo_of(lo.y, p.y), lo_of(lo.z, p.z)};
    hi = {hi_of(hi.x, p.x), hi_of(hi.y, p.y), hi_of(hi.z, p.z)};
  }
  void add(const Box &b) {   // AABB::grow(AABB&) (bvh.h:19-24): empty boxes are skipped
    if (b.lo.x != kLarge) {
      add(b.lo);
      add(b.hi);
    }
  }
  float half_area() const {  // AABB::area (bvh.h:26-29)
    float ex = hi.x - lo.x, ey = hi.y - lo.y, ez = hi.z - lo.z;
    return ex * ey + ey * ez + ez * ex;
  }
};

struct TNode {   // a node of the task tree
  Box box;
  uint32_t first = 0, count = 0;
  TNode *left = nullptr, *right = nullptr;
};

class Builder {
public:
  Builder(const Tri *tris, const Vec3 *cent, uint32_t *idx, int threads) : tris_(tris), cent_(cent), idx_(idx), threads_(threads < 1 ? 1 : threads) {
    arenas_.resize(threads_ + 1);
  }

  TNode *run(uint32_t count) {
    grain_ = threads_ > 1 ? std::max<uint32_t>(4096u, count / (uint32_t)(threads_ * 16)) : 0xFFFFFFFFu;
    TNode *root = alloc(threads_);
    root->first = 0, root->count = count;
    if (threads_ == 1) {
      grow(root, threads_);
      return root;
    }
    pending_ = 1;
    queue_.push_back(root);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads_; ++t) pool.emplace_back([this, t] { worker(t); });
    for (auto &th : pool) th.join();
    return root;
  }

private:
  TNode *alloc(int arena) {
    arenas_[arena].emplace_back();
    return &arenas_[arena].back();
  }

  void worker(int me) {
    for (;;) {
      TNode *n = nullptr;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [this] { return !queue_.empty() || pending_ == 0; });
        if (queue_.empty()) return;
        n = queue_.back();
        queue_.pop_back();
      }
      grow(n, me);
      bool done;
      {
        std::lock_guard<std::mutex> lk(mu_);
        done = (--pending_ == 0);
      }
      if (done) cv_.notify_all();
    }
  }

  void spawn(TNode *n) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      ++pending_;
      queue_.push_back(n);
    }
    cv_.notify_one();
  }

  // BVH::updateNodeBounds (bvh.cpp:142-161): node box over the vertices, centroid box over the centroids, in index order
  void bounds(TNode *n, Box &cbox) const {
    Box b, c;
    for (uint32_t i = 0; i < n->count; ++i) {
      uint32_t t = idx_[n->first + i];
      const Tri &tr = tris_[t];
      b.lo = {lo_of(b.lo.x, tr.v0.x), lo_of(b.lo.y, tr.v0.y), lo_of(b.lo.z, tr.v0.z)};
      b.lo = {lo_of(b.lo.x, tr.v1.x), lo_of(b.lo.y, tr.v1.y), lo_of(b.lo.z, tr.v1.z)};
      b.lo = {lo_of(b.lo.x, tr.v2.x), lo_of(b.lo.y, tr.v2.y), lo_of(b.lo.z, tr.v2.z)};
      b.hi = {hi_of(b.hi.x, tr.v0.x), hi_of(b.hi.y, tr.v0.y), hi_of(b.hi.z, tr.v0.z)};
      b.hi = {hi_of(b.hi.x, tr.v1.x), hi_of(b.hi.y, tr.v1.y), hi_of(b.hi.z, tr.v1.z)};
      b.hi = {hi_of(b.hi.x, tr.v2.x), hi_of(b.hi.y, tr.v2.y), hi_of(b.hi.z, tr.v2.z)};
      c.add(cent_[t]);
    }
    n->box = b;
    cbox = c;
  }

  // BVH::findBestSplitPlane (bvh.cpp:89-140): 8 bins per axis, 7 candidate planes, cost = n_left*A_left + n_right*A_right
  float best_plane(const TNode *n, const Box &cbox, uint32_t &axis, uint32_t &plane) const {
    float best = kLarge;
    for (int a = 0; a < 3; ++a) {
      float cmin = axis_of(cbox.lo, a), cmax = axis_of(cbox.hi, a);
      if (cmin == cmax) continue;
      float scale = kBins / (cmax - cmin);
      Box bb[kBins];
      int bn[kBins] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (uint32_t i = 0; i < n->count; ++i) {
        uint32_t t = idx_[n->first + i];
        int b = std::min(kBins - 1, (int)((axis_of(cent_[t], a) - cmin) * scale));
        bn[b]++;
        bb[b].add(tris_[t].v0);
        bb[b].add(tris_[t].v1);
        bb[b].add(tris_[t].v2);
      }
      float lcost[kBins - 1], rcost[kBins - 1];
      Box lbox, rbox;
      int lsum = 0, rsum = 0;
      for (int i = 0; i < kBins - 1; ++i) {
        lsum += bn[i];
        lbox.add(bb[i]);
        lcost[i] = lsum * lbox.half_area();
        rsum += bn[kBins - 1 - i];
        rbox.add(bb[kBins - 1 - i]);
        rcost[kBins - 2 - i] = rsum * rbox.half_area();
      }
      for (int i = 0; i < kBins - 1; ++i) {
        const float cost = lcost[i] + rcost[i];
        if (cost < best) axis = (uint32_t)a, plane = (uint32_t)(i + 1), best = cost;
      }
    }
    return best;
  }

  // BVH::partitionTriangles (bvh.cpp:68-87): in-place two-pointer partition on the bin index
  uint32_t partition(const TNode *n, const Box &cbox, uint32_t axis, uint32_t plane) const {
    float cmin = axis_of(cbox.lo, (int)axis);
    float scale = kBins / (axis_of(cbox.hi, (int)axis) - cmin);
    uint32_t *p = idx_ + n->first;
    uint32_t i = 0, j = n->count - 1;
    while (i <= j) {
      uint32_t b = (uint32_t)std::max(0, std::min((int)((axis_of(cent_[p[i]], (int)axis) - cmin) * scale), kBins - 1));
      if (b < plane) ++i;
      else std::swap(p[i], p[j--]);
    }
    return i;
  }

  // BVH::initializeNode + subdivide (bvh.cpp:29-66), as a loop over a private stack; large children become pool tasks
  void grow(TNode *root, int me) {
    std::vector<TNode *> todo{root};
    while (!todo.empty()) {
      TNode *n = todo.back();
      todo.pop_back();
      Box cbox;
      bounds(n, cbox);
      if (n->count <= 1) continue;
      uint32_t axis = 0, plane = 0;
      float split_cost = best_plane(n, cbox, axis, plane);
      float ex = n->box.hi.x - n->box.lo.x, ey = n->box.hi.y - n->box.lo.y, ez = n->box.hi.z - n->box.lo.z;
      float leaf_cost = (2 * (ex * ey + ey * ez + ez * ex)) * n->count;   // bvh_node_t::calculateNodeCost (common.h:41-43)
      if (split_cost >= leaf_cost) continue;
      uint32_t nl = partition(n, cbox, axis, plane);
      if (nl == 0 || nl == n->count) continue;
      TNode *l = alloc(me), *r = alloc(me);
      l->first = n->first, l->count = nl;
      r->first = n->first + nl, r->count = n->count - nl;
      n->left = l, n->right = r;
      for (TNode *c : {r, l}) {
        if (c->count >= grain_) spawn(c);
        else todo.push_back(c);
      }
    }
  }

  const Tri *tris_;
  const Vec3 *cent_;
  uint32_t *idx_;
  int threads_;
  uint32_t grain_ = 0xFFFFFFFFu;
  std::vector<std::deque<TNode>> arenas_;   // one per worker (+1 for the caller): stable addresses, no locking
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<TNode *> queue_;
  int pending_ = 0;
};

}  // namespace

uint32_t build_bvh2(const Tri *tris, const Vec3 *centroids, uint32_t count, uint32_t *idx, BvhNode *nodes, int threads) {
  Builder b(tris, centroids, idx, threads);
  TNode *root = b.run(count);
  // flatten in the reference's allocation order (bvh.cpp:52-65): a node's two children take the next two slots, then the
  // whole left subtree is laid out, then the right one
  uint32_t used = 1;
  std::vector<std::pair<const TNode *, uint32_t>> todo{{root, 0u}};
  while (!todo.empty()) {
    auto [n, at] = todo.back();
    todo.pop_back();
    BvhNode &o = nodes[at];
    o.lo = n->box.lo, o.hi = n->box.hi;
    if (n->left) {
      uint32_t l = used;
      used += 2;
      o.leftFirst = l, o.triCount = 0;
      todo.push_back({n->right, l + 1});   // popped after the whole left subtree ... but indices must follow DFS order:
      todo.push_back({n->left, l});        // the left subtree is expanded first because it is on top of the stack
    } else {
      o.leftFirst = n->first, o.triCount = n->count;
    }
  }
  return used;
}

}  // namespace vxh
