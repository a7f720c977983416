#include "om_bvh.h"

#include <math.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <limits>

namespace omb {
namespace {

constexpr int kBins = 16;
constexpr uint32_t kBalanceDepth = 34;
const float kInf = std::numeric_limits<float>::infinity();

struct Box {
    float lo[3] = {kInf, kInf, kInf}, hi[3] = {-kInf, -kInf, -kInf};
    void grow(const float* l, const float* h) {
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], l[a]); hi[a] = std::max(hi[a], h[a]); }
    }
    void grow_pt(const float* p) { grow(p, p); }
    float half_area() const {
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0.f) || !(dy >= 0.f) || !(dz >= 0.f)) return 0.f;
        return dx * dy + dy * dz + dz * dx;
    }
};

struct Item {
    uint32_t prim;  // index into refs/boxes
    float c[3];     // box centre
};

struct Builder {
    const std::vector<uint32_t>& refs;
    const std::vector<Aabb>& boxes;
    std::vector<Item>& items;  // shared; a builder only touches its own range
    std::vector<Node>& nodes;
    uint32_t max_depth = 0;

    Box bounds(size_t b, size_t e) const {
        Box r;
        for (size_t i = b; i < e; i++) r.grow(boxes[items[i].prim].lo, boxes[items[i].prim].hi);
        return r;
    }

    // picks the split of items[b,e) into [b,mid) and [mid,e)
    size_t split(size_t b, size_t e) {
        Box cb;
        for (size_t i = b; i < e; i++) cb.grow_pt(items[i].c);
        int best_axis = -1, best_bin = -1;
        float best_cost = kInf;
        for (int a = 0; a < 3; a++) {
            const float ext = cb.hi[a] - cb.lo[a];
            if (!(ext > 0.f)) continue;
            Box bb[kBins];
            uint32_t cnt[kBins] = {0};
            const float scale = float(kBins) / ext;
            for (size_t i = b; i < e; i++) {
                int k = int((items[i].c[a] - cb.lo[a]) * scale);
                k = std::min(std::max(k, 0), kBins - 1);
                cnt[k]++;
                bb[k].grow(boxes[items[i].prim].lo, boxes[items[i].prim].hi);
            }
            float right_area[kBins];
            uint32_t right_cnt[kBins];
            Box acc;
            uint32_t n = 0;
            for (int k = kBins - 1; k > 0; k--) {
                if (cnt[k]) acc.grow(bb[k].lo, bb[k].hi);
                n += cnt[k];
                right_area[k] = acc.half_area();
                right_cnt[k] = n;
            }
            acc = Box();
            n = 0;
            for (int k = 0; k < kBins - 1; k++) {
                if (cnt[k]) acc.grow(bb[k].lo, bb[k].hi);
                n += cnt[k];
                if (n == 0 || right_cnt[k + 1] == 0) continue;
                const float cost = acc.half_area() * float(n) + right_area[k + 1] * float(right_cnt[k + 1]);
                if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = k; }
            }
        }
        if (best_axis >= 0) {
            const float ext = cb.hi[best_axis] - cb.lo[best_axis];
            const float scale = float(kBins) / ext;
            const float lo = cb.lo[best_axis];
            const int a = best_axis, kb = best_bin;
            auto it = std::partition(items.begin() + b, items.begin() + e, [&](const Item& x) {
                int k = int((x.c[a] - lo) * scale);
                k = std::min(std::max(k, 0), kBins - 1);
                return k <= kb;
            });
            const size_t mid = size_t(it - items.begin());
            if (mid > b && mid < e) return mid;
        }
        // all centres coincide (or the partition degenerated): split the range in the middle
        return b + (e - b) / 2;
    }

    // returns the child reference for items[b,e)
    int32_t build_range(size_t b, size_t e, uint32_t depth) {
        max_depth = std::max(max_depth, depth);
        if (e - b == 1) return ~int32_t(refs[items[b].prim]);
        // SAH can peel one object per level off a badly spread set; the traversal stacks hold 64 entries, so below
        // kBalanceDepth the ranges are halved instead (at most 28 more levels for any count a ref can index)
        size_t mid;
        if (depth >= kBalanceDepth) {
            Box cb;
            for (size_t i = b; i < e; i++) cb.grow_pt(items[i].c);
            int ax = 0;
            for (int a = 1; a < 3; a++) if (cb.hi[a] - cb.lo[a] > cb.hi[ax] - cb.lo[ax]) ax = a;
            mid = b + (e - b) / 2;
            std::nth_element(items.begin() + b, items.begin() + mid, items.begin() + e, [ax](const Item& x, const Item& y) { return x.c[ax] < y.c[ax]; });
        } else mid = split(b, e);
        const int32_t me = int32_t(nodes.size());
        nodes.push_back(Node());
        const Box lb = bounds(b, mid), rb = bounds(mid, e);
        const int32_t l = build_range(b, mid, depth + 1);
        const int32_t r = build_range(mid, e, depth + 1);
        Node& n = nodes[size_t(me)];
        for (int a = 0; a < 3; a++) { n.lmin[a] = lb.lo[a]; n.lmax[a] = lb.hi[a]; n.rmin[a] = rb.lo[a]; n.rmax[a] = rb.hi[a]; }
        n.left = l; n.right = r; n.pad0 = n.pad1 = 0;
        return me;
    }
};

}  // namespace

// Large inputs: the top of the tree is split on the calling thread down to ranges of a few thousand objects, the
// subtrees under it are built by a pool of threads into their own arrays (disjoint item ranges), then appended with
// their indices shifted.  Same splits as the serial build, so the same tree (node order aside).
namespace {
struct Sub {
    size_t b, e;
    uint32_t depth;
    std::vector<Node> nodes;
    int32_t root = 0;
    uint32_t max_depth = 0;
};
struct TopBuilder {
    Builder& bld;
    std::vector<Sub>& subs;
    std::vector<int32_t> ltask, rtask;  // per top node: index of the Sub that hangs under the left / right child, or -1
    size_t grain;
    int32_t go(size_t b, size_t e, uint32_t depth, int32_t* task_out) {
        *task_out = -1;
        if (e - b == 1) return ~int32_t(bld.refs[bld.items[b].prim]);
        if (e - b <= grain || depth >= 12) {
            Sub s; s.b = b; s.e = e; s.depth = depth;
            subs.push_back(std::move(s));
            *task_out = int32_t(subs.size() - 1);
            return 0;
        }
        const size_t mid = bld.split(b, e);
        const int32_t me = int32_t(bld.nodes.size());
        bld.nodes.push_back(Node());
        ltask.push_back(-1); rtask.push_back(-1);
        const Box lb = bld.bounds(b, mid), rb = bld.bounds(mid, e);
        int32_t lt, rt;
        const int32_t l = go(b, mid, depth + 1, &lt);
        const int32_t r = go(mid, e, depth + 1, &rt);
        Node& n = bld.nodes[size_t(me)];
        for (int a = 0; a < 3; a++) { n.lmin[a] = lb.lo[a]; n.lmax[a] = lb.hi[a]; n.rmin[a] = rb.lo[a]; n.rmax[a] = rb.hi[a]; }
        n.left = l; n.right = r; n.pad0 = n.pad1 = 0;
        ltask[size_t(me)] = lt; rtask[size_t(me)] = rt;
        bld.max_depth = std::max(bld.max_depth, depth);
        return me;
    }
};
}  // namespace

void build(const std::vector<uint32_t>& refs, const std::vector<Aabb>& boxes, BuildResult& out) {
    out.nodes.clear();
    out.depth = 0;
    const size_t n = refs.size();
    if (n == 0) return;
    out.nodes.reserve(n);
    std::vector<Item> items(n);
    Builder bld{refs, boxes, items, out.nodes};
    for (size_t i = 0; i < n; i++) {
        items[i].prim = uint32_t(i);
        for (int a = 0; a < 3; a++) items[i].c[a] = 0.5f * (boxes[i].lo[a] + boxes[i].hi[a]);
    }
    if (n == 1) {  // a root with one real child and one empty child
        Node root;
        for (int a = 0; a < 3; a++) { root.lmin[a] = boxes[0].lo[a]; root.lmax[a] = boxes[0].hi[a]; root.rmin[a] = kInf; root.rmax[a] = -kInf; }
        root.left = ~int32_t(refs[0]); root.right = ~int32_t(refs[0]); root.pad0 = root.pad1 = 0;
        out.nodes.push_back(root);
        out.depth = 1;
        return;
    }
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    if (n < 65536 || hw < 2) {
        bld.build_range(0, n, 1);
        out.depth = bld.max_depth;
        return;
    }
    std::vector<Sub> subs;
    TopBuilder top{bld, subs, {}, {}, std::max<size_t>(4096, n / (8 * size_t(hw)))};
    int32_t root_task;
    top.go(0, n, 1, &root_task);  // (n >= 65536 > grain: the root is a top node)
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= subs.size()) return;
            Sub& s = subs[k];
            s.nodes.reserve(s.e - s.b);
            Builder sb{refs, boxes, items, s.nodes};
            s.root = sb.build_range(s.b, s.e, s.depth);
            s.max_depth = sb.max_depth;
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < std::min<unsigned>(hw, 32u); t++) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
    // append the subtrees; their inner-node indices move by the offset, leaves (negative) stay
    const size_t n_top = out.nodes.size();
    std::vector<int32_t> offset(subs.size());
    size_t at = n_top;
    for (size_t k = 0; k < subs.size(); k++) { offset[k] = int32_t(at); at += subs[k].nodes.size(); }
    out.nodes.resize(at);
    uint32_t depth = bld.max_depth;
    for (size_t k = 0; k < subs.size(); k++) {
        Node* dst = out.nodes.data() + offset[k];
        for (size_t i = 0; i < subs[k].nodes.size(); i++) {
            Node nd = subs[k].nodes[i];
            if (nd.left >= 0) nd.left += offset[k];
            if (nd.right >= 0) nd.right += offset[k];
            dst[i] = nd;
        }
        depth = std::max(depth, subs[k].max_depth);
    }
    for (size_t i = 0; i < n_top; i++) {
        if (top.ltask[i] >= 0) { const Sub& s = subs[size_t(top.ltask[i])]; out.nodes[i].left = s.root >= 0 ? s.root + offset[size_t(top.ltask[i])] : s.root; }
        if (top.rtask[i] >= 0) { const Sub& s = subs[size_t(top.rtask[i])]; out.nodes[i].right = s.root >= 0 ? s.root + offset[size_t(top.rtask[i])] : s.root; }
    }
    out.depth = depth;
}


namespace {
struct Kid {
    int32_t ref;  // binary child reference
    float lo[3], hi[3];
};
float kid_area(const Kid& k) {
    const float dx = k.hi[0] - k.lo[0], dy = k.hi[1] - k.lo[1], dz = k.hi[2] - k.lo[2];
    if (!(dx >= 0.f) || !(dy >= 0.f) || !(dz >= 0.f)) return -1.f;
    return dx * dy + dy * dz + dz * dx;
}
void kids_of(const Node& n, Kid& l, Kid& r) {
    l.ref = n.left; r.ref = n.right;
    for (int a = 0; a < 3; a++) { l.lo[a] = n.lmin[a]; l.hi[a] = n.lmax[a]; r.lo[a] = n.rmin[a]; r.hi[a] = n.rmax[a]; }
}
}  // namespace

void collapse4(const BuildResult& in, Wide4& out, uint32_t width) {
    out.nodes.clear();
    out.depth = 0;
    out.stack_need = 0;
    if (in.nodes.empty()) return;
    width = std::min(4u, std::max(2u, width));
    out.nodes.reserve(in.nodes.size());
    struct Task { int32_t bin; int32_t wide; uint32_t depth; uint32_t on_stack; };
    std::vector<Task> todo;
    out.nodes.push_back(Node4());
    todo.push_back({0, 0, 1u, 0u});
    while (!todo.empty()) {
        const Task t = todo.back();
        todo.pop_back();
        out.depth = std::max(out.depth, t.depth);
        Kid kids[4];
        uint32_t m = 2;
        kids_of(in.nodes[size_t(t.bin)], kids[0], kids[1]);
        if (kids[0].ref == kids[1].ref && kids[0].ref < 0 && kid_area(kids[1]) < 0.f) m = 1;  // the one-primitive root of build()
        while (m < width) {
            int best = -1;
            float best_area = -1.f;
            for (uint32_t j = 0; j < m; j++) {
                if (kids[j].ref < 0) continue;
                const float a = kid_area(kids[j]);
                if (a > best_area) { best_area = a; best = int(j); }
            }
            if (best < 0) break;
            Kid l, r;
            kids_of(in.nodes[size_t(kids[best].ref)], l, r);
            kids[best] = l;
            kids[m++] = r;
        }
        const uint32_t below = t.on_stack + (m - 1u);
        out.stack_need = std::max(out.stack_need, below);
        Node4 w;
        for (uint32_t j = 0; j < 4; j++) {
            if (j < m) {
                w.lox[j] = kids[j].lo[0]; w.hix[j] = kids[j].hi[0]; w.loy[j] = kids[j].lo[1]; w.hiy[j] = kids[j].hi[1]; w.loz[j] = kids[j].lo[2]; w.hiz[j] = kids[j].hi[2];
                if (kids[j].ref < 0) w.child[j] = kids[j].ref;
                else {
                    w.child[j] = int32_t(out.nodes.size());
                    out.nodes.push_back(Node4());
                    todo.push_back({kids[j].ref, w.child[j], t.depth + 1u, below});
                }
            } else {
                w.lox[j] = w.loy[j] = w.loz[j] = kInf;
                w.hix[j] = w.hiy[j] = w.hiz[j] = -kInf;
                w.child[j] = 0x7FFFFFFF;
            }
        }
        out.nodes[size_t(t.wide)] = w;
    }
}

}  // namespace omb
