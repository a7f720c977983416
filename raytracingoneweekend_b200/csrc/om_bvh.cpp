#include "om_bvh.h"

#include <math.h>

#include <algorithm>
#include <limits>

namespace omb {
namespace {

constexpr int kBins = 16;
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
    std::vector<Item> items;
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
        const size_t mid = split(b, e);
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

void build(const std::vector<uint32_t>& refs, const std::vector<Aabb>& boxes, BuildResult& out) {
    out.nodes.clear();
    out.depth = 0;
    const size_t n = refs.size();
    if (n == 0) return;
    out.nodes.reserve(n);
    Builder bld{refs, boxes, {}, out.nodes};
    bld.items.resize(n);
    for (size_t i = 0; i < n; i++) {
        bld.items[i].prim = uint32_t(i);
        for (int a = 0; a < 3; a++) bld.items[i].c[a] = 0.5f * (boxes[i].lo[a] + boxes[i].hi[a]);
    }
    if (n == 1) {  // a root with one real child and one empty child
        Node root;
        for (int a = 0; a < 3; a++) { root.lmin[a] = boxes[0].lo[a]; root.lmax[a] = boxes[0].hi[a]; root.rmin[a] = kInf; root.rmax[a] = -kInf; }
        root.left = ~int32_t(refs[0]); root.right = ~int32_t(refs[0]); root.pad0 = root.pad1 = 0;
        out.nodes.push_back(root);
        out.depth = 1;
        return;
    }
    bld.build_range(0, n, 1);
    out.depth = bld.max_depth;
}

}  // namespace omb
