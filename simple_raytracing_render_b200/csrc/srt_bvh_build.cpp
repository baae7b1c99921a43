// srt_bvh_build.cpp — see srt_bvh_build.h.
#include "srt_bvh_build.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <future>
#include <thread>

namespace {

struct Box {
    float lo[3], hi[3];
    void reset() {
        lo[0] = lo[1] = lo[2] = FLT_MAX;
        hi[0] = hi[1] = hi[2] = -FLT_MAX;
    }
    void grow(const float* b) {
        for (int a = 0; a < 3; a++) {
            lo[a] = std::min(lo[a], b[a]);
            hi[a] = std::max(hi[a], b[3 + a]);
        }
    }
    void grow(const Box& b) {
        for (int a = 0; a < 3; a++) {
            lo[a] = std::min(lo[a], b.lo[a]);
            hi[a] = std::max(hi[a], b.hi[a]);
        }
    }
    float half_area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (dx < 0 || dy < 0 || dz < 0) return 0.f;
        return dx * dy + dy * dz + dz * dx;
    }
};

struct BinNode {
    Box box;
    int32_t left, right;   // inner: child node indices; leaf: left = -1
    int32_t first, count;  // leaf: range in `order`
};

constexpr int kBins = 16;
constexpr int kMaxLeaf = 4;
constexpr float kTraversalCost = 1.0f;
constexpr float kIntersectCost = 1.0f;

struct Builder {
    const float* boxes;
    std::vector<int32_t>& order;
    std::vector<BinNode> nodes;
    std::atomic<int32_t> next_node{0};
    std::atomic<int> live_tasks{0};
    int max_tasks;

    Builder(const float* b, int n, std::vector<int32_t>& ord) : boxes(b), order(ord) {
        nodes.resize(std::max(1, 2 * n));
        max_tasks = (int)std::max(1u, std::thread::hardware_concurrency());
    }

    int32_t alloc() { return next_node.fetch_add(1); }

    void centroid(int item, float* c) const {
        const float* b = boxes + 6 * (size_t)item;
        for (int a = 0; a < 3; a++) c[a] = 0.5f * (b[a] + b[3 + a]);
    }

    int32_t build(int first, int count) {
        int32_t me = alloc();
        BinNode& nd = nodes[me];
        nd.box.reset();
        Box cb;
        cb.reset();
        for (int i = first; i < first + count; i++) {
            const float* b = boxes + 6 * (size_t)order[i];
            nd.box.grow(b);
            float c[3];
            centroid(order[i], c);
            float cc[6] = {c[0], c[1], c[2], c[0], c[1], c[2]};
            cb.grow(cc);
        }
        nd.left = nd.right = -1;
        nd.first = first;
        nd.count = count;
        if (count <= 1) return me;

        // binned SAH over the three axes
        float best_cost = FLT_MAX;
        int best_axis = -1, best_split = -1;
        float parent_area = nd.box.half_area();
        for (int axis = 0; axis < 3; axis++) {
            float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0.f)) continue;
            Box bin_box[kBins];
            int bin_cnt[kBins];
            for (int b = 0; b < kBins; b++) {
                bin_box[b].reset();
                bin_cnt[b] = 0;
            }
            float scale = (float)kBins / ext;
            for (int i = first; i < first + count; i++) {
                float c[3];
                centroid(order[i], c);
                int b = (int)((c[axis] - cb.lo[axis]) * scale);
                b = std::min(kBins - 1, std::max(0, b));
                bin_cnt[b]++;
                bin_box[b].grow(boxes + 6 * (size_t)order[i]);
            }
            float right_area[kBins];
            int right_cnt[kBins];
            Box acc;
            acc.reset();
            int cnt = 0;
            for (int b = kBins - 1; b > 0; b--) {
                acc.grow(bin_box[b]);
                cnt += bin_cnt[b];
                right_area[b] = acc.half_area();
                right_cnt[b] = cnt;
            }
            acc.reset();
            cnt = 0;
            for (int b = 0; b < kBins - 1; b++) {
                acc.grow(bin_box[b]);
                cnt += bin_cnt[b];
                if (cnt == 0 || right_cnt[b + 1] == 0) continue;
                float cost = acc.half_area() * cnt + right_area[b + 1] * right_cnt[b + 1];
                if (cost < best_cost) {
                    best_cost = cost;
                    best_axis = axis;
                    best_split = b;
                }
            }
        }
        int mid = -1;
        if (best_axis >= 0) {
            float split_cost = kTraversalCost + kIntersectCost * best_cost / std::max(parent_area, 1e-30f);
            float leaf_cost = kIntersectCost * count;
            if (count <= kMaxLeaf && leaf_cost <= split_cost) return me;
            float ext = cb.hi[best_axis] - cb.lo[best_axis];
            float scale = (float)kBins / ext;
            float lo = cb.lo[best_axis];
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](int32_t item) {
                float c[3];
                centroid(item, c);
                int b = (int)((c[best_axis] - lo) * scale);
                b = std::min(kBins - 1, std::max(0, b));
                return b <= best_split;
            });
            mid = (int)(it - order.begin());
        }
        if (mid <= first || mid >= first + count) {
            if (count <= kMaxLeaf) return me;
            // all centroids coincide (or binning degenerated): split by count
            mid = first + count / 2;
        }
        int lc = mid - first, rc = count - lc;
        int32_t l, r;
        bool spawn = lc > 65536 && rc > 65536 && live_tasks.load() < max_tasks;
        if (spawn) {
            live_tasks.fetch_add(1);
            auto fut = std::async(std::launch::async, [&, first, lc] { return build(first, lc); });
            r = build(mid, rc);
            l = fut.get();
            live_tasks.fetch_sub(1);
        } else {
            l = build(first, lc);
            r = build(mid, rc);
        }
        BinNode& nd2 = nodes[me];
        nd2.left = l;
        nd2.right = r;
        return me;
    }
};

struct Collapser {
    const std::vector<BinNode>& bn;
    std::vector<WideNode4>& out;
    double sah = 0.0;
    float root_area;

    Collapser(const std::vector<BinNode>& b, std::vector<WideNode4>& o) : bn(b), out(o) {}

    static void set_slot(WideNode4& w, int i, const Box* b, int32_t child, int32_t count) {
        float* lox = &w.lox.x;
        float* loy = &w.loy.x;
        float* loz = &w.loz.x;
        float* hix = &w.hix.x;
        float* hiy = &w.hiy.x;
        float* hiz = &w.hiz.x;
        if (b) {
            lox[i] = b->lo[0];
            loy[i] = b->lo[1];
            loz[i] = b->lo[2];
            hix[i] = b->hi[0];
            hiy[i] = b->hi[1];
            hiz[i] = b->hi[2];
        } else {  // inverted: never hit
            lox[i] = loy[i] = loz[i] = FLT_MAX;
            hix[i] = hiy[i] = hiz[i] = -FLT_MAX;
        }
        (&w.child.x)[i] = child;
        (&w.ref.x)[i] = child >= 0 ? child : (count == 0 ? (int32_t)0x80000000 : ~(((~child) << 3) | count));
    }

    // emit a wide node for binary node `b` (must be inner, or the single-leaf root)
    int32_t emit(int32_t b) {
        int32_t me = (int32_t)out.size();
        out.emplace_back();
        int32_t kids[4];
        int nk = 0;
        if (bn[b].left < 0) {
            kids[nk++] = b;
        } else {
            kids[nk++] = bn[b].left;
            kids[nk++] = bn[b].right;
            while (nk < 4) {  // open the inner child with the largest area
                int pick = -1;
                float pa = -1.f;
                for (int i = 0; i < nk; i++)
                    if (bn[kids[i]].left >= 0) {
                        float a = bn[kids[i]].box.half_area();
                        if (a > pa) {
                            pa = a;
                            pick = i;
                        }
                    }
                if (pick < 0) break;
                int32_t k = kids[pick];
                kids[pick] = bn[k].left;
                kids[nk++] = bn[k].right;
            }
        }
        sah += kTraversalCost * bn[b].box.half_area() / root_area;
        for (int i = 0; i < 4; i++) {
            if (i >= nk) {
                set_slot(out[me], i, nullptr, -1, 0);
                continue;
            }
            const BinNode& k = bn[kids[i]];
            if (k.left < 0) {
                set_slot(out[me], i, &k.box, ~k.first, k.count);
                sah += kIntersectCost * k.count * k.box.half_area() / root_area;
            } else {
                int32_t c = emit(kids[i]);
                set_slot(out[me], i, &k.box, c, 0);
            }
        }
        return me;
    }
};

}  // namespace

void build_wide_bvh(const float* boxes, int n, WideBvh& out) {
    out.order.resize(n);
    for (int i = 0; i < n; i++) out.order[i] = i;
    Builder b(boxes, n, out.order);
    b.build(0, n);
    out.nodes.clear();
    out.nodes.reserve((size_t)n / 2 + 4);
    Collapser c(b.nodes, out.nodes);
    c.root_area = std::max(b.nodes[0].box.half_area(), 1e-30f);
    c.emit(0);
    out.sah_cost = (float)c.sah;
}

namespace {
int32_t ref_build(const float* ib, int lo, int n, std::vector<RefNode>& out) {
    int32_t me = (int32_t)out.size();
    out.emplace_back();
    RefNode nd;
    std::memset(&nd, 0, sizeof(nd));
    nd.left = nd.right = -1;
    nd.mid = lo;
    if (n <= 2) {
        // n == 1: left = right = l[0]; n == 2: left = l[0], right = l[1] (bvh.h:104-110)
        const float* a = ib + 6 * (size_t)lo;
        const float* b = ib + 6 * (size_t)(lo + n - 1);
        for (int k = 0; k < 3; k++) {
            nd.lo[k] = a[k] < b[k] ? a[k] : b[k];               // ffmin, aabb.h:8
            nd.hi[k] = a[3 + k] > b[3 + k] ? a[3 + k] : b[3 + k];  // ffmax, aabb.h:9
        }
        out[me] = nd;
        return me;
    }
    int32_t l = ref_build(ib, lo, n / 2, out);
    int32_t r = ref_build(ib, lo + n / 2, n - n / 2, out);
    nd.left = l;
    nd.right = r;
    nd.mid = lo + n / 2;
    for (int k = 0; k < 3; k++) {
        float a = out[l].lo[k], b = out[r].lo[k];
        nd.lo[k] = a < b ? a : b;
        a = out[l].hi[k];
        b = out[r].hi[k];
        nd.hi[k] = a > b ? a : b;
    }
    out[me] = nd;
    return me;
}
}  // namespace

void build_ref_tree(const float* item_boxes, int n, std::vector<RefNode>& out) {
    out.clear();
    out.reserve((size_t)n);
    ref_build(item_boxes, 0, n, out);
}
