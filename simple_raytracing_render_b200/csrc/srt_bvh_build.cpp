// srt_bvh_build.cpp — see srt_bvh_build.h.
#include "srt_bvh_build.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <future>
#include <map>
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
    std::vector<float> cent;  // [n][3] box centroids, computed once
    std::atomic<int32_t> next_node{0};
    std::atomic<int> live_tasks{0};
    int max_tasks;
    int max_leaf = kMaxLeaf;

    Builder(const float* b, int n, std::vector<int32_t>& ord) : boxes(b), order(ord) {
        nodes.resize(std::max(1, 2 * n));
        max_tasks = (int)std::max(1u, std::thread::hardware_concurrency());
        cent.resize(3 * (size_t)n);
        for_chunks(0, n, [&](int f, int c, int) {
            for (int i = f; i < f + c; i++) {
                const float* bx = boxes + 6 * (size_t)i;
                for (int a = 0; a < 3; a++) cent[3 * (size_t)i + a] = 0.5f * (bx[a] + bx[3 + a]);
            }
        });
    }

    int32_t alloc() { return next_node.fetch_add(1); }

    // how many threads a pass over `count` items is worth (the top levels of the tree only)
    int chunks_for(int count) const { return count >= (1 << 15) ? std::min(max_tasks, std::max(1, count >> 13)) : 1; }

    // f(first, count, chunk_index) over [first, first + count) split into chunks_for(count) pieces
    template <class F>
    void for_chunks(int first, int count, F f) {
        const int T = chunks_for(count);
        if (T <= 1) {
            f(first, count, 0);
            return;
        }
        std::vector<std::thread> th;
        th.reserve(T - 1);
        const int per = (count + T - 1) / T;
        for (int t = 1; t < T; t++) {
            int lo = first + t * per, n = std::min(per, first + count - lo);
            if (n > 0) th.emplace_back([=, &f] { f(lo, n, t); });
        }
        f(first, std::min(per, count), 0);
        for (auto& x : th) x.join();
    }

    struct Bins {
        Box box[3][kBins];
        int cnt[3][kBins];
        void reset() {
            for (int a = 0; a < 3; a++)
                for (int b = 0; b < kBins; b++) {
                    box[a][b].reset();
                    cnt[a][b] = 0;
                }
        }
    };

    int32_t build(int first, int count) {
        int32_t me = alloc();
        BinNode& nd = nodes[me];
        // pass 1: node box and centroid box (per-chunk partial boxes, merged: min / max are order-free)
        const int T = chunks_for(count);
        Box part_one[2];
        std::vector<Box> part_many;
        if (T > 1) part_many.resize(2 * (size_t)T);
        Box* part = T > 1 ? part_many.data() : part_one;  // no heap traffic in the ~n small nodes
        for (int t = 0; t < 2 * T; t++) part[t].reset();
        for_chunks(first, count, [&](int f, int c, int t) {
            Box nb, cbx;
            nb.reset();
            cbx.reset();
            for (int i = f; i < f + c; i++) {
                const int item = order[i];
                nb.grow(boxes + 6 * (size_t)item);
                const float* cc = &cent[3 * (size_t)item];
                for (int a = 0; a < 3; a++) {
                    cbx.lo[a] = std::min(cbx.lo[a], cc[a]);
                    cbx.hi[a] = std::max(cbx.hi[a], cc[a]);
                }
            }
            part[2 * t] = nb;
            part[2 * t + 1] = cbx;
        });
        nd.box.reset();
        Box cb;
        cb.reset();
        for (int t = 0; t < T; t++) {
            nd.box.grow(part[2 * t]);
            cb.grow(part[2 * t + 1]);
        }
        nd.left = nd.right = -1;
        nd.first = first;
        nd.count = count;
        if (count <= 1) return me;

        // pass 2: binned SAH, the three axes in one sweep over the items
        float ext[3], scale[3];
        bool use_axis[3];
        for (int axis = 0; axis < 3; axis++) {
            ext[axis] = cb.hi[axis] - cb.lo[axis];
            use_axis[axis] = ext[axis] > 0.f;
            scale[axis] = use_axis[axis] ? (float)kBins / ext[axis] : 0.f;
        }
        Bins bins_one;
        std::vector<Bins> bins_many;
        if (T > 1) bins_many.resize((size_t)T);
        Bins* bins = T > 1 ? bins_many.data() : &bins_one;
        for_chunks(first, count, [&](int f, int c, int t) {
            Bins& B = bins[t];
            B.reset();
            for (int i = f; i < f + c; i++) {
                const int item = order[i];
                const float* bx = boxes + 6 * (size_t)item;
                const float* cc = &cent[3 * (size_t)item];
                for (int axis = 0; axis < 3; axis++) {
                    if (!use_axis[axis]) continue;
                    int b = (int)((cc[axis] - cb.lo[axis]) * scale[axis]);
                    b = std::min(kBins - 1, std::max(0, b));
                    B.cnt[axis][b]++;
                    B.box[axis][b].grow(bx);
                }
            }
        });
        for (int t = 1; t < T; t++)
            for (int axis = 0; axis < 3; axis++)
                for (int b = 0; b < kBins; b++) {
                    bins[0].cnt[axis][b] += bins[t].cnt[axis][b];
                    bins[0].box[axis][b].grow(bins[t].box[axis][b]);
                }
        float best_cost = FLT_MAX;
        int best_axis = -1, best_split = -1;
        float parent_area = nd.box.half_area();
        for (int axis = 0; axis < 3; axis++) {
            if (!use_axis[axis]) continue;
            const Box* bin_box = bins[0].box[axis];
            const int* bin_cnt = bins[0].cnt[axis];
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
            if (count <= max_leaf && leaf_cost <= split_cost) return me;
            const float bscale = scale[best_axis];
            const float lo = cb.lo[best_axis];
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](int32_t item) {
                int b = (int)((cent[3 * (size_t)item + best_axis] - lo) * bscale);
                b = std::min(kBins - 1, std::max(0, b));
                return b <= best_split;
            });
            mid = (int)(it - order.begin());
        }
        if (mid <= first || mid >= first + count) {
            if (count <= max_leaf) return me;
            // all centroids coincide (or binning degenerated): split by count
            mid = first + count / 2;
        }
        int lc = mid - first, rc = count - lc;
        int32_t l, r;
        bool spawn = lc > 512 && rc > 512 && live_tasks.load() < max_tasks;
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
    int max_depth = 0;

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
    int32_t emit(int32_t b, int depth = 1) {
        max_depth = std::max(max_depth, depth);
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
                int32_t c = emit(kids[i], depth + 1);
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
    auto T0 = std::chrono::steady_clock::now();
    Builder b(boxes, n, out.order);
    auto T1 = std::chrono::steady_clock::now();
    b.build(0, n);
    auto T2 = std::chrono::steady_clock::now();
    out.nodes.clear();
    out.nodes.reserve((size_t)n / 2 + 4);
    Collapser c(b.nodes, out.nodes);
    c.root_area = std::max(b.nodes[0].box.half_area(), 1e-30f);
    c.emit(0);
    out.sah_cost = (float)c.sah;
    out.max_depth = c.max_depth;
    if (getenv("SRT_SCENE_TIMING")) {
        auto T3 = std::chrono::steady_clock::now();
        auto ms = [](auto a, auto b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        fprintf(stderr, "[build_wide_bvh] n %d: setup %.1f ms, binary SAH build %.1f ms, collapse %.1f ms\n", n, ms(T0, T1),
                ms(T1, T2), ms(T2, T3));
    }
}

void build_binary_sah(const float* boxes, int n, int max_leaf, std::vector<SahBinNode>& nodes, std::vector<int32_t>& order) {
    order.resize(n);
    for (int i = 0; i < n; i++) order[i] = i;
    Builder b(boxes, n, order);
    b.max_leaf = std::max(1, max_leaf);
    b.build(0, n);
    const int used = b.next_node.load();
    nodes.resize((size_t)used);
    for (int k = 0; k < used; k++) {
        const BinNode& s = b.nodes[(size_t)k];
        SahBinNode& o = nodes[(size_t)k];
        for (int a = 0; a < 3; a++) {
            o.lo[a] = s.box.lo[a];
            o.hi[a] = s.box.hi[a];
        }
        o.left = s.left;
        o.right = s.right;
        o.first = s.first;
        o.count = s.count;
    }
}

namespace {
// nodes of the reference subtree over n leaves (one node for n <= 2); memoised, a level holds two sizes at most
struct NodeCount {
    std::map<long long, long long> memo;
    long long fill(long long n) {
        if (n <= 2) return 1;
        auto it = memo.find(n);
        if (it != memo.end()) return it->second;
        long long v = 1 + fill(n / 2) + fill(n - n / 2);
        memo[n] = v;
        return v;
    }
    long long operator()(long long n) const { return n <= 2 ? 1 : memo.at(n); }  // after fill(root): read-only
};

// Node indices are the depth-first numbers, known before the subtree is built (left = me + 1,
// right = me + 1 + nodes(left)), so the upper levels fork into independent tasks.
struct RefTree {
    const float* ib;
    RefNode* out;
    NodeCount nodes;

    void build(int lo, int n, int32_t me, int forks) {
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
            return;
        }
        const int nl = n / 2;
        const int32_t l = me + 1, r = me + 1 + (int32_t)nodes(nl);
        if (forks > 0 && n >= (1 << 15)) {
            auto other = std::async(std::launch::async, [&] { build(lo, nl, l, forks - 1); });
            build(lo + nl, n - nl, r, forks - 1);
            other.get();
        } else {
            build(lo, nl, l, 0);
            build(lo + nl, n - nl, r, 0);
        }
        nd.left = l;
        nd.right = r;
        nd.mid = lo + nl;
        for (int k = 0; k < 3; k++) {
            float a = out[l].lo[k], b = out[r].lo[k];
            nd.lo[k] = a < b ? a : b;
            a = out[l].hi[k];
            b = out[r].hi[k];
            nd.hi[k] = a > b ? a : b;
        }
        out[me] = nd;
    }
};
}  // namespace

void build_ref_tree(const float* item_boxes, int n, std::vector<RefNode>& out) {
    RefTree t;
    t.ib = item_boxes;
    out.resize((size_t)t.nodes.fill(n));
    t.out = out.data();
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    int forks = 0;
    while ((1u << forks) < std::min(hw, 32u)) forks++;
    t.build(0, n, 0, forks);
}

int ref_tree_levels(long long n, int size_lo[32], int nodes_lo[32], int nodes_hi[32]) {
    NodeCount nc;
    int levels = 0;
    for (int L = 0; L < 32; L++) {
        const long long m = n >> L;
        size_lo[L] = (int)m;
        nodes_lo[L] = (int)nc.fill(m);
        nodes_hi[L] = (int)nc.fill(m + 1);
        // nodes exist at level L when the largest subtree one level up, ceil(n / 2^(L-1)), has more than two leaves
        if (L == 0 || ((n + (1ll << (L - 1)) - 1) >> (L - 1)) > 2) levels = L + 1;
    }
    return levels;
}

// ---- reference bvh_node order replay (bvh.h:96-119) ----------------------------------------------------------
namespace {
struct RefOrder {
    const float* mins;
    int32_t* order;
    std::vector<uint8_t> axis;  // per node, depth-first numbering
    std::atomic<int> live{0};
    int max_tasks;

    NodeCount node_count;

    void sort_slice(long long lo, long long n, int ax, std::vector<std::pair<float, int32_t>>& tmp) {
        tmp.resize((size_t)n);
        for (long long i = 0; i < n; i++) tmp[(size_t)i] = {mins[3 * (size_t)order[lo + i] + ax], (int32_t)i};
        // ascending keys, equal keys in reverse input order (the never-equal comparator under glibc's merge sort)
        std::sort(tmp.begin(), tmp.end(), [](const std::pair<float, int32_t>& a, const std::pair<float, int32_t>& b) {
            if (a.first < b.first) return true;
            if (b.first < a.first) return false;
            return a.second > b.second;
        });
        std::vector<int32_t> moved((size_t)n);
        for (long long i = 0; i < n; i++) moved[(size_t)i] = order[lo + tmp[(size_t)i].second];
        std::memcpy(order + lo, moved.data(), sizeof(int32_t) * (size_t)n);
    }

    void build(long long lo, long long n, long long node) {
        std::vector<std::pair<float, int32_t>> tmp;
        build_rec(lo, n, node, tmp);
    }

    void build_rec(long long lo, long long n, long long node, std::vector<std::pair<float, int32_t>>& tmp) {
        sort_slice(lo, n, axis[(size_t)node], tmp);
        if (n <= 2) return;
        const long long nl = n / 2, nr = n - nl;
        const long long left_node = node + 1, right_node = node + 1 + node_count(nl);
        if (n > (1 << 15) && live.load() < max_tasks) {
            live.fetch_add(1);
            auto fut = std::async(std::launch::async, [=] { build(lo, nl, left_node); });
            build_rec(lo + nl, nr, right_node, tmp);
            fut.get();
            live.fetch_sub(1);
        } else {
            build_rec(lo, nl, left_node, tmp);
            build_rec(lo + nl, nr, right_node, tmp);
        }
    }
};
}  // namespace

void ref_bvh_leaf_order(const float* mins, long long n, unsigned long long* seed, int32_t* order_out) {
    for (long long i = 0; i < n; i++) order_out[i] = (int32_t)i;
    if (n <= 0) return;
    RefOrder R;
    R.mins = mins;
    R.order = order_out;
    R.max_tasks = (int)std::max(1u, std::thread::hardware_concurrency());
    const long long nodes = R.node_count.fill(n);
    R.axis.resize((size_t)nodes);
    unsigned long long s = *seed & 0xFFFFFFFFFFFFull;
    for (long long k = 0; k < nodes; k++) {  // mathf.h:14-19; the constructor draws before it recurses
        s = (0x5DEECE66Dull * s + 0xB16ull) & 0xFFFFFFFFFFFFull;  // __a, __c of mathf.h:6-8
        const double r = (double)(s >> 16) / 4294967296.0;
        R.axis[(size_t)k] = (uint8_t)(int)(3 * r);
    }
    *seed = s;
    R.build(0, n, 0);
}
