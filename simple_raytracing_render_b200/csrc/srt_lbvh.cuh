// srt_lbvh.cuh — device-side build of the wide BVH for large triangle meshes.
//
// Replaces, for OUR acceleration structure, the reference's O(n log^2 n) qsort recursion (bvh.h:96-119) and
// the host binned-SAH builder of srt_bvh_build.cpp (4.7 s for the 10 M triangles of C5).  It decides nothing
// about results: like the host tree it only orders and prunes the search of bvh_hit (srt_intersect.cuh);
// leaf tests, reference ranks and the reference-tree replay are unchanged, so every hit-parity test holds.
//
// Pipeline (HLBVH: Pantaleoni & Luebke 2010, Garanzha et al. 2011; radix tree: Karras 2012):
//   1. k_item_boxes      one thread per item: padded box of its triangle, scene bounds by ordered-int atomics
//   2. k_morton          63-bit Morton code of the box centroid (21 bits per axis)
//   3. cub radix sort    (code, item) pairs — the one library call of the build (set-up code, not the hot path)
//   4. k_radix_tree      binary radix tree over the sorted codes (ties broken by position), one thread per
//                        internal node; k_fit: bottom-up boxes with one atomic flag per node
//   5. k_mark_clusters   the cut through the radix tree where subtrees drop to <= kClusterSize items:
//                        n / ~100 clusters, whose boxes go to the HOST binned-SAH builder (srt_bvh_build.cpp,
//                        max_leaf = 1: a few ms for 100 k clusters).  The upper levels of the tree, where
//                        split quality matters most, are therefore SAH-built; the radix tree supplies the
//                        lower levels, where Morton order is as good as it gets.
//   6. k_collapse        level-synchronous collapse of the joined binary tree to WideNode4 (srt_types.h):
//                        a wide node opens the child of largest surface area until it has four, subtrees
//                        of <= 4 items become leaves (leaf slots are positions in Morton order, so a leaf is
//                        a contiguous slot range, as the traversal expects)
//   7. k_leaf_arrays     leaf_rank[slot] = item (the reference in-order rank), leaf_geom[3 slot ..] = vertices
#ifndef SRT_LBVH_CUH
#define SRT_LBVH_CUH

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <chrono>
#include <string>
#include <vector>

#include "srt_bvh_build.h"
#include "srt_types.h"

namespace srt_lbvh {

constexpr int kMaxLeaf = 4;        // items per wide-BVH leaf (WideNode4::ref packs a count <= 7)
constexpr int kClusterSize = 128;  // radix subtrees of at most this many items go to the host SAH as one item

// floats ordered as unsigned integers (for atomicMin / atomicMax on bounds)
__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// same padding as the host builder's item boxes (srt_api.cu: pad_box)
__device__ __forceinline__ void pad_axis(float& lo, float& hi) {
    float m = fmaxf(fabsf(lo), fabsf(hi));
    float pad = 1e-4f * fmaxf(m, hi - lo) + 1e-7f;
    lo -= pad;
    hi += pad;
}

// tri_pos: [rows][9]; item_row[k] = triangle row of item k.  bounds: 6 ordered uints (min xyz, max xyz of centroids)
__global__ void k_item_boxes(const float* __restrict__ tri_pos, const int32_t* __restrict__ item_row, int n,
                             float4* __restrict__ blo, float4* __restrict__ bhi, unsigned* __restrict__ bounds) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float cx = 0.f, cy = 0.f, cz = 0.f;
    bool live = i < n;
    if (live) {
        const float* t = tri_pos + 9 * (size_t)item_row[i];
        float lx = fminf(t[0], fminf(t[3], t[6])), hx = fmaxf(t[0], fmaxf(t[3], t[6]));
        float ly = fminf(t[1], fminf(t[4], t[7])), hy = fmaxf(t[1], fmaxf(t[4], t[7]));
        float lz = fminf(t[2], fminf(t[5], t[8])), hz = fmaxf(t[2], fmaxf(t[5], t[8]));
        pad_axis(lx, hx);
        pad_axis(ly, hy);
        pad_axis(lz, hz);
        blo[i] = make_float4(lx, ly, lz, 0.f);
        bhi[i] = make_float4(hx, hy, hz, 0.f);
        cx = 0.5f * (lx + hx);
        cy = 0.5f * (ly + hy);
        cz = 0.5f * (lz + hz);
    }
    // centroid bounds: warp reduce, one atomic per warp and component
    const unsigned FULL = 0xffffffffu;
    float mnx = live ? cx : 3.4e38f, mny = live ? cy : 3.4e38f, mnz = live ? cz : 3.4e38f;
    float mxx = live ? cx : -3.4e38f, mxy = live ? cy : -3.4e38f, mxz = live ? cz : -3.4e38f;
    for (int o = 16; o > 0; o >>= 1) {
        mnx = fminf(mnx, __shfl_xor_sync(FULL, mnx, o));
        mny = fminf(mny, __shfl_xor_sync(FULL, mny, o));
        mnz = fminf(mnz, __shfl_xor_sync(FULL, mnz, o));
        mxx = fmaxf(mxx, __shfl_xor_sync(FULL, mxx, o));
        mxy = fmaxf(mxy, __shfl_xor_sync(FULL, mxy, o));
        mxz = fmaxf(mxz, __shfl_xor_sync(FULL, mxz, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(bounds + 0, f2ord(mnx));
        atomicMin(bounds + 1, f2ord(mny));
        atomicMin(bounds + 2, f2ord(mnz));
        atomicMax(bounds + 3, f2ord(mxx));
        atomicMax(bounds + 4, f2ord(mxy));
        atomicMax(bounds + 5, f2ord(mxz));
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long v) {  // 21 bits -> every third bit
    v &= 0x1fffffull;
    v = (v | (v << 32)) & 0x1f00000000ffffull;
    v = (v | (v << 16)) & 0x1f0000ff0000ffull;
    v = (v | (v << 8)) & 0x100f00f00f00f00full;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}

__global__ void k_morton(const float4* __restrict__ blo, const float4* __restrict__ bhi, int n,
                         const unsigned* __restrict__ bounds, unsigned long long* __restrict__ keys,
                         int32_t* __restrict__ vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lo[3] = {ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2])};
    float hi[3] = {ord2f(bounds[3]), ord2f(bounds[4]), ord2f(bounds[5])};
    float4 a = blo[i], b = bhi[i];
    float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
    // one cell size for the three axes (the largest extent): cubical cells keep a flat scene (a terrain) from
    // being cut into slabs along its thin axis, which per-axis normalisation does
    const float e = fmaxf(fmaxf(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
    unsigned long long q[3];
    for (int k = 0; k < 3; k++) {
        float u = e > 0.f ? (c[k] - lo[k]) / e : 0.f;
        u = fminf(fmaxf(u, 0.f), 1.f);
        q[k] = (unsigned long long)fminf(u * 2097152.0f, 2097151.0f);
    }
    keys[i] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
    vals[i] = i;
}

// ---- binary radix tree (Karras 2012) ----------------------------------------------------------------------
// Internal node i in [0, n-1); a child reference >= 0 is an internal node, < 0 is the leaf at sorted position ~ref.
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned long long a = keys[i], b = keys[j];
    if (a != b) return __clzll((long long)(a ^ b));
    return 64 + __clz(i ^ j);  // equal codes: the position breaks the tie
}

__global__ void k_radix_tree(const unsigned long long* __restrict__ keys, int n, int32_t* __restrict__ left,
                             int32_t* __restrict__ right, int32_t* __restrict__ parent, int32_t* __restrict__ first,
                             int32_t* __restrict__ last) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int lc = (lo == gamma) ? ~gamma : gamma;
    int rc = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    left[i] = lc;
    right[i] = rc;
    first[i] = lo;
    last[i] = hi;
    if (lc >= 0) parent[lc] = i; else parent[n - 1 + gamma] = i;          // leaf parents live behind the inner ones
    if (rc >= 0) parent[rc] = i; else parent[n - 1 + gamma + 1] = i;
    if (i == 0) parent[0] = -1;
}

// bottom-up boxes: each leaf climbs; the second arrival at a node (atomic flag) computes its box and goes on
__global__ void k_fit(int n, const int32_t* __restrict__ vals, const float4* __restrict__ ilo, const float4* __restrict__ ihi,
                      const int32_t* __restrict__ left, const int32_t* __restrict__ right, const int32_t* __restrict__ parent,
                      float4* __restrict__ nlo, float4* __restrict__ nhi, int* __restrict__ flags) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int node = parent[n - 1 + j];
    while (node >= 0) {
        if (atomicAdd(flags + node, 1) == 0) return;  // the sibling subtree is not finished: it will continue
        __threadfence();
        int lc = left[node], rc = right[node];
        float4 al = lc >= 0 ? nlo[lc] : ilo[vals[~lc]], ah = lc >= 0 ? nhi[lc] : ihi[vals[~lc]];
        float4 bl = rc >= 0 ? nlo[rc] : ilo[vals[~rc]], bh = rc >= 0 ? nhi[rc] : ihi[vals[~rc]];
        nlo[node] = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0.f);
        nhi[node] = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.f);
        __threadfence();
        node = parent[node];
    }
}

// ---- the cut handed to the host SAH builder -----------------------------------------------------------------
// A child of an internal node larger than kClusterSize that is itself at most kClusterSize items (or a single
// leaf) is a cluster root.  out_ref[k] = its reference, out_box[6 k ..] = its box.
__global__ void k_mark_clusters(int n, const int32_t* __restrict__ vals, const float4* __restrict__ ilo,
                                const float4* __restrict__ ihi, const int32_t* __restrict__ left,
                                const int32_t* __restrict__ right, const int32_t* __restrict__ first,
                                const int32_t* __restrict__ last, const float4* __restrict__ nlo,
                                const float4* __restrict__ nhi, int32_t* __restrict__ out_ref, float* __restrict__ out_box,
                                int* __restrict__ counter) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    if (last[i] - first[i] + 1 <= kClusterSize) return;
    for (int side = 0; side < 2; side++) {
        int c = side ? right[i] : left[i];
        int size = c >= 0 ? last[c] - first[c] + 1 : 1;
        if (size > kClusterSize) continue;
        float4 lo = c >= 0 ? nlo[c] : ilo[vals[~c]], hi = c >= 0 ? nhi[c] : ihi[vals[~c]];
        int k = atomicAdd(counter, 1);
        out_ref[k] = c;
        float* b = out_box + 6 * (size_t)k;
        b[0] = lo.x;
        b[1] = lo.y;
        b[2] = lo.z;
        b[3] = hi.x;
        b[4] = hi.y;
        b[5] = hi.z;
    }
}

// ---- per-cluster binned SAH (one warp per cluster) ------------------------------------------------------------
// The radix subtree below a cluster root is REPLACED by a binned-SAH tree over the same <= kClusterSize items,
// built in shared memory with the rules of the host builder (16 bins, three axes, leaf when <= kMaxLeaf items
// cost less than a split).  The subtree keeps its node ids (a radix subtree over slots [f, l] owns the internal
// ids f .. l-1; the root keeps its own) and its slot range; items are permuted inside the range so that every
// node covers contiguous slots.  A leaf node is marked with left == right == kLeafMark.
constexpr int32_t kLeafMark = (int32_t)0x80000001;
constexpr int kSahBins = 16;

struct ClusterSmem {
    float lo[3][kClusterSize], hi[3][kClusterSize], cen[3][kClusterSize];
    int32_t item[kClusterSize];
    unsigned char perm[kClusterSize], tmp[kClusterSize];
    unsigned bins[3][kSahBins][6];
    int bcnt[3][kSahBins];
    int stack[kClusterSize][3];  // first, count, node
};

__device__ __forceinline__ float warp_min(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(128) k_cluster_sah(int n_clusters, const int32_t* __restrict__ c_ref, int32_t* __restrict__ vals,
                                                     const float4* __restrict__ ilo, const float4* __restrict__ ihi,
                                                     int32_t* __restrict__ left, int32_t* __restrict__ right,
                                                     int32_t* __restrict__ first, int32_t* __restrict__ last,
                                                     float4* __restrict__ nlo, float4* __restrict__ nhi,
                                                     int* __restrict__ n_split_total) {
    extern __shared__ unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    ClusterSmem& S = reinterpret_cast<ClusterSmem*>(smem_raw)[warp];
    const int ci = blockIdx.x * (blockDim.x >> 5) + warp;
    if (ci >= n_clusters) return;
    const int32_t root = c_ref[ci];
    if (root < 0) return;  // a single item: nothing to build
    const int f = first[root], l = last[root], m = l - f + 1;
    for (int k = lane; k < m; k += 32) {
        int it = vals[f + k];
        float4 a = ilo[it], b = ihi[it];
        S.item[k] = it;
        S.lo[0][k] = a.x; S.lo[1][k] = a.y; S.lo[2][k] = a.z;
        S.hi[0][k] = b.x; S.hi[1][k] = b.y; S.hi[2][k] = b.z;
        S.cen[0][k] = 0.5f * (a.x + b.x); S.cen[1][k] = 0.5f * (a.y + b.y); S.cen[2][k] = 0.5f * (a.z + b.z);
        S.perm[k] = (unsigned char)k;
    }
    __syncwarp();
    int next_id = f;  // ids f .. l-1 belong to this subtree; the root keeps `root`
    int sp = 0, n_split = 0;
    if (lane == 0) {
        S.stack[0][0] = 0;
        S.stack[0][1] = m;
        S.stack[0][2] = root;
    }
    sp = 1;
    __syncwarp();
    while (sp > 0) {
        sp--;
        const int a = S.stack[sp][0], cnt = S.stack[sp][1], node = S.stack[sp][2];
        __syncwarp();
        // node box and centroid box
        float nl[3] = {3.4e38f, 3.4e38f, 3.4e38f}, nh[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
        float cl[3] = {3.4e38f, 3.4e38f, 3.4e38f}, ch[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
        for (int k = a + lane; k < a + cnt; k += 32) {
            int p = S.perm[k];
#pragma unroll
            for (int ax = 0; ax < 3; ax++) {
                nl[ax] = fminf(nl[ax], S.lo[ax][p]);
                nh[ax] = fmaxf(nh[ax], S.hi[ax][p]);
                cl[ax] = fminf(cl[ax], S.cen[ax][p]);
                ch[ax] = fmaxf(ch[ax], S.cen[ax][p]);
            }
        }
#pragma unroll
        for (int ax = 0; ax < 3; ax++) {
            nl[ax] = warp_min(nl[ax]);
            nh[ax] = warp_max(nh[ax]);
            cl[ax] = warp_min(cl[ax]);
            ch[ax] = warp_max(ch[ax]);
        }
        if (lane == 0) {
            nlo[node] = make_float4(nl[0], nl[1], nl[2], 0.f);
            nhi[node] = make_float4(nh[0], nh[1], nh[2], 0.f);
            first[node] = f + a;
            last[node] = f + a + cnt - 1;
        }
        // binning, the three axes at once
        for (int k = lane; k < 3 * kSahBins; k += 32) {
            int ax = k / kSahBins, b = k % kSahBins;
            S.bcnt[ax][b] = 0;
#pragma unroll
            for (int c = 0; c < 3; c++) {
                S.bins[ax][b][c] = 0xffffffffu;
                S.bins[ax][b][3 + c] = 0u;
            }
        }
        __syncwarp();
        float scale[3];
#pragma unroll
        for (int ax = 0; ax < 3; ax++) scale[ax] = ch[ax] > cl[ax] ? (float)kSahBins / (ch[ax] - cl[ax]) : 0.f;
        for (int k = a + lane; k < a + cnt; k += 32) {
            int p = S.perm[k];
#pragma unroll
            for (int ax = 0; ax < 3; ax++) {
                if (scale[ax] == 0.f) continue;
                int b = min(kSahBins - 1, max(0, (int)((S.cen[ax][p] - cl[ax]) * scale[ax])));
                atomicAdd(&S.bcnt[ax][b], 1);
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    atomicMin(&S.bins[ax][b][c], f2ord(S.lo[c][p]));
                    atomicMax(&S.bins[ax][b][3 + c], f2ord(S.hi[c][p]));
                }
            }
        }
        __syncwarp();
        // candidates: 3 axes x (kSahBins - 1) split planes, spread over the lanes
        float best = 3.4e38f;
        int best_key = -1;
        for (int cand = lane; cand < 3 * (kSahBins - 1); cand += 32) {
            int ax = cand / (kSahBins - 1), sp_ = cand % (kSahBins - 1);
            if (scale[ax] == 0.f) continue;
            float l0[3] = {3.4e38f, 3.4e38f, 3.4e38f}, h0[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
            float l1[3] = {3.4e38f, 3.4e38f, 3.4e38f}, h1[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
            int c0 = 0, c1 = 0;
            for (int b = 0; b < kSahBins; b++) {
                int c = S.bcnt[ax][b];
                if (c == 0) continue;
                if (b <= sp_) {
                    c0 += c;
                    for (int q = 0; q < 3; q++) {
                        l0[q] = fminf(l0[q], ord2f(S.bins[ax][b][q]));
                        h0[q] = fmaxf(h0[q], ord2f(S.bins[ax][b][3 + q]));
                    }
                } else {
                    c1 += c;
                    for (int q = 0; q < 3; q++) {
                        l1[q] = fminf(l1[q], ord2f(S.bins[ax][b][q]));
                        h1[q] = fmaxf(h1[q], ord2f(S.bins[ax][b][3 + q]));
                    }
                }
            }
            if (c0 == 0 || c1 == 0) continue;
            float a0 = (h0[0] - l0[0]) * (h0[1] - l0[1]) + (h0[1] - l0[1]) * (h0[2] - l0[2]) + (h0[2] - l0[2]) * (h0[0] - l0[0]);
            float a1 = (h1[0] - l1[0]) * (h1[1] - l1[1]) + (h1[1] - l1[1]) * (h1[2] - l1[2]) + (h1[2] - l1[2]) * (h1[0] - l1[0]);
            float cost = a0 * c0 + a1 * c1;
            if (cost < best) {
                best = cost;
                best_key = cand;
            }
        }
        for (int o = 16; o > 0; o >>= 1) {  // warp argmin (ties: smaller candidate number)
            float ob = __shfl_xor_sync(0xffffffffu, best, o);
            int ok = __shfl_xor_sync(0xffffffffu, best_key, o);
            if (ok >= 0 && (best_key < 0 || ob < best || (ob == best && ok < best_key))) {
                best = ob;
                best_key = ok;
            }
        }
        const float parea = fmaxf((nh[0] - nl[0]) * (nh[1] - nl[1]) + (nh[1] - nl[1]) * (nh[2] - nl[2]) +
                                      (nh[2] - nl[2]) * (nh[0] - nl[0]), 1e-30f);
        if (cnt <= kMaxLeaf && (best_key < 0 || (float)cnt <= 1.0f + best / parea)) {
            if (lane == 0) left[node] = right[node] = kLeafMark;
            continue;
        }
        int nl_cnt;
        if (best_key < 0) {
            nl_cnt = cnt / 2;  // all centroids coincide: split by count, order as it is
        } else {
            const int ax = best_key / (kSahBins - 1), sp_ = best_key % (kSahBins - 1);
            // stable partition of perm[a .. a+cnt) by bin <= split, 32 items per step
            int nleft = 0, nright = 0;
            for (int base = a; base < a + cnt; base += 32) {
                int k = base + lane;
                bool valid = k < a + cnt;
                int p = valid ? S.perm[k] : 0;
                bool goes_left = valid && min(kSahBins - 1, max(0, (int)((S.cen[ax][p] - cl[ax]) * scale[ax]))) <= sp_;
                unsigned ml = __ballot_sync(0xffffffffu, goes_left), mv = __ballot_sync(0xffffffffu, valid);
                unsigned mr = mv & ~ml;
                unsigned lt = (1u << lane) - 1u;
                if (goes_left) S.tmp[nleft + __popc(ml & lt)] = (unsigned char)p;
                else if (valid) S.tmp[kClusterSize - 1 - (nright + __popc(mr & lt))] = (unsigned char)p;  // right part, reversed
                nleft += __popc(ml);
                nright += __popc(mr);
            }
            __syncwarp();
            for (int k = lane; k < cnt; k += 32)
                S.perm[a + k] = k < nleft ? S.tmp[k] : S.tmp[kClusterSize - 1 - (k - nleft)];
            __syncwarp();
            nl_cnt = nleft;
        }
        const int nr_cnt = cnt - nl_cnt;
        n_split++;
        int lref, rref;
        if (nl_cnt == 1) {
            lref = ~(f + a);
        } else {
            lref = next_id++;
            if (lref == root) lref = next_id++;
        }
        if (nr_cnt == 1) {
            rref = ~(f + a + nl_cnt);
        } else {
            rref = next_id++;
            if (rref == root) rref = next_id++;
        }
        if (lane == 0) {
            left[node] = lref;
            right[node] = rref;
            if (nr_cnt > 1) {
                S.stack[sp][0] = a + nl_cnt;
                S.stack[sp][1] = nr_cnt;
                S.stack[sp][2] = rref;
            }
            if (nl_cnt > 1) {
                int at = sp + (nr_cnt > 1 ? 1 : 0);
                S.stack[at][0] = a;
                S.stack[at][1] = nl_cnt;
                S.stack[at][2] = lref;
            }
        }
        sp += (nr_cnt > 1 ? 1 : 0) + (nl_cnt > 1 ? 1 : 0);
        __syncwarp();
    }
    for (int k = lane; k < m; k += 32) vals[f + k] = S.item[S.perm[k]];
    if (lane == 0 && n_split) atomicAdd(n_split_total, n_split);  // bounds the wide-node count
}

struct Tree {  // the joined binary tree: radix nodes [0, n-1), host-built top nodes behind them
    int n;
    const int32_t* vals;
    const float4 *ilo, *ihi;          // item boxes (item order)
    const int32_t *left, *right;      // per internal node
    const int32_t *first, *last;      // slot range (top nodes: 0 .. INT_MAX / 2, never a leaf bundle)
    const float4 *nlo, *nhi;          // internal node boxes
};

struct Kid {
    int32_t ref;
    float4 lo, hi;
    int first, count;  // count == 0: inner node
};

__device__ __forceinline__ Kid make_kid(const Tree& T, int32_t r) {
    Kid k;
    k.ref = r;
    if (r < 0) {
        int item = T.vals[~r];
        k.lo = T.ilo[item];
        k.hi = T.ihi[item];
        k.first = ~r;
        k.count = 1;
    } else {
        k.lo = T.nlo[r];
        k.hi = T.nhi[r];
        int size = T.last[r] - T.first[r] + 1;
        k.first = T.first[r];
        k.count = T.left[r] == kLeafMark ? size : 0;  // k_cluster_sah decided this subtree is one leaf
    }
    return k;
}
__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
    float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

// one thread per queued (binary node, wide node index); children that stay inner get the next free wide indices
__global__ void k_collapse(Tree T, const int2* __restrict__ q_in, int n_in, int2* __restrict__ q_out,
                           int* __restrict__ n_out, WideNode4* __restrict__ wide, int* __restrict__ wide_count,
                           float* __restrict__ sah, float inv_root_area) {
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_in) return;
    int2 item = q_in[w];
    Kid kids[4];
    int nk = 0;
    Kid self = make_kid(T, item.x);
    if (self.count > 0) {  // the whole tree is one leaf
        kids[nk++] = self;
    } else {
        kids[nk++] = make_kid(T, T.left[item.x]);
        kids[nk++] = make_kid(T, T.right[item.x]);
        while (nk < 4) {  // open the inner child of largest surface area
            int pick = -1;
            float pa = -1.f;
            for (int i = 0; i < nk; i++)
                if (kids[i].count == 0) {
                    float a = half_area(kids[i].lo, kids[i].hi);
                    if (a > pa) {
                        pa = a;
                        pick = i;
                    }
                }
            if (pick < 0) break;
            int32_t r = kids[pick].ref;
            kids[pick] = make_kid(T, T.left[r]);
            kids[nk++] = make_kid(T, T.right[r]);
        }
    }
    float cost = half_area(self.lo, self.hi) * inv_root_area;
    WideNode4 out;
    float* lox = &out.lox.x;
    float* loy = &out.loy.x;
    float* loz = &out.loz.x;
    float* hix = &out.hix.x;
    float* hiy = &out.hiy.x;
    float* hiz = &out.hiz.x;
    int* child = &out.child.x;
    int* ref = &out.ref.x;
    for (int i = 0; i < 4; i++) {
        if (i >= nk) {
            lox[i] = loy[i] = loz[i] = 3.402823466e+38f;
            hix[i] = hiy[i] = hiz[i] = -3.402823466e+38f;
            child[i] = -1;
            ref[i] = (int32_t)0x80000000;
            continue;
        }
        const Kid& k = kids[i];
        lox[i] = k.lo.x;
        loy[i] = k.lo.y;
        loz[i] = k.lo.z;
        hix[i] = k.hi.x;
        hiy[i] = k.hi.y;
        hiz[i] = k.hi.z;
        if (k.count > 0) {
            child[i] = ~k.first;
            ref[i] = ~((k.first << 3) | k.count);
            cost += (float)k.count * half_area(k.lo, k.hi) * inv_root_area;
        } else {
            int widx = atomicAdd(wide_count, 1);
            int slot = atomicAdd(n_out, 1);
            q_out[slot] = make_int2(k.ref, widx);
            child[i] = widx;
            ref[i] = widx;
        }
    }
    wide[item.y] = out;
    atomicAdd(sah, cost);
}

// leaf_rank[slot] = item (== the reference in-order rank); leaf_geom[3 slot + k] = vertex k of its triangle
__global__ void k_leaf_arrays(const float* __restrict__ tri_pos, const int32_t* __restrict__ item_row,
                              const int32_t* __restrict__ item_prim, const int32_t* __restrict__ vals, int n,
                              int32_t* __restrict__ leaf_rank, float4* __restrict__ leaf_geom) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int item = vals[s];
    leaf_rank[s] = item;
    const float* t = tri_pos + 9 * (size_t)item_row[item];
    // the spare words carry the rank and the primitive id (srt_types.h, DBvh)
    leaf_geom[3 * (size_t)s + 0] = make_float4(t[0], t[1], t[2], __int_as_float(item));
    leaf_geom[3 * (size_t)s + 1] = make_float4(t[3], t[4], t[5], __int_as_float(item_prim[item]));
    leaf_geom[3 * (size_t)s + 2] = make_float4(t[6], t[7], t[8], 0.f);
}

// ---- host driver ------------------------------------------------------------------------------------------------
struct DeviceBvh {
    WideNode4* nodes = nullptr;     // device, n_nodes records (caller frees with release())
    int32_t* leaf_rank = nullptr;   // device, n
    float4* leaf_geom = nullptr;    // device, 3 n
    int n_nodes = 0, n = 0, max_depth = 0, n_clusters = 0;
    float sah_cost = 0.f;
    double ms_device = 0.0, ms_host_top = 0.0;
    void release() {
        cudaFree(nodes);
        cudaFree(leaf_rank);
        cudaFree(leaf_geom);
        nodes = nullptr;
        leaf_rank = nullptr;
        leaf_geom = nullptr;
    }
};

struct Scratch {  // freed on scope exit; one arena for the whole build (a cudaMalloc / cudaFree pair per array costs
                  // more than the kernels of a 600 k-triangle build), single allocations past its end
    std::vector<void*> ptrs;
    char* arena = nullptr;
    size_t arena_bytes = 0, used = 0;
    cudaError_t reserve(size_t bytes) {
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, bytes);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            return cudaSuccess;  // no arena: alloc() falls back to single allocations
        }
        ptrs.push_back(q);
        arena = (char*)q;
        arena_bytes = bytes;
        return cudaSuccess;
    }
    template <class T>
    cudaError_t alloc(T** p, size_t count) {
        const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~(size_t)255;
        if (used + bytes <= arena_bytes) {
            *p = (T*)(arena + used);
            used += bytes;
            return cudaSuccess;
        }
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, bytes);
        if (e == cudaSuccess) ptrs.push_back(q);
        *p = (T*)q;
        return e;
    }
    ~Scratch() {
        for (void* p : ptrs) cudaFree(p);
    }
};

#define LBVH_TRY(expr)                                                         \
    do {                                                                       \
        cudaError_t e_ = (expr);                                               \
        if (e_ != cudaSuccess) {                                               \
            err = std::string(#expr) + ": " + cudaGetErrorString(e_);          \
            out.release();                                                     \
            return false;                                                      \
        }                                                                      \
    } while (0)

// d_tri_pos: device [rows][9]; d_item_row / d_item_prim: device [n] triangle row and primitive id of item k (items in
// reference leaf order).
inline bool build(cudaStream_t st, const float* d_tri_pos, const int32_t* d_item_row, const int32_t* d_item_prim, int n,
                  DeviceBvh& out, std::string& err) {
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    out = DeviceBvh();
    out.n = n;
    Scratch S;
    const int T = 256;
    const unsigned gb = (unsigned)((n + T - 1) / T);
    float4 *ilo, *ihi, *nlo, *nhi;
    unsigned* bounds;
    unsigned long long *keys, *keys2;
    int32_t *vals, *vals2, *left, *right, *parent, *first, *last;
    int* flags;
    const int top_cap = 2 * (n / 2 + 2);  // clusters <= n, in practice n / 60; sized for the worst cut we accept
    const size_t n_int = (size_t)(n - 1) + (size_t)top_cap;
    // arrays below + the sort's temporary storage (about one key/value copy) + the collapse queues
    S.reserve((size_t)n * (16 + 16 + 8 + 8 + 4 + 4 + 8 + 4 + 2 + 12) + n_int * (4 * 4 + 32) + (size_t)n * 24 + ((size_t)8 << 20));
    LBVH_TRY(S.alloc(&ilo, n));
    LBVH_TRY(S.alloc(&ihi, n));
    LBVH_TRY(S.alloc(&bounds, 8));
    LBVH_TRY(S.alloc(&keys, n));
    LBVH_TRY(S.alloc(&keys2, n));
    LBVH_TRY(S.alloc(&vals, n));
    LBVH_TRY(S.alloc(&vals2, n));
    LBVH_TRY(S.alloc(&left, n_int));
    LBVH_TRY(S.alloc(&right, n_int));
    LBVH_TRY(S.alloc(&first, n_int));
    LBVH_TRY(S.alloc(&last, n_int));
    LBVH_TRY(S.alloc(&nlo, n_int));
    LBVH_TRY(S.alloc(&nhi, n_int));
    LBVH_TRY(S.alloc(&parent, (size_t)2 * n));
    LBVH_TRY(S.alloc(&flags, n));
    const unsigned init_bounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    LBVH_TRY(cudaMemcpyAsync(bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, st));
    LBVH_TRY(cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)n, st));
    k_item_boxes<<<gb, T, 0, st>>>(d_tri_pos, d_item_row, n, ilo, ihi, bounds);
    k_morton<<<gb, T, 0, st>>>(ilo, ihi, n, bounds, keys, vals);
    {
        size_t tmp_bytes = 0;
        LBVH_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, vals, vals2, n, 0, 63, st));
        void* tmp = nullptr;
        LBVH_TRY(S.alloc((char**)&tmp, tmp_bytes));
        LBVH_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, vals, vals2, n, 0, 63, st));
    }
    k_radix_tree<<<gb, T, 0, st>>>(keys2, n, left, right, parent, first, last);
    k_fit<<<gb, T, 0, st>>>(n, vals2, ilo, ihi, left, right, parent, nlo, nhi, flags);
    // ---- the cut: clusters to the host SAH builder --------------------------------------------------------------
    int32_t* c_ref;
    float* c_box;
    int* counters;  // [0] clusters, [1] queue out, [2] wide nodes
    float* d_sah;
    const int c_cap = n / 2 + 2;
    LBVH_TRY(S.alloc(&c_ref, c_cap));
    LBVH_TRY(S.alloc(&c_box, (size_t)6 * c_cap));
    LBVH_TRY(S.alloc(&counters, 4));
    LBVH_TRY(S.alloc(&d_sah, 1));
    LBVH_TRY(cudaMemsetAsync(counters, 0, sizeof(int) * 4, st));
    LBVH_TRY(cudaMemsetAsync(d_sah, 0, sizeof(float), st));
    k_mark_clusters<<<gb, T, 0, st>>>(n, vals2, ilo, ihi, left, right, first, last, nlo, nhi, c_ref, c_box, counters);
    {
        const int wpb = 4;  // warps (clusters) per block
        const size_t smem = sizeof(ClusterSmem) * wpb;
        LBVH_TRY(cudaFuncSetAttribute(k_cluster_sah, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        // the cluster count is on the device; the launch covers the capacity and excess warps return at once
        LBVH_TRY(cudaMemcpyAsync(&out.n_clusters, counters, sizeof(int), cudaMemcpyDeviceToHost, st));
        LBVH_TRY(cudaStreamSynchronize(st));
        if (out.n_clusters < 1 || out.n_clusters > c_cap) {
            err = "device BVH build: unexpected cluster count";
            out.release();
            return false;
        }
        k_cluster_sah<<<(unsigned)((out.n_clusters + wpb - 1) / wpb), 32 * wpb, smem, st>>>(out.n_clusters, c_ref, vals2, ilo, ihi, left,
                                                                                        right, first, last, nlo, nhi, counters + 3);
    }
    int n_clusters = 0, n_big = 0;
    LBVH_TRY(cudaMemcpyAsync(&n_clusters, counters, sizeof(int), cudaMemcpyDeviceToHost, st));
    LBVH_TRY(cudaMemcpyAsync(&n_big, counters + 3, sizeof(int), cudaMemcpyDeviceToHost, st));
    LBVH_TRY(cudaStreamSynchronize(st));
    LBVH_TRY(cudaGetLastError());
    if (n_clusters < 1 || n_clusters > c_cap) {
        err = "device BVH build: unexpected cluster count";
        out.release();
        return false;
    }
    out.n_clusters = n_clusters;
    std::vector<int32_t> h_ref(n_clusters);
    std::vector<float> h_box((size_t)6 * n_clusters);
    LBVH_TRY(cudaMemcpyAsync(h_ref.data(), c_ref, sizeof(int32_t) * n_clusters, cudaMemcpyDeviceToHost, st));
    LBVH_TRY(cudaMemcpyAsync(h_box.data(), c_box, sizeof(float) * 6 * n_clusters, cudaMemcpyDeviceToHost, st));
    LBVH_TRY(cudaStreamSynchronize(st));
    auto t1 = clk::now();
    // host: binned SAH over the clusters, one cluster per leaf; its inner nodes become nodes (n-1) + t
    int32_t root = h_ref[0];
    int n_top = 0;
    if (n_clusters >= 2) {
        std::vector<SahBinNode> top;
        std::vector<int32_t> order;
        build_binary_sah(h_box.data(), n_clusters, 1, top, order);
        // inner top nodes get consecutive ids in the order they appear
        std::vector<int32_t> id(top.size(), -1);
        for (size_t k = 0; k < top.size(); k++)
            if (top[k].left >= 0) id[k] = (n - 1) + n_top++;
        if (n_top > top_cap) {
            err = "device BVH build: top tree larger than its reservation";
            out.release();
            return false;
        }
        std::vector<int32_t> h_left(n_top), h_right(n_top), h_first(n_top, 0), h_last(n_top, 0x3fffffff);
        std::vector<float4> h_lo(n_top), h_hi(n_top);
        auto ref_of = [&](int32_t k) { return top[k].left >= 0 ? id[k] : h_ref[order[top[k].first]]; };
        for (size_t k = 0; k < top.size(); k++) {
            if (top[k].left < 0) continue;
            int t = id[k] - (n - 1);
            h_left[t] = ref_of(top[k].left);
            h_right[t] = ref_of(top[k].right);
            h_lo[t] = float4{top[k].lo[0], top[k].lo[1], top[k].lo[2], 0.f};
            h_hi[t] = float4{top[k].hi[0], top[k].hi[1], top[k].hi[2], 0.f};
        }
        root = id[0];
        const size_t off = (size_t)(n - 1);
        LBVH_TRY(cudaMemcpyAsync(left + off, h_left.data(), sizeof(int32_t) * n_top, cudaMemcpyHostToDevice, st));
        LBVH_TRY(cudaMemcpyAsync(right + off, h_right.data(), sizeof(int32_t) * n_top, cudaMemcpyHostToDevice, st));
        LBVH_TRY(cudaMemcpyAsync(first + off, h_first.data(), sizeof(int32_t) * n_top, cudaMemcpyHostToDevice, st));
        LBVH_TRY(cudaMemcpyAsync(last + off, h_last.data(), sizeof(int32_t) * n_top, cudaMemcpyHostToDevice, st));
        LBVH_TRY(cudaMemcpyAsync(nlo + off, h_lo.data(), sizeof(float4) * n_top, cudaMemcpyHostToDevice, st));
        LBVH_TRY(cudaMemcpyAsync(nhi + off, h_hi.data(), sizeof(float4) * n_top, cudaMemcpyHostToDevice, st));
        LBVH_TRY(cudaStreamSynchronize(st));  // the host vectors go out of scope
    }
    auto t2 = clk::now();
    // ---- collapse --------------------------------------------------------------------------------------------------
    // wide nodes <= inner binary nodes that are not inside a leaf bundle (counted above) + the top nodes
    const size_t wide_cap = (size_t)n_big + (size_t)n_top + 16;
    LBVH_TRY(cudaMalloc(&out.nodes, sizeof(WideNode4) * wide_cap));
    int2 *qa, *qb;
    LBVH_TRY(S.alloc(&qa, wide_cap));
    LBVH_TRY(S.alloc(&qb, wide_cap));
    Tree tree{n, vals2, ilo, ihi, left, right, first, last, nlo, nhi};
    float4 rlo, rhi;
    if (root >= 0) {
        LBVH_TRY(cudaMemcpy(&rlo, nlo + root, sizeof rlo, cudaMemcpyDeviceToHost));
        LBVH_TRY(cudaMemcpy(&rhi, nhi + root, sizeof rhi, cudaMemcpyDeviceToHost));
    } else {
        rlo = float4{0, 0, 0, 0};
        rhi = float4{1, 1, 1, 0};
    }
    const float ra = std::max((rhi.x - rlo.x) * (rhi.y - rlo.y) + (rhi.y - rlo.y) * (rhi.z - rlo.z) +
                                  (rhi.z - rlo.z) * (rhi.x - rlo.x), 1e-30f);
    int2 seed = make_int2(root, 0);
    int h_counts[4] = {0, 0, 1, 0};
    LBVH_TRY(cudaMemcpyAsync(qa, &seed, sizeof seed, cudaMemcpyHostToDevice, st));
    LBVH_TRY(cudaMemcpyAsync(counters, h_counts, sizeof h_counts, cudaMemcpyHostToDevice, st));
    int n_in = 1, depth = 0;
    while (n_in > 0) {
        depth++;
        if ((size_t)n_in > wide_cap) {
            err = "device BVH build: collapse queue overflow";
            out.release();
            return false;
        }
        k_collapse<<<(unsigned)((n_in + 127) / 128), 128, 0, st>>>(tree, qa, n_in, qb, counters + 1, out.nodes,
                                                                     counters + 2, d_sah, 1.0f / ra);
        LBVH_TRY(cudaMemcpyAsync(h_counts, counters, sizeof h_counts, cudaMemcpyDeviceToHost, st));
        LBVH_TRY(cudaMemsetAsync(counters + 1, 0, sizeof(int), st));
        LBVH_TRY(cudaStreamSynchronize(st));
        n_in = h_counts[1];
        if ((size_t)h_counts[2] > wide_cap) {
            err = "device BVH build: wide node reservation exceeded";
            out.release();
            return false;
        }
        std::swap(qa, qb);
    }
    out.n_nodes = h_counts[2];
    out.max_depth = depth;
    LBVH_TRY(cudaMemcpyAsync(&out.sah_cost, d_sah, sizeof(float), cudaMemcpyDeviceToHost, st));
    LBVH_TRY(cudaMalloc(&out.leaf_rank, sizeof(int32_t) * (size_t)n));
    LBVH_TRY(cudaMalloc(&out.leaf_geom, sizeof(float4) * 3 * (size_t)n));
    k_leaf_arrays<<<gb, T, 0, st>>>(d_tri_pos, d_item_row, d_item_prim, vals2, n, out.leaf_rank, out.leaf_geom);
    LBVH_TRY(cudaStreamSynchronize(st));
    LBVH_TRY(cudaGetLastError());
    auto t3 = clk::now();
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    out.ms_device = ms(t0, t1) + ms(t2, t3);
    out.ms_host_top = ms(t1, t2);
    return true;
}
#undef LBVH_TRY

}  // namespace srt_lbvh
#endif
