// srt_bvh_build.h — host-side acceleration-structure builders (run once per srt_scene_create).
//
//   build_wide_bvh : binned-SAH binary tree over item boxes, collapsed to the 4-wide WideNode4
//                    layout the traversal kernels read (srt_types.h).  This replaces the
//                    reference's random-axis median split (bvh.h:96-119) as the tree that is
//                    actually WALKED; it decides nothing about results.
//   build_ref_tree : the reference's OWN tree shape over the leaves in their in-order sequence
//                    (median split n/2, bvh.h:104-114) with its un-padded surrounding_box unions
//                    (aabb.h:54-62), kept only as boxes: the kernels replay a candidate's ancestor
//                    chain through it so that exactly the leaves the reference can reach are
//                    reported.
#ifndef SRT_BVH_BUILD_H
#define SRT_BVH_BUILD_H

#include <vector>

#include "srt_types.h"

struct WideBvh {
    std::vector<WideNode4> nodes;  // nodes[0] is the root; child indices are relative to it
    std::vector<int32_t> order;    // leaf slot -> item index
    float sah_cost = 0.f;
    int max_depth = 0;  // inner-node levels above the deepest leaf (the traversal stack holds <= 3 per level)
};

// boxes: [n][6] = min xyz, max xyz (already padded by the caller).  n >= 1.
void build_wide_bvh(const float* boxes, int n, WideBvh& out);

// The binary binned-SAH tree itself (before the collapse), for callers that join it with other subtrees: the
// device builder (srt_lbvh.cuh) hands it the boxes of its radix-tree clusters with max_leaf = 1.
// nodes[0] is the root; a leaf has left < 0 and covers order[first .. first + count).
struct SahBinNode {
    float lo[3], hi[3];
    int32_t left, right;
    int32_t first, count;
};
void build_binary_sah(const float* boxes, int n, int max_leaf, std::vector<SahBinNode>& nodes, std::vector<int32_t>& order);

// item_boxes: [n][6] reference leaf boxes in the reference tree's in-order leaf sequence.
// out[0] is the root; child indices are relative to out[0].
void build_ref_tree(const float* item_boxes, int n, std::vector<RefNode>& out);

// The shape of that tree level by level, for the device build of large trees (k_ref_tree_level, srt_api.cu): level L
// holds subtrees of size_lo[L] or size_lo[L] + 1 leaves, which have nodes_lo[L] / nodes_hi[L] nodes.  Returns the
// number of levels that hold nodes; the tables are filled one level past that.  n < 2^30.
int ref_tree_levels(long long n, int size_lo[32], int nodes_lo[32], int nodes_hi[32]);

// The in-order leaf sequence of the reference's `new bvh_node(l, n, t0, t1)` (bvh.h:96-119): at every node, in
// depth-first order, axis = int(3 * drand48()) from the global 48-bit LCG (mathf.h:12-19), qsort of the node's
// slice by box minimum on that axis with a comparator that never returns 0 (bvh.h:21-55; glibc's merge sort
// then leaves equal keys in REVERSE input order), split at n / 2.  mins: [n][3]; order_out[k] = input index of
// the k-th leaf; *seed is the LCG state on entry and on return (one draw per node).  Multi-threaded: the axis
// of every node is fixed by the tree shape alone, so the draws are made up front and the sorts run as tasks.
void ref_bvh_leaf_order(const float* mins, long long n, unsigned long long* seed, int32_t* order_out);

#endif
