// Host-side scene flattening and BVH builders (plain C++17, no CUDA calls).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/bdpt.h"
#include "scene.cuh"

namespace bdpt {

struct HostScene {
    std::vector<TriGeom> tris;
    std::vector<TriShade> shade;
    std::vector<double> area;          // triangle.h:28-29
    std::vector<Mat> mats;
    std::vector<uint32_t> lights;      // prim IDs (scene_loader.h:190-202; main.cpp:67 fallback)
    std::vector<double> light_area;    // running sums of the light triangles' areas, list order
    double light_total_area = 0.0;

    // reference topology
    std::vector<RefNode> ref_nodes;    // preorder; root = 0
    std::vector<uint32_t> leaf_order;  // builder's final array == DFS leaf order
    std::vector<uint32_t> dfs_rank;    // prim -> position in leaf_order
    uint32_t ref_depth = 0;
    std::vector<int32_t> ref_parent;   // node -> parent (-1 root)
    std::vector<int32_t> prim_leaf;    // prim -> leaf node
    std::vector<double> chain_box;     // prim -> intersection of the boxes on its root-to-leaf chain: lo_x,hi_x,lo_y,hi_y,lo_z,hi_z

    // fast topology
    std::vector<FastNode> fast_nodes;
    std::vector<uint32_t> fast_prims;
    std::vector<TriGeom> fast_tris;    // tris[fast_prims[k]]: a leaf's triangles are contiguous, no id indirection on the walk
    uint32_t fast_depth = 0;
    std::vector<WideNode> wide_nodes;  // the fast tree collapsed to 4-wide nodes (same leaves)
    uint32_t wide_depth = 0;
    double extent = 0.0;               // largest side of the scene bounds
    double center[3] = {0, 0, 0};

    // per-triangle fp64 bounds (triangle.h:31-38), unpadded: lo xyz, hi xyz
    std::vector<double> tri_box;
};

// Throws std::runtime_error on bad input (caught at the C-ABI).
void flatten_scene(const double* verts, const uint32_t* mat_id, uint64_t n_tris, const bdpt_material* mats,
                   uint32_t n_mats, HostScene& out);
void build_ref_bvh(HostScene& s);
void build_fast_bvh(HostScene& s);
void collapse_wide(HostScene& s);   // fast_nodes -> wide_nodes
void scene_bounds(HostScene& s, double lo[3], double hi[3]);

}  // namespace bdpt
