// 4-wide BVH over all triangles and spheres of a scene — the replacement for Embree's
// rtcCommitScene (/root/reference/src/scene.cpp:19-34) and the data the traversal kernel
// (gdpt_trace.cuh) walks.  Geometry is rounded to float exactly as the reference hands it
// to Embree (src/shapes/triangle_mesh.inl:11-14, src/shapes/sphere.inl:1-10).
//
// Node (128 B, one L2/L1 line, 8 x 16-byte vector loads):
//   float4 lox, loy, loz, hix, hiy, hiz   bounds of the 4 children, SoA over children
//   int4   child                           >0: node index; <0: leaf, ~child = (first primitive << 3) | count;
//                                          0: empty slot (its bounds are inverted, never hit)
//   int4   count                           primitives in the leaf (informational; pads the node to 128 B)
// Primitive (48 B, 3 x 16-byte loads): v0.xyz | prim_id,  v1.xyz | shape_id,  v2.xyz | kind
//   kind 0 = triangle (float-rounded vertices, exactly what Embree would hold),
//   kind 1 = sphere (centre/radius stay double in DScene::spheres, as in sphere.inl)
#pragma once
#include <cstdint>
#include <vector>

namespace gdpt {

struct BvhNode {
    float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];
    int child[4];
    int count[4];
};
static_assert(sizeof(BvhNode) == 128, "node must be one 128-byte line");

struct BvhPrim {
    float v0[3]; int prim_id;
    float v1[3]; int shape_id;
    float v2[3]; int kind;
};
static_assert(sizeof(BvhPrim) == 48, "primitive record must be 48 bytes");

struct HostShape {
    int kind;                 // 0 sphere, 1 mesh
    const double *positions;  // mesh: [nverts][3]
    const int *indices;       // mesh: [ntris][3]
    int nverts, ntris;
    double sphere[4];
};

struct Bvh {
    std::vector<BvhNode> nodes;
    std::vector<BvhPrim> prims;
    float lo[3], hi[3];       // float scene bounds (what rtcGetSceneBounds reports)
    int max_depth = 0;
};

// Binned-SAH binary build collapsed to 4-wide. Deterministic.
void build_bvh(const std::vector<HostShape> &shapes, Bvh &out);
// Only Bvh::lo / hi (what rtcGetSceneBounds reports), without building anything: for a context whose tree arrives from
// another process (gdpt_create_ex + gdpt_bvh_adopt).  build_bvh computes its bounds with this function.
void scene_bounds(const std::vector<HostShape> &shapes, float lo[3], float hi[3]);

} // namespace gdpt
