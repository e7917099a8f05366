// CPU check of the host BVH builder (gradient_based_path_tracing_b200/csrc/gdpt_bvh.cpp): builds the tree for a
// geometry dump and verifies the invariants the traversal kernel relies on.  Prints one JSON line.
//   bvh_check shapes.bin offsets.bin spheres.bin positions.bin indices.bin
#include "gdpt_bvh.h"

#include <cstdio>
#include <fstream>
#include <set>
#include <vector>
using namespace gdpt;

template <typename T> static std::vector<T> rd(const char *p) {
    std::ifstream f(p, std::ios::binary); f.seekg(0, std::ios::end); size_t n = (size_t)f.tellg(); f.seekg(0);
    std::vector<T> v(n / sizeof(T)); f.read((char *)v.data(), (std::streamsize)n); return v;
}

int main(int argc, char **argv) {
    if (argc < 6) return 2;
    auto shapes = rd<int>(argv[1]); auto off = rd<long long>(argv[2]); auto sph = rd<double>(argv[3]);
    auto pos = rd<double>(argv[4]); auto idx = rd<int>(argv[5]);
    int n = (int)shapes.size() / 6;
    std::vector<HostShape> hs(n);
    size_t n_prims = 0;
    for (int i = 0; i < n; i++) {
        hs[i].kind = shapes[6 * i]; hs[i].nverts = shapes[6 * i + 3]; hs[i].ntris = shapes[6 * i + 4];
        hs[i].positions = pos.data() + 3 * off[4 * i]; hs[i].indices = idx.data() + 3 * off[4 * i + 1];
        for (int k = 0; k < 4; k++) hs[i].sphere[k] = sph[4 * i + k];
        n_prims += hs[i].kind == 1 ? hs[i].ntris : 1;
    }
    Bvh a, b;
    build_bvh(hs, a);
    build_bvh(hs, b);
    bool deterministic = a.nodes.size() == b.nodes.size() && a.prims.size() == b.prims.size();
    for (size_t i = 0; deterministic && i < a.nodes.size(); i++)
        for (int k = 0; k < 4; k++) if (a.nodes[i].child[k] != b.nodes[i].child[k] || a.nodes[i].lox[k] != b.nodes[i].lox[k]) deterministic = false;

    // walk the tree: every leaf range valid, every primitive of the scene referenced at least once, and every
    // referenced primitive piece-wise covered: the union of the leaf boxes that reference a triangle contains it
    long long bad_range = 0, bad_child = 0, leaves = 0, max_leaf = 0, uncovered = 0;
    std::set<std::pair<int, int>> seen;
    struct Acc { float lo[3], hi[3]; bool any; };
    std::vector<Acc> cover(a.prims.size());
    std::vector<std::pair<int, int>> stack;   // (node, depth)
    int depth = 0;
    if (!a.nodes.empty()) stack.push_back({0, 1});
    std::vector<char> visited(a.nodes.size(), 0);
    while (!stack.empty()) {
        auto [ni, d] = stack.back(); stack.pop_back();
        if (ni < 0 || ni >= (int)a.nodes.size() || visited[ni]) { bad_child++; continue; }
        visited[ni] = 1; depth = d > depth ? d : depth;
        const BvhNode &nd = a.nodes[ni];
        for (int k = 0; k < 4; k++) {
            int c = nd.child[k];
            if (c == 0) continue;
            if (c > 0) { stack.push_back({c, d + 1}); continue; }
            int enc = ~c, first = enc >> 3, cnt = enc & 7;
            leaves++; max_leaf = cnt > max_leaf ? cnt : max_leaf;
            if (cnt < 1 || first < 0 || first + cnt > (int)a.prims.size()) { bad_range++; continue; }
            for (int i = first; i < first + cnt; i++) {
                const BvhPrim &p = a.prims[i];
                seen.insert({p.shape_id, p.prim_id});
                if (p.kind != 0) continue;
                // the part of the triangle inside this leaf box must be non-empty for at least one vertex-or-edge sample;
                // cheap necessary check: the leaf box intersects the triangle's box
                float tlo[3], thi[3];
                for (int x = 0; x < 3; x++) { tlo[x] = std::min(p.v0[x], std::min(p.v1[x], p.v2[x])); thi[x] = std::max(p.v0[x], std::max(p.v1[x], p.v2[x])); }
                const float blo[3] = {nd.lox[k], nd.loy[k], nd.loz[k]}, bhi[3] = {nd.hix[k], nd.hiy[k], nd.hiz[k]};
                for (int x = 0; x < 3; x++) if (blo[x] > thi[x] || bhi[x] < tlo[x]) uncovered++;
            }
        }
    }
    long long unvisited = 0;
    for (char v : visited) if (!v) unvisited++;
    printf("{\"nodes\": %zu, \"prim_records\": %zu, \"scene_prims\": %zu, \"distinct_prims\": %zu, \"leaves\": %lld, \"max_leaf\": %lld, "
           "\"depth\": %d, \"reported_depth\": %d, \"bad_range\": %lld, \"bad_child\": %lld, \"unvisited\": %lld, \"uncovered\": %lld, \"deterministic\": %s, "
           "\"lo\": [%.9g, %.9g, %.9g], \"hi\": [%.9g, %.9g, %.9g]}\n",
           a.nodes.size(), a.prims.size(), n_prims, seen.size(), leaves, max_leaf, depth, a.max_depth, bad_range, bad_child, unvisited, uncovered,
           deterministic ? "true" : "false", a.lo[0], a.lo[1], a.lo[2], a.hi[0], a.hi[1], a.hi[2]);
    return 0;
}
