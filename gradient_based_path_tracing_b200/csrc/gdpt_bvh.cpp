// Host-side BVH construction (see gdpt_bvh.h).  Binned SAH (16 bins, 3 axes) binary tree,
// then a greedy surface-area collapse into 4-wide nodes.  The build is a one-off per scene
// (Sponza: 66k triangles, a few tens of ms); traversal quality, not build time, is what the
// render kernels feel.
#include "gdpt_bvh.h"

#include <algorithm>
#include <cmath>
#include <limits>

namespace gdpt {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<float>::infinity(); hi[k] = -lo[k]; } }
    void grow(const Box &b) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
    double area() const {
        double dx = (double)hi[0] - lo[0], dy = (double)hi[1] - lo[1], dz = (double)hi[2] - lo[2];
        if (dx < 0) return 0;
        return 2 * (dx * dy + dy * dz + dz * dx);
    }
};

struct Ref { Box b; float c[3]; int idx; };
struct BNode { Box b; int left = -1, right = -1, first = 0, count = 0; };

struct Builder {
    std::vector<Ref> refs;
    std::vector<BNode> nodes;

    int build(int first, int count) {
        BNode n; n.b.reset();
        Box cb; cb.reset();
        for (int i = first; i < first + count; i++) {
            n.b.grow(refs[i].b);
            for (int k = 0; k < 3; k++) { cb.lo[k] = std::min(cb.lo[k], refs[i].c[k]); cb.hi[k] = std::max(cb.hi[k], refs[i].c[k]); }
        }
        n.first = first; n.count = count;
        int id = (int)nodes.size();
        nodes.push_back(n);
        if (count <= 1) return id;
        const int NB = 16;
        double best = std::numeric_limits<double>::infinity(); int bax = -1, bbin = -1;
        for (int ax = 0; ax < 3; ax++) {
            double ext = (double)cb.hi[ax] - cb.lo[ax];
            if (!(ext > 0)) continue;
            int cnt[NB] = {0}; Box bb[NB];
            for (int b = 0; b < NB; b++) bb[b].reset();
            for (int i = first; i < first + count; i++) {
                int b = std::min(NB - 1, (int)(NB * ((double)refs[i].c[ax] - cb.lo[ax]) / ext));
                cnt[b]++; bb[b].grow(refs[i].b);
            }
            double ra[NB]; int rc[NB]; Box acc; acc.reset(); int c = 0;
            for (int b = NB - 1; b > 0; b--) { c += cnt[b]; acc.grow(bb[b]); ra[b] = acc.area(); rc[b] = c; }
            acc.reset(); c = 0;
            for (int b = 0; b < NB - 1; b++) {
                c += cnt[b]; acc.grow(bb[b]);
                if (c == 0 || rc[b + 1] == 0) continue;
                double cost = acc.area() * c + ra[b + 1] * rc[b + 1];
                if (cost < best) { best = cost; bax = ax; bbin = b; }
            }
        }
        int mid;
        if (bax < 0) {
            if (count <= 4) return id;
            mid = first + count / 2;
        } else {
            // leaf cost vs split cost (traversal cost 1, intersection cost 1 per primitive)
            double leaf = n.b.area() * count, split = n.b.area() * 1.0 + best;
            if (count <= 4 && leaf <= split) return id;
            double ext = (double)cb.hi[bax] - cb.lo[bax];
            auto it = std::partition(refs.begin() + first, refs.begin() + first + count, [&](const Ref &r) {
                int b = std::min(NB - 1, (int)(NB * ((double)r.c[bax] - cb.lo[bax]) / ext));
                return b <= bbin;
            });
            mid = (int)(it - refs.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        int l = build(first, mid - first);
        int r = build(mid, first + count - mid);
        nodes[id].left = l; nodes[id].right = r; nodes[id].count = 0;
        return id;
    }
};

} // namespace

void build_bvh(const std::vector<HostShape> &shapes, Bvh &out) {
    std::vector<BvhPrim> src;
    Builder B;
    for (int s = 0; s < (int)shapes.size(); s++) {
        const HostShape &sh = shapes[s];
        if (sh.kind == 1) {
            for (int t = 0; t < sh.ntris; t++) {
                const int *ix = sh.indices + 3 * t;
                float v[3][3];
                for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) v[c][k] = (float)sh.positions[3 * (size_t)ix[c] + k];
                BvhPrim p;
                Ref r; r.b.reset();
                for (int k = 0; k < 3; k++) {
                    p.v0[k] = v[0][k]; p.v1[k] = v[1][k]; p.v2[k] = v[2][k];
                    r.b.lo[k] = std::min(v[0][k], std::min(v[1][k], v[2][k]));
                    r.b.hi[k] = std::max(v[0][k], std::max(v[1][k], v[2][k]));
                    r.c[k] = 0.5f * (r.b.lo[k] + r.b.hi[k]);
                }
                p.prim_id = t; p.shape_id = s; p.kind = 0;
                r.idx = (int)src.size();
                src.push_back(p); B.refs.push_back(r);
            }
        } else {
            BvhPrim p;
            Ref r;
            for (int k = 0; k < 3; k++) {
                p.v0[k] = (float)sh.sphere[k]; p.v1[k] = 0; p.v2[k] = 0;
                // RTCBounds are floats: (float)(centre -/+ radius) as sphere_bounds_func stores them
                r.b.lo[k] = (float)(sh.sphere[k] - sh.sphere[3]);
                r.b.hi[k] = (float)(sh.sphere[k] + sh.sphere[3]);
                // keep the box conservative after rounding
                if ((double)r.b.lo[k] > sh.sphere[k] - sh.sphere[3]) r.b.lo[k] = std::nextafter(r.b.lo[k], -std::numeric_limits<float>::infinity());
                if ((double)r.b.hi[k] < sh.sphere[k] + sh.sphere[3]) r.b.hi[k] = std::nextafter(r.b.hi[k], std::numeric_limits<float>::infinity());
                r.c[k] = (float)sh.sphere[k];
            }
            p.v1[0] = (float)sh.sphere[3];
            p.prim_id = 0; p.shape_id = s; p.kind = 1;
            r.idx = (int)src.size();
            src.push_back(p); B.refs.push_back(r);
        }
    }
    out.nodes.clear(); out.prims.clear();
    for (int k = 0; k < 3; k++) { out.lo[k] = std::numeric_limits<float>::infinity(); out.hi[k] = -out.lo[k]; }
    if (B.refs.empty()) return;
    // scene bounds as rtcGetSceneBounds reports them: float min/max of the primitive bounds,
    // spheres contributing (float)(c -/+ r) exactly (no conservative nudge)
    for (int s = 0; s < (int)shapes.size(); s++) {
        const HostShape &sh = shapes[s];
        if (sh.kind == 1) {
            for (int t = 0; t < sh.ntris; t++) for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) {
                float f = (float)sh.positions[3 * (size_t)sh.indices[3 * t + c] + k];
                out.lo[k] = std::min(out.lo[k], f); out.hi[k] = std::max(out.hi[k], f);
            }
        } else {
            for (int k = 0; k < 3; k++) {
                out.lo[k] = std::min(out.lo[k], (float)(sh.sphere[k] - sh.sphere[3]));
                out.hi[k] = std::max(out.hi[k], (float)(sh.sphere[k] + sh.sphere[3]));
            }
        }
    }
    B.nodes.reserve(2 * B.refs.size());
    B.build(0, (int)B.refs.size());
    out.prims.resize(B.refs.size());
    for (size_t i = 0; i < B.refs.size(); i++) out.prims[i] = src[B.refs[i].idx];

    // collapse to 4-wide
    struct Item { int bnode; int wide; int depth; };
    std::vector<Item> work;
    auto new_wide = [&]() { BvhNode n; for (int c = 0; c < 4; c++) {
            n.lox[c] = n.loy[c] = n.loz[c] = std::numeric_limits<float>::infinity();
            n.hix[c] = n.hiy[c] = n.hiz[c] = -std::numeric_limits<float>::infinity();
            n.child[c] = 0; n.count[c] = 0; }
        out.nodes.push_back(n); return (int)out.nodes.size() - 1; };
    int root = new_wide();
    work.push_back({0, root, 1});
    while (!work.empty()) {
        Item it = work.back(); work.pop_back();
        out.max_depth = std::max(out.max_depth, it.depth);
        int kids[4]; int nk = 0;
        const BNode &bn = B.nodes[it.bnode];
        if (bn.left < 0) { kids[nk++] = it.bnode; }
        else { kids[nk++] = bn.left; kids[nk++] = bn.right; }
        while (nk < 4) {
            int pick = -1; double pa = -1;
            for (int i = 0; i < nk; i++) {
                const BNode &k = B.nodes[kids[i]];
                if (k.left >= 0 && k.b.area() > pa) { pa = k.b.area(); pick = i; }
            }
            if (pick < 0) break;
            int k = kids[pick];
            kids[pick] = B.nodes[k].left; kids[nk++] = B.nodes[k].right;
        }
        for (int i = 0; i < nk; i++) {
            const BNode &k = B.nodes[kids[i]];
            BvhNode &w = out.nodes[it.wide];
            w.lox[i] = k.b.lo[0]; w.loy[i] = k.b.lo[1]; w.loz[i] = k.b.lo[2];
            w.hix[i] = k.b.hi[0]; w.hiy[i] = k.b.hi[1]; w.hiz[i] = k.b.hi[2];
            if (k.left < 0) { w.child[i] = ~((k.first << 3) | k.count); w.count[i] = k.count; }
            else {
                int c = new_wide();
                out.nodes[it.wide].child[i] = c; out.nodes[it.wide].count[i] = 0;
                work.push_back({kids[i], c, it.depth + 1});
            }
        }
    }
}

} // namespace gdpt
