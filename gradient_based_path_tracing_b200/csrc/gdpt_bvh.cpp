// Host-side BVH construction (see gdpt_bvh.h).  Binned SAH (16 bins, 3 axes) binary tree,
// then a greedy surface-area collapse into 4-wide nodes.  The build is a one-off per scene
// (Sponza: 66k triangles, a few tens of ms); traversal quality, not build time, is what the
// render kernels feel.
#include "gdpt_bvh.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <thread>
#include <utility>
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
    std::vector<BNode> nodes;          // preallocated (2n)
    double ci = 1.0;      // SAH cost of one primitive test relative to one (binary) node step
    int max_leaf = 4;     // <= 7 (3-bit count in the leaf encoding)
    std::atomic<int> workers{0};   // extra build threads alive
    int max_workers = 0;

    // Node slots are handed out by position, not by a shared counter: the subtree over `count` references owns the
    // slots [slot, slot + 2*count - 1), its root sits at `slot`, the left child's range follows, then the right
    // child's.  Concurrent subtrees therefore never touch the same cache lines, and the layout is deterministic.
    int build(int first, int count, int slot, int depth = 0) {
        BNode n; n.b.reset();
        Box cb; cb.reset();
        for (int i = first; i < first + count; i++) {
            n.b.grow(refs[i].b);
            for (int k = 0; k < 3; k++) { cb.lo[k] = std::min(cb.lo[k], refs[i].c[k]); cb.hi[k] = std::max(cb.hi[k], refs[i].c[k]); }
        }
        n.first = first; n.count = count;
        const int id = slot;
        nodes[id] = n;
        if (count <= 1) return id;
        // 16 bins per axis; small nodes (most of the tree) use fewer, their cost is the bin bookkeeping, not the references
        constexpr int NBMAX = 16;
        const int NB = count > 32 ? 16 : (count > 8 ? 8 : 4);
        // one pass over the references fills the bins of all three axes
        float scale[3]; bool live[3];
        for (int ax = 0; ax < 3; ax++) { float ext = cb.hi[ax] - cb.lo[ax]; live[ax] = ext > 0.f; scale[ax] = live[ax] ? (float)NB / ext : 0.f; }
        auto bin_of = [&](const Ref &r, int ax) { int b = (int)((r.c[ax] - cb.lo[ax]) * scale[ax]); return b < 0 ? 0 : (b > NB - 1 ? NB - 1 : b); };
        int cnt[3][NBMAX]; Box bb[3][NBMAX];
        for (int ax = 0; ax < 3; ax++) for (int b = 0; b < NB; b++) { cnt[ax][b] = 0; bb[ax][b].reset(); }
        for (int i = first; i < first + count; i++) {
            const Ref &r = refs[i];
            for (int ax = 0; ax < 3; ax++) { int b = bin_of(r, ax); cnt[ax][b]++; bb[ax][b].grow(r.b); }
        }
        double best = std::numeric_limits<double>::infinity(); int bax = -1, bbin = -1;
        for (int ax = 0; ax < 3; ax++) {
            if (!live[ax]) continue;
            double ra[NBMAX]; int rc[NBMAX]; Box acc; acc.reset(); int c = 0;
            for (int b = NB - 1; b > 0; b--) { c += cnt[ax][b]; acc.grow(bb[ax][b]); ra[b] = acc.area(); rc[b] = c; }
            acc.reset(); c = 0;
            for (int b = 0; b < NB - 1; b++) {
                c += cnt[ax][b]; acc.grow(bb[ax][b]);
                if (c == 0 || rc[b + 1] == 0) continue;
                double cost = acc.area() * c + ra[b + 1] * rc[b + 1];
                if (cost < best) { best = cost; bax = ax; bbin = b; }
            }
        }
        int mid;
        if (bax < 0) {
            if (count <= max_leaf) return id;
            mid = first + count / 2;
        } else {
            // leaf cost vs split cost (traversal cost 1, intersection cost 1 per primitive)
            double leaf = n.b.area() * count * ci, split = n.b.area() * 1.0 + best * ci;
            if (count <= max_leaf && leaf <= split) return id;
            auto it = std::partition(refs.begin() + first, refs.begin() + first + count, [&](const Ref &r) { return bin_of(r, bax) <= bbin; });
            mid = (int)(it - refs.begin());
            if (mid == first || mid == first + count) mid = first + count / 2;
        }
        const int lc = mid - first, lslot = slot + 1, rslot = slot + 2 * lc;
        int l, r;
        // SAH splits of real scenes are lopsided (Sponza's first cuts peel off walls with a few percent of the
        // triangles), so forking by depth leaves one thread with nearly everything: fork wherever BOTH halves are
        // still large, as long as a worker slot is free.  Disjoint reference ranges, disjoint node slots.
        if (std::min(lc, count - lc) > 4096 && workers.fetch_add(1) < max_workers) {
            std::thread th([&] { l = build(first, lc, lslot, depth + 1); workers.fetch_sub(1); });
            r = build(mid, count - lc, rslot, depth + 1);
            th.join();
        } else {
            if (std::min(lc, count - lc) > 4096) workers.fetch_sub(1);
            l = build(first, lc, lslot, depth + 1);
            r = build(mid, count - lc, rslot, depth + 1);
        }
        nodes[id].left = l; nodes[id].right = r; nodes[id].count = 0;
        return id;
    }
};

// ---- triangle pre-splitting (spatial splits before the build, as in Embree's HIGH-quality mode the
// reference asks for, scene.cpp:22): the boxes of large triangles are cut at the spatial median of
// their longest axis and re-fitted to the clipped triangle, so a long wall or floor triangle stops
// inflating every node above it.  A triangle may then sit in several leaves; the closest-hit rule
// is unaffected (same triangle, same t, same ids).
struct Poly { double v[10][3]; int n; };

inline void clip_poly(const Poly &in, int axis, double plane, bool keep_below, Poly &out) {
    out.n = 0;
    for (int i = 0; i < in.n; i++) {
        const double *a = in.v[i], *b = in.v[(i + 1) % in.n];
        bool ia = keep_below ? a[axis] <= plane : a[axis] >= plane, ib = keep_below ? b[axis] <= plane : b[axis] >= plane;
        if (ia) { for (int k = 0; k < 3; k++) out.v[out.n][k] = a[k]; out.n++; }
        if (ia != ib) {
            double t = (plane - a[axis]) / (b[axis] - a[axis]);
            for (int k = 0; k < 3; k++) out.v[out.n][k] = a[k] + t * (b[k] - a[k]);
            out.v[out.n][axis] = plane; out.n++;
        }
    }
}

// tight float box of a clipped polygon, rounded outwards and kept inside `box`; false if empty
inline bool poly_box(const Poly &p, const Box &box, Box &out) {
    if (p.n == 0) return false;
    const float inf = std::numeric_limits<float>::infinity();
    for (int k = 0; k < 3; k++) {
        double lo = p.v[0][k], hi = p.v[0][k];
        for (int i = 1; i < p.n; i++) { lo = std::min(lo, p.v[i][k]); hi = std::max(hi, p.v[i][k]); }
        float flo = (float)lo, fhi = (float)hi;
        if ((double)flo > lo) flo = std::nextafter(flo, -inf);
        if ((double)fhi < hi) fhi = std::nextafter(fhi, inf);
        flo = std::nextafter(flo, -inf); fhi = std::nextafter(fhi, inf);        // one more ulp for the interpolation error
        out.lo[k] = std::max(flo, box.lo[k]); out.hi[k] = std::min(fhi, box.hi[k]);
        if (out.lo[k] > out.hi[k]) return false;
    }
    return true;
}

// Split the largest boxes first until the reference count reaches `budget`.  Every reference that has been cut
// carries the polygon (triangle ∩ its box) it covers, so a cut is one clip per side instead of re-clipping the
// triangle against all six planes of each half.
inline void presplit(std::vector<Ref> &refs, const std::vector<BvhPrim> &src, size_t budget) {
    // heap entries carry their key (area, then index for a deterministic order): comparing through refs[] would be
    // two cache misses per comparison
    typedef std::pair<double, int> Entry;
    std::vector<Entry> heap;
    for (int i = 0; i < (int)refs.size(); i++) if (src[refs[i].idx].kind == 0) heap.push_back(Entry(refs[i].b.area(), i));
    std::make_heap(heap.begin(), heap.end());
    auto push = [&](int i) { heap.push_back(Entry(refs[i].b.area(), i)); std::push_heap(heap.begin(), heap.end()); };
    std::vector<Poly> polys;                       // polygon of reference i at poly_of[i] (or -1: still the whole triangle)
    std::vector<int> poly_of(refs.size(), -1);
    polys.reserve(budget > refs.size() ? 2 * (budget - refs.size()) + 16 : 16);
    refs.reserve(budget + 1); poly_of.reserve(budget + 1);
    while (!heap.empty() && refs.size() < budget) {
        std::pop_heap(heap.begin(), heap.end());
        int i = heap.back().second; heap.pop_back();
        Ref r = refs[i];
        int ax = 0; float ext = r.b.hi[0] - r.b.lo[0];
        for (int k = 1; k < 3; k++) if (r.b.hi[k] - r.b.lo[k] > ext) { ext = r.b.hi[k] - r.b.lo[k]; ax = k; }
        if (!(ext > 0.f)) break;      // the largest box is a point: nothing left to split
        float mid = 0.5f * (r.b.lo[ax] + r.b.hi[ax]);
        if (!(mid > r.b.lo[ax] && mid < r.b.hi[ax])) continue;
        Poly whole;
        if (poly_of[i] < 0) {
            const BvhPrim &p = src[r.idx];
            whole.n = 3;
            for (int k = 0; k < 3; k++) { whole.v[0][k] = p.v0[k]; whole.v[1][k] = p.v1[k]; whole.v[2][k] = p.v2[k]; }
        } else whole = polys[poly_of[i]];
        Poly pl, pr;
        clip_poly(whole, ax, (double)mid, true, pl);
        clip_poly(whole, ax, (double)mid, false, pr);
        Box lb = r.b, rb = r.b; lb.hi[ax] = mid; rb.lo[ax] = mid;
        Box lt, rt;
        bool hl = poly_box(pl, lb, lt), hr = poly_box(pr, rb, rt);
        if (!hl || !hr) {             // the triangle only touches one half: just tighten
            if (hl || hr) {
                refs[i].b = hl ? lt : rt; for (int k = 0; k < 3; k++) refs[i].c[k] = 0.5f * (refs[i].b.lo[k] + refs[i].b.hi[k]);
                if (poly_of[i] < 0) { poly_of[i] = (int)polys.size(); polys.push_back(hl ? pl : pr); } else polys[poly_of[i]] = hl ? pl : pr;
                if (refs[i].b.area() < r.b.area()) push(i);
            }
            continue;
        }
        Ref a = r, b = r;
        a.b = lt; b.b = rt;
        for (int k = 0; k < 3; k++) { a.c[k] = 0.5f * (lt.lo[k] + lt.hi[k]); b.c[k] = 0.5f * (rt.lo[k] + rt.hi[k]); }
        refs[i] = a; refs.push_back(b);
        if (poly_of[i] < 0) { poly_of[i] = (int)polys.size(); polys.push_back(pl); } else polys[poly_of[i]] = pl;
        poly_of.push_back((int)polys.size()); polys.push_back(pr);
        push(i); push((int)refs.size() - 1);
    }
}

// The heap above is sequential; the references are therefore cut into a FIXED number of chunks (independent of the thread
// count, so the tree is deterministic), each chunk gets the share of the extra references that its share of the total box area
// calls for — the global rule "always split the largest box" ends with every box below one common area, i.e. with a piece
// count proportional to area — and the chunks run side by side.  Sponza: 62 ms -> under 10 ms on 16 host threads.
inline void presplit_parallel(std::vector<Ref> &refs, const std::vector<BvhPrim> &src, size_t budget) {
    constexpr int CHUNKS = 16;
    const size_t n = refs.size();
    if (n < 4096 || budget <= n) { presplit(refs, src, budget); return; }
    std::vector<std::vector<Ref>> part(CHUNKS);
    double area[CHUNKS], total = 0;
    for (int c = 0; c < CHUNKS; c++) {
        // dealt out round-robin: a mesh's large triangles (walls, floors) sit next to each other in the index order, and a
        // contiguous cut would leave one chunk with most of the area, i.e. most of the splits
        part[c].reserve(n / CHUNKS + 1);
        for (size_t i = c; i < n; i += CHUNKS) part[c].push_back(refs[i]);
        area[c] = 0;
        for (const Ref &r : part[c]) if (src[r.idx].kind == 0) area[c] += r.b.area();
        total += area[c];
    }
    if (!(total > 0)) return;
    std::vector<std::thread> th;
    for (int c = 0; c < CHUNKS; c++) {
        const size_t extra = (size_t)((double)(budget - n) * (area[c] / total));
        th.emplace_back([&part, &src, c, extra] { presplit(part[c], src, part[c].size() + extra); });
    }
    for (auto &t : th) t.join();
    refs.clear();
    for (int c = 0; c < CHUNKS; c++) refs.insert(refs.end(), part[c].begin(), part[c].end());
}

} // namespace

// scene bounds as rtcGetSceneBounds reports them: float min/max of the primitive bounds,
// spheres contributing (float)(c -/+ r) exactly (no conservative nudge)
void scene_bounds(const std::vector<HostShape> &shapes, float lo[3], float hi[3]) {
    for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<float>::infinity(); hi[k] = -lo[k]; }
    for (int s = 0; s < (int)shapes.size(); s++) {
        const HostShape &sh = shapes[s];
        if (sh.kind == 1) {
            for (int t = 0; t < sh.ntris; t++) for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) {
                float f = (float)sh.positions[3 * (size_t)sh.indices[3 * t + c] + k];
                lo[k] = std::min(lo[k], f); hi[k] = std::max(hi[k], f);
            }
        } else {
            for (int k = 0; k < 3; k++) {
                lo[k] = std::min(lo[k], (float)(sh.sphere[k] - sh.sphere[3]));
                hi[k] = std::max(hi[k], (float)(sh.sphere[k] + sh.sphere[3]));
            }
        }
    }
}

void build_bvh(const std::vector<HostShape> &shapes, Bvh &out) {
    const bool verbose = getenv("GDPT_VERBOSE") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        auto now = std::chrono::steady_clock::now();
        if (verbose) fprintf(stderr, "[build_bvh]   %-16s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
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
    scene_bounds(shapes, out.lo, out.hi);
    if (const char *e = getenv("GDPT_BVH_CI")) B.ci = atof(e);                                   // tuning knobs (tools/gpu_bvh_sweep.py)
    if (const char *e = getenv("GDPT_BVH_MAX_LEAF")) B.max_leaf = std::min(7, std::max(1, atoi(e)));
    double split_budget = 2.0;      // extra references allowed, as a fraction of the primitive count (B200 sweep: profiles/r1_bvh_sweep.txt)
    if (const char *e = getenv("GDPT_BVH_SPLIT")) split_budget = atof(e);
    lap("references");
    if (split_budget > 0.0) presplit_parallel(B.refs, src, (size_t)((1.0 + split_budget) * B.refs.size()));
    lap("pre-split");
    B.nodes.resize(2 * B.refs.size());
    B.max_workers = (int)std::min(15u, std::max(1u, std::thread::hardware_concurrency()) - 1);
    if (const char *e = getenv("GDPT_BVH_THREADS")) B.max_workers = std::max(0, atoi(e) - 1);
    B.build(0, (int)B.refs.size(), 0);
    lap("binned SAH");
    out.prims.resize(B.refs.size());
    for (size_t i = 0; i < B.refs.size(); i++) out.prims[i] = src[B.refs[i].idx];

    // collapse to 4-wide: a child is opened as long as fewer than four are held, largest surface area first.  The top of the
    // tree is collapsed serially until enough independent subtrees are open, the subtrees then side by side into private
    // node arrays that are appended in subtree order (so the result does not depend on the thread count).
    struct Item { int bnode; int wide; int depth; };
    auto blank = []() { BvhNode n; for (int c = 0; c < 4; c++) {
            n.lox[c] = n.loy[c] = n.loz[c] = std::numeric_limits<float>::infinity();
            n.hix[c] = n.hiy[c] = n.hiz[c] = -std::numeric_limits<float>::infinity();
            n.child[c] = 0; n.count[c] = 0; }
        return n; };
    // one collapse step of `it` inside `nodes`; children that are inner nodes are appended to `nodes` and to `work`
    auto expand = [&](const Item &it, std::vector<BvhNode> &nodes, std::vector<Item> &work) {
        int kids[4]; int nk = 0;
        const BNode &bn = B.nodes[it.bnode];
        if (bn.left < 0) { kids[nk++] = it.bnode; }
        else { kids[nk++] = bn.left; kids[nk++] = bn.right; }
        while (nk < 4) {
            int pick = -1; double pa = -1;
            for (int i = 0; i < nk; i++) {
                const BNode &k = B.nodes[kids[i]];
                if (k.left >= 0) { double ar = k.b.area(); if (ar > pa) { pa = ar; pick = i; } }
            }
            if (pick < 0) break;
            int k = kids[pick];
            kids[pick] = B.nodes[k].left; kids[nk++] = B.nodes[k].right;
        }
        for (int i = 0; i < nk; i++) {
            const BNode &k = B.nodes[kids[i]];
            BvhNode &w = nodes[it.wide];
            w.lox[i] = k.b.lo[0]; w.loy[i] = k.b.lo[1]; w.loz[i] = k.b.lo[2];
            w.hix[i] = k.b.hi[0]; w.hiy[i] = k.b.hi[1]; w.hiz[i] = k.b.hi[2];
            if (k.left < 0) { w.child[i] = ~((k.first << 3) | k.count); w.count[i] = k.count; }
            else {
                nodes.push_back(blank());
                const int c = (int)nodes.size() - 1;
                nodes[it.wide].child[i] = c; nodes[it.wide].count[i] = 0;
                work.push_back({kids[i], c, it.depth + 1});
            }
        }
    };
    out.nodes.push_back(blank());
    std::vector<Item> frontier(1, Item{0, 0, 1});
    out.max_depth = std::max(out.max_depth, 1);
    while (!frontier.empty() && frontier.size() < 256) {      // breadth first: the frontier items are roots of disjoint subtrees
        std::vector<Item> next;
        for (const Item &it : frontier) { out.max_depth = std::max(out.max_depth, it.depth); expand(it, out.nodes, next); }
        frontier.swap(next);
    }
    if (!frontier.empty()) {
        const int nitems = (int)frontier.size();
        std::vector<std::vector<BvhNode>> sub(nitems);
        std::vector<int> sub_depth(nitems, 0);
        const int nthreads = std::max(1, std::min(B.max_workers + 1, nitems));
        std::atomic<int> next_item{0};
        auto worker = [&]() {
            for (int q = next_item.fetch_add(1); q < nitems; q = next_item.fetch_add(1)) {
                std::vector<BvhNode> &nodes = sub[q];
                nodes.push_back(blank());                       // local 0 = the frontier item's own node
                std::vector<Item> work(1, Item{frontier[q].bnode, 0, frontier[q].depth});
                int md = 0;
                while (!work.empty()) { Item it = work.back(); work.pop_back(); md = std::max(md, it.depth); expand(it, nodes, work); }
                sub_depth[q] = md;
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; t++) th.emplace_back(worker);
        worker();
        for (auto &t : th) t.join();
        // append: local node j > 0 of item q becomes global base_q + j - 1; local 0 is written over the frontier node
        for (int q = 0; q < nitems; q++) {
            const int base = (int)out.nodes.size();
            auto fix = [&](BvhNode n) { for (int c = 0; c < 4; c++) if (n.child[c] > 0) n.child[c] = base + n.child[c] - 1; return n; };
            out.nodes[frontier[q].wide] = fix(sub[q][0]);
            for (size_t j = 1; j < sub[q].size(); j++) out.nodes.push_back(fix(sub[q][j]));
            out.max_depth = std::max(out.max_depth, sub_depth[q]);
        }
    }
    lap("collapse to 4");
}

} // namespace gdpt
