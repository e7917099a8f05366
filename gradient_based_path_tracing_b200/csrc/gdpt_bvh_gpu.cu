// BVH construction ON THE DEVICE — the replacement for rtcCommitScene (/root/reference/src/scene.cpp:19-34,
// src/shapes/triangle_mesh.inl:1-22, src/shapes/sphere.inl:151-162) that the traversal kernels (gdpt_trace.cuh) walk.
//
//   k_bvh_prims      one thread per triangle / sphere: the 48-byte primitive record (float-rounded vertices exactly as the
//                    reference hands them to Embree) and its box
//   k_bvh_split      triangle pre-splitting (the spatial splits of Embree's HIGH build quality, scene.cpp:22): every
//                    triangle whose box is larger than a threshold area is cut at the spatial median of its box's longest
//                    axis, the halves re-fitted to the clipped polygon, recursively; one thread per triangle, its pieces on
//                    a private stack.  Run twice: count, exclusive scan, emit.  The threshold comes from a short search so
//                    that the reference count lands near the budget (+200 %).
//   k_bvh_morton     24-bit Morton code of the box centre + histogram; the counting sort of gdpt_wavefront.cuh orders them
//   PLOC             parallel locally-ordered clustering (Meister & Bittner 2017): every cluster looks R = 16 neighbours
//                    to each side along the Morton order for the partner with the smallest merged surface area, mutual
//                    pairs merge into a new binary node, the survivors are compacted in order; ~25 rounds for Sponza
//   k_bvh_leaf_flags SAH leaf decision on subtrees of at most 4 primitives
//   k_bvh_collapse   binary tree -> 4-wide nodes (gdpt_bvh.h layout: 128 B, SoA child bounds), level by level: a child is
//                    opened largest surface area first until four are held; leaves copy their primitive records into one
//                    contiguous run
// The host builder (gdpt_bvh.cpp, binned SAH) stays as the fallback when the device tree is deeper than the traversal
// stack allows, and for GDPT_BVH=host.
#include "gdpt_context.cuh"
#include "gdpt_wavefront.cuh"

namespace gdpt {

namespace {

constexpr int PLOC_RADIUS_MAX = 64;     // shared-memory halo of the neighbour search; the radius itself is a run-time knob
constexpr int PLOC_BLOCK = 256;
constexpr int SPLIT_STACK = 24;         // pieces a thread may hold at once while it cuts one triangle

struct Poly { float v[9][3]; int n; };  // a triangle clipped by axis-aligned planes: at most 3 + 6 vertices

__device__ inline float box_area(const float lo[3], const float hi[3]) {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    return dx * dy + dy * dz + dz * dx;         // half the surface area: only compared and scaled
}
__device__ inline float area4(float4 lo, float4 hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

// ---- primitives
__global__ void k_bvh_prims(const DShape *shapes, int n_shapes, const double *positions, const int *indices, const double *spheres,
                            const int *prim_shape /* [n] shape of primitive i */, const int *prim_local /* [n] index inside the shape */, int n,
                            BvhPrim *prims, float4 *blo, float4 *bhi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = prim_shape[i], t = prim_local[i];
    const DShape sh = shapes[s];
    BvhPrim p;
    float lo[3], hi[3];
    if (sh.kind == 1) {
        const int *ix = indices + 3 * ((size_t)sh.tri_off + t);
        float v[3][3];
        for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) v[c][k] = (float)positions[3 * ((size_t)sh.vtx_off + ix[c]) + k];
        for (int k = 0; k < 3; k++) {
            p.v0[k] = v[0][k]; p.v1[k] = v[1][k]; p.v2[k] = v[2][k];
            lo[k] = fminf(v[0][k], fminf(v[1][k], v[2][k])); hi[k] = fmaxf(v[0][k], fmaxf(v[1][k], v[2][k]));
        }
        p.prim_id = t; p.shape_id = s; p.kind = 0;
    } else {
        const double *sp = spheres + 4 * (size_t)s;
        for (int k = 0; k < 3; k++) {
            p.v0[k] = (float)sp[k]; p.v1[k] = 0.f; p.v2[k] = 0.f;
            lo[k] = __double2float_rd(sp[k] - sp[3]); hi[k] = __double2float_ru(sp[k] + sp[3]);      // conservative after rounding
        }
        p.v1[0] = (float)sp[3];
        p.prim_id = 0; p.shape_id = s; p.kind = 1;
    }
    prims[i] = p;
    blo[i] = make_float4(lo[0], lo[1], lo[2], __int_as_float(i));
    bhi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
}

// ---- pre-splitting
__device__ inline void clip_poly(const Poly &in, int axis, float plane, bool keep_below, Poly &out) {
    out.n = 0;
    for (int i = 0; i < in.n; i++) {
        const float *a = in.v[i], *b = in.v[i + 1 == in.n ? 0 : i + 1];
        const bool ia = keep_below ? a[axis] <= plane : a[axis] >= plane, ib = keep_below ? b[axis] <= plane : b[axis] >= plane;
        if (ia && out.n < 9) { for (int k = 0; k < 3; k++) out.v[out.n][k] = a[k]; out.n++; }
        if (ia != ib && out.n < 9) {
            const float t = (plane - a[axis]) / (b[axis] - a[axis]);
            for (int k = 0; k < 3; k++) out.v[out.n][k] = a[k] + t * (b[k] - a[k]);
            out.v[out.n][axis] = plane; out.n++;
        }
    }
}
// tight box of a clipped polygon inside `lo/hi`, widened by two ulps for the interpolation error; false if empty
__device__ inline bool poly_box(const Poly &p, const float lo[3], const float hi[3], float olo[3], float ohi[3]) {
    if (p.n == 0) return false;
    for (int k = 0; k < 3; k++) {
        float a = p.v[0][k], b = p.v[0][k];
        for (int i = 1; i < p.n; i++) { a = fminf(a, p.v[i][k]); b = fmaxf(b, p.v[i][k]); }
        a = nextafterf(nextafterf(a, -INFINITY), -INFINITY); b = nextafterf(nextafterf(b, INFINITY), INFINITY);
        olo[k] = fmaxf(a, lo[k]); ohi[k] = fminf(b, hi[k]);
        if (olo[k] > ohi[k]) return false;
    }
    return true;
}

struct Piece { float lo[3], hi[3]; Poly poly; };

// emit == nullptr: count the references triangle i produces under `threshold`; otherwise write them at offsets[i]
__global__ void k_bvh_split(const BvhPrim *prims, const float4 *blo, const float4 *bhi, int n, float threshold, int max_pieces,
                            int *counts, const int *offsets, float4 *olo, float4 *ohi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const BvhPrim p = prims[i];
    const float4 l4 = blo[i], h4 = bhi[i];
    int produced = 0;
    const int base = offsets ? offsets[i] : 0;
    auto emit = [&](const float lo[3], const float hi[3]) {
        if (olo) { olo[base + produced] = make_float4(lo[0], lo[1], lo[2], __int_as_float(i)); ohi[base + produced] = make_float4(hi[0], hi[1], hi[2], 0.f); }
        produced++;
    };
    float lo0[3] = {l4.x, l4.y, l4.z}, hi0[3] = {h4.x, h4.y, h4.z};
    if (p.kind != 0 || !(box_area(lo0, hi0) > threshold)) { emit(lo0, hi0); if (counts) counts[i] = produced; return; }
    Piece stack[SPLIT_STACK];
    int sp = 0;
    {
        Piece &q = stack[sp++];
        for (int k = 0; k < 3; k++) { q.lo[k] = lo0[k]; q.hi[k] = hi0[k]; q.poly.v[0][k] = p.v0[k]; q.poly.v[1][k] = p.v1[k]; q.poly.v[2][k] = p.v2[k]; }
        q.poly.n = 3;
    }
    while (sp > 0) {
        Piece cur = stack[--sp];
        int ax = 0; float ext = cur.hi[0] - cur.lo[0];
        for (int k = 1; k < 3; k++) if (cur.hi[k] - cur.lo[k] > ext) { ext = cur.hi[k] - cur.lo[k]; ax = k; }
        const float mid = 0.5f * (cur.lo[ax] + cur.hi[ax]);
        // stop: small enough, cannot be cut, no room on the stack, or the triangle's share of references is used up
        const bool room = sp + 2 <= SPLIT_STACK && produced + sp + 2 <= max_pieces;
        if (!(box_area(cur.lo, cur.hi) > threshold) || !(ext > 0.f) || !(mid > cur.lo[ax] && mid < cur.hi[ax]) || !room) { emit(cur.lo, cur.hi); continue; }
        Piece a, b;
        clip_poly(cur.poly, ax, mid, true, a.poly);
        clip_poly(cur.poly, ax, mid, false, b.poly);
        float llo[3], lhi[3], rlo[3], rhi[3];
        for (int k = 0; k < 3; k++) { llo[k] = rlo[k] = cur.lo[k]; lhi[k] = rhi[k] = cur.hi[k]; }
        lhi[ax] = mid; rlo[ax] = mid;
        const bool hl = poly_box(a.poly, llo, lhi, a.lo, a.hi), hr = poly_box(b.poly, rlo, rhi, b.lo, b.hi);
        if (hl && hr) { stack[sp++] = b; stack[sp++] = a; }
        else if (hl || hr) {
            Piece &only = hl ? a : b;          // the triangle touches one half only: the box just got tighter
            if (box_area(only.lo, only.hi) < box_area(cur.lo, cur.hi)) stack[sp++] = only; else emit(only.lo, only.hi);
        } else emit(cur.lo, cur.hi);
    }
    if (counts) counts[i] = produced;
}

// exclusive scan of n ints by one block (n is a few hundred thousand; the build is a one-off)
__global__ void __launch_bounds__(1024) k_scan_one_block(const int *in, int *out, int n, int *total) {
    __shared__ int part[1024];
    const int t = threadIdx.x, seg = (n + 1023) / 1024, a = min(n, t * seg), b = min(n, a + seg);
    int s = 0;
    for (int i = a; i < b; i++) s += in[i];
    part[t] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) { int v = t >= d ? part[t - d] : 0; __syncthreads(); part[t] += v; __syncthreads(); }
    int run = part[t] - s;
    for (int i = a; i < b; i++) { int v = in[i]; out[i] = run; run += v; }
    if (t == 1023 && total) *total = part[1023];
}

// ---- Morton order
__global__ void k_bvh_centroid_bounds(const float4 *blo, const float4 *bhi, int n, int *bounds /* 6 ordered ints */) {
    auto ord = [](float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; };
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = blo[i], b = bhi[i];
        const float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
        for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], c[k]); hi[k] = fmaxf(hi[k], c[k]); }
    }
    for (int k = 0; k < 3; k++) {
        for (int s = 16; s > 0; s >>= 1) { lo[k] = fminf(lo[k], __shfl_down_sync(0xffffffffu, lo[k], s)); hi[k] = fmaxf(hi[k], __shfl_down_sync(0xffffffffu, hi[k], s)); }
        if ((threadIdx.x & 31) == 0) { atomicMin(bounds + k, ord(lo[k])); atomicMax(bounds + 3 + k, ord(hi[k])); }
    }
}
__device__ inline unsigned spread8(unsigned x) {      // 8 bits -> every third bit
    x = (x | (x << 16)) & 0x030000ffu; x = (x | (x << 8)) & 0x0300f00fu; x = (x | (x << 4)) & 0x030c30c3u; x = (x | (x << 2)) & 0x09249249u;
    return x;
}
__global__ void k_bvh_morton(const float4 *blo, const float4 *bhi, int n, const int *bounds, int *keys, int *hist) {
    auto unord = [](int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); };
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = blo[i], b = bhi[i];
    const float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
    unsigned q[3];
    for (int k = 0; k < 3; k++) {
        const float lo = unord(bounds[k]), hi = unord(bounds[3 + k]), ext = hi - lo;
        const float f = ext > 0.f ? (c[k] - lo) / ext : 0.f;
        q[k] = (unsigned)min(255, max(0, (int)(f * 256.f)));
    }
    const int key = (int)(spread8(q[0]) | (spread8(q[1]) << 1) | (spread8(q[2]) << 2));
    keys[i] = key;
    atomicAdd(hist + key, 1);
}

// The scatter of the counting sort places equal keys in the order its atomics happen to run.  Each run of equal keys is put
// into ascending reference order here (one thread per run; runs are short: a 256^3 grid over the scene), which makes the
// Morton order — and with it the whole tree — the same on every build.
__global__ void k_bvh_order_runs(const int *keys, int *order, int n) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int key = keys[order[p]];
    if (p > 0 && keys[order[p - 1]] == key) return;        // not the first of its run
    int e = p + 1;
    while (e < n && keys[order[e]] == key) e++;
    for (int i = p + 1; i < e; i++) {                       // insertion sort of order[p, e)
        const int v = order[i];
        int j = i - 1;
        while (j >= p && order[j] > v) { order[j + 1] = order[j]; j--; }
        order[j + 1] = v;
    }
}

// ---- PLOC.  Binary nodes: ids 0..n-1 are the leaves (references in Morton order), n.. are merged nodes.
__global__ void k_ploc_leaves(const float4 *blo, const float4 *bhi, const int *order, int n, float4 *nlo, float4 *nhi, int *count, int *cluster) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int r = order[i];
    nlo[i] = blo[r]; nhi[i] = bhi[r];        // .w of nlo keeps the primitive index
    count[i] = 1; cluster[i] = i;
}
__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_nn(const int *cluster, int c, const float4 *nlo, const float4 *nhi, int *nn, int PLOC_RADIUS, int n_leaves, int no_self) {
    __shared__ float4 slo[PLOC_BLOCK + 2 * PLOC_RADIUS_MAX], shi[PLOC_BLOCK + 2 * PLOC_RADIUS_MAX];
    const int start = blockIdx.x * PLOC_BLOCK - PLOC_RADIUS;
    for (int k = threadIdx.x; k < PLOC_BLOCK + 2 * PLOC_RADIUS; k += PLOC_BLOCK) {
        const int j = start + k;
        // .w of the staged lower corner: the primitive of a still unmerged reference (pieces of one pre-split triangle), else -1
        if (j >= 0 && j < c) { const int id = cluster[j]; float4 l = nlo[id]; if (id >= n_leaves) l.w = __int_as_float(-1); slo[k] = l; shi[k] = nhi[id]; }
    }
    __syncthreads();
    const int i = blockIdx.x * PLOC_BLOCK + threadIdx.x;
    if (i >= c) return;
    const float4 a = slo[threadIdx.x + PLOC_RADIUS], b = shi[threadIdx.x + PLOC_RADIUS];
    float best = INFINITY; int bj = -1;
    for (int d = -PLOC_RADIUS; d <= PLOC_RADIUS; d++) {
        const int j = i + d;
        if (d == 0 || j < 0 || j >= c) continue;
        const float4 l = slo[threadIdx.x + PLOC_RADIUS + d], h = shi[threadIdx.x + PLOC_RADIUS + d];
        // two pieces of the same triangle would only rebuild the box the split removed: they may meet higher up, not as a pair
        if (no_self && __float_as_int(a.w) >= 0 && __float_as_int(a.w) == __float_as_int(l.w)) continue;
        const float4 ul = make_float4(fminf(a.x, l.x), fminf(a.y, l.y), fminf(a.z, l.z), 0.f), uh = make_float4(fmaxf(b.x, h.x), fmaxf(b.y, h.y), fmaxf(b.z, h.z), 0.f);
        const float ar = area4(ul, uh);
        if (ar < best) { best = ar; bj = j; }       // ties: the smaller index (d runs upwards)
    }
    nn[i] = bj;
}
__global__ void k_ploc_merge(const int *cluster, const int *nn, int c, float4 *nlo, float4 *nhi, int2 *child, int *count, int n_leaves, int *next_node,
                             int *cluster_out, int *valid) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c) return;
    const int j = nn[i];
    int out = cluster[i], keep = 1;
    if (j >= 0 && nn[j] == i) {
        if (i < j) {
            const int l = cluster[i], r = cluster[j], id = atomicAdd(next_node, 1);
            const float4 al = nlo[l], ah = nhi[l], bl = nlo[r], bh = nhi[r];
            nlo[id] = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), 0.f);
            nhi[id] = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.f);
            child[id - n_leaves] = make_int2(l, r);
            count[id] = count[l] + count[r];
            out = id;
        } else keep = 0;
    }
    cluster_out[i] = out; valid[i] = keep;
}
__global__ void k_ploc_compact(const int *cluster_in, const int *valid, const int *pos, int c, int *cluster_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < c && valid[i]) cluster_out[pos[i]] = cluster_in[i];
}

// ---- leaf decision + collapse
// a subtree of at most 4 primitives becomes a leaf when testing them all is not dearer than one more node step plus its two
// halves (the same one-level SAH comparison as the host builder, primitive cost 1)
__global__ void k_bvh_leaf_flags(const float4 *nlo, const float4 *nhi, const int2 *child, const int *count, int n_leaves, int n_nodes, unsigned char *is_leaf, int max_leaf) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_nodes) return;
    if (id < n_leaves) { is_leaf[id] = 1; return; }
    const int cnt = count[id];
    unsigned char f = 0;
    if (cnt <= max_leaf) {
        const int2 ch = child[id - n_leaves];
        const float a = area4(nlo[id], nhi[id]);
        const float leaf = a * cnt, split = a + area4(nlo[ch.x], nhi[ch.x]) * count[ch.x] + area4(nlo[ch.y], nhi[ch.y]) * count[ch.y];
        f = leaf <= split;
    }
    is_leaf[id] = f;
}

struct CollapseItem { int bnode, wide; };

__global__ void k_bvh_collapse(const CollapseItem *items, int n_items, const float4 *nlo, const float4 *nhi, const int2 *child, const int *count,
                               const unsigned char *is_leaf, int n_leaves, const BvhPrim *src, BvhNode *wide, int *n_wide, BvhPrim *out_prims, int *n_prims,
                               CollapseItem *next_items, int *n_next) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_items) return;
    const CollapseItem it = items[q];
    int kids[4], nk = 0;
    if (is_leaf[it.bnode]) kids[nk++] = it.bnode;
    else { const int2 ch = child[it.bnode - n_leaves]; kids[nk++] = ch.x; kids[nk++] = ch.y; }
    while (nk < 4) {
        int pick = -1; float pa = -1.f;
        for (int i = 0; i < nk; i++) if (!is_leaf[kids[i]]) { const float a = area4(nlo[kids[i]], nhi[kids[i]]); if (a > pa) { pa = a; pick = i; } }
        if (pick < 0) break;
        const int2 ch = child[kids[pick] - n_leaves];
        kids[pick] = ch.x; kids[nk++] = ch.y;
    }
    BvhNode w;
    for (int i = 0; i < 4; i++) { w.lox[i] = w.loy[i] = w.loz[i] = INFINITY; w.hix[i] = w.hiy[i] = w.hiz[i] = -INFINITY; w.child[i] = 0; w.count[i] = 0; }
    for (int i = 0; i < nk; i++) {
        const int k = kids[i];
        const float4 l = nlo[k], h = nhi[k];
        w.lox[i] = l.x; w.loy[i] = l.y; w.loz[i] = l.z; w.hix[i] = h.x; w.hiy[i] = h.y; w.hiz[i] = h.z;
        if (is_leaf[k]) {
            const int cnt = count[k], first = atomicAdd(n_prims, cnt);
            int stack[8], sp = 0, written = 0;
            stack[sp++] = k;
            while (sp > 0) {          // the subtree holds at most 4 primitives
                const int id = stack[--sp];
                if (id < n_leaves) out_prims[first + written++] = src[__float_as_int(nlo[id].w)];
                else { const int2 ch = child[id - n_leaves]; stack[sp++] = ch.y; stack[sp++] = ch.x; }
            }
            w.child[i] = ~((first << 3) | cnt); w.count[i] = cnt;
        } else {
            const int c = atomicAdd(n_wide, 1);
            w.child[i] = c;
            next_items[atomicAdd(n_next, 1)] = CollapseItem{k, c};
        }
    }
    wide[it.wide] = w;
}

}  // namespace

#define BV_LAUNCH(kern, n, block) kern<<<((n) + (block) - 1) / (block), (block), 0, st>>>

// Builds the tree of the context's scene on its device.  Returns 0 and fills c->bvh_nodes / bvh_prims / bvh_depth, or a
// positive value when the caller should fall back to the host builder (tree too deep for the traversal stack).
int build_bvh_device(gdpt_ctx *c, const std::vector<DShape> &shapes, double split_budget) {
    cudaStream_t st = c->stream;
    // primitive list: all triangles of all meshes, one entry per sphere
    std::vector<int> prim_shape, prim_local;
    for (int s = 0; s < (int)shapes.size(); s++) {
        if (shapes[s].kind == 1) for (int t = 0; t < shapes[s].n_tris; t++) { prim_shape.push_back(s); prim_local.push_back(t); }
        else { prim_shape.push_back(s); prim_local.push_back(0); }
    }
    const int n = (int)prim_shape.size();
    if (n == 0) { c->bvh.n_nodes = 0; c->bvh.n_prims = 0; c->bvh_depth = 0; return 0; }
    DevBuf d_ps, d_pl, d_src, d_blo, d_bhi, d_cnt, d_off, d_rlo, d_rhi, d_misc;
    struct Release { std::vector<DevBuf *> b; ~Release() { for (auto *x : b) x->release(); } } rel;
    rel.b = {&d_ps, &d_pl, &d_src, &d_blo, &d_bhi, &d_cnt, &d_off, &d_rlo, &d_rhi, &d_misc};
    if (int r = upload(d_ps, prim_shape)) return r;
    if (int r = upload(d_pl, prim_local)) return r;
    if (int r = d_src.ensure((size_t)n * sizeof(BvhPrim))) return r;
    if (int r = d_blo.ensure((size_t)n * sizeof(float4))) return r;
    if (int r = d_bhi.ensure((size_t)n * sizeof(float4))) return r;
    if (int r = d_misc.ensure(256)) return r;
    int *misc = d_misc.as<int>();              // [0..5] centroid bounds, [8] scan total / counters
    BV_LAUNCH(k_bvh_prims, n, 256)(c->shapes.as<DShape>(), (int)shapes.size(), c->pos_d.as<double>(), c->indices.as<int>(), c->spheres.as<double>(),
                                   d_ps.as<int>(), d_pl.as<int>(), n, d_src.as<BvhPrim>(), d_blo.as<float4>(), d_bhi.as<float4>());
    // ---- pre-split: search the area threshold that spends the budget, then emit
    const float4 *rlo = d_blo.as<float4>(), *rhi = d_bhi.as<float4>();
    int nref = n;
    if (split_budget > 0.0 && n >= 64) {
        if (int r = d_cnt.ensure((size_t)n * sizeof(int))) return r;
        if (int r = d_off.ensure((size_t)n * sizeof(int))) return r;
        const long long budget = (long long)((1.0 + split_budget) * n);
        const int max_pieces = 4096;
        // total box area of the scene's triangles sets the scale: start at (sum of areas) / budget and correct geometrically
        std::vector<float4> hlo(n), hhi(n);
        GD_CUDA(cudaMemcpyAsync(hlo.data(), d_blo.p, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaMemcpyAsync(hhi.data(), d_bhi.p, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        double sum = 0;
        for (int i = 0; i < n; i++) { double dx = hhi[i].x - hlo[i].x, dy = hhi[i].y - hlo[i].y, dz = hhi[i].z - hlo[i].z; sum += dx * dy + dy * dz + dz * dx; }
        double thr = sum / (double)budget;
        int total = n;
        for (int iter = 0; iter < 6 && thr > 0; iter++) {
            BV_LAUNCH(k_bvh_split, n, 128)(d_src.as<BvhPrim>(), d_blo.as<float4>(), d_bhi.as<float4>(), n, (float)thr, max_pieces, d_cnt.as<int>(), nullptr, nullptr, nullptr);
            k_scan_one_block<<<1, 1024, 0, st>>>(d_cnt.as<int>(), d_off.as<int>(), n, misc + 8);
            GD_CUDA(cudaMemcpyAsync(&total, misc + 8, sizeof(int), cudaMemcpyDeviceToHost, st));
            GD_CUDA(cudaStreamSynchronize(st));
            c->stats.kernel_launches += 2;
            const double ratio = (double)(total - n + 1) / (double)(budget - n + 1);      // extra references made / wanted
            if (ratio > 0.9 && ratio < 1.1) break;
            thr *= ratio;                                                                  // pieces scale about inversely with the threshold
        }
        if (total > n) {
            if (int r = d_rlo.ensure((size_t)total * sizeof(float4))) return r;
            if (int r = d_rhi.ensure((size_t)total * sizeof(float4))) return r;
            BV_LAUNCH(k_bvh_split, n, 128)(d_src.as<BvhPrim>(), d_blo.as<float4>(), d_bhi.as<float4>(), n, (float)thr, max_pieces, nullptr, d_off.as<int>(), d_rlo.as<float4>(), d_rhi.as<float4>());
            c->stats.kernel_launches++;
            rlo = d_rlo.as<float4>(); rhi = d_rhi.as<float4>(); nref = total;
        }
    }
    // ---- Morton order of the references
    constexpr int MORTON_BINS = 1 << 24;
    DevBuf d_keys, d_order, d_hist, d_nlo, d_nhi, d_child, d_count, d_cl[2], d_nn, d_valid, d_pos, d_flags, d_items[2];
    rel.b.insert(rel.b.end(), {&d_keys, &d_order, &d_hist, &d_nlo, &d_nhi, &d_child, &d_count, &d_cl[0], &d_cl[1], &d_nn, &d_valid, &d_pos, &d_flags, &d_items[0], &d_items[1]});
    if (int r = d_keys.ensure((size_t)nref * sizeof(int))) return r;
    if (int r = d_order.ensure((size_t)nref * sizeof(int))) return r;
    if (int r = d_hist.ensure((size_t)(MORTON_BINS + 1024) * sizeof(int))) return r;
    const int init_bounds[6] = {0x7f800000, 0x7f800000, 0x7f800000, (int)(0xff800000u ^ 0x7fffffffu), (int)(0xff800000u ^ 0x7fffffffu), (int)(0xff800000u ^ 0x7fffffffu)};
    GD_CUDA(cudaMemcpyAsync(misc, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
    GD_CUDA(cudaMemsetAsync(d_hist.p, 0, (size_t)MORTON_BINS * sizeof(int), st));
    k_bvh_centroid_bounds<<<148 * 4, 256, 0, st>>>(rlo, rhi, nref, misc);
    BV_LAUNCH(k_bvh_morton, nref, 256)(rlo, rhi, nref, misc, d_keys.as<int>(), d_hist.as<int>());
    const int n_chunks = MORTON_BINS / SORT_CHUNK;
    k_sort_scan<<<n_chunks, 1024, 0, st>>>(d_hist.as<int>(), d_hist.as<int>() + MORTON_BINS);
    k_sort_chunks<<<1, 1024, 0, st>>>(d_hist.as<int>() + MORTON_BINS, n_chunks);
    BV_LAUNCH(k_sort_scatter, nref, 256)(d_keys.as<int>(), d_hist.as<int>(), d_hist.as<int>() + MORTON_BINS, nref, d_order.as<int>(), nullptr);
    BV_LAUNCH(k_bvh_order_runs, nref, 256)(d_keys.as<int>(), d_order.as<int>(), nref);
    c->stats.kernel_launches += 7;
    // ---- PLOC
    const int max_nodes = 2 * nref;
    if (int r = d_nlo.ensure((size_t)max_nodes * sizeof(float4))) return r;
    if (int r = d_nhi.ensure((size_t)max_nodes * sizeof(float4))) return r;
    if (int r = d_child.ensure((size_t)nref * sizeof(int2))) return r;
    if (int r = d_count.ensure((size_t)max_nodes * sizeof(int))) return r;
    for (auto &b : d_cl) if (int r = b.ensure((size_t)nref * sizeof(int))) return r;
    if (int r = d_nn.ensure((size_t)nref * sizeof(int))) return r;
    if (int r = d_valid.ensure((size_t)nref * sizeof(int))) return r;
    if (int r = d_pos.ensure((size_t)nref * sizeof(int))) return r;
    float4 *nlo = d_nlo.as<float4>(), *nhi = d_nhi.as<float4>();
    BV_LAUNCH(k_ploc_leaves, nref, 256)(rlo, rhi, d_order.as<int>(), nref, nlo, nhi, d_count.as<int>(), d_cl[0].as<int>());
    GD_CUDA(cudaMemcpyAsync(misc + 16, &nref, sizeof(int), cudaMemcpyHostToDevice, st));     // next binary node id
    int radius = 8, max_leaf = 4;           // B200 sweep: profiles/r2_bvh_device.txt (tuning knobs: tools/gpu_r2_probe.py bvh)
    if (const char *e = getenv("GDPT_PLOC_RADIUS")) radius = std::min(PLOC_RADIUS_MAX, std::max(1, atoi(e)));
    if (const char *e = getenv("GDPT_BVH_MAX_LEAF")) max_leaf = std::min(4, std::max(1, atoi(e)));
    int no_self = nref > n ? 1 : 0;
    if (const char *e = getenv("GDPT_PLOC_NOSELF")) no_self = atoi(e);
    int cnum = nref, cur = 0, rounds = 0;
    while (cnum > 1) {
        k_ploc_nn<<<(cnum + PLOC_BLOCK - 1) / PLOC_BLOCK, PLOC_BLOCK, 0, st>>>(d_cl[cur].as<int>(), cnum, nlo, nhi, d_nn.as<int>(), radius, nref, no_self);
        BV_LAUNCH(k_ploc_merge, cnum, 256)(d_cl[cur].as<int>(), d_nn.as<int>(), cnum, nlo, nhi, d_child.as<int2>(), d_count.as<int>(), nref, misc + 16, d_cl[cur ^ 1].as<int>(), d_valid.as<int>());
        k_scan_one_block<<<1, 1024, 0, st>>>(d_valid.as<int>(), d_pos.as<int>(), cnum, misc + 8);
        BV_LAUNCH(k_ploc_compact, cnum, 256)(d_cl[cur ^ 1].as<int>(), d_valid.as<int>(), d_pos.as<int>(), cnum, d_cl[cur].as<int>());
        int next = 0;
        GD_CUDA(cudaMemcpyAsync(&next, misc + 8, sizeof(int), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        c->stats.kernel_launches += 4;
        if (next >= cnum || next < 1) {
            if (no_self) { no_self = 0; continue; }      // only pieces of one triangle are left near each other: let them pair
            return 1;                                    // no mutual pair (cannot happen with the index tie rule; be safe): host builder
        }
        cnum = next; rounds++;
    }
    int n_bin = 0, root = 0;
    GD_CUDA(cudaMemcpyAsync(&n_bin, misc + 16, sizeof(int), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaMemcpyAsync(&root, d_cl[cur].p, sizeof(int), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));
    // ---- leaf flags, collapse to 4-wide
    if (int r = d_flags.ensure((size_t)n_bin)) return r;
    BV_LAUNCH(k_bvh_leaf_flags, n_bin, 256)(nlo, nhi, d_child.as<int2>(), d_count.as<int>(), nref, n_bin, d_flags.as<unsigned char>(), max_leaf);
    for (auto &b : d_items) if (int r = b.ensure((size_t)nref * sizeof(CollapseItem))) return r;
    if (int r = c->bvh_nodes.ensure((size_t)std::max(nref, 1) * sizeof(BvhNode))) return r;
    if (int r = c->bvh_prims.ensure((size_t)nref * sizeof(BvhPrim))) return r;
    const CollapseItem first{root, 0};
    const int counters0[3] = {1, 0, 0};                   // wide nodes allocated, primitive records written, items of the next level
    GD_CUDA(cudaMemcpyAsync(d_items[0].p, &first, sizeof(first), cudaMemcpyHostToDevice, st));
    GD_CUDA(cudaMemcpyAsync(misc + 24, counters0, sizeof(counters0), cudaMemcpyHostToDevice, st));
    int n_items = 1, depth = 0, qi = 0;
    while (n_items > 0) {
        depth++;
        GD_CUDA(cudaMemsetAsync(misc + 26, 0, sizeof(int), st));
        BV_LAUNCH(k_bvh_collapse, n_items, 128)(d_items[qi].as<CollapseItem>(), n_items, nlo, nhi, d_child.as<int2>(), d_count.as<int>(), d_flags.as<unsigned char>(), nref,
                                                d_src.as<BvhPrim>(), c->bvh_nodes.as<BvhNode>(), misc + 24, c->bvh_prims.as<BvhPrim>(), misc + 25,
                                                d_items[qi ^ 1].as<CollapseItem>(), misc + 26);
        GD_CUDA(cudaMemcpyAsync(&n_items, misc + 26, sizeof(int), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        c->stats.kernel_launches++;
        qi ^= 1;
        if (depth > 4096) return fail(GDPT_ERR_CUDA, "BVH collapse does not terminate");
    }
    int fin[2] = {0, 0};
    GD_CUDA(cudaMemcpyAsync(fin, misc + 24, sizeof(fin), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));
    GD_CUDA(cudaGetLastError());
    if (getenv("GDPT_VERBOSE")) fprintf(stderr, "[build_bvh_device] %d primitives, %d references, %d PLOC rounds, %d wide nodes, %d primitive records, depth %d\n", n, nref, rounds, fin[0], fin[1], depth);
    if (3 * depth + 1 > STACK_TOTAL) return 1;            // deeper than the traversal stack: the caller falls back to the host builder
    c->bvh.nodes = c->bvh_nodes.as<float4>(); c->bvh.prims = c->bvh_prims.as<float4>();
    c->bvh.n_nodes = fin[0]; c->bvh.n_prims = fin[1];
    c->bvh_depth = depth;
    return 0;
}

}  // namespace gdpt
