// TEST INFRASTRUCTURE — not part of the product.
//
// Stand-in for the 21 Embree 4.3.0 entry points lajolla calls (Embree itself is
// absent: the reference's `embree` submodule is empty and the prebuilt
// libembree4.so is a missing blob).  It lets the UNMODIFIED reference sources
// under /root/reference/src link and run as the CPU oracle / CPU baseline
// (oracle/_ref).  Call sites it serves:
//   rtcNewDevice/rtcReleaseDevice      src/main.cpp:30,49
//   scene build                        src/scene.cpp:20-31,58
//   triangle geometry                  src/shapes/triangle_mesh.inl:2-20
//   user (sphere) geometry             src/shapes/sphere.inl:152-160
//   rtcIntersect1 / rtcOccluded1       src/intersection.cpp:32,83
//
// Semantics (documented tie rule, since nothing in the reference pins it):
//   * triangles are tested with Moeller-Trumbore evaluated in DOUBLE on the
//     float-rounded vertices / float-rounded ray Embree would receive;
//     accepted iff tnear < t <= tfar; u,v follow Embree: p=(1-u-v)v0+u v1+v v2;
//     Ng = cross(v1-v0, v2-v0), unnormalised.
//   * closest hit = smallest t, ties broken by smallest geomID then primID.
//   * user geometry goes through the registered callbacks (the sphere maths
//     lives in the reference: src/shapes/sphere.inl:40-106).
//   * the returned ray.tfar is the hit distance rounded to float, as Embree does.
// Acceleration (so the timed CPU baseline is not handicapped by the stand-in): a binned-SAH
// binary tree collapsed into 4-wide nodes with float bounds stored SoA, walked one ray at a
// time with SSE slab tests of the four children at once and front-to-back ordering — the
// layout and traversal style of Embree's BVH4 single-ray kernels.  The boxes are exact (they
// are min/max of float data) and the slab test is widened by a few ulp, so the traversal is
// conservative and the accepted hit is decided only by the double-precision primitive tests.
#include <embree4/rtcore.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <vector>
#include <x86intrin.h>

namespace {

// Time spent inside rtcIntersect1 / rtcOccluded1, per thread (TSC ticks), so the bench can state which fraction of
// the reference's CPU time is this stand-in's ray casting rather than the reference's own code (bench.py: rtc_fraction).
struct RtcCounter { unsigned long long ticks = 0, calls = 0; };
std::mutex g_counter_mutex;
std::vector<RtcCounter *> g_counters;
RtcCounter *my_counter() {
    thread_local RtcCounter *c = nullptr;
    if (!c) { c = new RtcCounter(); std::lock_guard<std::mutex> lk(g_counter_mutex); g_counters.push_back(c); }
    return c;
}
struct RtcTimer {
    RtcCounter *c; unsigned long long t0;
    RtcTimer() : c(my_counter()), t0(__rdtsc()) {}
    ~RtcTimer() { c->ticks += __rdtsc() - t0; c->calls++; }
};

struct Geom {
    RTCGeometryType type;
    int refcount = 1;
    // triangle data
    char *vbuf = nullptr; size_t vstride = 0, vcount = 0;
    char *ibuf = nullptr; size_t istride = 0, icount = 0;
    // user data
    void *user = nullptr;
    unsigned nuser = 0;
    RTCBoundsFunction bounds_fn = nullptr;
    void *bounds_ptr = nullptr;
    RTCIntersectFunctionN isect_fn = nullptr;
    RTCOccludedFunctionN occl_fn = nullptr;
};

struct Prim { unsigned geom, prim; double lo[3], hi[3]; };

struct Node {
    double lo[3], hi[3];
    int left, right;      // children when count == 0
    int first, count;     // leaf range in prim order
};

// 4-wide node, SoA over children.  child > 0: wide node index; child < 0: leaf, ~child indexes `leaves`; 0: empty slot
struct alignas(16) Wide { float lo[3][4], hi[3][4]; int child[4]; int pad[4]; };
struct Leaf { int first, count; };
// leaf-order primitive record: float vertices exactly as the geometry buffers hold them
struct PackedPrim { float v[3][3]; unsigned geom, prim; int is_tri; };

struct SceneImpl {
    int refcount = 1;
    std::vector<Geom *> geoms;
    std::vector<Prim> prims;
    std::vector<Node> nodes;
    std::vector<Wide> wide;
    std::vector<Leaf> leaves;
    std::vector<PackedPrim> packed;
    float blo[3], bhi[3];
};

struct DeviceImpl { int refcount = 1; };

inline const float *vtx(const Geom *g, unsigned i) { return (const float *)(g->vbuf + g->vstride * i); }
inline const unsigned *tri(const Geom *g, unsigned i) { return (const unsigned *)(g->ibuf + g->istride * i); }

void release_geom(Geom *g) {
    if (--g->refcount == 0) { free(g->vbuf); free(g->ibuf); delete g; }
}

int build_node(SceneImpl *s, int first, int count) {
    Node n;
    for (int k = 0; k < 3; k++) { n.lo[k] = std::numeric_limits<double>::infinity(); n.hi[k] = -n.lo[k]; }
    double clo[3], chi[3];
    for (int k = 0; k < 3; k++) { clo[k] = n.lo[k]; chi[k] = n.hi[k]; }
    for (int i = first; i < first + count; i++) {
        const Prim &p = s->prims[i];
        for (int k = 0; k < 3; k++) {
            n.lo[k] = std::min(n.lo[k], p.lo[k]); n.hi[k] = std::max(n.hi[k], p.hi[k]);
            double c = 0.5 * (p.lo[k] + p.hi[k]);
            clo[k] = std::min(clo[k], c); chi[k] = std::max(chi[k], c);
        }
    }
    n.left = n.right = -1; n.first = first; n.count = count;
    int id = (int)s->nodes.size();
    s->nodes.push_back(n);
    if (count <= 2) return id;
    // binned SAH over the three axes
    const int NB = 16;
    double best_cost = std::numeric_limits<double>::infinity();
    int best_axis = -1, best_bin = -1;
    auto area = [](const double *lo, const double *hi) {
        double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return 2 * (dx * dy + dy * dz + dz * dx);
    };
    for (int ax = 0; ax < 3; ax++) {
        double ext = chi[ax] - clo[ax];
        if (!(ext > 0)) continue;
        int cnt[NB] = {0};
        double blo[NB][3], bhi[NB][3];
        for (int b = 0; b < NB; b++) for (int k = 0; k < 3; k++) { blo[b][k] = std::numeric_limits<double>::infinity(); bhi[b][k] = -blo[b][k]; }
        for (int i = first; i < first + count; i++) {
            const Prim &p = s->prims[i];
            double c = 0.5 * (p.lo[ax] + p.hi[ax]);
            int b = std::min(NB - 1, (int)(NB * (c - clo[ax]) / ext));
            cnt[b]++;
            for (int k = 0; k < 3; k++) { blo[b][k] = std::min(blo[b][k], p.lo[k]); bhi[b][k] = std::max(bhi[b][k], p.hi[k]); }
        }
        double rarea[NB]; int rcnt[NB];
        double lo[3], hi[3];
        for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<double>::infinity(); hi[k] = -lo[k]; }
        int c = 0;
        for (int b = NB - 1; b > 0; b--) {
            c += cnt[b];
            for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], blo[b][k]); hi[k] = std::max(hi[k], bhi[b][k]); }
            rarea[b] = c ? area(lo, hi) : 0; rcnt[b] = c;
        }
        for (int k = 0; k < 3; k++) { lo[k] = std::numeric_limits<double>::infinity(); hi[k] = -lo[k]; }
        c = 0;
        for (int b = 0; b < NB - 1; b++) {
            c += cnt[b];
            for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], blo[b][k]); hi[k] = std::max(hi[k], bhi[b][k]); }
            if (c == 0 || rcnt[b + 1] == 0) continue;
            double cost = area(lo, hi) * c + rarea[b + 1] * rcnt[b + 1];
            if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = b; }
        }
    }
    int mid;
    if (best_axis < 0) {
        if (count <= 8) return id;
        mid = first + count / 2;  // identical centroids: split in the middle
    } else {
        double leaf_cost = area(n.lo, n.hi) * count;
        if (count <= 4 && best_cost >= leaf_cost) return id;
        double ext = chi[best_axis] - clo[best_axis];
        auto it = std::partition(s->prims.begin() + first, s->prims.begin() + first + count, [&](const Prim &p) {
            double c = 0.5 * (p.lo[best_axis] + p.hi[best_axis]);
            int b = std::min(NB - 1, (int)(NB * (c - clo[best_axis]) / ext));
            return b <= best_bin;
        });
        mid = (int)(it - s->prims.begin());
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    int l = build_node(s, first, mid - first);
    int r = build_node(s, mid, first + count - mid);
    s->nodes[id].left = l; s->nodes[id].right = r; s->nodes[id].count = 0;
    return id;
}

struct RayD { double o[3], d[3], inv[3]; };

// Moeller-Trumbore in double on float vertices. Returns true and (t,u,v,Ng) when the line hits.
inline bool tri_hit(const float *a, const float *b, const float *c, const RayD &r, double &t, double &u, double &v, double Ng[3]) {
    double e1[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]};
    double e2[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
    double p[3] = {r.d[1] * e2[2] - r.d[2] * e2[1], r.d[2] * e2[0] - r.d[0] * e2[2], r.d[0] * e2[1] - r.d[1] * e2[0]};
    double det = e1[0] * p[0] + e1[1] * p[1] + e1[2] * p[2];
    if (det == 0 || !(det == det)) return false;
    double inv = 1.0 / det;
    double s[3] = {r.o[0] - a[0], r.o[1] - a[1], r.o[2] - a[2]};
    u = (s[0] * p[0] + s[1] * p[1] + s[2] * p[2]) * inv;
    if (!(u >= 0 && u <= 1)) return false;
    double q[3] = {s[1] * e1[2] - s[2] * e1[1], s[2] * e1[0] - s[0] * e1[2], s[0] * e1[1] - s[1] * e1[0]};
    v = (r.d[0] * q[0] + r.d[1] * q[1] + r.d[2] * q[2]) * inv;
    if (!(v >= 0 && u + v <= 1)) return false;
    t = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
    Ng[0] = e1[1] * e2[2] - e1[2] * e2[1];
    Ng[1] = e1[2] * e2[0] - e1[0] * e2[2];
    Ng[2] = e1[0] * e2[1] - e1[1] * e2[0];
    return true;
}

// The float ray as SSE broadcasts, with the near/far plane of every axis picked once from the sign of the direction.
struct RayF4 {
    __m128 o[3], inv[3];
    int nearhi[3];      // 1: the near plane of this axis is `hi`
};
inline RayF4 make_ray4(const RTCRay &ray) {
    RayF4 r;
    const float o[3] = {ray.org_x, ray.org_y, ray.org_z}, d[3] = {ray.dir_x, ray.dir_y, ray.dir_z};
    for (int k = 0; k < 3; k++) {
        r.o[k] = _mm_set1_ps(o[k]); r.inv[k] = _mm_set1_ps(1.0f / d[k]);
        r.nearhi[k] = std::signbit(d[k]) ? 1 : 0;
    }
    return r;
}

// Slab test of the four children of `n` against [tnear, tfar]: bit k of the result is set when child k may hold a hit,
// dist[k] is its entry distance.  A NaN plane distance (0 * inf: the ray runs inside a bounding plane) is ignored by
// putting the candidate FIRST in min/max (SSE returns the second operand when either is NaN).
inline int slab4(const Wide &n, const RayF4 &r, float tnear, float tfar, float dist[4]) {
    __m128 t0 = _mm_set1_ps(tnear), t1 = _mm_set1_ps(tfar);
    for (int k = 0; k < 3; k++) {
        const __m128 lo = _mm_load_ps(n.lo[k]), hi = _mm_load_ps(n.hi[k]);
        const __m128 pn = r.nearhi[k] ? hi : lo, pf = r.nearhi[k] ? lo : hi;
        t0 = _mm_max_ps(_mm_mul_ps(_mm_sub_ps(pn, r.o[k]), r.inv[k]), t0);
        t1 = _mm_min_ps(_mm_mul_ps(_mm_sub_ps(pf, r.o[k]), r.inv[k]), t1);
    }
    // conservative: each distance carries two float roundings (t0 >= tnear >= 0 here)
    const __m128 ok = _mm_cmple_ps(_mm_mul_ps(t0, _mm_set1_ps(0.999999f)), _mm_add_ps(_mm_mul_ps(t1, _mm_set1_ps(1.000001f)), _mm_set1_ps(1e-30f)));
    _mm_storeu_ps(dist, t0);
    int m = _mm_movemask_ps(ok);
    for (int k = 0; k < 4; k++) if (n.child[k] == 0) m &= ~(1 << k);
    return m;
}

inline RayD make_ray(const RTCRay &ray) {
    RayD r;
    r.o[0] = ray.org_x; r.o[1] = ray.org_y; r.o[2] = ray.org_z;
    r.d[0] = ray.dir_x; r.d[1] = ray.dir_y; r.d[2] = ray.dir_z;
    for (int k = 0; k < 3; k++) r.inv[k] = 1.0 / r.d[k];
    return r;
}

} // namespace

extern "C" {

RTCDevice rtcNewDevice(const char *) { return (RTCDevice) new DeviceImpl(); }
void rtcReleaseDevice(RTCDevice d) { delete (DeviceImpl *)d; }

RTCScene rtcNewScene(RTCDevice) { return (RTCScene) new SceneImpl(); }
void rtcReleaseScene(RTCScene hs) {
    SceneImpl *s = (SceneImpl *)hs;
    if (--s->refcount == 0) { for (Geom *g : s->geoms) release_geom(g); delete s; }
}
void rtcSetSceneBuildQuality(RTCScene, enum RTCBuildQuality) {}
void rtcSetSceneFlags(RTCScene, enum RTCSceneFlags) {}

RTCGeometry rtcNewGeometry(RTCDevice, enum RTCGeometryType type) {
    Geom *g = new Geom(); g->type = type; return (RTCGeometry)g;
}
unsigned int rtcAttachGeometry(RTCScene hs, RTCGeometry hg) {
    SceneImpl *s = (SceneImpl *)hs; Geom *g = (Geom *)hg;
    g->refcount++; s->geoms.push_back(g);
    return (unsigned)s->geoms.size() - 1;
}
void rtcCommitGeometry(RTCGeometry) {}
void rtcReleaseGeometry(RTCGeometry hg) { release_geom((Geom *)hg); }
void rtcSetGeometryVertexAttributeCount(RTCGeometry, unsigned int) {}
void *rtcSetNewGeometryBuffer(RTCGeometry hg, enum RTCBufferType type, unsigned int, enum RTCFormat,
                              size_t byteStride, size_t itemCount) {
    Geom *g = (Geom *)hg;
    char *buf = (char *)calloc(byteStride * itemCount + 16, 1);
    if (type == RTC_BUFFER_TYPE_VERTEX) { g->vbuf = buf; g->vstride = byteStride; g->vcount = itemCount; }
    else if (type == RTC_BUFFER_TYPE_INDEX) { g->ibuf = buf; g->istride = byteStride; g->icount = itemCount; }
    else { free(buf); return nullptr; }
    return buf;
}
void rtcSetGeometryUserData(RTCGeometry hg, void *ptr) { ((Geom *)hg)->user = ptr; }
void rtcSetGeometryUserPrimitiveCount(RTCGeometry hg, unsigned int n) { ((Geom *)hg)->nuser = n; }
void rtcSetGeometryBoundsFunction(RTCGeometry hg, RTCBoundsFunction f, void *p) { ((Geom *)hg)->bounds_fn = f; ((Geom *)hg)->bounds_ptr = p; }
void rtcSetGeometryIntersectFunction(RTCGeometry hg, RTCIntersectFunctionN f) { ((Geom *)hg)->isect_fn = f; }
void rtcSetGeometryOccludedFunction(RTCGeometry hg, RTCOccludedFunctionN f) { ((Geom *)hg)->occl_fn = f; }

void rtcCommitScene(RTCScene hs) {
    SceneImpl *s = (SceneImpl *)hs;
    s->prims.clear(); s->nodes.clear();
    for (int k = 0; k < 3; k++) { s->blo[k] = std::numeric_limits<float>::infinity(); s->bhi[k] = -s->blo[k]; }
    for (unsigned gi = 0; gi < s->geoms.size(); gi++) {
        Geom *g = s->geoms[gi];
        if (g->type == RTC_GEOMETRY_TYPE_TRIANGLE) {
            for (unsigned t = 0; t < g->icount; t++) {
                Prim p; p.geom = gi; p.prim = t;
                const unsigned *ix = tri(g, t);
                for (int k = 0; k < 3; k++) { p.lo[k] = std::numeric_limits<double>::infinity(); p.hi[k] = -p.lo[k]; }
                for (int c = 0; c < 3; c++) {
                    const float *v = vtx(g, ix[c]);
                    for (int k = 0; k < 3; k++) { p.lo[k] = std::min(p.lo[k], (double)v[k]); p.hi[k] = std::max(p.hi[k], (double)v[k]); }
                }
                s->prims.push_back(p);
            }
        } else if (g->type == RTC_GEOMETRY_TYPE_USER) {
            for (unsigned t = 0; t < g->nuser; t++) {
                RTCBounds b; memset(&b, 0, sizeof(b));
                RTCBoundsFunctionArguments a; a.geometryUserPtr = g->user; a.primID = t; a.timeStep = 0; a.bounds_o = &b;
                g->bounds_fn(&a);
                Prim p; p.geom = gi; p.prim = t;
                p.lo[0] = b.lower_x; p.lo[1] = b.lower_y; p.lo[2] = b.lower_z;
                p.hi[0] = b.upper_x; p.hi[1] = b.upper_y; p.hi[2] = b.upper_z;
                s->prims.push_back(p);
            }
        }
    }
    for (const Prim &p : s->prims)
        for (int k = 0; k < 3; k++) { s->blo[k] = std::min(s->blo[k], (float)p.lo[k]); s->bhi[k] = std::max(s->bhi[k], (float)p.hi[k]); }
    s->wide.clear(); s->leaves.clear(); s->packed.clear();
    if (s->prims.empty()) return;
    s->nodes.reserve(2 * s->prims.size()); build_node(s, 0, (int)s->prims.size());
    // leaf-order primitive records
    s->packed.resize(s->prims.size());
    for (size_t i = 0; i < s->prims.size(); i++) {
        const Prim &p = s->prims[i]; PackedPrim &q = s->packed[i];
        const Geom *g = s->geoms[p.geom];
        q.geom = p.geom; q.prim = p.prim; q.is_tri = g->type == RTC_GEOMETRY_TYPE_TRIANGLE;
        memset(q.v, 0, sizeof(q.v));
        if (q.is_tri) { const unsigned *ix = tri(g, p.prim); for (int c = 0; c < 3; c++) memcpy(q.v[c], vtx(g, ix[c]), 12); }
    }
    // collapse the binary tree into 4-wide nodes: open the child with the largest surface area until four are held
    auto area = [](const Node &n) { double dx = n.hi[0] - n.lo[0], dy = n.hi[1] - n.lo[1], dz = n.hi[2] - n.lo[2]; return dx * dy + dy * dz + dz * dx; };
    auto new_wide = [&]() {
        Wide w; memset(&w, 0, sizeof(w));
        for (int k = 0; k < 3; k++) for (int c = 0; c < 4; c++) { w.lo[k][c] = std::numeric_limits<float>::infinity(); w.hi[k][c] = -w.lo[k][c]; }
        s->wide.push_back(w); return (int)s->wide.size() - 1; };
    struct Item { int bnode, wide; };
    std::vector<Item> work;
    work.push_back({0, new_wide()});
    while (!work.empty()) {
        Item it = work.back(); work.pop_back();
        int kids[4], nk = 0;
        const Node &bn = s->nodes[it.bnode];
        if (bn.count != 0) kids[nk++] = it.bnode; else { kids[nk++] = bn.left; kids[nk++] = bn.right; }
        while (nk < 4) {
            int pick = -1; double pa = -1;
            for (int i = 0; i < nk; i++) { const Node &k = s->nodes[kids[i]]; if (k.count == 0 && area(k) > pa) { pa = area(k); pick = i; } }
            if (pick < 0) break;
            int k = kids[pick]; kids[pick] = s->nodes[k].left; kids[nk++] = s->nodes[k].right;
        }
        for (int i = 0; i < nk; i++) {
            const Node &k = s->nodes[kids[i]];
            for (int a = 0; a < 3; a++) {   // the double boxes hold float values (min/max of float data): the casts are exact
                float lo = (float)k.lo[a], hi = (float)k.hi[a];
                if ((double)lo > k.lo[a]) lo = std::nextafter(lo, -std::numeric_limits<float>::infinity());
                if ((double)hi < k.hi[a]) hi = std::nextafter(hi, std::numeric_limits<float>::infinity());
                s->wide[it.wide].lo[a][i] = lo; s->wide[it.wide].hi[a][i] = hi;
            }
            if (k.count != 0) { s->leaves.push_back({k.first, k.count}); s->wide[it.wide].child[i] = ~((int)s->leaves.size() - 1); }
            else { int c = new_wide(); s->wide[it.wide].child[i] = c; work.push_back({kids[i], c}); }
        }
    }
}

void rtcGetSceneBounds(RTCScene hs, struct RTCBounds *b) {
    SceneImpl *s = (SceneImpl *)hs;
    b->lower_x = s->blo[0]; b->lower_y = s->blo[1]; b->lower_z = s->blo[2]; b->align0 = 0;
    b->upper_x = s->bhi[0]; b->upper_y = s->bhi[1]; b->upper_z = s->bhi[2]; b->align1 = 0;
}

void rtcIntersect1(RTCScene hs, struct RTCRayHit *rh, struct RTCIntersectArguments *args) {
    SceneImpl *s = (SceneImpl *)hs;
    if (s->wide.empty()) return;
    RtcTimer timer;
    const float tnear_f = rh->ray.tnear;
    if (!(rh->ray.tfar > tnear_f)) return;          // a hit needs tnear < t <= tfar
    RTCRayQueryContext local_ctx; rtcInitRayQueryContext(&local_ctx);
    RTCRayQueryContext *ctx = (args && args->context) ? args->context : &local_ctx;
    const RayD r = make_ray(rh->ray);
    const RayF4 r4 = make_ray4(rh->ray);
    const double tnear = rh->ray.tnear;
    double best = rh->ray.tfar;
    unsigned bg = RTC_INVALID_GEOMETRY_ID, bp = RTC_INVALID_GEOMETRY_ID;
    struct Entry { int node; float dist; };
    Entry stack[192]; int sp = 0;
    int cur = 0;                                     // > 0 or the root: wide node; < 0: leaf
    bool have = true;
    while (have) {
        if (cur >= 0) {
            float dist[4];
            const Wide &n = s->wide[cur];
            int m = slab4(n, r4, tnear_f, (float)std::min(best, (double)std::numeric_limits<float>::max()), dist);
            // order the surviving children front to back (insertion sort of at most 4)
            int idx[4], cnt = 0;
            for (int k = 0; k < 4; k++) if (m >> k & 1) { int j = cnt++; while (j > 0 && dist[idx[j - 1]] > dist[k]) { idx[j] = idx[j - 1]; j--; } idx[j] = k; }
            for (int j = cnt - 1; j >= 1; j--) stack[sp++] = {n.child[idx[j]], dist[idx[j]]};
            if (cnt) { cur = n.child[idx[0]]; continue; }
        } else {
            const Leaf lf = s->leaves[~cur];
            for (int i = lf.first; i < lf.first + lf.count; i++) {
                const PackedPrim &p = s->packed[i];
                if (p.is_tri) {
                    double t, u, v, Ng[3];
                    if (!tri_hit(p.v[0], p.v[1], p.v[2], r, t, u, v, Ng)) continue;
                    if (!(t > tnear && t <= best)) continue;
                    if (t == best && bg != RTC_INVALID_GEOMETRY_ID && !(p.geom < bg || (p.geom == bg && p.prim < bp))) continue;
                    best = t; bg = p.geom; bp = p.prim;
                    rh->ray.tfar = (float)t;
                    rh->hit.Ng_x = (float)Ng[0]; rh->hit.Ng_y = (float)Ng[1]; rh->hit.Ng_z = (float)Ng[2];
                    rh->hit.u = (float)u; rh->hit.v = (float)v;
                    rh->hit.primID = p.prim; rh->hit.geomID = p.geom;
                    rh->hit.instID[0] = RTC_INVALID_GEOMETRY_ID;
                } else {
                    Geom *g = s->geoms[p.geom];
                    if (!g->isect_fn) continue;
                    int valid = -1;
                    float before = rh->ray.tfar;
                    RTCIntersectFunctionNArguments a;
                    a.valid = &valid; a.geometryUserPtr = g->user; a.primID = p.prim; a.context = ctx;
                    a.rayhit = (RTCRayHitN *)rh; a.N = 1; a.geomID = p.geom;
                    g->isect_fn(&a);
                    if (rh->ray.tfar != before || (rh->hit.geomID == p.geom && rh->hit.primID == p.prim && bg != p.geom)) {
                        best = rh->ray.tfar; bg = p.geom; bp = p.prim;
                    }
                }
            }
        }
        // pop the next entry that can still hold a closer (or equally close: tie rule) hit
        have = false;
        while (sp) { Entry e = stack[--sp]; if ((double)e.dist * 0.999999 <= best) { cur = e.node; have = true; break; } }
    }
}

void rtcOccluded1(RTCScene hs, struct RTCRay *ray, struct RTCOccludedArguments *args) {
    SceneImpl *s = (SceneImpl *)hs;
    if (s->wide.empty()) return;
    RtcTimer timer;
    if (!(ray->tfar > ray->tnear)) return;
    RTCRayQueryContext local_ctx; rtcInitRayQueryContext(&local_ctx);
    RTCRayQueryContext *ctx = (args && args->context) ? args->context : &local_ctx;
    const RayD r = make_ray(*ray);
    const RayF4 r4 = make_ray4(*ray);
    const double tnear = ray->tnear, tfar = ray->tfar;
    const float tfar_f = std::min(ray->tfar, std::numeric_limits<float>::max());
    int stack[192]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const int cur = stack[--sp];
        if (cur >= 0) {
            float dist[4];
            const Wide &n = s->wide[cur];
            int m = slab4(n, r4, ray->tnear, tfar_f, dist);
            for (int k = 0; k < 4; k++) if (m >> k & 1) stack[sp++] = n.child[k];
            continue;
        }
        const Leaf lf = s->leaves[~cur];
        for (int i = lf.first; i < lf.first + lf.count; i++) {
            const PackedPrim &p = s->packed[i];
            if (p.is_tri) {
                double t, u, v, Ng[3];
                if (tri_hit(p.v[0], p.v[1], p.v[2], r, t, u, v, Ng) && t > tnear && t <= tfar) {
                    ray->tfar = -std::numeric_limits<float>::infinity(); return;
                }
            } else {
                Geom *g = s->geoms[p.geom];
                if (!g->occl_fn) continue;
                int valid = -1;
                RTCOccludedFunctionNArguments a;
                a.valid = &valid; a.geometryUserPtr = g->user; a.primID = p.prim; a.context = ctx;
                a.ray = (RTCRayN *)ray; a.N = 1; a.geomID = p.geom;
                g->occl_fn(&a);
                if (ray->tfar < 0) return;
            }
        }
    }
}

// out[0] = TSC ticks spent inside rtcIntersect1 + rtcOccluded1 summed over all threads, out[1] = calls; reset != 0 clears
void gdpt_rtc_standin_stats(unsigned long long *out, int reset) {
    std::lock_guard<std::mutex> lk(g_counter_mutex);
    unsigned long long t = 0, n = 0;
    for (RtcCounter *c : g_counters) { t += c->ticks; n += c->calls; if (reset) { c->ticks = 0; c->calls = 0; } }
    if (out) { out[0] = t; out[1] = n; }
}
unsigned long long gdpt_rtc_standin_tsc(void) { return __rdtsc(); }

} // extern "C"
