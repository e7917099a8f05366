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
// Acceleration: a binned-SAH binary BVH so the timed CPU baseline is not
// handicapped by the stand-in.
#include <embree4/rtcore.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

namespace {

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

struct SceneImpl {
    int refcount = 1;
    std::vector<Geom *> geoms;
    std::vector<Prim> prims;
    std::vector<Node> nodes;
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

inline bool slab(const Node &n, const RayD &r, double tnear, double tfar, double &tentry) {
    double t0 = tnear, t1 = tfar;
    for (int k = 0; k < 3; k++) {
        double a = (n.lo[k] - r.o[k]) * r.inv[k];
        double b = (n.hi[k] - r.o[k]) * r.inv[k];
        if (a > b) std::swap(a, b);
        // NaN (0 * inf) must not cull: comparisons with NaN are false.
        if (a > t0) t0 = a;
        if (b < t1) t1 = b;
    }
    tentry = t0;
    return t0 <= t1 * (1 + 1e-12) + 1e-300;
}

// Moeller-Trumbore in double. Returns true and (t,u,v,Ng) when the line hits.
inline bool tri_hit(const Geom *g, unsigned prim, const RayD &r, double &t, double &u, double &v, double Ng[3]) {
    const unsigned *ix = tri(g, prim);
    const float *a = vtx(g, ix[0]), *b = vtx(g, ix[1]), *c = vtx(g, ix[2]);
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
    if (!s->prims.empty()) { s->nodes.reserve(2 * s->prims.size()); build_node(s, 0, (int)s->prims.size()); }
}

void rtcGetSceneBounds(RTCScene hs, struct RTCBounds *b) {
    SceneImpl *s = (SceneImpl *)hs;
    b->lower_x = s->blo[0]; b->lower_y = s->blo[1]; b->lower_z = s->blo[2]; b->align0 = 0;
    b->upper_x = s->bhi[0]; b->upper_y = s->bhi[1]; b->upper_z = s->bhi[2]; b->align1 = 0;
}

void rtcIntersect1(RTCScene hs, struct RTCRayHit *rh, struct RTCIntersectArguments *args) {
    SceneImpl *s = (SceneImpl *)hs;
    if (s->nodes.empty()) return;
    RTCRayQueryContext local_ctx; rtcInitRayQueryContext(&local_ctx);
    RTCRayQueryContext *ctx = (args && args->context) ? args->context : &local_ctx;
    RayD r = make_ray(rh->ray);
    const double tnear = rh->ray.tnear;
    double best = rh->ray.tfar;
    unsigned bg = RTC_INVALID_GEOMETRY_ID, bp = RTC_INVALID_GEOMETRY_ID;
    int stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const Node &n = s->nodes[stack[--sp]];
        double te;
        if (!slab(n, r, tnear, best, te)) continue;
        if (n.count == 0) {
            double tl, tr;
            bool hl = slab(s->nodes[n.left], r, tnear, best, tl);
            bool hr = slab(s->nodes[n.right], r, tnear, best, tr);
            if (hl && hr) {
                if (tl <= tr) { stack[sp++] = n.right; stack[sp++] = n.left; }
                else { stack[sp++] = n.left; stack[sp++] = n.right; }
            } else if (hl) stack[sp++] = n.left;
            else if (hr) stack[sp++] = n.right;
            continue;
        }
        for (int i = n.first; i < n.first + n.count; i++) {
            const Prim &p = s->prims[i];
            Geom *g = s->geoms[p.geom];
            if (g->type == RTC_GEOMETRY_TYPE_TRIANGLE) {
                double t, u, v, Ng[3];
                if (!tri_hit(g, p.prim, r, t, u, v, Ng)) continue;
                if (!(t > tnear && t <= best)) continue;
                if (t == best && bg != RTC_INVALID_GEOMETRY_ID &&
                    !(p.geom < bg || (p.geom == bg && p.prim < bp))) continue;
                best = t; bg = p.geom; bp = p.prim;
                rh->ray.tfar = (float)t;
                rh->hit.Ng_x = (float)Ng[0]; rh->hit.Ng_y = (float)Ng[1]; rh->hit.Ng_z = (float)Ng[2];
                rh->hit.u = (float)u; rh->hit.v = (float)v;
                rh->hit.primID = p.prim; rh->hit.geomID = p.geom;
                rh->hit.instID[0] = RTC_INVALID_GEOMETRY_ID;
            } else if (g->isect_fn) {
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
}

void rtcOccluded1(RTCScene hs, struct RTCRay *ray, struct RTCOccludedArguments *args) {
    SceneImpl *s = (SceneImpl *)hs;
    if (s->nodes.empty()) return;
    RTCRayQueryContext local_ctx; rtcInitRayQueryContext(&local_ctx);
    RTCRayQueryContext *ctx = (args && args->context) ? args->context : &local_ctx;
    RayD r = make_ray(*ray);
    const double tnear = ray->tnear, tfar = ray->tfar;
    int stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const Node &n = s->nodes[stack[--sp]];
        double te;
        if (!slab(n, r, tnear, tfar, te)) continue;
        if (n.count == 0) { stack[sp++] = n.left; stack[sp++] = n.right; continue; }
        for (int i = n.first; i < n.first + n.count; i++) {
            const Prim &p = s->prims[i];
            Geom *g = s->geoms[p.geom];
            if (g->type == RTC_GEOMETRY_TYPE_TRIANGLE) {
                double t, u, v, Ng[3];
                if (tri_hit(g, p.prim, r, t, u, v, Ng) && t > tnear && t <= tfar) {
                    ray->tfar = -std::numeric_limits<float>::infinity(); return;
                }
            } else if (g->occl_fn) {
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

} // extern "C"
