// The reference's stand-alone gradient-domain path tracer (/root/reference/small_gdpt.cpp) restated
// for the device: its own scene of eight spheres (two of them a perfect mirror and a perfect dielectric, which lajolla's
// materials cannot express), erand48 streams, base path by BSDF sampling, four offset paths by random-number replay with a
// reconnection at the first diffuse-diffuse edge, balance-heuristic weights, the five accumulation images.  It is the
// independent pin of the shift-mapping algorithm: the compiled reference program (oracle/smallgdpt/) is compared with this
// restatement image for image (tests/test_smallgdpt_*.py), in double, with the reference's order of operations kept
// expression by expression so that the host compilation of this header (tests/host/smallgdpt_check.cpp) is BIT-IDENTICAL
// to the reference's outputs and the device differs only by its libm's last-bit differences in sin / cos.
//
// Two behaviours of the reference program that the restatement keeps on purpose (found by comparing outputs):
//   * the offset path's first-vertex incoming direction in pathContrib / pathProb comes from offsetPath.rnds[0..1], which
//     shiftPath never writes (small_gdpt.cpp:200-285): the program reads whatever its stack holds there.  The compiled
//     reference behaves as if both were 0 (stale stack words read as doubles are denormal or zero and vanish when added to
//     the pixel index), i.e. the direction through the pixel's corner; it only matters when the first vertex is glass.
//   * `Ray ray(cam.o + d*140, d.norm())` (:160, :213): norm() normalises d in place and the order in which the two
//     arguments are evaluated is the compiler's; SG_ORIGIN_AFTER_NORM states what gcc does here.
// Every function below cites the lines it follows.  Compiles for the host when GDPT_SMALLGDPT_HOST is defined.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__) && !defined(GDPT_SMALLGDPT_HOST)
#define SG_HD __host__ __device__ inline
#else
#define SG_HD inline
#endif

namespace gdpt {
namespace sg {

constexpr int SG_MAX_DEPTH = 32;            // small_gdpt.cpp:14
constexpr int SG_SPHERES = 8;               // :60-69
constexpr double SG_PI = 3.14159265358979323846;
#ifndef SG_ORIGIN_AFTER_NORM
#define SG_ORIGIN_AFTER_NORM 1
#endif

struct V { double x, y, z; };
SG_HD V mk(double x, double y, double z) { V r; r.x = x; r.y = y; r.z = z; return r; }
SG_HD V operator+(const V &a, const V &b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
SG_HD V operator-(const V &a, const V &b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
SG_HD V operator-(const V &a) { return mk(-a.x, -a.y, -a.z); }
SG_HD V operator*(const V &a, double s) { return mk(a.x * s, a.y * s, a.z * s); }
SG_HD V mul(const V &a, const V &b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
SG_HD double dot(const V &a, const V &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
SG_HD V cross(const V &a, const V &b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
SG_HD V unit(const V &a) { return a * (1 / sqrt(a.x * a.x + a.y * a.y + a.z * a.z)); }      // Vec::norm, :25
SG_HD double vmax(const V &a) { return a.x > a.y && a.x > a.z ? a.x : a.y > a.z ? a.y : a.z; } // Vec::max, :28

enum { SG_DIFF = 0, SG_SPEC = 1, SG_REFR = 2 };
struct Sphere { double rad; V p, e, c; int refl; };
struct Scene {
    Sphere s[SG_SPHERES];
    V cam_o, cam_d, cx, cy;
    int width, height;
};

// the scene and camera of small_gdpt.cpp:55-69 at a film of w x h
inline Scene make_scene(int w, int h) {
    Scene sc;
    auto set = [&](int i, double rad, V p, V e, V c, int refl) { sc.s[i].rad = rad; sc.s[i].p = p; sc.s[i].e = e; sc.s[i].c = c; sc.s[i].refl = refl; };
    const V zero = mk(0, 0, 0);
    set(0, 1e5, mk(1e5 + 1, 40.8, 81.6), zero, mk(.75, .25, .25), SG_DIFF);
    set(1, 1e5, mk(-1e5 + 99, 40.8, 81.6), zero, mk(.25, .25, .75), SG_DIFF);
    set(2, 1e5, mk(50, 40.8, 1e5), zero, mk(.75, .75, .75), SG_DIFF);
    set(3, 1e5, mk(50, 1e5, 81.6), zero, mk(.75, .75, .75), SG_DIFF);
    set(4, 1e5, mk(50, -1e5 + 81.6, 81.6), zero, mk(.75, .75, .75), SG_DIFF);
    set(5, 16.5, mk(27, 16.5, 47), zero, mk(1.0, 1.0, 1.0) * .999, SG_SPEC);
    set(6, 16.5, mk(73, 16.5, 78), zero, mk(1.0, 1.0, 1.0) * .999, SG_REFR);
    set(7, 600, mk(50, 681.6 - .27, 81.6), mk(12, 12, 12), zero, SG_DIFF);
    sc.width = w; sc.height = h;
    sc.cam_o = mk(50, 50, 295.6);
    sc.cam_d = unit(mk(0, -0.042612, -1));
    sc.cx = mk(w * .5135 / h, 0, 0);
    sc.cy = unit(cross(sc.cx, sc.cam_d)) * .5135;
    return sc;
}

struct Ray { V o, d; };
struct PathVert { V p, n; int id; };                                  // :73-75
struct Path {                                                          // :78-86
    PathVert verts[SG_MAX_DEPTH];
    double rnds[2 * SG_MAX_DEPTH];
    int vertCount, x, y;
};

// erand48: X <- (0x5DEECE66D X + 0xB) mod 2^48, value X / 2^48 (POSIX; glibc builds exactly this double)
SG_HD double erand(uint64_t &X) {
    X = (X * 0x5DEECE66DULL + 0xBULL) & 0xFFFFFFFFFFFFULL;
    return (double)X * (1.0 / 281474976710656.0);
}
// the row's stream: unsigned short rng[3] = {0, 0, y*y*y} (:462)
SG_HD uint64_t row_seed(int y) { return (uint64_t)(uint16_t)(y * y * y) << 32; }

SG_HD double hit_sphere(const Sphere &s, const Ray &r) {              // Sphere::intersect, :44-49
    const V op = s.p - r.o;
    double t; const double eps = 1e-4, b = dot(op, r.d);
    double det = b * b - dot(op, op) + s.rad * s.rad;
    if (det < 0) return 0; else det = sqrt(det);
    return (t = b - det) > eps ? t : ((t = b + det) > eps ? t : 0);
}
SG_HD bool intersect(const Scene &sc, const Ray &r, double &t, int &id) {   // :89-93
    const double inf = t = 1e20;
    for (int i = SG_SPHERES; i--;) { const double d = hit_sphere(sc.s[i], r); if (d != 0 && d < t) { t = d; id = i; } }
    return t < inf;
}
SG_HD V reflect(const V &d, const V &n) { return d - n * 2.0 * dot(n, d); }  // :95-97

SG_HD V film_dir(const Scene &sc, double r0, double r1, int x, int y) {      // :158-159
    return sc.cx * ((r0 + x) / sc.width - .5) + sc.cy * ((r1 + y) / sc.height - .5) + sc.cam_d;
}
SG_HD Ray camera_ray(const Scene &sc, const V &d) {                          // :160, :213
    Ray r; r.d = unit(d);
    r.o = sc.cam_o + (SG_ORIGIN_AFTER_NORM ? r.d : d) * 140;
    return r;
}

// Refraction at the glass sphere for direction d arriving at normal n (:109-117, :136-146, :313-320): returns cos2t; the
// refracted direction and Schlick's reflectance are formed when it is not negative (the sampler's branch is `< 0`, the
// two evaluators' is `> 0`, :111 / :138 / :314)
SG_HD double refract(const V &d, const V &n, bool into, V &tdir, double &Re) {
    const V nl = into ? n : n * -1;
    const double nc = 1, nt = 1.5, nnt = into ? nc / nt : nt / nc, ddn = dot(d, nl);
    const double cos2t = 1 - nnt * nnt * (1 - ddn * ddn);
    if (cos2t < 0) return cos2t;
    tdir = unit(d * nnt - n * ((into ? 1 : -1) * (ddn * nnt + sqrt(cos2t))));
    const double a = nt - nc, b = nt + nc, R0 = a * a / (b * b), c = 1 - (into ? -ddn : dot(tdir, n));
    Re = R0 + (1 - R0) * c * c * c * c * c;
    return cos2t;
}

SG_HD Ray sample_bsdf(const Ray &ray, const Sphere &obj, const PathVert &vert, double u0, double u1) {   // :99-126
    Ray out; out.o = vert.p;
    if (obj.refl == SG_DIFF) {
        const double r1 = 2 * SG_PI * u0, r2 = u1, r2s = sqrt(r2);
        const V nl = dot(vert.n, ray.d) < 0 ? vert.n : vert.n * -1;
        const V w = nl, u = unit(cross(fabs(w.x) > .1 ? mk(0, 1, 0) : mk(1, 0, 0), w)), v = cross(w, u);
        out.d = unit(u * cos(r1) * r2s + v * sin(r1) * r2s + w * sqrt(1 - r2));
    } else if (obj.refl == SG_SPEC) {
        out.d = reflect(ray.d, vert.n);
    } else {
        const V rd = reflect(ray.d, vert.n);
        const bool into = dot(vert.n, ray.d) < 0;
        V tdir; double Re;
        if (refract(ray.d, vert.n, into, tdir, Re) < 0) { out.d = rd; return out; }   // total internal reflection
        const double P = .25 + .5 * Re;
        out.d = u0 < P ? rd : tdir;
    }
    return out;
}

SG_HD double bsdf_prob(int refl_t, const V &wi, const V &n, const V &wo) {           // :128-150
    if (refl_t == SG_DIFF) return fabs(dot(wo, n)) / SG_PI;
    if (refl_t == SG_SPEC) return 1.0;
    const bool refl = dot(wi, n) * dot(wo, n) > 0.0, into = dot(n, wi) > 0;
    double P = refl ? 1.0 : 0.0, Re; V tdir;
    if (refract(-wi, n, into, tdir, Re) > 0) { P = .25 + .5 * Re; if (!refl) P = 1.0 - P; }
    return P;
}

SG_HD bool generate_path(const Scene &sc, int x, int y, uint64_t &rng, Path &path) {   // :153-197
    path.x = x; path.y = y;
    path.rnds[0] = erand(rng); path.rnds[1] = erand(rng);
    Ray ray = camera_ray(sc, film_dir(sc, path.rnds[0], path.rnds[1], x, y));
    path.vertCount = 0;
    for (int depth = 1; depth <= SG_MAX_DEPTH; depth++) {
        double t; int id = -1;
        if (!intersect(sc, ray, t, id)) return false;
        const Sphere &obj = sc.s[id];
        PathVert vert;
        vert.p = ray.o + ray.d * t; vert.n = unit(vert.p - obj.p); vert.id = id;
        path.verts[depth - 1] = vert;
        path.vertCount++;
        if (vmax(obj.c) <= 0.0) return true;                  // the light
        if (depth == SG_MAX_DEPTH) return false;
        path.rnds[2 * depth] = erand(rng);
        path.rnds[2 * depth + 1] = erand(rng);
        ray = sample_bsdf(ray, obj, vert, path.rnds[2 * depth], path.rnds[2 * depth + 1]);
    }
    return false;
}

SG_HD double geom_term(const V &from, const V &n_from, const V &to, const V &n_to) {    // :258-268
    V dir = to - from;
    const double dist2 = dot(dir, dir);
    dir = dir * (1.0 / sqrt(dist2));
    return fabs(dot(dir, n_to)) * fabs(dot(dir, n_from)) / dist2;
}

SG_HD bool shift_path(const Scene &sc, int x, int y, const Path &base, Path &off, double &jacobian) {   // :200-285
    off.x = x; off.y = y;
    V baseWi = -unit(film_dir(sc, base.rnds[0], base.rnds[1], base.x, base.y));
    Ray ray = camera_ray(sc, film_dir(sc, base.rnds[0], base.rnds[1], x, y));
    V wi = -ray.d;
    off.vertCount = base.vertCount;
    for (int k = 0; k < base.vertCount; k++) off.verts[k] = base.verts[k];
    jacobian = 1.0;
    for (int vertId = 0; vertId < base.vertCount; vertId++) {
        const int depth = vertId + 1;
        double t; int id = -1;
        if (!intersect(sc, ray, t, id)) return false;
        const Sphere &obj = sc.s[id];
        const Sphere &baseObj = sc.s[base.verts[vertId].id];
        if (obj.refl != baseObj.refl) return false;
        PathVert vert;
        vert.p = ray.o + ray.d * t; vert.n = unit(vert.p - obj.p); vert.id = id;
        off.verts[vertId] = vert;
        if (vertId == base.vertCount - 1) break;
        if (obj.refl == SG_DIFF && sc.s[base.verts[vertId + 1].id].refl == SG_DIFF) {
            Ray con; con.o = vert.p; con.d = unit(base.verts[depth].p - vert.p);
            if (!intersect(sc, con, t, id) || id != base.verts[vertId + 1].id) return false;
            const double baseGeom = geom_term(base.verts[depth - 1].p, base.verts[depth - 1].n, base.verts[depth].p, base.verts[depth].n);
            const double shiftGeom = geom_term(vert.p, vert.n, base.verts[depth].p, base.verts[depth].n);
            jacobian *= (shiftGeom / baseGeom);
            return true;
        }
        ray = sample_bsdf(ray, obj, vert, base.rnds[2 * depth], base.rnds[2 * depth + 1]);
        const V baseWo = unit(base.verts[vertId + 1].p - base.verts[vertId].p);
        const double basePDF = bsdf_prob(baseObj.refl, baseWi, base.verts[vertId].n, baseWo);
        const double shiftPDF = bsdf_prob(obj.refl, wi, vert.n, ray.d);
        if (shiftPDF <= 0.0) return false;
        jacobian *= (basePDF / shiftPDF);
        baseWi = -baseWo; wi = -ray.d;
    }
    return vmax(sc.s[off.verts[off.vertCount - 1].id].c) <= 0.0;
}

// pathContrib (:288-333) and pathProb (:336-352) in one walk; r0, r1 are the path's first two random numbers as the
// reference reads them (the base path's own; 0, 0 for an offset path — see the header)
SG_HD void path_contrib_prob(const Scene &sc, const Path &path, double r0, double r1, V &contrib, double &prob) {
    V throughput = mk(1, 1, 1);
    V wi = -unit(film_dir(sc, r0, r1, path.x, path.y));
    prob = 1.0;
    bool contrib_dead = false, prob_dead = false;
    for (int v = 0; v < path.vertCount - 1; v++) {
        const PathVert &cur = path.verts[v], &nxt = path.verts[v + 1];
        const V wo = unit(nxt.p - cur.p);
        const double cosTheta = fabs(dot(wo, cur.n));
        const Sphere &obj = sc.s[cur.id];
        if (cosTheta <= 1e-6) { contrib_dead = true; prob_dead = true; break; }
        if (obj.refl == SG_DIFF) throughput = mul(throughput, obj.c * (cosTheta / SG_PI));
        else if (obj.refl == SG_SPEC) throughput = mul(throughput, obj.c);
        else {
            const bool refl = dot(wi, cur.n) * dot(wo, cur.n) > 0.0, into = dot(cur.n, wi) > 0;
            double f = refl ? 1.0 : 0.0, Re; V tdir;
            if (refract(-wi, cur.n, into, tdir, Re) > 0) { const double Tr = 1 - Re; f = refl ? Re : Tr; }
            throughput = mul(throughput, obj.c * f);
        }
        if (!prob_dead) {
            prob *= bsdf_prob(obj.refl, wi, cur.n, wo);
            if (prob <= 0.0) prob_dead = true;                 // pathProb returns 0 here; pathContrib walks on
        }
        wi = -wo;
    }
    contrib = contrib_dead ? mk(0, 0, 0) : mul(throughput, sc.s[path.verts[path.vertCount - 1].id].e);
    if (prob_dead) prob = 0.0;
}

struct PixelSums { V r, dx0, dy0, dx1, dy1; };

// one pixel-sample of main()'s loop (:470-520): base path, four shifts, the five running sums
SG_HD void sample_pixel(const Scene &sc, int x, int y, int samps, uint64_t &rng, Path &path, Path &oPath, PixelSums &a) {
    if (!generate_path(sc, x, y, rng, path)) return;
    V contrib; double prob;
    path_contrib_prob(sc, path, path.rnds[0], path.rnds[1], contrib, prob);
    if (!(prob > 0.0)) return;
    V cS[4]; double wS[4];
    const int sx[4] = {x - 1, x, x + 1, x}, sy[4] = {y, y + 1, y, y - 1};      // X0, Y0, X1, Y1 (:490-510)
    a.r = a.r + (contrib * (1.0 / prob)) * (1.0 / (double)samps);
    for (int k = 0; k < 4; k++) {
        cS[k] = mk(0, 0, 0); wS[k] = 1;
        double jacobian;
        if (shift_path(sc, sx[k], sy[k], path, oPath, jacobian)) {
            V c; double p;
            path_contrib_prob(sc, oPath, 0.0, 0.0, c, p);
            cS[k] = c * jacobian;
            const double pk = p * jacobian;
            wS[k] = prob / (prob + pk);
        }
    }
    a.dx0 = a.dx0 + (contrib - cS[0]) * (wS[0] / (prob * (double)samps));
    a.dy0 = a.dy0 + (contrib - cS[1]) * (wS[1] / (prob * (double)samps));
    a.dx1 = a.dx1 + (cS[2] - contrib) * (wS[2] / (prob * (double)samps));
    a.dy1 = a.dy1 + (cS[3] - contrib) * (wS[3] / (prob * (double)samps));
}

// main()'s pixel (:463-527): samps samples, then the five images at the flipped row
SG_HD void render_pixel(const Scene &sc, int x, int y, int samps, uint64_t &rng, Path &path, Path &oPath,
                        double *c, double *cx0, double *cy0, double *cx1, double *cy1) {
    PixelSums a; a.r = a.dx0 = a.dy0 = a.dx1 = a.dy1 = mk(0, 0, 0);
    for (int s = 0; s < samps; s++) sample_pixel(sc, x, y, samps, rng, path, oPath, a);
    const size_t i = ((size_t)(sc.height - y - 1) * sc.width + x) * 3;
    c[i] = a.r.x; c[i + 1] = a.r.y; c[i + 2] = a.r.z;
    cx0[i] = a.dx0.x; cx0[i + 1] = a.dx0.y; cx0[i + 2] = a.dx0.z;
    cy0[i] = a.dy0.x; cy0[i + 1] = a.dy0.y; cy0[i + 2] = a.dy0.z;
    cx1[i] = a.dx1.x; cx1[i + 1] = a.dx1.y; cx1[i + 2] = a.dx1.z;
    cy1[i] = a.dy1.x; cy1[i + 1] = a.dy1.y; cy1[i + 2] = a.dy1.z;
}

// a stream of its own per pixel, for the mode that gives every pixel a thread (no counterpart in the reference, whose
// stream runs along the row): splitmix64 of the pixel index, 48 bits
SG_HD uint64_t pixel_seed(int x, int y, int w) {
    uint64_t z = ((uint64_t)y * (uint64_t)w + (uint64_t)x + 1) * 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; z ^= z >> 31;
    return z & 0xFFFFFFFFFFFFULL;
}

}  // namespace sg
}  // namespace gdpt
