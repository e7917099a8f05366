// TEST INFRASTRUCTURE — not part of the product.
//
// C entry points around the REAL reference (compiled from /root/reference/src
// into oracle/_ref/liblajolla_ref.so by oracle/build_ref.sh) so that tests, the
// golden-vector generator and bench.py's CPU arm can drive it through ctypes.
// Nothing here re-implements the reference's path; it only calls the reference's own
// functions (the one exception, gdpt_ref_shift_sample at the end, is the checker of the product's
// opt-in shift-mapped mode, which the reference does not have: a statement of that algorithm
// built from the reference's own routines, and marked as such):
//   parse_scene      src/parsers/parse_scene.cpp:1615-1630
//   render           src/render.cpp:374-394  (GradPath arm -> gradient_path_render :257-370)
//   fourierSolve     src/render.cpp:172-254
//   grad_path_tracing src/path_tracing.h:354-1050
//   path_tracing     src/path_tracing.h:13-348; sample_light / sample_point_on_light / occluded (scene.cpp:61-67, light.cpp:57-71, intersection.cpp:66-85)
//   sample_primary   src/camera.cpp:23-47
//   intersect        src/intersection.cpp:7-65
//   eval / pdf_sample_bsdf / sample_bsdf   src/material.cpp:90-119
// plus an exporter that flattens the reference's parsed `Scene` into the
// "GDPT1" array container the product's host side consumes (so the product is
// fed by the reference's own XML/OBJ/.serialized/JPG loaders, byte for byte).
#include "scene.h"
#include "render.h"
#include "parallel.h"
#include "intersection.h"
#include "material.h"
#include "camera.h"
#include "image.h"
#include "parsers/parse_scene.h"

#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

// same layout as src/pcg.h:16-19 (pcg.h defines non-inline specialisations, so it
// cannot be included by a second translation unit)
struct pcg32_state { uint64_t state; uint64_t inc; };
GraidentPTRadiance grad_path_tracing(const Scene &scene, int x, int y, pcg32_state &rng);
Spectrum path_tracing(const Scene &scene, int x, int y, pcg32_state &rng);   // src/path_tracing.h:13 (non-inline, defined in the render.cpp TU)
void fourierSolve(int width, int height, const double *imgData, const double *imgGradX,
                  const double *imgGradY, double dataCost, double *imgOut);

extern "C" {
typedef struct { uint64_t state, inc; } gdpt_pcg32;
void gdpt_oracle_pcg_init(gdpt_pcg32 *r, uint64_t stream, uint64_t seed);
double gdpt_oracle_pcg_real(gdpt_pcg32 *r);   // next_pcg32_real<double> (src/pcg.h:52-68), bit-exact restatement (oracle/pcg_oracle.c)
}

namespace {
RTCDevice g_device = nullptr;
bool g_parallel = false;
// capture buffers filled by the dump hook that build_ref.sh splices after the
// fourierSolve call at src/render.cpp:353
thread_local double *g_c = nullptr, *g_cx = nullptr, *g_cy = nullptr, *g_out = nullptr;

struct BlobWriter {
    FILE *f;
    void arr(const std::string &name, const char *dtype, const std::vector<int64_t> &shape, const void *data, size_t bytes) {
        uint32_t nl = (uint32_t)name.size();
        fwrite(&nl, 4, 1, f); fwrite(name.data(), 1, nl, f);
        char dt[8] = {0}; strncpy(dt, dtype, 7); fwrite(dt, 1, 8, f);
        uint32_t nd = (uint32_t)shape.size(); fwrite(&nd, 4, 1, f);
        for (int64_t s : shape) fwrite(&s, 8, 1, f);
        uint64_t nb = bytes; fwrite(&nb, 8, 1, f);
        if (bytes) fwrite(data, 1, bytes, f);
    }
};

template <typename T> void tex_entry(const Texture<T> &t, std::vector<double> &out);

inline void put3(std::vector<double> &o, const Vector3 &v) { o.push_back(v.x); o.push_back(v.y); o.push_back(v.z); }
inline void put3(std::vector<double> &o, const Real &v) { o.push_back(v); o.push_back(v); o.push_back(v); }

// 14 doubles: kind, is_spectrum, v0[3], v1[3], image_id, uscale, vscale, uoffset, voffset, pad
template <typename T> void tex_entry(const Texture<T> &t, std::vector<double> &o) {
    const double is_spec = std::is_same<T, Vector3>::value ? 1.0 : 0.0;
    if (auto *c = std::get_if<ConstantTexture<T>>(&t)) {
        o.push_back(0); o.push_back(is_spec); put3(o, c->value); put3(o, T{}); o.push_back(-1);
        o.push_back(1); o.push_back(1); o.push_back(0); o.push_back(0); o.push_back(0);
    } else if (auto *im = std::get_if<ImageTexture<T>>(&t)) {
        o.push_back(1); o.push_back(is_spec); put3(o, T{}); put3(o, T{}); o.push_back(im->texture_id);
        o.push_back(im->uscale); o.push_back(im->vscale); o.push_back(im->uoffset); o.push_back(im->voffset); o.push_back(0);
    } else if (auto *cb = std::get_if<CheckerboardTexture<T>>(&t)) {
        o.push_back(2); o.push_back(is_spec); put3(o, cb->color0); put3(o, cb->color1); o.push_back(-1);
        o.push_back(cb->uscale); o.push_back(cb->vscale); o.push_back(cb->uoffset); o.push_back(cb->voffset); o.push_back(0);
    }
}

struct MatFlat {
    std::vector<double> tex;      // 14 per texture
    std::vector<double> mats;     // per material: type, eta, 12 texture ids (-1 unused)
    int ntex = 0;
    template <typename T> int add(const Texture<T> &t) { tex_entry(t, tex); return ntex++; }
    void mat(int type, double eta, std::vector<int> ids) {
        mats.push_back(type); mats.push_back(eta);
        for (int i = 0; i < 12; i++) mats.push_back(i < (int)ids.size() ? ids[i] : -1);
    }
};

template <typename T> void export_images(BlobWriter &w, const char *prefix, const std::vector<Mipmap<T>> &pool) {
    constexpr int C = std::is_same<T, Vector3>::value ? 3 : 1;
    for (size_t i = 0; i < pool.size(); i++) {
        const Image<T> &img = pool[i].images[0];
        std::string base = std::string(prefix) + "." + std::to_string(i);
        int32_t sz[2] = {img.width, img.height};
        w.arr(base + ".size", "i4", {2}, sz, sizeof(sz));
        const double *d = (const double *)img.data.data();
        size_t n = (size_t)img.width * img.height * C;
        std::map<double, int> pal;
        bool ok = true;
        for (size_t k = 0; k < n && ok; k++) { pal.emplace(d[k], 0); if (pal.size() > 256) ok = false; }
        if (ok) {
            std::vector<double> values; int idx = 0;
            for (auto &kv : pal) { kv.second = idx++; values.push_back(kv.first); }
            std::vector<uint8_t> index(n);
            for (size_t k = 0; k < n; k++) index[k] = (uint8_t)pal[d[k]];
            w.arr(base + ".palette", "f8", {(int64_t)values.size()}, values.data(), values.size() * 8);
            w.arr(base + ".index", "u1", {img.height, img.width, C}, index.data(), n);
        } else {
            w.arr(base + ".data", "f8", {img.height, img.width, C}, d, n * 8);
        }
    }
}
} // namespace

extern "C" {

// called from the spliced line in render.cpp
void gdpt_ref_dump_hook(int w, int h, const double *c, const double *cx, const double *cy, const double *out) {
    size_t n = (size_t)w * h * 3 * sizeof(double);
    if (g_c) memcpy(g_c, c, n);
    if (g_cx) memcpy(g_cx, cx, n);
    if (g_cy) memcpy(g_cy, cy, n);
    if (g_out) memcpy(g_out, out, n);
}

int gdpt_ref_init(int num_threads) {
    if (!g_device) g_device = rtcNewDevice(nullptr);
    if (g_parallel) { parallel_cleanup(); g_parallel = false; }
    parallel_init(num_threads);
    g_parallel = true;
    return 0;
}

void gdpt_ref_shutdown() {
    if (g_parallel) { parallel_cleanup(); g_parallel = false; }
}

void *gdpt_ref_load(const char *xml) {
    if (!g_device) g_device = rtcNewDevice(nullptr);
    std::unique_ptr<Scene> s = parse_scene(fs::path(xml), g_device);
    return s.release();
}
void gdpt_ref_free(void *s) { delete (Scene *)s; }

void gdpt_ref_info(void *hs, int *out /* w,h,spp,max_depth,rr_depth,integrator,nshapes,nmats */) {
    Scene *s = (Scene *)hs;
    out[0] = s->camera.width; out[1] = s->camera.height;
    out[2] = s->options.samples_per_pixel; out[3] = s->options.max_depth; out[4] = s->options.rr_depth;
    out[5] = (int)s->options.integrator; out[6] = (int)s->shapes.size(); out[7] = (int)s->materials.size();
}

// values < -1 keep the current setting
void gdpt_ref_set_options(void *hs, int spp, int max_depth, int rr_depth) {
    Scene *s = (Scene *)hs;
    if (spp >= 0) s->options.samples_per_pixel = spp;
    if (max_depth >= -1) s->options.max_depth = max_depth;
    if (rr_depth >= -1) s->options.rr_depth = rr_depth;
}

// The reference's own render() with the GradPath integrator (tile-seeded RNG,
// src/render.cpp:281).  c/cx/cy/out (each W*H*3 doubles, may be NULL) receive the
// arrays of src/render.cpp:336-353.
int gdpt_ref_render(void *hs, double *c, double *cx, double *cy, double *out) {
    Scene *s = (Scene *)hs;
    Integrator saved = s->options.integrator;
    s->options.integrator = Integrator::GradPath;
    g_c = c; g_cx = cx; g_cy = cy; g_out = out;
    Image3 img = render(*s);
    g_c = g_cx = g_cy = g_out = nullptr;
    s->options.integrator = saved;
    return 0;
}

// The reference's plain `path` integrator (src/render.cpp:74-117), for the
// equal-time comparator.  out: W*H*3.
int gdpt_ref_render_path(void *hs, double *out) {
    Scene *s = (Scene *)hs;
    Integrator saved = s->options.integrator;
    s->options.integrator = Integrator::Path;
    Image3 img = render(*s);
    memcpy(out, img.data.data(), sizeof(double) * 3 * img.data.size());
    s->options.integrator = saved;
    return 0;
}

// The reference's auxiliary integrators (aux_render, src/render.cpp:15-72) through render(): `integrator` is the enum
// value 0..4 (depth, shadingNormal, meanCurvature, rayDifferential, mipmapLevel).  out: W*H*3.
int gdpt_ref_render_aux(void *hs, int integrator, double *out) {
    Scene *s = (Scene *)hs;
    Integrator saved = s->options.integrator;
    s->options.integrator = (Integrator)integrator;
    Image3 img = render(*s);
    memcpy(out, img.data.data(), sizeof(double) * 3 * img.data.size());
    s->options.integrator = saved;
    return 0;
}

void gdpt_ref_fourier_solve(int w, int h, const double *u, const double *gx, const double *gy, double alpha, double *out) {
    fourierSolve(w, h, u, gx, gy, alpha, out);
}

// One grad_path_tracing evaluation with rng = init_pcg32(stream, seed).
// rec[23]: radiance3, contrib3, contribX0 3, contribX1 3, contribY0 3, contribY1 3, prob, wX0, wY0, wX1, wY1
void gdpt_ref_grad_sample(void *hs, int x, int y, uint64_t stream, uint64_t seed, double *rec) {
    Scene *s = (Scene *)hs;
    gdpt_pcg32 r0; gdpt_oracle_pcg_init(&r0, stream, seed);
    pcg32_state rng{r0.state, r0.inc};
    GraidentPTRadiance g = grad_path_tracing(*s, x, y, rng);
    const Vector3 *v[6] = {&g.radiance, &g.contrib, &g.contribX0, &g.contribX1, &g.contribY0, &g.contribY1};
    for (int i = 0; i < 6; i++) { rec[3 * i] = v[i]->x; rec[3 * i + 1] = v[i]->y; rec[3 * i + 2] = v[i]->z; }
    rec[18] = g.prob; rec[19] = g.wX0; rec[20] = g.wY0; rec[21] = g.wX1; rec[22] = g.wY1;
}

// imread3 (src/image.cpp:80-133; tinyexr / PFM): lets the tests read the product's output files with the reference's
// own loader.  Two calls: data == NULL returns the size.
int gdpt_ref_imread3(const char *path, int *w, int *h, double *data) {
    Image3 img = imread3(fs::path(path));
    *w = img.width; *h = img.height;
    if (data) memcpy(data, img.data.data(), sizeof(double) * 3 * img.data.size());
    return 0;
}

// One path_tracing evaluation (src/path_tracing.h:13-348) with rng = init_pcg32(stream, seed). rad[3]
void gdpt_ref_path_sample(void *hs, int x, int y, uint64_t stream, uint64_t seed, double *rad) {
    Scene *s = (Scene *)hs;
    gdpt_pcg32 r0; gdpt_oracle_pcg_init(&r0, stream, seed);
    pcg32_state rng{r0.state, r0.inc};
    Spectrum v = path_tracing(*s, x, y, rng);
    rad[0] = v.x; rad[1] = v.y; rad[2] = v.z;
}

// sample_light + sample_point_on_light + pdf_point_on_light + light_pmf (src/scene.cpp:61-67, src/light.cpp:57-71)
// rnd[n][4] = light_w, uv0, uv1, shape_w; out[n][9] = light id, position3, normal3, pdf of the point, pmf of the light
void gdpt_ref_sample_light(void *hs, int n, const double *ref_points, const double *rnd, double *out) {
    Scene *s = (Scene *)hs;
    for (int i = 0; i < n; i++) {
        Vector3 ref{ref_points[3 * i], ref_points[3 * i + 1], ref_points[3 * i + 2]};
        const double *q = rnd + 4 * i;
        int id = sample_light(*s, q[0]);
        const Light &light = s->lights[id];
        PointAndNormal p = sample_point_on_light(light, ref, Vector2{q[1], q[2]}, q[3], *s);
        double *o = out + 9 * i;
        o[0] = id; o[1] = p.position.x; o[2] = p.position.y; o[3] = p.position.z; o[4] = p.normal.x; o[5] = p.normal.y; o[6] = p.normal.z;
        o[7] = pdf_point_on_light(light, p, ref, *s); o[8] = light_pmf(*s, id);
    }
}

// occluded (src/intersection.cpp:66-85). rays[n][8] = org, dir, tnear, tfar; out[n]
void gdpt_ref_occluded(void *hs, int n, const double *rays, int *out) {
    Scene *s = (Scene *)hs;
    for (int i = 0; i < n; i++) {
        const double *q = rays + 8 * i;
        Ray r{Vector3{q[0], q[1], q[2]}, Vector3{q[3], q[4], q[5]}, q[6], q[7]};
        out[i] = occluded(*s, r) ? 1 : 0;
    }
}

// Pixel-seeded path accumulation (stream = sample * W*H + y*W + x, like the GPU product): img = sum radiance / spp
int gdpt_ref_render_path_pixelseed(void *hs, uint64_t seed, int spp, int sample0, double *img) {
    Scene *s = (Scene *)hs;
    const int w = s->camera.width, h = s->camera.height;
    parallel_for([&](Vector2i rc) {
        int y = rc[0];
        for (int x = 0; x < w; x++) {
            Vector3 r{0, 0, 0};
            for (int k = 0; k < spp; k++) {
                uint64_t stream = (uint64_t)(sample0 + k) * (uint64_t)w * h + (uint64_t)y * w + x;
                gdpt_pcg32 r0; gdpt_oracle_pcg_init(&r0, stream, seed);
                pcg32_state rng{r0.state, r0.inc};
                r += path_tracing(*s, x, y, rng);
            }
            r = r / Real(spp);
            size_t i = ((size_t)y * w + x) * 3;
            img[i] = r.x; img[i + 1] = r.y; img[i + 2] = r.z;
        }
    }, Vector2i(h, 1));
    return 0;
}

// Pixel-seeded gradpath accumulation: every (pixel, sample) gets its own stream
//   stream = sample * W*H + y*W + x
// (the seeding the GPU product uses, since a serial per-tile stream cannot be
// reproduced in parallel).  The per-sample arithmetic is the reference's
// grad_path_tracing; the accumulation restates src/render.cpp:311-327 and the
// merge :340-350.  Outputs c,cx,cy: W*H*3 each.  Rows [y0,y1) only.
int gdpt_ref_render_pixelseed(void *hs, uint64_t seed, int spp, int sample0, int y0, int y1,
                              double *c, double *cx0, double *cy0, double *cx1, double *cy1) {
    Scene *s = (Scene *)hs;
    const int w = s->camera.width, h = s->camera.height;
    if (y0 < 0) y0 = 0;
    if (y1 > h) y1 = h;
    parallel_for([&](Vector2i rc) {
        int y = y0 + rc[0];
        for (int x = 0; x < w; x++) {
            Vector3 r{0, 0, 0}, dX0{0, 0, 0}, dX1{0, 0, 0}, dY0{0, 0, 0}, dY1{0, 0, 0};
            for (int k = 0; k < spp; k++) {
                uint64_t stream = (uint64_t)(sample0 + k) * (uint64_t)w * h + (uint64_t)y * w + x;
                gdpt_pcg32 r0; gdpt_oracle_pcg_init(&r0, stream, seed);
                pcg32_state rng{r0.state, r0.inc};
                GraidentPTRadiance g = grad_path_tracing(*s, x, y, rng);
                if (g.prob > 0.0) {
                    r = r + g.radiance / Real(spp);
                    dX0 = dX0 + (g.contrib - g.contribX0) * (g.wX0 / (g.prob * Real(spp)));
                    dY0 = dY0 + (g.contrib - g.contribY0) * (g.wY0 / (g.prob * Real(spp)));
                    dX1 = dX1 + (g.contribX1 - g.contrib) * (g.wX1 / (g.prob * Real(spp)));
                    dY1 = dY1 + (g.contribY1 - g.contrib) * (g.wY1 / (g.prob * Real(spp)));
                }
            }
            size_t i = ((size_t)y * w + x) * 3;
            const Vector3 *src[5] = {&r, &dX0, &dY0, &dX1, &dY1};
            double *dst[5] = {c, cx0, cy0, cx1, cy1};
            for (int b = 0; b < 5; b++) if (dst[b]) { dst[b][i] = src[b]->x; dst[b][i + 1] = src[b]->y; dst[b][i + 2] = src[b]->z; }
        }
    }, Vector2i(y1 - y0, 1));
    return 0;
}

// sample_primary on n screen positions. rays: n*6 (org, dir)
void gdpt_ref_sample_primary(void *hs, int n, const double *screen_pos, double *rays) {
    Scene *s = (Scene *)hs;
    for (int i = 0; i < n; i++) {
        Ray r = sample_primary(s->camera, Vector2{screen_pos[2 * i], screen_pos[2 * i + 1]});
        rays[6 * i] = r.org.x; rays[6 * i + 1] = r.org.y; rays[6 * i + 2] = r.org.z;
        rays[6 * i + 3] = r.dir.x; rays[6 * i + 4] = r.dir.y; rays[6 * i + 5] = r.dir.z;
    }
}

// intersect() on n rays. rays: n*8 (org3, dir3, tnear, tfar); rdiff: n*2 (radius, spread) or NULL.
// ids: n*3 (shape, prim, material) (-1 on miss); vtx: n*24 doubles:
//   position3, geometric_normal3, frame.x3, frame.y3, frame.n3, st2, uv2, uv_screen_size, mean_curvature, ray_radius, t(float tfar)
void gdpt_ref_intersect(void *hs, int n, const double *rays, const double *rdiff, int *ids, double *vtx) {
    Scene *s = (Scene *)hs;
    parallel_for([&](Vector2i rc) {
        int64_t i = rc[0];
        const double *q = rays + 8 * i;
        Ray r{Vector3{q[0], q[1], q[2]}, Vector3{q[3], q[4], q[5]}, q[6], q[7]};
        RayDifferential rd{};
        if (rdiff) { rd.radius = rdiff[2 * i]; rd.spread = rdiff[2 * i + 1]; }
        std::optional<PathVertex> v = intersect(*s, r, rd);
        double *o = vtx + 24 * i;
        if (!v) { ids[3 * i] = ids[3 * i + 1] = ids[3 * i + 2] = -1; for (int k = 0; k < 24; k++) o[k] = 0; return; }
        ids[3 * i] = v->shape_id; ids[3 * i + 1] = v->primitive_id; ids[3 * i + 2] = v->material_id;
        const Vector3 *vs[5] = {&v->position, &v->geometric_normal, &v->shading_frame.x, &v->shading_frame.y, &v->shading_frame.n};
        for (int k = 0; k < 5; k++) { o[3 * k] = vs[k]->x; o[3 * k + 1] = vs[k]->y; o[3 * k + 2] = vs[k]->z; }
        o[15] = v->st.x; o[16] = v->st.y; o[17] = v->uv.x; o[18] = v->uv.y;
        o[19] = v->uv_screen_size; o[20] = v->mean_curvature; o[21] = v->ray_radius;
        o[22] = distance(r.org, v->position) / length(r.dir);
        o[23] = 0;
    }, Vector2i(n, 1));
}

// BSDF triple on material `mat_id` for n configurations.
//   vtx: n*15 (geometric_normal3, frame.x3, frame.y3, frame.n3, uv2, uv_screen_size)
//   dirs: n*6 (dir_in3, dir_out3), rnd: n*3 (u, v, w)
//   out: n*10 (eval3, pdf, ok, sdir3, seta, srough)
void gdpt_ref_bsdf(void *hs, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out) {
    Scene *s = (Scene *)hs;
    const Material &m = s->materials[mat_id];
    for (int i = 0; i < n; i++) {
        const double *q = vtx + 15 * i;
        PathVertex v;
        v.position = Vector3{0, 0, 0};
        v.geometric_normal = Vector3{q[0], q[1], q[2]};
        v.shading_frame = Frame(Vector3{q[3], q[4], q[5]}, Vector3{q[6], q[7], q[8]}, Vector3{q[9], q[10], q[11]});
        v.uv = Vector2{q[12], q[13]}; v.uv_screen_size = q[14];
        v.st = v.uv; v.mean_curvature = 0; v.ray_radius = 0; v.material_id = mat_id;
        Vector3 din{dirs[6 * i], dirs[6 * i + 1], dirs[6 * i + 2]}, dout{dirs[6 * i + 3], dirs[6 * i + 4], dirs[6 * i + 5]};
        double *o = out + 10 * i;
        Spectrum f = eval(m, din, dout, v, s->texture_pool);
        o[0] = f.x; o[1] = f.y; o[2] = f.z;
        o[3] = pdf_sample_bsdf(m, din, dout, v, s->texture_pool);
        std::optional<BSDFSampleRecord> r = sample_bsdf(m, din, v, s->texture_pool, Vector2{rnd[3 * i], rnd[3 * i + 1]}, rnd[3 * i + 2]);
        if (r) { o[4] = 1; o[5] = r->dir_out.x; o[6] = r->dir_out.y; o[7] = r->dir_out.z; o[8] = r->eta; o[9] = r->roughness; }
        else { o[4] = 0; o[5] = o[6] = o[7] = o[8] = o[9] = 0; }
    }
}

// Flatten the parsed reference Scene into a GDPT1 container file.
int gdpt_ref_export(void *hs, const char *path) {
    Scene *s = (Scene *)hs;
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    fwrite("GDPT1\0\0\0", 1, 8, f);
    BlobWriter w{f};
    const Camera &cam = s->camera;
    w.arr("camera.sample_to_cam", "f8", {4, 4}, cam.sample_to_cam.data, 128);
    w.arr("camera.cam_to_world", "f8", {4, 4}, cam.cam_to_world.data, 128);
    int32_t size[2] = {cam.width, cam.height};
    w.arr("camera.size", "i4", {2}, size, 8);
    double filt[2] = {0, 0};
    if (auto *b = std::get_if<Box>(&cam.filter)) { filt[0] = 0; filt[1] = b->width; }
    else if (auto *t = std::get_if<Tent>(&cam.filter)) { filt[0] = 1; filt[1] = t->width; }
    else if (auto *g = std::get_if<Gaussian>(&cam.filter)) { filt[0] = 2; filt[1] = g->stddev; }
    w.arr("camera.filter", "f8", {2}, filt, 16);
    int32_t opt[4] = {s->options.samples_per_pixel, s->options.max_depth, s->options.rr_depth, (int)s->options.integrator};
    w.arr("options", "i4", {4}, opt, 16);
    double bnd[4] = {s->bounds.radius, s->bounds.center.x, s->bounds.center.y, s->bounds.center.z};
    w.arr("bounds", "f8", {4}, bnd, 32);

    // shapes
    std::vector<int32_t> shapes; std::vector<int64_t> offs; std::vector<double> spheres;
    std::vector<double> pos, nrm, uv; std::vector<int32_t> idx;
    for (const Shape &sh : s->shapes) {
        if (auto *sp = std::get_if<Sphere>(&sh)) {
            shapes.insert(shapes.end(), {0, sp->material_id, sp->area_light_id, 0, 0, 0});
            spheres.insert(spheres.end(), {sp->position.x, sp->position.y, sp->position.z, sp->radius});
            offs.insert(offs.end(), {(int64_t)pos.size() / 3, (int64_t)idx.size() / 3, (int64_t)nrm.size() / 3, (int64_t)uv.size() / 2});
        } else if (auto *m = std::get_if<TriangleMesh>(&sh)) {
            int flags = (m->normals.size() ? 1 : 0) | (m->uvs.size() ? 2 : 0);
            shapes.insert(shapes.end(), {1, m->material_id, m->area_light_id, (int)m->positions.size(), (int)m->indices.size(), flags});
            spheres.insert(spheres.end(), {0, 0, 0, 0});
            offs.insert(offs.end(), {(int64_t)pos.size() / 3, (int64_t)idx.size() / 3, (int64_t)nrm.size() / 3, (int64_t)uv.size() / 2});
            for (auto &p : m->positions) { pos.push_back(p.x); pos.push_back(p.y); pos.push_back(p.z); }
            for (auto &t : m->indices) { idx.push_back(t[0]); idx.push_back(t[1]); idx.push_back(t[2]); }
            for (auto &p : m->normals) { nrm.push_back(p.x); nrm.push_back(p.y); nrm.push_back(p.z); }
            for (auto &p : m->uvs) { uv.push_back(p.x); uv.push_back(p.y); }
        }
    }
    int64_t ns = (int64_t)s->shapes.size();
    w.arr("shapes", "i4", {ns, 6}, shapes.data(), shapes.size() * 4);
    w.arr("shape_offsets", "i8", {ns, 4}, offs.data(), offs.size() * 8);
    w.arr("spheres", "f8", {ns, 4}, spheres.data(), spheres.size() * 8);
    w.arr("positions", "f8", {(int64_t)pos.size() / 3, 3}, pos.data(), pos.size() * 8);
    w.arr("indices", "i4", {(int64_t)idx.size() / 3, 3}, idx.data(), idx.size() * 4);
    w.arr("normals", "f8", {(int64_t)nrm.size() / 3, 3}, nrm.data(), nrm.size() * 8);
    w.arr("uvs", "f8", {(int64_t)uv.size() / 2, 2}, uv.data(), uv.size() * 8);

    // materials
    MatFlat mf;
    for (const Material &m : s->materials) {
        if (auto *p = std::get_if<Lambertian>(&m)) mf.mat(0, 0, {mf.add(p->reflectance)});
        else if (auto *p = std::get_if<RoughPlastic>(&m)) mf.mat(1, p->eta, {mf.add(p->diffuse_reflectance), mf.add(p->specular_reflectance), mf.add(p->roughness)});
        else if (auto *p = std::get_if<RoughDielectric>(&m)) mf.mat(2, p->eta, {mf.add(p->specular_reflectance), mf.add(p->specular_transmittance), mf.add(p->roughness)});
        else if (auto *p = std::get_if<DisneyDiffuse>(&m)) mf.mat(3, 0, {mf.add(p->base_color), mf.add(p->roughness), mf.add(p->subsurface)});
        else if (auto *p = std::get_if<DisneyMetal>(&m)) mf.mat(4, 0, {mf.add(p->base_color), mf.add(p->roughness), mf.add(p->anisotropic)});
        else if (auto *p = std::get_if<DisneyGlass>(&m)) mf.mat(5, p->eta, {mf.add(p->base_color), mf.add(p->roughness), mf.add(p->anisotropic)});
        else if (auto *p = std::get_if<DisneyClearcoat>(&m)) mf.mat(6, 0, {mf.add(p->clearcoat_gloss)});
        else if (auto *p = std::get_if<DisneySheen>(&m)) mf.mat(7, 0, {mf.add(p->base_color), mf.add(p->sheen_tint)});
        else if (auto *p = std::get_if<DisneyBSDF>(&m))
            mf.mat(8, p->eta, {mf.add(p->base_color), mf.add(p->specular_transmission), mf.add(p->metallic), mf.add(p->subsurface),
                               mf.add(p->specular), mf.add(p->roughness), mf.add(p->specular_tint), mf.add(p->anisotropic),
                               mf.add(p->sheen), mf.add(p->sheen_tint), mf.add(p->clearcoat), mf.add(p->clearcoat_gloss)});
    }
    // lights: kind (0 area, 1 envmap); area: shape_id, intensity3; envmap: texture id of `values` (appended to the
    // texture table), scale — its to_world / to_local matrices go into their own arrays (src/light.h:23-30)
    std::vector<double> lights;
    for (const Light &l : s->lights) {
        if (auto *a = std::get_if<DiffuseAreaLight>(&l)) lights.insert(lights.end(), {0, (double)a->shape_id, a->intensity.x, a->intensity.y, a->intensity.z});
        else if (auto *e = std::get_if<Envmap>(&l)) {
            lights.insert(lights.end(), {1, (double)mf.add(e->values), e->scale, 0, 0});
            w.arr("envmap.to_world", "f8", {4, 4}, e->to_world.data, 128);
            w.arr("envmap.to_local", "f8", {4, 4}, e->to_local.data, 128);
        }
    }
    w.arr("materials", "f8", {(int64_t)s->materials.size(), 14}, mf.mats.data(), mf.mats.size() * 8);
    w.arr("textures", "f8", {(int64_t)mf.ntex, 14}, mf.tex.data(), mf.tex.size() * 8);
    w.arr("lights", "f8", {(int64_t)s->lights.size(), 5}, lights.data(), lights.size() * 8);

    export_images<Vector3>(w, "img3", s->texture_pool.image3s);
    export_images<Real>(w, "img1", s->texture_pool.image1s);
    fclose(f);
    return 0;
}

// Shift-mapped gradpath sample (the product's opt-in offset mode, csrc/gdpt_shift.cuh) restated on the reference's OWN
// primitives — sample_primary, intersect, occluded, sample_bsdf / eval / pdf_sample_bsdf, emission — in double, operation
// for operation, so the GPU's FP64 records can be compared bit for bit.  It follows small_gdpt.cpp:153-352 (base path that
// must end on a light, random-number replay, reconnection at the first diffuse pair, J = p/p' resp. the geometry ratio,
// w = P/(P + P_s J)); it is NOT code of the reference, which has no such mode on lajolla scenes: this pins the GPU kernel
// to an independent CPU statement of the same algorithm, not to a golden of the reference.
// rec[23]: radiance3, contrib3, contribX0 3, contribX1 3, contribY0 3, contribY1 3, prob, wX0, wY0, wX1, wY1
// with X0 = (x-1, y), Y0 = (x, y-1), X1 = (x+1, y), Y1 = (x, y+1).
void gdpt_ref_shift_sample(void *hs, int x, int y, uint64_t stream, uint64_t seed, double *rec) {
    const Scene &scene = *(Scene *)hs;
    const int MAX_VERTS = 24, MAX_PREFIX = 12;
    const double SURVIVAL = 0.75;
    const int w = scene.camera.width, h = scene.camera.height;
    const int max_depth = scene.options.max_depth, rr_depth = scene.options.rr_depth;
    const Real eps = get_intersection_epsilon(scene);
    auto diffuse = [&](int material_id) {
        const Material &m = scene.materials[material_id];
        return std::holds_alternative<Lambertian>(m) || std::holds_alternative<DisneyDiffuse>(m);
    };
    auto cls_of = [&](const PathVertex &v) {
        bool light = is_light(scene.shapes[v.shape_id]);
        return ((light || diffuse(v.material_id)) ? 1 : 0) | (light ? 2 : 0);
    };
    double cS[4][3], wS[4], contrib[3] = {0, 0, 0}, rad[3] = {0, 0, 0}, prob = 1.0;
    for (int k = 0; k < 4; k++) { cS[k][0] = cS[k][1] = cS[k][2] = 0.0; wS[k] = 1.0; }
    auto write = [&]() {
        rec[0] = rad[0]; rec[1] = rad[1]; rec[2] = rad[2]; rec[3] = contrib[0]; rec[4] = contrib[1]; rec[5] = contrib[2];
        const int slot[4] = {0, 2, 1, 3};     // record order X0, X1, Y0, Y1; offsets here X0, Y0, X1, Y1
        for (int k = 0; k < 4; k++) for (int c = 0; c < 3; c++) rec[6 + 3 * slot[k] + c] = cS[k][c];
        rec[18] = prob; rec[19] = wS[0]; rec[20] = wS[1]; rec[21] = wS[2]; rec[22] = wS[3];
    };
    gdpt_pcg32 rng; gdpt_oracle_pcg_init(&rng, stream, seed);
    const Real rx = gdpt_oracle_pcg_real(&rng), ry = gdpt_oracle_pcg_real(&rng);
    const RayDifferential rd0 = init_ray_differential(w, h);

    // ---- base path
    Real pu0[MAX_PREFIX], pu1[MAX_PREFIX], pw[MAX_PREFIX], pp[MAX_PREFIX];
    int pclass[MAX_PREFIX + 1];
    Ray ray = sample_primary(scene.camera, Vector2((x + rx) / w, (y + ry) / h));
    std::optional<PathVertex> hit = intersect(scene, ray, rd0);
    bool base_ok = false, offsets_ok = true, prev_diffuse = false;
    int L = -1, dr = -1;
    double C[3] = {1, 1, 1}, P = 1, inv_rr = 1, Sf[3] = {1, 1, 1}, Sp = 1;
    Spectrum Le{0, 0, 0};
    Vector3 wr{0, 0, 1}, pos_dr = ray.org, prev_pos = ray.org, dir_in = ray.dir;
    PathVertex vr;
    if (hit) {
        PathVertex v = *hit;
        for (int j = 0; j < MAX_VERTS; j++) {
            const int cls = cls_of(v);
            if (dr < 0 && j > 0 && prev_diffuse && (cls & 1)) { dr = j - 1; vr = v; pos_dr = prev_pos; }
            if (dr < 0 || j <= dr) { if (j <= MAX_PREFIX) pclass[j] = cls; else offsets_ok = false; }
            if (cls & 2) { Le = emission(v, -dir_in, scene); L = j; base_ok = true; break; }
            if (j + 1 >= MAX_VERTS) break;
            if (max_depth != -1 && j + 2 > max_depth) break;
            const Real u0 = gdpt_oracle_pcg_real(&rng), u1 = gdpt_oracle_pcg_real(&rng), uw = gdpt_oracle_pcg_real(&rng);
            const Material &mat = scene.materials[v.material_id];
            std::optional<BSDFSampleRecord> s = sample_bsdf(mat, -dir_in, v, scene.texture_pool, Vector2{u0, u1}, uw);
            if (!s) break;
            const Spectrum f = eval(mat, -dir_in, s->dir_out, v, scene.texture_pool);
            const Real p = pdf_sample_bsdf(mat, -dir_in, s->dir_out, v, scene.texture_pool);
            if (!(p > 0) || !(f.x > 0 || f.y > 0 || f.z > 0)) break;
            if (dr < 0) { if (j < MAX_PREFIX) { pu0[j] = u0; pu1[j] = u1; pw[j] = uw; pp[j] = p; } else offsets_ok = false; }
            else if (j == dr + 1) wr = s->dir_out;
            else { Sf[0] *= f.x; Sf[1] *= f.y; Sf[2] *= f.z; Sp *= p; }
            C[0] *= f.x; C[1] *= f.y; C[2] *= f.z; P *= p;
            if (j + 2 >= rr_depth) {
                if (!(gdpt_oracle_pcg_real(&rng) <= Real(SURVIVAL))) break;
                inv_rr /= SURVIVAL;
            }
            Ray next{v.position, s->dir_out, eps, infinity<Real>()};
            hit = intersect(scene, next);
            if (!hit) break;
            prev_pos = v.position; prev_diffuse = (cls & 1) != 0; dir_in = s->dir_out;
            v = *hit;
        }
    }
    if (!base_ok) { write(); return; }
    contrib[0] = C[0] * Le.x * inv_rr; contrib[1] = C[1] * Le.y * inv_rr; contrib[2] = C[2] * Le.z * inv_rr;
    for (int c = 0; c < 3; c++) rad[c] = contrib[c] / P;
    prob = P;

    // ---- offsets
    const int ox[4] = {-1, 0, 1, 0}, oy[4] = {0, -1, 0, 1};
    const int jend = dr >= 0 ? dr : L;
    for (int k = 0; k < 4 && offsets_ok; k++) {
        Ray ro = sample_primary(scene.camera, Vector2(((x + ox[k]) + rx) / w, ((y + oy[k]) + ry) / h));
        std::optional<PathVertex> ho = intersect(scene, ro, rd0);
        if (!ho) continue;
        PathVertex vo = *ho;
        Vector3 din = ro.dir;
        double Cs[3] = {1, 1, 1}, Ps = 1, J = 1;
        bool ok = true;
        for (int j = 0; ; j++) {
            if (cls_of(vo) != pclass[j]) { ok = false; break; }
            if (j == jend) break;
            const Material &mat = scene.materials[vo.material_id];
            std::optional<BSDFSampleRecord> s = sample_bsdf(mat, -din, vo, scene.texture_pool, Vector2{pu0[j], pu1[j]}, pw[j]);
            if (!s) { ok = false; break; }
            const Spectrum f = eval(mat, -din, s->dir_out, vo, scene.texture_pool);
            const Real p = pdf_sample_bsdf(mat, -din, s->dir_out, vo, scene.texture_pool);
            if (!(p > 0)) { ok = false; break; }
            Cs[0] *= f.x; Cs[1] *= f.y; Cs[2] *= f.z; Ps *= p; J *= pp[j] / p;
            Ray next{vo.position, s->dir_out, eps, infinity<Real>()};
            ho = intersect(scene, next);
            if (!ho) { ok = false; break; }
            din = s->dir_out;
            vo = *ho;
        }
        if (!ok) continue;
        if (dr < 0) {
            const Spectrum e = emission(vo, -din, scene);
            Cs[0] *= e.x; Cs[1] *= e.y; Cs[2] *= e.z;
        } else {
            const Vector3 dv = vr.position - vo.position;
            const Real d2s = dot(dv, dv);
            if (!(d2s > 0)) continue;
            const Real dist = sqrt(d2s);
            const Vector3 ds = dv / dist;
            const Material &mat = scene.materials[vo.material_id];
            const Spectrum f = eval(mat, -din, ds, vo, scene.texture_pool);
            const Real p = pdf_sample_bsdf(mat, -din, ds, vo, scene.texture_pool);
            const Vector3 bv = vr.position - pos_dr;
            const Real d2b = dot(bv, bv);
            const Real cos_s = fabs(dot(ds, vr.geometric_normal)), cos_b = fabs(dot(bv, vr.geometric_normal)) / sqrt(d2b);
            if (!(p > 0 && cos_b > 0 && d2b > 0)) continue;
            J *= (cos_s / d2s) / (cos_b / d2b);
            Cs[0] *= f.x; Cs[1] *= f.y; Cs[2] *= f.z; Ps *= p;
            if (L == dr + 1) {
                const Spectrum e = emission(vr, -ds, scene);
                Cs[0] *= e.x; Cs[1] *= e.y; Cs[2] *= e.z;
            } else {
                const Material &mr = scene.materials[vr.material_id];
                const Spectrum ft = eval(mr, -ds, wr, vr, scene.texture_pool);
                const Real pt = pdf_sample_bsdf(mr, -ds, wr, vr, scene.texture_pool);
                if (!(pt > 0)) continue;
                Cs[0] *= ft.x * Sf[0] * Le.x; Cs[1] *= ft.y * Sf[1] * Le.y; Cs[2] *= ft.z * Sf[2] * Le.z;
                Ps *= pt * Sp;
            }
            if (occluded(scene, Ray{vo.position, ds, eps, (1 - eps) * dist})) continue;
        }
        if (!(J > 0.0) || !std::isfinite(J) || !std::isfinite(Ps)) continue;
        for (int c = 0; c < 3; c++) cS[k][c] = Cs[c] * J * inv_rr;
        wS[k] = P / (P + Ps * J);
    }
    write();
}

} // extern "C"
