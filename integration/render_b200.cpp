// Reference-side binding of libgdpt.so: the file a lajolla maintainer adds to the reference tree as src/render_b200.cpp
// (INTEGRATION.md).  It is compiled HERE, against the reference's own headers under /root/reference/src, by
// integration/build_lajolla_gdpt.sh, and linked with the reference's own main.cpp + scene parser + loaders into
// oracle/_ref/lajolla_gdpt — `./lajolla_gdpt scene.xml` is then the reference's CLI and XML parser
// (/root/reference/src/main.cpp:11-51, src/parsers/parse_scene.cpp:1615-1630) with the `gradpath` arm of render()
// (src/render.cpp:387-388) running on the GPU through the C ABI of include/gdpt.h.
//
// Nothing of the product lives here: this file only flattens the fields of `Scene` that gradpath reads
// (src/scene.h:43-83) into gdpt_scene_desc and wraps the result in an Image3.
#include "render.h"
#include "scene.h"
#include "image.h"
#include "flexception.h"
#include "gdpt.h"

#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <variant>
#include <vector>

namespace {

struct FlatScene {
    gdpt_scene_desc d;
    std::vector<int32_t> shapes, indices;
    std::vector<int64_t> offsets;
    std::vector<double> spheres, positions, normals, uvs, materials, textures, lights;
    std::vector<gdpt_image> images3, images1;
    double env_to_world[16], env_to_local[16];
    int ntex = 0;
};

void copy4x4(const Matrix4x4 &m, double *out) {
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) out[4 * i + j] = m(i, j);
}

void put3(std::vector<double> &o, const Vector3 &v) { o.push_back(v.x); o.push_back(v.y); o.push_back(v.z); }
void put3(std::vector<double> &o, const Real &v) { o.push_back(v); o.push_back(v); o.push_back(v); }

// one row of gdpt_scene_desc::textures: kind, is_spectrum, v0[3], v1[3], image id, uscale, vscale, uoffset, voffset, pad
// (src/texture.h:80-113: ConstantTexture / ImageTexture / CheckerboardTexture)
template <typename T> int add_texture(FlatScene &f, const Texture<T> &t) {
    std::vector<double> &o = f.textures;
    const double spec = std::is_same<T, Vector3>::value ? 1.0 : 0.0;
    if (auto *c = std::get_if<ConstantTexture<T>>(&t)) {
        o.push_back(0); o.push_back(spec); put3(o, c->value); put3(o, T{}); o.push_back(-1);
        o.push_back(1); o.push_back(1); o.push_back(0); o.push_back(0); o.push_back(0);
    } else if (auto *im = std::get_if<ImageTexture<T>>(&t)) {
        o.push_back(1); o.push_back(spec); put3(o, T{}); put3(o, T{}); o.push_back(im->texture_id);
        o.push_back(im->uscale); o.push_back(im->vscale); o.push_back(im->uoffset); o.push_back(im->voffset); o.push_back(0);
    } else {
        const auto &cb = std::get<CheckerboardTexture<T>>(t);
        o.push_back(2); o.push_back(spec); put3(o, cb.color0); put3(o, cb.color1); o.push_back(-1);
        o.push_back(cb.uscale); o.push_back(cb.vscale); o.push_back(cb.uoffset); o.push_back(cb.voffset); o.push_back(0);
    }
    return f.ntex++;
}

// one row of gdpt_scene_desc::materials: type (index in the std::variant of src/material.h:112-120), eta, 12 texture ids
void add_material(FlatScene &f, int type, double eta, std::initializer_list<int> ids) {
    f.materials.push_back(type); f.materials.push_back(eta);
    int n = 0;
    for (int id : ids) { f.materials.push_back(id); n++; }
    for (; n < 12; n++) f.materials.push_back(-1);
}

template <typename T> void add_images(std::vector<gdpt_image> &out, const std::vector<Mipmap<T>> &pool) {
    for (const Mipmap<T> &m : pool) {       // level 0 only: gdpt_create rebuilds the chain like make_mipmap (src/mipmap.h:28-44)
        const Image<T> &img = m.images[0];
        gdpt_image gi; memset(&gi, 0, sizeof(gi));
        gi.width = img.width; gi.height = img.height; gi.channels = std::is_same<T, Vector3>::value ? 3 : 1;
        gi.data = reinterpret_cast<const double *>(img.data.data());
        out.push_back(gi);
    }
}

void flatten(const Scene &s, FlatScene &f) {
    gdpt_scene_desc &d = f.d;
    memset(&d, 0, sizeof(d));
    // camera (src/camera.h:10-24) and pixel filter (src/filter.h:31-43)
    copy4x4(s.camera.sample_to_cam, d.sample_to_cam);
    copy4x4(s.camera.cam_to_world, d.cam_to_world);
    d.width = s.camera.width; d.height = s.camera.height;
    if (auto *b = std::get_if<Box>(&s.camera.filter)) { d.filter_kind = 0; d.filter_param = b->width; }
    else if (auto *t = std::get_if<Tent>(&s.camera.filter)) { d.filter_kind = 1; d.filter_param = t->width; }
    else { d.filter_kind = 2; d.filter_param = std::get<Gaussian>(s.camera.filter).stddev; }
    // shapes (src/shape.h:26-53); shape id == index == Embree geomID (src/scene.cpp:24-26)
    for (const Shape &sh : s.shapes) {
        f.offsets.insert(f.offsets.end(), {(int64_t)f.positions.size() / 3, (int64_t)f.indices.size() / 3, (int64_t)f.normals.size() / 3, (int64_t)f.uvs.size() / 2});
        if (auto *sp = std::get_if<Sphere>(&sh)) {
            f.shapes.insert(f.shapes.end(), {0, sp->material_id, sp->area_light_id, 0, 0, 0});
            f.spheres.insert(f.spheres.end(), {sp->position.x, sp->position.y, sp->position.z, sp->radius});
        } else {
            const TriangleMesh &m = std::get<TriangleMesh>(sh);
            const int flags = (m.normals.empty() ? 0 : 1) | (m.uvs.empty() ? 0 : 2);
            f.shapes.insert(f.shapes.end(), {1, m.material_id, m.area_light_id, (int)m.positions.size(), (int)m.indices.size(), flags});
            f.spheres.insert(f.spheres.end(), {0, 0, 0, 0});
            for (const Vector3 &p : m.positions) put3(f.positions, p);
            for (const Vector3i &t : m.indices) f.indices.insert(f.indices.end(), {t[0], t[1], t[2]});
            for (const Vector3 &n : m.normals) put3(f.normals, n);
            for (const Vector2 &u : m.uvs) { f.uvs.push_back(u.x); f.uvs.push_back(u.y); }
        }
    }
    // materials (src/material.h:12-110), in the order of the Material variant
    for (const Material &m : s.materials) {
        if (auto *p = std::get_if<Lambertian>(&m)) add_material(f, 0, 0, {add_texture(f, p->reflectance)});
        else if (auto *p = std::get_if<RoughPlastic>(&m)) { int a = add_texture(f, p->diffuse_reflectance), b = add_texture(f, p->specular_reflectance), c = add_texture(f, p->roughness); add_material(f, 1, p->eta, {a, b, c}); }
        else if (auto *p = std::get_if<RoughDielectric>(&m)) { int a = add_texture(f, p->specular_reflectance), b = add_texture(f, p->specular_transmittance), c = add_texture(f, p->roughness); add_material(f, 2, p->eta, {a, b, c}); }
        else if (auto *p = std::get_if<DisneyDiffuse>(&m)) { int a = add_texture(f, p->base_color), b = add_texture(f, p->roughness), c = add_texture(f, p->subsurface); add_material(f, 3, 0, {a, b, c}); }
        else if (auto *p = std::get_if<DisneyMetal>(&m)) { int a = add_texture(f, p->base_color), b = add_texture(f, p->roughness), c = add_texture(f, p->anisotropic); add_material(f, 4, 0, {a, b, c}); }
        else if (auto *p = std::get_if<DisneyGlass>(&m)) { int a = add_texture(f, p->base_color), b = add_texture(f, p->roughness), c = add_texture(f, p->anisotropic); add_material(f, 5, p->eta, {a, b, c}); }
        else if (auto *p = std::get_if<DisneyClearcoat>(&m)) add_material(f, 6, 0, {add_texture(f, p->clearcoat_gloss)});
        else if (auto *p = std::get_if<DisneySheen>(&m)) { int a = add_texture(f, p->base_color), b = add_texture(f, p->sheen_tint); add_material(f, 7, 0, {a, b}); }
        else {
            const DisneyBSDF &q = std::get<DisneyBSDF>(m);
            int t[12] = {add_texture(f, q.base_color), add_texture(f, q.specular_transmission), add_texture(f, q.metallic), add_texture(f, q.subsurface),
                         add_texture(f, q.specular), add_texture(f, q.roughness), add_texture(f, q.specular_tint), add_texture(f, q.anisotropic),
                         add_texture(f, q.sheen), add_texture(f, q.sheen_tint), add_texture(f, q.clearcoat), add_texture(f, q.clearcoat_gloss)};
            add_material(f, 8, q.eta, {t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7], t[8], t[9], t[10], t[11]});
        }
    }
    // lights (src/light.h:15-34): area lights, and the environment map (ignored by gradpath, used by `path`)
    d.envmap_to_world = d.envmap_to_local = nullptr;
    for (const Light &l : s.lights) {
        if (auto *a = std::get_if<DiffuseAreaLight>(&l)) f.lights.insert(f.lights.end(), {0, (double)a->shape_id, a->intensity.x, a->intensity.y, a->intensity.z});
        else {
            const Envmap &e = std::get<Envmap>(l);
            f.lights.insert(f.lights.end(), {1, (double)add_texture(f, e.values), e.scale, 0, 0});
            copy4x4(e.to_world, f.env_to_world); copy4x4(e.to_local, f.env_to_local);
            d.envmap_to_world = f.env_to_world; d.envmap_to_local = f.env_to_local;
        }
    }
    add_images(f.images3, s.texture_pool.image3s);
    add_images(f.images1, s.texture_pool.image1s);
    d.n_shapes = (int)s.shapes.size(); d.shapes = f.shapes.data(); d.shape_offsets = f.offsets.data(); d.spheres = f.spheres.data();
    d.n_positions = (int64_t)f.positions.size() / 3; d.positions = f.positions.data();
    d.n_triangles = (int64_t)f.indices.size() / 3; d.indices = f.indices.data();
    d.n_normals = (int64_t)f.normals.size() / 3; d.normals = f.normals.data();
    d.n_uvs = (int64_t)f.uvs.size() / 2; d.uvs = f.uvs.data();
    d.n_materials = (int)s.materials.size(); d.materials = f.materials.data();
    d.n_textures = f.ntex; d.textures = f.textures.data();
    d.n_images3 = (int)f.images3.size(); d.images3 = f.images3.data();
    d.n_images1 = (int)f.images1.size(); d.images1 = f.images1.data();
    d.n_lights = (int)s.lights.size(); d.lights = f.lights.data();
    d.max_depth = s.options.max_depth; d.rr_depth = s.options.rr_depth; d.samples_per_pixel = s.options.samples_per_pixel;
}

Image3 wrap(int w, int h, const std::vector<double> &v) {
    Image3 img(w, h);
    for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) { const double *q = &v[3 * ((size_t)y * w + x)]; img(x, y) = Vector3{q[0], q[1], q[2]}; }
    return img;
}

int env_int(const char *name, int dflt) { const char *e = getenv(name); return e ? atoi(e) : dflt; }

}  // namespace

// Replaces gradient_path_render (src/render.cpp:257-370) in the GradPath arm of render() (src/render.cpp:387-388).
// Environment (all optional): GDPT_DEVICE n; GDPT_PRECISION f32|f64 (default f32; f64 = the reference's Real = double,
// bit-comparable records); GDPT_OFFSETS reference|shift; GDPT_WRITE_BUFFERS prefix -> <prefix>_primal/_dx/_dy.pfm beside
// the Poisson image main() writes (render() drops them: src/render.cpp:355-369).
// The sample count is the XML's <sampler sampleCount> (the shipped gradient_path_render hard-codes 1000, src/render.cpp:293).
Image3 gradient_path_render_b200(const Scene &scene) {
    FlatScene f; flatten(scene, f);
    gdpt_ctx *ctx = nullptr;
    if (gdpt_create(&f.d, env_int("GDPT_DEVICE", 0), &ctx) != GDPT_OK) Error(std::string("gdpt_create: ") + gdpt_last_error());
    const int w = f.d.width, h = f.d.height;
    const size_t n = (size_t)3 * w * h;
    std::vector<double> c(n), cx(n), cy(n), out(n);          // c / cx / cy / out of src/render.cpp:336-338
    gdpt_render_params p; memset(&p, 0, sizeof(p));
    p.spp = scene.options.samples_per_pixel; p.max_depth = -2; p.rr_depth = -2;
    p.data_cost = 0.04;                                      // src/render.cpp:353
    const char *prec = getenv("GDPT_PRECISION");
    p.precision = (prec && std::string(prec) == "f64") ? GDPT_PRECISION_F64 : GDPT_PRECISION_F32;
    const char *off = getenv("GDPT_OFFSETS");
    p.offset_mode = (off && std::string(off) == "shift") ? GDPT_OFFSETS_SHIFT : GDPT_OFFSETS_REFERENCE;
    gdpt_images img{c.data(), cx.data(), cy.data(), out.data()};
    const int rc = gdpt_render(ctx, &p, &img);
    const std::string err = rc == GDPT_OK ? "" : gdpt_last_error();
    gdpt_destroy(ctx);
    if (rc != GDPT_OK) Error("gdpt_render: " + err);
    if (const char *prefix = getenv("GDPT_WRITE_BUFFERS")) {
        imwrite(std::string(prefix) + "_primal.pfm", wrap(w, h, c));
        imwrite(std::string(prefix) + "_dx.pfm", wrap(w, h, cx));
        imwrite(std::string(prefix) + "_dy.pfm", wrap(w, h, cy));
    }
    return wrap(w, h, out);
}
