#include "gdpt_container.h"

#include <zlib.h>

#include <cstdio>
#include <cstring>
#include <fstream>

namespace gdpt_host {

static bool ends_with(const std::string &s, const char *suffix) {
    size_t n = strlen(suffix);
    return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}

static std::string inflate_all(const std::vector<uint8_t> &in, std::vector<uint8_t> &out) {
    z_stream zs; memset(&zs, 0, sizeof(zs));
    if (inflateInit(&zs) != Z_OK) return "inflateInit failed";
    zs.next_in = const_cast<Bytef *>(in.data()); zs.avail_in = (uInt)in.size();
    out.clear();
    std::vector<uint8_t> chunk(1 << 22);
    int rc = Z_OK;
    while (rc != Z_STREAM_END) {
        zs.next_out = chunk.data(); zs.avail_out = (uInt)chunk.size();
        rc = inflate(&zs, Z_NO_FLUSH);
        if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&zs); return "corrupt zlib stream"; }
        out.insert(out.end(), chunk.data(), chunk.data() + (chunk.size() - zs.avail_out));
    }
    inflateEnd(&zs);
    return "";
}

std::string read_container(const std::string &path, Container &out) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return "cannot open " + path;
    std::vector<uint8_t> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    if (ends_with(path, ".z")) {
        std::vector<uint8_t> plain;
        std::string e = inflate_all(raw, plain);
        if (!e.empty()) return path + ": " + e;
        raw.swap(plain);
    }
    if (raw.size() < 8 || memcmp(raw.data(), "GDPT1", 5) != 0) return path + " is not a GDPT1 container";
    size_t pos = 8;
    auto need = [&](size_t n) { return pos + n <= raw.size(); };
    while (pos < raw.size()) {
        uint32_t nl, nd; uint64_t nb;
        if (!need(4)) return path + ": truncated record";
        memcpy(&nl, &raw[pos], 4); pos += 4;
        if (!need(nl + 8 + 4)) return path + ": truncated record";
        std::string name((const char *)&raw[pos], nl); pos += nl;
        Array a;
        a.dtype = std::string((const char *)&raw[pos], strnlen((const char *)&raw[pos], 8)); pos += 8;
        memcpy(&nd, &raw[pos], 4); pos += 4;
        if (nd > 8 || !need(8 * (size_t)nd + 8)) return path + ": bad record " + name;
        a.shape.resize(nd);
        if (nd) memcpy(a.shape.data(), &raw[pos], 8 * (size_t)nd);
        pos += 8 * (size_t)nd;
        memcpy(&nb, &raw[pos], 8); pos += 8;
        if (!need(nb)) return path + ": truncated payload of " + name;
        a.bytes.assign(raw.begin() + pos, raw.begin() + pos + nb); pos += nb;
        out[name] = std::move(a);
    }
    return "";
}

// ---- flattening ---------------------------------------------------------------------------
static const Array *find(const Container &c, const std::string &name, const char *dtype, std::string &err) {
    auto it = c.find(name);
    if (it == c.end()) { err = "container has no array '" + name + "'"; return nullptr; }
    if (it->second.dtype != dtype) { err = "array '" + name + "' has dtype " + it->second.dtype + ", expected " + dtype; return nullptr; }
    return &it->second;
}

template <typename T> static bool take(const Container &c, const std::string &name, const char *dtype, std::vector<T> &dst, std::string &err) {
    const Array *a = find(c, name, dtype, err);
    if (!a) return false;
    if (a->bytes.size() != a->count() * sizeof(T)) { err = "array '" + name + "' has an inconsistent size"; return false; }
    dst.assign(a->as<T>(), a->as<T>() + a->count());
    return true;
}

std::string FlatScene::build(const Container &c) {
    std::string err;
    std::vector<double> s2c, c2w, filt; std::vector<int32_t> size, opts;
    if (!take(c, "camera.sample_to_cam", "f8", s2c, err) || !take(c, "camera.cam_to_world", "f8", c2w, err) ||
        !take(c, "camera.size", "i4", size, err) || !take(c, "camera.filter", "f8", filt, err) || !take(c, "options", "i4", opts, err)) return err;
    if (s2c.size() != 16 || c2w.size() != 16 || size.size() != 2 || filt.size() != 2 || opts.size() != 4) return "camera/options arrays have the wrong shape";
    memcpy(desc.sample_to_cam, s2c.data(), sizeof(desc.sample_to_cam));
    memcpy(desc.cam_to_world, c2w.data(), sizeof(desc.cam_to_world));
    desc.width = size[0]; desc.height = size[1];
    desc.filter_kind = (int)filt[0]; desc.filter_param = filt[1];
    desc.samples_per_pixel = opts[0]; desc.max_depth = opts[1]; desc.rr_depth = opts[2]; integrator = opts[3];   // options = spp, maxDepth, rrDepth, integrator

    if (!take(c, "shapes", "i4", shapes, err) || !take(c, "shape_offsets", "i8", offsets, err) || !take(c, "spheres", "f8", spheres, err) ||
        !take(c, "positions", "f8", positions, err) || !take(c, "indices", "i4", indices, err) || !take(c, "normals", "f8", normals, err) ||
        !take(c, "uvs", "f8", uvs, err) || !take(c, "materials", "f8", materials, err) || !take(c, "textures", "f8", textures, err) ||
        !take(c, "lights", "f8", lights, err)) return err;
    desc.n_shapes = (int)(shapes.size() / 6);
    if (offsets.size() != 4 * (size_t)desc.n_shapes || spheres.size() != 4 * (size_t)desc.n_shapes) return "shape tables disagree in length";
    desc.n_positions = (int64_t)(positions.size() / 3); desc.n_triangles = (int64_t)(indices.size() / 3);
    desc.n_normals = (int64_t)(normals.size() / 3); desc.n_uvs = (int64_t)(uvs.size() / 2);
    desc.n_materials = (int)(materials.size() / 14); desc.n_textures = (int)(textures.size() / 14); desc.n_lights = (int)(lights.size() / 5);
    auto pad = [](auto &v, size_t n) { if (v.size() < n) v.resize(n); };   // never hand out a null pointer for an empty table
    pad(positions, 3); pad(indices, 3); pad(normals, 3); pad(uvs, 2); pad(materials, 14); pad(textures, 14); pad(lights, 5); pad(spheres, 4);
    desc.shapes = shapes.data(); desc.shape_offsets = offsets.data(); desc.spheres = spheres.data();
    desc.positions = positions.data(); desc.indices = indices.data(); desc.normals = normals.data(); desc.uvs = uvs.data();
    desc.materials = materials.data(); desc.textures = textures.data(); desc.lights = lights.data();

    // images: "<prefix>.<i>.size" + either ".data" (doubles) or ".palette" + ".index" (bytes)
    for (int pass = 0; pass < 2; pass++) {
        const char *prefix = pass == 0 ? "img3" : "img1";
        std::vector<gdpt_image> &dst = pass == 0 ? images3 : images1;
        for (int i = 0;; i++) {
            std::string base = std::string(prefix) + "." + std::to_string(i);
            if (!c.count(base + ".size")) break;
            std::vector<int32_t> wh;
            if (!take(c, base + ".size", "i4", wh, err) || wh.size() != 2) return err.empty() ? base + ".size has the wrong shape" : err;
            gdpt_image g{}; g.width = wh[0]; g.height = wh[1]; g.channels = pass == 0 ? 3 : 1;
            const size_t n = (size_t)g.width * g.height * g.channels;
            if (c.count(base + ".data")) {
                image_data.emplace_back();
                if (!take(c, base + ".data", "f8", image_data.back(), err)) return err;
                if (image_data.back().size() != n) return base + ".data has the wrong size";
                g.data = image_data.back().data();
            } else {
                palettes.emplace_back(); image_index.emplace_back();
                if (!take(c, base + ".palette", "f8", palettes.back(), err) || !take(c, base + ".index", "u1", image_index.back(), err)) return err;
                if (image_index.back().size() != n) return base + ".index has the wrong size";
                g.palette = palettes.back().data(); g.palette_size = (int)palettes.back().size(); g.index = image_index.back().data();
            }
            dst.push_back(g);
        }
    }
    desc.n_images3 = (int)images3.size(); desc.n_images1 = (int)images1.size();
    pad(images3, 1); pad(images1, 1);
    desc.images3 = images3.data(); desc.images1 = images1.data();
    return "";
}

std::string write_pfm(const std::string &path, int w, int h, const double *rgb) {
    std::ofstream f(path, std::ios::binary);
    if (!f) return "cannot write " + path;
    f << "PF\n" << w << " " << h << "\n-1\n";
    std::vector<float> row((size_t)w * 3);
    for (int y = 0; y < h; y++) {
        for (int i = 0; i < w * 3; i++) row[i] = (float)rgb[(size_t)y * w * 3 + i];
        f.write((const char *)row.data(), (std::streamsize)(row.size() * sizeof(float)));
    }
    return f ? "" : "short write to " + path;
}

} // namespace gdpt_host
