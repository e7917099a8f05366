#include "gdpt_container.h"

#include <zlib.h>
#include <algorithm>

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
    auto need = [&](size_t n) { return pos <= raw.size() && n <= raw.size() - pos; };   // no wrap-around on a file-controlled length
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
    // Envmap::to_world / to_local (src/light.h:23-30), present when the scene has an environment map
    desc.envmap_to_world = desc.envmap_to_local = nullptr;
    if (c.count("envmap.to_world") && c.count("envmap.to_local")) {
        if (!take(c, "envmap.to_world", "f8", envmap_to_world, err) || !take(c, "envmap.to_local", "f8", envmap_to_local, err)) return err;
        if (envmap_to_world.size() != 16 || envmap_to_local.size() != 16) return "envmap matrices have the wrong shape";
        desc.envmap_to_world = envmap_to_world.data(); desc.envmap_to_local = envmap_to_local.data();
    }

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

// ---- OpenEXR output, the reference's default film format (image.exr; imwrite -> tinyexr SaveEXR(data, w, h, 3,
// save_as_fp16 = 1), /root/reference/src/image.cpp:150-171): single-part scan-line file, channels B, G, R of type HALF,
// ZIP compression in blocks of 16 scan lines (NONE below 16x16 like SaveEXR), increasing-Y line order.
static unsigned short float_to_half(float f) {      // IEEE round to nearest even, with denormals, inf and NaN
    unsigned int x; std::memcpy(&x, &f, 4);
    unsigned int sign = (x >> 16) & 0x8000u, mant = x & 0x007fffffu;
    int exp = (int)((x >> 23) & 0xff);
    if (exp == 0xff) return (unsigned short)(sign | 0x7c00u | (mant ? 0x0200u | (mant >> 13) : 0));
    int e = exp - 127 + 15;
    if (e >= 0x1f) return (unsigned short)(sign | 0x7c00u);
    if (e <= 0) {
        if (e < -10) return (unsigned short)sign;
        mant |= 0x00800000u;
        unsigned int shift = (unsigned int)(14 - e), h = mant >> shift, rem = mant & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) h++;
        return (unsigned short)(sign | h);
    }
    unsigned int h = ((unsigned int)e << 10) | (mant >> 13), rem = mant & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) h++;      // a carry into the exponent is the correct rounding
    return (unsigned short)(sign | h);
}

std::string write_exr(const std::string &path, int w, int h, const double *rgb) {
    std::vector<unsigned char> out;
    auto put = [&](const void *p, size_t n) { const unsigned char *b = (const unsigned char *)p; out.insert(out.end(), b, b + n); };
    auto put_i32 = [&](int v) { put(&v, 4); };
    auto put_f32 = [&](float v) { put(&v, 4); };
    auto put_str = [&](const char *s) { put(s, std::strlen(s) + 1); };
    auto attr = [&](const char *name, const char *type, int size) { put_str(name); put_str(type); put_i32(size); };
    const unsigned char magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};
    put(magic, 8);
    attr("channels", "chlist", 3 * 18 + 1);
    for (const char *c : {"B", "G", "R"}) { put_str(c); put_i32(1 /* HALF */); unsigned char lin[4] = {0, 0, 0, 0}; put(lin, 4); put_i32(1); put_i32(1); }
    out.push_back(0);
    const bool zip = !(w < 16 && h < 16);
    attr("compression", "compression", 1); out.push_back(zip ? 3 : 0);
    attr("dataWindow", "box2i", 16); put_i32(0); put_i32(0); put_i32(w - 1); put_i32(h - 1);
    attr("displayWindow", "box2i", 16); put_i32(0); put_i32(0); put_i32(w - 1); put_i32(h - 1);
    attr("lineOrder", "lineOrder", 1); out.push_back(0);
    attr("pixelAspectRatio", "float", 4); put_f32(1.0f);
    attr("screenWindowCenter", "v2f", 8); put_f32(0.0f); put_f32(0.0f);
    attr("screenWindowWidth", "float", 4); put_f32(1.0f);
    out.push_back(0);
    const int lines = zip ? 16 : 1, nblocks = (h + lines - 1) / lines;
    const size_t table = out.size();
    out.resize(out.size() + 8 * (size_t)nblocks);
    std::vector<unsigned char> raw, tmp, comp;
    for (int b = 0; b < nblocks; b++) {
        const int y0 = b * lines, y1 = std::min(h, y0 + lines);
        raw.clear();
        for (int y = y0; y < y1; y++)
            for (int c = 2; c >= 0; c--)             // B, G, R
                for (int x = 0; x < w; x++) { unsigned short hv = float_to_half((float)rgb[((size_t)y * w + x) * 3 + c]); raw.push_back((unsigned char)(hv & 0xff)); raw.push_back((unsigned char)(hv >> 8)); }
        const unsigned char *payload = raw.data(); size_t psize = raw.size();
        if (zip) {
            const size_t n = raw.size();
            tmp.resize(n);
            size_t t1 = 0, t2 = (n + 1) / 2;         // even bytes first, then odd bytes
            for (size_t i = 0; i < n; i++) { if (i & 1) tmp[t2++] = raw[i]; else tmp[t1++] = raw[i]; }
            unsigned char prev = tmp[0];
            for (size_t i = 1; i < n; i++) { unsigned char cur = tmp[i]; tmp[i] = (unsigned char)((int)cur - (int)prev + 128 + 256); prev = cur; }
            uLongf clen = compressBound((uLong)n);
            comp.resize(clen);
            if (compress(comp.data(), &clen, tmp.data(), (uLong)n) != Z_OK) return "zlib compress failed for " + path;
            if (clen < n) { payload = comp.data(); psize = clen; }   // a block that does not shrink is stored raw
        }
        unsigned long long off = out.size();
        std::memcpy(out.data() + table + 8 * (size_t)b, &off, 8);
        put_i32(y0); put_i32((int)psize); put(payload, psize);
    }
    std::ofstream f(path, std::ios::binary);
    if (!f) return "cannot write " + path;
    f.write((const char *)out.data(), (std::streamsize)out.size());
    return f ? "" : "short write to " + path;
}

std::string write_image(const std::string &path, int w, int h, const double *rgb) {
    auto ends = [&](const char *suf) { size_t n = std::strlen(suf); return path.size() >= n && path.compare(path.size() - n, n, suf) == 0; };
    if (ends(".pfm")) return write_pfm(path, w, h, rgb);
    if (ends(".exr")) return write_exr(path, w, h, rgb);
    return "Unsupported image format: " + path;       // imwrite writes nothing for other names (image.cpp:134-173)
}

} // namespace gdpt_host
