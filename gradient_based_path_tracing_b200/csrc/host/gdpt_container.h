// "GDPT1" flat scene container (the host data model of Appendix A of SURVEY.md, written by the
// reference's own parser through oracle/ref_shims/ref_hooks.cpp::gdpt_ref_export) and its
// flattening into gdpt_scene_desc — the C++ twin of gradient_based_path_tracing_b200/scene.py.
//
//   8-byte magic "GDPT1\0\0\0", then records
//     u32 name_len, name, char dtype[8], u32 ndim, i64 shape[ndim], u64 nbytes, payload
//   a file ending in ".z" is the same stream zlib-compressed.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "gdpt.h"

namespace gdpt_host {

struct Array {
    std::string dtype;              // "f8", "f4", "i4", "i8", "u1", "u2"
    std::vector<int64_t> shape;
    std::vector<uint8_t> bytes;
    size_t count() const { size_t n = 1; for (int64_t s : shape) n *= (size_t)s; return n; }
    template <typename T> const T *as() const { return reinterpret_cast<const T *>(bytes.data()); }
};

using Container = std::map<std::string, Array>;

// Reads `path` (optionally ".z") into `out`; arrays already present are replaced (camera overlays).
// Returns an empty string on success, else the error.
std::string read_container(const std::string &path, Container &out);

// POD mirror of lajolla's Scene built from a container; owns every buffer gdpt_scene_desc points to.
struct FlatScene {
    gdpt_scene_desc desc{};
    int integrator = 7;             // Integrator::GradPath (src/scene.h:14-23)
    std::vector<int32_t> shapes, indices;
    std::vector<int64_t> offsets;
    std::vector<double> spheres, positions, normals, uvs, materials, textures, lights, envmap_to_world, envmap_to_local;
    std::vector<std::vector<double>> image_data, palettes;
    std::vector<std::vector<uint8_t>> image_index;
    std::vector<gdpt_image> images3, images1;
    // Returns an empty string on success, else the error.
    std::string build(const Container &c);
};

// PFM as lajolla's imwrite writes it (src/image.cpp:135-150): "PF", "w h", "-1", float RGB rows top-down.
std::string write_pfm(const std::string &path, int w, int h, const double *rgb);
// OpenEXR as imwrite -> tinyexr SaveEXR(..., 3, fp16) writes it (src/image.cpp:150-171): B, G, R HALF channels, ZIP.
std::string write_exr(const std::string &path, int w, int h, const double *rgb);
// imwrite's dispatch on the file name (.pfm / .exr)
std::string write_image(const std::string &path, int w, int h, const double *rgb);

} // namespace gdpt_host
