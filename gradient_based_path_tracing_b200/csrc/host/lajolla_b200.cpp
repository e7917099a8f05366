// lajolla_b200 — the C++ host of the B200 gradpath renderer: same command line shape and console
// output as `./lajolla [-t n] [-o file] scene.xml` (/root/reference/src/main.cpp:11-51), with the scene
// read from a flat GDPT1 container instead of XML (the XML front end is the reference's; see
// tools/export_scenes.py) and the GradPath arm of render() (src/render.cpp:387-388) going through the
// C ABI of include/gdpt.h.  Writes the Poisson reconstruction to the -o file like the reference,
// plus <stem>_primal / <stem>_dx / <stem>_dy (SURVEY.md 8b "Outputs on disk").
//
//   lajolla_b200 [-o out.pfm] [-overlay camera.gdpt] [-spp N] [-precision f32|f64] [-integrator path|gradpath] [-device D] [-seed S] scene.gdpt[.z] ...
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <string>
#include <vector>

#include "gdpt_container.h"

using namespace gdpt_host;

static double seconds_since(std::chrono::steady_clock::time_point &t) {
    auto now = std::chrono::steady_clock::now();
    double s = std::chrono::duration<double>(now - t).count();
    t = now;
    return s;
}

static unsigned long long fnv(const void *p, size_t n, unsigned long long h = 1469598103934665603ull) {
    const unsigned char *b = (const unsigned char *)p;
    for (size_t i = 0; i < n; i++) h = (h ^ b[i]) * 1099511628211ull;
    return h;
}

// One JSON line: the sizes and content hashes of everything gdpt_create would be handed (tests compare it with the Python host's).
static void describe_scene(const FlatScene &s) {
    const gdpt_scene_desc &d = s.desc;
    unsigned long long himg = 1469598103934665603ull;
    for (int pass = 0; pass < 2; pass++) {
        const gdpt_image *im = pass == 0 ? d.images3 : d.images1;
        for (int i = 0; i < (pass == 0 ? d.n_images3 : d.n_images1); i++) {
            size_t n = (size_t)im[i].width * im[i].height * im[i].channels;
            for (size_t k = 0; k < n; k++) { double v = im[i].data ? im[i].data[k] : im[i].palette[im[i].index[k]]; himg = fnv(&v, 8, himg); }
        }
    }
    std::printf("{\"width\": %d, \"height\": %d, \"filter_kind\": %d, \"filter_param\": %.17g, \"spp\": %d, \"max_depth\": %d, \"rr_depth\": %d, \"integrator\": %d, "
                "\"n_shapes\": %d, \"n_positions\": %lld, \"n_triangles\": %lld, \"n_normals\": %lld, \"n_uvs\": %lld, \"n_materials\": %d, \"n_textures\": %d, "
                "\"n_images3\": %d, \"n_images1\": %d, \"n_lights\": %d, "
                "\"h_camera\": %llu, \"h_shapes\": %llu, \"h_offsets\": %llu, \"h_spheres\": %llu, \"h_positions\": %llu, \"h_indices\": %llu, \"h_normals\": %llu, "
                "\"h_uvs\": %llu, \"h_materials\": %llu, \"h_textures\": %llu, \"h_lights\": %llu, \"h_images\": %llu}\n",
                d.width, d.height, d.filter_kind, d.filter_param, d.samples_per_pixel, d.max_depth, d.rr_depth, s.integrator,
                d.n_shapes, (long long)d.n_positions, (long long)d.n_triangles, (long long)d.n_normals, (long long)d.n_uvs, d.n_materials, d.n_textures,
                d.n_images3, d.n_images1, d.n_lights,
                fnv(d.cam_to_world, 128, fnv(d.sample_to_cam, 128)), fnv(d.shapes, 24 * (size_t)d.n_shapes), fnv(d.shape_offsets, 32 * (size_t)d.n_shapes),
                fnv(d.spheres, 32 * (size_t)d.n_shapes), fnv(d.positions, 24 * (size_t)d.n_positions), fnv(d.indices, 12 * (size_t)d.n_triangles),
                fnv(d.normals, 24 * (size_t)d.n_normals), fnv(d.uvs, 16 * (size_t)d.n_uvs), fnv(d.materials, 112 * (size_t)d.n_materials),
                fnv(d.textures, 112 * (size_t)d.n_textures), fnv(d.lights, 40 * (size_t)d.n_lights), himg);
}

int main(int argc, char *argv[]) {
    if (argc <= 1) {
        std::cout << "[Usage] ./lajolla_b200 [-o output_file_name] [-overlay camera.gdpt] [-spp n] [-precision f32|f64] [-offsets reference|shift] [-reconstruct l2|l1] [-device d] [-gpus n] [-seed s] filename.gdpt[.z]" << std::endl;
        return 0;
    }
    std::string outputfile, overlay, precision = "f32", offsets = "reference", reconstruct = "l2";
    int spp = 0, device = 0, gpus = 1; unsigned long long seed = 0;
    bool describe = false;
    int integrator_override = -1;   // Integrator enum value (src/scene.h:14-23): 5 Path, 7 GradPath
    std::vector<std::string> filenames;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto value = [&]() -> std::string { if (i + 1 >= argc) { std::cerr << "missing value after " << a << std::endl; exit(2); } return argv[++i]; };
        if (a == "-t") value();                                  // accepted for compatibility: the device schedules itself
        else if (a == "-o") outputfile = value();
        else if (a == "-overlay") overlay = value();
        else if (a == "-spp") spp = std::stoi(value());
        else if (a == "-precision") precision = value();
        else if (a == "-offsets") offsets = value();              // gradpath only: the shipped offset semantics, or shift-mapped offsets (gdpt.h)
        else if (a == "-reconstruct") reconstruct = value();      // gradpath only: l2 = fourierSolve (render.cpp:353), l1 = IRLS (gdpt_poisson_l1) written beside it as *_l1
        else if (a == "-device") device = std::stoi(value());
        else if (a == "-gpus") gpus = std::stoi(value());         // gradpath over n GPUs of this box (gdpt_multi_*: one host thread per device)
        else if (a == "-seed") seed = std::stoull(value());
        else if (a == "-integrator") { std::string v = value(); const char *names[] = {"depth", "shadingNormal", "meanCurvature", "rayDifferential", "mipmapLevel", "path", "volpath", "gradpath"};   // parse_scene.cpp:520-566
            for (int k = 0; k < 8; k++) if (v == names[k]) integrator_override = k;
            if (integrator_override < 0) { std::cerr << "unknown integrator " << v << std::endl; return 2; } }
        else if (a == "-describe") describe = true;             // print a digest of the flattened scene and exit (no device needed)
        else filenames.push_back(a);
    }
    if (precision != "f32" && precision != "f64") { std::cerr << "-precision must be f32 or f64" << std::endl; return 2; }
    if (offsets != "reference" && offsets != "shift") { std::cerr << "-offsets must be reference or shift" << std::endl; return 2; }
    if (reconstruct != "l2" && reconstruct != "l1") { std::cerr << "-reconstruct must be l2 or l1" << std::endl; return 2; }
    if (!describe && gdpt_device_count() < 1) { std::cerr << "no CUDA device: lajolla_b200 has no CPU fallback" << std::endl; return 1; }

    for (const std::string &filename : filenames) {
        auto timer = std::chrono::steady_clock::now();
        std::cout << "Parsing and constructing scene " << filename << "." << std::endl;
        Container c;
        std::string err = read_container(filename, c);
        if (err.empty() && !overlay.empty()) err = read_container(overlay, c);
        FlatScene scene;
        if (err.empty()) err = scene.build(c);
        if (!err.empty()) { std::cerr << err << std::endl; return 1; }
        if (integrator_override >= 0) scene.integrator = integrator_override;
        const bool aux = scene.integrator >= 0 && scene.integrator <= 4;   // depth .. mipmapLevel -> aux_render (render.cpp:374-380)
        if (scene.integrator != 7 && scene.integrator != 5 && !aux) {   // render() asserts on integrators it does not know (render.cpp:390-393); VolPath is not built
            std::cerr << "integrator " << scene.integrator << " is outside the gradpath hot path" << std::endl;
            return 1;
        }
        if (describe) { describe_scene(scene); continue; }
        gdpt_ctx *ctx = nullptr;
        gdpt_multi *multi = nullptr;
        const bool use_multi = gpus > 1 && scene.integrator == 7;
        if (gpus > 1 && !use_multi) { std::cerr << "-gpus applies to the gradpath integrator" << std::endl; return 2; }
        if ((use_multi ? gdpt_multi_create(&scene.desc, gpus, nullptr, &multi) : gdpt_create(&scene.desc, device, &ctx)) != GDPT_OK) { std::cerr << gdpt_last_error() << std::endl; return 1; }
        std::cout << "Done. Took " << seconds_since(timer) << " seconds." << std::endl;

        std::cout << "Rendering..." << std::endl;
        const int w = scene.desc.width, h = scene.desc.height;
        std::vector<double> primal((size_t)w * h * 3), dx(primal.size()), dy(primal.size()), out(primal.size());
        gdpt_render_params p{};
        p.spp = spp > 0 ? spp : scene.desc.samples_per_pixel;
        p.max_depth = -2; p.rr_depth = -2; p.seed = seed; p.data_cost = 0.04;
        p.precision = precision == "f64" ? GDPT_PRECISION_F64 : GDPT_PRECISION_F32;
        p.offset_mode = offsets == "shift" ? GDPT_OFFSETS_SHIFT : GDPT_OFFSETS_REFERENCE;
        gdpt_images img{primal.data(), dx.data(), dy.data(), out.data()};
        const bool path = scene.integrator == 5;          // Integrator::Path -> path_render (render.cpp:385-386): one image
        int rc = use_multi ? gdpt_multi_render(multi, &p, &img)
                 : aux ? gdpt_render_aux(ctx, scene.integrator, p.precision, out.data()) : path ? gdpt_render_path(ctx, &p, out.data()) : gdpt_render(ctx, &p, &img);
        gdpt_stats st{};
        if (use_multi) { gdpt_multi_get_stats(multi, &st); gdpt_multi_destroy(multi); }
        else { gdpt_get_stats(ctx, &st); gdpt_destroy(ctx); }
        if (rc != GDPT_OK) { std::cerr << gdpt_last_error() << std::endl; return 1; }
        double took = seconds_since(timer);
        std::cout << "Done. Took " << took << " seconds." << std::endl;
        if (aux) std::printf("auxiliary buffer %d (aux_render)\n", scene.integrator);
        else if (path) std::printf("%.1f Mpaths/s (path, NEE + MIS; %d spp, %lld kernel launches, trace %.1f ms, shade %.1f ms, copies %.1f ms)\n",
                              1.0 * w * h * p.spp / took / 1e6, p.spp, (long long)st.kernel_launches, st.ms_trace, st.ms_shade, st.ms_copy);
        else std::printf("%.1f Mpaths/s (base + 4 shifts; %d spp, %lld kernel launches, trace %.1f ms, shade %.1f ms, Poisson %.1f ms, copies %.1f ms)\n",
                    5.0 * w * h * p.spp / took / 1e6, p.spp, (long long)st.kernel_launches, st.ms_trace, st.ms_shade, st.ms_poisson, st.ms_copy);

        // the film name: -o, else the reference's default image.exr (parse_scene.cpp:15); .exr (fp16) or .pfm like imwrite
        std::string of = outputfile.empty() ? "image.exr" : outputfile;
        auto ends = [&](const char *suf) { size_t n = std::char_traits<char>::length(suf); return of.size() >= n && of.compare(of.size() - n, n, suf) == 0; };
        if (!ends(".pfm") && !ends(".exr")) { std::cerr << "Unsupported image format: " << of << std::endl; return 1; }
        const std::string stem = of.substr(0, of.size() - 4), ext = of.substr(of.size() - 4);
        const std::pair<std::string, const double *> files[] = {{of, out.data()}, {stem + "_primal" + ext, primal.data()}, {stem + "_dx" + ext, dx.data()}, {stem + "_dy" + ext, dy.data()}};
        for (const auto &f : files) {
            if ((path || aux) && f.second != out.data()) continue;
            err = write_image(f.first, w, h, f.second);
            if (!err.empty()) { std::cerr << err << std::endl; return 1; }
        }
        if (reconstruct == "l1" && !path && !aux) {
            std::vector<double> l1(primal.size());
            gdpt_l1_params lp{}; lp.irls_iterations = 10; lp.cg_iterations = 100; lp.start = GDPT_L1_START_FOURIER;
            double info[4] = {0, 0, 0, 0};
            if (gdpt_poisson_l1(device, w, h, primal.data(), dx.data(), dy.data(), &lp, l1.data(), info) != GDPT_OK) { std::cerr << gdpt_last_error() << std::endl; return 1; }
            err = write_image(stem + "_l1" + ext, w, h, l1.data());
            if (!err.empty()) { std::cerr << err << std::endl; return 1; }
            std::printf("L1 reconstruction: %d reweightings, %d conjugate-gradient steps, %.1f ms on the device -> %s\n", lp.irls_iterations, (int)info[1], info[2], (stem + "_l1" + ext).c_str());
        }
        std::cout << "Image written to " << of << std::endl;
    }
    return 0;
}
