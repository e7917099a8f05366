// smallgdpt_b200 — the reference's second program (/root/reference/small_gdpt.cpp) as a host over the C ABI: same command
// line (`./smallgdpt [number of samples per pixel]`, small_gdpt.cpp:7,442), same four output files in the working directory
// (image.pfm, image_dx.pfm, image_dy.pfm, image_poisson.pfm: float32 "PF" files with scale -1, gradients as absolute values,
// :543-561), the film of :55 unless -size is given.  All arithmetic runs in libgdpt.so (gdpt_smallgdpt_render); no CPU path.
//   smallgdpt_b200 [spp] [-size W H] [-rng rows|pixels] [-device d]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "gdpt.h"

static bool write_pfm(const char *name, int w, int h, const std::vector<float> &px) {
    FILE *f = fopen(name, "w");
    if (!f) return false;
    fprintf(f, "PF\n%d %d\n%d\n", w, h, -1);
    const bool ok = fwrite(px.data(), sizeof(float), px.size(), f) == px.size();
    return fclose(f) == 0 && ok;
}

int main(int argc, char *argv[]) {
    int samps = 100, w = 1024, h = 768, device = 0, mode = GDPT_SMALLGDPT_ROWS;     // small_gdpt.cpp:442, :55
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if (a == "-size" && i + 2 < argc) { w = atoi(argv[++i]); h = atoi(argv[++i]); }
        else if (a == "-rng" && i + 1 < argc) { const std::string v = argv[++i]; if (v == "rows") mode = GDPT_SMALLGDPT_ROWS; else if (v == "pixels") mode = GDPT_SMALLGDPT_PIXELS; else { fprintf(stderr, "-rng must be rows or pixels\n"); return 2; } }
        else if (a == "-device" && i + 1 < argc) device = atoi(argv[++i]);
        else if (!a.empty() && a[0] != '-') samps = atoi(a.c_str());
        else { fprintf(stderr, "usage: %s [number of samples per pixel] [-size W H] [-rng rows|pixels] [-device d]\n", argv[0]); return 2; }
    }
    if (gdpt_device_count() < 1) { fprintf(stderr, "no CUDA device: smallgdpt_b200 has no CPU fallback\n"); return 1; }
    const size_t n = (size_t)(w > 0 ? w : 0) * (size_t)(h > 0 ? h : 0) * 3;
    std::vector<double> c(n), cx(n), cy(n), out(n);
    gdpt_images img{c.data(), cx.data(), cy.data(), out.data()};
    double ms = 0.0;
    fprintf(stderr, "Rendering (%d spp)\n", samps);
    if (gdpt_smallgdpt_render(device, w, h, samps, mode, &img, &ms) != GDPT_OK) { fprintf(stderr, "%s\n", gdpt_last_error()); return 1; }
    fprintf(stderr, "%.1f ms on the device, %.1f Msamples/s\n", ms, ms > 0 ? (double)w * h * samps / ms / 1e3 : 0.0);
    std::vector<float> fc(n), fout(n), fcx(n), fcy(n);                                // pfm requires single precision (:545)
    for (size_t i = 0; i < n; i++) { fc[i] = (float)c[i]; fout[i] = (float)out[i]; fcx[i] = (float)std::fabs(cx[i]); fcy[i] = (float)std::fabs(cy[i]); }
    if (!write_pfm("image.pfm", w, h, fc) || !write_pfm("image_dx.pfm", w, h, fcx) || !write_pfm("image_dy.pfm", w, h, fcy) || !write_pfm("image_poisson.pfm", w, h, fout)) {
        perror("writing the PFM files"); return 1;
    }
    return 0;
}
