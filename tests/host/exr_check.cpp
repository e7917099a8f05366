// Test driver of the C++ host's image writers: exr_check W H in.f64 out.{exr,pfm}
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "gdpt_container.h"
int main(int argc, char **argv) {
    if (argc != 5) return 2;
    int w = atoi(argv[1]), h = atoi(argv[2]);
    std::vector<double> img((size_t)w * h * 3);
    FILE *f = fopen(argv[3], "rb");
    if (!f || fread(img.data(), 8, img.size(), f) != img.size()) return 3;
    fclose(f);
    std::string err = gdpt_host::write_image(argv[4], w, h, img.data());
    if (!err.empty()) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    return 0;
}
