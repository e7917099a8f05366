// Host compilation of csrc/gdpt_smallgdpt.cuh (the device functions of the smallgdpt restatement, one emulated thread per
// row): writes the five accumulation images as raw doubles so that tests/test_smallgdpt_cpu.py can compare them, after the
// reference's merge and solve, with the outputs of the compiled reference program.
//   smallgdpt_check W H SPP OUT.bin      -> c, cx0, cy0, cx1, cy1, each [H][W][3] doubles
//   smallgdpt_check erand Y N            -> the first N numbers of row Y's stream, as hex floats
#define GDPT_SMALLGDPT_HOST 1
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "gdpt_smallgdpt.cuh"

int main(int argc, char **argv) {
    if (argc == 4 && std::string(argv[1]) == "erand") {
        uint64_t rng = gdpt::sg::row_seed(atoi(argv[2]));
        for (int i = 0, n = atoi(argv[3]); i < n; i++) printf("%a\n", gdpt::sg::erand(rng));
        return 0;
    }
    if (argc != 5) { fprintf(stderr, "usage: %s W H SPP OUT.bin\n", argv[0]); return 2; }
    using namespace gdpt::sg;
    const int w = atoi(argv[1]), h = atoi(argv[2]), spp = atoi(argv[3]);
    const Scene sc = make_scene(w, h);
    const size_t n = (size_t)w * h * 3;
    std::vector<double> img(5 * n, 0.0);
    Path *path = new Path, *off = new Path;
    for (int y = 0; y < h; y++) {
        uint64_t rng = row_seed(y);
        for (int x = 0; x < w; x++)
            render_pixel(sc, x, y, spp, rng, *path, *off, &img[0], &img[n], &img[2 * n], &img[3 * n], &img[4 * n]);
    }
    FILE *f = fopen(argv[4], "wb");
    if (!f || fwrite(img.data(), sizeof(double), img.size(), f) != img.size()) { perror("write"); return 1; }
    fclose(f);
    return 0;
}
