// TEST INFRASTRUCTURE — not part of the product.
// Runs the reference's second program, /root/reference/small_gdpt.cpp (its own sphere scene, erand48 streams, shift
// mapping, MIS, screened-Poisson solve), UNMODIFIED at a film size given on the command line: the source is included
// from where it lies (the recipe passes -I<reference dir>), its main() renamed by the preprocessor; width / height and
// the two camera vectors derived from them (small_gdpt.cpp:55-57) are re-assigned before it runs.  FFTW is replaced by
// oracle/ref_shims/fftw_standin.c + oracle/dct1.c, as for the lajolla reference (oracle/build_ref.sh).
//   smallgdpt_ref W H SPP OUTDIR   -> OUTDIR/image.pfm, image_dx.pfm, image_dy.pfm, image_poisson.pfm (small_gdpt.cpp:546-561)
#include <unistd.h>
#define main smallgdpt_reference_main
#include "small_gdpt.cpp"
#undef main

int main(int argc, char **argv) {
    if (argc != 5) { fprintf(stderr, "usage: %s W H SPP OUTDIR\n", argv[0]); return 2; }
    width = atoi(argv[1]); height = atoi(argv[2]);
    if (width < 2 || height < 2 || width > 65535) { fprintf(stderr, "bad film size\n"); return 2; }
    cx = Vec(width * .5135 / height);
    cy = (cx % cam.d).norm() * .5135;
    if (chdir(argv[4]) != 0) { perror("chdir"); return 2; }
    char *args[] = {argv[0], argv[3], nullptr};
    return smallgdpt_reference_main(2, args);
}
