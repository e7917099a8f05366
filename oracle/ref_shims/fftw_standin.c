/* TEST INFRASTRUCTURE — not part of the product.
 * Stand-in for the 5 FFTW entry points the reference's fourierSolve uses
 * (/root/reference/src/render.cpp:178-192,227,242,250-253) so the unmodified
 * reference links without FFTW binaries.  Only the one plan shape the reference
 * creates is supported: in-place 2-D REDFT00 x REDFT00.
 */
#include <stdlib.h>
#include <fftw3.h>

void gdpt_oracle_dct1_2d(double *buf, int h, int w);

struct fftw_plan_s { int n0, n1; double *buf; };

void *fftw_malloc(size_t n) { return malloc(n); }
void fftw_free(void *p) { free(p); }

fftw_plan fftw_plan_r2r_2d(int n0, int n1, double *in, double *out, fftw_r2r_kind k0, fftw_r2r_kind k1, unsigned flags) {
    (void)flags;
    if (in != out || k0 != FFTW_REDFT00 || k1 != FFTW_REDFT00) abort();
    struct fftw_plan_s *p = (struct fftw_plan_s *)malloc(sizeof(*p));
    p->n0 = n0; p->n1 = n1; p->buf = in;
    return (fftw_plan)p;
}
void fftw_execute(const fftw_plan p) { gdpt_oracle_dct1_2d(p->buf, p->n0, p->n1); }
void fftw_destroy_plan(fftw_plan p) { free(p); }
