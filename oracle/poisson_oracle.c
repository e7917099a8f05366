/* TEST INFRASTRUCTURE — CPU oracle, not part of the product.
 *
 * Restatement of the reference's screened-Poisson reconstruction
 *   fourierSolve(width, height, imgData, imgGradX, imgGradY, dataCost, imgOut)
 *   /root/reference/src/render.cpp:172-254   (identical copy: small_gdpt.cpp:355-436)
 * Inputs/outputs: interleaved-RGB double arrays, row-major, W*H*3 each.
 * Per channel:  rhs = a*u - Dx - Dy  (render.cpp:202-225), 2-D DCT-I (:227),
 * divide by (a - (float)(lapY[y]+lapX[x])) (:231-235, note the float narrowing),
 * DC := sum m(x) m(y) u (:205-211,239), 2-D DCT-I (:242), scale 1/(4(W-1)(H-1)) (:243-247).
 */
#include <math.h>
#include <stdlib.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

void gdpt_oracle_dct1_2d(double *buf, int h, int w);

void gdpt_oracle_poisson(int width, int height, const double *img, const double *gx, const double *gy,
                         double data_cost, double *out) {
    size_t n = (size_t)width * height;
    double *buf = (double *)malloc(sizeof(double) * n);
    double *lapx = (double *)malloc(sizeof(double) * width);
    double *lapy = (double *)malloc(sizeof(double) * height);
    for (int x = 0; x < width; x++) lapx[x] = 2.0 * cos(M_PI * x / (width - 1));
    for (int y = 0; y < height; y++) lapy[y] = -4.0 + (2.0 * cos(M_PI * y / (height - 1)));
    for (int ch = 0; ch < 3; ch++) {
        double dc = 0.0;
        for (int y = 0; y < height; y++)
            for (int x = 0; x < width; x++) {
                size_t node = (size_t)y * width + x;
                size_t px = node * 3 + ch;
                int ix = (x > 0) && (x < width - 1), iy = (y > 0) && (y < height - 1);
                double m = 1.0;
                if (ix) m *= 2.0;
                if (iy) m *= 2.0;
                dc += m * img[px];
                double v = data_cost * img[px];
                if (ix) v -= (gx[px + 3] - gx[px]); else v -= (-2.0 * gx[px]);
                if (iy) v -= (gy[px + (size_t)width * 3] - gy[px]); else v -= (-2.0 * gy[px]);
                buf[node] = v;
            }
        gdpt_oracle_dct1_2d(buf, height, width);
        for (int y = 0; y < height; y++)
            for (int x = 0; x < width; x++) {
                float resp = (float)(lapy[y] + lapx[x]);
                buf[(size_t)y * width + x] /= (data_cost - resp);
            }
        buf[0] = dc;
        gdpt_oracle_dct1_2d(buf, height, width);
        double denom = 4.0 * (width - 1) * (height - 1);
        for (size_t i = 0; i < n; i++) out[i * 3 + ch] = buf[i] / denom;
    }
    free(buf); free(lapx); free(lapy);
}

/* gradient merge of gradient_path_render, /root/reference/src/render.cpp:340-350 */
void gdpt_oracle_merge(int w, int h, const double *cx0, const double *cx1, const double *cy0, const double *cy1,
                       double *cx, double *cy) {
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++)
            for (int c = 0; c < 3; c++) {
                size_t i = ((size_t)y * w + x) * 3 + c;
                cx[i] = x == 0 ? cx0[i] : cx0[i] + cx1[i - 3];
                cy[i] = y == 0 ? cy0[i] : cy0[i] + cy1[i - (size_t)w * 3];
            }
}
