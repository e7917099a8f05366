/* TEST INFRASTRUCTURE — CPU oracle, not part of the product.
 * Restatement of PCG32 XSH-RR as used by lajolla: /root/reference/src/pcg.h:16-68
 * (init_pcg32 :33-41, next_pcg32 :22-30, next_pcg32_real<double> :60-68).
 * Known-answer vectors: the pcg32 reference generator (wjakob/pcg32 = O'Neill's
 * pcg32_srandom_r/pcg32_random_r) — checked in tests/test_oracle_pcg.py.
 */
#include <stdint.h>

typedef struct { uint64_t state, inc; } gdpt_pcg32;

uint32_t gdpt_oracle_pcg_next(gdpt_pcg32 *r) {
    uint64_t old = r->state;
    r->state = old * 6364136223846793005ULL + (r->inc | 1);
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((-rot) & 31));
}

void gdpt_oracle_pcg_init(gdpt_pcg32 *r, uint64_t stream, uint64_t seed) {
    r->state = 0u; r->inc = (stream << 1u) | 1u;
    gdpt_oracle_pcg_next(r);
    r->state += seed;
    gdpt_oracle_pcg_next(r);
}

double gdpt_oracle_pcg_real(gdpt_pcg32 *r) {
    union { uint64_t u; double d; } x;
    x.u = ((uint64_t)gdpt_oracle_pcg_next(r) << 20) | 0x3ff0000000000000ULL;
    return x.d - 1.0;
}

/* fill out[n] with the first n doubles of stream `stream` */
void gdpt_oracle_pcg_fill(uint64_t stream, uint64_t seed, int n, double *out, uint32_t *raw) {
    gdpt_pcg32 r; gdpt_oracle_pcg_init(&r, stream, seed);
    for (int i = 0; i < n; i++) {
        gdpt_pcg32 c = r;
        if (raw) raw[i] = gdpt_oracle_pcg_next(&c);
        out[i] = gdpt_oracle_pcg_real(&r);
    }
}
