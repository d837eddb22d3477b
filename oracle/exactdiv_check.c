/* oracle/exactdiv_check.c -- TEST INFRASTRUCTURE ONLY.
 * Host-side exhaustive proof obligations of csrc/exactdiv.cuh (same FMA sequences, evaluated with fmaf):
 *   mode 0: one-step sequence, divisors 1..600 and 32767, numerators = EVERY significand of one binade
 *           (the sequence is invariant under scaling the numerator by 2^k while nothing underflows);
 *   mode 1: quick subset of mode 0 (every 257th significand) for the default CPU test run;
 *   mode 2: the claim behind dropping the LOWER operand window of the run-time-divisor path (capi.cu: make_fuse_params): for a
 *           numerator below 2^-84 (denormals included), a divisor in [2^-20, 2^20] and a reciprocal as inexact as
 *           rcp.approx + one Newton step can be, the two-step sequence may differ from IEEE `/`, but `q + principal point`
 *           -- the only use of the quotient -- is bit-identical for every principal point >= 2^-10, so a pixel
 *           coordinate, the `u >= 1` image test and everything after them cannot tell the difference;
 * prints the number of quotients that differ from IEEE `/` (must be 0). Build: gcc -O2 -ffp-contract=off -fopenmp. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline float step1(float a, float b, float y) { float q = a * y; float r = fmaf(-b, q, a); return fmaf(r, y, q); }
static inline float step2(float a, float b, float y) { float q = a * y; float r = fmaf(-b, q, a); q = fmaf(r, y, q); r = fmaf(-b, q, a); return fmaf(r, y, q); }
static inline int differs(float x, float y) { return memcmp(&x, &y, 4) != 0; }

static uint32_t lcg(uint32_t *s) { *s = *s * 1664525u + 1013904223u; return *s; }

static int below_window(void)
{
    const float points[] = {0.0009765625f, 0.37f, 0.5f, 0.99999994f, 1.0f, 1.5f, 2.0f, 255.314f, 319.5f, 609.5593f, 4095.5f, -3.25f};
    long bad = 0, inexact = 0, n = 0;
    uint32_t s = 12345u;
    for (long it = 0; it < 20000000; it++) {
        /* numerator: sign, exponent field 0 (denormal) .. 43 (2^-84), any significand */
        uint32_t ua = (lcg(&s) & 0x80000000u) | ((lcg(&s) % 44u) << 23) | (lcg(&s) & 0x7fffffu);
        uint32_t ub = ((107u + lcg(&s) % 41u) << 23) | (lcg(&s) & 0x7fffffu);          /* 2^-20 .. 2^20, positive (a depth) */
        float a, b; memcpy(&a, &ua, 4); memcpy(&b, &ub, 4);
        /* a reciprocal with up to ~2 ulp of error in either direction, then the device's Newton step */
        float y = 1.0f / b; uint32_t uy; memcpy(&uy, &y, 4); uy += (lcg(&s) % 5u) - 2u; memcpy(&y, &uy, 4);
        y = fmaf(y, fmaf(-b, y, 1.0f), y);
        float q = step2(a, b, y), t = a / b;
        inexact += differs(q, t);
        for (unsigned k = 0; k < sizeof(points) / sizeof(points[0]); k++) bad += differs(q + points[k], t + points[k]);
        n++;
    }
    printf("%ld %ld %ld\n", bad, inexact, n);
    return bad ? 1 : 0;
}

int main(int argc, char **argv)
{
    int mode = argc > 1 ? atoi(argv[1]) : 1;
    if (mode == 2) return below_window();
    uint32_t stride = mode == 0 ? 1u : 257u;
    long bad1 = 0, bad2 = 0, n = 0;
    int divisors[602], nd = 0;
    for (int w = 1; w <= 600; w++) divisors[nd++] = w;
    divisors[nd++] = 32767;
#pragma omp parallel for reduction(+ : bad1, bad2, n) schedule(dynamic)
    for (int k = 0; k < nd; k++) {
        float b = (float)divisors[k], y = 1.0f / b;
        for (uint32_t m = 0; m < (1u << 23); m += stride) {
            uint32_t u = 0x3f800000u | m; float a; memcpy(&a, &u, 4);
            float t = a / b;
            bad1 += differs(step1(a, b, y), t); bad2 += differs(step2(a, b, y), t);
            bad1 += differs(step1(-a, b, y), -t);
            n += 2;
        }
    }
    /* every int16 / 32767 and every byte / 255 */
    for (int a = -32768; a <= 32767; a++) bad1 += differs(step1((float)a, 32767.0f, 1.0f / 32767.0f), (float)a / 32767.0f);
    for (int a = 0; a < 256; a++) bad1 += differs(step1((float)a, 255.0f, 1.0f / 255.0f), (float)a / 255.0f);
    printf("%ld %ld %ld\n", bad1, bad2, n);
    return (bad1 || bad2) ? 1 : 0;
}
