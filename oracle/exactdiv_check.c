/* oracle/exactdiv_check.c -- TEST INFRASTRUCTURE ONLY.
 * Host-side exhaustive proof obligations of csrc/exactdiv.cuh (same FMA sequences, evaluated with fmaf):
 *   mode 0: one-step sequence, divisors 1..600 and 32767, numerators = EVERY significand of one binade
 *           (the sequence is invariant under scaling the numerator by 2^k while nothing underflows);
 *   mode 1: quick subset of mode 0 (every 257th significand) for the default CPU test run;
 * prints the number of quotients that differ from IEEE `/` (must be 0). Build: gcc -O2 -ffp-contract=off -fopenmp. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline float step1(float a, float b, float y) { float q = a * y; float r = fmaf(-b, q, a); return fmaf(r, y, q); }
static inline float step2(float a, float b, float y) { float q = a * y; float r = fmaf(-b, q, a); q = fmaf(r, y, q); r = fmaf(-b, q, a); return fmaf(r, y, q); }
static inline int differs(float x, float y) { return memcmp(&x, &y, 4) != 0; }

int main(int argc, char **argv)
{
    int mode = argc > 1 ? atoi(argv[1]) : 1;
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
