/* Test helper (not part of the product): the three-instruction division by a compile-time count that the AGG block
 * mean uses on the GPU (csrc/srnn_device.cuh, div_count<N>), restated with the C library's fma, against the IEEE
 * division the oracle and the reference use.  Prints "checked <n> mismatches <m>". */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t state = 88172645463325252ULL;
static uint64_t rnd(void) { state ^= state << 13; state ^= state >> 7; state ^= state << 17; return state; }

static double div_count(double a, double n) {
    const double y = 1.0 / n;
    const double q0 = a * y;
    const double r = fma(-n, q0, a);
    const double q = fma(r, y, q0);
    return isinf(q0) ? q0 : q;
}

int main(int argc, char **argv) {
    const long per_n = argc > 1 ? atol(argv[1]) : 200000;
    long bad = 0, tot = 0;
    for (int n = 1; n <= 64; ++n) {
        for (long it = 0; it < per_n; ++it) {
            double a;
            const int mode = (int)(it & 3);
            if (mode == 0) {            /* any double far enough from the subnormal range */
                const uint64_t r = rnd();
                memcpy(&a, &r, 8);
                if (isnan(a) || (fabs(a) < 1e-290 && a != 0.0)) continue;
            } else {                    /* what the kernel sees: a sequential double sum of float32 values */
                const int cnt = mode == 1 ? n : 5;
                double t = 0.0;
                for (int k = 0; k < cnt; ++k) {
                    uint32_t u = (uint32_t)rnd();
                    float f;
                    if (mode == 3) u = (u & 0x807fffffu) | ((120u + (u >> 23) % 16u) << 23);   /* near-cancellation */
                    memcpy(&f, &u, 4);
                    if (isnan(f)) f = 1.0f;
                    if (fabsf(f) < 1.17549435e-38f) f = 0.0f;   /* stored weights are flushed results */
                    t += (double)f;
                }
                a = t;
            }
            const double q = div_count(a, (double)n), ref = a / (double)n;
            ++tot;
            if (memcmp(&q, &ref, 8) != 0 && !(isnan(q) && isnan(ref))) {
                if (bad < 10) printf("n=%d a=%a got=%a want=%a\n", n, a, q, ref);
                ++bad;
            }
        }
    }
    printf("checked %ld mismatches %ld\n", tot, bad);
    return bad != 0;
}
