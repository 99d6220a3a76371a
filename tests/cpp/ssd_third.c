/* Identity behind search_ssd's first term (csrc/dvo_kernels.cuh): for a non-negative finite float a,
 *   (float)(0.0 + (double)(1.0 * 1 / 3) * (double)a)      -- implement.cpp:134 as the reference evaluates it
 * equals Markstein's two-FMA refinement of a * RN(1/3), i.e. the correctly rounded a / 3.
 * usage: ssd_third <stride>   (stride 1 = all 2^31 - 2^23 bit patterns, ~20 s on one core); prints the mismatch count. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char** argv)
{
    const long long stride = argc > 1 ? atoll(argv[1]) : 1;
    const float y = 0.3333333432674407958984375f;
    long long bad = 0, n = 0;
    for (long long i = 0; i < 0x7F800000LL; i += stride, n++) {
        const uint32_t u = (uint32_t)i;
        float a;
        memcpy(&a, &u, 4);
        volatile double prod = (1.0 * 1 / 3) * (double)a;
        const float ref = (float)(0.0 + prod);
        const float q = a * y;
        const float got = fmaf(fmaf(-3.0f, q, a), y, q);
        if (memcmp(&ref, &got, 4) != 0) bad++;
    }
    printf("%lld %lld\n", n, bad);
    return bad != 0;
}
