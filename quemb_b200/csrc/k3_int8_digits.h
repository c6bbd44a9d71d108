// k3_int8_digits.h -- digit extraction of the int8-slice K3, shared by the device slicer (k3_int8.cu) and the host
// entry point qemb_k3_int8_slice_host (so the CPU suite can check the very code the kernel runs against the numpy
// restatement, oracle/k3_int8_oracle.py).
//
//   r = x 2^-e, |r| < 1;   N = nearest integer of r 2^(8S-2) (ties to even), |N| <= 2^(8S-2) < 2^63
//   N = sum_{a<S} D_a 256^(S-1-a),  D_a in [-128, 127] (balanced base-256 digits, taken from the low end),
//   so  x = 2^e ( sum_a D_a 2^-(6+8a) + R ),  |R| <= 2^-(8S-1),  |D_0| <= 65.
// Rounding to nearest and balanced digits keep remainder and dropped digit products free of a preferred sign.
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define QEMB_HD __host__ __device__
#else
#define QEMB_HD
#endif

namespace qemb {

template <int S>
QEMB_HD inline void i8_digits(double x, int e, int8_t *out, int64_t stride)
{
    static_assert(S >= 1 && S <= 8, "N has to fit 63 bits");
    long long N = llrint(ldexp(x, 8 * S - 2 - e));
#pragma unroll
    for (int a = S - 1; a >= 0; --a) {
        const int d = (int)((N + 128) & 255) - 128; // N = d mod 256, d in [-128, 127]
        out[a * stride] = (int8_t)d;
        N = (N - d) / 256; // exact
    }
}

} // namespace qemb
