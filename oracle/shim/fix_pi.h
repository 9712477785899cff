// TEST INFRASTRUCTURE (oracle/_ref build only) -- not part of the product.
// Pre-included when compiling the untouched reference sources for sm_100.
// glibc defines M_PIf32 as a _Float32 literal, which nvcc 12.9 cannot lower
// ("unsupported float variant"); the reference uses it at SiftOps.cu:318.
// Same float value, plain float literal.
#pragma once
#include <cmath>
#undef M_PIf32
#define M_PIf32 3.14159265358979323846f
