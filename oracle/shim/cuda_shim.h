// cuda_shim.h — TEST INFRASTRUCTURE. Lets the unmodified reference headers
// under /root/reference/CudaTest/src compile as host C++.
// It defines the CUDA decorators away and supplies a curandState stand-in
// whose curand_uniform() is the build's counter-based generator
// (include/crt_rng.h), or a caller-injected sequence for scatter vectors.
#pragma once
#ifndef __host__
#define __host__
#endif
#ifndef __device__
#define __device__
#endif
#ifndef __global__
#define __global__
#endif
#ifndef _USE_MATH_DEFINES
#define _USE_MATH_DEFINES
#endif
#include <math.h>
#include <cmath>
#include <cfloat>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <algorithm>
using std::min;
using std::max;

#include "crt_rng.h"

struct curandState {
    uint64_t key = 0;
    uint32_t ctr = 0;
    const float* inject = nullptr;  // when set, values are served from here
    int inject_n = 0;
    int inject_pos = 0;
};

// (0, 1] like the CUDA device API the reference calls.
inline float curand_uniform(curandState* s)
{
    if (s->inject) {
        float v = s->inject[s->inject_pos % s->inject_n];
        s->inject_pos++;
        return v;
    }
    return crt_rng_uniform(s->key, s->ctr++);
}
