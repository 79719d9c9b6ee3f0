// Host-side set-up of the sample streams of wf_sampler.cuh.
//
//   HaltonSampler::HaltonSampler (src/samplers/halton.cpp:64-93): base scales / exponents covering min(res, 128), the
//   sample stride and the two multiplicative inverses;
//   ComputeRadicalInversePermutations (src/core/lowdiscrepancy.cpp:2490-2504): one random digit permutation per prime
//   base, drawn with Shuffle (src/core/sampling.h:151-157) from a default-constructed RNG (PCG32, src/core/rng.h) — the
//   table is a process-wide static in the reference (halton.cpp:67-72,95), i.e. the same for every render.
#include <mutex>
#include <utility>
#include <vector>
#include "wf_sampler_host.h"

namespace b200rt {

namespace {
// PCG32 as src/core/rng.h:52-146 (default state / stream)
struct Pcg32 {
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    uint32_t next() {
        const uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        const uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    uint32_t below(uint32_t b) {   // RNG::UniformUInt32(b)
        const uint32_t threshold = (~b + 1u) % b;
        for (;;) {
            const uint32_t r = next();
            if (r >= threshold) return r % b;
        }
    }
};

int64_t mod64(int64_t a, int64_t b) { const int64_t r = a - (a / b) * b; return r < 0 ? r + b : r; }
void extended_gcd(uint64_t a, uint64_t b, int64_t *x, int64_t *y) {
    if (b == 0) { *x = 1; *y = 0; return; }
    const int64_t d = (int64_t)(a / b);
    int64_t xp, yp;
    extended_gcd(b, a % b, &xp, &yp);
    *x = yp;
    *y = xp - d * yp;
}
uint64_t multiplicative_inverse(int64_t a, int64_t n) {
    int64_t x, y;
    extended_gcd((uint64_t)a, (uint64_t)n, &x, &y);
    return (uint64_t)mod64(x, n);
}
}  // namespace

const HaltonTables &halton_tables() {
    static HaltonTables T;
    static std::once_flag once;
    std::call_once(once, [] {
        // the first 1000 primes (the reference's Primes[] table, lowdiscrepancy.cpp:40-; PrimeTableSize = 1000)
        std::vector<uint32_t> primes;
        const uint32_t limit = 8000;   // the 1000th prime is 7919
        std::vector<uint8_t> composite(limit + 1, 0);
        for (uint32_t n = 2; n <= limit && primes.size() < (size_t)kHaltonDims; ++n) {
            if (composite[n]) continue;
            primes.push_back(n);
            for (uint32_t m = n * n; m <= limit; m += n) composite[m] = 1;
        }
        size_t total = 0;
        T.dims.resize(primes.size());
        for (size_t i = 0; i < primes.size(); ++i) {
            T.dims[i].prime = primes[i];
            T.dims[i].offset = (uint32_t)total;
            total += primes[i];
            // always-add invariant division (see div_magic): l = floor(log2 d), m = floor(2^(33 + l) / d) + 1 (low 32 bits).
            // d = 2 is the one power of two among the primes: m = 2^32 does not fit, and dimension 0 never takes this path
            // (HaltonSampler::SampleDimension reverses bits for base 2); give it a working pair anyway: a / 2 = (((a - 0) >> 1) + 0) >> 0
            const uint32_t d = primes[i];
            uint32_t l = 0;
            while ((2u << l) <= d) ++l;
            if (d == 2) { T.dims[i].magic = 0; T.dims[i].shift = 0; continue; }
            const unsigned __int128 num = (unsigned __int128)1 << (33 + l);
            T.dims[i].magic = (uint32_t)(uint64_t)(num / d + 1);
            T.dims[i].shift = l;
        }
        T.perms.resize(total);
        Pcg32 rng;
        uint16_t *p = T.perms.data();
        for (size_t i = 0; i < primes.size(); ++i) {
            const int count = (int)primes[i];
            for (int j = 0; j < count; ++j) p[j] = (uint16_t)j;
            for (int j = 0; j < count; ++j) {   // Shuffle(p, count, 1, rng)
                const int other = j + (int)rng.below((uint32_t)(count - j));
                std::swap(p[j], p[other]);
            }
            p += count;
        }
    });
    return T;
}

void halton_setup(int resx, int resy, WfSampler *S) {
    const int kMaxResolution = 128;
    const int res[2] = {resx, resy};
    int scales[2], exps[2];
    for (int i = 0; i < 2; ++i) {
        const int base = i == 0 ? 2 : 3;
        int scale = 1, e = 0;
        while (scale < (res[i] < kMaxResolution ? res[i] : kMaxResolution)) { scale *= base; ++e; }
        scales[i] = scale; exps[i] = e;
    }
    S->baseScale0 = (uint32_t)scales[0]; S->baseScale1 = (uint32_t)scales[1];
    S->baseExp0 = (uint32_t)exps[0]; S->baseExp1 = (uint32_t)exps[1];
    S->stride = (uint32_t)(scales[0] * scales[1]);
    S->multInv0 = (uint32_t)multiplicative_inverse(scales[1], scales[0]);
    S->multInv1 = (uint32_t)multiplicative_inverse(scales[0], scales[1]);
}

void sobol_setup(int resx, int resy, int sbMinX, int sbMinY, WfSampler *S) {
    // SobolSampler::SobolSampler (samplers/sobol.h): resolution = RoundUpPow2(max(extent)), log2Resolution = Log2Int
    int m = resx > resy ? resx : resy;
    uint32_t r = 1, l = 0;
    while ((int)r < m) { r <<= 1; ++l; }
    S->res = r; S->log2Res = l; S->sbMinX = sbMinX; S->sbMinY = sbMinY;
}

}  // namespace b200rt
