// The reference's sample streams, as pure functions of (pixel, sample number, dimension) — which is what a GlobalSampler
// is (src/core/sampler.h:108-128, sampler.cpp:137-195): value = SampleDimension(GetIndexForSample(sampleNum), dimension).
// A wavefront path therefore draws exactly the reference's numbers in any execution order, on any rank.
//
//   Sampler "halton"  src/samplers/halton.cpp:64-127; RadicalInverse / ScrambledRadicalInverse src/core/lowdiscrepancy.cpp:389-424,
//                     the digit permutations of ComputeRadicalInversePermutations (:2490-2504) built on the host by
//                     wf_sampler_host.cpp from the reference's default RNG (PCG32, rng.h)
//   Sampler "sobol"   src/samplers/sobol.cpp:43-60; SobolIntervalToIndex / SobolSampleFloat src/core/lowdiscrepancy.h:229-272 over
//                     the generator matrices the HOST hands in (SobolMatrices32, VdCSobolMatrices[Inv], sobolmatrices.cpp)
//   kind 0            the round-1 counter hash (order independent too, but no reference stream)
#pragma once
#include "wf_core.cuh"

namespace b200rt {

enum { WF_SAMPLER_HASH = 0, WF_SAMPLER_HALTON = 1, WF_SAMPLER_SOBOL = 2 };
constexpr int kSobolMatrixSize = 52;     // sobolmatrices.h:48
constexpr int kSobolDims = 1024;         // sobolmatrices.h:47
constexpr int kHaltonDims = 1000;        // PrimeTableSize (lowdiscrepancy.h)

// magic / shift: floor(a / prime) for every 32-bit a as ((a - q) >> 1) + q) >> shift with q = mulhi(a, magic) — the
// divisor-invariant multiplication of Granlund & Montgomery in its always-add form; integer results, so the digits (and every
// sample value) are the ones `a / prime` gives.  A runtime division costs ~20 instructions on the SM and a path vertex takes
// up to 8 dimensions x ~6 digits of them.
struct HaltonDim { uint32_t prime, offset, magic, shift; };
WF_HD uint32_t div_magic(uint32_t a, uint32_t magic, uint32_t shift) {
#ifdef __CUDA_ARCH__
    const uint32_t q = __umulhi(a, magic);
#else
    const uint32_t q = (uint32_t)(((uint64_t)a * magic) >> 32);
#endif
    return (((a - q) >> 1) + q) >> shift;
}

WF_HD uint32_t lowest_bit32(uint32_t a) {   // index of the lowest set bit, a != 0
#ifdef __CUDA_ARCH__
    return (uint32_t)(__ffs((int)a) - 1);
#else
    return (uint32_t)__builtin_ctz(a);
#endif
}
WF_HD int lowest_bit(uint64_t a) {   // index of the lowest set bit, a != 0
#ifdef __CUDA_ARCH__
    return __ffsll((long long)a) - 1;
#else
    return __builtin_ctzll(a);
#endif
}

struct WfSampler {
    uint32_t kind, seed;
    // halton
    const uint16_t *perms;       // radicalInversePermutations
    const HaltonDim *dimInfo;    // per dimension: (prime, offset of its permutation = PrimeSums[dim])
    uint32_t baseScale0, baseScale1, baseExp0, baseExp1, stride, multInv0, multInv1, nDims;
    // sobol
    const uint32_t *sobolBytes;  // device only (k_sobol_bytes), else nullptr: kSobolDims x 4 x 256 byte-wise XOR tables of sobol32
    const uint32_t *sobol32;     // kSobolDims x 52
    const uint64_t *vdc, *vdcInv;   // row log2Res - 1 of VdCSobolMatrices / VdCSobolMatricesInv (52 entries each)
    uint32_t log2Res, res;
    int32_t sbMinX, sbMinY;      // sampleBounds.pMin
};

WF_HD uint32_t hash_mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
WF_HD float hash_sample(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t dim) {
    uint32_t h = hash_mix32(seed ^ 0x9e3779b9U);
    h = hash_mix32(h ^ pixel);
    h = hash_mix32(h + sample * 0x85ebca6bU);
    h = hash_mix32(h ^ (dim * 0xc2b2ae35U + 0x27d4eb2fU));
    return (h >> 8) * 5.9604644775390625e-08f;  // [0,1)
}

WF_HD uint64_t reverse_bits64(uint64_t n) {
#ifdef __CUDA_ARCH__
    return __brevll(n);
#else
    uint64_t r = 0;
    for (int i = 0; i < 64; ++i) { r = (r << 1) | (n & 1); n >>= 1; }
    return r;
#endif
}

// InverseRadicalInverse<base>(inverse, nDigits) (lowdiscrepancy.h:82-91)
WF_HD uint64_t inverse_radical_inverse(uint32_t base, uint64_t inverse, uint32_t nDigits) {
    uint64_t index = 0;
    for (uint32_t i = 0; i < nDigits; ++i) {
        const uint64_t digit = inverse % base;
        inverse /= base;
        index = index * base + digit;
    }
    return index;
}

// GlobalSampler::GetIndexForSample: the index of sample `s` of full-frame pixel (px, py) in the global sequence.
WF_HD uint64_t sampler_index(const WfSampler &S, int32_t px, int32_t py, uint32_t s) {
    if (S.kind == WF_SAMPLER_HALTON) {
        // HaltonSampler::GetIndexForSample (halton.cpp:96-116); kMaxResolution = 128
        // sampleStride = baseScales[0] * baseScales[1] <= 128 * 243, so sampleStride / baseScales[i] is the other scale and
        // both products stay below 2^23: the reference's 64-bit sum and modulo are exact in 32 bits
        uint32_t offset = 0;
        if (S.stride > 1) {
            const uint32_t pmx = (uint32_t)(((px % 128) + 128) % 128), pmy = (uint32_t)(((py % 128) + 128) % 128);
            const uint32_t d0 = (uint32_t)inverse_radical_inverse(2, pmx, S.baseExp0);
            const uint32_t d1 = (uint32_t)inverse_radical_inverse(3, pmy, S.baseExp1);
            offset += d0 * S.baseScale1 * S.multInv0;
            offset += d1 * S.baseScale0 * S.multInv1;
            offset %= S.stride;
        }
        return offset + (uint64_t)s * S.stride;
    }
    if (S.kind == WF_SAMPLER_SOBOL) {
        // SobolIntervalToIndex(log2Resolution, sampleNum, currentPixel - sampleBounds.pMin) (lowdiscrepancy.h:229-250)
        const uint32_t m = S.log2Res;
        if (m == 0) return 0;
        // the reference walks the bits from the lowest and XORs a matrix column for every set one: visiting only the set bits
        // gives the same XOR (lowest_bit)
        uint64_t frame = s;
        uint64_t index = frame << (m << 1);
        uint64_t delta = 0;
        for (; frame; frame &= frame - 1) delta ^= S.vdc[lowest_bit(frame)];
        uint64_t b = (((uint64_t)((uint32_t)(px - S.sbMinX)) << m) | ((uint32_t)(py - S.sbMinY))) ^ delta;
        for (; b; b &= b - 1) index ^= S.vdcInv[lowest_bit(b)];
        return index;
    }
    return s;
}

// RadicalInverseSpecialized<base> (lowdiscrepancy.cpp:389-403); the digits of `a` come off with 32-bit divisions whenever
// a < 2^32 (every frame this library is asked for: stride <= 128 * 243), the reversed number is kept in 64 bits.
WF_HD float radical_inverse(uint32_t base, uint64_t a) {
    const float invBase = 1.f / (float)base;
    uint64_t rev = 0;
    float invBaseN = 1.f;
    while (a >> 32) {
        const uint64_t next = a / base, digit = a - next * base;
        rev = rev * base + digit;
        invBaseN *= invBase;
        a = next;
    }
    uint32_t a32 = (uint32_t)a;
    while (a32) {
        const uint32_t next = a32 / base, digit = a32 - next * base;
        rev = rev * base + digit;
        invBaseN *= invBase;
        a32 = next;
    }
    return smin((float)rev * invBaseN, kOneMinusEps);
}
// ScrambledRadicalInverseSpecialized<base> (lowdiscrepancy.cpp:405-424)
WF_HD float scrambled_radical_inverse(uint32_t base, uint32_t magic, uint32_t shift, const uint16_t *perm, uint64_t a) {
    const float invBase = 1.f / (float)base;
    uint64_t rev = 0;
    float invBaseN = 1.f;
    while (a >> 32) {
        const uint64_t next = a / base, digit = a - next * base;
        rev = rev * base + perm[digit];
        invBaseN *= invBase;
        a = next;
    }
    uint32_t a32 = (uint32_t)a;
    while (a32) {
        const uint32_t next = div_magic(a32, magic, shift), digit = a32 - next * base;
        rev = rev * base + perm[digit];
        invBaseN *= invBase;
        a32 = next;
    }
    return smin(invBaseN * ((float)rev + invBase * (float)perm[0] / (1.f - invBase)), kOneMinusEps);
}

// SobolSampler::SampleDimension (sobol.cpp:48-60) over SobolSampleFloat (lowdiscrepancy.h:252-266), scramble 0.  Out of line: the
// shade kernels draw at ten places, and the lean instantiation (Halton in every thesis template) proved sensitive to what is
// inlined around its hot path — the 32-bit loop below, inlined, cost the HEADLINE frame 6 % (107.4 against 101.3 ms, same box).
WF_HD_NOINLINE float sobol_dim(const WfSampler &S, uint64_t index, uint32_t dim, int32_t px, int32_t py) {
    // SobolSampler::SampleDimension (sobol.cpp:48-60) over SobolSampleFloat (lowdiscrepancy.h:252-266), scramble 0
    if (dim >= (uint32_t)kSobolDims) dim = kSobolDims - 1;
    uint32_t v = 0;
    const uint32_t *col = S.sobol32 + dim * kSobolMatrixSize;
    // (the index of every frame this library is asked for fits 32 bits: 2 * log2(resolution) + log2(spp) bits — the 32-bit
    // find-first-set is one instruction, the 64-bit one several, and this loop is a quarter of k_shade<true>'s instructions)
    uint32_t lo = (uint32_t)index;
    if (S.sobolBytes) {   // the same XOR, a byte of the index at a time
        const uint32_t *T = S.sobolBytes + (size_t)dim * 1024u;
        v = T[lo & 255u] ^ T[256u + ((lo >> 8) & 255u)] ^ T[512u + ((lo >> 16) & 255u)] ^ T[768u + (lo >> 24)];
        lo = 0;
    }
    for (; lo != 0; lo &= lo - 1) v ^= col[lowest_bit32(lo)];
    for (uint64_t a = index >> 32; a != 0; a &= a - 1) v ^= col[32 + lowest_bit(a)];
    float sv = smin((float)v * 0x1p-32f, kOneMinusEps);
    if (dim < 2) {
        sv = sv * (float)(int32_t)S.res + (float)(dim == 0 ? S.sbMinX : S.sbMinY);
        sv = clampf(sv - (float)(dim == 0 ? px : py), 0.f, kOneMinusEps);
    }
    return sv;
}

// GlobalSampler::SampleDimension of the sample with global index `index` at full-frame pixel (px, py).
WF_HD float sampler_dim(const WfSampler &S, uint64_t index, uint32_t dim, int32_t px, int32_t py, uint32_t pixOrd, uint32_t s) {
    if (S.kind == WF_SAMPLER_HALTON) {
        // HaltonSampler::SampleDimension (halton.cpp:118-127)
        if (dim == 0) return (float)((double)reverse_bits64(index >> S.baseExp0) * 0x1p-64);   // RadicalInverse base 2 (:427-435)
        if (dim == 1) return radical_inverse(3, index / S.baseScale1);
        if (dim >= S.nDims) dim = S.nDims - 1;   // host rejects frames that can get here (max_depth vs. table size)
        const HaltonDim di = S.dimInfo[dim];
        return scrambled_radical_inverse(di.prime, di.magic, di.shift, S.perms + di.offset, index);
    }
    if (S.kind == WF_SAMPLER_SOBOL) return sobol_dim(S, index, dim, px, py);
    return hash_sample(S.seed, pixOrd, s, dim);
}

}  // namespace b200rt
