// Host-side set-up of the sample streams (wf_sampler_host.cpp).
#pragma once
#include <stdint.h>
#include <vector>
#include "wf_sampler.cuh"

namespace b200rt {

struct HaltonTables {
    std::vector<uint16_t> perms;      // ComputeRadicalInversePermutations(RNG())
    std::vector<HaltonDim> dims;      // (Primes[d], PrimeSums[d]) for d < 1000
};
const HaltonTables &halton_tables();
// HaltonSampler's constructor for a sample-bounds extent of resx x resy pixels
void halton_setup(int resx, int resy, WfSampler *S);
// SobolSampler's constructor (resolution, log2Resolution, sampleBounds.pMin)
void sobol_setup(int resx, int resy, int sbMinX, int sbMinY, WfSampler *S);

}  // namespace b200rt
