// Host-side set-up shared by the wavefront driver (wavefront.cu) and the host-compiled check of the shade stage
// (tests/hostshade): tile layout of a render description and the material record -> lobe list conversion.
#pragma once
#include <vector>

#include "wf_shade.cuh"

namespace b200rt {
int set_error(int code, const char *fmt, ...);
int tiling_of(const b200rt_render_desc *d, WfTiling *t);
// One material record -> its lobes, in the order the material's ComputeScatteringFunctions adds BxDFs.
// rec: `stride` floats (b200rt_wavefront.h).  out: kMatFloats floats.  Returns false (with the error set) on a bad record.
bool material_lobes(const float *rec, uint32_t stride, uint32_t index, float *out, bool *hasSpecular, uint32_t nTextures = 0);
// Light block of WfParams from the description; returns an error code for unsupported / inconsistent lights.
int light_of(const b200rt_render_desc *d, uint32_t n_prims, WfLight *Lt);
int lights_of(const b200rt_render_desc *d, uint32_t n_prims, std::vector<WfLight> &out);   // Scene::lights + the light distribution (ABI v10)
// RadicalInverse(0..4, i), i < 128: the 3-D Halton points and 2-D light samples of SpatialLightDistribution::ComputeDistribution
// (lightdistrib.cpp:248-257); 5 floats per i.
void spatial_halton_table(float *out640);
}  // namespace b200rt
