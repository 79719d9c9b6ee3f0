// TEST INFRASTRUCTURE.  src/textures/ptex.cpp needs the vendored Ptex library (cmake build); the full-scene test binary
// (oracle/_ref/pbrt_b200) links these two factories instead: `Texture "..." "ptex"` is reported as unavailable.
#include "std_first.h"
#include "textures/ptex.h"
#include "core/paramset.h"
namespace pbrt {
PtexTexture<Float> *CreatePtexFloatTexture(const Transform &, const TextureParams &) { Error("ptex textures are not part of this test build"); return nullptr; }
PtexTexture<Spectrum> *CreatePtexSpectrumTexture(const Transform &, const TextureParams &) { Error("ptex textures are not part of this test build"); return nullptr; }
}  // namespace pbrt
