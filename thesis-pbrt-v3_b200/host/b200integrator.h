// b200integrator.h — `Integrator "b200wavefront"` for Thesis-pbrt-v3: the pbrt-side entry of the wavefront driver.
//
// Mirrors the reference's integrator interface for the hot path:
//   class Integrator { virtual void Render(const Scene &scene) = 0; }            src/core/integrator.h:52-58
//   SamplerIntegrator::Render (tile / pixel / sample loops)                       src/core/integrator.cpp:230-360
//   PathIntegrator(maxDepth, camera, sampler, pixelBounds, ...) + CreatePathIntegrator   src/integrators/path.cpp:206-240
// and hands the frame to b200rt_render (include/b200rt_wavefront.h).  The camera, a point light, matte materials and
// the film are read / written through pbrt's PUBLIC interfaces (Camera::GenerateRay, Light::Sample_Li,
// Material::ComputeScatteringFunctions + BSDF::f, Film::SetImage / WriteImage); the sphere area light's geometry and
// PlasticMaterial's parameters have no accessors and are read as b200integrator.cpp explains.
//
// The branch a maintainer adds to RenderOptions::MakeIntegrator (api.cpp:1895-1945, before the final `else`):
//     else if (IntegratorName == "b200wavefront")
//         integrator = CreateB200WavefrontIntegrator(IntegratorParams, sampler, camera);
#ifndef B200RT_HOST_B200INTEGRATOR_H
#define B200RT_HOST_B200INTEGRATOR_H

#include <memory>
#include <string>

#include "pbrt.h"
#include "integrator.h"
#include "b200rt_wavefront.h"

namespace pbrt {

class B200WavefrontIntegrator : public Integrator {
  public:
    // rank/world: this process renders the 16x16 tiles t with t % world == rank (integrator.cpp:237-239 tile grid);
    // with world > 1 the ranks' films are summed into rank 0's through b200rt_film_reduce (CUDA IPC peer memory; `rendezvous`
    // names the directory the handles travel through) and rank 0 alone writes the image and the statistic matrices.
    B200WavefrontIntegrator(int maxDepth, std::shared_ptr<const Camera> camera, std::shared_ptr<Sampler> sampler,
                            int rank = 0, int world = 1, int seed = 0, bool generalStats = false, std::string rendezvous = "",
                            std::string lightSampleStrategy = "spatial")
        : maxDepth(maxDepth), camera(std::move(camera)), sampler(std::move(sampler)), rank(rank), world(world), seed(seed),
          generalStats(generalStats), rendezvous(std::move(rendezvous)), lightSampleStrategy(std::move(lightSampleStrategy)) {}
    void Render(const Scene &scene);
    const b200rt_render_stats &stats() const { return lastStats; }

  private:
    const int maxDepth;
    std::shared_ptr<const Camera> camera;
    std::shared_ptr<Sampler> sampler;
    const int rank, world, seed;
    const bool generalStats;
    const std::string rendezvous;
    const std::string lightSampleStrategy;   // with several lights: "spatial" (default) | "uniform" | "power"
    b200rt_render_stats lastStats{};
};

// Parameters: "integer maxdepth" (path.cpp:209, default 5), "integer rank" / "integer world" (explicit only, default 0 / 1;
// world > 1 requires "string rendezvous", a directory all ranks of the node can write), "integer seed" (hash sampler only),
// "string lightsamplestrategy" (path.cpp:218-219, default "spatial"; it only matters when the scene has several lights),
// "bool generalstats" (default false:
// true renders through the instrumented kernels, b200rt_render_heatmaps, fills the Film's per-pixel GeneralStats and
// writes the thesis's statistic matrices with Film::WriteGeneralStats like SamplerIntegrator::Render does,
// integrator.cpp:327-328,359).
B200WavefrontIntegrator *CreateB200WavefrontIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler,
                                                       std::shared_ptr<const Camera> camera);

}  // namespace pbrt
#endif
