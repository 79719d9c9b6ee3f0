/*
 * b200rt_wavefront.h — C ABI of the wavefront path-tracing driver.
 *
 * Restructures the reference's per-tile, per-pixel, per-sample loop
 *   SamplerIntegrator::Render            src/core/integrator.cpp:230-360
 *   PathIntegrator::Li                   src/integrators/path.cpp:64-204
 *   UniformSampleOneLight/EstimateDirect src/core/integrator.cpp:86-217
 *   VisibilityTester::Unoccluded         src/core/light.cpp:59-64
 * into device-resident ray batches: camera rays -> closest-hit queue -> shade -> (shadow any-hit
 * queue, bounce closest-hit queue) -> ... up to maxdepth, with warp-ballot queue compaction.
 *
 * Shading subset (what the config scenes' templates need around the traversal): perspective camera
 * (src/cameras/perspective.cpp), point lights (src/lights/point.cpp), distant lights (src/lights/distant.cpp), sphere area
 * lights (src/lights/diffuse.cpp, Sphere::Sample/Pdf) and constant infinite lights (src/lights/infinite.cpp with the 1x1
 * map) — the last two with multiple importance sampling (EstimateDirect's two strategies); one of them through the
 * description's own fields or several through its `lights` array (ABI v10, UniformSampleOneLight) —, "matte" and
 * "plastic", "mirror", smooth "glass", "metal" and "substrate" materials per primitive (src/materials/matte.cpp, plastic.cpp, mirror.cpp, glass.cpp, metal.cpp, substrate.cpp), box pixel filter, Russian roulette of
 * path.cpp:190-199, path depths up to 68 (scenes/living-room uses 65).  Samples come
 * from a counter-based hash of (pixel, sample, dimension), so the image does not depend on the
 * execution order, the batch size or the number of GPUs.
 *
 * Image tiles are the reference's 16x16 tile grid (integrator.cpp:237-239); a rank renders the tiles
 * with `tile_index % world == rank`.
 */
#ifndef B200RT_WAVEFRONT_H
#define B200RT_WAVEFRONT_H
#include "b200rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    float eye[3];
    float dir[3];    /* unit view direction */
    float right[3];  /* screen-x half extent at distance 1 (already scaled by tan(fov/2) and aspect) */
    float up[3];     /* screen-y half extent at distance 1 */
    uint32_t xres, yres;
} b200rt_camera;

/* One image texture: the reference's MIPMap pyramid (b200rt_mipmap_build makes it from the image as ReadImage returns it; the
 * C++ drop-in hands over the pyramid of the reference's own ImageTexture) and its UVMapping2D (core/texture.cpp).
 * texels: HOST pointer, RGB floats, level 0 (width x height, both powers of two, row-major, s fastest, row 0 = t 0) followed by
 * the coarser levels max(1, width >> l) x max(1, height >> l). */
typedef struct {
    const float *texels;
    uint32_t width, height, n_levels;
    float su, sv, du, dv;   /* "float uscale" / "vscale" / "udelta" / "vdelta": 1, 1, 0, 0 */
    uint32_t filter;        /* B200RT_TEXFILTER_TRILINEAR (0) = "bool trilinear" true: MIPMap::Lookup(st, width) (mipmap.h:244-260);
                               B200RT_TEXFILTER_EWA (1) = the ImageTexture default: the elliptically weighted average of
                               MIPMap::Lookup(st, dst0, dst1) / EWA (mipmap.h:277-353) over the two pyramid levels around the
                               ellipse's minor axis, Gaussian weights from MIPMap::weightLut (b200rt_ewa_weights) */
    float max_anisotropy;   /* "float maxanisotropy": 0 = the default 8 */
} b200rt_texture;
enum { B200RT_TEXFILTER_TRILINEAR = 0, B200RT_TEXFILTER_EWA = 1 };

/* One light of a scene with several (ABI v10; `n_lights` below).  The fields mean what the desc's single-light fields
 * light_kind / light_pos / light_I / light_radius / light_prim / light_to_world mean for that kind. */
typedef struct {
    uint32_t kind;        /* B200RT_LIGHT_* */
    uint32_t prim;        /* B200RT_LIGHT_SPHERE: the sphere's primitive number */
    float pos[3];         /* point: "from"; sphere: centre; distant: unit direction TOWARDS the light */
    float I[3];           /* point: "I"; the others: radiance "L" */
    float radius;         /* sphere: its radius; distant / infinite: the scene's bounding-sphere radius */
    float to_world[9];    /* infinite: rotation of LightToWorld, row-major; all zero = identity */
    float weight;         /* this light's entry in the light-sampling distribution (below) */
} b200rt_light;

typedef struct {
    b200rt_camera camera;
    uint32_t spp;        /* Sampler "pixelsamples" */
    uint32_t max_depth;  /* Integrator "path" "maxdepth" (path.cpp:110) */
    float light_pos[3];  /* LightSource "point" "from" */
    float light_I[3];    /* "I" */
    float kd[3];         /* Material "matte" "Kd" */
    uint32_t tile_size;  /* 16 in the reference (integrator.cpp:237) */
    uint32_t rank, world; /* this process renders tiles t with t % world == rank */
    uint32_t seed;
    uint32_t max_batch;  /* paths in flight per wavefront batch (0 = default 1<<25; about 250 bytes of queue buffers per path and batch in flight) */
    /* ---- appended in ABI v4; all-zero = the v3 behaviour (one point light, one Kd for every primitive) ---- */
    uint32_t light_kind;   /* B200RT_LIGHT_POINT, or B200RT_LIGHT_SPHERE = AreaLightSource "area" on a full `Shape "sphere"`
                              (lights/diffuse.cpp:68-87 + Sphere::Sample / Pdf, shapes/sphere.cpp:232-305; one-sided),
                              sampled with multiple importance sampling as EstimateDirect does (integrator.cpp:109-217):
                              a light sample (shadow any-hit ray) AND a BSDF sample (an extra closest-hit "MIS" ray) per bounce,
                              plus the emitted radiance seen directly by camera rays (path.cpp:97-107). */
    float light_radius;    /* sphere radius; light_pos = its centre, light_I = its emitted radiance "L".
                              B200RT_LIGHT_DISTANT (LightSource "distant", lights/distant.cpp:44-69): light_pos = the unit
                              direction TOWARDS the light (wLight = Normalize(from - to)), light_I = its radiance "L", and
                              light_radius = the scene's bounding-sphere radius (DistantLight::Preprocess); shadow rays end
                              at p + wLight * 2 * radius.  A delta light: no MIS rays. */
    uint32_t light_prim;   /* primitive number of that sphere inside the aggregate (scene-desc prim_kind == 1) */
    /* Materials: NULL = Material "matte" "Kd" = kd for all.  Else n_materials records of 8 floats
     * (kind, Kd.r, Kd.g, Kd.b, Ks.r, Ks.g, Ks.b, roughness), kind B200RT_MAT_MATTE (materials/matte.cpp, sigma 0), B200RT_MAT_PLASTIC (materials/plastic.cpp), B200RT_MAT_MIRROR, B200RT_MAT_GLASS, B200RT_MAT_METAL or B200RT_MAT_SUBSTRATE (records below),
     * and one material index per primitive.  HOST pointers, read during the call. */
    const float *materials;
    uint32_t n_materials;
    const uint8_t *prim_material; /* n_prims bytes */
    /* ---- appended in ABI v5 ---- */
    uint32_t crop[4];      /* Film "cropwindow" in PIXELS of the camera's frame: x0, x1, y0, y1 (= Film::croppedPixelBounds,
                              film.cpp:49-57: ceil(res * window)); all zero = the whole frame.  Tiles are laid from the
                              window's first pixel (integrator.cpp:237-246 over GetSampleBounds), and `film` / `heat` hold
                              (y1-y0) x (x1-x0) pixels, like the reference's Film. */
    /* ---- appended in ABI v6 ---- */
    float light_to_world[9]; /* B200RT_LIGHT_INFINITE (LightSource "infinite", lights/infinite.cpp:44-150) with a CONSTANT
                              radiance — what the reference builds when "mapname" is absent or unreadable (a 1x1 map,
                              infinite.cpp:52-60; scenes/living-room's "textures/Sky 19.pfm" is not in the tree):
                              row-major 3x3 rotation part of LightToWorld, light_I = its radiance "L" (x "scale"),
                              light_radius = the scene's bounding-sphere radius (Preprocess).  Sampled with multiple
                              importance sampling as EstimateDirect does: Sample_Li over (theta, phi) = (u1 pi, u0 2pi)
                              with pdf 1 / (2 pi^2 sin theta) -> shadow ray to p + wi * 2 * radius; a BSDF-sampled
                              closest-hit "MIS" ray that contributes Le when it ESCAPES (integrator.cpp:196-212); camera
                              rays that escape add Le (path.cpp:97-107).  All zero = identity. */
    /* ---- appended in ABI v8: the reference's sample streams and its camera arithmetic ---- */
    uint32_t sampler;      /* B200RT_SAMPLER_HASH (the v7 counter hash), B200RT_SAMPLER_HALTON = Sampler "halton"
                              (samplers/halton.cpp: GetIndexForSample / SampleDimension over the sample bounds of the film's
                              crop window, digit permutations from the reference's default RNG) or B200RT_SAMPLER_SOBOL =
                              Sampler "sobol" (samplers/sobol.cpp) over the generator matrices below.  Both are GlobalSamplers:
                              a value depends on (pixel, sample number, dimension) only, so the image is still independent
                              of batch size and rank count.  Dimensions are consumed in PathIntegrator::Li's order (camera
                              0-4; per bounce: light choice, uLight, uScattering, BSDF sample, Russian roulette if drawn). */
    uint32_t camera_exact; /* != 0: camera rays are generated from the two matrices below with the reference's float
                              operations (PerspectiveCamera::GenerateRayDifferential, cameras/perspective.cpp:97-143, lensradius 0;
                              Transform::operator()(Ray), transform.h:251-264) instead of camera.eye/dir/right/up */
    float raster_to_camera[16]; /* ProjectiveCamera::RasterToCamera.m, row-major (camera.h:101-107) */
    float camera_to_world[16];  /* Camera::CameraToWorld at the shutter-open time, row-major */
    const uint32_t *sobol_matrices32; /* B200RT_SAMPLER_SOBOL: SobolMatrices32 (1024 x 52, sobolmatrices.cpp), HOST pointer */
    const uint64_t *sobol_vdc;        /* VdCSobolMatrices, sobol_vdc_rows x 52 */
    const uint64_t *sobol_vdc_inv;    /* VdCSobolMatricesInv, sobol_vdc_rows x 52 */
    uint32_t sobol_vdc_rows;
    uint32_t material_stride; /* floats per material record: 0 or 8 = the v7 records; 20 = v8 records
                              (kind, c0.rgb, c1.rgb, roughness, Kr.rgb, Kt.rgb, opacity.rgb, eta, kdTex, opacityTex) which B200RT_MAT_UBER needs */
    /* ---- ABI v9: `Texture "name" "spectrum" "imagemap" ... "bool trilinear" "true"` (textures/imagemap.cpp, core/mipmap.h) ---- */
    const b200rt_texture *textures; /* HOST array; slot 18 of a 20-float material record = 1 + index of the texture its Kd reads
                              (matte, plastic, substrate, uber), slot 19 = 1 + index of uber's opacity texture; 0 = the constant */
    uint32_t n_textures;
    /* ---- ABI v10: several lights, UniformSampleOneLight over a light distribution (integrator.cpp:86-107) ---- */
    const b200rt_light *lights; /* HOST array of n_lights records, in the order of Scene::lights; when n_lights > 0 it REPLACES the
                              single-light fields above (n_lights = 1 is the same render as those fields).  Per path vertex one
                              light is chosen as Distribution1D::SampleDiscrete (sampling.h:90-100) chooses it from the sample
                              the path integrator draws for the choice: the distribution is built, as the reference builds it
                              (sampling.h:57-69), from the records' `weight`s — all equal = Integrator "path" "string
                              lightsamplestrategy" "uniform" (UniformLightDistribution, lightdistrib.cpp:68-75), Light::Power().y()
                              of each light = "power" (PowerLightDistribution).  The estimate is divided by the choice's
                              probability; a light of weight 0 is never chosen.  "spatial" (the default when a scene has
                              more than one light): light_strategy below.  Emitted radiance seen by camera rays and after specular
                              bounces: the hit sphere's own light (path.cpp:100) and the SUM over every infinite light for rays
                              that escape (path.cpp:103-104).  At most B200RT_MAX_LIGHTS. */
    uint32_t n_lights;
    uint32_t light_strategy; /* with n_lights > 1: B200RT_LIGHTS_WEIGHTS (0) = the one distribution over the records' weights described
                              above; B200RT_LIGHTS_SPATIAL (1) = "lightsamplestrategy" "spatial", the path integrator's default
                              (SpatialLightDistribution, lightdistrib.cpp:95-300): the scene's bound is cut into voxels (64 along its
                              widest side), each voxel gets its own Distribution1D from 128 Halton points inside it at which every
                              light is sampled (Li.y() / pdf, floored at 0.1 % of the average), and a path vertex chooses from the
                              distribution of the voxel it lies in.  The reference fills its hash table of voxels lazily; here every
                              voxel's distribution is computed on the device before the frame (a voxel's values do not depend on when
                              it is computed).  The weights are ignored. */
} b200rt_render_desc;
enum { B200RT_LIGHTS_WEIGHTS = 0, B200RT_LIGHTS_SPATIAL = 1 };
#define B200RT_MAX_LIGHTS 64u
enum { B200RT_SAMPLER_HASH = 0, B200RT_SAMPLER_HALTON = 1, B200RT_SAMPLER_SOBOL = 2 };
enum { B200RT_LIGHT_POINT = 0, B200RT_LIGHT_SPHERE = 1, B200RT_LIGHT_DISTANT = 2, B200RT_LIGHT_INFINITE = 3 };
enum {
    B200RT_MAT_MATTE = 0,   /* materials/matte.cpp:44-62; record = (kind, Kd.rgb, sigma, -, -, -): sigma (degrees) > 0 = OrenNayar */
    B200RT_MAT_PLASTIC = 1, /* materials/plastic.cpp: Kd + Ks microfacet lobe, roughness remapped */
    /* ---- ABI v7: the perfectly specular materials (path.cpp's specularBounce / etaScale bookkeeping comes with them) ---- */
    B200RT_MAT_MIRROR = 2,  /* materials/mirror.cpp:44-56: SpecularReflection(Kr, FresnelNoOp); record = (kind, Kr.rgb, -, -, -, -) */
    B200RT_MAT_GLASS = 3,   /* materials/glass.cpp:44-63 with roughness 0 under the path integrator (allowMultipleLobes):
                               FresnelSpecular(Kr, Kt, 1, index); record = (kind, Kr.rgb, Kt.rgb, index) */
    B200RT_MAT_METAL = 4,   /* materials/metal.cpp:60-82: MicrofacetReflection(1, TrowbridgeReitz, FresnelConductor(1, eta, k)),
                               isotropic; record = (kind, eta.rgb, k.rgb, roughness) */
    B200RT_MAT_SUBSTRATE = 5, /* materials/substrate.cpp:45-66: FresnelBlend(Kd, Ks, TrowbridgeReitz), isotropic;
                               record = (kind, Kd.rgb, Ks.rgb, roughness) */
    /* ---- ABI v8 ---- */
    B200RT_MAT_UBER = 6,    /* materials/uber.cpp:45-100 (20-float records): SpecularTransmission(1 - opacity) + Lambertian(op Kd) +
                               MicrofacetReflection(op Ks, TrowbridgeReitz, FresnelDielectric(1, eta)) + SpecularReflection(op Kr) +
                               SpecularTransmission(op Kt, 1, eta); c0 = Kd, c1 = Ks */
    B200RT_MAT_NO_REMAP = 32, /* added to `kind`: "bool remaproughness" false — roughness IS the Trowbridge-Reitz alpha
                               (plastic, metal, substrate; scenes/living-room's materials are written that way) */
    B200RT_MAT_REVERSE_ORIENTATION = 16 /* added to `kind`: the shape's reverseOrientation ^ transformSwapsHandedness
                               (triangle.cpp:421-424) — flips the side of a triangle its normal (= the glass's outside) is on */
};

typedef struct {
    uint64_t camera_rays;   /* = paths started */
    uint64_t closest_rays;  /* rays through Aggregate::Intersect (camera + bounce) */
    uint64_t shadow_rays;   /* rays through Aggregate::IntersectP */
    uint64_t launches;      /* kernels launched by this call */
    float ms_closest;       /* CUDA-event time inside the closest-hit kernels */
    float ms_any;           /* ... inside the any-hit kernels */
    float ms_total;         /* whole call on the stream */
    uint32_t n_tiles;       /* tiles rendered by this rank */
    uint32_t launches_closest, launches_any; /* traversal-kernel launches behind ms_closest / ms_any */
    uint64_t mis_rays;      /* of closest_rays: the BSDF-sampled MIS rays of EstimateDirect (integrator.cpp:203) */
} b200rt_render_stats;

/* Number of tiles / pixels rank `rank` of `world` owns for this camera and tile size. */
int b200rt_render_tiles(const b200rt_render_desc *desc, uint32_t *n_tiles, uint64_t *n_pixels);

/* Renders this rank's tiles.  film: DEVICE or HOST buffer of xres*yres*4 floats (sum R, sum G,
 * sum B, sum weight per pixel — Film's Pixel accumulators, src/core/film.h:85-92); pixels of other
 * ranks' tiles are left untouched.  stats may be NULL. */
int b200rt_render(b200rt_scene *scene, const b200rt_render_desc *desc, float *film, b200rt_render_stats *stats,
                  void *stream);

/* Wavefront batches in flight per render call (default 2, or B200RT_WF_STREAMS): consecutive batches are independent, so the
 * next batch's camera / shade / film kernels and the head of its traversal fill the gaps the current batch's traversal tails
 * leave (one B200: 114.5 -> 105.3 ms per headline frame, profiles/r02_ab_wf_streams.log).  The image does not depend on it.
 * With n > 1 the per-kernel times in b200rt_render_stats overlap each other; measure kernels with n = 1. */
int b200rt_render_set_streams(int n);

/* b200rt_render + the thesis's per-pixel traversal statistics: every ray of a path (camera, bounce and shadow rays)
 * adds its Ray::stats to the path's pixel, as PathIntegrator::Li does with `r.stats += ray.stats`
 * (src/integrators/path.cpp:92-200, light.cpp:62) before SamplerIntegrator::Render adds the sample's total to the
 * film pixel (integrator.cpp:327-328); Film::WriteGeneralStats (film.cpp:170-264) dumps one matrix per counter.
 * heat: HOST or DEVICE buffer of B200RT_HEAT_PLANES planes of xres*yres uint32, accumulated into (+=).
 * Runs the instrumented (one-ray-per-thread) traversal kernels, so it is an analysis mode, not the fast path. */
#define B200RT_HEAT_PLANES 6
enum {
    B200RT_HEAT_PRIM = 0,   /* primitiveIntersections                                       */
    B200RT_HEAT_PRIM_P = 1, /* primitiveIntersectionsP                                      */
    B200RT_HEAT_NODE = 2,   /* bvh- / kd- / bspTreeNodeTraversals (interior nodes, closest) */
    B200RT_HEAT_NODE_P = 3, /* ... of shadow rays                                           */
    B200RT_HEAT_LEAF = 4,   /* leafNodeTraversals                                           */
    B200RT_HEAT_LEAF_P = 5  /* leafNodeTraversalsP                                          */
};
int b200rt_render_heatmaps(b200rt_scene *scene, const b200rt_render_desc *desc, float *film, uint32_t *heat,
                           b200rt_render_stats *stats, void *stream);

/* Dumps the ray queues of the LAST b200rt_render batch of this scene (the real closest/shadow mix)
 * for kernel benchmarking: queue q = 2*depth (closest) or 2*depth+1 (shadow).  Pass rays = NULL to
 * query the count.  rays: HOST or DEVICE buffer of *n b200rt_ray. */
int b200rt_render_queue(b200rt_scene *scene, uint32_t queue, b200rt_ray *rays, uint64_t *n);
int b200rt_render_keep_queues(b200rt_scene *scene, int keep); /* enable recording (off by default) */
/* Path slot (index into the batch; see b200rt_render_slot_pixel) of every entry of queue q, and — for shadow (odd q) and
 * MIS queues — the radiance the ray carries if it is unoccluded / reaches the light.  MIS queues are q = 2*(max_depth+1)+depth.
 * path / contrib: HOST buffers of *n entries (contrib: 4 floats each; NULL to skip). */
int b200rt_render_queue_paths(b200rt_scene *scene, uint32_t queue, uint32_t *path, float *contrib, uint64_t *n);

/* The camera matrices of `LookAt eye look up` (or of an explicit world-to-camera matrix, `Transform [...]`, when w2c16 != NULL)
 * + `Camera "perspective" "float fov"` on an xres x yres film (frame aspect default screen window), computed with the
 * reference's own float operations: LookAt / Perspective / Scale / Translate / Inverse / operator* (src/core/transform.cpp:82-136,
 * 203-236,303-311; camera.h:87-108; api.cpp pbrtCamera).  Fills desc->camera (xres, yres and the legacy frame),
 * raster_to_camera, camera_to_world and sets camera_exact = 1.  lookat9 = eye, look, up. */
int b200rt_camera_perspective(const float *lookat9, const float *w2c16, float fov, uint32_t xres, uint32_t yres,
                              b200rt_render_desc *desc);

/* The one exchange step of a multi-process render (one process per GPU, ONE node): sums the films of all `world` ranks into
 * rank 0's buffer, in rank order (deterministic).  Replaces what Film::MergeFilmTile (src/core/film.cpp:118-135) does for the
 * reference's threads: each rank's film holds its own tiles, the boundary samples' contributions to the one-pixel halo around
 * them, and zeros elsewhere, so the frame is the sum.  Rank 0 maps each peer's device buffer through CUDA IPC (NVLink / NVSwitch
 * peer access between GPUs; also works for ranks sharing a GPU) and adds it with one kernel; there is no launcher on the C++
 * side (`pbrt scene.pbrt` per rank), so the handles travel through files `film.<rank>` / `done.<rank>` in the directory
 * `rendezvous`, which every rank must reach and which must not hold files of an earlier, aborted call.
 * film: HOST or DEVICE, n_floats = 4 * pixels; heat: optional (NULL) B200RT_HEAT_PLANES planes of uint32, summed likewise.
 * On rank 0 film / heat hold the frame afterwards; on other ranks they are unchanged.  Blocks until rank 0 has read this
 * rank's buffer; timeout_s <= 0 means 120 s.  world <= 1 returns 0 at once. */
/* MIPMap::MIPMap (core/mipmap.h:113-209) + ImageTexture::GetTexture's y flip and convertIn (textures/imagemap.cpp:73-90) on the
 * host: rgb = w x h x 3 floats as ReadImage returns them (row 0 = top of the image; 8-bit formats: byte / 255.f), gamma != 0 =
 * InverseGammaCorrect (the default for .tga / .png), scale = "float scale".  b200rt_mipmap_dims gives the power-of-two level-0
 * size, the level count and the total texel count; out = n_texels x 3 floats.  Texels equal the reference's bit for bit. */
int b200rt_mipmap_dims(uint32_t w, uint32_t h, uint32_t *w0, uint32_t *h0, uint32_t *n_levels, uint64_t *n_texels);
int b200rt_mipmap_build(const float *rgb, uint32_t w, uint32_t h, int gamma, float scale, float *out);
/* MIPMap::weightLut (mipmap.h:94-101): weightLut[i] = exp(-2 r2) - exp(-2), r2 = i / 127, evaluated like the reference does
 * (std::exp on floats, this machine's libm).  out: 128 floats. */
void b200rt_ewa_weights(float *out128);

int b200rt_film_reduce(const char *rendezvous, uint32_t rank, uint32_t world, float *film, uint64_t n_floats,
                       uint32_t *heat, uint64_t n_heat, double timeout_s);

#ifdef __cplusplus
}
#endif
#endif
