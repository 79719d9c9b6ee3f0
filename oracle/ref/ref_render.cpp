// Shim TU: drives the reference's OWN renderer — SamplerIntegrator::Render (core/integrator.cpp:230-360) with
// PathIntegrator::Li (integrators/path.cpp:64-204), PerspectiveCamera, Film + BoxFilter, RandomSampler /
// HaltonSampler, PointLight, MatteMaterial and the reference's accelerators — over a programmatically built
// scene (no parser, no image libraries: core/imageio.cpp is not compiled, the two functions Film needs from it
// are defined here and capture the pixels instead of writing a file).  TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#include "ref_internal.h"
#define private public
#define protected public
#include "core/sampler.h"
#include "core/camera.h"
#include "core/primitive.h"
#include "core/film.h"
#include "core/mipmap.h"
#include "textures/imagemap.h"
#undef private
#undef protected
#include "core/api.h"
#include "core/camera.h"
#include "core/imageio.h"
#include "core/light.h"
#include "core/parallel.h"
#include "core/paramset.h"
#include "core/scene.h"
#include "core/spectrum.h"
#include "core/stats.h"
#include "cameras/perspective.h"
#include "filters/box.h"
#include "integrators/path.h"
#include "lights/point.h"
#include "lights/diffuse.h"
#include "lights/distant.h"
#include "lights/infinite.h"
#include "materials/matte.h"
#include "materials/plastic.h"
#include "materials/mirror.h"
#include "materials/glass.h"
#include "materials/metal.h"
#include "materials/substrate.h"
#include "materials/uber.h"
#include "samplers/halton.h"
#include "samplers/sobol.h"
#include "samplers/random.h"
#include "textures/constant.h"

namespace pbrt {
static std::mutex g_imgMutex;
static std::vector<float> g_lastImage;
static Point2i g_lastRes;
// core/imageio.h:55 — Film::WriteImage (film.cpp:266-305) ends here with the final RGB pixels.
void WriteImage(const std::string &, const Float *rgb, const Bounds2i &outputBounds, const Point2i &) {
    std::lock_guard<std::mutex> l(g_imgMutex);
    Vector2i res = outputBounds.Diagonal();
    g_lastRes = Point2i(res.x, res.y);
    g_lastImage.assign(rgb, rgb + 3 * (size_t)res.x * res.y);
}
// ReadImage is the reference's own (core/imageio.cpp compiled with its WriteImage renamed, oracle/ref/Makefile): image textures
}  // namespace pbrt
using namespace pbrt;

// The pixels of the last Film::WriteImage in this process (used by tests/hostplugin for the C++ integrator).
extern "C" int ref_last_image(float *out_rgb, int *xres, int *yres) {
    std::lock_guard<std::mutex> l(g_imgMutex);
    *xres = g_lastRes.x; *yres = g_lastRes.y;
    if (out_rgb) std::memcpy(out_rgb, g_lastImage.data(), sizeof(float) * g_lastImage.size());
    return 0;
}

namespace {
// Counts the calls that reach the aggregate (what Scene::Intersect/IntersectP count, scene.cpp:41-52) in per-thread,
// cache-line-sized slots, so that counting does not slow the multithreaded render down.
struct alignas(64) CountSlot { uint64_t closest = 0, any = 0; };
static std::mutex g_slotMutex;
static std::vector<CountSlot *> g_slots;
static CountSlot *my_slot() {
    static thread_local CountSlot *tl = nullptr;
    if (!tl) {
        tl = new CountSlot;   // lives until process exit: pool threads may outlive the render
        std::lock_guard<std::mutex> l(g_slotMutex);
        g_slots.push_back(tl);
    }
    return tl;
}
struct CountingAggregate : public Aggregate {
    std::shared_ptr<Primitive> inner;
    explicit CountingAggregate(std::shared_ptr<Primitive> p) : inner(std::move(p)) {
        std::lock_guard<std::mutex> l(g_slotMutex);
        for (CountSlot *c : g_slots) c->closest = c->any = 0;
    }
    Bounds3f WorldBound() const override { return inner->WorldBound(); }
    bool Intersect(const Ray &r, SurfaceInteraction *si) const override {
        my_slot()->closest++;
        return inner->Intersect(r, si);
    }
    bool IntersectP(const Ray &r) const override {
        my_slot()->any++;
        return inner->IntersectP(r);
    }
    void totals(uint64_t *c, uint64_t *a) const {
        std::lock_guard<std::mutex> l(g_slotMutex);
        *c = *a = 0;
        for (CountSlot *s : g_slots) { *c += s->closest; *a += s->any; }
    }
};
}  // namespace

// ---- event trace of a single-threaded render (tests/: the shade stage's sample dimensions and spawned rays are pinned
// against it).  One 16-word record per event, in execution order:
//   1 StartPixel        [1] x [2] y
//   2 StartNextSample   [1] returned bool
//   3 Get1D             [1] dimension the value came from [2] value
//   4 Get2D             [1] dimension [2] [3] values
//   5 Intersect         [1..3] o [4..6] d [7] tMax before [8] primitive (-1 miss) [9] tMax after [10..12] isect.n [13..15] isect.shading.n
//   6 IntersectP        [1..3] o [4..6] d [7] tMax [8] occluded
static bool g_trace = false;
static std::vector<uint32_t> g_traceLog;
static inline uint32_t fbits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static void trace_push(const uint32_t (&r)[16]) { g_traceLog.insert(g_traceLog.end(), r, r + 16); }
extern "C" void ref_trace_enable(int on) { g_trace = on != 0; g_traceLog.clear(); }
extern "C" uint64_t ref_trace_records(void) { return g_traceLog.size() / 16; }
extern "C" void ref_trace_copy(uint32_t *out) { std::memcpy(out, g_traceLog.data(), g_traceLog.size() * 4); }

namespace {
template <class Base>
struct TracingSampler : public Base {
    template <class... A> explicit TracingSampler(A &&... a) : Base(std::forward<A>(a)...) {}
    void StartPixel(const Point2i &p) override {
        uint32_t r[16] = {1, (uint32_t)p.x, (uint32_t)p.y};
        trace_push(r);
        Base::StartPixel(p);
    }
    bool StartNextSample() override {
        const bool b = Base::StartNextSample();
        uint32_t r[16] = {2, b ? 1u : 0u};
        trace_push(r);
        return b;
    }
    Float Get1D() override {
        const Float v = Base::Get1D();
        uint32_t r[16] = {3, (uint32_t)(this->dimension - 1), fbits(v)};
        trace_push(r);
        return v;
    }
    Point2f Get2D() override {
        const Point2f v = Base::Get2D();
        uint32_t r[16] = {4, (uint32_t)(this->dimension - 2), fbits(v.x), fbits(v.y)};
        trace_push(r);
        return v;
    }
    std::unique_ptr<Sampler> Clone(int) override { return std::unique_ptr<Sampler>(new TracingSampler<Base>(*this)); }
};
struct TracingAggregate : public Aggregate {
    std::shared_ptr<Primitive> inner;
    std::unordered_map<const Primitive *, int> index;
    explicit TracingAggregate(std::shared_ptr<Primitive> p) : inner(std::move(p)) {}
    Bounds3f WorldBound() const override { return inner->WorldBound(); }
    bool Intersect(const Ray &ray, SurfaceInteraction *si) const override {
        uint32_t r[16] = {5, fbits(ray.o.x), fbits(ray.o.y), fbits(ray.o.z), fbits(ray.d.x), fbits(ray.d.y), fbits(ray.d.z), fbits(ray.tMax)};
        const bool hit = inner->Intersect(ray, si);
        int prim = -1;
        if (hit) { auto it = index.find(si->primitive); prim = it == index.end() ? -2 : it->second; }
        r[8] = (uint32_t)prim; r[9] = fbits(ray.tMax);
        if (hit) {
            r[10] = fbits(si->n.x); r[11] = fbits(si->n.y); r[12] = fbits(si->n.z);
            r[13] = fbits(si->shading.n.x); r[14] = fbits(si->shading.n.y); r[15] = fbits(si->shading.n.z);
        }
        trace_push(r);
        return hit;
    }
    bool IntersectP(const Ray &ray) const override {
        uint32_t r[16] = {6, fbits(ray.o.x), fbits(ray.o.y), fbits(ray.o.z), fbits(ray.d.x), fbits(ray.d.y), fbits(ray.d.z), fbits(ray.tMax)};
        const bool occ = inner->IntersectP(ray);
        r[8] = occ ? 1u : 0u;
        trace_push(r);
        return occ;
    }
};
}  // namespace

// The matrices of the camera ref_render builds for (cam, xres, yres): ProjectiveCamera::RasterToCamera and CameraToWorld
// (row-major 4x4 each) — what tests compare b200rt_camera_perspective with.
extern "C" void ref_camera_matrices(const float *cam, int xres, int yres, const float *w2c16, float *r2c_out, float *c2w_out) {
    std::unique_ptr<Filter> filter(new BoxFilter(Vector2f(0.5f, 0.5f)));
    Film *film = new Film(Point2i(xres, yres), Bounds2f(Point2f(0, 0), Point2f(1, 1)), std::move(filter), 35.f, "x.exr", 1.f);
    Transform w2c = LookAt(Point3f(cam[0], cam[1], cam[2]), Point3f(cam[3], cam[4], cam[5]), Vector3f(cam[6], cam[7], cam[8]));
    if (w2c16)
        w2c = Transform(Matrix4x4(w2c16[0], w2c16[1], w2c16[2], w2c16[3], w2c16[4], w2c16[5], w2c16[6], w2c16[7], w2c16[8], w2c16[9],
                                  w2c16[10], w2c16[11], w2c16[12], w2c16[13], w2c16[14], w2c16[15]));
    Transform c2w = Inverse(w2c);
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    Float frame = Float(xres) / Float(yres);
    Bounds2f screen;
    if (frame > 1.f) { screen.pMin = Point2f(-frame, -1.f); screen.pMax = Point2f(frame, 1.f); }
    else { screen.pMin = Point2f(-1.f, -1.f / frame); screen.pMax = Point2f(1.f, 1.f / frame); }
    PerspectiveCamera pc(ac2w, screen, 0.f, 1.f, 0.f, 1e6f, cam[9], film, nullptr);
    std::memcpy(r2c_out, pc.RasterToCamera.GetMatrix().m, 64);
    std::memcpy(c2w_out, c2w.GetMatrix().m, 64);
}

// LightSource "distant" "point from" "point to" for the NEXT ref_render calls (NULL = back to the point light):
// from_to = from[3], to[3]; its radiance "L" is ref_render's light_I (lights/distant.cpp:117-125, CreateDistantLight).
static bool g_distant = false;
static float g_from_to[6];
extern "C" void ref_render_set_distant(const float *from_to) {
    g_distant = from_to != nullptr;
    if (from_to) std::memcpy(g_from_to, from_to, sizeof(g_from_to));
}

// LightSource "infinite" "rgb L" (no readable "mapname": the constant 1x1 map of lights/infinite.cpp:52-60) for the NEXT
// renders, under the LightToWorld whose row-major 4x4 is l2w16 (NULL arguments = off).  ReadImage above returns null,
// exactly what the reference gets for scenes/living-room's missing "textures/Sky 19.pfm".
static bool g_infinite = false;
static float g_infL[3], g_infM[16];
extern "C" void ref_render_set_infinite(const float *L_rgb, const float *l2w16) {
    g_infinite = L_rgb != nullptr && l2w16 != nullptr;
    if (g_infinite) { std::memcpy(g_infL, L_rgb, sizeof(g_infL)); std::memcpy(g_infM, l2w16, sizeof(g_infM)); }
}
// More LightSource statements after the first one, for the NEXT renders (n = 0: off): Scene::lights = the render's own light
// followed by these, in order.  A record is 24 floats: kind (0 "point": a = "from", rgb = "I"; 2 "distant": a = from - to,
// rgb = "L"; 3 "infinite": rgb = "L", m = row-major 4x4 LightToWorld, constant map), a[3], rgb[3], m[16], unused.
// strategy = Integrator "path" "string lightsamplestrategy" (path.cpp:218-219; lightdistrib.cpp:48-66).
static std::vector<float> g_extraLights, g_lastPowers;
static std::string g_lightStrategy = "spatial";
extern "C" void ref_render_set_extra_lights(const float *recs, int n, const char *strategy) {
    g_extraLights.assign(recs && n > 0 ? recs : nullptr, recs && n > 0 ? recs + (size_t)24 * n : nullptr);
    g_lightStrategy = (strategy && n > 0) ? strategy : "spatial";
}
extern "C" int ref_render_get_extra_lights(float *recs, int cap, char *strategy, int strategy_cap) {   // for tests/hostplugin's driver
    const int n = (int)(g_extraLights.size() / 24);
    if (recs && n <= cap) std::memcpy(recs, g_extraLights.data(), g_extraLights.size() * sizeof(float));
    if (strategy && strategy_cap > 0) { std::strncpy(strategy, g_lightStrategy.c_str(), (size_t)strategy_cap - 1); strategy[strategy_cap - 1] = 0; }
    return n;
}
// Light::Power().y() of every light of the LAST render's scene, in Scene::lights order (after Scene's constructor ran their
// Preprocess): what PowerLightDistribution weighs the lights by (integrator.cpp:219-227).
extern "C" int ref_render_last_light_powers(float *out, int cap) {
    const int n = (int)g_lastPowers.size();
    for (int i = 0; i < n && i < cap; ++i) out[i] = g_lastPowers[(size_t)i];
    return n;
}
// `Transform [16 floats]` before `Camera` (scenes/living-room/scene.pbrt:2): the world-to-camera CTM as a row-major
// 4x4 for the NEXT renders instead of the LookAt in `cam` (NULL = LookAt again); cam[9] stays the fov.
static bool g_camMatrix = false;
static float g_w2c[16];
extern "C" void ref_render_set_camera_matrix(const float *w2c16) {
    g_camMatrix = w2c16 != nullptr;
    if (w2c16) std::memcpy(g_w2c, w2c16, sizeof(g_w2c));
}

extern "C" int ref_render_get_infinite(float *L_rgb, float *l2w16) {   // for tests/hostplugin's driver (same scene)
    if (g_infinite) { std::memcpy(L_rgb, g_infL, sizeof(g_infL)); std::memcpy(l2w16, g_infM, sizeof(g_infM)); }
    return g_infinite ? 1 : 0;
}
extern "C" int ref_render_get_camera_matrix(float *w2c16) {
    if (g_camMatrix) std::memcpy(w2c16, g_w2c, sizeof(g_w2c));
    return g_camMatrix ? 1 : 0;
}

// floats per prim_mat record for the NEXT renders: 8 (default) or 20 (the device's long record, which adds Material "uber":
// [0] = 6 (+32: remaproughness false), Kd 1-3, Ks 4-6, roughness 7, Kr 8-10, Kt 11-13, opacity 14-16, eta 17)
static int g_matStride = 8;
extern "C" void ref_render_set_mat_stride(int n) { g_matStride = n == 20 ? 20 : 8; }
extern "C" int ref_render_get_mat_stride(void) { return g_matStride; }

// `Texture "t<i>" "spectrum" "imagemap" "string filename" <file> "bool trilinear" "true"` for the NEXT renders: slot 18 / 19 of a
// 20-float prim_mat record name (1-based) the texture the material's Kd / uber's opacity reads.  The textures are made by the
// reference's own CreateImageSpectrumTexture (textures/imagemap.cpp:146-188) -> ReadImage -> MIPMap.
static std::vector<std::string> g_texFiles;
// "bool trilinear" / "float maxanisotropy" of the textures made for the NEXT renders and lookups (defaults of this harness:
// trilinear, 8; the reference's own default is trilinear false = EWA)
static bool g_texTrilinear = true;
static float g_texMaxAniso = 8.f;
extern "C" void ref_render_set_texture_filter(int trilinear, float maxaniso) { g_texTrilinear = trilinear != 0; g_texMaxAniso = maxaniso > 0 ? maxaniso : 8.f; }
extern "C" void ref_render_get_texture_filter(int *trilinear, float *maxaniso) { *trilinear = g_texTrilinear ? 1 : 0; *maxaniso = g_texMaxAniso; }
static std::shared_ptr<Texture<Spectrum>> make_imagemap(const std::string &file) {
    ParamSet ps;
    std::unique_ptr<std::string[]> f(new std::string[1]);
    f[0] = file;
    ps.AddString("filename", std::move(f), 1);
    std::unique_ptr<bool[]> tri(new bool[1]);
    tri[0] = g_texTrilinear;
    ps.AddBool("trilinear", std::move(tri), 1);
    std::unique_ptr<Float[]> ma(new Float[1]);
    ma[0] = g_texMaxAniso;
    ps.AddFloat("maxanisotropy", std::move(ma), 1);
    std::map<std::string, std::shared_ptr<Texture<Float>>> ft;
    std::map<std::string, std::shared_ptr<Texture<Spectrum>>> st;
    TextureParams tp(ps, ps, ft, st);
    // MIPMap's constructor runs ParallelFor loops; the thread pool is not up here (pbrtInit starts it before any texture is made):
    // with one thread the loops run inline (parallel.cpp:186)
    const int nThreads = PbrtOptions.nThreads;
    PbrtOptions.nThreads = 1;
    std::shared_ptr<Texture<Spectrum>> t(CreateImageSpectrumTexture(Transform(), tp));
    PbrtOptions.nThreads = nThreads;
    return t;
}
extern "C" void ref_render_set_textures(int n, const char *const *files) {
    g_texFiles.clear();
    ImageTexture<RGBSpectrum, Spectrum>::ClearCache();
    for (int i = 0; i < n; ++i) g_texFiles.push_back(files[i]);
}
extern "C" int ref_render_get_textures(int i, char *out, int cap) {   // for tests/hostplugin's driver; returns the count
    if (i >= 0 && i < (int)g_texFiles.size() && out) std::snprintf(out, cap, "%s", g_texFiles[i].c_str());
    return (int)g_texFiles.size();
}
// The MIPMap the reference builds for `file`: level count and level-0 size; with out != NULL also the texels of `level`
// (RGB, row-major, s fastest).  Returns the number of levels, 0 if the file cannot be read.
extern "C" int ref_texture_pyramid(const char *file, int level, float *out, int *w, int *h) {
    ImageTexture<RGBSpectrum, Spectrum>::ClearCache();
    std::shared_ptr<Texture<Spectrum>> t = make_imagemap(file);
    auto *it = static_cast<ImageTexture<RGBSpectrum, Spectrum> *>(t.get());
    const MIPMap<RGBSpectrum> *mm = it->mipmap;
    const int nL = mm->Levels();
    if (level < 0 || level >= nL) return 0;
    const BlockedArray<RGBSpectrum> &a = *mm->pyramid[level];
    *w = a.uSize(); *h = a.vSize();
    if (out)
        for (int t2 = 0; t2 < a.vSize(); ++t2)
            for (int s2 = 0; s2 < a.uSize(); ++s2) {
                Float rgb[3];
                a(s2, t2).ToRGB(rgb);
                for (int c = 0; c < 3; ++c) out[3 * ((size_t)t2 * a.uSize() + s2) + c] = rgb[c];
            }
    return nL;
}
// MIPMap::Lookup(st, width) of the reference's texture for n (s, t, width) triples
extern "C" int ref_texture_lookup(const char *file, const float *stw, int n, float *out) {
    std::shared_ptr<Texture<Spectrum>> t = make_imagemap(file);
    auto *it = static_cast<ImageTexture<RGBSpectrum, Spectrum> *>(t.get());
    for (int i = 0; i < n; ++i) {
        Float rgb[3];
        it->mipmap->Lookup(Point2f(stw[3 * i], stw[3 * i + 1]), stw[3 * i + 2]).ToRGB(rgb);
        for (int c = 0; c < 3; ++c) out[3 * i + c] = rgb[c];
    }
    return 0;
}

// MIPMap::Lookup(st, dst0, dst1) of the reference's texture (the filter set by ref_render_set_texture_filter) for n lookups of
// 6 floats: s, t, dst0.x, dst0.y, dst1.x, dst1.y
extern "C" int ref_texture_lookup_diff(const char *file, const float *q, int n, float *out) {
    std::shared_ptr<Texture<Spectrum>> t = make_imagemap(file);
    auto *it = static_cast<ImageTexture<RGBSpectrum, Spectrum> *>(t.get());
    for (int i = 0; i < n; ++i) {
        Float rgb[3];
        it->mipmap->Lookup(Point2f(q[6 * i], q[6 * i + 1]), Vector2f(q[6 * i + 2], q[6 * i + 3]), Vector2f(q[6 * i + 4], q[6 * i + 5])).ToRGB(rgb);
        for (int c = 0; c < 3; ++c) out[3 * i + c] = rgb[c];
    }
    return 0;
}

// Film "float cropwindow" [x0 x1 y0 y1] for the NEXT renders (NULL = whole frame); out_rgb / out_heat then hold the
// window's pixels only (Film::croppedPixelBounds, film.cpp:49-57).
static float g_crop[4] = {0.f, 1.f, 0.f, 1.f};
extern "C" void ref_render_set_crop(const float *cw) {
    const float full[4] = {0.f, 1.f, 0.f, 1.f};
    std::memcpy(g_crop, cw ? cw : full, sizeof(g_crop));
}
extern "C" void ref_render_get_crop(float *cw) { std::memcpy(cw, g_crop, sizeof(g_crop)); }
extern "C" int ref_render_get_distant(float *from_to) {   // for tests/hostplugin's driver, which builds the same scene
    if (g_distant) std::memcpy(from_to, g_from_to, sizeof(g_from_to));
    return g_distant ? 1 : 0;
}

// cam: eye[3], look[3], up[3], fov — exactly a scene file's `LookAt` + `Camera "perspective" "float fov"`.
// sampler: "random" | "halton".  out_rgb: yres*xres*3 floats (Film::WriteImage's buffer).  Returns 0 on success.
extern "C" int ref_render(ref_scene *s, const char *accel_kind, const float *cam, int xres, int yres, int spp,
                          int maxdepth, const float *light_pos, const float *light_I, const float *kd,
                          const char *sampler_name, int n_threads, const char *tmp_prefix, float *out_rgb,
                          uint64_t *n_closest, uint64_t *n_shadow, uint32_t *out_heat /* nullable: 6 planes of yres*xres */,
                          const float *prim_mat /* nullable: 8 floats per primitive (kind 0 matte / 1 plastic, Kd rgb, Ks rgb,
                                                   roughness; kind 2 mirror: Kr; kind 3 glass: Kr, Kt, index; kind 4 metal: eta, k,
                                                   roughness; kind 5 substrate: Kd, Ks, roughness; +32 = remaproughness false) instead of the
                                                   single matte `kd` */,
                          const float *emitter_L /* nullable: rgb; the scene's LAST primitive (a sphere) becomes
                                                    AreaLightSource "area" "color L" and replaces the point light */,
                          double *render_seconds /* nullable: wall time of integrator.Render() alone (no tree build) */) {
    PbrtOptions.quiet = true;
    PbrtOptions.nThreads = n_threads;
    // material: Material "matte" "rgb Kd" (materials/matte.cpp:44-60), sigma 0
    Float rgb[3] = {kd[0], kd[1], kd[2]};
    auto Kd = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(rgb));
    auto sigma = std::make_shared<ConstantTexture<Float>>(0.f);
    auto matte = std::make_shared<MatteMaterial>(Kd, sigma, nullptr);
    MediumInterface mi;
    std::vector<std::shared_ptr<Primitive>> prims;
    std::shared_ptr<AreaLight> area;
    std::map<std::array<float, 20>, std::shared_ptr<Material>> byKd;   // one Material per distinct record
    for (size_t i = 0; i < s->prims.size(); ++i) {
        auto *gp = static_cast<GeometricPrimitive *>(s->prims[i].get());
        std::shared_ptr<Material> m = matte;
        if (prim_mat) {
            std::array<float, 20> k{};
            std::memcpy(k.data(), prim_mat + (size_t)g_matStride * i, sizeof(float) * g_matStride);
            auto it = byKd.find(k);
            if (it == byKd.end()) {
                Float c[3] = {k[1], k[2], k[3]}, cs[3] = {k[4], k[5], k[6]};
                std::shared_ptr<Texture<Spectrum>> tex = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(c));
                if (g_matStride == 20 && k[18] > 0 && (size_t)k[18] <= g_texFiles.size()) tex = make_imagemap(g_texFiles[(size_t)k[18] - 1]);
                std::shared_ptr<Material> mm;
                const int kind = (int)k[0] & 15;
                const bool remap = !((int)k[0] & 32);   // "bool remaproughness"
                if (kind == 1) {   // Material "plastic" "color Kd" "color Ks" "float roughness" (materials/plastic.cpp:72-84)
                    auto texS = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs));
                    auto rough = std::make_shared<ConstantTexture<Float>>(k[7]);
                    mm = std::make_shared<PlasticMaterial>(tex, texS, rough, nullptr, remap);
                } else if (kind == 4) {   // Material "metal" "rgb eta" "rgb k" "float roughness" (materials/metal.cpp:111-128)
                    auto texK = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs));
                    auto rough = std::make_shared<ConstantTexture<Float>>(k[7]);
                    mm = std::make_shared<MetalMaterial>(tex, texK, rough, nullptr, nullptr, nullptr, remap);
                } else if (kind == 5) {   // Material "substrate" "rgb Kd" "rgb Ks" "float uroughness" = "vroughness" (materials/substrate.cpp:68-83)
                    auto texS = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs));
                    auto rough = std::make_shared<ConstantTexture<Float>>(k[7]);
                    mm = std::make_shared<SubstrateMaterial>(tex, texS, rough, rough, nullptr, remap);
                } else if (kind == 6) {   // Material "uber" (materials/uber.cpp:102-128): constant textures, isotropic roughness
                    Float cr[3] = {k[8], k[9], k[10]}, ct[3] = {k[11], k[12], k[13]}, co[3] = {k[14], k[15], k[16]};
                    auto texS = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs));
                    auto texR = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cr));
                    auto texT = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(ct));
                    std::shared_ptr<Texture<Spectrum>> texO = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(co));
                    if (k[19] > 0 && (size_t)k[19] <= g_texFiles.size()) texO = make_imagemap(g_texFiles[(size_t)k[19] - 1]);
                    auto rough = std::make_shared<ConstantTexture<Float>>(k[7]);
                    auto eta = std::make_shared<ConstantTexture<Float>>(k[17]);
                    mm = std::make_shared<UberMaterial>(tex, texS, texR, texT, rough, nullptr, nullptr, texO, eta, nullptr, remap);
                } else if (kind == 2) {   // Material "mirror" "color Kr" (materials/mirror.cpp:58-63)
                    mm = std::make_shared<MirrorMaterial>(tex, nullptr);
                } else if (kind == 3) {   // Material "glass" "color Kr" "color Kt" "float index" (materials/glass.cpp:97-110), roughness 0
                    auto texT = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs));
                    auto zero = std::make_shared<ConstantTexture<Float>>(0.f);
                    auto index = std::make_shared<ConstantTexture<Float>>(k[7]);
                    mm = std::make_shared<GlassMaterial>(tex, texT, zero, zero, index, nullptr, true);
                } else   // Material "matte" "rgb Kd" "float sigma" (sigma > 0: Oren-Nayar, materials/matte.cpp:55-61)
                    mm = std::make_shared<MatteMaterial>(tex, std::make_shared<ConstantTexture<Float>>(k[4]), nullptr);
                it = byKd.emplace(k, mm).first;
            }
            m = it->second;
        }
        std::shared_ptr<AreaLight> al;
        if (emitter_L && i + 1 == s->prims.size()) {
            // AttributeBegin Material "matte" Kd 0 ; AreaLightSource "area" L ; Shape "sphere" (scenes/killeroo-simple:18-24)
            Float c[3] = {0, 0, 0}, le[3] = {emitter_L[0], emitter_L[1], emitter_L[2]};
            auto tex = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(c));
            m = std::make_shared<MatteMaterial>(tex, sigma, nullptr);
            area = std::make_shared<DiffuseAreaLight>(*gp->shape->ObjectToWorld, mi, Spectrum::FromRGB(le), 1, gp->shape, false);
            al = area;
        }
        prims.push_back(std::make_shared<GeometricPrimitive>(gp->shape, m, al, mi));
    }
    ParamSet ps;
    std::shared_ptr<Primitive> accel;
    {
        ref_scene tmp;
        tmp.prims = prims;
        const char *none[] = {nullptr};
        ref_accel *a = ref_accel_create(&tmp, accel_kind, none, nullptr, none, nullptr);
        if (!a) return 1;
        accel = a->agg;
        delete a;
    }
    std::shared_ptr<Primitive> traced = accel;
    if (g_trace) {
        if (n_threads != 1) return 3;   // the trace is an ordered log: single-threaded renders only
        auto ta = std::make_shared<TracingAggregate>(accel);
        for (size_t i = 0; i < prims.size(); ++i) ta->index[prims[i].get()] = (int)i;
        traced = ta;
    }
    auto counting = std::make_shared<CountingAggregate>(traced);
    // LightSource "point" "point from" "rgb I" (lights/point.cpp:81-89)
    Float li[3] = {light_I[0], light_I[1], light_I[2]};
    Transform l2w = Translate(Vector3f(light_pos[0], light_pos[1], light_pos[2]));
    std::vector<std::shared_ptr<Light>> lights;
    if (n_threads != 1) ParallelInit();   // as pbrtInit does before any object is made (InfiniteAreaLight's constructor runs a ParallelFor)
    if (area) lights.push_back(area);
    else if (g_infinite) {
        Matrix4x4 m(g_infM[0], g_infM[1], g_infM[2], g_infM[3], g_infM[4], g_infM[5], g_infM[6], g_infM[7], g_infM[8], g_infM[9],
                    g_infM[10], g_infM[11], g_infM[12], g_infM[13], g_infM[14], g_infM[15]);
        Float lrgb[3] = {g_infL[0], g_infL[1], g_infL[2]};
        lights.push_back(std::make_shared<InfiniteAreaLight>(Transform(m), Spectrum::FromRGB(lrgb), 1, std::string("")));
    } else if (g_distant)
        lights.push_back(std::make_shared<DistantLight>(Transform(), Spectrum::FromRGB(li),
                                                        Point3f(g_from_to[0], g_from_to[1], g_from_to[2]) -
                                                            Point3f(g_from_to[3], g_from_to[4], g_from_to[5])));
    else lights.push_back(std::make_shared<PointLight>(l2w, mi, Spectrum::FromRGB(li)));
    for (size_t k = 0; k + 24 <= g_extraLights.size(); k += 24) {
        const float *r = g_extraLights.data() + k;
        Float c[3] = {r[4], r[5], r[6]};
        const int kind = (int)r[0];
        if (kind == 0) lights.push_back(std::make_shared<PointLight>(Translate(Vector3f(r[1], r[2], r[3])), mi, Spectrum::FromRGB(c)));
        else if (kind == 2) lights.push_back(std::make_shared<DistantLight>(Transform(), Spectrum::FromRGB(c), Vector3f(r[1], r[2], r[3])));
        else if (kind == 3) {
            const float *q = r + 7;
            Matrix4x4 m(q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7], q[8], q[9], q[10], q[11], q[12], q[13], q[14], q[15]);
            lights.push_back(std::make_shared<InfiniteAreaLight>(Transform(m), Spectrum::FromRGB(c), 1, std::string("")));
        } else return 4;
    }
    Scene scene(counting, lights);
    g_lastPowers.clear();
    for (const auto &l : scene.lights) g_lastPowers.push_back(l->Power().y());

    // Film "image" + PixelFilter "box" (film.cpp:307-351, filters/box.cpp) — defaults of every thesis template
    std::unique_ptr<Filter> filter(new BoxFilter(Vector2f(0.5f, 0.5f)));
    std::string fname = std::string(tmp_prefix) + ".exr";
    Film *film = new Film(Point2i(xres, yres), Bounds2f(Point2f(g_crop[0], g_crop[2]), Point2f(g_crop[1], g_crop[3])),
                          std::move(filter), 35.f, fname, 1.f);
    const Vector2i cres = film->croppedPixelBounds.Diagonal();
    // LookAt + Camera "perspective" (api.cpp:1338-1370 pbrtCamera; cameras/perspective.cpp:231-279)
    Transform w2c = LookAt(Point3f(cam[0], cam[1], cam[2]), Point3f(cam[3], cam[4], cam[5]), Vector3f(cam[6], cam[7], cam[8]));
    if (g_camMatrix)
        w2c = Transform(Matrix4x4(g_w2c[0], g_w2c[1], g_w2c[2], g_w2c[3], g_w2c[4], g_w2c[5], g_w2c[6], g_w2c[7], g_w2c[8], g_w2c[9],
                                  g_w2c[10], g_w2c[11], g_w2c[12], g_w2c[13], g_w2c[14], g_w2c[15]));
    Transform c2w = Inverse(w2c);
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    Float frame = Float(xres) / Float(yres);
    Bounds2f screen;
    if (frame > 1.f) { screen.pMin = Point2f(-frame, -1.f); screen.pMax = Point2f(frame, 1.f); }
    else { screen.pMin = Point2f(-1.f, -1.f / frame); screen.pMax = Point2f(1.f, 1.f / frame); }
    std::shared_ptr<const Camera> camera(new PerspectiveCamera(ac2w, screen, 0.f, 1.f, 0.f, 1e6f, cam[9], film, nullptr));
    std::shared_ptr<Sampler> sampler;
    if (std::string(sampler_name) == "halton") {
        if (g_trace) sampler.reset(new TracingSampler<HaltonSampler>(spp, film->GetSampleBounds()));
        else sampler.reset(new HaltonSampler(spp, film->GetSampleBounds()));
    } else if (std::string(sampler_name) == "sobol") {
        if (g_trace) sampler.reset(new TracingSampler<SobolSampler>(spp, film->GetSampleBounds()));
        else sampler.reset(new SobolSampler(spp, film->GetSampleBounds()));
    } else sampler.reset(new RandomSampler(spp));
    PathIntegrator integ(maxdepth, camera, sampler, film->GetSampleBounds(), 1.f, g_lightStrategy);

    const auto t0 = std::chrono::steady_clock::now();
    integ.Render(scene);     // -> film->WriteImage() -> the WriteImage above
    if (render_seconds) *render_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (n_threads != 1) ParallelCleanup();
    {
        std::lock_guard<std::mutex> l(g_imgMutex);
        if (g_lastRes.x != cres.x || g_lastRes.y != cres.y) return 2;
        std::memcpy(out_rgb, g_lastImage.data(), sizeof(float) * g_lastImage.size());
    }
    if (out_heat) {
        // the thesis's per-pixel GeneralStats (integrator.cpp:327-328), the matrices Film::WriteGeneralStats dumps
        // (film.cpp:170-264): planes = primitiveIntersections, ...P, {bvh+kd+bsp}TreeNodeTraversals, ...P,
        // leafNodeTraversals, ...P
        const size_t npx = (size_t)cres.x * cres.y;
        for (size_t i = 0; i < npx; ++i) {
            const GeneralStats &g = film->pixels[i].stats;
            out_heat[0 * npx + i] = (uint32_t)g.primitiveIntersections;
            out_heat[1 * npx + i] = (uint32_t)g.primitiveIntersectionsP;
            out_heat[2 * npx + i] = (uint32_t)(g.bvhTreeNodeTraversals + g.kdTreeNodeTraversals + g.bspTreeNodeTraversals);
            out_heat[3 * npx + i] = (uint32_t)(g.bvhTreeNodeTraversalsP + g.kdTreeNodeTraversalsP + g.bspTreeNodeTraversalsP);
            out_heat[4 * npx + i] = (uint32_t)g.leafNodeTraversals;
            out_heat[5 * npx + i] = (uint32_t)g.leafNodeTraversalsP;
        }
    }
    uint64_t tc, ta;
    counting->totals(&tc, &ta);
    if (n_closest) *n_closest = tc;
    if (n_shadow) *n_shadow = ta;
    // film is owned (and deleted) by the Camera (core/camera.cpp)
    return 0;
}
