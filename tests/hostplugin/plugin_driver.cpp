// Test driver for the C++ drop-in (thesis-pbrt-v3_b200/host/b200accel.cpp): creates `Accelerator "b200"` over
// the reference's own Primitives and exposes it behind the same ref_accel handle the reference accelerators
// use in oracle/ref/ref_scene.cpp, so ONE harness (ref_intersect / ref_intersect_p, which only see the
// Primitive interface) drives both sides.  TEST INFRASTRUCTURE ONLY.
#include "std_first.h"
#include "ref_internal.h"
#include "core/primitive.h"
#include "core/paramset.h"
#include "core/interaction.h"
#include "b200accel.h"

using namespace pbrt;

extern "C" ref_accel *plugin_accel_create(ref_scene *s, const char *tree, const char *builder,
                                          const char *const *int_names, const int *int_vals,
                                          const char *const *float_names, const float *float_vals) {
    ParamSet ps;
    for (int i = 0; int_names && int_names[i]; ++i) {
        std::unique_ptr<int[]> v(new int[1]);
        v[0] = int_vals[i];
        ps.AddInt(int_names[i], std::move(v), 1);
    }
    for (int i = 0; float_names && float_names[i]; ++i) {
        std::unique_ptr<Float[]> v(new Float[1]);
        v[0] = float_vals[i];
        ps.AddFloat(float_names[i], std::move(v), 1);
    }
    std::unique_ptr<std::string[]> t(new std::string[1]), b(new std::string[1]);
    t[0] = tree; b[0] = builder;
    ps.AddString("tree", std::move(t), 1);
    ps.AddString("builder", std::move(b), 1);
    // what MakeAccelerator's new branch does (api.cpp:790-1015)
    std::shared_ptr<Primitive> acc = CreateB200Accelerator(s->prims, ps);
    if (!acc) return nullptr;
    ref_accel *a = new ref_accel;
    a->agg = acc;
    a->scene = s;
    return a;
}

extern "C" int plugin_is_current(const ref_accel *a) { return B200Accel::current() == a->agg.get(); }

extern "C" void plugin_world_bound(const ref_accel *a, float *out6) {
    Bounds3f b = a->agg->WorldBound();
    out6[0] = b.pMin.x; out6[1] = b.pMin.y; out6[2] = b.pMin.z; out6[3] = b.pMax.x; out6[4] = b.pMax.y; out6[5] = b.pMax.z;
}

// Batched path: B200Accel::IntersectBatch + FillInteraction -> the same ref_hit records ref_intersect writes.
extern "C" int plugin_intersect_batch(const ref_accel *a, const float *rays, uint64_t n, ref_hit *hits) {
    const B200Accel *acc = static_cast<const B200Accel *>(a->agg.get());
    std::vector<Ray> rs;
    rs.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        rs.emplace_back(Point3f(r[0], r[1], r[2]), Vector3f(r[4], r[5], r[6]), r[3], r[7]);
    }
    std::vector<b200rt_hit> dh(n);
    std::vector<Float> tIn(n);
    for (uint64_t i = 0; i < n; ++i) tIn[i] = rs[i].tMax;
    if (!acc->IntersectBatch(rs.data(), n, dh.data())) return 1;
    for (uint64_t i = 0; i < n; ++i) {
        ref_hit h;
        std::memset(&h, 0, sizeof(h));
        h.prim = -1;
        h.t = rs[i].tMax;
        if (dh[i].prim != B200RT_MISS) {
            SurfaceInteraction isect;
            rs[i].tMax = tIn[i];
            if (!acc->FillInteraction(rs[i], dh[i], &isect)) return 2;
            auto it = a->scene->index.find(isect.primitive);
            h.prim = it == a->scene->index.end() ? -2 : it->second;
            h.t = rs[i].tMax;
            h.uv[0] = isect.uv.x; h.uv[1] = isect.uv.y;
            h.p[0] = isect.p.x; h.p[1] = isect.p.y; h.p[2] = isect.p.z;
        }
        hits[i] = h;
    }
    return 0;
}

extern "C" int plugin_intersect_p_batch(const ref_accel *a, const float *rays, uint64_t n, uint8_t *occ) {
    const B200Accel *acc = static_cast<const B200Accel *>(a->agg.get());
    std::vector<Ray> rs;
    rs.reserve(n);
    for (uint64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        rs.emplace_back(Point3f(r[0], r[1], r[2]), Vector3f(r[4], r[5], r[6]), r[3], r[7]);
    }
    return acc->IntersectPBatch(rs.data(), n, occ) ? 0 : 1;
}

// ---- Integrator "b200wavefront" (thesis-pbrt-v3_b200/host/b200integrator.cpp) ------------------------------------
// Same programmatic scene as oracle/ref/ref_render.cpp (matte material, one point light, perspective camera, box
// filter), but the aggregate is Accelerator "b200" and the integrator is the C++ wavefront plugin.  The image comes
// back through Film::WriteImage like the reference's.
#include "core/camera.h"
#include "core/film.h"
#include "core/scene.h"
#include "core/spectrum.h"
#include "cameras/perspective.h"
#include "filters/box.h"
#include "lights/point.h"
#include "materials/matte.h"
#include "materials/plastic.h"
#include "materials/mirror.h"
#include "materials/glass.h"
#include "materials/metal.h"
#include "materials/substrate.h"
#include "materials/uber.h"
#include "textures/imagemap.h"
#include "core/paramset.h"
#include "lights/diffuse.h"
#include "lights/distant.h"
#include "core/sampling.h"
#include "core/parallel.h"
#include "lights/infinite.h"
#include "samplers/random.h"
#include "samplers/halton.h"
#include "samplers/sobol.h"
#include "textures/constant.h"
#include "b200integrator.h"
extern "C" int ref_last_image(float *out_rgb, int *xres, int *yres);
extern "C" int ref_render_get_distant(float *from_to);
extern "C" int ref_render_get_infinite(float *L_rgb, float *l2w16);      // set through ref_render_set_infinite
extern "C" int ref_render_get_camera_matrix(float *w2c16);             // set through ref_render_set_camera_matrix
extern "C" int ref_render_get_textures(int i, char *out, int cap);   // set through ref_render_set_textures: image files of slots 18 / 19
extern "C" int ref_render_get_extra_lights(float *recs, int cap, char *strategy, int strategy_cap);   // set through ref_render_set_extra_lights
extern "C" void ref_render_get_texture_filter(int *trilinear, float *maxaniso);
extern "C" int ref_render_get_mat_stride(void);             // set through ref_render_set_mat_stride: 8 or 20 floats per record
extern "C" void ref_render_get_crop(float *cw);           // set through ref_render_set_crop   // set through ref_render_set_distant (oracle/ref/ref_render.cpp)
std::shared_ptr<pbrt::Shape> shape_of(const pbrt::Primitive *p);   // plugin_shape_of.cpp (GeometricPrimitive::shape is private)

// "bool generalstats" of the NEXT plugin_render calls (the thesis's per-pixel statistic matrices, Film::WriteGeneralStats)
static bool g_generalStats = false;
static std::string g_samplerName = "halton";
extern "C" void plugin_set_sampler(const char *name) { g_samplerName = name ? name : "halton"; }
extern "C" void plugin_set_generalstats(int on) { g_generalStats = on != 0; }

extern "C" int plugin_render(ref_scene *s, const char *tree, const float *cam, int xres, int yres, int spp, int maxdepth,
                             const float *light_pos, const float *light_I, const float *kd, int seed, int rank, int world,
                             const char *tmp_prefix, float *out_rgb, uint64_t *n_closest, uint64_t *n_shadow,
                             const float *prim_mat /* nullable: 8 floats per primitive, as in ref_render */,
                             const float *emitter_L /* nullable: the LAST primitive (a sphere) is an area light */) {
    PbrtOptions.quiet = true;
    Float rgb[3] = {kd[0], kd[1], kd[2]};
    auto Kd = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(rgb));
    auto sigma = std::make_shared<ConstantTexture<Float>>(0.f);
    auto matte = std::make_shared<MatteMaterial>(Kd, sigma, nullptr);
    MediumInterface mi;
    std::vector<std::shared_ptr<Primitive>> prims;
    std::shared_ptr<AreaLight> area;
    std::map<std::array<float, 20>, std::shared_ptr<Material>> byRec;
    const int stride = ref_render_get_mat_stride();
    for (size_t i = 0; i < s->prims.size(); ++i) {
        std::shared_ptr<Shape> shape = shape_of(s->prims[i].get());
        std::shared_ptr<Material> m = matte;
        if (prim_mat) {
            std::array<float, 20> k{};
            std::memcpy(k.data(), prim_mat + (size_t)stride * i, sizeof(float) * stride);
            auto it = byRec.find(k);
            if (it == byRec.end()) {
                Float c[3] = {k[1], k[2], k[3]}, cs[3] = {k[4], k[5], k[6]};
                std::shared_ptr<Texture<Spectrum>> tex = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(c));
                auto imagemap = [](int idx) -> std::shared_ptr<Texture<Spectrum>> {   // `Texture ... "imagemap" "bool trilinear" "true"`
                    char file[1024];
                    if (idx < 1 || idx > ref_render_get_textures(idx - 1, file, sizeof file)) return nullptr;
                    ParamSet ps;
                    std::unique_ptr<std::string[]> f(new std::string[1]);
                    f[0] = file;
                    ps.AddString("filename", std::move(f), 1);
                    std::unique_ptr<bool[]> tri(new bool[1]);
                    int triFlag = 1;
                    float maxAniso = 8.f;
                    ref_render_get_texture_filter(&triFlag, &maxAniso);   // set through ref_render_set_texture_filter
                    tri[0] = triFlag != 0;
                    ps.AddBool("trilinear", std::move(tri), 1);
                    std::unique_ptr<Float[]> ma(new Float[1]);
                    ma[0] = maxAniso;
                    ps.AddFloat("maxanisotropy", std::move(ma), 1);
                    std::map<std::string, std::shared_ptr<Texture<Float>>> ft;
                    std::map<std::string, std::shared_ptr<Texture<Spectrum>>> st;
                    TextureParams tp(ps, ps, ft, st);
                    const int nThreads = PbrtOptions.nThreads;
                    PbrtOptions.nThreads = 1;   // MIPMap's ParallelFor loops run inline: the pool is not up yet
                    std::shared_ptr<Texture<Spectrum>> t(CreateImageSpectrumTexture(Transform(), tp));
                    PbrtOptions.nThreads = nThreads;
                    return t;
                };
                if (stride == 20 && k[18] > 0) if (auto t = imagemap((int)k[18])) tex = t;
                std::shared_ptr<Material> mm;
                const int kind = (int)k[0] & 15;
                const bool remap = !((int)k[0] & 32);
                auto texS = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs));
                auto rough = std::make_shared<ConstantTexture<Float>>(k[7]);
                if (kind == 1) mm = std::make_shared<PlasticMaterial>(tex, texS, rough, nullptr, remap);
                else if (kind == 4) mm = std::make_shared<MetalMaterial>(tex, texS, rough, nullptr, nullptr, nullptr, remap);
                else if (kind == 5) mm = std::make_shared<SubstrateMaterial>(tex, texS, rough, rough, nullptr, remap);
                else if (kind == 2) mm = std::make_shared<MirrorMaterial>(tex, nullptr);
                else if (kind == 6) {   // Material "uber": Kr 8-10, Kt 11-13, opacity 14-16, eta 17
                    Float cr[3] = {k[8], k[9], k[10]}, ct[3] = {k[11], k[12], k[13]}, co[3] = {k[14], k[15], k[16]};
                    std::shared_ptr<Texture<Spectrum>> texO = std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(co));
                    if (k[19] > 0) if (auto t = imagemap((int)k[19])) texO = t;
                    mm = std::make_shared<UberMaterial>(tex, texS, std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cr)),
                                                        std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(ct)), rough, nullptr, nullptr,
                                                        texO,
                                                        std::make_shared<ConstantTexture<Float>>(k[17]), nullptr, remap);
                } else if (kind == 3)
                    mm = std::make_shared<GlassMaterial>(tex, std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(cs)),
                                                         std::make_shared<ConstantTexture<Float>>(0.f), std::make_shared<ConstantTexture<Float>>(0.f),
                                                         std::make_shared<ConstantTexture<Float>>(k[7]), nullptr, true);
                else mm = std::make_shared<MatteMaterial>(tex, std::make_shared<ConstantTexture<Float>>(k[4]), nullptr);
                it = byRec.emplace(k, mm).first;
            }
            m = it->second;
        }
        std::shared_ptr<AreaLight> al;
        if (emitter_L && i + 1 == s->prims.size()) {
            Float z[3] = {0, 0, 0}, le[3] = {emitter_L[0], emitter_L[1], emitter_L[2]};
            m = std::make_shared<MatteMaterial>(std::make_shared<ConstantTexture<Spectrum>>(Spectrum::FromRGB(z)), sigma, nullptr);
            area = std::make_shared<DiffuseAreaLight>(*shape->ObjectToWorld, mi, Spectrum::FromRGB(le), 1, shape, false);
            al = area;
        }
        prims.push_back(std::make_shared<GeometricPrimitive>(shape, m, al, mi));
    }
    ParamSet aps;
    std::unique_ptr<std::string[]> t(new std::string[1]);
    t[0] = tree;
    aps.AddString("tree", std::move(t), 1);
    std::shared_ptr<Primitive> accel = CreateB200Accelerator(prims, aps);
    if (!accel) return 1;
    Float li[3] = {light_I[0], light_I[1], light_I[2]};
    Transform l2w = Translate(Vector3f(light_pos[0], light_pos[1], light_pos[2]));
    std::vector<std::shared_ptr<Light>> lights;
    float ft[6];
    float infL[3], m16[16];
    if (area) lights.push_back(area);
    else if (ref_render_get_infinite(infL, m16)) {   // LightSource "infinite" with the constant map (no readable "mapname")
        Float lrgb[3] = {infL[0], infL[1], infL[2]};
        ParallelInit();   // InfiniteAreaLight's constructor runs a ParallelFor (pbrtInit starts the pool before any object is made)
        lights.push_back(std::make_shared<InfiniteAreaLight>(
            Transform(Matrix4x4(m16[0], m16[1], m16[2], m16[3], m16[4], m16[5], m16[6], m16[7], m16[8], m16[9], m16[10], m16[11],
                                m16[12], m16[13], m16[14], m16[15])), Spectrum::FromRGB(lrgb), 1, std::string("")));
        ParallelCleanup();
    } else if (ref_render_get_distant(ft))
        lights.push_back(std::make_shared<DistantLight>(Transform(), Spectrum::FromRGB(li),
                                                        Point3f(ft[0], ft[1], ft[2]) - Point3f(ft[3], ft[4], ft[5])));
    else lights.push_back(std::make_shared<PointLight>(l2w, mi, Spectrum::FromRGB(li)));
    // more LightSource statements (24-float records, oracle/ref/ref_render.cpp) and Integrator "string lightsamplestrategy"
    std::vector<float> extra((size_t)24 * 64);
    char strategy[32] = "spatial";
    const int nExtra = ref_render_get_extra_lights(extra.data(), 64, strategy, sizeof strategy);
    for (int k = 0; k < nExtra && k < 64; ++k) {
        const float *r = extra.data() + (size_t)24 * k;
        Float c[3] = {r[4], r[5], r[6]};
        const int kind = (int)r[0];
        if (kind == 0) lights.push_back(std::make_shared<PointLight>(Translate(Vector3f(r[1], r[2], r[3])), mi, Spectrum::FromRGB(c)));
        else if (kind == 2) lights.push_back(std::make_shared<DistantLight>(Transform(), Spectrum::FromRGB(c), Vector3f(r[1], r[2], r[3])));
        else if (kind == 3) {
            const float *q = r + 7;
            ParallelInit();
            lights.push_back(std::make_shared<InfiniteAreaLight>(
                Transform(Matrix4x4(q[0], q[1], q[2], q[3], q[4], q[5], q[6], q[7], q[8], q[9], q[10], q[11], q[12], q[13], q[14], q[15])),
                Spectrum::FromRGB(c), 1, std::string("")));
            ParallelCleanup();
        } else return 4;
    }
    Scene scene(accel, lights);
    std::unique_ptr<Filter> filter(new BoxFilter(Vector2f(0.5f, 0.5f)));
    float cw[4];
    ref_render_get_crop(cw);
    Film *film = new Film(Point2i(xres, yres), Bounds2f(Point2f(cw[0], cw[2]), Point2f(cw[1], cw[3])), std::move(filter), 35.f,
                          std::string(tmp_prefix) + ".exr", 1.f);
    const Vector2i cres = film->croppedPixelBounds.Diagonal();
    Transform w2c = LookAt(Point3f(cam[0], cam[1], cam[2]), Point3f(cam[3], cam[4], cam[5]), Vector3f(cam[6], cam[7], cam[8]));
    if (ref_render_get_camera_matrix(m16))   // `Transform [...]` before `Camera` (scenes/living-room/scene.pbrt:2)
        w2c = Transform(Matrix4x4(m16[0], m16[1], m16[2], m16[3], m16[4], m16[5], m16[6], m16[7], m16[8], m16[9], m16[10], m16[11],
                                  m16[12], m16[13], m16[14], m16[15]));
    Transform c2w = Inverse(w2c);
    AnimatedTransform ac2w(&c2w, 0.f, &c2w, 1.f);
    Float frame = Float(xres) / Float(yres);
    Bounds2f screen;
    if (frame > 1.f) { screen.pMin = Point2f(-frame, -1.f); screen.pMax = Point2f(frame, 1.f); }
    else { screen.pMin = Point2f(-1.f, -1.f / frame); screen.pMax = Point2f(1.f, 1.f / frame); }
    std::shared_ptr<const Camera> camera(new PerspectiveCamera(ac2w, screen, 0.f, 1.f, 0.f, 1e6f, cam[9], film, nullptr));
    // Sampler "halton" (every thesis template's sampler) unless a test asks for another one: the C++ integrator then runs the
    // reference's own Halton stream on the device, exactly what the Python mirror's default does
    std::shared_ptr<Sampler> sampler;
    if (g_samplerName == "halton") sampler.reset(new HaltonSampler(spp, film->GetSampleBounds()));
    else if (g_samplerName == "sobol") sampler.reset(new SobolSampler(spp, film->GetSampleBounds()));
    else sampler.reset(new RandomSampler(spp));
    // what MakeIntegrator's new branch does (api.cpp:1895-1945)
    ParamSet ips;
    auto addInt = [&](const char *name, int v) { std::unique_ptr<int[]> a(new int[1]); a[0] = v; ips.AddInt(name, std::move(a), 1); };
    addInt("maxdepth", maxdepth); addInt("seed", seed); addInt("rank", rank); addInt("world", world);
    if (g_generalStats) { std::unique_ptr<bool[]> b(new bool[1]); b[0] = true; ips.AddBool("generalstats", std::move(b), 1); }
    if (nExtra > 0) { std::unique_ptr<std::string[]> v(new std::string[1]); v[0] = strategy; ips.AddString("lightsamplestrategy", std::move(v), 1); }
    std::unique_ptr<B200WavefrontIntegrator> integ(CreateB200WavefrontIntegrator(ips, sampler, camera));
    integ->Render(scene);
    int w = 0, h = 0;
    ref_last_image(nullptr, &w, &h);
    if (w != cres.x || h != cres.y) return 2;
    ref_last_image(out_rgb, &w, &h);
    if (n_closest) *n_closest = integ->stats().closest_rays;
    if (n_shadow) *n_shadow = integ->stats().shadow_rays;
    return 0;
}
