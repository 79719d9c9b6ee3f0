// b200integrator.cpp — see b200integrator.h.  Host-side glue only: the frame is rendered by b200rt_render on the B200.
#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <utility>
#include <vector>
#include <glog/logging.h>

// The point-light / matte case reads the scene through pbrt's public interfaces only.  The sphere area light and
// "plastic" need three things that have no accessor (inside the reference tree: `friend class B200WavefrontIntegrator;`):
// GeometricPrimitive::shape + Sphere::radius (the emitter's geometry) and PlasticMaterial::Kd/Ks/roughness (likewise
// MirrorMaterial::Kr, GlassMaterial::Kr/Kt/index/uRoughness/vRoughness, MetalMaterial::eta/k/roughness and
// SubstrateMaterial::Kd/Ks/nu/nv).
#define private public
#define protected public
#include "core/primitive.h"
#include "shapes/sphere.h"
#include "materials/plastic.h"
#include "materials/matte.h"
#include "materials/mirror.h"
#include "materials/glass.h"
#include "materials/metal.h"
#include "materials/substrate.h"
#include "materials/uber.h"
#include "core/texture.h"         // UVMapping2D::su / sv / du / dv
#include "core/mipmap.h"          // MIPMap::pyramid / doTrilinear / wrapMode
#include "textures/imagemap.h"    // ImageTexture::mapping / mipmap
#include "core/film.h"   // Film::pixels / Pixel::stats for "bool generalstats" (no accessor either)
#include "core/transform.h"       // AnimatedTransform::startTransform
#include "core/camera.h"          // ProjectiveCamera::RasterToCamera / lensRadius
#include "cameras/perspective.h"
#include "core/sampler.h"
#include "samplers/halton.h"      // HaltonSampler::sampleAtPixelCenter
#include "samplers/sobol.h"
#undef private
#undef protected
#include "core/sobolmatrices.h"

#include "b200integrator.h"
#include "b200accel.h"
#include "camera.h"
#include "film.h"
#include "filter.h"
#include "interaction.h"
#include "light.h"
#include "material.h"
#include "memory.h"
#include "paramset.h"
#include "reflection.h"
#include "sampler.h"
#include "scene.h"

namespace pbrt {

namespace {

// The pinhole frame of a perspective camera, recovered through Camera::GenerateRay (cameras/perspective.cpp:97-143):
// d(fx, fy) = dir + (2 fx - 1) right + (1 - 2 fy) up for film coordinates fx, fy in [0,1].
bool describeCamera(const Camera &cam, b200rt_camera *out) {
    const Point2i res = cam.film->fullResolution;
    auto dirAt = [&](Float x, Float y, Point3f *o) {
        CameraSample cs;
        cs.pFilm = Point2f(x, y);
        cs.pLens = Point2f(0.5f, 0.5f);
        cs.time = 0;
        Ray r;
        cam.GenerateRay(cs, &r);
        if (o) *o = r.o;
        return Normalize(r.d);
    };
    Point3f eye, eye2;
    const Vector3f dc = dirAt(res.x * 0.5f, res.y * 0.5f, &eye);
    const Vector3f d00 = dirAt(0, 0, &eye2), d10 = dirAt((Float)res.x, 0, nullptr), d01 = dirAt(0, (Float)res.y, nullptr);
    if ((eye - eye2).Length() > 1e-6f * (1 + (eye - Point3f(0, 0, 0)).Length())) {
        Error("b200wavefront: only pinhole perspective cameras (lensradius 0) are supported");
        return false;
    }
    // put the corner directions on the plane at distance 1 along the view direction
    const Vector3f v00 = d00 / Dot(d00, dc), v10 = d10 / Dot(d10, dc), v01 = d01 / Dot(d01, dc);
    const Vector3f right = (v10 - v00) * 0.5f, up = (v00 - v01) * 0.5f;
    const Float e[3] = {eye.x, eye.y, eye.z}, d[3] = {dc.x, dc.y, dc.z}, r[3] = {right.x, right.y, right.z}, u[3] = {up.x, up.y, up.z};
    std::memcpy(out->eye, e, 12); std::memcpy(out->dir, d, 12); std::memcpy(out->right, r, 12); std::memcpy(out->up, u, 12);
    out->xres = (uint32_t)res.x; out->yres = (uint32_t)res.y;
    return true;
}

// Kd of a Lambertian material, measured through the public BSDF: f(n, n) * Pi (reflection.h LambertianReflection).
bool lambertianKd(const Material *m, Float kd[3]) {
    MemoryArena arena;
    const Normal3f n(0, 0, 1);
    SurfaceInteraction si(Point3f(0, 0, 0), Vector3f(0, 0, 0), Point2f(0.5f, 0.5f), Vector3f(0, 0, 1), Vector3f(1, 0, 0),
                          Vector3f(0, 1, 0), Normal3f(0, 0, 0), Normal3f(0, 0, 0), 0, nullptr);
    m->ComputeScatteringFunctions(&si, arena, TransportMode::Radiance, true);
    if (!si.bsdf) return false;
    const int all = si.bsdf->NumComponents(BSDF_ALL);
    if (all == 0) { kd[0] = kd[1] = kd[2] = 0; return true; }   // Material "matte" Kd 0 adds no lobe (matte.cpp:52-58)
    if (all != 1 || si.bsdf->NumComponents(BxDFType(BSDF_REFLECTION | BSDF_DIFFUSE)) != 1) return false;
    const Vector3f w(0, 0, 1);
    Float rgb[3];
    Spectrum(si.bsdf->f(w, w) * Pi).ToRGB(rgb);
    // a Lambertian lobe is direction-independent; an Oren-Nayar one (sigma > 0) is not
    Float rgb2[3];
    Spectrum(si.bsdf->f(w, Normalize(Vector3f(0.6f, 0.3f, 0.5f))) * Pi).ToRGB(rgb2);
    for (int i = 0; i < 3; ++i) {
        if (std::abs(rgb[i] - rgb2[i]) > 1e-6f) return false;
        kd[i] = rgb[i];
    }
    return true;
}

// One light of the scene as a b200rt_light: LightSource "point" (through Light::Sample_Li, lights/point.cpp:44-53), LightSource
// "distant", AreaLightSource "area" on a full Shape "sphere" that is a primitive of the aggregate (lights/diffuse.cpp,
// scenes/killeroo-simple:18-24) or a constant LightSource "infinite".
bool describeLight(const Scene &scene, const B200Accel *accel, const Light *light, b200rt_light *out) {
    std::memset(out, 0, sizeof(*out));
    if (light->flags & (int)LightFlags::DeltaPosition) {
        Interaction ref(Point3f(0, 0, 0), Normal3f(0, 0, 1), Vector3f(0, 0, 0), Vector3f(0, 0, 1), 0, MediumInterface());
        Vector3f wi;
        Float pdf;
        VisibilityTester vis;
        const Spectrum Li = light->Sample_Li(ref, Point2f(0.5f, 0.5f), &wi, &pdf, &vis);
        const Point3f pl = vis.P1().p;
        const Float d2 = (pl - Point3f(0, 0, 0)).LengthSquared();
        Float rgb[3];
        Spectrum(Li * d2).ToRGB(rgb);
        out->pos[0] = pl.x; out->pos[1] = pl.y; out->pos[2] = pl.z;
        for (int i = 0; i < 3; ++i) out->I[i] = rgb[i];
    } else if (light->flags & (int)LightFlags::DeltaDirection) {
        // LightSource "distant" (lights/distant.cpp:49-59): wi = wLight, Li = L for every reference point; shadow rays
        // end 2 * worldRadius away, worldRadius as DistantLight::Preprocess takes it from the scene's bound (distant.h)
        Interaction ref(Point3f(0, 0, 0), Normal3f(0, 0, 1), Vector3f(0, 0, 0), Vector3f(0, 0, 1), 0, MediumInterface());
        Vector3f wi;
        Float pdf;
        VisibilityTester vis;
        const Spectrum Li = light->Sample_Li(ref, Point2f(0.5f, 0.5f), &wi, &pdf, &vis);
        Point3f wc;
        Float wr;
        scene.WorldBound().BoundingSphere(&wc, &wr);
        Float rgb[3];
        Li.ToRGB(rgb);
        out->kind = B200RT_LIGHT_DISTANT;
        out->radius = wr;
        out->pos[0] = wi.x; out->pos[1] = wi.y; out->pos[2] = wi.z;
        for (int i = 0; i < 3; ++i) out->I[i] = rgb[i];
    } else if (light->flags & (int)LightFlags::Area) {
        int lp = -1;
        for (size_t i = 0; i < accel->primitiveCount(); ++i)
            if (accel->primitive((uint32_t)i)->GetAreaLight() == light) {
                if (lp >= 0) { Error("b200wavefront: the area light must sit on ONE sphere primitive"); return false; }
                lp = (int)i;
            }
        const GeometricPrimitive *gp = lp < 0 ? nullptr : dynamic_cast<const GeometricPrimitive *>(accel->primitive((uint32_t)lp).get());
        const Sphere *sp = gp ? dynamic_cast<const Sphere *>(gp->shape.get()) : nullptr;
        if (!sp) { Error("b200wavefront: the area light's shape must be a sphere inside the b200 aggregate"); return false; }
        const AreaLight *al = static_cast<const AreaLight *>(light);
        Interaction it;
        it.n = Normal3f(0, 0, 1);
        Float le[3];
        al->L(it, Vector3f(0, 0, 1)).ToRGB(le);
        if (!al->L(it, Vector3f(0, 0, -1)).IsBlack()) { Error("b200wavefront: two-sided area lights are not supported"); return false; }
        const Point3f c = (*sp->ObjectToWorld)(Point3f(0, 0, 0));
        out->kind = B200RT_LIGHT_SPHERE;
        out->radius = sp->radius;
        out->prim = (uint32_t)lp;
        out->pos[0] = c.x; out->pos[1] = c.y; out->pos[2] = c.z;
        for (int i = 0; i < 3; ++i) out->I[i] = le[i];
    } else if (light->flags & (int)LightFlags::Infinite) {
        // LightSource "infinite" (lights/infinite.cpp) with a CONSTANT map — what the reference builds when "mapname" is
        // absent or unreadable (scenes/living-room's sky map is not in the tree).  Read through the public interface:
        // with the 1x1 map Sample_Li(u) is (theta, phi) = (u[1] pi, u[0] 2 pi) -> LightToWorld's columns, pdf 1/(2 pi^2 sin theta).
        Point3f c; Float wr;
        scene.WorldBound().BoundingSphere(&c, &wr);
        Interaction ref(c, Normal3f(), Vector3f(), Vector3f(0, 0, 1), 0.f, MediumInterface());
        Vector3f col[3]; Float pdf; VisibilityTester vis;
        const Point2f us[3] = {Point2f(0.f, 0.5f), Point2f(0.25f, 0.5f), Point2f(0.f, 0.f)};
        Float le[3];
        for (int k = 0; k < 3; ++k) {
            const Spectrum Li = light->Sample_Li(ref, us[k], &col[k], &pdf, &vis);
            if (k == 0) Li.ToRGB(le);
            if (k < 2 && std::abs(pdf * 2 * Pi * Pi - 1) > 1e-4f) { Error("b200wavefront: only a constant infinite light (no usable \"mapname\") is supported"); return false; }
        }
        for (int k = 0; k < 26; ++k) {   // Le must not depend on the direction
            const Vector3f w = Normalize(Vector3f(Float(k % 3) - 1, Float(k / 3 % 3) - 1, Float(k / 9) - 1 + 0.01f));
            Float q[3];
            light->Le(RayDifferential(Ray(c, w))).ToRGB(q);
            for (int i = 0; i < 3; ++i)
                if (std::abs(q[i] - le[i]) > 1e-5f * (1 + std::abs(le[i]))) { Error("b200wavefront: only a constant infinite light is supported"); return false; }
        }
        out->kind = B200RT_LIGHT_INFINITE;
        out->radius = wr;
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) out->to_world[3 * r + k] = col[k][r];
        for (int i = 0; i < 3; ++i) out->I[i] = le[i];
    } else {
        Error("b200wavefront: only a point light, a distant light, a sphere area light or a constant infinite light is supported");
        return false;
    }
    return true;
}

}  // namespace

void B200WavefrontIntegrator::Render(const Scene &scene) {
    const B200Accel *accel = B200Accel::current();
    if (!accel || !accel->ok()) {
        Error("b200wavefront: the scene's aggregate is not Accelerator \"b200\"; nothing rendered");
        return;
    }
    Film *film = camera->film;
    b200rt_render_desc d;
    std::memset(&d, 0, sizeof(d));
    if (!describeCamera(*camera, &d.camera)) return;
    // The reference camera's own matrices: the device then generates PerspectiveCamera::GenerateRayDifferential's rays bit
    // for bit (cameras/perspective.cpp:97-143 with lensradius 0; csrc/wf_core.cuh camera_ray_ref)
    if (const PerspectiveCamera *pc = dynamic_cast<const PerspectiveCamera *>(camera.get())) {
        if (pc->lensRadius == 0 && !pc->CameraToWorld.actuallyAnimated) {
            std::memcpy(d.raster_to_camera, pc->RasterToCamera.GetMatrix().m, 64);
            std::memcpy(d.camera_to_world, pc->CameraToWorld.startTransform->GetMatrix().m, 64);
            d.camera_exact = 1;
        }
    }
    // The scene's Sampler: "halton" and "sobol" are restated on the device (same values, same dimension order as
    // PathIntegrator::Li); any other sampler's stream is not reproducible out of order -> the counter hash
    if (const HaltonSampler *hs = dynamic_cast<const HaltonSampler *>(sampler.get())) {
        if (hs->sampleAtPixelCenter) { Error("b200wavefront: Sampler \"halton\" \"bool samplepixelcenter\" is not supported"); return; }
        d.sampler = B200RT_SAMPLER_HALTON;
    } else if (dynamic_cast<const SobolSampler *>(sampler.get())) {
        d.sampler = B200RT_SAMPLER_SOBOL;
        d.sobol_matrices32 = SobolMatrices32;
        d.sobol_vdc = &VdCSobolMatrices[0][0];
        d.sobol_vdc_inv = &VdCSobolMatricesInv[0][0];
        d.sobol_vdc_rows = 25;
    } else {
        Warning("b200wavefront: only Sampler \"halton\" and \"sobol\" are reproduced sample for sample; using the counter-hash stream");
        d.sampler = B200RT_SAMPLER_HASH;
    }
    const Bounds2i cb = film->croppedPixelBounds;   // Film "cropwindow" (film.cpp:49-57); the film's pixels are the window's
    if (cb.pMin != Point2i(0, 0) || cb.pMax != film->fullResolution) {
        d.crop[0] = (uint32_t)cb.pMin.x; d.crop[1] = (uint32_t)cb.pMax.x;
        d.crop[2] = (uint32_t)cb.pMin.y; d.crop[3] = (uint32_t)cb.pMax.y;
    }
    if (film->filter->radius.x > 0.5f || film->filter->radius.y > 0.5f) {
        Error("b200wavefront: pixel filters wider than one pixel are not supported (use PixelFilter \"box\")");
        return;
    }
    // the lights (describeLight above), in Scene::lights order, and the distribution UniformSampleOneLight draws from
    if (scene.lights.empty()) { Error("b200wavefront: the scene has no light"); return; }
    if (scene.lights.size() > B200RT_MAX_LIGHTS) { Error("b200wavefront: %d lights, at most %d are supported", (int)scene.lights.size(), (int)B200RT_MAX_LIGHTS); return; }
    std::vector<b200rt_light> lights(scene.lights.size());
    for (size_t k = 0; k < scene.lights.size(); ++k) {
        if (!describeLight(scene, accel, scene.lights[k].get(), &lights[k])) return;
        lights[k].weight = 1.f;   // UniformLightDistribution (lightdistrib.cpp:68-71); also what one light gets under every strategy
    }
    if (lights.size() > 1) {
        // CreateLightSampleDistribution (lightdistrib.cpp:48-66): "uniform" and "power" are one Distribution1D for the whole
        // scene; "spatial" (the default, and what an unknown name falls back to) is one per voxel of the scene's bound, which the
        // device computes for every voxel before the frame (include/b200rt_wavefront.h light_strategy)
        if (lightSampleStrategy == "power") {
            for (size_t k = 0; k < lights.size(); ++k) lights[k].weight = scene.lights[k]->Power().y();   // ComputeLightPowerDistribution
        } else if (lightSampleStrategy != "uniform") {
            if (lightSampleStrategy != "spatial")
                Error("Light sample distribution type \"%s\" unknown. Using \"spatial\".", lightSampleStrategy.c_str());   // lightdistrib.cpp:59-64
            d.light_strategy = B200RT_LIGHTS_SPATIAL;
        }
        d.lights = lights.data();
        d.n_lights = (uint32_t)lights.size();
    }
    {   // one light: the description's own fields (every ABI version reads them)
        const b200rt_light &l = lights[0];
        d.light_kind = l.kind; d.light_prim = l.prim; d.light_radius = l.radius;
        std::memcpy(d.light_pos, l.pos, 12); std::memcpy(d.light_I, l.I, 12); std::memcpy(d.light_to_world, l.to_world, 36);
    }
    // materials: Material "matte" (constant Kd, sigma 0; measured through the public BSDF) and Material "plastic"
    // (constant Kd / Ks / roughness, remaproughness true), one record per distinct Material object
    std::vector<float> mats;
    std::vector<uint8_t> primMat(accel->primitiveCount(), 0);
    // `Texture ... "imagemap"` on a Kd or on uber's opacity: the pyramid of the reference's own MIPMap goes to the device as it is
    // (ABI v9); trilinear filtering, UV mapping and Repeat wrap only — what scenes/living-room and scenes/sponza ask for
    std::vector<std::vector<float>> texData;
    std::vector<b200rt_texture> texRecs;
    std::map<const void *, int> texSeen;
    bool texBad = false;
    auto textureIndex = [&](const std::shared_ptr<Texture<Spectrum>> &t) -> int {   // 1-based; 0 = not an image texture
        const auto *im = dynamic_cast<const ImageTexture<RGBSpectrum, Spectrum> *>(t.get());
        if (!im) return 0;
        auto it = texSeen.find(im);
        if (it != texSeen.end()) return it->second;
        const UVMapping2D *uv = dynamic_cast<const UVMapping2D *>(im->mapping.get());
        const MIPMap<RGBSpectrum> *mm = im->mipmap;
        if (!uv || !mm || mm->wrapMode != ImageWrap::Repeat || mm->Levels() > 16) {
            Error("b200wavefront: image textures are supported with \"uv\" mapping and \"repeat\" wrap only");
            texBad = true;
            return 0;
        }
        std::vector<float> px;
        for (int l = 0; l < mm->Levels(); ++l) {
            const BlockedArray<RGBSpectrum> &a = *mm->pyramid[l];
            for (int tt = 0; tt < a.vSize(); ++tt)
                for (int ss = 0; ss < a.uSize(); ++ss) {
                    Float rgb[3];
                    a(ss, tt).ToRGB(rgb);
                    px.insert(px.end(), rgb, rgb + 3);
                }
        }
        texData.push_back(std::move(px));
        b200rt_texture r;
        r.texels = nullptr;   // set below, once texData stops moving
        r.width = (uint32_t)mm->pyramid[0]->uSize(); r.height = (uint32_t)mm->pyramid[0]->vSize(); r.n_levels = (uint32_t)mm->Levels();
        r.su = uv->su; r.sv = uv->sv; r.du = uv->du; r.dv = uv->dv;
        r.filter = mm->doTrilinear ? B200RT_TEXFILTER_TRILINEAR : B200RT_TEXFILTER_EWA;   // "bool trilinear" (default false: EWA)
        r.max_anisotropy = mm->maxAnisotropy;
        texRecs.push_back(r);
        texSeen[im] = (int)texRecs.size();
        return (int)texRecs.size();
    };
    {
        // one record per distinct (Material object, orientation flag): glass needs to know on which side of a triangle the
        // shape's normal is (Shape::reverseOrientation ^ transformSwapsHandedness, triangle.cpp:421-424)
        std::map<std::pair<const Material *, bool>, int> seen;
        SurfaceInteraction si(Point3f(0, 0, 0), Vector3f(0, 0, 0), Point2f(0.5f, 0.5f), Vector3f(0, 0, 1), Vector3f(1, 0, 0),
                              Vector3f(0, 1, 0), Normal3f(0, 0, 0), Normal3f(0, 0, 0), 0, nullptr);
        for (size_t i = 0; i < accel->primitiveCount(); ++i) {
            const Material *m = accel->primitive((uint32_t)i)->GetMaterial();
            if (!m) { Error("b200wavefront: primitive %d has no material", (int)i); return; }
            bool flip = false;
            if (dynamic_cast<const GlassMaterial *>(m))
                if (const GeometricPrimitive *gpm = dynamic_cast<const GeometricPrimitive *>(accel->primitive((uint32_t)i).get()))
                    flip = gpm->shape->reverseOrientation ^ gpm->shape->transformSwapsHandedness;
            auto it = seen.find(std::make_pair(m, flip));
            if (it == seen.end()) {
                if (seen.size() >= 256) { Error("b200wavefront: more than 256 materials"); return; }
                float rec[20] = {0};
                if (const PlasticMaterial *pm = dynamic_cast<const PlasticMaterial *>(m)) {
                    if (pm->bumpMap) { Error("b200wavefront: plastic with a bump map is not supported"); return; }
                    Float kd[3], ks[3];
                    Spectrum(pm->Kd->Evaluate(si).Clamp()).ToRGB(kd);
                    Spectrum(pm->Ks->Evaluate(si).Clamp()).ToRGB(ks);
                    rec[0] = (float)(B200RT_MAT_PLASTIC + (pm->remapRoughness ? 0 : B200RT_MAT_NO_REMAP));
                    for (int c = 0; c < 3; ++c) { rec[1 + c] = kd[c]; rec[4 + c] = ks[c]; }
                    rec[7] = pm->roughness->Evaluate(si);
                    rec[18] = (float)textureIndex(pm->Kd);
                } else if (const MetalMaterial *me = dynamic_cast<const MetalMaterial *>(m)) {
                    const Float ur = me->uRoughness ? me->uRoughness->Evaluate(si) : me->roughness->Evaluate(si);
                    const Float vr = me->vRoughness ? me->vRoughness->Evaluate(si) : me->roughness->Evaluate(si);
                    if (me->bumpMap || ur != vr) { Error("b200wavefront: only isotropic metal without a bump map is supported"); return; }
                    Float e[3], kk[3];
                    me->eta->Evaluate(si).ToRGB(e);
                    me->k->Evaluate(si).ToRGB(kk);
                    rec[0] = (float)(B200RT_MAT_METAL + (me->remapRoughness ? 0 : B200RT_MAT_NO_REMAP));
                    for (int c = 0; c < 3; ++c) { rec[1 + c] = e[c]; rec[4 + c] = kk[c]; }
                    rec[7] = ur;
                } else if (const SubstrateMaterial *sm = dynamic_cast<const SubstrateMaterial *>(m)) {
                    if (sm->bumpMap || sm->nu->Evaluate(si) != sm->nv->Evaluate(si)) { Error("b200wavefront: only isotropic substrate without a bump map is supported"); return; }
                    Float kd[3], ks[3];
                    Spectrum(sm->Kd->Evaluate(si).Clamp()).ToRGB(kd);
                    Spectrum(sm->Ks->Evaluate(si).Clamp()).ToRGB(ks);
                    rec[0] = (float)(B200RT_MAT_SUBSTRATE + (sm->remapRoughness ? 0 : B200RT_MAT_NO_REMAP));
                    for (int c = 0; c < 3; ++c) { rec[1 + c] = kd[c]; rec[4 + c] = ks[c]; }
                    rec[7] = sm->nu->Evaluate(si);
                    rec[18] = (float)textureIndex(sm->Kd);
                } else if (const UberMaterial *um = dynamic_cast<const UberMaterial *>(m)) {
                    // Material "uber" (uber.cpp:45-100): constant textures only, isotropic roughness
                    if (um->bumpMap || um->roughnessu || um->roughnessv) { Error("b200wavefront: only isotropic uber without a bump map is supported"); return; }
                    Float kd[3], ks[3], kr[3], kt[3], op[3];
                    Spectrum(um->Kd->Evaluate(si).Clamp()).ToRGB(kd);
                    Spectrum(um->Ks->Evaluate(si).Clamp()).ToRGB(ks);
                    Spectrum(um->Kr->Evaluate(si).Clamp()).ToRGB(kr);
                    Spectrum(um->Kt->Evaluate(si).Clamp()).ToRGB(kt);
                    Spectrum(um->opacity->Evaluate(si).Clamp()).ToRGB(op);
                    rec[0] = (float)(B200RT_MAT_UBER + (um->remapRoughness ? 0 : B200RT_MAT_NO_REMAP));
                    for (int c = 0; c < 3; ++c) { rec[1 + c] = kd[c]; rec[4 + c] = ks[c]; rec[8 + c] = kr[c]; rec[11 + c] = kt[c]; rec[14 + c] = op[c]; }
                    rec[7] = um->roughness->Evaluate(si);
                    rec[17] = um->eta->Evaluate(si);
                    rec[18] = (float)textureIndex(um->Kd);
                    rec[19] = (float)textureIndex(um->opacity);
                } else if (const MirrorMaterial *mm = dynamic_cast<const MirrorMaterial *>(m)) {
                    if (mm->bumpMap) { Error("b200wavefront: mirror with a bump map is not supported"); return; }
                    Float kr[3];
                    Spectrum(mm->Kr->Evaluate(si).Clamp()).ToRGB(kr);
                    rec[0] = (float)B200RT_MAT_MIRROR;
                    for (int c = 0; c < 3; ++c) rec[1 + c] = kr[c];
                } else if (const GlassMaterial *gm = dynamic_cast<const GlassMaterial *>(m)) {
                    if (gm->bumpMap || gm->uRoughness->Evaluate(si) != 0 || gm->vRoughness->Evaluate(si) != 0) {
                        Error("b200wavefront: only smooth glass (roughness 0, no bump map) is supported");
                        return;
                    }
                    Float kr[3], kt[3];
                    Spectrum(gm->Kr->Evaluate(si).Clamp()).ToRGB(kr);
                    Spectrum(gm->Kt->Evaluate(si).Clamp()).ToRGB(kt);
                    rec[0] = (float)(B200RT_MAT_GLASS + (flip ? B200RT_MAT_REVERSE_ORIENTATION : 0));
                    for (int c = 0; c < 3; ++c) { rec[1 + c] = kr[c]; rec[4 + c] = kt[c]; }
                    rec[7] = gm->index->Evaluate(si);
                } else if (const MatteMaterial *mt = dynamic_cast<const MatteMaterial *>(m);
                           mt && !mt->bumpMap && (mt->sigma->Evaluate(si) != 0 || dynamic_cast<const ImageTexture<RGBSpectrum, Spectrum> *>(mt->Kd.get()))) {
                    // Material "matte" with "float sigma" > 0: an Oren-Nayar lobe (matte.cpp:55-61); sigma has no accessor.
                    // Or with an image texture on Kd
                    Float kd[3];
                    Spectrum(mt->Kd->Evaluate(si).Clamp()).ToRGB(kd);
                    for (int c = 0; c < 3; ++c) rec[1 + c] = kd[c];
                    rec[4] = Clamp(mt->sigma->Evaluate(si), 0, 90);
                    rec[18] = (float)textureIndex(mt->Kd);
                } else {
                    Float kd[3];
                    if (!lambertianKd(m, kd)) { Error("b200wavefront: only \"matte\", \"plastic\", \"mirror\", smooth \"glass\", \"metal\", \"substrate\" and \"uber\" materials are supported"); return; }
                    for (int c = 0; c < 3; ++c) rec[1 + c] = kd[c];
                }
                if (texBad) return;
                it = seen.emplace(std::make_pair(m, flip), (int)seen.size()).first;
                mats.insert(mats.end(), rec, rec + 20);
            }
            primMat[i] = (uint8_t)it->second;
        }
        if (mats.size() == 20 && mats[0] == 0.f && mats[4] == 0.f && mats[18] == 0.f && d.light_kind == B200RT_LIGHT_POINT) {
            for (int c = 0; c < 3; ++c) d.kd[c] = mats[1 + c];   // one Lambertian Kd (sigma 0): the lean kernel path
        } else {
            d.materials = mats.data(); d.n_materials = (uint32_t)(mats.size() / 20); d.prim_material = primMat.data();
            d.material_stride = 20;
            for (size_t k = 0; k < texRecs.size(); ++k) texRecs[k].texels = texData[k].data();
            if (!texRecs.empty()) { d.textures = texRecs.data(); d.n_textures = (uint32_t)texRecs.size(); }
        }
    }
    d.spp = (uint32_t)sampler->samplesPerPixel;
    d.max_depth = (uint32_t)maxDepth;
    d.tile_size = 16;   // integrator.cpp:237
    d.rank = (uint32_t)rank; d.world = (uint32_t)world; d.seed = (uint32_t)seed;

    const size_t nPix = (size_t)cb.Area();
    std::vector<float> rgbw(4 * nPix, 0.f);
    std::vector<uint32_t> heat;
    if (generalStats) heat.assign((size_t)B200RT_HEAT_PLANES * nPix, 0u);
    if (generalStats ? b200rt_render_heatmaps(accel->deviceScene(), &d, rgbw.data(), heat.data(), &lastStats, nullptr)
                     : b200rt_render(accel->deviceScene(), &d, rgbw.data(), &lastStats, nullptr)) {
        Error("b200wavefront: %s", b200rt_last_error());
        return;
    }
    if (world > 1) {
        // the frame is the sum of the ranks' films (each holds its own tiles + boundary-sample halos); rank 0 writes it
        if (b200rt_film_reduce(rendezvous.c_str(), (uint32_t)rank, (uint32_t)world, rgbw.data(), rgbw.size(),
                               generalStats ? heat.data() : nullptr, heat.size(), 0)) {
            Error("b200wavefront: %s", b200rt_last_error());
            return;
        }
        if (rank != 0) return;
    }
    // Film: pixel = sum / weight (film.cpp:266-305); SetImage stores it with filterWeightSum = 1
    std::unique_ptr<Spectrum[]> img(new Spectrum[nPix]);
    for (size_t i = 0; i < nPix; ++i) {
        const float w = rgbw[4 * i + 3];
        Float rgb[3] = {0, 0, 0};
        if (w > 0) for (int c = 0; c < 3; ++c) rgb[c] = rgbw[4 * i + c] / w;
        img[i] = Spectrum::FromRGB(rgb);
    }
    film->SetImage(img.get());
    film->WriteImage();
    if (generalStats) {
        // pixel.stats += ray.stats over the pixel's samples (integrator.cpp:327-328); the node counter is the one the
        // uploaded tree's family increments (bvh.cpp / kdtreeaccel.cpp / rbsp.cpp, BSP.cpp); rays = samples taken
        const int k = accel->treeKind();
        for (size_t i = 0; i < nPix; ++i) {
            GeneralStats &g = film->pixels[i].stats;
            g = GeneralStats();
            g.rays = (uint64_t)rgbw[4 * i + 3];
            g.primitiveIntersections = heat[B200RT_HEAT_PRIM * nPix + i];
            g.primitiveIntersectionsP = heat[B200RT_HEAT_PRIM_P * nPix + i];
            uint64_t &n = k == B200RT_BVH ? g.bvhTreeNodeTraversals : k == B200RT_KDTREE ? g.kdTreeNodeTraversals : g.bspTreeNodeTraversals;
            uint64_t &np = k == B200RT_BVH ? g.bvhTreeNodeTraversalsP : k == B200RT_KDTREE ? g.kdTreeNodeTraversalsP : g.bspTreeNodeTraversalsP;
            n = heat[B200RT_HEAT_NODE * nPix + i];
            np = heat[B200RT_HEAT_NODE_P * nPix + i];
            g.leafNodeTraversals = heat[B200RT_HEAT_LEAF * nPix + i];
            g.leafNodeTraversalsP = heat[B200RT_HEAT_LEAF_P * nPix + i];
        }
        film->WriteGeneralStats();   // film.cpp:170-188: <image name>-primitiveIntersections.txt, ...
    }
}

B200WavefrontIntegrator *CreateB200WavefrontIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler,
                                                       std::shared_ptr<const Camera> camera) {
    const int maxDepth = params.FindOneInt("maxdepth", 5);
    const int rank = params.FindOneInt("rank", 0);
    // "integer rank" / "integer world" are explicit: a stray WORLD_SIZE in the environment must not make a render partial
    const int world = params.FindOneInt("world", 1);
    const int seed = params.FindOneInt("seed", 0);
    const bool generalStats = params.FindOneBool("generalstats", false);
    const std::string rendezvous = params.FindOneString("rendezvous", "");
    if (world > 1 && rendezvous.empty()) {
        Error("b200wavefront: \"integer world\" > 1 needs \"string rendezvous\" (a directory every rank can write)");
        return nullptr;
    }
    if (rank < 0 || rank >= std::max(world, 1)) { Error("b200wavefront: rank %d outside world %d", rank, world); return nullptr; }
    // Integrator "path" "string lightsamplestrategy" (path.cpp:218-219, default "spatial"): it only matters with several lights
    const std::string strategy = params.FindOneString("lightsamplestrategy", "spatial");
    return new B200WavefrontIntegrator(maxDepth, std::move(camera), std::move(sampler), rank, world, seed, generalStats, rendezvous, strategy);
}

}  // namespace pbrt
