// b200rt_mipmap_build: the MIPMap pyramid of an `imagemap` texture, built on the host as the reference builds it
// (ImageTexture::GetTexture, textures/imagemap.cpp:52-97: y flip, convertIn = scale * InverseGammaCorrect; MIPMap::MIPMap,
// core/mipmap.h:113-209: Lanczos resampling to the next power of two with ImageWrap::Repeat, clamp to >= 0, then box-filtered
// levels).  Same float operations in the same order (compiled with -ffp-contract=off; powf / sinf are this host's libm, as they
// are for the reference), so the texels equal the reference's bit for bit (tests/test_texture.py).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/b200rt_wavefront.h"
#include "wf_setup_host.h"

namespace {

float inverse_gamma_correct(float value) {   // pbrt.h:298-301
    if (value <= 0.04045f) return value * 1.f / 12.92f;
    return std::pow((value + 0.055f) * 1.f / 1.055f, 2.4f);
}
float lanczos(float x, float tau = 2.f) {   // texture.cpp:254-262
    x = std::abs(x);
    if (x < 1e-5f) return 1;
    if (x > 1.f) return 0;
    x *= 3.14159265358979323846f;
    const float s = std::sin(x * tau) / (x * tau);
    const float l = std::sin(x) / x;
    return s * l;
}
struct ResampleWeight { int firstTexel; float weight[4]; };
std::vector<ResampleWeight> resample_weights(int oldRes, int newRes) {   // mipmap.h:76-96
    std::vector<ResampleWeight> wt(newRes);
    const float filterwidth = 2.f;
    for (int i = 0; i < newRes; ++i) {
        const float center = (i + .5f) * oldRes / newRes;
        wt[i].firstTexel = (int)std::floor((center - filterwidth) + 0.5f);
        for (int j = 0; j < 4; ++j) {
            const float pos = wt[i].firstTexel + j + .5f;
            wt[i].weight[j] = lanczos((pos - center) / filterwidth);
        }
        const float invSumWts = 1 / (wt[i].weight[0] + wt[i].weight[1] + wt[i].weight[2] + wt[i].weight[3]);
        for (int j = 0; j < 4; ++j) wt[i].weight[j] *= invSumWts;
    }
    return wt;
}
int mod(int a, int b) { const int r = a - (a / b) * b; return r < 0 ? r + b : r; }
uint32_t round_up_pow2(uint32_t v) { uint32_t p = 1; while (p < v) p <<= 1; return p; }

}  // namespace

extern "C" int b200rt_mipmap_dims(uint32_t w, uint32_t h, uint32_t *w0, uint32_t *h0, uint32_t *n_levels, uint64_t *n_texels) {
    if (!w || !h || w > 16384 || h > 16384) return b200rt::set_error(B200RT_EINVAL, "b200rt_mipmap_dims: image of %u x %u", w, h);
    const uint32_t W = round_up_pow2(w), H = round_up_pow2(h);
    uint32_t n = 1;
    while ((1u << n) <= (W > H ? W : H)) ++n;   // 1 + Log2Int(max)
    uint64_t total = 0;
    for (uint32_t l = 0; l < n; ++l) total += (uint64_t)((W >> l) ? (W >> l) : 1u) * ((H >> l) ? (H >> l) : 1u);
    if (w0) *w0 = W;
    if (h0) *h0 = H;
    if (n_levels) *n_levels = n;
    if (n_texels) *n_texels = total;
    return B200RT_OK;
}

// rgb: w x h x 3 floats as ReadImage returns them (first row = top of the image); out: n_texels x 3 floats, level after level.
extern "C" int b200rt_mipmap_build(const float *rgb, uint32_t w, uint32_t h, int gamma, float scale, float *out) {
    uint32_t W, H, nL;
    uint64_t total;
    if (!rgb || !out) return b200rt::set_error(B200RT_EINVAL, "b200rt_mipmap_build: null argument");
    if (int rc = b200rt_mipmap_dims(w, h, &W, &H, &nL, &total)) return rc;
    const int rw = (int)w, rh = (int)h;
    // flip in y (texture space has (0,0) at the lower left) and convertIn
    std::vector<float> img((size_t)rw * rh * 3);
    for (int y = 0; y < rh; ++y)
        for (int x = 0; x < rw; ++x)
            for (int c = 0; c < 3; ++c) {
                const float v = rgb[((size_t)(rh - 1 - y) * rw + x) * 3 + c];
                img[((size_t)y * rw + x) * 3 + c] = scale * (gamma ? inverse_gamma_correct(v) : v);
            }
    float *l0 = out;
    if (W != w || H != h) {
        const std::vector<ResampleWeight> sW = resample_weights(rw, (int)W);
        std::vector<float> res((size_t)W * H * 3, 0.f);
        for (int t = 0; t < rh; ++t)
            for (int s = 0; s < (int)W; ++s)
                for (int c = 0; c < 3; ++c) {
                    float acc = 0.f;
                    for (int j = 0; j < 4; ++j) {
                        const int origS = mod(sW[s].firstTexel + j, rw);
                        acc += sW[s].weight[j] * img[((size_t)t * rw + origS) * 3 + c];
                    }
                    res[((size_t)t * W + s) * 3 + c] = acc;
                }
        const std::vector<ResampleWeight> tW = resample_weights(rh, (int)H);
        std::vector<float> work((size_t)H);
        for (int s = 0; s < (int)W; ++s)
            for (int c = 0; c < 3; ++c) {
                for (int t = 0; t < (int)H; ++t) {
                    float acc = 0.f;
                    for (int j = 0; j < 4; ++j) {
                        const int off = mod(tW[t].firstTexel + j, rh);
                        acc += tW[t].weight[j] * res[((size_t)off * W + s) * 3 + c];
                    }
                    work[t] = acc;
                }
                for (int t = 0; t < (int)H; ++t) l0[((size_t)t * W + s) * 3 + c] = work[t] < 0.f ? 0.f : work[t];   // clamp(0, Infinity)
            }
    } else {
        std::memcpy(l0, img.data(), img.size() * sizeof(float));
    }
    // coarser levels: .25f * (four texels of the finer level), Texel() with Repeat wrap (mipmap.h:184-196)
    const float *prev = l0;
    uint32_t pw = W, ph = H;
    float *dst = out + (size_t)W * H * 3;
    for (uint32_t l = 1; l < nL; ++l) {
        const uint32_t sw = pw / 2 ? pw / 2 : 1u, sh = ph / 2 ? ph / 2 : 1u;
        for (uint32_t t = 0; t < sh; ++t)
            for (uint32_t s = 0; s < sw; ++s)
                for (int c = 0; c < 3; ++c) {
                    auto tex = [&](int ss, int tt) { return prev[((size_t)mod(tt, (int)ph) * pw + mod(ss, (int)pw)) * 3 + c]; };
                    dst[((size_t)t * sw + s) * 3 + c] = .25f * (tex(2 * s, 2 * t) + tex(2 * s + 1, 2 * t) + tex(2 * s, 2 * t + 1) + tex(2 * s + 1, 2 * t + 1));
                }
        prev = dst;
        dst += (size_t)sw * sh * 3;
        pw = sw; ph = sh;
    }
    return B200RT_OK;
}

extern "C" void b200rt_ewa_weights(float *out) {
    for (int i = 0; i < 128; ++i) {
        const float alpha = 2;
        const float r2 = float(i) / float(128 - 1);
        out[i] = std::exp(-alpha * r2) - std::exp(-alpha);
    }
}
