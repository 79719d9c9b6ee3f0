// b200rt_film_reduce: the ONE exchange step of a multi-process render.  Every rank's film holds its own tiles
// (plus the boundary-sample contributions to the one-pixel halo around them) and zeros elsewhere, so the frame is the SUM of
// the ranks' films (what Film::MergeFilmTile does tile by tile under the film mutex, src/core/film.cpp:118-135, for the
// reference's threads).  One process per GPU on one node: rank 0 maps every peer's device buffer through CUDA IPC
// (peer-to-peer over NVLink / NVSwitch when the ranks sit on different GPUs) and adds the films in RANK ORDER, so the sum
// is deterministic whatever the arrival order.  The handles travel through small files in a directory every rank can reach
// (the C++ Integrator "b200wavefront" has no launcher and no communicator: `"string rendezvous"` names that directory).
#include <cuda_runtime.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>

#include "../../include/b200rt_wavefront.h"
#include "scene_internal.h"

namespace {

struct PeerRecord {            // what rank r > 0 publishes
    uint32_t magic, rank, device, has_heat;
    uint64_t n_floats, n_heat;
    cudaIpcMemHandle_t film, heat;
};
constexpr uint32_t kMagic = 0xb200f11au;

__global__ void k_film_add(float4 *__restrict__ dst, const float4 *__restrict__ src, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 a = dst[i];
        const float4 b = __ldg(src + i);
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        dst[i] = a;
    }
}
__global__ void k_heat_add(uint32_t *__restrict__ dst, const uint32_t *__restrict__ src, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] += __ldg(src + i);
}

bool is_device_ptr(const void *p) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

bool write_file_atomic(const std::string &path, const void *data, size_t n) {
    const std::string tmp = path + ".tmp";
    FILE *f = std::fopen(tmp.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(data, 1, n, f) == n;
    std::fclose(f);
    return ok && std::rename(tmp.c_str(), path.c_str()) == 0;
}

bool wait_for_file(const std::string &path, void *data, size_t n, double timeout_s) {
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        if (FILE *f = std::fopen(path.c_str(), "rb")) {
            const size_t got = data ? std::fread(data, 1, n, f) : 0;
            std::fclose(f);
            if (!data || got == n) return true;
        }
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) return false;
        std::this_thread::sleep_for(std::chrono::milliseconds(2));
    }
}

// device copy of a host-or-device buffer (nullptr when p is already on the device and no copy is needed)
template <typename T>
int stage(T *p, size_t n, T **dev, bool *owned) {
    *owned = false;
    *dev = p;
    if (!p || is_device_ptr(p)) return 0;
    if (cudaMalloc(dev, n * sizeof(T)) != cudaSuccess) return b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: cudaMalloc of %zu bytes failed", n * sizeof(T));
    *owned = true;
    if (cudaMemcpy(*dev, p, n * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) return b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: upload failed");
    return 0;
}

}  // namespace

extern "C" int b200rt_film_reduce(const char *rendezvous, uint32_t rank, uint32_t world, float *film, uint64_t n_floats,
                                  uint32_t *heat, uint64_t n_heat, double timeout_s) {
    if (world <= 1) return 0;
    if (!rendezvous || !*rendezvous) return b200rt::set_error(B200RT_EINVAL, "b200rt_film_reduce: world > 1 needs a rendezvous directory");
    if (rank >= world || !film || n_floats == 0 || n_floats % 4) return b200rt::set_error(B200RT_EINVAL, "b200rt_film_reduce: bad rank / film (n_floats must be 4 per pixel)");
    if (timeout_s <= 0) timeout_s = 120;
    struct stat sb;
    if (stat(rendezvous, &sb) != 0 || !S_ISDIR(sb.st_mode)) return b200rt::set_error(B200RT_EINVAL, "b200rt_film_reduce: %s is not a directory", rendezvous);
    const std::string dir(rendezvous);
    auto peer_path = [&](uint32_t r) { return dir + "/film." + std::to_string(r); };
    auto done_path = [&](uint32_t r) { return dir + "/done." + std::to_string(r); };

    float *dFilm = nullptr;
    uint32_t *dHeat = nullptr;
    bool ownFilm = false, ownHeat = false;
    int rc = stage(film, (size_t)n_floats, &dFilm, &ownFilm);
    if (!rc) rc = stage(heat, (size_t)n_heat, &dHeat, &ownHeat);
    auto cleanup = [&]() { if (ownFilm) cudaFree(dFilm); if (ownHeat) cudaFree(dHeat); };
    if (rc) { cleanup(); return rc; }

    if (rank > 0) {
        // IPC handles need the allocation's base: a caller's device pointer may sit inside a larger allocation, so publish a
        // private copy in that case too
        float *pubFilm = dFilm;
        uint32_t *pubHeat = dHeat;
        bool pf = false, ph = false;
        if (!ownFilm) { if (cudaMalloc(&pubFilm, n_floats * 4) != cudaSuccess) { cleanup(); return b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: cudaMalloc failed"); } pf = true; cudaMemcpy(pubFilm, dFilm, n_floats * 4, cudaMemcpyDeviceToDevice); }
        if (heat && !ownHeat) { if (cudaMalloc(&pubHeat, n_heat * 4) != cudaSuccess) { cleanup(); return b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: cudaMalloc failed"); } ph = true; cudaMemcpy(pubHeat, dHeat, n_heat * 4, cudaMemcpyDeviceToDevice); }
        cudaDeviceSynchronize();
        PeerRecord rec{};
        rec.magic = kMagic; rec.rank = rank; rec.has_heat = heat ? 1u : 0u; rec.n_floats = n_floats; rec.n_heat = heat ? n_heat : 0;
        int dev = 0;
        cudaGetDevice(&dev);
        rec.device = (uint32_t)dev;
        cudaError_t e = cudaIpcGetMemHandle(&rec.film, pubFilm);
        if (e == cudaSuccess && heat) e = cudaIpcGetMemHandle(&rec.heat, pubHeat);
        if (e != cudaSuccess) rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
        std::remove(done_path(rank).c_str());
        if (!rc && !write_file_atomic(peer_path(rank), &rec, sizeof rec)) rc = b200rt::set_error(B200RT_EINVAL, "b200rt_film_reduce: cannot write %s", peer_path(rank).c_str());
        // the buffer must stay mapped until rank 0 has read it
        if (!rc && !wait_for_file(done_path(rank), nullptr, 0, timeout_s)) rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: rank 0 did not collect rank %u's film within %.0f s", rank, timeout_s);
        std::remove(done_path(rank).c_str());
        if (pf) cudaFree(pubFilm);
        if (ph) cudaFree(pubHeat);
        cleanup();
        return rc;
    }

    const int blocks = 148 * 4;
    for (uint32_t r = 1; r < world && !rc; ++r) {
        PeerRecord rec{};
        if (!wait_for_file(peer_path(r), &rec, sizeof rec, timeout_s)) { rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: rank %u's film did not arrive within %.0f s", r, timeout_s); break; }
        std::remove(peer_path(r).c_str());
        if (rec.magic != kMagic || rec.rank != r || rec.n_floats != n_floats || (heat && rec.n_heat != n_heat)) {
            rc = b200rt::set_error(B200RT_EINVAL, "b200rt_film_reduce: rank %u published a film of another shape (stale rendezvous directory?)", r);
            break;
        }
        void *pf = nullptr, *ph = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&pf, rec.film, cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess && heat && rec.has_heat) e = cudaIpcOpenMemHandle(&ph, rec.heat, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: cudaIpcOpenMemHandle(rank %u): %s", r, cudaGetErrorString(e)); }
        if (!rc) {
            k_film_add<<<blocks, 256>>>((float4 *)dFilm, (const float4 *)pf, (size_t)n_floats / 4);
            if (ph) k_heat_add<<<blocks, 256>>>(dHeat, (const uint32_t *)ph, (size_t)n_heat);
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: add kernel: %s", cudaGetErrorString(e));
        }
        if (pf) cudaIpcCloseMemHandle(pf);
        if (ph) cudaIpcCloseMemHandle(ph);
        write_file_atomic(done_path(r), "ok", 2);   // releases rank r even on error
    }
    if (!rc && ownFilm && cudaMemcpy(film, dFilm, n_floats * 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: download failed");
    if (!rc && ownHeat && cudaMemcpy(heat, dHeat, n_heat * 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = b200rt::set_error(B200RT_ECUDA, "b200rt_film_reduce: download failed");
    cleanup();
    return rc;
}
