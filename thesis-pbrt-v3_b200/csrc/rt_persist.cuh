// Persistent-warp traversal kernels (sm_100a): the production path of b200rt_intersect_closest/any.
//
// Why (ncu, profiles/r01_ncu_full_v1_summary.txt): the one-ray-per-thread baseline kernel is
// ISSUE-bound (issue slots 75-80 % busy, 2-3 eligible warps per issue, DRAM < 2 %) while only 8-10 of
// 32 lanes execute each instruction on path-traced ray queues.  So the levers are (1) lanes per
// instruction and (2) instructions per traversal step, not bytes:
//   * persistent warps: a warp keeps running and REFILLS idle lanes with new rays as soon as fewer
//     than `thresh` lanes are still traversing (ballot + popc + one chunked atomic per warp), instead
//     of waiting for the slowest ray of a fixed 32-ray packet;
//   * while-while scheduling: all lanes first walk interior nodes until each holds a leaf (or is
//     done), then all lanes run the (much longer) triangle code together, instead of paying
//     interior + leaf code on every iteration of one mixed loop;
//   * a branch-free slab test (2 x FMNMX3 per box) that is bit-identical to the reference's
//     sequential test for every ray whose 1/d.x is finite; the rare other rays take the reference's
//     literal sequence (slab_geom).
//
// Results are bit-identical to the baseline kernel and to the reference: same node order (near child
// first by dirIsNeg[axis], far child re-tested against the CURRENT ray.tMax when popped, bvh.cpp:367),
// leaf primitives in array order, strict comparisons.
#pragma once
#include "rt_kernels.cuh"

namespace b200rt {

// The persistent kernels' per-round vote: the warp runs an interior round when nInterior * wI >= nLeaf * wL, else a primitive
// round (one triangle per lane).  wI = wL = 1 is the plain majority; wI > 1 keeps the cheap interior rounds going until most
// lanes wait at a leaf, so that the expensive primitive rounds run fuller.  Schedules only: every lane still sees its own
// nodes and primitives in the reference's order.
inline uint32_t pack_steps(uint32_t steps, uint32_t wI, uint32_t wL) { return (steps & 0xffu) | ((wI & 0xffu) << 8) | ((wL & 0xffu) << 16); }


#define B200RT_DONE 0x7FFFFFFFu  // "no node": an interior ref that no tree can contain
#define B200RT_PBLOCK 128

// Bounds3::IntersectP(ray, invDir, dirIsNeg) (geometry.h:1754-1780) for rays with finite 1/d.x and
// finite origin.  Equivalence with the sequential form: comparisons with NaN are false there, so a
// NaN entry value acts as -inf and a NaN exit value as +inf — exactly what fmaxf/fminf do by
// ignoring a NaN operand; NaNs come only from 0*inf on an axis with d == 0, and on the x axis they
// are sticky in the reference (the ray misses), which is why such rays do not come here.  The
// pairwise early-outs test every (entry_a, exit_b) pair with a != b; same-axis pairs can only fail
// when that exit is < 0, which the final `tMax > 0` rejects as well.  x*k is monotone, so
// min(a*k, b*k, c*k) == min(a, b, c)*k bit for bit.
__device__ __forceinline__ bool slab_fast(const RayCtx &r, float lox, float loy, float loz, float hix, float hiy,
                                          float hiz, float &tEntry) {
    const float t0x = (lox - r.ox) * r.ix, t0y = (loy - r.oy) * r.iy, t0z = (loz - r.oz) * r.iz;
    const float t1x = (hix - r.ox) * r.ix, t1y = (hiy - r.oy) * r.iy, t1z = (hiz - r.oz) * r.iz;
    const float en = fmaxf(fmaxf(t0x, t0y), t0z);
    const float ex = fminf(fminf(t1x, t1y), t1z) * kSlabK;
    tEntry = en;
    return en <= ex && ex > 0;
}

// 256-bit read-only load (sm_100: LDG.E.256): one request instead of two LDG.128 — halves the L1
// data-pipe wavefronts of the scattered node fetch, which is what bounds the kernel once the lanes are
// kept busy (ncu l1tex__data_pipe_lsu_wavefronts 86 % of peak, profiles/r01_ncu_full_v3_vote_summary.txt).
__device__ __forceinline__ void ldg256(const void *p, float4 &lo, float4 &hi) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
                 : "l"(p));
}

struct PersistCounters {  // one slot per in-flight launch, zeroed by the last warp to leave
    unsigned int next;    // next unclaimed ray
    unsigned int exited;  // warps that have left the kernel
};

// Warp-level work queue: a warp claims `chunk` consecutive rays with one global atomic and hands
// them to its idle lanes by ballot rank (work stealing at warp granularity: fast warps simply claim
// more chunks).  Returns this lane's new ray index, or 0xFFFFFFFF if it gets none this round.
struct WarpQueue {
    uint32_t next = 0, end = 0;
    bool drained = false;  // the global queue has no more rays
};
__device__ __forceinline__ uint32_t warp_claim(WarpQueue &q, PersistCounters *pc, uint32_t n, uint32_t chunk,
                                               unsigned idleMask, bool laneIdle, int lane) {
    const unsigned FULL = 0xFFFFFFFFu;
    const uint32_t want = __popc(idleMask);
    if (q.end == q.next) {  // the remainder of a chunk (< 32 rays) is handed out before a new claim
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&pc->next, chunk);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n) { q.drained = true; q.next = q.end = 0; }
        else { q.next = base; q.end = base + chunk < n ? base + chunk : n; }
    }
    const uint32_t avail = q.end - q.next;
    const uint32_t rank = __popc(idleMask & ((1u << lane) - 1u));
    const uint32_t idx = (laneIdle && rank < avail) ? q.next + rank : 0xFFFFFFFFu;
    q.next += want < avail ? want : avail;
    return idx;
}
// the last warp out re-arms the counter slot for the next launch that uses it
__device__ __forceinline__ void warp_exit(PersistCounters *pc, int lane) {
    if (lane == 0) {
        const unsigned total = gridDim.x * (blockDim.x >> 5);
        __threadfence();
        if (atomicAdd(&pc->exited, 1u) == total - 1) { pc->next = 0; pc->exited = 0; __threadfence(); }
    }
}

template <bool ANY, int MINB, bool TOP = false>
__global__ void __launch_bounds__(B200RT_PBLOCK, MINB)
bvh_persistent_kernel(const __grid_constant__ BvhView S, const float4 *__restrict__ rays, uint64_t n64,
                      const uint32_t *__restrict__ n_dev, b200rt_hit *__restrict__ hits,
                      uint8_t *__restrict__ occluded, PersistCounters *pc, uint32_t thresh,
                      uint32_t chunk /* rays a warp claims per global atomic */,
                      uint32_t stepsW /* bits 0-7: interior steps a lane may take per vote; 8-15 / 16-23: the vote's weights (pack_steps) */,
                      uint32_t sdepth /* stack entries [0, sdepth) live in shared memory, the rest in local */,
                      uint32_t topN /* the first topN node records (= the top levels, BvhView::topCount) are staged in shared memory */) {
    // Shared-memory bottom of the traversal stack: entry d of thread t sits at sst[d * blockDim + t], i.e. in bank
    // t % 32 whatever d is, so a push/pop by 32 lanes at 32 DIFFERENT depths is still one conflict-free access
    // (the per-thread local array would touch up to 32 different 128-byte lines for the same push).
    extern __shared__ uint2 sst[];
    uint32_t n = (uint32_t)n64;
    if (n_dev) { const uint32_t m = *n_dev; n = m < n ? m : n; }
    // the host sized `chunk` for the queue's CAPACITY; a queue that cannot give every warp one chunk is handed out in 32-ray
    // chunks, so that a thin deep-bounce queue spreads over the machine instead of running 128 rays a warp, four after another
    if (chunk > 32u && n < gridDim.x * (blockDim.x >> 5) * chunk) chunk = 32u;
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    // Top of the tree in shared memory (north-star item (3)): one cooperative copy per block, behind the stack area.
    // A separate instantiation (TOP), off by default: the staged records were L1-resident anyway, so the frame time does not
    // move (profiles/r02_ab_bvh_toplevels_v2.log), while the extra branch in the node fetch costs the default path 5 % when it
    // is compiled in (profiles/r02_ab_bench_old_vs_new.log).
    const float4 *stop = reinterpret_cast<const float4 *>(sst + (size_t)sdepth * B200RT_PBLOCK);
    if (TOP && topN) {
        float4 *w = reinterpret_cast<float4 *>(sst + (size_t)sdepth * B200RT_PBLOCK);
        for (uint32_t i = threadIdx.x; i < topN * 4u; i += blockDim.x) w[i] = __ldg(reinterpret_cast<const float4 *>(S.nodes) + i);
        __syncthreads();
    }

    // per-lane traversal state
    RayCtx r;
    HitRec h;
    uint32_t cur = B200RT_DONE, rayIdx = 0;
    bool nx = false, ny = false, nz = false, slow = false, hit = false;
    bool live = false;  // this lane holds a ray whose result has not been written yet
    uint32_t negMask = 0;
    uint2 stack[64];  // (node ref, entry distance of its box): one 8-byte local store / load per push / pop
    int sp = 0;
    WarpQueue wq;

    for (;;) {
        // ---- refill idle lanes (warp-uniform control flow) -------------------------------------
        const unsigned idle = __ballot_sync(FULL, cur == B200RT_DONE);
        if (idle && !wq.drained) {
            const uint32_t got = warp_claim(wq, pc, n, chunk, idle, cur == B200RT_DONE, lane);
            if (got != 0xFFFFFFFFu) {
                rayIdx = got;
                ray_setup(r, __ldg(rays + 2 * (size_t)rayIdx), __ldg(rays + 2 * (size_t)rayIdx + 1));
                nx = r.ix < 0; ny = r.iy < 0; nz = r.iz < 0;  // dirIsNeg (bvh.cpp:359)
                negMask = (nx ? 1u : 0u) | (ny ? 2u : 0u) | (nz ? 4u : 0u);
                slow = !(fabsf(r.ix) <= 3.402823466e+38f) || !(fabsf(r.iy) <= 3.402823466e+38f) ||
                       !(fabsf(r.iz) <= 3.402823466e+38f) || !(fabsf(r.ox) <= 3.402823466e+38f) ||
                       !(fabsf(r.oy) <= 3.402823466e+38f) || !(fabsf(r.oz) <= 3.402823466e+38f);
                h.prim = B200RT_MISS; h.t = r.tMax; h.b0 = h.b1 = h.b2 = 0.f;
                hit = false;
                sp = 0;
                // root box (node 0 of the reference): literal sequence, once per ray
                float te;
                const bool ok = slab_geom(r, nx ? S.root[3] : S.root[0], ny ? S.root[4] : S.root[1],
                                          nz ? S.root[5] : S.root[2], nx ? S.root[0] : S.root[3],
                                          ny ? S.root[1] : S.root[4], nz ? S.root[2] : S.root[5], te);
                if (ok && te < r.tMax) { cur = S.rootRef; live = true; }
                else {  // misses everything
                    if (ANY) occluded[rayIdx] = 0;
                    else {
                        b200rt_hit o; o.prim = B200RT_MISS; o.t = h.t; o.b0 = o.b1 = o.b2 = 0.f;
                        hits[rayIdx] = o;
                    }
                }
            }
        }
        if (__ballot_sync(FULL, cur != B200RT_DONE) == 0) {
            if (wq.drained) break;
            continue;
        }
        // ---- vote-scheduled traversal until too few lanes are busy -------------------------------
        // Each round the warp runs ONE kind of step, the one more lanes are waiting for: an interior
        // step (fetch a 64-byte node, two slab tests, order, push/pop) or a primitive step (one
        // triangle of the lane's current leaf).  Lanes in the other state wait; they become the
        // majority within a few rounds.  Every lane still sees its own nodes and primitives in the
        // reference's order, so results do not depend on the schedule.
        for (;;) {
            const unsigned mI = __ballot_sync(FULL, !(cur & B200RT_LEAF_BIT) && cur != B200RT_DONE);
            const unsigned mL = __ballot_sync(FULL, (cur & B200RT_LEAF_BIT) != 0);
            const uint32_t nI = __popc(mI), nL = __popc(mL);
            if (nI + nL == 0 || (!wq.drained && nI + nL < thresh)) break;
            if (nI * ((stepsW >> 8) & 0xffu) >= nL * (stepsW >> 16)) {
#pragma unroll 1
                for (uint32_t k = 0; k < (stepsW & 0xffu) && !(cur & B200RT_LEAF_BIT) && cur != B200RT_DONE; ++k) {
                    const BvhNode2 *nd = S.nodes + cur;
                    float4 a, b, c, mf;
                    if (TOP && cur < topN) { const float4 *q = stop + 4 * (size_t)cur; a = q[0]; b = q[1]; c = q[2]; mf = q[3]; }
                    else { ldg256(&nd->a, a, b); ldg256(&nd->c, c, mf); }
                    const uint4 m = make_uint4(__float_as_uint(mf.x), __float_as_uint(mf.y), __float_as_uint(mf.z), 0u);
                    float t0, t1;
                    bool g0, g1;
                    const float l0x = nx ? a.w : a.x, l0y = ny ? b.x : a.y, l0z = nz ? b.y : a.z;
                    const float h0x = nx ? a.x : a.w, h0y = ny ? a.y : b.x, h0z = nz ? a.z : b.y;
                    const float l1x = nx ? c.y : b.z, l1y = ny ? c.z : b.w, l1z = nz ? c.w : c.x;
                    const float h1x = nx ? b.z : c.y, h1y = ny ? b.w : c.z, h1z = nz ? c.x : c.w;
                    if (!slow) {
                        g0 = slab_fast(r, l0x, l0y, l0z, h0x, h0y, h0z, t0);
                        g1 = slab_fast(r, l1x, l1y, l1z, h1x, h1y, h1z, t1);
                    } else {
                        g0 = slab_geom(r, l0x, l0y, l0z, h0x, h0y, h0z, t0);
                        g1 = slab_geom(r, l1x, l1y, l1z, h1x, h1y, h1z, t1);
                    }
                    const bool p0 = g0 && t0 < r.tMax, p1 = g1 && t1 < r.tMax;
                    // near = second child when dirIsNeg[axis] (bvh.cpp:382-388)
                    // (nearer-box-first ordering was measured: +0.4 % and a handful of near-tie differences per 4 M rays —
                    // profiles/r01_ab_bvh_child_order.log — so the reference's order stays)
                    const bool negAxis = (negMask >> m.z) & 1u;
                    const uint32_t nearRef = negAxis ? m.y : m.x, farRef = negAxis ? m.x : m.y;
                    const bool hitNear = negAxis ? p1 : p0, hitFar = negAxis ? p0 : p1;
                    if (hitNear) {
                        if (hitFar) {
                            const uint2 e = make_uint2(farRef, __float_as_uint(negAxis ? t0 : t1));
                            if ((uint32_t)sp < sdepth) sst[sp * B200RT_PBLOCK + threadIdx.x] = e;
                            else stack[sp - sdepth] = e;
                            ++sp;
                        }
                        cur = nearRef;
                    } else if (hitFar) cur = farRef;
                    else {
                        // pop: the reference re-tests the far box against the CURRENT ray.tMax (bvh.cpp:367)
                        cur = B200RT_DONE;
                        while (sp > 0) {
                            --sp;
                            const uint2 e = (uint32_t)sp < sdepth ? sst[sp * B200RT_PBLOCK + threadIdx.x] : stack[sp - sdepth];
                            if (ANY || __uint_as_float(e.y) < r.tMax) { cur = e.x; break; }
                        }
                    }
                }
            } else {
                if (cur & B200RT_LEAF_BIT) {
                    // next primitive of the leaf, in array order (bvh.cpp:372-375 / 412-418)
                    const float4 *rec = S.tris + 3 * (size_t)(cur & ~B200RT_LEAF_BIT);
                    const float4 v0 = __ldg(rec), v1 = __ldg(rec + 1), v2 = __ldg(rec + 2);
                    bool stop = false;
                    if (prim_closest<ANY>(r, v0, v1, v2, h, S.spheres, rays + 2 * (size_t)rayIdx + 1)) {
                        hit = true;
                        if (ANY) stop = true;
                    }
                    if (stop) cur = B200RT_DONE;
                    else if (__float_as_uint(v1.w) & B200RT_TRI_LAST) {
                        cur = B200RT_DONE;
                        while (sp > 0) {
                            --sp;
                            const uint2 e = (uint32_t)sp < sdepth ? sst[sp * B200RT_PBLOCK + threadIdx.x] : stack[sp - sdepth];
                            if (ANY || __uint_as_float(e.y) < r.tMax) { cur = e.x; break; }
                        }
                    } else cur += 1;
                }
            }
            if (live && cur == B200RT_DONE) {  // ray finished (in either kind of step): write its result
                live = false;
                if (ANY) occluded[rayIdx] = hit ? 1 : 0;
                else {
                    b200rt_hit o; o.prim = h.prim; o.t = h.t; o.b0 = h.b0; o.b1 = h.b1; o.b2 = h.b2;
                    hits[rayIdx] = o;
                }
            }
        }
    }
    warp_exit(pc, lane);
}

// ---- BVH, two levels per step (BvhNode4, rt_kernels.cuh) --------------------------------------------------------------
// Same persistent-warp / vote skeleton as bvh_persistent_kernel; an interior step fetches one 128-byte record (four
// 256-bit loads issued together), tests the four grandchild boxes and continues with the first one that passes in the
// reference's visiting order, pushing the others (up to three) in reverse order with their entry distances.
template <bool ANY, int MINB>
__global__ void __launch_bounds__(B200RT_PBLOCK, MINB)
bvh4_persistent_kernel(const __grid_constant__ BvhView S, const float4 *__restrict__ rays, uint64_t n64,
                       const uint32_t *__restrict__ n_dev, b200rt_hit *__restrict__ hits,
                       uint8_t *__restrict__ occluded, PersistCounters *pc, uint32_t thresh, uint32_t chunk,
                       uint32_t stepsW, uint32_t sdepth, uint32_t /*topN: unused*/) {
    extern __shared__ uint2 sst[];
    uint32_t n = (uint32_t)n64;
    if (n_dev) { const uint32_t m = *n_dev; n = m < n ? m : n; }
    // the host sized `chunk` for the queue's CAPACITY; a queue that cannot give every warp one chunk is handed out in 32-ray
    // chunks, so that a thin deep-bounce queue spreads over the machine instead of running 128 rays a warp, four after another
    if (chunk > 32u && n < gridDim.x * (blockDim.x >> 5) * chunk) chunk = 32u;
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    RayCtx r;
    HitRec h;
    uint32_t cur = B200RT_DONE, rayIdx = 0;
    bool nx = false, ny = false, nz = false, slow = false, hit = false;
    bool live = false;
    uint32_t negMask = 0;
    uint2 stack[96];   // up to three deferred slots per two reference levels (upload checks depth <= 64)
    int sp = 0;
    WarpQueue wq;
#define B4_PUSH(refv, tv)                                                                   \
    do {                                                                                    \
        const uint2 e_ = make_uint2((refv), __float_as_uint(tv));                           \
        if ((uint32_t)sp < sdepth) sst[sp * B200RT_PBLOCK + threadIdx.x] = e_;              \
        else stack[sp - sdepth] = e_;                                                       \
        ++sp;                                                                               \
    } while (0)
#define B4_POP()                                                                            \
    do {                                                                                    \
        cur = B200RT_DONE;                                                                  \
        while (sp > 0) {                                                                    \
            --sp;                                                                           \
            const uint2 e_ = (uint32_t)sp < sdepth ? sst[sp * B200RT_PBLOCK + threadIdx.x] : stack[sp - sdepth]; \
            if (ANY || __uint_as_float(e_.y) < r.tMax) { cur = e_.x; break; }               \
        }                                                                                   \
    } while (0)
    for (;;) {
        const unsigned idle = __ballot_sync(FULL, cur == B200RT_DONE);
        if (idle && !wq.drained) {
            const uint32_t got = warp_claim(wq, pc, n, chunk, idle, cur == B200RT_DONE, lane);
            if (got != 0xFFFFFFFFu) {
                rayIdx = got;
                ray_setup(r, __ldg(rays + 2 * (size_t)rayIdx), __ldg(rays + 2 * (size_t)rayIdx + 1));
                nx = r.ix < 0; ny = r.iy < 0; nz = r.iz < 0;
                negMask = (nx ? 1u : 0u) | (ny ? 2u : 0u) | (nz ? 4u : 0u);
                slow = !(fabsf(r.ix) <= 3.402823466e+38f) || !(fabsf(r.iy) <= 3.402823466e+38f) ||
                       !(fabsf(r.iz) <= 3.402823466e+38f) || !(fabsf(r.ox) <= 3.402823466e+38f) ||
                       !(fabsf(r.oy) <= 3.402823466e+38f) || !(fabsf(r.oz) <= 3.402823466e+38f);
                h.prim = B200RT_MISS; h.t = r.tMax; h.b0 = h.b1 = h.b2 = 0.f;
                hit = false;
                sp = 0;
                float te;
                const bool ok = slab_geom(r, nx ? S.root[3] : S.root[0], ny ? S.root[4] : S.root[1],
                                          nz ? S.root[5] : S.root[2], nx ? S.root[0] : S.root[3],
                                          ny ? S.root[1] : S.root[4], nz ? S.root[2] : S.root[5], te);
                if (ok && te < r.tMax) { cur = S.rootRef4; live = true; }
                else {
                    if (ANY) occluded[rayIdx] = 0;
                    else {
                        b200rt_hit o; o.prim = B200RT_MISS; o.t = h.t; o.b0 = o.b1 = o.b2 = 0.f;
                        hits[rayIdx] = o;
                    }
                }
            }
        }
        if (__ballot_sync(FULL, cur != B200RT_DONE) == 0) {
            if (wq.drained) break;
            continue;
        }
        for (;;) {
            const unsigned mI = __ballot_sync(FULL, !(cur & B200RT_LEAF_BIT) && cur != B200RT_DONE);
            const unsigned mL = __ballot_sync(FULL, (cur & B200RT_LEAF_BIT) != 0);
            const uint32_t nI = __popc(mI), nL = __popc(mL);
            if (nI + nL == 0 || (!wq.drained && nI + nL < thresh)) break;
            if (nI * ((stepsW >> 8) & 0xffu) >= nL * (stepsW >> 16)) {
#pragma unroll 1
                for (uint32_t k = 0; k < (stepsW & 0xffu) && !(cur & B200RT_LEAF_BIT) && cur != B200RT_DONE; ++k) {
                    const float4 *nd = S.nodes4[cur].q;
                    float4 a0, b0, c0, m0, a1, b1, c1, m1;
                    ldg256(nd, a0, b0); ldg256(nd + 2, c0, m0); ldg256(nd + 4, a1, b1); ldg256(nd + 6, c1, m1);
                    const uint32_t meta = __float_as_uint(m0.z);
                    float t[4];
                    bool g[4];
#define B4_PAIR(a, b, c, i0)                                                                                       \
                    {                                                                                              \
                        const float l0x = nx ? a.w : a.x, l0y = ny ? b.x : a.y, l0z = nz ? b.y : a.z;              \
                        const float h0x = nx ? a.x : a.w, h0y = ny ? a.y : b.x, h0z = nz ? a.z : b.y;              \
                        const float l1x = nx ? c.y : b.z, l1y = ny ? c.z : b.w, l1z = nz ? c.w : c.x;              \
                        const float h1x = nx ? b.z : c.y, h1y = ny ? b.w : c.z, h1z = nz ? c.x : c.w;              \
                        if (!slow) {                                                                               \
                            g[i0] = slab_fast(r, l0x, l0y, l0z, h0x, h0y, h0z, t[i0]);                             \
                            g[i0 + 1] = slab_fast(r, l1x, l1y, l1z, h1x, h1y, h1z, t[i0 + 1]);                     \
                        } else {                                                                                   \
                            g[i0] = slab_geom(r, l0x, l0y, l0z, h0x, h0y, h0z, t[i0]);                             \
                            g[i0 + 1] = slab_geom(r, l1x, l1y, l1z, h1x, h1y, h1z, t[i0 + 1]);                     \
                        }                                                                                          \
                    }
                    B4_PAIR(a0, b0, c0, 0)
                    B4_PAIR(a1, b1, c1, 2)
#undef B4_PAIR
                    const uint32_t rf[4] = {__float_as_uint(m0.x), __float_as_uint(m0.y), __float_as_uint(m1.x), __float_as_uint(m1.y)};
                    bool p[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) p[q] = g[q] && t[q] < r.tMax && rf[q] != B200RT_EMPTY_SLOT;
                    const bool negTop = (negMask >> (meta & 3u)) & 1u, neg0 = (negMask >> ((meta >> 2) & 3u)) & 1u,
                               neg1 = (negMask >> ((meta >> 4) & 3u)) & 1u;
                    // visiting order: near pair first (by the node's axis), inside a pair the near slot first (by that child's axis)
                    const int i0 = negTop ? (neg1 ? 3 : 2) : (neg0 ? 1 : 0), i1 = negTop ? (neg1 ? 2 : 3) : (neg0 ? 0 : 1);
                    const int i2 = negTop ? (neg0 ? 1 : 0) : (neg1 ? 3 : 2), i3 = negTop ? (neg0 ? 0 : 1) : (neg1 ? 2 : 3);
#define B4_SEL(arr, i) ((i) == 0 ? arr[0] : (i) == 1 ? arr[1] : (i) == 2 ? arr[2] : arr[3])
                    const bool q0 = B4_SEL(p, i0), q1 = B4_SEL(p, i1), q2 = B4_SEL(p, i2), q3 = B4_SEL(p, i3);
                    const uint32_t r0 = B4_SEL(rf, i0), r1 = B4_SEL(rf, i1), r2 = B4_SEL(rf, i2), r3 = B4_SEL(rf, i3);
                    const float t0 = B4_SEL(t, i0), t1 = B4_SEL(t, i1), t2 = B4_SEL(t, i2), t3 = B4_SEL(t, i3);
#undef B4_SEL
                    (void)t0;
                    if (q3 && (q0 || q1 || q2)) B4_PUSH(r3, t3);
                    if (q2 && (q0 || q1)) B4_PUSH(r2, t2);
                    if (q1 && q0) B4_PUSH(r1, t1);
                    if (q0) cur = r0;
                    else if (q1) cur = r1;
                    else if (q2) cur = r2;
                    else if (q3) cur = r3;
                    else B4_POP();
                }
            } else {
                if (cur & B200RT_LEAF_BIT) {
                    const float4 *rec = S.tris + 3 * (size_t)(cur & ~B200RT_LEAF_BIT);
                    const float4 v0 = __ldg(rec), v1 = __ldg(rec + 1), v2 = __ldg(rec + 2);
                    bool stop = false;
                    if (prim_closest<ANY>(r, v0, v1, v2, h, S.spheres, rays + 2 * (size_t)rayIdx + 1)) {
                        hit = true;
                        if (ANY) stop = true;
                    }
                    if (stop) cur = B200RT_DONE;
                    else if (__float_as_uint(v1.w) & B200RT_TRI_LAST) B4_POP();
                    else cur += 1;
                }
            }
            if (live && cur == B200RT_DONE) {
                live = false;
                if (ANY) occluded[rayIdx] = hit ? 1 : 0;
                else {
                    b200rt_hit o; o.prim = h.prim; o.t = h.t; o.b0 = h.b0; o.b1 = h.b1; o.b2 = h.b2;
                    hits[rayIdx] = o;
                }
            }
        }
    }
#undef B4_PUSH
#undef B4_POP
    warp_exit(pc, lane);
}

// ---- kd-tree / RBSP / RBSPKd ---------------------------------------------------------------------
// Same persistent-warp + vote skeleton over the reference's 8-byte nodes (kdtreeaccel.cpp:381-521,
// rbsp.cpp:405-547, rbspKd.cpp).  Lane states: walking interior nodes (cheap step: one 8-byte fetch, a
// plane distance, a push at most), or iterating the primitives of a leaf (expensive step: one
// triangle).  Interior steps are cheap, so a lane takes up to B200RT_BSP_STEPS of them per vote.
#define B200RT_BSP_STEPS 4
#define B200RT_ST_DONE 0u
#define B200RT_ST_NODE 1u
#define B200RT_ST_LEAF 2u

// PF (8-byte node kinds only): the below child is the next record (i + 1), so its words are fetched together with node i's —
// two independent loads, usually of the same 32-byte sector — and a step that goes below costs no dependent load
// (B200RT_KD_PREFETCH=1; A/B in profiles/r02_ab_kd_prefetch.log).  The node array is padded by one record for it.
template <int KIND, bool ANY, bool PF = false>
__global__ void __launch_bounds__(B200RT_PBLOCK, 8)
bsp_persistent_kernel(const __grid_constant__ BspView S, const float4 *__restrict__ rays, uint64_t n64,
                      const uint32_t *__restrict__ n_dev, b200rt_hit *__restrict__ hits,
                      uint8_t *__restrict__ occluded, PersistCounters *pc, uint32_t thresh, uint32_t chunk,
                      uint32_t sdepth /* todo entries [0, sdepth) live in shared memory (see bvh_persistent_kernel) */,
                      uint32_t stepsW /* bits 0-7: interior steps a lane may take per vote; 8-15 / 16-23: the vote's weights (pack_steps) */) {
    extern __shared__ uint32_t sst32[];  // 3 planes (node, tMin, tMax) of sdepth x blockDim words: bank = lane
    uint32_t n = (uint32_t)n64;
    if (n_dev) { const uint32_t m = *n_dev; n = m < n ? m : n; }
    // the host sized `chunk` for the queue's CAPACITY; a queue that cannot give every warp one chunk is handed out in 32-ray
    // chunks, so that a thin deep-bounce queue spreads over the machine instead of running 128 rays a warp, four after another
    if (chunk > 32u && n < gridDim.x * (blockDim.x >> 5) * chunk) chunk = 32u;
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const uint32_t off = KIND == B200RT_KDTREE ? 2u : S.bitOffset;
    const uint32_t mask = KIND == B200RT_KDTREE ? 3u : S.bitMask;
    const uint32_t leafTag = KIND == B200RT_KDTREE ? 3u : S.leafTag;

    RayCtx r;
    HitRec h;
    uint32_t state = B200RT_ST_DONE, rayIdx = 0;
    uint32_t cur = 0;            // node index (ST_NODE)
    uint32_t li = 0, lend = 0;   // ST_LEAF: next / end position; single-primitive leaves use lend = li + 1
    bool lsingle = false;        // ST_LEAF: li is the primitive itself (onePrimitive), not an index position
    float tMin = 0, tMax = 0;
    bool hit = false;
    uint4 todo[64];              // (node, tMin, tMax, -): one 16-byte local store / load per push / pop
    int todoPos = 0;
    WarpQueue wq;
    uint2 wPre = make_uint2(0u, 0u);   // PF: the words of node `cur` when havePre
    bool havePre = false;

    for (;;) {
        const unsigned idle = __ballot_sync(FULL, state == B200RT_ST_DONE);
        if (idle && !wq.drained) {
            const uint32_t got = warp_claim(wq, pc, n, chunk, idle, state == B200RT_ST_DONE, lane);
            if (got != 0xFFFFFFFFu) {
                rayIdx = got;
                ray_setup(r, __ldg(rays + 2 * (size_t)rayIdx), __ldg(rays + 2 * (size_t)rayIdx + 1));
                h.prim = B200RT_MISS; h.t = r.tMax; h.b0 = h.b1 = h.b2 = 0.f;
                hit = false;
                todoPos = 0;
                cur = 0;
                havePre = false;
                if (root_clip(r, S.bounds, tMin, tMax)) state = B200RT_ST_NODE;  // kdtreeaccel.cpp:385
                else {
                    if (ANY) occluded[rayIdx] = 0;
                    else {
                        b200rt_hit o; o.prim = B200RT_MISS; o.t = h.t; o.b0 = o.b1 = o.b2 = 0.f;
                        hits[rayIdx] = o;
                    }
                }
            }
        }
        if (__ballot_sync(FULL, state != B200RT_ST_DONE) == 0) {
            if (wq.drained) break;
            continue;
        }
        for (;;) {
            const unsigned mN = __ballot_sync(FULL, state == B200RT_ST_NODE);
            const unsigned mL = __ballot_sync(FULL, state == B200RT_ST_LEAF);
            const uint32_t nN = __popc(mN), nL = __popc(mL);
            if (nN + nL == 0 || (!wq.drained && nN + nL < thresh)) break;
            bool finished = false;
            if (nN * ((stepsW >> 8) & 0xffu) >= nL * (stepsW >> 16)) {
#pragma unroll 1
                for (uint32_t k = 0; k < (stepsW & 0xffu) && state == B200RT_ST_NODE; ++k) {
                    if (!ANY && r.tMax < tMin) { finished = true; break; }  // kdtreeaccel.cpp:400 (closest-hit only)
                    BspStep nd;
                    uint2 wBelow = make_uint2(0u, 0u);
                    if (PF && !BspCodec<KIND>::wide) {
                        const uint2 w = havePre ? wPre : __ldg(S.nodes + cur);
                        wBelow = __ldg(S.nodes + cur + 1);
                        nd = bsp_decode<KIND>(S, r, w, off, mask, leafTag);
                        havePre = false;
                    } else nd = bsp_fetch<KIND>(S, r, cur, off, mask, leafTag);
                    if (!nd.leaf) {
                        const float tPlane = nd.tPlane;
                        const uint32_t above = nd.payload;
                        const uint32_t first = nd.belowFirst ? cur + 1 : above, second = nd.belowFirst ? above : cur + 1;
                        if (PF && !BspCodec<KIND>::wide) {   // the node visited next is the below child in these two cases
                            const bool goesFirst = !(tPlane < tMin) || tPlane > tMax || tPlane <= 0;
                            if ((goesFirst ? first : second) == cur + 1) { wPre = wBelow; havePre = true; }
                        }
                        if (tPlane > tMax || tPlane <= 0) cur = first;
                        else if (tPlane < tMin) cur = second;
                        else {
                            if ((uint32_t)todoPos < sdepth) {
                                uint32_t *q = sst32 + todoPos * B200RT_PBLOCK + threadIdx.x;
                                q[0] = second; q[sdepth * B200RT_PBLOCK] = __float_as_uint(tPlane);
                                q[2 * sdepth * B200RT_PBLOCK] = __float_as_uint(tMax);
                            } else todo[todoPos - sdepth] = make_uint4(second, __float_as_uint(tPlane), __float_as_uint(tMax), 0u);
                            ++todoPos;
                            cur = first;
                            tMax = tPlane;
                        }
                    } else {
                        const uint32_t np = nd.payload;
                        if (np == 0) {
                            if (todoPos > 0) {
                                --todoPos;
                            uint4 e;
                            if ((uint32_t)todoPos < sdepth) {
                                const uint32_t *q = sst32 + todoPos * B200RT_PBLOCK + threadIdx.x;
                                e = make_uint4(q[0], q[sdepth * B200RT_PBLOCK], q[2 * sdepth * B200RT_PBLOCK], 0u);
                            } else e = todo[todoPos - sdepth];
                                cur = e.x; tMin = __uint_as_float(e.y); tMax = __uint_as_float(e.z);
                            } else finished = true;
                        } else {
                            state = B200RT_ST_LEAF;
                            lsingle = np == 1;
                            li = nd.w0; lend = nd.w0 + np;  // np == 1: li IS the primitive (kdtreeaccel.cpp:112-118)
                        }
                    }
                    if (finished) break;
                }
            } else if (state == B200RT_ST_LEAF) {
                const uint32_t p = lsingle ? li : __ldg(S.indices + li);
                if (prim_closest<ANY>(r, S.tris + 3 * (size_t)p, h, S.spheres)) {
                    hit = true;
                    if (ANY) finished = true;
                }
                if (!finished && ++li == lend) {
                    if (todoPos > 0) {
                        --todoPos;
                            uint4 e;
                            if ((uint32_t)todoPos < sdepth) {
                                const uint32_t *q = sst32 + todoPos * B200RT_PBLOCK + threadIdx.x;
                                e = make_uint4(q[0], q[sdepth * B200RT_PBLOCK], q[2 * sdepth * B200RT_PBLOCK], 0u);
                            } else e = todo[todoPos - sdepth];
                        cur = e.x; tMin = __uint_as_float(e.y); tMax = __uint_as_float(e.z);
                        state = B200RT_ST_NODE;
                    } else finished = true;
                }
            }
            if (finished) {
                state = B200RT_ST_DONE;
                if (ANY) occluded[rayIdx] = hit ? 1 : 0;
                else {
                    b200rt_hit o; o.prim = h.prim; o.t = h.t; o.b0 = h.b0; o.b1 = h.b1; o.b2 = h.b2;
                    hits[rayIdx] = o;
                }
            }
        }
    }
    warp_exit(pc, lane);
}

}  // namespace b200rt
