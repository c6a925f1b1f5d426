// rt_wavefront.cu -- the wavefront pipeline: ray queues in HBM, three kinds of kernels.
//
//   wave_gen           primary-ray generation (OpenCL-Raytracer.cpp:18-26,68-72 on the device) into queue 0
//   wave_trace_*       nearest hit / any hit of every ray of a queue:
//                        _bvh     persistent warps, dynamic ray fetch, while-while traversal, deferred exact tests
//                        _filter  N-body style: rays in registers, object tiles staged in shared memory
//                        _brute   the reference's arithmetic on every pair (correctness baseline)
//   wave_advance       consumes the hits, advances every path's state machine (rt_wave.cuh), compacts the
//                      surviving rays into the next queue with a warp ballot, writes finished pixels as
//                      128-bit stores
// All generations of a frame are enqueued back to back; queue sizes live on the device and an empty
// generation costs an early-exit launch, never a host round trip.
#include "rt_wavefront.cuh"

#include <stdlib.h>

#include <algorithm>

#include "rt_grid3.cuh"
#include "rt_pgrid.cuh"
#include "rt_wave.cuh"

namespace rt {
namespace {

constexpr int kBlock = 256;
constexpr int kTileW = 32, kTileH = 8;
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

__device__ void flush_local(const LocalCounters& lc, Counters* c) {
    const unsigned lane = threadIdx.x & 31u;
#define RT_FLUSH(field)                                    \
    {                                                      \
        const unsigned long long s = warp_sum64(lc.field); \
        if (lane == 0 && s) atomicAdd(&c->field, s);       \
    }
    RT_FLUSH(ref_primary)
    RT_FLUSH(ref_shadow)
    RT_FLUSH(ref_reflect)
    RT_FLUSH(traced_shadow)
    RT_FLUSH(traced_reflect)
    RT_FLUSH(pair_tests)
    RT_FLUSH(exact_tests)
    RT_FLUSH(node_visits)
    RT_FLUSH(hit_pixels)
#undef RT_FLUSH
}

// Append one entry per participating thread of the block to a queue: ONE atomic per block, slots of a block
// contiguous, order within the block kept.  (One atomic per warp -- ballot + popc -- put 1 M same-address atomics per
// 8K generation on a single counter; they serialise in L2 and cost 6 ms of the 67 ms frame.)  Every thread of the
// block must call it; sh holds kBlock/32 words.
__device__ __forceinline__ uint32_t block_append(bool emit, uint32_t* counter, uint32_t* sh) {
    constexpr unsigned kWarps = kBlock / 32;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned m = __ballot_sync(kFull, emit);
    if (lane == 0) sh[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (warp == 0) {
        const uint32_t c = lane < kWarps ? sh[lane] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (unsigned o = 1; o < kWarps; o <<= 1) {
            const uint32_t v = __shfl_up_sync(kFull, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t total = __shfl_sync(kFull, incl, kWarps - 1);
        uint32_t base = 0;
        if (lane == 0 && total) base = atomicAdd(counter, total);
        base = __shfl_sync(kFull, base, 0);
        if (lane < kWarps) sh[lane] = base + incl - c;
    }
    __syncthreads();
    return sh[warp] + (uint32_t)__popc(m & lanemask_lt());
}

// The same for the two ends of a queue at once (class B: front, class A: back): both per-warp counts travel through the
// scan packed into one word (a block emits at most 256 of each), lanes 0 and 1 of warp 0 do the two atomics.  Returns the
// slot in the queue end the thread emits to (emit_a wins if both are set).  sh holds 2 * kBlock/32 words.
template <int BLOCK = kBlock>
__device__ __forceinline__ uint32_t block_append2(bool emit_b, bool emit_a, uint32_t* count_b, uint32_t* count_a, uint32_t* sh) {
    constexpr unsigned kWarps = BLOCK / 32;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned ma = __ballot_sync(kFull, emit_a), mb = __ballot_sync(kFull, emit_b && !emit_a);
    if (lane == 0) sh[warp] = (uint32_t)__popc(mb) | ((uint32_t)__popc(ma) << 16);
    __syncthreads();
    if (warp == 0) {
        const uint32_t c = lane < kWarps ? sh[lane] : 0u;
        uint32_t incl = c;
#pragma unroll
        for (unsigned o = 1; o < kWarps; o <<= 1) {
            const uint32_t v = __shfl_up_sync(kFull, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t total = __shfl_sync(kFull, incl, kWarps - 1);
        uint32_t base = 0;
        if (lane == 0 && (total & 0xffffu)) base = atomicAdd(count_b, total & 0xffffu);
        if (lane == 1 && (total >> 16)) base = atomicAdd(count_a, total >> 16);
        const uint32_t base_b = __shfl_sync(kFull, base, 0), base_a = __shfl_sync(kFull, base, 1);
        const uint32_t excl = incl - c;
        // two 32-bit bases do not fit the packed word: store B's slot base here, A's in the second half of sh
        if (lane < kWarps) {
            sh[lane] = base_b + (excl & 0xffffu);
            sh[kWarps + lane] = base_a + (excl >> 16);
        }
    }
    __syncthreads();
    if (emit_a) return sh[kWarps + warp] + (uint32_t)__popc(ma & lanemask_lt());
    return sh[warp] + (uint32_t)__popc(mb & lanemask_lt());
}

// ------------------------------------------------------------------------------------------------ generation 0
// to_grid != 0: the camera has a perspective grid, the primary rays are class A tasks (back of the queue, counter count[1])
__global__ void __launch_bounds__(kBlock) wave_gen(const Frame fr, const TaskQueue q, const PathArrays pa, uint32_t* __restrict__ count, const uint32_t strip,
                                                   const uint32_t to_grid, const uint32_t qcap) {
    bool active;
    uint64_t oi = 0;
    Ray ray;
    if (fr.rays) {
        oi = (uint64_t)blockIdx.x * kBlock + threadIdx.x;
        active = oi < fr.n_local;
        if (active) {
            const float4* rp = reinterpret_cast<const float4*>(fr.rays + oi);
            ray.s = __ldg(rp);
            ray.d = __ldg(rp + 1);
        }
    } else {
        const uint32_t tiles_x = (fr.width + kTileW - 1) / kTileW;
        uint32_t tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
        if (strip > 1u) {  // tiles run down columns inside strips of `strip` tile rows: the queue order is compact in 2-D
            const uint32_t tiles_y = (fr.local_rows + kTileH - 1) / kTileH;
            const uint32_t per = tiles_x * strip, st = blockIdx.x / per, in = blockIdx.x % per;
            const uint32_t h = min(strip, tiles_y - st * strip);
            tx = in / h;
            ty = st * strip + in % h;
        }
        const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
        const uint32_t lx = tx * kTileW + (warp & 3u) * 8u + (lane & 7u);
        const uint32_t ly = ty * kTileH + (warp >> 2) * 4u + (lane >> 3);
        active = lx < fr.width && ly < fr.local_rows;
        if (active) {
            const uint32_t lrow = fr.lrow0 + ly;
            const uint32_t blk = lrow / fr.block_rows, rib = lrow % fr.block_rows;
            const uint32_t row = fr.row0 + (blk * fr.n_shards + fr.shard) * fr.block_rows + rib;
            active = row < fr.height;
            if (active) {
                oi = (uint64_t)row * fr.width + lx;
                ray = camera_ray(lx, row, (float)fr.width / 2.0f, (float)fr.height / 2.0f, (float)fr.height, fr.zdir);
            }
        }
    }
    __shared__ uint32_t sh_append[kBlock / 32];
    const uint32_t slot = block_append(active, count + (to_grid ? 1 : 0), sh_append);
    if (active) {
        const uint32_t at = to_grid ? qcap - 1u - slot : slot;
        q.s[at] = ray.s;
        q.d[at] = ray.d;
        q.path[at] = slot | (RT_TASK_NO_GRID << RT_TASK_LIGHT_SHIFT);  // path id == slot of the primary task
        pa.out[slot] = (uint32_t)oi;
    }
}

// ------------------------------------------------------------------------------------------------ trace: BVH
#ifndef RT_PEND_CAP
#define RT_PEND_CAP 8
#endif
constexpr int kPendCap = RT_PEND_CAP;  // queued candidates per lane (a lane stops traversing while it holds one)

// The reference arithmetic for one (ray, object) pair, kept out of line: it is long (two 4x4 transforms,
// IEEE divides and a square root) and rare, and inlining it into the hot loops thrashes the instruction cache.
// The current best hit travels by VALUE: a reference parameter of a function that is not inlined forces the caller to keep
// best.t in local memory, and the traversal loop then reloads it (LDL) in every node step.
struct BestHit {
    float t;
    int32_t id;
    uint32_t type;
};
__device__ __noinline__ BestHit exact_pair_value(const Scene& sc, const float4 s, const float4 d, uint32_t obj, uint32_t type, BestHit cur) {
    Ray r;
    r.s = s;
    r.d = d;
    Hit best;
    best.t = cur.t;
    best.id = cur.id;
    uint32_t btype = cur.type;
    exact_pair(sc, r, obj, type, best, btype);
    cur.t = best.t;
    cur.id = best.id;
    cur.type = btype;
    return cur;
}
__device__ __forceinline__ void exact_pair_call(const Scene& sc, const float4 s, const float4 d, uint32_t obj, uint32_t type, Hit& best,
                                                uint32_t& btype) {
    BestHit b;
    b.t = best.t;
    b.id = best.id;
    b.type = btype;
    b = exact_pair_value(sc, s, d, obj, type, b);
    best.t = b.t;
    best.id = b.id;
    btype = b.type;
}

// (Requesting a survivor's inverse matrix -- the first thing its exact test loads -- with a prefetch when the survivor is queued
// was measured on the 8K cfg5 frame and dropped: 23.5 ms against 23.2 without.)
// Persistent warps.  Every lane owns one ray of the queue and is in one of four states:
//   T  traversing (stands on an inner node)           S  traversing, but its candidate buffer is nearly full
//   F  traversal finished, candidates still queued    I  idle (no ray)
// A node step tests the two child boxes; a child that is a leaf is not visited -- its objects go straight into
// the lane's candidate buffer (shared memory), without a per-object filter: with the default one object per leaf
// the child box IS the object's padded bounding box -- so the traversal loop consists of node steps only.  The warp
// then schedules whole-warp rounds of the three activities so that each runs with most lanes taking part:
//   node steps   while >= 12 lanes are in T (or nothing else is worth doing)
//   exact tests  when >= 4 lanes have a candidate queued or nobody can traverse
//   refill       when >= 16 lanes are idle (or nobody is traversing) and the queue is not exhausted
// (thresholds swept on the B200 with tools/sweep_sched.sh, again after the queue became tile-contiguous: 16/8/12 ->
// 12/4/16, 58.8 -> 57.6 ms per 8K frame; RT_WAVE_MIN_T / _DRAIN / _REFILL override them)
// Deferring exact tests never changes the result: every queued candidate is tested before its ray retires, and
// pruning against the best t found so far only ever skips nodes that cannot contain a better hit.
#ifndef RT_BVH_MINB
#define RT_BVH_MINB 3  // resident CTAs per SM of the persistent trace kernel (80 registers)
#endif
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, RT_BVH_MINB) wave_trace_bvh(const Scene sc, const TaskQueue q, const uint32_t* __restrict__ n_ptr,
                                                            uint32_t* __restrict__ fetch, float* __restrict__ res_t,
                                                            int32_t* __restrict__ res_id, const int min_t, const int drain_t,
                                                            const int refill_t, Counters* ctr) {
    const uint32_t n = *n_ptr;
    if (n == 0u) return;
    __shared__ uint32_t pend[kPendCap][kBlock];
    const unsigned lane = threadIdx.x & 31u, tid = threadIdx.x;
    LocalCounters lc;

    // lane state: 0 idle, 1 traversing, 2 traversal finished (candidates may still be queued) -- one register, one compare per
    // vote; tlim: the pruning limit (best t, capped at 1 for a shadow ray), kept up to date where best.t changes
    int st = 0;
    bool anyhit = false;
    float tlim = RT_MAX_FLOAT;
    uint32_t task = 0;
    float ox = 0.f, oy = 0.f, oz = 0.f, idx = 0.f, idy = 0.f, idz = 0.f;
    Hit best;
    best.t = RT_MAX_FLOAT;
    best.id = -1;
    uint32_t btype = 0;
    int32_t cur = 0;
    int sp = 0, npend = 0;
    int32_t stack[RT_BVH_STACK];
    bool exhausted = false;

    for (;;) {
        // ---- refill
        {
            const unsigned idle = __ballot_sync(kFull, st == 0);
            const unsigned trav = __ballot_sync(kFull, st == 1);
            if (idle && !exhausted && (__popc(idle) >= refill_t || trav == 0u)) {
                // (one same-address atomic per refill; reserving 32..256-ray chunks per warp instead -- guided
                // self-scheduling -- was measured equal on the 8K frame, fixed chunks slower by their tails)
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(fetch, (uint32_t)__popc(idle));
                base = __shfl_sync(kFull, base, 0);
                if (st == 0) {
                    const uint32_t i = base + (uint32_t)__popc(idle & lanemask_lt());
                    if (i < n) {
                        task = i;
                        const float4 s = q.s[i], d = q.d[i];
                        anyhit = (q.path[i] & RT_TASK_ANYHIT) != 0u;
                        ox = s.x; oy = s.y; oz = s.z;
                        // conservative use only: approximate reciprocals are well inside the 1e-5 slack
                        idx = fabsf(d.x) < 1e-30f ? copysignf(1e30f, d.x) : __fdividef(1.f, d.x);
                        idy = fabsf(d.y) < 1e-30f ? copysignf(1e30f, d.y) : __fdividef(1.f, d.y);
                        idz = fabsf(d.z) < 1e-30f ? copysignf(1e30f, d.z) : __fdividef(1.f, d.z);
                        best.t = RT_MAX_FLOAT;
                        best.id = -1;
                        btype = 0;
                        cur = sc.root;
                        sp = 0;
                        npend = 0;
                        st = 1;
                        tlim = anyhit ? 1.f : RT_MAX_FLOAT;
                        if (cur < 0) {  // the whole scene is one leaf (tiny scenes): test its objects in place
                            const uint32_t code = (uint32_t)(~cur);
                            const uint32_t first = code >> 4, cnt = sc.n_objs ? (code & 15u) + 1u : 0u;
                            for (uint32_t j = 0; j < cnt; ++j) {
                                exact_pair_call(sc, s, d, __ldg(sc.leaf_obj + first + j), __ldg(&sc.leaf_cull[first + j].type), best, btype);
                                if (COUNT) lc.exact_tests++;
                            }
                            st = 2;
                        }
                    }
                }
                if (base + (uint32_t)__popc(idle) >= n) exhausted = true;
            }
            if (!__any_sync(kFull, st != 0)) break;
        }

        // ---- node steps
        for (;;) {
            // a lane that holds a candidate waits for its exact test: nothing is traversed with a stale best t
            const bool go = st == 1 && npend == 0;
            const unsigned T = __ballot_sync(kFull, go);
            if (T == 0u) break;
            if (__popc(T) < min_t) {
                if (__popc(__ballot_sync(kFull, npend > 0)) >= drain_t) break;
                if (!exhausted && __popc(__ballot_sync(kFull, st == 0 || (st == 2 && npend == 0))) >= refill_t) break;
            }
            if (go) {
                const float4* np = reinterpret_cast<const float4*>(sc.nodes + cur);
                const float4 q0 = __ldg(np + 0), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
                if (COUNT) lc.node_visits++;
                const float tlimit = tlim;
                float tn0, tn1;
                // child 0: ctr (q0.x q0.y q0.z) ext (q0.w q1.x q1.y); child 1: ctr (q1.z q1.w q2.x) ext (q2.y q2.z q2.w)
                bool h0 = cv::box_may_hit(ox, oy, oz, idx, idy, idz, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q3.z, tlimit, tn0);
                bool h1 = cv::box_may_hit(ox, oy, oz, idx, idy, idz, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, q3.w, tlimit, tn1);
                const int32_t c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
                if ((h0 && c0 < 0) || (h1 && c1 < 0)) {  // leaf children: their objects become candidates
#pragma unroll 1
                    for (int side = 0; side < 2; ++side) {
                        const int32_t c = side ? c1 : c0;
                        if (!(side ? h1 : h0) || c >= 0) continue;
                        const uint32_t code = (uint32_t)(~c);
                        const uint32_t first = code >> 4, cnt = (code & 15u) + 1u;
                        for (uint32_t j = 0; j < cnt; ++j) {
                            if (npend < kPendCap) {
                                pend[npend++][tid] = first + j;
                            } else {  // buffer full (leaves larger than the buffer): test in place
                                exact_pair_call(sc, q.s[task], q.d[task], __ldg(sc.leaf_obj + first + j),
                                                __ldg(&sc.leaf_cull[first + j].type), best, btype);
                                tlim = anyhit ? fminf(best.t, 1.f) : best.t;
                                if (COUNT) lc.exact_tests++;
                            }
                        }
                    }
                    h0 = h0 && c0 >= 0;
                    h1 = h1 && c1 >= 0;
                }
                if (h0 && h1) {
                    const bool first0 = tn0 <= tn1;
                    cur = first0 ? c0 : c1;
                    stack[sp++] = first0 ? c1 : c0;
                } else if (h0 || h1) {
                    cur = h0 ? c0 : c1;
                } else if (sp > 0) {
                    cur = stack[--sp];
                } else {
                    st = 2;
                }
            }
        }

        // ---- exact tests: one queued candidate per lane and round (a dense variant -- all candidates of the warp spread
        // over its 32 lanes, owners merging the results -- was measured 1.5 % slower: the scan, the owner search and the
        // staging through shared memory cost more than the idle lanes)
        for (;;) {
            const bool has = npend > 0;
            const unsigned P = __ballot_sync(kFull, has);
            if (P == 0u) break;
            const unsigned T = __ballot_sync(kFull, st == 1 && npend == 0);
            if (!(__popc(P) >= drain_t || T == 0u)) break;
            if (has) {
                const uint32_t slot = pend[--npend][tid];
                exact_pair_call(sc, q.s[task], q.d[task], __ldg(sc.leaf_obj + slot), __ldg(&sc.leaf_cull[slot].type), best, btype);
                if (COUNT) lc.exact_tests++;
                tlim = anyhit ? fminf(best.t, 1.f) : best.t;
                if (anyhit && best.t < 1.f) {  // the shadow ray is decided (SAR:229)
                    st = 2;
                    npend = 0;
                }
            }
        }

        // ---- retire
        if (st == 2 && npend == 0) {
            res_t[task] = best.id >= 0 ? best.t : RT_MAX_FLOAT;
            res_id[task] = best.id;
            st = 0;
        }
    }
    if (COUNT) flush_local(lc, ctr);
}

// ------------------------------------------------------------------------------------------------ trace: perspective grids
// Class A tasks (rt_wave.cuh): camera rays and shadow rays of point lights.  Same persistent-warp scheme as the BVH
// kernel, but a lane's "traversal" is a walk along ONE sorted list -- the cell of its direction as seen from the apex
// (rt_pgrid.cuh): load the next entry, stop if it lies beyond the current hit (camera) or behind the ray start (shadow),
// bounding-sphere filter, and a survivor waits for the warp's next round of exact tests.  Rays of one 8x4 pixel tile
// share their cell, so a warp of primary rays walks its list in lockstep with broadcast loads.
#ifndef RT_GRID_MINB
#define RT_GRID_MINB 5  // 48 registers (4: 56); the list walk waits on L2, a fifth CTA is worth 0.4 ms per frame
#endif
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, RT_GRID_MINB) wave_trace_grid(const Scene sc, const GridSet cam, const GridSet lit, const TaskQueue q,
                                                                       const uint32_t* __restrict__ n_ptr, uint32_t* __restrict__ fetch,
                                                                       float* __restrict__ res_t, int32_t* __restrict__ res_id, const uint32_t qcap,
                                                                       const int min_t, const int drain_t, const int refill_t, Counters* ctr) {
    const uint32_t n = *n_ptr;
    if (n == 0u) return;
    const unsigned lane = threadIdx.x & 31u;
    LocalCounters lc;
    const uint32_t kNone = 0xffffffffu;
    const float4* cull4 = reinterpret_cast<const float4*>(sc.cull);

    bool have = false, anyhit = false;
    uint32_t at = 0;  // index of the lane's task in the queue arrays
    float ox = 0.f, oy = 0.f, oz = 0.f, ux = 0.f, uy = 0.f, uz = 0.f, dlen = 0.f, cutoff = 0.f;
    Hit best;
    best.t = RT_MAX_FLOAT;
    best.id = -1;
    uint32_t btype = 0, cur = 0, end = 0, pend = kNone, ptype = 0;
    const GridEntry* ent = cam.entries;
    bool exhausted = false;

    for (;;) {
        // ---- refill
        {
            const unsigned idle = __ballot_sync(kFull, !have);
            const unsigned walk = __ballot_sync(kFull, have && pend == kNone && cur < end);
            if (idle && !exhausted && (__popc(idle) >= refill_t || walk == 0u)) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(fetch, (uint32_t)__popc(idle));
                base = __shfl_sync(kFull, base, 0);
                if (!have) {
                    const uint32_t i = base + (uint32_t)__popc(idle & lanemask_lt());
                    if (i < n) {
                        at = qcap - 1u - i;
                        const float4 s = q.s[at], d = q.d[at];
                        const uint32_t pw = q.path[at];
                        anyhit = (pw & RT_TASK_ANYHIT) != 0u;
                        ox = s.x; oy = s.y; oz = s.z;
                        const float l2 = fmaf(d.z, d.z, fmaf(d.y, d.y, d.x * d.x));
                        dlen = sqrtf(l2);
                        const float il = 1.f / dlen;
                        ux = d.x * il; uy = d.y * il; uz = d.z * il;
                        PGrid g;
                        float wx = d.x, wy = d.y, wz = d.z;
                        const uint32_t* cs;
                        if (anyhit) {
                            wx = -wx; wy = -wy; wz = -wz;
                            const uint32_t li = (pw >> RT_TASK_LIGHT_SHIFT) & 15u;
                            g = lit.grids[6u * li + pgrid_face(wx, wy, wz)];
                            cs = lit.cell_start;
                            ent = lit.entries;
                            const float ax = s.x - g.ax, ay = s.y - g.ay, az = s.z - g.az;
                            cutoff = sqrtf(fmaf(az, az, fmaf(ay, ay, ax * ax))) * 1.0001f + 1e-4f;
                        } else {
                            g = cam.grids[0];
                            cs = cam.cell_start;
                            ent = cam.entries;
                            cutoff = RT_MAX_FLOAT;
                        }
                        const uint32_t cell = pgrid_cell(g, wx, wy, wz);
                        cur = __ldg(cs + cell);
                        end = __ldg(cs + cell + 1);
                        best.t = RT_MAX_FLOAT;
                        best.id = -1;
                        btype = 0;
                        pend = kNone;
                        have = true;
                    }
                }
                if (base + (uint32_t)__popc(idle) >= n) exhausted = true;
            }
            if (!__any_sync(kFull, have)) break;
        }

        // ---- list steps
        for (;;) {
            const bool go = have && pend == kNone && cur < end;
            const unsigned T = __ballot_sync(kFull, go);
            if (T == 0u) break;
            if (__popc(T) < min_t) {
                if (__popc(__ballot_sync(kFull, have && pend != kNone)) >= drain_t) break;
                if (!exhausted && __popc(__ballot_sync(kFull, !have || (pend == kNone && cur >= end))) >= refill_t) break;
            }
            if (go) {
                const GridEntry e = ent[cur];
                ++cur;
                if (e.dnear > cutoff) {
                    cur = end;
                } else {
                    const float4 a = __ldg(cull4 + 2 * (size_t)e.obj), b = __ldg(cull4 + 2 * (size_t)e.obj + 1);
                    if (COUNT) lc.pair_tests++;
                    const float x = a.x - ox, y = a.y - oy, z = a.z - oz;
                    const float oc2 = fmaf(z, z, fmaf(y, y, x * x));
                    const float dt = fmaf(z, uz, fmaf(y, uy, x * ux));
                    if (!(fmaf(-dt, dt, oc2 * b.x) > a.w)) {  // cv::sphere_may_hit
                        pend = e.obj;
                        ptype = __float_as_uint(b.w);
                    }
                }
            }
        }

        // ---- exact tests: when enough lanes hold a survivor, or nobody can walk on
        {
            const bool has = have && pend != kNone;
            const unsigned P = __ballot_sync(kFull, has);
            const unsigned W = __ballot_sync(kFull, have && pend == kNone && cur < end);
            if (P != 0u && (__popc(P) >= drain_t || W == 0u)) {
                if (has) {
                    exact_pair_call(sc, q.s[at], q.d[at], pend, ptype, best, btype);
                    if (COUNT) lc.exact_tests++;
                    pend = kNone;
                    if (anyhit) {
                        if (best.t < 1.f) cur = end;  // the shadow ray is decided (SAR:229)
                    } else if (best.id >= 0) {
                        cutoff = best.t * dlen * 1.0001f + 1e-4f;  // distance apex -> hit point (the ray starts at the apex)
                    }
                }
            }
        }

        // ---- retire
        if (have && pend == kNone && cur >= end) {
            res_t[at] = best.id >= 0 ? best.t : RT_MAX_FLOAT;
            res_id[at] = best.id;
            have = false;
        }
    }
    if (COUNT) flush_local(lc, ctr);
}

// ------------------------------------------------------------------------------------------------ trace: uniform grid, DDA
// Class B tasks (everything without a common apex) on the uniform 3-D grid of rt_grid3.cuh instead of the BVH: per lane a
// DDA walk, per cell a short object list through the bounding-sphere filter, survivors wait for the warp's next round of
// exact tests.  An iteration of the walk loop moves a lane whose list is used up into the next cell AND tests one entry, so
// the two short code paths do not alternate between warp votes.
#ifndef RT_DDA_MINB
#define RT_DDA_MINB 4  // 64 registers: the walk is latency bound, a fourth resident CTA is worth 9 % (27.2 -> 24.9 ms per frame; 5: 25.8)
#endif
template <bool COUNT>
__global__ void __launch_bounds__(kBlock, RT_DDA_MINB) wave_trace_dda(const Scene sc, const Grid3 g, const TaskQueue q, const uint32_t* __restrict__ n_ptr,
                                                                     uint32_t* __restrict__ fetch, float* __restrict__ res_t, int32_t* __restrict__ res_id,
                                                                     const int min_t, const int drain_t, const int refill_t, Counters* ctr) {
    const uint32_t n = *n_ptr;
    if (n == 0u) return;
    const unsigned lane = threadIdx.x & 31u;
    LocalCounters lc;
    const uint32_t kNone = 0xffffffffu;
    const float4* cull4 = reinterpret_cast<const float4*>(sc.cull);

    int st = 0;  // 0 idle, 1 walking, 2 walk finished
    bool anyhit = false;
    uint32_t at = 0;
    float ox = 0.f, oy = 0.f, oz = 0.f, ux = 0.f, uy = 0.f, uz = 0.f;
    Dda w;
    w.tmx = w.tmy = w.tmz = w.dtx = w.dty = w.dtz = w.t_end = 0.f;
    w.rx = w.ry = w.rz = w.lx = w.ly = w.lz = 0;
    w.cell = 0u;
    Hit best;
    best.t = RT_MAX_FLOAT;
    best.id = -1;
    uint32_t btype = 0, cur = 0, end = 0, pend = kNone, ptype = 0;
    uint32_t mb0 = kNone, mb1 = kNone, mb2 = kNone;  // the last objects this ray ran the exact test on
    bool exhausted = false;

    for (;;) {
        // ---- refill
        {
            const unsigned idle = __ballot_sync(kFull, st == 0);
            const unsigned walk = __ballot_sync(kFull, st == 1 && pend == kNone);
            if (idle && !exhausted && (__popc(idle) >= refill_t || walk == 0u)) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(fetch, (uint32_t)__popc(idle));
                base = __shfl_sync(kFull, base, 0);
                if (st == 0) {
                    const uint32_t i = base + (uint32_t)__popc(idle & lanemask_lt());
                    if (i < n) {
                        at = i;
                        Ray ray;
                        ray.s = q.s[i];
                        ray.d = q.d[i];
                        anyhit = (q.path[i] & RT_TASK_ANYHIT) != 0u;
                        best.t = RT_MAX_FLOAT;
                        best.id = -1;
                        btype = 0;
                        pend = kNone;
                        mb0 = mb1 = mb2 = kNone;
                        st = 1;
                        cur = end = 0;
                        const float chk = ray.s.x + ray.s.y + ray.s.z + ray.d.x + ray.d.y + ray.d.z;
                        if (!(fabsf(chk) < __int_as_float(0x7f800000))) {
                            // not finite (a hit point at infinity): the exact test on every object, as the reference does
                            for (uint32_t k = 0; k < sc.n_objs; ++k) {
                                exact_pair_call(sc, ray.s, ray.d, k, __ldg(&sc.cull[k].type), best, btype);
                                if (COUNT) lc.exact_tests++;
                            }
                            st = 2;
                        } else if (!dda_start(g, ray, anyhit ? 1.f : RT_MAX_FLOAT, w)) {
                            st = 2;  // misses the grid
                        } else {
                            ox = ray.s.x; oy = ray.s.y; oz = ray.s.z;
                            const float l = sqrtf(fmaf(ray.d.z, ray.d.z, fmaf(ray.d.y, ray.d.y, ray.d.x * ray.d.x)));
                            const float il = 1.f / l;
                            ux = ray.d.x * il; uy = ray.d.y * il; uz = ray.d.z * il;
                            cur = __ldg(g.cell_start + w.cell);
                            end = __ldg(g.cell_start + w.cell + 1);
                            if (COUNT) lc.node_visits++;
                        }
                    }
                }
                if (base + (uint32_t)__popc(idle) >= n) exhausted = true;
            }
            if (!__any_sync(kFull, st != 0)) break;
        }

        // ---- walk: a lane whose list is used up moves on by one cell (a dozen instructions), then every lane with an entry
        // left tests one
        for (;;) {
            const bool go = st == 1 && pend == kNone;
            const unsigned T = __ballot_sync(kFull, go);
            if (T == 0u) break;
            if (__popc(T) < min_t) {
                if (__popc(__ballot_sync(kFull, pend != kNone)) >= drain_t) break;
                if (!exhausted && __popc(__ballot_sync(kFull, st == 0 || (st == 2 && pend == kNone))) >= refill_t) break;
            }
            if (go) {
                if (cur >= end) {  // next cell
                    bool out;
                    const float t_exit = dda_advance(w, out);
                    // every object whose hit point lies before the exit of the cell just left has been seen
                    if ((best.id >= 0 && best.t < t_exit * (1.f - 1e-5f)) || t_exit > w.t_end || out) {
                        st = 2;
                    } else {
                        cur = __ldg(g.cell_start + w.cell);
                        end = __ldg(g.cell_start + w.cell + 1);
                        if (COUNT) lc.node_visits++;
                    }
                }
                if (st == 1 && cur < end) {
                    const uint32_t obj = __ldg(g.entries + cur);
                    ++cur;
                    const float4 a = __ldg(cull4 + 2 * (size_t)obj), b = __ldg(cull4 + 2 * (size_t)obj + 1);
                    if (COUNT) lc.pair_tests++;
                    const float x = a.x - ox, y = a.y - oy, z = a.z - oz;
                    const float oc2 = fmaf(z, z, fmaf(y, y, x * x));
                    const float dt = fmaf(z, uz, fmaf(y, uy, x * ux));
                    if (!(fmaf(-dt, dt, oc2 * b.x) > a.w) && obj != mb0 && obj != mb1 && obj != mb2) {  // cv::sphere_may_hit
                        pend = obj;
                        ptype = __float_as_uint(b.w);
                    }
                }
            }
        }

        // ---- exact tests: when enough lanes hold a survivor, or nobody can walk on
        {
            const bool has = pend != kNone;
            const unsigned P = __ballot_sync(kFull, has);
            const unsigned W = __ballot_sync(kFull, st == 1 && pend == kNone);
            if (P != 0u && (__popc(P) >= drain_t || W == 0u)) {
                if (has) {
                    exact_pair_call(sc, q.s[at], q.d[at], pend, ptype, best, btype);
                    if (COUNT) lc.exact_tests++;
                    mb2 = mb1; mb1 = mb0; mb0 = pend;
                    pend = kNone;
                    if (anyhit && best.t < 1.f) st = 2;  // the shadow ray is decided (SAR:229)
                }
            }
        }

        // ---- retire
        if (st == 2 && pend == kNone) {
            res_t[at] = best.id >= 0 ? best.t : RT_MAX_FLOAT;
            res_id[at] = best.id;
            st = 0;
        }
    }
    if (COUNT) flush_local(lc, ctr);
}

// --------------------------------------------------------------------------- trace: brute force, two flavours
// Exact arithmetic on every pair (reference order).
template <bool COUNT>
__global__ void __launch_bounds__(kBlock) wave_trace_brute(const Scene sc, const TaskQueue q, const uint32_t* __restrict__ n_ptr,
                                                           float* __restrict__ res_t, int32_t* __restrict__ res_id, Counters* ctr) {
    const uint32_t n = *n_ptr;
    LocalCounters lc;
    for (uint32_t i = blockIdx.x * kBlock + threadIdx.x; i < n; i += gridDim.x * kBlock) {
        Ray r;
        r.s = q.s[i];
        r.d = q.d[i];
        const Hit h = trace_brute(sc, r, nullptr);
        if (COUNT) {
            lc.pair_tests += sc.n_objs;
            lc.exact_tests += sc.n_objs;
        }
        res_t[i] = h.id >= 0 ? h.t : RT_MAX_FLOAT;
        res_id[i] = h.id;
    }
    if (COUNT) flush_local(lc, ctr);
}

// N-body layout: RPT rays per thread in registers, the filter records of a tile of objects in shared memory
// (every lane of a warp reads the same record: broadcast LDS.128), survivors run the exact test.
//
// The bounding-sphere filter  h*|c-o|^2 - ((c-o).u)^2 > r2  (reject) is evaluated in expanded form with the
// scene-wide smallest h (conservative: a smaller h only lets more pairs through) folded into per-ray constants:
//     dt = c.u - o.u                            3 FMA      (o.u per ray)
//     w  = 2h c.o - h*|o|^2                     3 FMA      (h*|o|^2 and 2h*o per ray)
//     l  = dt*dt + w                            1 FMA
//     reject iff l < e',  e' = h*|c|^2 - r2     (e' per object, computed when the tile is staged)
// The expanded form cancels when |c|, |o| >> |c-o|; its rounding error is below 44 u (|c|^2 + |o|^2), which is added
// to the accept side (kExpandPad on e' per object and on |o|^2 per ray).
//
// sm_100a packed arithmetic: two rays share one 64-bit register pair and every FMA above is one FFMA2
// (__ffma2_rn, two IEEE fmas per issue slot), so 8 rays x 1 object cost 28 FFMA2 = 56 FMA-pipe cycles but only 28
// issue slots; the tile stores every value twice ({cx,cx,cy,cy} {cz,cz,e',e'}) so that the broadcast operand arrives
// as an aligned pair.  "Does any of my rays pass" is a 3-input max tree over the eight l values and ONE compare, the
// warp then votes; the per-ray compares run only on the rare slow path.
//
// Survivors (about 1e-4 of the pairs) do not run the exact test where they are found -- that would be ~400
// instructions with one or two lanes active -- but are appended to a per-warp list in shared memory; whenever the
// list holds 32 entries the warp runs 32 exact tests at once, one per lane, and every owner lane then merges the
// results of its own rays (ex::replaces is order independent, so the merge order does not matter).
constexpr float kExpandPad = 4.7684e-06f;  // 80 u

// ---- TMA (bulk async copy) + mbarrier plumbing of the tile pipeline.  The filter records are packed ONCE per scene, in the
// layout the hot loop reads ({cx,cx,cy,cy} {cz,cz,e',e'}: every value twice, so that the packed FFMA2 gets its broadcast operand
// as an aligned register pair), padded to whole tiles; a tile is then one contiguous 32 KB chunk that a single thread moves
// with cp.async.bulk (UBLKCP in SASS) into one of two shared-memory buffers while the block still computes on the other.
// full[b] completes when the bytes of buffer b have landed (complete_tx), empty[b] when every warp has finished reading it.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)), "l"(src_gmem),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Packs the filter records of the scene into the tile layout (2 float4 per object, n_pad objects, n_pad a multiple of the
// tile): e' = h_min |c|^2 - r2 - pad folds the scene-wide smallest h; r2 < 0 ("never hit") and the padding get e' = +inf.
__global__ void filter_pack_tiles(const CullRec* __restrict__ cull, const uint32_t n_objs, const uint32_t n_pad, const float hmin,
                                  float4* __restrict__ out) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_pad) return;
    const float inf = __int_as_float(0x7f800000);
    float4 a = make_float4(0.f, 0.f, 0.f, -1.f), b = make_float4(1.f, 0.f, 0.f, 0.f);
    if (j < n_objs) {
        const float4* c4 = reinterpret_cast<const float4*>(cull);
        a = __ldg(c4 + 2 * (size_t)j);      // cx cy cz r2
        b = __ldg(c4 + 2 * (size_t)j + 1);  // h cc e0 type
    }
    const float e = a.w < 0.f ? inf : -fmaf(-hmin, b.y, fmaf(kExpandPad, b.y, a.w));
    out[2 * (size_t)j] = make_float4(a.x, a.x, a.y, a.y);
    out[2 * (size_t)j + 1] = make_float4(a.z, a.z, e, e);
}

template <int CAP>     // CAP: per warp, 31 left over + (rays per thread x 32 lanes) of one object
struct FilterShared {  // static part; the tile is dynamic shared memory
    uint32_t cand_obj[kBlock / 32][CAP];
    uint16_t cand_who[kBlock / 32][CAP];  // (owner lane << 3) | ray slot
    float res_t[kBlock / 32][32];              // results of one batch of exact tests (< 0: no candidate)
    uint32_t res_type[kBlock / 32][32];
    uint32_t cand_n[kBlock / 32];
};

// One batch: lanes [0, k) each run the exact arithmetic of one listed (ray, object) pair, then every lane merges the
// results that belong to its rays into best_t / best_i.  Whole warp, converged.
template <int RPT, int CAP>
__device__ __noinline__ void filter_flush(const Scene& sc, const TaskQueue& q, FilterShared<CAP>& fs, float (*best_t)[kBlock],
                                          int32_t (*best_i)[kBlock], uint32_t warp_base, uint32_t first, uint32_t k) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, tid = threadIdx.x;
    uint32_t obj = 0;
    if (lane < k) {
        obj = fs.cand_obj[warp][first + lane];
        const uint32_t who = fs.cand_who[warp][first + lane];
        const uint32_t i = warp_base + (who >> 3) + (who & 7u) * kBlock;
        const uint32_t type = __ldg(&sc.cull[obj].type);
        const float4* mi = sc.minv + 4 * (size_t)obj;
        const float4 c0 = __ldg(mi + 0), c1 = __ldg(mi + 1), c2 = __ldg(mi + 2), c3 = __ldg(mi + 3);
        const float4 s = ex::xform(c0, c1, c2, c3, q.s[i]);
        const float4 d = ex::xform(c0, c1, c2, c3, q.d[i]);
        float t = 0.f;
        const bool hit = ex::candidate(type, s, d, t);
        fs.res_t[warp][lane] = hit ? t : -1.f;  // candidates have t >= 0 (a NaN t is dropped by ex::replaces, see there)
        fs.res_type[warp][lane] = type;
    }
    __syncwarp();
    for (uint32_t e = 0; e < k; ++e) {
        const uint32_t who = fs.cand_who[warp][first + e];
        if ((who >> 3) == lane) {
            const float tc = fs.res_t[warp][e];
            if (!(tc < 0.f)) {
                const uint32_t r = who & 7u;
                const int32_t c = (int32_t)fs.cand_obj[warp][first + e];
                const int32_t kcur = best_i[r][tid];
                const uint32_t type_k = kcur >= 0 ? __ldg(&sc.cull[kcur].type) : 0u;
                if (ex::replaces(tc, c, fs.res_type[warp][e], best_t[r][tid], kcur, type_k)) {
                    best_t[r][tid] = tc;
                    best_i[r][tid] = c;
                }
            }
        }
    }
    __syncwarp();
}

template <int OPI, int P>
struct FilterSurvivors {
    float2 l[OPI][P];
    float thr[OPI];
};

// The rare path of the filter loop: queue the (ray, object) pairs that passed the filter, run full 32-wide batches.
// obj0: scene index of the iteration's first object; n_valid: objects of the iteration that are not padding.
template <int RPT, int OPI, int CAP>
__device__ __noinline__ uint32_t filter_collect(const Scene& sc, const TaskQueue& q, FilterShared<CAP>& fs, float (*best_t)[kBlock],
                                            int32_t (*best_i)[kBlock], uint32_t warp_base, const FilterSurvivors<OPI, RPT / 2> sv,
                                            uint32_t obj0, uint32_t n_valid) {
    constexpr int P = RPT / 2;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t batches = 0;  // returned for the exact-test counter (counted per lane, as filter_flush works per lane)
#pragma unroll
    for (int o = 0; o < OPI; ++o) {
        if ((uint32_t)o >= n_valid) continue;  // padding of the last iteration
#pragma unroll
        for (int p = 0; p < P; ++p) {
            if (sv.l[o][p].x >= sv.thr[o]) {
                const uint32_t slot = atomicAdd(&fs.cand_n[warp], 1u);
                fs.cand_obj[warp][slot] = obj0 + o;
                fs.cand_who[warp][slot] = (uint16_t)((lane << 3) | (2 * p));
            }
            if (sv.l[o][p].y >= sv.thr[o]) {
                const uint32_t slot = atomicAdd(&fs.cand_n[warp], 1u);
                fs.cand_obj[warp][slot] = obj0 + o;
                fs.cand_who[warp][slot] = (uint16_t)((lane << 3) | (2 * p + 1));
            }
        }
        __syncwarp();
        uint32_t have = fs.cand_n[warp];
        if (have >= 32u) {  // full batches from the tail of the list (the list holds at most CAP entries)
            do {
                filter_flush<RPT, CAP>(sc, q, fs, best_t, best_i, warp_base, have - 32u, 32u);
                ++batches;
                have -= 32u;
            } while (have >= 32u);
            __syncwarp();
            if (lane == 0) fs.cand_n[warp] = have;
        }
        __syncwarp();
    }
    return batches;
}

#ifndef RT_FILTER_RPT  // the configuration wave_render launches (build-time knobs for the A/B sweeps)
#define RT_FILTER_RPT 4
#define RT_FILTER_OPI 4
#define RT_FILTER_MINB 2
#define RT_FILTER_TILE 1024  // two buffers of 32 KB per block
#endif
// RPT rays per thread (even), OPI objects per loop iteration (independent FMA chains for latency hiding), MINB resident
// blocks per SM, TILE objects per shared-memory tile.
template <int RPT, int OPI, int MINB, int TILE, bool COUNT>
__global__ void __launch_bounds__(kBlock, MINB) wave_trace_filter(const Scene sc, const TaskQueue q, const uint32_t* __restrict__ n_ptr,
                                                                  const uint32_t n_min, const uint32_t n_max, float* __restrict__ res_t,
                                                                  int32_t* __restrict__ res_id, const float4* __restrict__ packed, Counters* ctr) {
    static_assert(RPT % 2 == 0 && RPT <= 8, "rays come in pairs; the ray slot is packed into 3 bits");
    static_assert(TILE % OPI == 0, "tile is a whole number of iterations");
    constexpr int P = RPT / 2;
    constexpr int CAP = 32 + RPT * 32;  // 31 left over + every ray of the warp on one object
    const uint32_t n = *n_ptr;
    if (n == 0u || n < n_min || n >= n_max) return;  // the other instantiation handles this queue size
    extern __shared__ __align__(128) float4 tiles[];  // 2 buffers x TILE objects x 2 float4: {cx, cx, cy, cy} {cz, cz, e', e'}
    __shared__ __align__(8) uint64_t bar_full[2], bar_empty[2];
    __shared__ float best_t[RPT][kBlock];
    __shared__ int32_t best_i[RPT][kBlock];
    __shared__ FilterShared<CAP> fs;
    LocalCounters lc;
    const uint32_t M = sc.n_objs;
    const float hmin = sc.filter_h;
    const uint32_t per_block = kBlock * RPT;
    const uint32_t n_groups = (n + per_block - 1) / per_block;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const float inf = __int_as_float(0x7f800000);
    const uint32_t n_tiles = (M + TILE - 1) / TILE;
    constexpr uint32_t kTileBytes = TILE * 2 * sizeof(float4);
    // tile pipeline: `it` counts the tiles this block has consumed (over all its ray groups -- every group walks the same
    // tile sequence, so the copy of the next tile is always known); tile `it` lives in buffer it & 1, fill number it >> 1
    const uint32_t my_groups = blockIdx.x < n_groups ? (n_groups - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    const uint64_t total_tiles = (uint64_t)my_groups * n_tiles;
    uint64_t it = 0;
    if (tid == 0) {
        mbar_init(&bar_full[0], 1);
        mbar_init(&bar_full[1], 1);
        mbar_init(&bar_empty[0], kBlock / 32);
        mbar_init(&bar_empty[1], kBlock / 32);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0 && total_tiles > 0) {  // prologue: tile 0 into buffer 0
        mbar_expect_tx(&bar_full[0], kTileBytes);
        bulk_copy_g2s(tiles, packed, kTileBytes, &bar_full[0]);
    }

    for (uint32_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
        float2 UX[P], UY[P], UZ[P], NQ[P], PX[P], PY[P], PZ[P], OO[P];
        const uint32_t base = g * per_block + tid;  // ray r of this thread: base + r * kBlock (coalesced loads)
        const uint32_t warp_base = g * per_block + warp * 32u;
        if (lane == 0) fs.cand_n[warp] = 0u;
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const uint32_t i = base + r * kBlock;
            best_t[r][tid] = RT_MAX_FLOAT;
            best_i[r][tid] = -1;
            float ux = 0.f, uy = 0.f, uz = 0.f, nq = 0.f, px = 0.f, py = 0.f, pz = 0.f, oo = -inf;  // padding: never passes
            if (i < n) {
                Ray ry;
                ry.s = q.s[i];
                ry.d = q.d[i];
                const cv::RayK rk = cv::prepare(ry);
                ux = rk.ux; uy = rk.uy; uz = rk.uz;
                px = 2.f * hmin * rk.ox; py = 2.f * hmin * rk.oy; pz = 2.f * hmin * rk.oz;
                nq = -(rk.ox * rk.ux + rk.oy * rk.uy + rk.oz * rk.uz);
                const float o2 = rk.ox * rk.ox + rk.oy * rk.oy + rk.oz * rk.oz;
                oo = -(hmin * (o2 - 2.f * kExpandPad * o2));
                // a degenerate ray (zero or non-finite direction / origin) must reach the exact test on every object,
                // as a NaN did in the compare form  !(lhs > e):  l = +inf passes every e'
                const float chk = ux + uy + uz + nq + px + py + pz + oo;
                if (!(fabsf(chk) < inf)) {
                    ux = uy = uz = nq = px = py = pz = 0.f;
                    oo = inf;
                }
            }
            if (r & 1) {
                UX[r / 2].y = ux; UY[r / 2].y = uy; UZ[r / 2].y = uz; NQ[r / 2].y = nq;
                PX[r / 2].y = px; PY[r / 2].y = py; PZ[r / 2].y = pz; OO[r / 2].y = oo;
            } else {
                UX[r / 2].x = ux; UY[r / 2].x = uy; UZ[r / 2].x = uz; NQ[r / 2].x = nq;
                PX[r / 2].x = px; PY[r / 2].x = py; PZ[r / 2].x = pz; OO[r / 2].x = oo;
            }
        }
        for (uint32_t t0 = 0; t0 < M; t0 += TILE, ++it) {
            const uint32_t cnt = min((uint32_t)TILE, M - t0);
            const uint32_t cnt_pad = (cnt + OPI - 1) / OPI * OPI;
            const uint32_t buf = (uint32_t)(it & 1u);
            if (tid == 0 && it + 1 < total_tiles) {
                // next tile into the other buffer, once every warp has left that buffer's previous contents (fill (it+1)>>1 needs
                // the consumption of fill ((it+1)>>1) - 1 to be complete)
                const uint64_t nx = it + 1;
                const uint32_t nb = (uint32_t)(nx & 1u);
                if (nx >= 2) mbar_wait(&bar_empty[nb], (uint32_t)(((nx >> 1) - 1) & 1u));
                mbar_expect_tx(&bar_full[nb], kTileBytes);
                bulk_copy_g2s(tiles + (size_t)nb * TILE * 2, packed + 2 * (size_t)((uint32_t)(nx % n_tiles) * TILE), kTileBytes, &bar_full[nb]);
            }
            mbar_wait(&bar_full[buf], (uint32_t)((it >> 1) & 1u));  // the bytes of this tile have landed
            const float4* tile = tiles + (size_t)buf * TILE * 2;
            for (uint32_t j = 0; j < cnt_pad;) {
                float2 l[OPI][P];
                float thr[OPI];
                bool any;
                // The hot loop: OPI objects against the RPT rays of every lane, until somebody in the warp has a survivor
                // (rare: a few percent of the iterations) or the tile ends -- straight-line code with ONE branch, the
                // backward one; the survivors are dealt with outside.
                do {
                    bool mine = false;
#pragma unroll
                    for (int o = 0; o < OPI; ++o) {
                        const float4 A = tile[2 * (j + o)], B = tile[2 * (j + o) + 1];
                        const float2 ax = make_float2(A.x, A.y), ay = make_float2(A.z, A.w), az = make_float2(B.x, B.y);
                        thr[o] = B.z;
#pragma unroll
                        for (int p = 0; p < P; ++p) {
                            const float2 dt = __ffma2_rn(ax, UX[p], __ffma2_rn(ay, UY[p], __ffma2_rn(az, UZ[p], NQ[p])));
                            const float2 w = __ffma2_rn(ax, PX[p], __ffma2_rn(ay, PY[p], __ffma2_rn(az, PZ[p], OO[p])));
                            l[o][p] = __ffma2_rn(dt, dt, w);
                        }
                    }
#pragma unroll
                    for (int o = 0; o < OPI; ++o) {
                        float m = fmaxf(l[o][0].x, l[o][0].y);
#pragma unroll
                        for (int p = 1; p < P; ++p) m = fmaxf(fmaxf(m, l[o][p].x), l[o][p].y);
                        mine |= m >= thr[o];
                    }
                    any = __any_sync(kFull, mine);
                    j += OPI;
                } while (!any && j < cnt_pad);
                if (any) {  // the iteration that just ran (objects j - OPI ...) left survivors
                    FilterSurvivors<OPI, P> sv;
#pragma unroll
                    for (int o = 0; o < OPI; ++o) {
                        sv.thr[o] = thr[o];
#pragma unroll
                        for (int p = 0; p < P; ++p) sv.l[o][p] = l[o][p];
                    }
                    const uint32_t batches = filter_collect<RPT, OPI, CAP>(sc, q, fs, best_t, best_i, warp_base, sv, t0 + j - OPI, cnt - (j - OPI));
                    if (COUNT) lc.exact_tests += batches;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_empty[buf]);  // this warp is done with the buffer
        }
        {  // the remainder of the list
            __syncwarp();
            const uint32_t have = fs.cand_n[warp];
            if (have) filter_flush<RPT, CAP>(sc, q, fs, best_t, best_i, warp_base, 0u, have);
            if (COUNT && lane < have) lc.exact_tests++;
            __syncwarp();
        }
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const uint32_t i = base + r * kBlock;
            if (i < n) {
                if (COUNT) lc.pair_tests += M;
                const int32_t id = best_i[r][tid];
                res_t[i] = id >= 0 ? best_t[r][tid] : RT_MAX_FLOAT;
                res_id[i] = id;
            }
        }
    }
    if (COUNT) flush_local(lc, ctr);
}

// One configuration of the filter kernel: launches the big-queue and the small-queue instantiation.
template <int RPT, int OPI, int MINB, int TILE>
cudaError_t launch_filter(const Scene& sc, const TaskQueue& qin, const uint32_t* n_ptr, float* res_t, int32_t* res_id, const float4* packed,
                          Counters* ctr, bool count, int sm_count, cudaStream_t stream) {
    const unsigned fgrid = (unsigned)sm_count * MINB;
    const uint32_t split = fgrid * kBlock * RPT;
    const int bytes = 2 * TILE * 2 * (int)sizeof(float4);  // two buffers
    if (!packed) return cudaErrorInvalidValue;
    {  // 64 KB of dynamic shared memory need the opt-in, and function attributes are per DEVICE: one process may drive
       // several GPUs (CudaRaytracer::UseDevices), so the opt-in is made once for every device that launches the kernel
        static bool done[64] = {};
        int dev = 0;
        cudaError_t a = cudaGetDevice(&dev);
        if (a != cudaSuccess) return a;
        if (dev < 0 || dev >= 64 || !done[dev]) {
            a = cudaFuncSetAttribute(wave_trace_filter<RPT, OPI, MINB, TILE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (a == cudaSuccess) a = cudaFuncSetAttribute(wave_trace_filter<RPT, OPI, MINB, TILE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (a == cudaSuccess) a = cudaFuncSetAttribute(wave_trace_filter<2, 1, MINB, TILE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (a == cudaSuccess) a = cudaFuncSetAttribute(wave_trace_filter<2, 1, MINB, TILE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (a != cudaSuccess) return a;
            if (dev >= 0 && dev < 64) done[dev] = true;
        }
    }
    // two instantiations, each takes the queue sizes it is efficient for (the size lives on the device)
    if (count) {
        wave_trace_filter<RPT, OPI, MINB, TILE, true><<<fgrid, kBlock, bytes, stream>>>(sc, qin, n_ptr, split, 0xffffffffu, res_t, res_id, packed, ctr);
        wave_trace_filter<2, 1, MINB, TILE, true><<<fgrid, kBlock, bytes, stream>>>(sc, qin, n_ptr, 0u, split, res_t, res_id, packed, ctr);
    } else {
        wave_trace_filter<RPT, OPI, MINB, TILE, false><<<fgrid, kBlock, bytes, stream>>>(sc, qin, n_ptr, split, 0xffffffffu, res_t, res_id, packed, ctr);
        wave_trace_filter<2, 1, MINB, TILE, false><<<fgrid, kBlock, bytes, stream>>>(sc, qin, n_ptr, 0u, split, res_t, res_id, packed, ctr);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ advance
#ifndef RT_ADV_MINB
#define RT_ADV_MINB 4  // resident CTAs per SM the advance kernel is compiled for (64 registers; latency bound: 3 -> 4 is 24.9 -> 24.0 ms per
                       // frame in round 2, 5 the same as 4)
#endif
#ifndef RT_ADV_BLOCK
#define RT_ADV_BLOCK 256  // threads per block of wave_advance.  Every loop iteration ends in a block-wide scan behind two barriers (ncu:
                          // `stalled_barrier` 3.5-5 per issue beside `long_scoreboard` 9-14) and two atomics on the queue counters:
                          // smaller blocks wait less and hammer the counters more.  8K cfg5 frame, 64 / 128 / 256 / 512 / 1024
                          // threads: 24.25 / 23.45 / 23.17 / 23.31 / 23.65 ms
#endif
constexpr int kAdvBlock = RT_ADV_BLOCK;
#ifndef RT_ADV_PREFETCH
#define RT_ADV_PREFETCH 1  // 0 off, 1 into L1, 2 into L2.  8K cfg5 frame: 23.50 / 23.21 / 23.19 ms of kernels, Render() 26.4 / 26.0 / 26.1
#endif
#if RT_ADV_PREFETCH == 1
__device__ __forceinline__ void prefetch_state(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
#elif RT_ADV_PREFETCH == 2
__device__ __forceinline__ void prefetch_state(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#endif

template <bool COUNT>
__global__ void __launch_bounds__(kAdvBlock, (RT_ADV_MINB * kBlock / kAdvBlock) > 0 ? (RT_ADV_MINB * kBlock / kAdvBlock) : 1) wave_advance(const Scene sc, const WaveParams wp, const PathArrays pa, const TaskQueue qin,
                                                       const TaskQueue qout, const uint32_t* __restrict__ n_in, uint32_t* __restrict__ n_out,
                                                       const float* __restrict__ res_t, const int32_t* __restrict__ res_id,
                                                       const int first, float4* __restrict__ fb, int32_t* __restrict__ aov_id,
                                                       float* __restrict__ aov_t, Counters* ctr) {
    const uint32_t nb = n_in[0], na = n_in[1];  // class B tasks at [0, nb), class A tasks at qcap - 1 - j
    const uint32_t n = nb + na;
    if (n == 0u) return;
    LocalCounters lc;
    __shared__ uint32_t sh_append[2 * kAdvBlock / 32];
    const uint32_t back = wp.qcap - 1u;
    const uint64_t n_round = ((uint64_t)n + kAdvBlock - 1u) / kAdvBlock * kAdvBlock;  // whole blocks take part in the append
    for (uint64_t i = (uint64_t)blockIdx.x * kAdvBlock + threadIdx.x; i < n_round; i += (uint64_t)gridDim.x * kAdvBlock) {
        StepOut o;
        o.emit = o.write = o.anyhit = false;
        o.light = RT_TASK_NO_GRID;
        uint32_t p = 0;
        if (i < n) {
            const uint32_t at = i < nb ? (uint32_t)i : back - ((uint32_t)i - nb);
            const uint32_t pw = qin.path[at];
            p = pw & RT_TASK_PATH_MASK;
            Ray ray;
            if (pw & RT_TASK_ANYHIT) {  // a shadow ray came back: the state machine needs its result, not the ray (32 B less per event)
                ray.s = ray.d = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                ray.s = qin.s[at];
                ray.d = qin.d[at];
            }
            Hit res;
            res.t = res_t[at];
            res.id = res_id[at];
#if RT_ADV_PREFETCH
            // the state machine walks a chain of dependent gathers (ctrl word -> path state -> object records); what the path
            // word and the hit already tell is requested now, so that the chain finds it on its way in
            if (!first) {
#define RT_PF prefetch_state
                if (pw & RT_TASK_ANYHIT) {
                    RT_PF(pa.P + p);
                    RT_PF(pa.Nm + p);
                    RT_PF(pa.acc + p);
                    RT_PF(pa.refl + p);
                } else if (res.id < 0) {
                    RT_PF(pa.acc + p);
                    RT_PF(pa.refl + p);
                } else {
                    RT_PF(sc.minv + 4 * (size_t)res.id);
                    RT_PF(sc.minv + 4 * (size_t)res.id + 2);
                    RT_PF(sc.mv + 4 * (size_t)res.id);
                    RT_PF(sc.mv + 4 * (size_t)res.id + 2);
                    RT_PF(sc.mat + 3 * (size_t)res.id);
                }
#undef RT_PF
            }
#endif
            if (first && aov_id) {
                const uint32_t oi = pa.out[p];
                aov_id[oi] = res.id;
                aov_t[oi] = res.t;
            }
            wave::step<COUNT>(sc, wp, pa, p, first != 0, ray, res, o, lc);
        }
        // a shadow ray of a light with grids is a class A task -- unless it is not finite (a hit point at infinity): the BVH
        // kernel lets such a ray reach the exact test on every object, as the reference does
        bool to_a = false;
        if (o.emit && o.anyhit && o.light < RT_TASK_NO_GRID && ((wp.grid_lights >> o.light) & 1u)) {
            const float chk = o.ray.s.x + o.ray.s.y + o.ray.s.z + o.ray.d.x + o.ray.d.y + o.ray.d.z;
            to_a = fabsf(chk) < __int_as_float(0x7f800000);
        }
        const uint32_t slot = block_append2<kAdvBlock>(o.emit, to_a, n_out, n_out + 1, sh_append);
        if (o.emit) {
            const uint32_t at = to_a ? back - slot : slot;
            qout.s[at] = o.ray.s;
            qout.d[at] = o.ray.d;
            qout.path[at] = p | ((to_a ? o.light : RT_TASK_NO_GRID) << RT_TASK_LIGHT_SHIFT) | (o.anyhit ? RT_TASK_ANYHIT : 0u);
        }
        if (o.write && fb) fb[pa.out[p]] = o.px;  // one 128-bit store per finished pixel
    }
    if (COUNT) flush_local(lc, ctr);
}

// ------------------------------------------------------------------------------------------------ tail
// The late generations of a frame are thin: a few thousand rays each, every kernel of the chain bound by the latency of its
// longest ray, ~0.05-0.3 ms per generation whatever the GPU could do in that time -- the fixed cost that limits small frames
// (a shard of an 8-GPU job, a readback band).  From generation `tail_gen` on, ONE launch finishes every path that is still
// alive: a thread takes a task of that generation's queue and runs trace -> state machine -> trace ... in place until the
// path writes its pixel (the megakernel's per-ray searches: rt_device.cuh trace_bvh, rt_pgrid.cuh trace_grid; the same
// wave::step as wave_advance, so the same bits).  Warps fetch 32 tasks at a time.
template <bool COUNT>
__global__ void __launch_bounds__(128) wave_tail(const Scene sc, const GridSet lit, const Grid3 g3, const WaveParams wp, const PathArrays pa, const TaskQueue qin,
                                                  const uint32_t* __restrict__ n_in, uint32_t* __restrict__ fetch, float4* __restrict__ fb, Counters* ctr) {
    const uint32_t nb = n_in[0], na = n_in[1];
    const uint32_t n = nb + na;
    if (n == 0u) return;
    const unsigned lane = threadIdx.x & 31u;
    LocalCounters lc;
    const uint32_t back = wp.qcap - 1u;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(fetch, 32u);
        base = __shfl_sync(kFull, base, 0);
        if (base >= n) break;
        const uint32_t i = base + lane;
        if (i >= n) continue;
        const uint32_t at = i < nb ? i : back - (i - nb);
        Ray ray;
        ray.s = qin.s[at];
        ray.d = qin.d[at];
        uint32_t pw = qin.path[at];
        const uint32_t p = pw & RT_TASK_PATH_MASK;
        for (;;) {
            unsigned long long ne = 0, nn = 0, np = 0;
            Hit res;
            const uint32_t li = (pw >> RT_TASK_LIGHT_SHIFT) & 15u;
            if ((pw & RT_TASK_ANYHIT) && li != RT_TASK_NO_GRID)
                res = trace_grid<true>(sc, lit, 6u * li, ray, COUNT ? &np : nullptr, COUNT ? &ne : nullptr);
            else if (g3.valid && (pw & RT_TASK_ANYHIT))
                res = trace_grid3<true>(sc, g3, ray, 1.f, COUNT ? &np : nullptr, COUNT ? &ne : nullptr, COUNT ? &nn : nullptr);
            else if (g3.valid)
                res = trace_grid3<false>(sc, g3, ray, RT_MAX_FLOAT, COUNT ? &np : nullptr, COUNT ? &ne : nullptr, COUNT ? &nn : nullptr);
            else if (pw & RT_TASK_ANYHIT)
                res = trace_bvh<true>(sc, ray, 1.f, COUNT ? &ne : nullptr, COUNT ? &nn : nullptr);
            else
                res = trace_bvh<false>(sc, ray, RT_MAX_FLOAT, COUNT ? &ne : nullptr, COUNT ? &nn : nullptr);
            if (COUNT) {
                lc.exact_tests += ne;
                lc.node_visits += nn;
                lc.pair_tests += np;
            }
            if (res.id < 0) res.t = RT_MAX_FLOAT;
            StepOut o;
            wave::step<COUNT>(sc, wp, pa, p, false, ray, res, o, lc);
            if (o.write && fb) fb[pa.out[p]] = o.px;
            if (!o.emit) break;
            bool to_a = false;
            if (o.anyhit && o.light < RT_TASK_NO_GRID && ((wp.grid_lights >> o.light) & 1u)) {
                const float chk = o.ray.s.x + o.ray.s.y + o.ray.s.z + o.ray.d.x + o.ray.d.y + o.ray.d.z;
                to_a = fabsf(chk) < __int_as_float(0x7f800000);
            }
            ray = o.ray;
            pw = p | ((to_a ? o.light : RT_TASK_NO_GRID) << RT_TASK_LIGHT_SHIFT) | (o.anyhit ? RT_TASK_ANYHIT : 0u);
        }
    }
    if (COUNT) flush_local(lc, ctr);
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host driver
uint32_t wave_max_generations(uint32_t max_bounces, uint32_t n_lights, int mode) {
    if (mode == 0) return 1;
    if (mode == 1) return 1 + n_lights;
    const uint64_t g = 1ull + n_lights + (uint64_t)max_bounces * (1ull + n_lights);
    return (uint32_t)(g > 0x7fffffffull ? 0x7fffffffull : g);
}

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

uint32_t wave_filter_tile_pad(uint32_t n_objs) { return (n_objs + RT_FILTER_TILE - 1) / RT_FILTER_TILE * RT_FILTER_TILE; }

cudaError_t wave_pack_filter_tiles(const Scene& sc, float4* out, cudaStream_t stream) {
    const uint32_t n_pad = wave_filter_tile_pad(sc.n_objs);
    if (n_pad) filter_pack_tiles<<<(n_pad + 255) / 256, 256, 0, stream>>>(sc.cull, sc.n_objs, n_pad, sc.filter_h, out);
    return cudaGetLastError();
}

cudaError_t wave_render(const Scene& sc, const Frame& fr, const WaveBuffers& wb, int accel, const WaveGrids* grids, int sm_count,
                        uint32_t* launches, cudaStream_t stream, int tail_gen) {
    // warp-scheduler thresholds of wave_trace_bvh (tunable for experiments)
    static const int min_t = env_int("RT_WAVE_MIN_T", 12), drain_t = env_int("RT_WAVE_DRAIN", 4), refill_t = env_int("RT_WAVE_REFILL", 16);
    static const int g_min_t = env_int("RT_GRID_MIN_T", 12), g_drain_t = env_int("RT_GRID_DRAIN", 4), g_refill_t = env_int("RT_GRID_REFILL", 16);
    const uint32_t gens = wave_max_generations(fr.max_bounces, sc.n_lights, fr.mode);
    if (gens + 1 > wb.n_counts) return cudaErrorInvalidValue;
    if (wb.qcap == 0u || wb.qcap > RT_TASK_PATH_MASK) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(wb.counts, 0, sizeof(uint32_t) * 4 * (size_t)wb.n_counts, stream);
    if (e != cudaSuccess) return e;
    uint32_t* qcount = wb.counts;                    // qcount[2 g], qcount[2 g + 1]: class B / class A tasks of generation g
    uint32_t* fetch = wb.counts + 2 * wb.n_counts;   // fetch[2 g + class]: dynamic-fetch cursors
    uint32_t nl = 0;
    // perspective grids: the BVH search with the rays through a common apex taken out (accel 3)
    const bool use_grids = accel == 3 && grids != nullptr;
    const bool cam_grid = use_grids && !fr.rays && (grids->cam.valid_mask & 1u);
    const uint32_t light_grids = use_grids ? grids->lit.valid_mask : 0u;
    const int accel_b = accel == 3 ? 2 : accel;

    unsigned grid0;
    if (fr.rays) {
        grid0 = (unsigned)((fr.n_local + kBlock - 1) / kBlock);
    } else {
        grid0 = ((fr.width + kTileW - 1) / kTileW) * ((fr.local_rows + kTileH - 1) / kTileH);
    }
    if (grid0 == 0) {
        if (launches) *launches = 0;
        return cudaSuccess;
    }
    static const int gen_strip = env_int("RT_GEN_STRIP", 64);
    wave_gen<<<grid0, kBlock, 0, stream>>>(fr, wb.q[0], wb.pa, qcount + 0, (uint32_t)gen_strip, cam_grid ? 1u : 0u, wb.qcap);
    ++nl;
    WaveParams wp;
    wp.max_bounces = fr.max_bounces;
    wp.mode = fr.mode;
    wp.grid_lights = light_grids;
    wp.qcap = wb.qcap;
    const unsigned pgrid = (unsigned)sm_count * RT_BVH_MINB;  // persistent trace: RT_BVH_MINB CTAs of 8 warps per SM
    const unsigned ggrid = (unsigned)sm_count * RT_GRID_MINB;
    const unsigned dgrid = (unsigned)sm_count * RT_DDA_MINB;
    static const int d_min_t = env_int("RT_DDA_MIN_T", 12), d_drain_t = env_int("RT_DDA_DRAIN", 4), d_refill_t = env_int("RT_DDA_REFILL", 16);
    const unsigned sgrid = (unsigned)sm_count * 8u;    // grid-stride kernels
    static const int adv_mult = env_int("RT_ADV_GRID", RT_ADV_MINB);  // = resident CTAs per SM; 3..32 measured within 1 %
    const unsigned agrid = std::max(1u, (unsigned)sm_count * (unsigned)adv_mult * kBlock / kAdvBlock);  // wave_advance
    const bool count = fr.counters != nullptr;
    for (uint32_t g = 0; g < gens; ++g) {
        const TaskQueue& qin = wb.q[g & 1u];
        const TaskQueue& qout = wb.q[(g + 1u) & 1u];
        uint32_t* nq = qcount + 2 * g;
        if (tail_gen > 0 && g >= (uint32_t)tail_gen && accel >= 2 && !fr.hit_id) {  // one launch finishes what is left
            const unsigned tgrid = (unsigned)sm_count * 8u;
            const GridSet lit = use_grids ? grids->lit : GridSet{};
            const Grid3 g3 = use_grids ? grids->g3 : Grid3{};
            if (count)
                wave_tail<true><<<tgrid, 128, 0, stream>>>(sc, lit, g3, wp, wb.pa, qin, nq, fetch + 2 * g, fr.out, fr.counters);
            else
                wave_tail<false><<<tgrid, 128, 0, stream>>>(sc, lit, g3, wp, wb.pa, qin, nq, fetch + 2 * g, fr.out, fr.counters);
            ++nl;
            break;
        }
        // class A: generation 0 when the camera has a grid, later generations when a light has
        if ((g == 0 && cam_grid) || (g > 0 && light_grids)) {
            if (count)
                wave_trace_grid<true><<<ggrid, kBlock, 0, stream>>>(sc, grids->cam, grids->lit, qin, nq + 1, fetch + 2 * g + 1, wb.res_t, wb.res_id, wb.qcap, g_min_t, g_drain_t, g_refill_t, fr.counters);
            else
                wave_trace_grid<false><<<ggrid, kBlock, 0, stream>>>(sc, grids->cam, grids->lit, qin, nq + 1, fetch + 2 * g + 1, wb.res_t, wb.res_id, wb.qcap, g_min_t, g_drain_t, g_refill_t, fr.counters);
            ++nl;
        }
        if (g == 0 && cam_grid) {
            // every primary ray is a class A task: no class B launch in this generation
        } else if (accel_b == 2 && use_grids && grids->g3.valid) {
            if (count)
                wave_trace_dda<true><<<dgrid, kBlock, 0, stream>>>(sc, grids->g3, qin, nq, fetch + 2 * g, wb.res_t, wb.res_id, d_min_t, d_drain_t, d_refill_t, fr.counters);
            else
                wave_trace_dda<false><<<dgrid, kBlock, 0, stream>>>(sc, grids->g3, qin, nq, fetch + 2 * g, wb.res_t, wb.res_id, d_min_t, d_drain_t, d_refill_t, fr.counters);
            ++nl;
        } else if (accel_b == 2) {
            if (count)
                wave_trace_bvh<true><<<pgrid, kBlock, 0, stream>>>(sc, qin, nq, fetch + 2 * g, wb.res_t, wb.res_id, min_t, drain_t, refill_t, fr.counters);
            else
                wave_trace_bvh<false><<<pgrid, kBlock, 0, stream>>>(sc, qin, nq, fetch + 2 * g, wb.res_t, wb.res_id, min_t, drain_t, refill_t, fr.counters);
            ++nl;
        } else if (accel_b == 1) {
            // 4 rays per thread x 4 objects per iteration, 2 blocks per SM, 2048-object tiles: the fastest of the
            // configurations swept on the B200 (tools/build_filter_variants.sh; history in DESIGN.md)
            const cudaError_t fe = launch_filter<RT_FILTER_RPT, RT_FILTER_OPI, RT_FILTER_MINB, RT_FILTER_TILE>(sc, qin, nq, wb.res_t, wb.res_id,
                                                                                                          grids ? grids->filter_tiles : nullptr, fr.counters, count, sm_count, stream);
            if (fe != cudaSuccess) return fe;
            nl += 2;
        } else {
            if (count)
                wave_trace_brute<true><<<sgrid, kBlock, 0, stream>>>(sc, qin, nq, wb.res_t, wb.res_id, fr.counters);
            else
                wave_trace_brute<false><<<sgrid, kBlock, 0, stream>>>(sc, qin, nq, wb.res_t, wb.res_id, fr.counters);
            ++nl;
        }
        if (count)
            wave_advance<true><<<agrid, kAdvBlock, 0, stream>>>(sc, wp, wb.pa, qin, qout, nq, nq + 2, wb.res_t, wb.res_id,
                                                             g == 0 ? 1 : 0, fr.out, fr.hit_id, fr.hit_t, fr.counters);
        else
            wave_advance<false><<<agrid, kAdvBlock, 0, stream>>>(sc, wp, wb.pa, qin, qout, nq, nq + 2, wb.res_t, wb.res_id,
                                                              g == 0 ? 1 : 0, fr.out, fr.hit_id, fr.hit_t, fr.counters);
        ++nl;
    }
    if (launches) *launches = nl;
    return cudaGetLastError();
}

}  // namespace rt
