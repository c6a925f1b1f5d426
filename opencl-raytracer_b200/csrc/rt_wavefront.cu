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

// Append one entry per participating lane to a queue: one atomic per warp, order within the warp kept.
__device__ __forceinline__ uint32_t warp_append(bool emit, uint32_t* counter) {
    const unsigned m = __ballot_sync(kFull, emit);
    if (m == 0u) return 0u;
    const unsigned lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(kFull, base, __ffs(m) - 1);
    return base + (uint32_t)__popc(m & lanemask_lt());
}

// ------------------------------------------------------------------------------------------------ generation 0
__global__ void __launch_bounds__(kBlock) wave_gen(const Frame fr, const TaskQueue q, const PathArrays pa, uint32_t* __restrict__ count) {
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
        const uint32_t tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
        const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
        const uint32_t lx = tx * kTileW + (warp & 3u) * 8u + (lane & 7u);
        const uint32_t ly = ty * kTileH + (warp >> 2) * 4u + (lane >> 3);
        active = lx < fr.width && ly < fr.local_rows;
        if (active) {
            const uint32_t blk = ly / fr.block_rows, rib = ly % fr.block_rows;
            const uint32_t row = (blk * fr.n_shards + fr.shard) * fr.block_rows + rib;
            active = row < fr.height;
            if (active) {
                oi = (uint64_t)row * fr.width + lx;
                ray = camera_ray(lx, row, (float)fr.width / 2.0f, (float)fr.height / 2.0f, (float)fr.height, fr.zdir);
            }
        }
    }
    const uint32_t slot = warp_append(active, count);
    if (active) {
        q.s[slot] = ray.s;
        q.d[slot] = ray.d;
        q.path[slot] = slot;  // path id == slot of the primary task
        pa.out[slot] = (uint32_t)oi;
    }
}

// ------------------------------------------------------------------------------------------------ trace: BVH
constexpr int kPendCap = 8;

struct LaneRay {  // what a lane keeps in registers for its ray
    float ox, oy, oz, ux, uy, uz, idx, idy, idz, onorm;
};

__device__ __forceinline__ bool node_box(const LaneRay& k, float lx, float ly, float lz, float hx, float hy, float hz, float infl,
                                         float tlimit, float& tnear) {
    const float ax = (lx - k.ox) * k.idx, bx = (hx - k.ox) * k.idx;
    const float ay = (ly - k.oy) * k.idy, by = (hy - k.oy) * k.idy;
    const float az = (lz - k.oz) * k.idz, bz = (hz - k.oz) * k.idz;
    const float nx = fmaf(-infl, fabsf(k.idx), fminf(ax, bx)), fx = fmaf(infl, fabsf(k.idx), fmaxf(ax, bx));
    const float ny = fmaf(-infl, fabsf(k.idy), fminf(ay, by)), fy = fmaf(infl, fabsf(k.idy), fmaxf(ay, by));
    const float nz = fmaf(-infl, fabsf(k.idz), fminf(az, bz)), fz = fmaf(infl, fabsf(k.idz), fmaxf(az, bz));
    const float tn = fmaxf(fmaxf(nx, ny), nz);
    const float tf = fminf(fminf(fx, fy), fz);
    tnear = tn;
    const float slack = 1e-5f * fmaxf(fabsf(tn), fabsf(tf));
    return !(tn - slack > fminf(tf, tlimit)) && !(tf + slack < 0.f);
}

template <bool COUNT>
__global__ void __launch_bounds__(kBlock, 2) wave_trace_bvh(const Scene sc, const TaskQueue q, const uint32_t* __restrict__ n_ptr,
                                                            uint32_t* __restrict__ fetch, float* __restrict__ res_t,
                                                            int32_t* __restrict__ res_id, const uint32_t leaf_max, Counters* ctr) {
    const uint32_t n = *n_ptr;
    if (n == 0u) return;
    __shared__ uint32_t pend[kPendCap][kBlock];
    const unsigned lane = threadIdx.x & 31u, tid = threadIdx.x;
    LocalCounters lc;

    bool have = false, tdone = false;
    uint32_t task = 0, path_bits = 0;
    LaneRay k;
    k.ox = k.oy = k.oz = k.ux = k.uy = k.uz = k.idx = k.idy = k.idz = k.onorm = 0.f;
    float tcap = RT_MAX_FLOAT;
    Hit best;
    best.t = RT_MAX_FLOAT;
    best.id = -1;
    uint32_t btype = 0;
    int32_t cur = 0;
    int sp = 0, npend = 0;
    int32_t stack[RT_BVH_STACK];
    bool exhausted = false;
    const float4* lcull = reinterpret_cast<const float4*>(sc.leaf_cull);
    const int must_level = kPendCap - (int)min(leaf_max, (uint32_t)kPendCap);

    for (;;) {
        // A. idle lanes fetch the next rays of the queue
        const unsigned idle = __ballot_sync(kFull, !have);
        if (idle && !exhausted) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(fetch, (uint32_t)__popc(idle));
            base = __shfl_sync(kFull, base, 0);
            if (!have) {
                const uint32_t i = base + (uint32_t)__popc(idle & lanemask_lt());
                if (i < n) {
                    task = i;
                    const float4 s = q.s[i], d = q.d[i];
                    path_bits = q.path[i];
                    Ray r;
                    r.s = s;
                    r.d = d;
                    const cv::RayK rk = cv::prepare(r);
                    k.ox = rk.ox; k.oy = rk.oy; k.oz = rk.oz;
                    k.ux = rk.ux; k.uy = rk.uy; k.uz = rk.uz;
                    k.idx = rk.idx; k.idy = rk.idy; k.idz = rk.idz;
                    k.onorm = rk.onorm;
                    tcap = (path_bits & RT_TASK_ANYHIT) ? 1.f : RT_MAX_FLOAT;
                    best.t = RT_MAX_FLOAT;
                    best.id = -1;
                    btype = 0;
                    cur = sc.root;
                    sp = 0;
                    npend = 0;
                    tdone = sc.n_objs == 0u;
                    have = true;
                }
            }
            if (base + (uint32_t)__popc(idle) >= n) exhausted = true;
        }
        if (!__any_sync(kFull, have)) break;

        // B. inner nodes until this lane stands on a leaf (or runs out of nodes)
        while (have && !tdone && cur >= 0) {
            const float4* np = reinterpret_cast<const float4*>(sc.nodes + cur);
            const float4 q0 = __ldg(np + 0), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
            const float2 ff = __ldg(reinterpret_cast<const float2*>(sc.nodef + cur));
            if (COUNT) lc.node_visits++;
            const float tlimit = fminf(best.t, tcap);
            float tn0, tn1;
            const bool h0 = node_box(k, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, fmaf(ff.x, k.onorm, q3.z), tlimit, tn0);
            const bool h1 = node_box(k, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, fmaf(ff.y, k.onorm, q3.w), tlimit, tn1);
            const int32_t c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            if (h0 && h1) {
                const bool first0 = tn0 <= tn1;
                cur = first0 ? c0 : c1;
                stack[sp++] = first0 ? c1 : c0;
            } else if (h0) {
                cur = c0;
            } else if (h1) {
                cur = c1;
            } else if (sp > 0) {
                cur = stack[--sp];
            } else {
                tdone = true;
            }
        }

        // C. leaf: bounding-sphere filter, survivors are queued for the exact test
        if (have && !tdone && cur < 0) {
            const uint32_t code = (uint32_t)(~cur);
            const uint32_t first = code >> 4, cnt = (code & 15u) + 1u;
            for (uint32_t i = 0; i < cnt; ++i) {
                const float4 a = __ldg(lcull + 2 * (size_t)(first + i));
                const float h = __ldg(&sc.leaf_cull[first + i].h);
                const float x = a.x - k.ox, y = a.y - k.oy, z = a.z - k.oz;
                const float oc2 = fmaf(z, z, fmaf(y, y, x * x));
                const float dt = fmaf(z, k.uz, fmaf(y, k.uy, x * k.ux));
                if (!(fmaf(-dt, dt, oc2 * h) > a.w)) {
                    if (npend == kPendCap) {  // overflow (only with leaves larger than the buffer): test in place
                        Ray r;
                        r.s = q.s[task];
                        r.d = q.d[task];
                        exact_pair(sc, r, __ldg(sc.leaf_obj + first + i), __ldg(&sc.leaf_cull[first + i].type), best, btype);
                        if (COUNT) lc.exact_tests++;
                    } else {
                        pend[npend++][tid] = first + i;
                    }
                }
            }
            if (sp > 0)
                cur = stack[--sp];
            else
                tdone = true;
        }

        // D. exact tests, run when enough lanes have one queued (or a lane cannot go on without)
        for (;;) {
            const bool has = have && npend > 0;
            const unsigned hasp = __ballot_sync(kFull, has);
            if (hasp == 0u) break;
            const unsigned must = __ballot_sync(kFull, has && (tdone || npend > must_level));
            if (must == 0u && __popc(hasp) < 16) break;
            if (has) {
                const uint32_t slot = pend[--npend][tid];
                Ray r;
                r.s = q.s[task];
                r.d = q.d[task];
                exact_pair(sc, r, __ldg(sc.leaf_obj + slot), __ldg(&sc.leaf_cull[slot].type), best, btype);
                if (COUNT) lc.exact_tests++;
                if ((path_bits & RT_TASK_ANYHIT) && best.t < tcap) {  // shadow ray decided (SAR:229)
                    tdone = true;
                    npend = 0;
                }
            }
        }

        // E. retire finished rays
        if (have && tdone && npend == 0) {
            res_t[task] = best.id >= 0 ? best.t : RT_MAX_FLOAT;
            res_id[task] = best.id;
            have = false;
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
// (every lane of a warp reads the same record: one broadcast wavefront per load), survivors run the exact test.
constexpr int kFilterTile = 1024;  // objects per tile: 1024 x 32 B = 32 KB

template <int RPT, bool COUNT>
__global__ void __launch_bounds__(kBlock) wave_trace_filter(const Scene sc, const TaskQueue q, const uint32_t* __restrict__ n_ptr,
                                                            float* __restrict__ res_t, int32_t* __restrict__ res_id, Counters* ctr) {
    const uint32_t n = *n_ptr;
    if (n == 0u) return;
    __shared__ float4 tileA[kFilterTile];  // cx, cy, cz, r2
    __shared__ float tileH[kFilterTile];   // h
    LocalCounters lc;
    const uint32_t M = sc.n_objs;
    const float4* cull4 = reinterpret_cast<const float4*>(sc.cull);
    const uint32_t per_block = kBlock * RPT;
    const uint32_t n_groups = (n + per_block - 1) / per_block;

    for (uint32_t g = blockIdx.x; g < n_groups; g += gridDim.x) {
        float ox[RPT], oy[RPT], oz[RPT], ux[RPT], uy[RPT], uz[RPT];
        Hit best[RPT];
        uint32_t btype[RPT];
        uint32_t idx[RPT];
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            // consecutive lanes take consecutive rays: coalesced loads, and neighbours stay neighbours
            idx[r] = g * per_block + r * kBlock + threadIdx.x;
            best[r].t = RT_MAX_FLOAT;
            best[r].id = -1;
            btype[r] = 0;
            if (idx[r] < n) {
                Ray ry;
                ry.s = q.s[idx[r]];
                ry.d = q.d[idx[r]];
                const cv::RayK rk = cv::prepare(ry);
                ox[r] = rk.ox; oy[r] = rk.oy; oz[r] = rk.oz;
                ux[r] = rk.ux; uy[r] = rk.uy; uz[r] = rk.uz;
            } else {  // padding lane: a ray that passes every filter test as "miss" cheaply
                ox[r] = oy[r] = oz[r] = 0.f;
                ux[r] = uy[r] = uz[r] = 0.f;
            }
        }
        for (uint32_t t0 = 0; t0 < M; t0 += kFilterTile) {
            const uint32_t cnt = min((uint32_t)kFilterTile, M - t0);
            __syncthreads();
            for (uint32_t j = threadIdx.x; j < cnt; j += kBlock) {
                tileA[j] = __ldg(cull4 + 2 * (size_t)(t0 + j));
                tileH[j] = __ldg(&sc.cull[t0 + j].h);
            }
            __syncthreads();
#pragma unroll 4
            for (uint32_t j = 0; j < cnt; ++j) {
                const float4 a = tileA[j];
                const float h = tileH[j];
#pragma unroll
                for (int r = 0; r < RPT; ++r) {
                    const float x = a.x - ox[r], y = a.y - oy[r], z = a.z - oz[r];
                    const float oc2 = fmaf(z, z, fmaf(y, y, x * x));
                    const float dt = fmaf(z, uz[r], fmaf(y, uy[r], x * ux[r]));
                    if (!(fmaf(-dt, dt, oc2 * h) > a.w)) {
                        if (idx[r] < n) {
                            Ray ry;
                            ry.s = q.s[idx[r]];
                            ry.d = q.d[idx[r]];
                            exact_pair(sc, ry, t0 + j, __ldg(&sc.cull[t0 + j].type), best[r], btype[r]);
                            if (COUNT) lc.exact_tests++;
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < RPT; ++r)
            if (idx[r] < n) {
                if (COUNT) lc.pair_tests += M;
                res_t[idx[r]] = best[r].id >= 0 ? best[r].t : RT_MAX_FLOAT;
                res_id[idx[r]] = best[r].id;
            }
    }
    if (COUNT) flush_local(lc, ctr);
}

// ------------------------------------------------------------------------------------------------ advance
template <bool COUNT>
__global__ void __launch_bounds__(kBlock) wave_advance(const Scene sc, const WaveParams wp, const PathArrays pa, const TaskQueue qin,
                                                       const TaskQueue qout, const uint32_t* __restrict__ n_in, uint32_t* __restrict__ n_out,
                                                       const float* __restrict__ res_t, const int32_t* __restrict__ res_id,
                                                       const int first, float4* __restrict__ fb, int32_t* __restrict__ aov_id,
                                                       float* __restrict__ aov_t, Counters* ctr) {
    const uint32_t n = *n_in;
    if (n == 0u) return;
    LocalCounters lc;
    const uint32_t n_round = (n + 31u) & ~31u;  // whole warps take part in the ballots
    for (uint32_t i = blockIdx.x * kBlock + threadIdx.x; i < n_round; i += gridDim.x * kBlock) {
        StepOut o;
        o.emit = o.write = o.anyhit = false;
        uint32_t p = 0;
        if (i < n) {
            Ray ray;
            ray.s = qin.s[i];
            ray.d = qin.d[i];
            p = qin.path[i] & ~RT_TASK_ANYHIT;
            Hit res;
            res.t = res_t[i];
            res.id = res_id[i];
            if (first && aov_id) {
                const uint32_t oi = pa.out[p];
                aov_id[oi] = res.id;
                aov_t[oi] = res.t;
            }
            wave::step<COUNT>(sc, wp, pa, p, first != 0, ray, res, o, lc);
        }
        const uint32_t slot = warp_append(o.emit, n_out);
        if (o.emit) {
            qout.s[slot] = o.ray.s;
            qout.d[slot] = o.ray.d;
            qout.path[slot] = p | (o.anyhit ? RT_TASK_ANYHIT : 0u);
        }
        if (o.write) fb[pa.out[p]] = o.px;  // one 128-bit store per finished pixel
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

cudaError_t wave_render(const Scene& sc, const Frame& fr, const WaveBuffers& wb, int accel, uint32_t leaf_max, int sm_count,
                        uint32_t* launches, cudaStream_t stream) {
    const uint32_t gens = wave_max_generations(fr.max_bounces, sc.n_lights, fr.mode);
    if (gens + 1 > wb.n_counts) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(wb.counts, 0, sizeof(uint32_t) * 2 * (size_t)wb.n_counts, stream);
    if (e != cudaSuccess) return e;
    uint32_t* qcount = wb.counts;                // qcount[g]: size of the queue of generation g
    uint32_t* fetch = wb.counts + wb.n_counts;   // fetch[g]: dynamic-fetch cursor of generation g
    uint32_t nl = 0;

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
    wave_gen<<<grid0, kBlock, 0, stream>>>(fr, wb.q[0], wb.pa, qcount + 0);
    ++nl;
    WaveParams wp;
    wp.max_bounces = fr.max_bounces;
    wp.mode = fr.mode;
    const unsigned pgrid = (unsigned)sm_count * 2u;    // persistent trace: 2 CTAs of 8 warps per SM
    const unsigned sgrid = (unsigned)sm_count * 8u;    // grid-stride kernels
    const bool count = fr.counters != nullptr;
    for (uint32_t g = 0; g < gens; ++g) {
        const TaskQueue& qin = wb.q[g & 1u];
        const TaskQueue& qout = wb.q[(g + 1u) & 1u];
        if (accel == 2) {
            if (count)
                wave_trace_bvh<true><<<pgrid, kBlock, 0, stream>>>(sc, qin, qcount + g, fetch + g, wb.res_t, wb.res_id, leaf_max, fr.counters);
            else
                wave_trace_bvh<false><<<pgrid, kBlock, 0, stream>>>(sc, qin, qcount + g, fetch + g, wb.res_t, wb.res_id, leaf_max, fr.counters);
        } else if (accel == 1) {
            if (count)
                wave_trace_filter<4, true><<<sgrid / 4, kBlock, 0, stream>>>(sc, qin, qcount + g, wb.res_t, wb.res_id, fr.counters);
            else
                wave_trace_filter<4, false><<<sgrid / 4, kBlock, 0, stream>>>(sc, qin, qcount + g, wb.res_t, wb.res_id, fr.counters);
        } else {
            if (count)
                wave_trace_brute<true><<<sgrid, kBlock, 0, stream>>>(sc, qin, qcount + g, wb.res_t, wb.res_id, fr.counters);
            else
                wave_trace_brute<false><<<sgrid, kBlock, 0, stream>>>(sc, qin, qcount + g, wb.res_t, wb.res_id, fr.counters);
        }
        if (count)
            wave_advance<true><<<sgrid, kBlock, 0, stream>>>(sc, wp, wb.pa, qin, qout, qcount + g, qcount + g + 1, wb.res_t, wb.res_id,
                                                             g == 0 ? 1 : 0, fr.out, fr.hit_id, fr.hit_t, fr.counters);
        else
            wave_advance<false><<<sgrid, kBlock, 0, stream>>>(sc, wp, wb.pa, qin, qout, qcount + g, qcount + g + 1, wb.res_t, wb.res_id,
                                                              g == 0 ? 1 : 0, fr.out, fr.hit_id, fr.hit_t, fr.counters);
        nl += 2;
    }
    if (launches) *launches = nl;
    return cudaGetLastError();
}

}  // namespace rt
