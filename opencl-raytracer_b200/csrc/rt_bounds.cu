// rt_bounds.cu -- the per-object part of rt_upload_scene on the device (scenes that change every frame).
//
// rt_scene.cpp does, per object and in double precision on host threads: AoS -> SoA repack (the cl_* marshalling of
// OpenCLRaytracer.cpp:16-24,108-146), the inverse of the linear part of mvInverse, its extreme singular values (cyclic
// Jacobi), the view-space bounds, the noise coefficients and the padded filter record -- 12-45 ms for 10^5 objects.  With
// RT_OPT_BVH_BUILDER = 1 the raw rt_object array is uploaded instead and the same arithmetic runs here, one thread per
// object, followed by the device tree builder (rt_lbvh.cu) and the device grid builder (rt_pgrid.cu): a re-upload is a
// few milliseconds and touches no host loop over the objects.  Same formulas, same constants, same outward roundings as
// rt_scene.cpp (the tests render the same frame from both); unknown object types (> 2) become leaves that are never hit.
#include "rt_bounds.cuh"

#include <math.h>

namespace rt {
namespace {

constexpr int kB = 128;
constexpr double U = 5.9604644775390625e-08;  // 2^-24
constexpr double K_FUZZ = 32.0, C_LIN = 24.0, LIN_EPS = 1.0 / 64.0;  // rt_scene.cpp

struct Geo {
    double c[3], ext[3], R, smin, smax;
};

__device__ bool inverse3(const double a[3][3], double out[3][3]) {
    const double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
    const double c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
    const double c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    const double det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
    if (!(fabs(det) > 0.0) || !isfinite(det)) return false;
    const double id = 1.0 / det;
    out[0][0] = c00 * id;
    out[1][0] = c01 * id;
    out[2][0] = c02 * id;
    out[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) * id;
    out[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) * id;
    out[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) * id;
    out[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) * id;
    out[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) * id;
    out[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) * id;
    return true;
}

// extreme singular values: sqrt of the eigenvalues of M^T M, cyclic Jacobi (rt_scene.cpp singular_range)
__device__ void singular_range(const double m[3][3], double& smin, double& smax) {
    double s[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            s[i][j] = 0;
            for (int k = 0; k < 3; ++k) s[i][j] += m[k][i] * m[k][j];
        }
    for (int sweep = 0; sweep < 32; ++sweep) {
        const double off = fabs(s[0][1]) + fabs(s[0][2]) + fabs(s[1][2]);
        if (off < 1e-300 || off < 1e-18 * (fabs(s[0][0]) + fabs(s[1][1]) + fabs(s[2][2]))) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (s[p][q] == 0.0) continue;
                const double theta = (s[q][q] - s[p][p]) / (2.0 * s[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), sn = t * c;
                for (int k = 0; k < 3; ++k) {
                    const double skp = s[k][p], skq = s[k][q];
                    s[k][p] = c * skp - sn * skq;
                    s[k][q] = sn * skp + c * skq;
                }
                for (int k = 0; k < 3; ++k) {
                    const double spk = s[p][k], sqk = s[q][k];
                    s[p][k] = c * spk - sn * sqk;
                    s[q][k] = sn * spk + c * sqk;
                }
            }
    }
    const double e0 = fmax(s[0][0], 0.0), e1 = fmax(s[1][1], 0.0), e2 = fmax(s[2][2], 0.0);
    smax = sqrt(fmax(e0, fmax(e1, e2)));
    smin = sqrt(fmin(e0, fmin(e1, e2)));
}

__device__ float up(double v) {  // round to float, then one ulp outward (rt_scene.cpp)
    float f = (float)v;
    if ((double)f < v) f = nextafterf(f, INFINITY);
    return nextafterf(f, INFINITY);
}
__device__ float down(double v) {
    float f = (float)v;
    if ((double)f > v) f = nextafterf(f, -INFINITY);
    return nextafterf(f, -INFINITY);
}

__device__ void fail_at(BoundsGlobals* g, uint32_t i, int code) {
    const unsigned long long key = ((unsigned long long)i << 8) | (unsigned)code;
    atomicMin(&g->first_bad, key);
}

__device__ void atomic_max_pos(double* addr, double v) {  // v >= 0: the bit patterns order like the values
    atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}
// ordered-integer image of a float: monotone for all finite values
__device__ __forceinline__ int ordered_key(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : (int)(0x80000000u - (unsigned)i);
}
__device__ void atomic_minmax_float(float* lo, float* hi, float v) {
    atomicMin(reinterpret_cast<int*>(lo), ordered_key(v));
    atomicMax(reinterpret_cast<int*>(hi), ordered_key(v));
}

// ---- pass 1: repack + geometry (rt_scene.cpp pass 1)
__global__ void __launch_bounds__(kB) bounds_pass1(const rt_object* __restrict__ objs, const uint32_t n, float4* __restrict__ minv,
                                                   float4* __restrict__ mv, float4* __restrict__ mat, Geo* __restrict__ geo, BoundsGlobals* g) {
    const uint32_t i = blockIdx.x * kB + threadIdx.x;
    if (i >= n) return;
    const rt_object& o = objs[i];
    float mi[16], mm[16];
    bool finite = true;
    for (int k = 0; k < 16; ++k) {
        mi[k] = o.mvInverse[k];
        mm[k] = o.mv[k];
        finite = finite && isfinite(mi[k]) && isfinite(mm[k]);
    }
    for (int c = 0; c < 4; ++c) {
        minv[4 * (size_t)i + c] = make_float4(mi[4 * c], mi[4 * c + 1], mi[4 * c + 2], mi[4 * c + 3]);
        mv[4 * (size_t)i + c] = make_float4(mm[4 * c], mm[4 * c + 1], mm[4 * c + 2], mm[4 * c + 3]);
    }
    mat[3 * (size_t)i + 0] = make_float4(o.mat.ambient[0], o.mat.ambient[1], o.mat.ambient[2], o.mat.absorption);
    mat[3 * (size_t)i + 1] = make_float4(o.mat.diffuse[0], o.mat.diffuse[1], o.mat.diffuse[2], o.mat.shininess);
    mat[3 * (size_t)i + 2] = make_float4(o.mat.specular[0], o.mat.specular[1], o.mat.specular[2], 0.f);
    Geo ge;
    for (int k = 0; k < 3; ++k) ge.c[k] = ge.ext[k] = 0;
    ge.R = 0;
    ge.smin = ge.smax = 1;
    if (!finite) {
        fail_at(g, i, RT_BOUNDS_NONFINITE);
        geo[i] = ge;
        return;
    }
    const double brow = fabs((double)mi[3]) + fabs((double)mi[7]) + fabs((double)mi[11]) + fabs((double)mm[3]) + fabs((double)mm[7]) + fabs((double)mm[11]);
    const double wd = fmax(fabs((double)mi[15] - 1.0), fabs((double)mm[15] - 1.0));
    if (brow > 0.0 || wd > 1e-4) {
        fail_at(g, i, RT_BOUNDS_PROJECTIVE);
        geo[i] = ge;
        return;
    }
    atomic_max_pos(&g->wdev_max, wd);
    const uint32_t type = o.type;
    if (type <= 2u) {
        double A[3][3], F[3][3];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) A[r][c] = mi[4 * c + r];
        if (!inverse3(A, F)) {
            fail_at(g, i, RT_BOUNDS_SINGULAR);
            geo[i] = ge;
            return;
        }
        for (int r = 0; r < 3; ++r) ge.c[r] = -(F[r][0] * mi[12] + F[r][1] * mi[13] + F[r][2] * mi[14]);
        singular_range(F, ge.smin, ge.smax);
        if (!(ge.smin > 0) || !isfinite(ge.smax)) {
            fail_at(g, i, RT_BOUNDS_DEGENERATE);
            geo[i] = ge;
            return;
        }
        if (type == 0u) {
            for (int r = 0; r < 3; ++r) ge.ext[r] = sqrt(F[r][0] * F[r][0] + F[r][1] * F[r][1] + F[r][2] * F[r][2]);
            ge.R = ge.smax;
        } else if (type == 2u) {
            const double vx[3] = {-0.5, 0.5, -0.5}, vy[3] = {-0.5, -0.5, 0.5};
            double rmax = 0;
            for (int v = 0; v < 3; ++v) {
                double v2 = 0;
                for (int r = 0; r < 3; ++r) {
                    const double c = F[r][0] * vx[v] + F[r][1] * vy[v];
                    ge.ext[r] = fmax(ge.ext[r], fabs(c));
                    v2 += c * c;
                }
                rmax = fmax(rmax, sqrt(v2));
            }
            ge.R = rmax;
        } else {
            for (int r = 0; r < 3; ++r) ge.ext[r] = 0.5 * (fabs(F[r][0]) + fabs(F[r][1]) + fabs(F[r][2]));
            double rmax = 0;
            for (int sy = -1; sy <= 1; sy += 2)
                for (int sz = -1; sz <= 1; sz += 2) {
                    double v2 = 0;
                    for (int r = 0; r < 3; ++r) {
                        const double v = 0.5 * (F[r][0] + sy * F[r][1] + sz * F[r][2]);
                        v2 += v * v;
                    }
                    rmax = fmax(rmax, sqrt(v2));
                }
            ge.R = rmax;
        }
        const double cn = sqrt(ge.c[0] * ge.c[0] + ge.c[1] * ge.c[1] + ge.c[2] * ge.c[2]);
        atomic_max_pos(&g->r0, cn + ge.R);
    }
    geo[i] = ge;
}

// ---- pass 2: padded bounds, filter record, tree leaf (rt_scene.cpp pass 2 + the lbvh_leaves block)
__global__ void __launch_bounds__(kB) bounds_pass2(const rt_object* __restrict__ objs, const uint32_t n, const Geo* __restrict__ geo, const double omax,
                                                   CullRec* __restrict__ cull, LbvhLeaf* __restrict__ leaves, float4* __restrict__ aabb, BoundsGlobals* g) {
    const uint32_t i = blockIdx.x * kB + threadIdx.x;
    if (i >= n) return;
    const uint32_t type = objs[i].type;
    const Geo ge = geo[i];
    CullRec cr;
    cr.type = type;
    LbvhLeaf lf;
    lf.obj = i;
    if (type > 2u) {  // never hit: a point leaf at the origin, a filter record that rejects everything
        cr.cx = cr.cy = cr.cz = 0.f;
        cr.r2 = -1.f;
        cr.h = 1.f;
        cr.cc = 0.f;
        cr.e0 = 1.f;
        cull[i] = cr;
        lf.c[0] = lf.c[1] = lf.c[2] = 0.f;
        lf.half[0] = lf.half[1] = lf.half[2] = 0.f;
        lf.f = 0.f;
        leaves[i] = lf;
        aabb[i] = make_float4(0.f, 0.f, 0.f, -1.f);
        atomic_minmax_float(&g->lo[0], &g->hi[0], 0.f);
        atomic_minmax_float(&g->lo[1], &g->hi[1], 0.f);
        atomic_minmax_float(&g->lo[2], &g->hi[2], 0.f);
        return;
    }
    const double wdev_max = g->wdev_max;
    const double ratio = ge.smax / ge.smin;
    double gg, ff, rscale = 1.0;
    if (type == 0u) {
        gg = K_FUZZ * U * ratio * ratio + 24.0 * U;
        ff = sqrt(gg);
    } else {
        ff = C_LIN * U * ratio * ratio;
        gg = ff * ff * (1.0 + 1.0 / LIN_EPS) + 24.0 * U;
        rscale = sqrt(1.0 + LIN_EPS);
    }
    if (gg > 0.25 || ff > 0.5) {
        fail_at(g, i, RT_BOUNDS_ANISOTROPY);
        return;
    }
    const double cn = sqrt(ge.c[0] * ge.c[0] + ge.c[1] * ge.c[1] + ge.c[2] * ge.c[2]);
    const double pad = ratio * (omax + cn) * (32.0 * U + 40.0 * wdev_max) + 1e-6 * ge.R;
    double half[3];
    for (int a = 0; a < 3; ++a) half[a] = ge.ext[a] * (1.0 + 1e-6) + pad;
    const double R = (ge.R * (1.0 + 1e-6) + pad) * rscale;
    cr.cx = (float)ge.c[0];
    cr.cy = (float)ge.c[1];
    cr.cz = (float)ge.c[2];
    const double Rf = R + 2.0 * U * cn;
    cr.r2 = up(Rf * Rf);
    cr.h = down(1.0 - gg);
    const double ccf = (double)cr.cx * cr.cx + (double)cr.cy * cr.cy + (double)cr.cz * cr.cz;
    cr.cc = (float)ccf;
    cr.e0 = down((double)cr.h * ccf * (1.0 - 8.0 * U) - (double)cr.r2);
    cull[i] = cr;
    for (int k = 0; k < 3; ++k) {
        lf.c[k] = (float)ge.c[k];
        lf.half[k] = up(half[k] + fabs(ge.c[k] - (double)lf.c[k]));
        atomic_minmax_float(&g->lo[k], &g->hi[k], lf.c[k]);
    }
    lf.f = up(ff);
    leaves[i] = lf;
    {   // the box a ray from anywhere in the scene can need (rt_scene.cpp pass 2: HostScene::aabb) and the bounds of all of them
        const double infl = ff * (omax + cn) * 1.0001;
        float hb[3];
        const float ctr[3] = {cr.cx, cr.cy, cr.cz};
        for (int k = 0; k < 3; ++k) {
            hb[k] = up(half[k] + infl + 2.0 * U * cn);
            atomicMin(reinterpret_cast<int*>(&g->box_lo[k]), ordered_key(__fsub_rd(ctr[k], hb[k])));
            atomicMax(reinterpret_cast<int*>(&g->box_hi[k]), ordered_key(__fadd_ru(ctr[k], hb[k])));
        }
        aabb[i] = make_float4(hb[0], hb[1], hb[2], 1.f);
        atomicAdd(&g->live, 1u);
    }
    atomicMin(reinterpret_cast<unsigned*>(&g->h_min), __float_as_uint(cr.h));   // 0 < h <= 1: bit patterns order like the values
    atomicMax(reinterpret_cast<unsigned*>(&g->f_max), __float_as_uint(lf.f));
}

}  // namespace

size_t bounds_scratch_bytes(uint32_t n) { return (size_t)n * sizeof(Geo); }

void bounds_globals_init(BoundsGlobals* h) {
    h->r0 = 0.0;
    h->wdev_max = 0.0;
    h->first_bad = ~0ull;
    h->h_min = 1.f;
    h->f_max = 0.f;
    const int big = 0x7f7fffff;  // ordered-integer image of +MAX_FLOAT
    for (int k = 0; k < 3; ++k) {
        reinterpret_cast<int*>(h->lo)[k] = big;
        reinterpret_cast<int*>(h->hi)[k] = -big;
        reinterpret_cast<int*>(h->box_lo)[k] = big;
        reinterpret_cast<int*>(h->box_hi)[k] = -big;
    }
    h->live = 0;
}

float bounds_decode_ordered(float stored) {
    int i;
    memcpy(&i, &stored, 4);
    const int bits = i >= 0 ? i : (int)(0x80000000u - (unsigned)i);
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

cudaError_t bounds_pass1_launch(const rt_object* d_objs, uint32_t n, float4* minv, float4* mv, float4* mat, void* scratch, BoundsGlobals* d_g,
                                cudaStream_t stream) {
    if (n) bounds_pass1<<<(n + kB - 1) / kB, kB, 0, stream>>>(d_objs, n, minv, mv, mat, reinterpret_cast<Geo*>(scratch), d_g);
    return cudaGetLastError();
}

cudaError_t bounds_pass2_launch(const rt_object* d_objs, uint32_t n, const void* scratch, double omax, CullRec* cull, LbvhLeaf* leaves, float4* aabb,
                                BoundsGlobals* d_g, cudaStream_t stream) {
    if (n) bounds_pass2<<<(n + kB - 1) / kB, kB, 0, stream>>>(d_objs, n, reinterpret_cast<const Geo*>(scratch), omax, cull, leaves, aabb, d_g);
    return cudaGetLastError();
}

}  // namespace rt
