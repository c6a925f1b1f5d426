// rt_scene.cpp -- host side of rt_upload_scene: AoS -> SoA repack (replaces the cl_* marshalling of
// OpenCLRaytracer.cpp:16-24,108-146), conservative per-object bounds, BVH build.
//
// Why the bounds need care.  The reference decides hit / miss in binary32 with the arithmetic of
// shade_and_reflect_kernel.cl:78-144, and that arithmetic is *noisy*: for a sphere the discriminant
// B*B - 4*A*C cancels catastrophically when the ray origin is far from the object in object-space
// units (|s| >> 1), so the silhouette the reference actually renders differs from the geometric one
// by up to ~15 u |s|^2 in squared object-space radius (u = 2^-24; derivation in DESIGN.md).  The CUDA
// path reproduces that arithmetic bit for bit (ex:: in rt_device.cuh), but it only runs it on pairs
// that survive a cheap conservative filter, so every bound produced here must contain the *noisy*
// object, not just the geometric one:
//
//      fuzzy squared radius (object space)   rho^2 <= 1 + K u |s|^2,          |s| <= |oc| / sigma_min
//      => view space, any direction          extent <= sigma_max * rho
//      => bounding sphere test               p^2 <= R^2 + g |oc|^2,           g = K u (sigma_max/sigma_min)^2
//      => box inflation                      + sqrt(g) * |oc|,                |oc| <= distance origin -> farthest box point
//
// plus static pads for the rounding of the transform itself.  K = 32 is twice the first-order worst case.
// tests/test_emul.py (test_filters_never_drop_a_noisy_hit, test_perspective_grids_never_drop_a_noisy_hit) checks on
// adversarial scenes that no exact hit is ever filtered.
#include "rt_scene.hpp"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <numeric>
#include <thread>

namespace rt {
namespace {

const double U = 5.9604644775390625e-08;  // 2^-24
const double K_FUZZ = 32.0;
// Boxes (and the triangle extension) decide hit / miss without a cancelling discriminant: their noise is LINEAR in the
// distance.  The ray direction in object space carries a relative error <= 3 sqrt(3) u ratio (three products summed,
// |A dir| >= |dir| / sigma_max), i.e. the object appears moved sideways by <= 5.2 u ratio^2 |oc| in view space; every
// slab parameter t = (+-0.5 - s) / d carries <= 2u relative error (a face moved by <= 2u ratio |oc|); the triangle's
// in-plane point s + t d <= 8u ratio |oc|.  Together <= 13.2 u ratio^2 |oc|; C_LIN = 24 leaves a factor 1.8.
const double C_LIN = 24.0;
// bounding-sphere filter for a linear-noise object:  (R + lin |oc|)^2 <= R^2 (1 + eps) + lin^2 (1 + 1/eps) |oc|^2
const double LIN_EPS = 1.0 / 64.0;

struct M3 {
    double a[3][3];
};

bool inverse3(const M3& m, M3& out) {
    const double (*a)[3] = m.a;
    const double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
    const double c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
    const double c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    const double det = a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02;
    if (!(fabs(det) > 0.0) || !std::isfinite(det)) return false;
    const double id = 1.0 / det;
    out.a[0][0] = c00 * id;
    out.a[1][0] = c01 * id;
    out.a[2][0] = c02 * id;
    out.a[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) * id;
    out.a[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) * id;
    out.a[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) * id;
    out.a[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) * id;
    out.a[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) * id;
    out.a[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) * id;
    return true;
}

// Extreme singular values of a 3x3 matrix: sqrt of the eigenvalues of M^T M (cyclic Jacobi).
void singular_range(const M3& m, double& smin, double& smax) {
    double s[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            s[i][j] = 0;
            for (int k = 0; k < 3; ++k) s[i][j] += m.a[k][i] * m.a[k][j];
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
                for (int k = 0; k < 3; ++k) {  // S <- S * J
                    const double skp = s[k][p], skq = s[k][q];
                    s[k][p] = c * skp - sn * skq;
                    s[k][q] = sn * skp + c * skq;
                }
                for (int k = 0; k < 3; ++k) {  // S <- J^T * S
                    const double spk = s[p][k], sqk = s[q][k];
                    s[p][k] = c * spk - sn * sqk;
                    s[q][k] = sn * spk + c * sqk;
                }
            }
    }
    const double e0 = std::max(s[0][0], 0.0), e1 = std::max(s[1][1], 0.0), e2 = std::max(s[2][2], 0.0);
    smax = sqrt(std::max(e0, std::max(e1, e2)));
    smin = sqrt(std::min(e0, std::min(e1, e2)));
}

float up(double v) {  // round to float, then one ulp outward (towards +inf)
    float f = (float)v;
    if ((double)f < v) f = nextafterf(f, INFINITY);
    return nextafterf(f, INFINITY);
}
float down(double v) {
    float f = (float)v;
    if ((double)f > v) f = nextafterf(f, -INFINITY);
    return nextafterf(f, -INFINITY);
}

// Runs body(i, worker) for i in [0, n) on up to 16 host threads (contiguous chunks); returns the smallest index whose
// body returned non-zero (with that code), or n.  The bounds of 10^5 objects are a Jacobi SVD each: ~90 ms on one core.
// RT_BUILD_TIMING=1: stage times of build_scene on stderr
struct StageTimer {
    bool on = getenv("RT_BUILD_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        const auto n = std::chrono::steady_clock::now();
        fprintf(stderr, "build_scene: %-28s %7.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};

int g_build_threads = 0;  // 0: as many as the host has, at most 16 (set through build_scene's `threads`)
int worker_count() {
    unsigned hw = std::thread::hardware_concurrency();
    const int dflt = (int)std::min<unsigned>(hw ? hw : 1u, 16u);
    return g_build_threads > 0 ? std::min(g_build_threads, 16) : dflt;
}

template <typename Body>
uint32_t parallel_objects(uint32_t n, int* code_out, Body body, int* n_workers_out) {
    int nw = worker_count();
    if (n < 4096u) nw = 1;
    *n_workers_out = nw;
    std::vector<uint32_t> first_bad(nw, n);
    std::vector<int> codes(nw, 0);
    auto run = [&](int w) {
        const uint32_t b = (uint32_t)((uint64_t)n * w / nw), e = (uint32_t)((uint64_t)n * (w + 1) / nw);
        for (uint32_t i = b; i < e; ++i) {
            const int rc = body(i, w);
            if (rc != 0) {
                first_bad[w] = i;
                codes[w] = rc;
                return;
            }
        }
    };
    if (nw == 1) {
        run(0);
    } else {
        std::vector<std::thread> pool;
        for (int w = 0; w < nw; ++w) pool.emplace_back(run, w);
        for (auto& t : pool) t.join();
    }
    uint32_t bad = n;
    *code_out = 0;
    for (int w = 0; w < nw; ++w)
        if (first_bad[w] < bad) {
            bad = first_bad[w];
            *code_out = codes[w];
        }
    return bad;
}

struct ObjBound {
    double c[3];     // centre
    double half[3];  // padded half extents of the AABB
    double R;        // padded bounding-sphere radius
    double g;        // bounding-sphere filter: hit => p^2 <= R^2 + g |oc|^2
    double f;        // boxes: hit => the ray passes within (half + f |oc|) of the centre per axis
};

struct BuildNode {
    double lo[3], hi[3];
    double f;
    int left = -1, right = -1;  // BuildNode indices
    uint32_t first = 0, count = 0;
    int depth = 0;
};

// What the builder needs to know about a set of objects: the box around their padded boxes, the box around their
// centres, the largest noise coefficient, the count.  Unions of these are exact (min / max / sum), so a child's record
// taken from the parent's bins equals the one a pass over the child's objects would give.
struct Extent {
    double lo[3], hi[3], clo[3], chi[3], f;
    uint32_t n;
    Extent() { reset(); }
    void reset() {
        for (int k = 0; k < 3; ++k) {
            lo[k] = clo[k] = INFINITY;
            hi[k] = chi[k] = -INFINITY;
        }
        f = 0;
        n = 0;
    }
    void add(const ObjBound& b) {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b.c[k] - b.half[k]);
            hi[k] = std::max(hi[k], b.c[k] + b.half[k]);
            clo[k] = std::min(clo[k], b.c[k]);
            chi[k] = std::max(chi[k], b.c[k]);
        }
        f = std::max(f, b.f);
        ++n;
    }
    void merge(const Extent& o) {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], o.lo[k]);
            hi[k] = std::max(hi[k], o.hi[k]);
            clo[k] = std::min(clo[k], o.clo[k]);
            chi[k] = std::max(chi[k], o.chi[k]);
        }
        f = std::max(f, o.f);
        n += o.n;
    }
};

// A subtree whose construction was put off: node `node` of the builder that deferred it covers idx[first, first+count).
struct BuildTask {
    int node;
    uint32_t first, count;
    int depth;
    Extent ext;
};

constexpr int kBins = 16;
struct BinSet {
    Extent bin[3][kBins];
};

// Binned-SAH builder (16 bins per axis over the centres).  One pass over a node's objects fills the bins of all three
// axes; the children's records come out of the bins, so every object is touched once per tree level.
struct Builder {
    const std::vector<ObjBound>& ob;
    std::vector<uint32_t>& idx;  // object permutation, shared by all builders of a tree (they work on disjoint ranges)
    std::vector<BuildNode> bn;
    int leaf_size;
    uint32_t max_depth = 0;
    int workers = 1;  // > 1: passes over >= 16384 objects are split over this many threads (top builder only)
    // top builder only: subtrees at this depth (or smaller than defer_count objects) are recorded, not built
    std::vector<BuildTask>* defer = nullptr;
    int defer_depth = 0;
    uint32_t defer_count = 0;

    Builder(const std::vector<ObjBound>& o, std::vector<uint32_t>& ix, int ls) : ob(o), idx(ix), leaf_size(ls) {}

    template <typename Acc, typename Body>
    void pass(uint32_t first, uint32_t count, Acc& total, Body body) const {  // body(acc, object); total.merge(acc) per chunk
        const int nw = (workers > 1 && count >= 16384u) ? workers : 1;
        if (nw == 1) {
            for (uint32_t i = first; i < first + count; ++i) body(total, ob[idx[i]]);
            return;
        }
        std::vector<Acc> part(nw);
        std::vector<std::thread> pool;
        for (int w = 0; w < nw; ++w)
            pool.emplace_back([&, w] {
                const uint32_t b = first + (uint32_t)((uint64_t)count * w / nw), e = first + (uint32_t)((uint64_t)count * (w + 1) / nw);
                for (uint32_t i = b; i < e; ++i) body(part[w], ob[idx[i]]);
            });
        for (auto& t : pool) t.join();
        for (int w = 0; w < nw; ++w) total.merge(part[w]);
    }

    Extent measure(uint32_t first, uint32_t count) const {
        Extent e;
        pass(first, count, e, [](Extent& acc, const ObjBound& o) { acc.add(o); });
        return e;
    }

    static double area(const double* lo, const double* hi) {
        const double x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return 2.0 * (x * y + y * z + z * x);
    }

    int build(uint32_t first, uint32_t count, int depth, const Extent& e) {
        const int me = (int)bn.size();
        bn.emplace_back();
        {
            BuildNode n;
            for (int k = 0; k < 3; ++k) {
                n.lo[k] = e.lo[k];
                n.hi[k] = e.hi[k];
            }
            n.f = e.f;
            n.first = first;
            n.count = count;
            n.depth = depth;
            bn[me] = n;
        }
        max_depth = std::max<uint32_t>(max_depth, (uint32_t)depth);
        if (count <= (uint32_t)leaf_size) return me;
        if (defer && (depth >= defer_depth || count <= defer_count)) {
            defer->push_back(BuildTask{me, first, count, depth, e});
            return me;
        }

        // bins of all three axes in one pass
        double scale[3];
        bool splittable[3];
        for (int a = 0; a < 3; ++a) {
            const double ext = e.chi[a] - e.clo[a];
            splittable[a] = ext > 0;
            scale[a] = splittable[a] ? kBins / ext : 0.0;
        }
        struct Bins : BinSet {
            void merge(const Bins& o) {
                for (int a = 0; a < 3; ++a)
                    for (int b = 0; b < kBins; ++b) bin[a][b].merge(o.bin[a][b]);
            }
        } bins;
        pass(first, count, bins, [&](Bins& acc, const ObjBound& o) {
            for (int a = 0; a < 3; ++a) {
                if (!splittable[a]) continue;
                int b = (int)((o.c[a] - e.clo[a]) * scale[a]);
                b = std::min(std::max(b, 0), kBins - 1);
                acc.bin[a][b].add(o);
            }
        });
        int best_axis = -1, best_bin = -1;
        double best_cost = INFINITY;
        for (int a = 0; a < 3; ++a) {
            if (!splittable[a]) continue;
            const Extent* bin = bins.bin[a];
            double rl[kBins][3], rh[kBins][3];
            uint32_t rn[kBins];
            double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
            uint32_t n = 0;
            for (int b = kBins - 1; b > 0; --b) {
                for (int k = 0; k < 3; ++k) {
                    lo[k] = std::min(lo[k], bin[b].lo[k]);
                    hi[k] = std::max(hi[k], bin[b].hi[k]);
                    rl[b][k] = lo[k];
                    rh[b][k] = hi[k];
                }
                n += bin[b].n;
                rn[b] = n;
            }
            for (int k = 0; k < 3; ++k) {
                lo[k] = INFINITY;
                hi[k] = -INFINITY;
            }
            n = 0;
            for (int b = 0; b < kBins - 1; ++b) {
                for (int k = 0; k < 3; ++k) {
                    lo[k] = std::min(lo[k], bin[b].lo[k]);
                    hi[k] = std::max(hi[k], bin[b].hi[k]);
                }
                n += bin[b].n;
                if (n == 0 || rn[b + 1] == 0) continue;
                const double cost = area(lo, hi) * n + area(rl[b + 1], rh[b + 1]) * rn[b + 1];
                if (cost < best_cost) {
                    best_cost = cost;
                    best_axis = a;
                    best_bin = b;
                }
            }
        }
        uint32_t mid;
        Extent el, er;
        bool have_extents = false;
        if (best_axis < 0 || depth >= 32) {  // depth cap: median splits from here on keep the tree below RT_BVH_STACK
            mid = first + count / 2;  // (also when all centroids coincide)
        } else {
            const int a = best_axis;
            const double clo_a = e.clo[a], sc = scale[a];
            auto it = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t id) {
                int b = (int)((ob[id].c[a] - clo_a) * sc);
                b = std::min(std::max(b, 0), kBins - 1);
                return b <= best_bin;
            });
            mid = (uint32_t)(it - idx.begin());
            if (mid == first || mid == first + count) {
                mid = first + count / 2;
            } else {
                for (int b = 0; b <= best_bin; ++b) el.merge(bins.bin[a][b]);
                for (int b = best_bin + 1; b < kBins; ++b) er.merge(bins.bin[a][b]);
                have_extents = true;
            }
        }
        if (!have_extents) {
            el = measure(first, mid - first);
            er = measure(mid, first + count - mid);
        }
        const int l = build(first, mid - first, depth + 1, el);
        const int r = build(mid, first + count - mid, depth + 1, er);
        bn[me].left = l;
        bn[me].right = r;
        return me;
    }
};

}  // namespace

int build_scene(const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights, float omax_user,
                int leaf_size, HostScene& out, std::string& err, int threads) {
    g_build_threads = threads;
    out = HostScene();
    out.n_objs = n_objs;
    out.n_lights = n_lights;
    if ((n_objs && !objs) || (n_lights && !lights)) {
        err = "null scene arrays";
        return RT_ERR_INVALID;
    }
    if (n_objs >= (1u << 27)) {
        err = "too many objects";
        return RT_ERR_INVALID;
    }
    const bool device_tree = leaf_size == 0;
    leaf_size = std::min(std::max(leaf_size, 1), 16);

    out.minv.resize(4 * (size_t)n_objs);
    out.mv.resize(4 * (size_t)n_objs);
    out.mat.resize(3 * (size_t)n_objs);
    out.cull.resize(n_objs);
    out.aabb.assign(n_objs, F4{0.f, 0.f, 0.f, -1.f});
    out.lights.resize(4 * (size_t)n_lights);

    for (uint32_t i = 0; i < n_lights; ++i) {
        const rt_light& l = lights[i];
        out.lights[4 * i + 0] = F4{l.ambient[0], l.ambient[1], l.ambient[2], 0.f};
        out.lights[4 * i + 1] = F4{l.diffuse[0], l.diffuse[1], l.diffuse[2], 0.f};
        out.lights[4 * i + 2] = F4{l.specular[0], l.specular[1], l.specular[2], 0.f};
        out.lights[4 * i + 3] = F4{l.position[0], l.position[1], l.position[2], l.position[3]};
    }

    StageTimer timer;
    timer.lap("repack lights, resize");
    // ---- pass 1: geometry of every object in double precision
    struct Geo {
        double c[3], ext[3], R, smin, smax, wdev;
    };
    std::vector<Geo> geo(n_objs);
    double r0 = 0.0, wdev_max = 0.0;
    constexpr int kMaxWorkers = 16;
    double r0_w[kMaxWorkers] = {}, wdev_w[kMaxWorkers] = {};
    std::string err_w[kMaxWorkers];
    auto pass1 = [&](uint32_t i, int worker) -> int {
        std::string& err = err_w[worker];
        double& r0 = r0_w[worker];
        double& wdev_max = wdev_w[worker];
        const rt_object& o = objs[i];
        for (int k = 0; k < 16; ++k)
            if (!std::isfinite(o.mv[k]) || !std::isfinite(o.mvInverse[k])) {
                err = "object " + std::to_string(i) + ": non-finite matrix";
                return RT_ERR_UNSUPPORTED;
            }
        for (int c = 0; c < 4; ++c) {
            out.minv[4 * (size_t)i + c] = F4{o.mvInverse[4 * c], o.mvInverse[4 * c + 1], o.mvInverse[4 * c + 2], o.mvInverse[4 * c + 3]};
            out.mv[4 * (size_t)i + c] = F4{o.mv[4 * c], o.mv[4 * c + 1], o.mv[4 * c + 2], o.mv[4 * c + 3]};
        }
        out.mat[3 * (size_t)i + 0] = F4{o.mat.ambient[0], o.mat.ambient[1], o.mat.ambient[2], o.mat.absorption};
        out.mat[3 * (size_t)i + 1] = F4{o.mat.diffuse[0], o.mat.diffuse[1], o.mat.diffuse[2], o.mat.shininess};
        out.mat[3 * (size_t)i + 2] = F4{o.mat.specular[0], o.mat.specular[1], o.mat.specular[2], 0.f};

        // affine check: bottom rows (0,0,0,1) up to rounding of glm::inverse
        const float* mi = o.mvInverse;
        const float* mv = o.mv;
        const double brow = fabs(mi[3]) + fabs(mi[7]) + fabs(mi[11]) + fabs(mv[3]) + fabs(mv[7]) + fabs(mv[11]);
        const double wd = std::max(fabs((double)mi[15] - 1.0), fabs((double)mv[15] - 1.0));
        if (brow > 0.0 || wd > 1e-4) {
            err = "object " + std::to_string(i) + ": projective (non-affine) matrices are outside the supported domain";
            return RT_ERR_UNSUPPORTED;
        }
        Geo& g = geo[i];
        g.wdev = wd;
        wdev_max = std::max(wdev_max, wd);
        M3 A, F;
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) A.a[r][c] = mi[4 * c + r];
        if (o.type > 2u) {  // never hit: keep a degenerate record far away
            g.c[0] = g.c[1] = g.c[2] = 0;
            g.ext[0] = g.ext[1] = g.ext[2] = 0;
            g.R = 0;
            g.smin = g.smax = 1;
            return 0;
        }
        if (!inverse3(A, F)) {
            err = "object " + std::to_string(i) + ": singular mvInverse";
            return RT_ERR_UNSUPPORTED;
        }
        for (int r = 0; r < 3; ++r) g.c[r] = -(F.a[r][0] * mi[12] + F.a[r][1] * mi[13] + F.a[r][2] * mi[14]);
        singular_range(F, g.smin, g.smax);
        if (!(g.smin > 0) || !std::isfinite(g.smax)) {
            err = "object " + std::to_string(i) + ": degenerate transform";
            return RT_ERR_UNSUPPORTED;
        }
        if (o.type == 0u) {
            for (int r = 0; r < 3; ++r) g.ext[r] = sqrt(F.a[r][0] * F.a[r][0] + F.a[r][1] * F.a[r][1] + F.a[r][2] * F.a[r][2]);
            g.R = g.smax;
        } else if (o.type == 2u) {
            // unit triangle (-.5,-.5,0) (.5,-.5,0) (-.5,.5,0): bounded around the image of the object origin (the midpoint
            // of the hypotenuse).  Its hit decision rounds like the box's slabs (error linear in |s|), same pads.
            const double vx[3] = {-0.5, 0.5, -0.5}, vy[3] = {-0.5, -0.5, 0.5};
            double rmax = 0;
            for (int r = 0; r < 3; ++r) g.ext[r] = 0;
            for (int v = 0; v < 3; ++v) {
                double v2 = 0;
                for (int r = 0; r < 3; ++r) {
                    const double c = F.a[r][0] * vx[v] + F.a[r][1] * vy[v];
                    g.ext[r] = std::max(g.ext[r], fabs(c));
                    v2 += c * c;
                }
                rmax = std::max(rmax, sqrt(v2));
            }
            g.R = rmax;
        } else {
            for (int r = 0; r < 3; ++r) g.ext[r] = 0.5 * (fabs(F.a[r][0]) + fabs(F.a[r][1]) + fabs(F.a[r][2]));
            double rmax = 0;
            for (int sy = -1; sy <= 1; sy += 2)
                for (int sz = -1; sz <= 1; sz += 2) {
                    double v2 = 0;
                    for (int r = 0; r < 3; ++r) {
                        const double v = 0.5 * (F.a[r][0] + sy * F.a[r][1] + sz * F.a[r][2]);
                        v2 += v * v;
                    }
                    rmax = std::max(rmax, sqrt(v2));
                }
            g.R = rmax;
        }
        const double cn = sqrt(g.c[0] * g.c[0] + g.c[1] * g.c[1] + g.c[2] * g.c[2]);
        r0 = std::max(r0, cn + g.R);
        return 0;
    };
    {
        int code = 0, nw = 1;
        const uint32_t bad = parallel_objects(n_objs, &code, pass1, &nw);
        if (bad < n_objs) {
            for (int w = 0; w < nw; ++w)
                if (!err_w[w].empty() && err_w[w].find("object " + std::to_string(bad) + ":") == 0) err = err_w[w];
            return code;
        }
        for (int w = 0; w < nw; ++w) {
            r0 = std::max(r0, r0_w[w]);
            wdev_max = std::max(wdev_max, wdev_w[w]);
        }
    }
    // Ray origins: the camera (0), caller rays (omax_user) and points on object surfaces + skin.
    const double omax = std::max(1.05 * r0 + 1.0, (double)omax_user * 1.0001 + 1e-3);
    out.omax = (float)omax;
    out.scene_radius = (float)r0;

    timer.lap("pass 1 (geometry, SVD)");
    // ---- pass 2: padded bounds
    std::vector<ObjBound> ob(n_objs);
    std::vector<uint32_t> live;
    live.reserve(n_objs);
    float hmin_w[kMaxWorkers];
    for (int w = 0; w < kMaxWorkers; ++w) hmin_w[w] = 1.f;
    auto pass2 = [&](uint32_t i, int worker) -> int {
        std::string& err = err_w[worker];
        const Geo& g = geo[i];
        CullRec& cr = out.cull[i];
        cr.type = objs[i].type;
        if (objs[i].type > 2u) {
            cr.cx = cr.cy = cr.cz = 0.f;
            cr.r2 = -1.f;  // h*|oc|^2 - dt^2 >= ... never <= -1 unless NaN
            cr.h = 1.f;
            cr.cc = 0.f;
            cr.e0 = 1.f;
            return 0;
        }
        const double ratio = g.smax / g.smin;
        // 24u: rounding of the filter's own arithmetic (h |oc|^2 - dt^2 in binary32)
        double gg, ff, rscale = 1.0;
        if (objs[i].type == 0u) {  // sphere: noisy discriminant, radius^2 grows by K u ratio^2 |oc|^2
            gg = K_FUZZ * U * ratio * ratio + 24.0 * U;
            ff = sqrt(gg);
        } else {  // box / triangle: linear noise
            ff = C_LIN * U * ratio * ratio;
            gg = ff * ff * (1.0 + 1.0 / LIN_EPS) + 24.0 * U;
            rscale = sqrt(1.0 + LIN_EPS);
        }
        if (gg > 0.25 || ff > 0.5) {
            err = "object " + std::to_string(i) + ": anisotropy too extreme for conservative bounds";
            return RT_ERR_UNSUPPORTED;
        }
        const double cn = sqrt(g.c[0] * g.c[0] + g.c[1] * g.c[1] + g.c[2] * g.c[2]);
        // rounding of s = mvInverse * (o, w): 4 products / 3 sums of magnitude <= (|o| + |c|) / sigma_min each,
        // plus the drift of w away from 1 (mvInverse[15] != 1 after glm::inverse), mapped back to view space.
        const double pad = ratio * (omax + cn) * (32.0 * U + 40.0 * wdev_max) + 1e-6 * g.R;
        ObjBound& b = ob[i];
        for (int a = 0; a < 3; ++a) {
            b.c[a] = g.c[a];
            b.half[a] = g.ext[a] * (1.0 + 1e-6) + pad;
        }
        b.R = (g.R * (1.0 + 1e-6) + pad) * rscale;
        b.g = gg;
        b.f = ff;
        {
            const double infl = ff * (omax + cn) * 1.0001;
            out.aabb[i] = F4{up(b.half[0] + infl + 2.0 * U * cn), up(b.half[1] + infl + 2.0 * U * cn), up(b.half[2] + infl + 2.0 * U * cn), 1.f};
        }
        cr.cx = (float)g.c[0];
        cr.cy = (float)g.c[1];
        cr.cz = (float)g.c[2];
        // centre rounding to float moves the sphere by <= u*|c|: add it to the radius
        const double Rf = b.R + 2.0 * U * cn;
        cr.r2 = up(Rf * Rf);
        cr.h = down(1.0 - gg);
        const double ccf = (double)cr.cx * cr.cx + (double)cr.cy * cr.cy + (double)cr.cz * cr.cz;
        cr.cc = (float)ccf;
        cr.e0 = down((double)cr.h * ccf * (1.0 - 8.0 * U) - (double)cr.r2);
        hmin_w[worker] = std::min(hmin_w[worker], cr.h);
        return 0;
    };
    {
        int code = 0, nw = 1;
        for (int w = 0; w < kMaxWorkers; ++w) err_w[w].clear();
        const uint32_t bad = parallel_objects(n_objs, &code, pass2, &nw);
        if (bad < n_objs) {
            for (int w = 0; w < nw; ++w)
                if (!err_w[w].empty() && err_w[w].find("object " + std::to_string(bad) + ":") == 0) err = err_w[w];
            return code;
        }
        for (int w = 0; w < nw; ++w) out.h_min = std::min(out.h_min, hmin_w[w]);
        for (uint32_t i = 0; i < n_objs; ++i)
            if (objs[i].type <= 2u) live.push_back(i);
    }

    timer.lap("pass 2 (padded bounds)");
    // ---- BVH over the live objects
    out.root = 0;
    if (live.empty()) {
        out.root = ~0;  // empty leaf; never dereferenced because n_objs == 0 or count handled below
        BvhNode n;
        memset(&n, 0, sizeof n);
        out.nodes.push_back(n);
        out.leaf_obj.push_back(0);
        CullRec dead;
        memset(&dead, 0, sizeof dead);
        dead.r2 = -1.f;
        dead.h = 1.f;
        dead.type = 0xffffffffu;  // no primitive: the kernels' switch skips it (n_objs may be > 0 with unknown types only)
        out.leaf_cull.push_back(dead);
        return RT_OK;
    }
    std::vector<ObjBound> lob(live.size());
    for (size_t i = 0; i < live.size(); ++i) lob[i] = ob[live[i]];
    if (device_tree && live.size() > 1) {  // the tree is built on the device (rt_lbvh.cu) from these records
        out.lbvh_leaves.resize(live.size());
        for (int k = 0; k < 3; ++k) {
            out.lbvh_lo[k] = INFINITY;
            out.lbvh_hi[k] = -INFINITY;
        }
        double fmax = 0;
        for (size_t i = 0; i < live.size(); ++i) {
            LbvhLeaf& l = out.lbvh_leaves[i];
            for (int k = 0; k < 3; ++k) {
                l.c[k] = (float)lob[i].c[k];
                l.half[k] = up(lob[i].half[k] + fabs(lob[i].c[k] - (double)l.c[k]));  // the centre moved by its rounding
                out.lbvh_lo[k] = std::min(out.lbvh_lo[k], l.c[k]);
                out.lbvh_hi[k] = std::max(out.lbvh_hi[k], l.c[k]);
            }
            l.f = up(lob[i].f);
            l.obj = live[i];
            fmax = std::max(fmax, lob[i].f);
        }
        out.f_global = up(fmax);
        out.root = 0;
        return RT_OK;
    }
    // The tree is built in rounds of three levels: the root's three levels here (their passes over the objects split
    // over the host threads), then every subtree that was put off as a task of the next round, all tasks of a round in
    // parallel on worker threads with private node lists that are spliced in afterwards.  The node numbering below is a
    // preorder walk, so the result does not depend on the thread count.
    std::vector<uint32_t> perm(live.size());
    std::iota(perm.begin(), perm.end(), 0u);
    Builder bld(lob, perm, leaf_size);
    bld.bn.reserve(2 * live.size());
    std::vector<BuildTask> frontier;
    const int nw = live.size() >= 8192 ? worker_count() : 1;
    if (nw > 1) {
        bld.defer = &frontier;
        bld.defer_depth = 3;
        bld.defer_count = 2048;
    }
    bld.workers = nw;
    const int root = bld.build(0, (uint32_t)live.size(), 0, bld.measure(0, (uint32_t)live.size()));
    bld.workers = 1;
    while (!frontier.empty()) {
        std::vector<BuildTask> tasks;
        tasks.swap(frontier);
        std::vector<Builder> sub;
        sub.reserve(tasks.size());
        std::vector<std::vector<BuildTask>> more(tasks.size());
        for (size_t t = 0; t < tasks.size(); ++t) {
            sub.emplace_back(lob, perm, leaf_size);
            if (tasks[t].count > 4096u) {  // large: three more levels, the rest next round
                sub[t].defer = &more[t];
                sub[t].defer_depth = tasks[t].depth + 3;
                sub[t].defer_count = 2048;
            }
        }
        std::atomic<size_t> next{0};
        auto run = [&]() {
            for (size_t t; (t = next.fetch_add(1)) < tasks.size();) sub[t].build(tasks[t].first, tasks[t].count, tasks[t].depth, tasks[t].ext);
        };
        std::vector<std::thread> pool;
        for (int w = 0; w < std::min<int>(nw, (int)tasks.size()); ++w) pool.emplace_back(run);
        for (auto& th : pool) th.join();
        for (size_t t = 0; t < tasks.size(); ++t) {  // splice: local node 0 is the deferred node itself
            const int base = (int)bld.bn.size() - 1;
            auto remap = [&](int j) { return j < 0 ? j : (j == 0 ? tasks[t].node : base + j); };
            const std::vector<BuildNode>& l = sub[t].bn;
            bld.bn[tasks[t].node].left = remap(l[0].left);
            bld.bn[tasks[t].node].right = remap(l[0].right);
            for (size_t j = 1; j < l.size(); ++j) {
                BuildNode n = l[j];
                n.left = remap(n.left);
                n.right = remap(n.right);
                bld.bn.push_back(n);
            }
            bld.max_depth = std::max(bld.max_depth, sub[t].max_depth);
            for (BuildTask m : more[t]) {
                m.node = remap(m.node);
                frontier.push_back(m);
            }
        }
    }
    out.bvh_depth = bld.max_depth;
    // the traversal kernels keep RT_BVH_STACK (64) pending siblings per ray and push without a bound: SAH to depth 32, then
    // median splits of at most 2^27 objects, stays below 60 -- enforced here rather than assumed
    if (out.bvh_depth >= 62u) {
        err = "BVH deeper than the traversal stack";
        return RT_ERR_UNSUPPORTED;
    }

    out.leaf_obj.resize(live.size());
    out.leaf_cull.resize(live.size());
    for (size_t i = 0; i < live.size(); ++i) {
        out.leaf_obj[i] = live[perm[i]];
        out.leaf_cull[i] = out.cull[out.leaf_obj[i]];
    }
    timer.lap("SAH build");
    // flatten: inner BuildNodes get consecutive indices in preorder (node, left subtree, right subtree)
    std::vector<int32_t> inner_index(bld.bn.size(), -1);
    int32_t n_inner = 0;
    {
        std::vector<int> todo;
        todo.push_back(root);
        while (!todo.empty()) {
            const int i = todo.back();
            todo.pop_back();
            if (bld.bn[i].left < 0) continue;
            inner_index[i] = n_inner++;
            todo.push_back(bld.bn[i].right);
            todo.push_back(bld.bn[i].left);
        }
    }
    out.nodes.resize(std::max(n_inner, 1));
    memset(out.nodes.data(), 0, out.nodes.size() * sizeof(BvhNode));
    auto ref_of = [&](int bi) -> int32_t {
        const BuildNode& n = bld.bn[bi];
        if (n.left >= 0) return inner_index[bi];
        return ~(int32_t)((n.first << 4) | (n.count - 1));
    };
    auto child = [&](int bi, float* ctr, float* ext, float& f) {
        const BuildNode& n = bld.bn[bi];
        double rf2 = 0;
        for (int k = 0; k < 3; ++k) {
            const double m = std::max(fabs(n.lo[k]), fabs(n.hi[k]));
            rf2 += m * m;
        }
        // slab arithmetic of cv::box_may_hit: (ctr - o) rounds with error <= u * max(|ctr|, |o|); folded into ext
        const double a = 8.0 * U * (omax + sqrt(rf2));
        // 1.00001: rounding of f * (|x| + |y| + |z|) on the device; RT_SLAB_REL (2e-6): relative rounding of the slab test
        const double fd = n.f * 1.00001 + 2e-6;
        f = up(fd);
        double h[3], hsum = 0;
        for (int k = 0; k < 3; ++k) {
            const double c = 0.5 * (n.lo[k] + n.hi[k]);
            ctr[k] = (float)c;
            // half extent measured from the ROUNDED centre
            h[k] = std::max(n.hi[k] - (double)ctr[k], (double)ctr[k] - n.lo[k]) + a;
            hsum += h[k];
        }
        // static part of the inflation f * |farthest point - origin|_1 = f * sum(|c - o|) + f * sum(half extents)
        for (int k = 0; k < 3; ++k) ext[k] = up(h[k] + (double)f * hsum * (1.0 + 1e-6));
    };
    {
        auto record = [&](uint32_t i, int) -> int {
            const BuildNode& n = bld.bn[i];
            if (n.left < 0) return 0;
            BvhNode& d = out.nodes[inner_index[i]];
            child(n.left, d.ctr0, d.ext0, d.f0);
            child(n.right, d.ctr1, d.ext1, d.f1);
            d.c0 = ref_of(n.left);
            d.c1 = ref_of(n.right);
            return 0;
        };
        int code = 0, nw_rec = 1;
        parallel_objects((uint32_t)bld.bn.size(), &code, record, &nw_rec);
    }
    timer.lap("flatten + node records");
    out.root = ref_of(root);
    out.f_global = up(bld.bn[root].f);
    return RT_OK;
}

}  // namespace rt
