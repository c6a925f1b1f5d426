// rt_lbvh.cu -- device-side BVH build: Morton order + the Karras (2012) binary radix tree + a
// bottom-up refit, for scenes that change between frames (the default builder is the host's binned SAH in
// rt_scene.cpp: a better tree, ~150 ms for 10^5 objects; this one takes about a millisecond).
//
// Input: one LbvhLeaf per live object (padded bounds, rounded outwards on the host).  Output: the same 64-byte
// BvhNode array, leaf_obj / leaf_cull orders and leaf references the traversal kernels already consume -- one
// object per leaf, inner node 0 is the root.  Every float operation that widens a box rounds towards the outside
// (__f*_ru / __f*_rd), so the boxes contain what the host builder's double-precision boxes would contain.
#include "rt_lbvh.cuh"

#include "rt_device.cuh"

#include <cub/cub.cuh>

namespace rt {
namespace {

constexpr int kB = 256;

__device__ __forceinline__ uint32_t expand10(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// 64-bit key: 30-bit Morton code of the centre | position in the input (makes every key unique)
__global__ void lbvh_keys(const LbvhLeaf* __restrict__ leaves, uint32_t n, float3 lo, float3 inv, uint64_t* __restrict__ keys) {
    const uint32_t i = blockIdx.x * kB + threadIdx.x;
    if (i >= n) return;
    const LbvhLeaf l = leaves[i];
    const float x = fminf(fmaxf((l.c[0] - lo.x) * inv.x, 0.f), 1023.f);
    const float y = fminf(fmaxf((l.c[1] - lo.y) * inv.y, 0.f), 1023.f);
    const float z = fminf(fmaxf((l.c[2] - lo.z) * inv.z, 0.f), 1023.f);
    const uint32_t m = (expand10((uint32_t)x) << 2) | (expand10((uint32_t)y) << 1) | expand10((uint32_t)z);
    // a radix tree is at most as deep as its keys have significant bits: 30 (Morton) + 27 (index; build_scene rejects
    // 2^27 objects and more) = 57 levels, below the traversal stack of RT_BVH_STACK entries
    static_assert(30 + 27 < RT_BVH_STACK, "LBVH depth bound exceeds the traversal stack");
    keys[i] = ((uint64_t)m << 32) | i;
}

__device__ __forceinline__ int delta(const uint64_t* __restrict__ k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(k[i] ^ k[j]));
}

// Karras 2012, one thread per inner node: children and parent links.  Child reference: >= 0 inner node, < 0 leaf
// ~((position << 4) | 0) exactly as rt_scene.cpp encodes a one-object leaf.
__global__ void lbvh_topology(const uint64_t* __restrict__ keys, int n, BvhNode* __restrict__ nodes, int32_t* __restrict__ parent_inner,
                              int32_t* __restrict__ parent_leaf) {
    const int i = blockIdx.x * kB + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) > 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    int32_t c0, c1;
    if (lo == gamma) {
        c0 = ~(int32_t)((uint32_t)gamma << 4);
        parent_leaf[gamma] = i;
    } else {
        c0 = gamma;
        parent_inner[gamma] = i;
    }
    if (hi == gamma + 1) {
        c1 = ~(int32_t)((uint32_t)(gamma + 1) << 4);
        parent_leaf[gamma + 1] = i;
    } else {
        c1 = gamma + 1;
        parent_inner[gamma + 1] = i;
    }
    nodes[i].c0 = c0;
    nodes[i].c1 = c1;
    if (i == 0) parent_inner[0] = -1;
}

struct Box {
    float lo[3], hi[3], f;
};

__device__ __forceinline__ Box leaf_box(const LbvhLeaf& l) {
    Box b;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        b.lo[k] = __fsub_rd(l.c[k], l.half[k]);
        b.hi[k] = __fadd_ru(l.c[k], l.half[k]);
    }
    b.f = l.f;
    return b;
}

// a box another thread wrote before its __threadfence(): read around L1
__device__ __forceinline__ Box load_box(const Box* p) {
    Box b;
    const float* f = reinterpret_cast<const float*>(p);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        b.lo[k] = __ldcg(f + k);
        b.hi[k] = __ldcg(f + 3 + k);
    }
    b.f = __ldcg(f + 6);
    return b;
}

// The child record of a node, as rt_scene.cpp's child(): centre, half extent (+ the pad for the rounding of (c - o) in
// the slab test + the static part f * sum(half) of the 1-norm inflation), noise coefficient.
__device__ __forceinline__ void write_child(const Box& b, float omax, float* ctr, float* ext, float& f_out) {
    const float U = 5.9604644775390625e-08f;
    float rf2 = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float m = fmaxf(fabsf(b.lo[k]), fabsf(b.hi[k]));
        rf2 = __fadd_ru(rf2, __fmul_ru(m, m));
    }
    const float a = __fmul_ru(8.f * U, __fadd_ru(omax, __fsqrt_ru(rf2)));
    const float f = __fadd_ru(__fmul_ru(b.f, 1.00001f), 2e-6f);
    float h[3], hsum = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float c = 0.5f * (b.lo[k] + b.hi[k]);
        ctr[k] = c;
        h[k] = __fadd_ru(fmaxf(__fsub_ru(b.hi[k], c), __fsub_ru(c, b.lo[k])), a);
        hsum = __fadd_ru(hsum, h[k]);
    }
    const float stat = __fmul_ru(__fmul_ru(f, hsum), 1.000001f);
#pragma unroll
    for (int k = 0; k < 3; ++k) ext[k] = __fadd_ru(h[k], stat);
    f_out = f;
}

// Leaf order arrays + bottom-up refit: one thread per leaf climbs towards the root; the second thread to arrive at a
// node (atomic counter) owns it, writes its two child records and its own box, and climbs on.
__global__ void lbvh_refit(const uint64_t* __restrict__ keys, const LbvhLeaf* __restrict__ leaves, const CullRec* __restrict__ cull, int n,
                           float omax, BvhNode* __restrict__ nodes, const int32_t* __restrict__ parent_inner,
                           const int32_t* __restrict__ parent_leaf, uint32_t* __restrict__ arrived, Box* boxes,
                           uint32_t* __restrict__ leaf_obj, CullRec* __restrict__ leaf_cull) {
    const int p = blockIdx.x * kB + threadIdx.x;
    if (p >= n) return;
    {
        const LbvhLeaf l = leaves[(uint32_t)keys[p]];
        leaf_obj[p] = l.obj;
        leaf_cull[p] = cull[l.obj];
    }
    int node = parent_leaf[p];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&arrived[node], 1u) == 0u) return;  // the sibling subtree is not finished yet
        const int32_t c0 = nodes[node].c0, c1 = nodes[node].c1;
        Box b0, b1;
        b0 = c0 < 0 ? leaf_box(leaves[(uint32_t)keys[(uint32_t)(~c0) >> 4]]) : load_box(boxes + c0);
        b1 = c1 < 0 ? leaf_box(leaves[(uint32_t)keys[(uint32_t)(~c1) >> 4]]) : load_box(boxes + c1);
        BvhNode nd;
        nd.c0 = c0;
        nd.c1 = c1;
        write_child(b0, omax, nd.ctr0, nd.ext0, nd.f0);
        write_child(b1, omax, nd.ctr1, nd.ext1, nd.f1);
        nodes[node] = nd;
        Box u;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            u.lo[k] = fminf(b0.lo[k], b1.lo[k]);
            u.hi[k] = fmaxf(b0.hi[k], b1.hi[k]);
        }
        u.f = fmaxf(b0.f, b1.f);
        boxes[node] = u;
        node = parent_inner[node];
    }
}

}  // namespace

size_t lbvh_scratch_bytes(uint32_t n) {
    size_t cub_bytes = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, cub_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)n, 0, 62);
    const size_t a = 256;
    auto up = [&](size_t b) { return (b + a - 1) / a * a; };
    return up(cub_bytes) + 2 * up(sizeof(uint64_t) * n) + 2 * up(sizeof(int32_t) * n) + up(sizeof(uint32_t) * n) + up(sizeof(Box) * n);
}

cudaError_t build_lbvh_device(const LbvhLeaf* d_leaves, uint32_t n, const CullRec* d_cull, const float lo[3], const float hi[3],
                              float omax, void* scratch, size_t scratch_bytes, BvhNode* d_nodes, uint32_t* d_leaf_obj,
                              CullRec* d_leaf_cull, cudaStream_t stream) {
    if (n < 2) return cudaErrorInvalidValue;
    if (scratch_bytes < lbvh_scratch_bytes(n)) return cudaErrorInvalidValue;
    size_t cub_bytes = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, cub_bytes, (const uint64_t*)nullptr, (uint64_t*)nullptr, (int)n, 0, 62);
    const size_t a = 256;
    auto up = [&](size_t b) { return (b + a - 1) / a * a; };
    char* p = static_cast<char*>(scratch);
    void* cub_tmp = p;                                         p += up(cub_bytes);
    uint64_t* keys_in = reinterpret_cast<uint64_t*>(p);        p += up(sizeof(uint64_t) * n);
    uint64_t* keys = reinterpret_cast<uint64_t*>(p);           p += up(sizeof(uint64_t) * n);
    int32_t* parent_inner = reinterpret_cast<int32_t*>(p);     p += up(sizeof(int32_t) * n);
    int32_t* parent_leaf = reinterpret_cast<int32_t*>(p);      p += up(sizeof(int32_t) * n);
    uint32_t* arrived = reinterpret_cast<uint32_t*>(p);        p += up(sizeof(uint32_t) * n);
    Box* boxes = reinterpret_cast<Box*>(p);

    float3 flo = make_float3(lo[0], lo[1], lo[2]), inv;
    inv.x = hi[0] > lo[0] ? 1023.999f / (hi[0] - lo[0]) : 0.f;
    inv.y = hi[1] > lo[1] ? 1023.999f / (hi[1] - lo[1]) : 0.f;
    inv.z = hi[2] > lo[2] ? 1023.999f / (hi[2] - lo[2]) : 0.f;
    const unsigned g = (n + kB - 1) / kB;
    lbvh_keys<<<g, kB, 0, stream>>>(d_leaves, n, flo, inv, keys_in);
    cudaError_t e = cub::DeviceRadixSort::SortKeys(cub_tmp, cub_bytes, keys_in, keys, (int)n, 0, 62, stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(arrived, 0, sizeof(uint32_t) * n, stream);
    if (e != cudaSuccess) return e;
    lbvh_topology<<<g, kB, 0, stream>>>(keys, (int)n, d_nodes, parent_inner, parent_leaf);
    lbvh_refit<<<g, kB, 0, stream>>>(keys, d_leaves, d_cull, (int)n, omax, d_nodes, parent_inner, parent_leaf, arrived, boxes, d_leaf_obj,
                                     d_leaf_cull);
    return cudaGetLastError();
}

}  // namespace rt
