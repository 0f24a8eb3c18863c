/* @file vcrt_lbvh.cuh

    VCRT_TRAVERSAL_LBVH, part 1: the GPU build of a linear BVH over the sphere list
    (new scope of BASELINE.json configs[3]; the reference has only the linear scan of
    functions.glsl:77-81).
    SPDX-License-Identifier: WTFPL

    Build (all on the device, one stream):
      1. 30-bit Morton code of every sphere centre inside the centre bounding box, made unique by
         appending the sphere's device index: 64-bit key = code << 32 | index;
      2. radix sort of the keys (cub::DeviceRadixSort, a plain library call outside the hot path);
      3. one thread per internal node: Karras' construction (HPG 2012) finds the node's key range
         and split from common-prefix lengths, so the whole hierarchy is built in one pass;
      4. bottom-up refit with one atomic counter per node: the second child to arrive computes the
         parent's box.
    A node keeps the boxes of BOTH children (64 bytes, four 16-byte loads) so one visit decides two
    subtrees.  Leaves are single spheres, stored in sorted order next to their original index.

    Spheres with a radius above kAlwaysRadius (the ground) and a few outliers much larger than the rest are not
    in the tree; every ray tests them first (ensureBvh in vcrt_cabi.cu picks them).
*/
#pragma once
#include "vcrt_shade.cuh"

namespace vcrt {

struct BvhNode {
    float4 a;   // child0 min.xyz, child0 max.x
    float4 b;   // child0 max.yz, child1 min.xy
    float4 c;   // child1 min.z, child1 max.xyz
    int child0, child1;   // >= 0: internal node index; < 0: leaf, ~position in the sorted sphere arrays
    int pad0, pad1;
};

// The traversal reads octant copies of the node array: copy o (bit a set = direction component a < 0) holds each child's
// NEAR planes where BvhNode has the minimum and its FAR planes where BvhNode has the maximum, so a ray picks its planes once
// (one base pointer) instead of with 12 selects per visit.  In the copies the two children are INTERLEAVED,
//     a = (near0.x, near1.x, near0.y, near1.y)   b = (near0.z, near1.z, far0.x, far1.x)   c = (far0.y, far1.y, far0.z, far1.z)
// so that one packed FFMA2 evaluates a slab plane of both children: 6 per visit instead of 12 FFMA.
struct __align__(32) LeafRecord { float4 shape; uint32_t device, orig, pad0, pad1; };

struct BvhBuildArgs {
    const float4* shape;        // device-order (cx, cy, cz, r)
    const uint32_t* origIndex;  // device index -> index in the caller's array
    const uint32_t* treeIndex;  // [count] device indices of the spheres that go into the tree (ascending)
    uint32_t count;
    float3 boxMin, boxInvExtent;   // centre bounding box -> [0,1)^3
    unsigned long long* keys;      // [count]
    float4* leafShape;             // [count] sorted (cx, cy, cz, r)
    uint32_t* leafOrig;            // [count] sorted original index
    uint32_t* leafDevice;          // [count] sorted device index (surf / kind lookup)
    struct LeafRecord* leafRecord; // [count] the three above in one 32-byte sector: what a leaf visit of the traversal loads
    BvhNode* nodes;                // [count-1]
    BvhNode* octNodes;             // [8][count-1], see above
    int* parent;                   // [2*count-1]: internal nodes 0..count-2, then leaves
    float* bounds;                 // [2*count-1][6] min.xyz max.xyz of every node / leaf
    unsigned int* arrivals;        // [count-1]
};

VCRT_DEV uint32_t expand_bits10(uint32_t v) {   // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void lbvh_keys_kernel(BvhBuildArgs g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.count) return;
    const uint32_t device = g.treeIndex[i];
    float4 s = g.shape[device];
    float x = fminf(fmaxf((s.x - g.boxMin.x) * g.boxInvExtent.x * 1024.f, 0.f), 1023.f);
    float y = fminf(fmaxf((s.y - g.boxMin.y) * g.boxInvExtent.y * 1024.f, 0.f), 1023.f);
    float z = fminf(fmaxf((s.z - g.boxMin.z) * g.boxInvExtent.z * 1024.f, 0.f), 1023.f);
    uint32_t code = (expand_bits10((uint32_t)x) << 2) | (expand_bits10((uint32_t)y) << 1) | expand_bits10((uint32_t)z);
    g.keys[i] = ((unsigned long long)code << 32) | device;
}

// Leaf boxes: centre -+ radius, rounded outward.  (The per-ray margin that makes the box test conservative against
// the exact test's rounding is added at traversal time, see vcrt_kernel_lbvh.cuh.)
__global__ void lbvh_leaves_kernel(BvhBuildArgs g) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.count) return;
    uint32_t j = (uint32_t)(g.keys[i] & 0xffffffffull);
    float4 s = g.shape[j];
    g.leafShape[i] = s;
    g.leafOrig[i] = g.origIndex[j];
    g.leafDevice[i] = j;
    LeafRecord rec; rec.shape = s; rec.device = j; rec.orig = g.origIndex[j]; rec.pad0 = rec.pad1 = 0u;
    g.leafRecord[i] = rec;
    float r = fabsf(s.w);
    float* bb = g.bounds + 6 * (size_t)(g.count - 1 + i);
    bb[0] = __fsub_rd(s.x, r); bb[1] = __fsub_rd(s.y, r); bb[2] = __fsub_rd(s.z, r);
    bb[3] = __fadd_ru(s.x, r); bb[4] = __fadd_ru(s.y, r); bb[5] = __fadd_ru(s.z, r);
}

VCRT_DEV int lbvh_delta(const unsigned long long* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(keys[i] ^ keys[j]));   // keys are unique, so the xor is never 0
}

__global__ void lbvh_hierarchy_kernel(BvhBuildArgs g) {
    const int n = (int)g.count;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const unsigned long long* keys = g.keys;
    // direction of the range
    const int d = lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int deltaMin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > deltaMin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > deltaMin) l += t;
    const int j = i + l * d;
    const int deltaNode = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > deltaNode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? ~gamma : gamma;             // leaf ids are ~position
    const int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    g.nodes[i].child0 = left; g.nodes[i].child1 = right; g.nodes[i].pad0 = lo; g.nodes[i].pad1 = hi;
    g.parent[left >= 0 ? left : (n - 1 + ~left)] = i;
    g.parent[right >= 0 ? right : (n - 1 + ~right)] = i;
    if (i == 0) g.parent[0] = -1;
}

__global__ void lbvh_refit_kernel(BvhBuildArgs g) {
    const int n = (int)g.count;
    int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n) return;
    int node = g.parent[n - 1 + leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&g.arrivals[node], 1u) == 0u) return;   // the first child to arrive stops; the second continues upward
        BvhNode& nd = g.nodes[node];
        const float* b0 = g.bounds + 6 * (size_t)(nd.child0 >= 0 ? nd.child0 : (n - 1 + ~nd.child0));
        const float* b1 = g.bounds + 6 * (size_t)(nd.child1 >= 0 ? nd.child1 : (n - 1 + ~nd.child1));
        float v0[6], v1[6];
        for (int k = 0; k < 6; k++) { v0[k] = __ldcg(b0 + k); v1[k] = __ldcg(b1 + k); }
        nd.a = make_float4(v0[0], v0[1], v0[2], v0[3]);
        nd.b = make_float4(v0[4], v0[5], v1[0], v1[1]);
        nd.c = make_float4(v1[2], v1[3], v1[4], v1[5]);
        for (int o = 0; o < 8; o++) {
            float w0[6], w1[6];
            for (int a = 0; a < 3; a++) {
                const bool neg = (o >> a) & 1;
                w0[a] = neg ? v0[3 + a] : v0[a]; w0[3 + a] = neg ? v0[a] : v0[3 + a];
                w1[a] = neg ? v1[3 + a] : v1[a]; w1[3 + a] = neg ? v1[a] : v1[3 + a];
            }
            BvhNode& on = g.octNodes[(size_t)o * (size_t)(n - 1) + node];
            on.a = make_float4(w0[0], w1[0], w0[1], w1[1]);
            on.b = make_float4(w0[2], w1[2], w0[3], w1[3]);
            on.c = make_float4(w0[4], w1[4], w0[5], w1[5]);
            on.child0 = nd.child0; on.child1 = nd.child1; on.pad0 = 0; on.pad1 = 0;
        }
        float* bb = g.bounds + 6 * (size_t)node;
        for (int k = 0; k < 3; k++) { __stcg(bb + k, fminf(v0[k], v1[k])); __stcg(bb + 3 + k, fmaxf(v0[3 + k], v1[3 + k])); }
        node = g.parent[node];
    }
}

}  // namespace vcrt
