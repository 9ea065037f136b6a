// sphere_fast.cuh -- the swept-sphere cast of the character controller
// (mesh_bvh.inl:743-883, called from applyVelocitySystem / fallSystem) on a
// node format made for it.
//
// Every cast of the step sweeps the SAME sphere (radius kAgentRadius), so the
// per-child work the reference repeats at every visit
//     bmin = node.min + 2^exp * q_min - r        bmax = node.min + 2^exp * q_max + r
// (two int8 -> float conversions, a multiply and two adds per bound, 24 bounds
// per node: ~340 instructions per node visit, 45 % of k_move) is done ONCE when
// the BVH is staged into shared memory, with the same operations in the same
// order -- the floats are the reference's, bit for bit.  A visit is then
//     (b - o) * inv_d
// on 24 precomputed bounds.  As in ray_fast.cuh the bounds are stored as
// (near, far) pairs once per sign of the direction component, so the
// reference's sign test (mesh_bvh.inl:832-840) is an address, not a select.
//
// The reference's running   t_min = i_min > t_min ? i_min : t_min   and
// t_max = i_max < t_max ? i_max : t_max   keep the accumulator when the new
// term is NaN (0 * inf: a box face through the origin of an axis-parallel
// cast); fmaxf / fminf drop a NaN operand the same way, and the accumulators
// start from 0 and the finite running hit distance, so they are never NaN
// themselves.  Only the sign of a zero can differ, which `t_min < t_max`
// cannot see.
#pragma once

#include "mp_geom.h"

namespace mpenv {

struct __align__(16) SphereNode {
    int32_t child[4];          // kBVHSentinel, leaf (bit 31) or node index
    float2 b[3][2][4];         // [axis][direction sign][child] = (near, far) bound, expanded by the radius
};
static_assert(sizeof(SphereNode) == 208, "SphereNode is 13 x 16 bytes");

struct SphereBVH {
    const SphereNode *nodes;
    const Vec3 *verts;         // 3 per triangle, de-indexed, leaf order (+ one pad triangle)
};

__host__ __device__ inline size_t sphereBVHBytes(uint32_t num_nodes, uint32_t vert_words)
{
    return (size_t)num_nodes * sizeof(SphereNode) + (((size_t)vert_words * 4 + 15) / 16) * 16;
}

// Re-encodes the reference-format BVH into the image at `dst` (SphereNode[numNodes], then the vertices)
// for casts of radius `sphere_r`; the threads of ONE CTA.  Done once per simulator
// (k_encode_sphere_bvh); k_move copies the image into its shared memory.
__device__ __forceinline__ void encodeSphereBVH(const BVHView &g, uint32_t vert_words, float sphere_r, void *dst)
{
    SphereNode *sn = reinterpret_cast<SphereNode *>(dst);
    uint32_t *sv = reinterpret_cast<uint32_t *>(sn + g.numNodes);
    for (uint32_t i = threadIdx.x; i < g.numNodes * 4u; i += blockDim.x) {
        const uint32_t n = i >> 2, k = i & 3u;
        const BVHNode &node = g.nodes[n];
        const float s[3] = { u2f(((uint32_t)(int32_t)node.expX + 127u) << 23),
                             u2f(((uint32_t)(int32_t)node.expY + 127u) << 23),
                             u2f(((uint32_t)(int32_t)node.expZ + 127u) << 23) };
        const float mn[3] = { node.minX, node.minY, node.minZ };
        const float qlo[3] = { (float)node.qMinX[k], (float)node.qMinY[k], (float)node.qMinZ[k] };
        const float qhi[3] = { (float)node.qMaxX[k], (float)node.qMaxY[k], (float)node.qMaxZ[k] };
        for (int ax = 0; ax < 3; ax++) {
            // mesh_bvh.inl:775-786 then :822-827
            float lo = mn[ax] + s[ax] * qlo[ax];
            float hi = mn[ax] + s[ax] * qhi[ax];
            lo -= sphere_r;
            hi += sphere_r;
            sn[n].b[ax][0][k] = make_float2(lo, hi);      // inv_d >= +0: enters at bmin
            sn[n].b[ax][1][k] = make_float2(hi, lo);      // sign bit set: enters at bmax
        }
        sn[n].child[k] = node.children[k];
    }
    const uint32_t *gv = reinterpret_cast<const uint32_t *>(g.verts);
    for (uint32_t i = threadIdx.x; i < vert_words; i += blockDim.x) sv[i] = __ldg(gv + i);
    // the image is copied in 16-byte units: define the tail
    const uint32_t padded = (uint32_t)((((size_t)vert_words * 4 + 15) / 16) * 4);
    for (uint32_t i = vert_words + threadIdx.x; i < padded; i += blockDim.x) sv[i] = 0;
}

__global__ void __launch_bounds__(256) k_encode_sphere_bvh(BVHView g, uint32_t vert_words, float sphere_r, void *dst)
{
    encodeSphereBVH(g, vert_words, sphere_r, dst);
}

// Copies the SphereBVH image (global memory, 16-byte units) into `smem`.
__device__ __forceinline__ SphereBVH stageSphereBVH(const uint4 *image, uint32_t num_nodes, uint32_t vert_words, void *smem)
{
    const uint32_t n16 = (uint32_t)(((size_t)num_nodes * sizeof(SphereNode) + (((size_t)vert_words * 4 + 15) / 16) * 16) / 16);
    uint4 *d = reinterpret_cast<uint4 *>(smem);
    for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) d[i] = __ldg(image + i);
    __syncthreads();
    SphereBVH f;
    f.nodes = reinterpret_cast<const SphereNode *>(smem);
    f.verts = reinterpret_cast<const Vec3 *>(f.nodes + num_nodes);
    return f;
}

// The grown box of child i (the staged bounds are expanded by the radius already): is it entered before
// t_bound, and does the sweep meet the child's own box at all?  Detection only (kSweepSlack, mp_geom.h):
// plain arithmetic, no parity with anything.
__device__ __forceinline__ bool sphereChildGrownHit(const char *nd, int i, uint32_t offX, uint32_t offY, uint32_t offZ,
                                                    Vec3 o, Vec3 inv_d, float t_bound, bool need_own_box)
{
    const float2 bx = *reinterpret_cast<const float2 *>(nd + offX + 8 * i);
    const float2 by = *reinterpret_cast<const float2 *>(nd + offY + 8 * i);
    const float2 bz = *reinterpret_cast<const float2 *>(nd + offZ + 8 * i);
    if (need_own_box) {
        const float tn = fmaxf(fmaxf(fmaxf(0.f, (bx.x - o.x) * inv_d.x), (by.x - o.y) * inv_d.y), (bz.x - o.z) * inv_d.z);
        const float tf = fminf(fminf((bx.y - o.x) * inv_d.x, (by.y - o.y) * inv_d.y), (bz.y - o.z) * inv_d.z);
        if (!(tn < tf)) return false;
    }
    // near moves outwards against the direction of travel, far along it
    const float sx = copysignf(kSweepSlack, inv_d.x), sy = copysignf(kSweepSlack, inv_d.y), sz = copysignf(kSweepSlack, inv_d.z);
    const float tn = fmaxf(fmaxf(fmaxf(0.f, ((bx.x - sx) - o.x) * inv_d.x), ((by.x - sy) - o.y) * inv_d.y), ((bz.x - sz) - o.z) * inv_d.z);
    const float tf = fminf(fminf(((bx.y + sx) - o.x) * inv_d.x, ((by.y + sy) - o.y) * inv_d.y), ((bz.y + sz) - o.z) * inv_d.z);
    return tn < fminf(tf, t_bound);
}

// bvhSphereCast (mp_geom.h) on the staged nodes: same visiting order, same t_bound / risky contract.
// kDetect: the bounded form with detection (two or three of an agent's casts); the plain form -- the
// vertical casts, and the unbounded repeat of a risky one -- carries none of that code.
template <bool kDetect>
__device__ __noinline__ float sphereCastStaged(const SphereBVH &bvh, Vec3 ray_o, Vec3 ray_d, float sphere_r,
                                               Vec3 *out_n, float t_bound, bool *risky)
{
    MP_COUNT_RAYS(3, 1);
    // the staged image lives in shared memory: say so, or this (non-inlined) function reads it through
    // generic loads (LD + R2UR instead of LDS: 2 M more instructions per launch)
    const SphereNode *const nodes = bvh.nodes;
    const Vec3 *const verts = bvh.verts;
    __builtin_assume(__isShared(nodes));
    __builtin_assume(__isShared(verts));
    const Vec3 inv_d = v3(1.f / ray_d.x, 1.f / ray_d.y, 1.f / ray_d.z);
    // byte offset of each axis's (near, far) order inside a node
    const uint32_t offX = (uint32_t)offsetof(SphereNode, b) + (f2u(inv_d.x) >> 31) * 32u;
    const uint32_t offY = (uint32_t)offsetof(SphereNode, b) + 64u + (f2u(inv_d.y) >> 31) * 32u;
    const uint32_t offZ = (uint32_t)offsetof(SphereNode, b) + 128u + (f2u(inv_d.z) >> 31) * 32u;
    constexpr int32_t kDetectOnly = 0x40000000;
    bool flagged = false;            // kDetect: a (mis)reported feature hit below the bound is possible
    int32_t stack[32];
    int32_t stack_size = 0;
    stack[stack_size++] = 0;
    Vec3 closest_n = v3(0.f, 0.f, 0.f);
    const float t_max = t_bound;
    float hit_t = t_max;
    while (stack_size > 0) {
        const int32_t entry = stack[--stack_size];
        const char *nd = reinterpret_cast<const char *>(nodes + (entry & ~kDetectOnly));
        const int4 ch = *reinterpret_cast<const int4 *>(nd);
        const int32_t child[4] = { ch.x, ch.y, ch.z, ch.w };
        if (kDetect && (entry & kDetectOnly)) {
            // a subtree the bound culled: only whether a feature test inside could fire before the bound
            for (int i = 0; i < 4; i++) {
                if (child[i] == kBVHSentinel || flagged) continue;
                if (!sphereChildGrownHit(nd, i, offX, offY, offZ, ray_o, inv_d, t_max, true)) continue;
                if (child[i] & 0x80000000) {
                    const int32_t leaf_idx = child[i] & ~0x80000000;
                    for (int k = 0; k < 2; k++) {
                        const Vec3 *tv = verts + (size_t)(leaf_idx + k) * 3;
                        if (sphereCastFeaturesBefore(tv[0], tv[1], tv[2], ray_o, ray_d, t_max, sphere_r)) flagged = true;
                    }
                } else {
                    stack[stack_size++] = child[i] | kDetectOnly;
                }
            }
            continue;
        }
        float tn[4], tf[4];
        uint32_t grown = 0;          // kDetect: bit i = child i's box, grown by the slack, is entered before the bound
        const float gx = copysignf(kSweepSlack, inv_d.x), gy = copysignf(kSweepSlack, inv_d.y), gz = copysignf(kSweepSlack, inv_d.z);
        {
            const float4 *px = reinterpret_cast<const float4 *>(nd + offX);
            const float4 *py = reinterpret_cast<const float4 *>(nd + offY);
            const float4 *pz = reinterpret_cast<const float4 *>(nd + offZ);
#pragma unroll
            for (int h = 0; h < 2; h++) {          // children 2h, 2h + 1
                const float4 bx = px[h], by = py[h], bz = pz[h];
                // mesh_bvh.inl:842-850, axis order x, y, z
                const float nx0 = (bx.x - ray_o.x) * inv_d.x, fx0 = (bx.y - ray_o.x) * inv_d.x;
                const float nx1 = (bx.z - ray_o.x) * inv_d.x, fx1 = (bx.w - ray_o.x) * inv_d.x;
                const float ny0 = (by.x - ray_o.y) * inv_d.y, fy0 = (by.y - ray_o.y) * inv_d.y;
                const float ny1 = (by.z - ray_o.y) * inv_d.y, fy1 = (by.w - ray_o.y) * inv_d.y;
                const float nz0 = (bz.x - ray_o.z) * inv_d.z, fz0 = (bz.y - ray_o.z) * inv_d.z;
                const float nz1 = (bz.z - ray_o.z) * inv_d.z, fz1 = (bz.w - ray_o.z) * inv_d.z;
                tn[2 * h] = fmaxf(fmaxf(fmaxf(0.f, nx0), ny0), nz0);
                tn[2 * h + 1] = fmaxf(fmaxf(fmaxf(0.f, nx1), ny1), nz1);
                // the far side is folded onto the running hit distance at the child's turn (below)
                tf[2 * h] = fminf(fminf(fx0, fy0), fz0);
                tf[2 * h + 1] = fminf(fminf(fx1, fy1), fz1);
                if (kDetect) {
                    const float g0 = fmaxf(fmaxf(fmaxf(0.f, ((bx.x - gx) - ray_o.x) * inv_d.x), ((by.x - gy) - ray_o.y) * inv_d.y),
                                           ((bz.x - gz) - ray_o.z) * inv_d.z);
                    const float g1 = fmaxf(fmaxf(fmaxf(0.f, ((bx.z - gx) - ray_o.x) * inv_d.x), ((by.z - gy) - ray_o.y) * inv_d.y),
                                           ((bz.z - gz) - ray_o.z) * inv_d.z);
                    grown |= (g0 < t_max ? 1u : 0u) << (2 * h) | (g1 < t_max ? 2u : 0u) << (2 * h);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (child[i] == kBVHSentinel) continue;
            if (!(tn[i] < fminf(tf[i], hit_t))) {
                // culled by the bound alone (the sweep does meet the box, and nothing was hit yet)?
                if (kDetect && ((grown >> i) & 1u) && hit_t == t_max && tn[i] < tf[i] && !flagged) {
                    if (child[i] & 0x80000000) {
                        const int32_t leaf_idx = child[i] & ~0x80000000;
                        for (int k = 0; k < 2; k++) {
                            const Vec3 *tv = verts + (size_t)(leaf_idx + k) * 3;
                            if (sphereCastFeaturesBefore(tv[0], tv[1], tv[2], ray_o, ray_d, t_max, sphere_r)) flagged = true;
                        }
                    } else {
                        stack[stack_size++] = child[i] | kDetectOnly;
                    }
                }
                continue;
            }
            if (child[i] & 0x80000000) {
                const int32_t leaf_idx = child[i] & ~0x80000000;
                Vec3 leaf_n = v3(0.f, 0.f, 0.f);
                float leaf_t = hit_t;
                for (int k = 0; k < 2; k++) {
                    const Vec3 *tv = verts + (size_t)(leaf_idx + k) * 3;
                    const float t = sphereCastTriangle(tv[0], tv[1], tv[2], ray_o, ray_d, leaf_t, sphere_r, &leaf_n, kDetect);
                    if (kDetect && t < 0.f) flagged = true;       // kSweepRisky: leaf_t stays
                    else leaf_t = t;
                }
                if (leaf_t < hit_t) { hit_t = leaf_t; closest_n = leaf_n; }
            } else {
                stack[stack_size++] = child[i];
            }
        }
    }
    if (kDetect && flagged) *risky = true;
    if (hit_t < t_max) *out_n = closest_n;
    return hit_t < t_max ? hit_t : FLT_MAX;
}

__device__ __forceinline__ float bvhSphereCast(const SphereBVH &bvh, Vec3 ray_o, Vec3 ray_d, float sphere_r,
                                               Vec3 *out_n, float t_bound = FLT_MAX, bool *risky = nullptr)
{
    if (risky != nullptr && t_bound < FLT_MAX) return sphereCastStaged<true>(bvh, ray_o, ray_d, sphere_r, out_n, t_bound, risky);
    return sphereCastStaged<false>(bvh, ray_o, ray_d, sphere_r, out_n, t_bound, nullptr);
}

}  // namespace mpenv
