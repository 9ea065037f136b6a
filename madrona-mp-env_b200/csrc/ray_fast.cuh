// ray_fast.cuh -- warp-pooled closest-hit ray engine for the ray-heavy kernels
// (k_lidar, k_los).  Device only.
//
// Same results as mp_geom.h's bvhTraceRay / traceRayWorld (reference
// mesh_bvh.inl:110-208, 360-554; utils.cpp:10-72), bit for bit: every float
// expression is the reference's, and each ray visits nodes, children and
// triangles in the reference's ORDER (LIFO stack, children 0..3, leaves tested
// inline) -- the order decides which of two coplanar triangles' 1-ulp-apart t
// values wins.  What changes is the execution shape:
//
//  * the BVH is re-encoded once per CTA into shared memory as FastNode /
//    FastTri: quantised bounds pre-converted to float (exact), the 2^exp scales
//    materialised, one 32-byte record per child, triangles padded to 48 bytes:
//    traversal is LDS.128 only -- no generic loads, no I2F, no byte extracts;
//  * a node visit computes all four child slab tests in one branch-free block.
//    The reference's test  t_near <= min(t_far_xyz, t_max)  is split into
//    (t_near <= t_far_xyz), evaluated at the visit, and (t_near <= t_max),
//    re-evaluated when the child's turn comes (t_max may have shrunk since);
//    for non-NaN operands the two forms are identical;
//  * rays are pulled from a per-warp pool in shared memory: a lane that
//    finishes its ray fetches the next one, so a warp's 32 lanes stay busy
//    until the pool drains instead of idling behind the longest ray;
//  * per-ray constants (reciprocals, Woop shear, axis permutation) are computed
//    once per unique direction by converged lanes and parked in the pool.
#pragma once

#include <cuda_runtime.h>

#include "mp_geom.h"

namespace mpenv {

struct FastChild {
    float qMinX, qMinY, qMinZ, qMaxX, qMaxY, qMaxZ;
    int32_t child;          // kBVHSentinel, leaf (bit 31) or node index
    int32_t triSize;
};
struct __align__(16) FastNode {
    float minX, minY, minZ, sX;
    float sY, sZ;
    uint32_t leafMask;      // bit i: child i is a leaf
    uint32_t triSizes;      // triSize of child i in byte i
    FastChild c[4];
};
static_assert(sizeof(FastNode) == 160, "FastNode is 10 x 16 bytes");
struct __align__(16) FastTri { float v[12]; };   // a.xyz b.xyz c.xyz + pad

struct FastBVH {
    const FastNode *nodes;
    const FastTri *tris;
};

__host__ __device__ inline size_t fastBVHBytes(uint32_t num_nodes, uint32_t num_tris)
{
    return (size_t)num_nodes * sizeof(FastNode) + (size_t)(num_tris + 1) * sizeof(FastTri);
}

// Re-encodes the reference-format BVH (global memory) into `smem`.
__device__ __forceinline__ FastBVH stageFastBVH(const BVHView &g, void *smem)
{
    FastNode *sn = reinterpret_cast<FastNode *>(smem);
    FastTri *st = reinterpret_cast<FastTri *>(sn + g.numNodes);
    for (uint32_t i = threadIdx.x; i < g.numNodes * 4u; i += blockDim.x) {
        const uint32_t n = i >> 2, k = i & 3u;
        const BVHNode &src = g.nodes[n];
        FastChild c;
        c.qMinX = (float)src.qMinX[k]; c.qMinY = (float)src.qMinY[k]; c.qMinZ = (float)src.qMinZ[k];
        c.qMaxX = (float)src.qMaxX[k]; c.qMaxY = (float)src.qMaxY[k]; c.qMaxZ = (float)src.qMaxZ[k];
        c.child = src.children[k];
        c.triSize = src.triSize[k];
        if (src.children[k] == kBVHSentinel) {
            // absent child: NaN bounds make the slab test fail by itself (fminf / fmaxf drop the
            // NaN operand, so t_near = 0 and t_far = NaN, and 0 <= NaN is false) -- no sentinel
            // compare in the visit
            const float qnan = __int_as_float(0x7FC00000);
            c.qMinX = c.qMinY = c.qMinZ = c.qMaxX = c.qMaxY = c.qMaxZ = qnan;
        }
        sn[n].c[k] = c;
        if (k == 0) {
            sn[n].minX = src.minX; sn[n].minY = src.minY; sn[n].minZ = src.minZ;
            sn[n].sX = u2f((uint32_t)((int32_t)src.expX + 127) << 23);
            sn[n].sY = u2f((uint32_t)((int32_t)src.expY + 127) << 23);
            sn[n].sZ = u2f((uint32_t)((int32_t)src.expZ + 127) << 23);
            uint32_t lm = 0, ts = 0;
            for (int q = 0; q < 4; q++) {
                if (src.children[q] != kBVHSentinel && (src.children[q] & 0x80000000)) lm |= 1u << q;
                ts |= (uint32_t)src.triSize[q] << (8 * q);
            }
            sn[n].leafMask = lm; sn[n].triSizes = ts;
        }
    }
    const float *gv = reinterpret_cast<const float *>(g.verts);
    for (uint32_t i = threadIdx.x; i < (g.numTris + 1u) * 12u; i += blockDim.x) {
        const uint32_t t = i / 12u, k = i - t * 12u;
        st[t].v[k] = k < 9u ? gv[t * 9u + k] : 0.f;
    }
    __syncthreads();
    FastBVH f;
    f.nodes = sn;
    f.tris = st;
    return f;
}

// Per-direction constants parked in the pool (48 bytes).
struct __align__(16) RayDirRec {
    float dx, dy, dz, invX;      // invX/Y/Z: the slab-test reciprocals (mesh_bvh.inl:125-135)
    float invY, invZ, Sx, Sy;    // Woop shear (mesh_bvh.inl:584-741)
    float Sz;
    int32_t kperm;               // kx | ky << 2 | kz << 4
    float a2;                    // dx*dx + dy*dy        (capsule side quadratic)
    float dd;                    // a2 + dz*dz           (capsule cap quadratic)
};
static_assert(sizeof(RayDirRec) == 48, "RayDirRec is 3 x 16 bytes");

__device__ __forceinline__ RayDirRec makeRayDirRec(Vec3 d)
{
    constexpr float diveps = 0.0000001f;
    const Vec3 inv_d = v3(1.f / d.x, 1.f / d.y, 1.f / d.z);
    const RayTriTxfm tx = computeRayTriTxfm(d, inv_d);
    RayDirRec r;
    r.dx = d.x; r.dy = d.y; r.dz = d.z;
    r.invX = copysignf(d.x == 0 ? 1 / diveps : 1 / d.x, d.x);
    r.invY = copysignf(d.y == 0 ? 1 / diveps : 1 / d.y, d.y);
    r.invZ = copysignf(d.z == 0 ? 1 / diveps : 1 / d.z, d.z);
    r.Sx = tx.Sx; r.Sy = tx.Sy; r.Sz = tx.Sz;
    r.kperm = tx.kx | (tx.ky << 2) | (tx.kz << 4);
    r.a2 = d.x * d.x + d.y * d.y;
    r.dd = r.a2 + d.z * d.z;
    return r;
}

// The ray a lane is currently traversing.
struct RayRegs {
    float ox, oy, oz;
    float invX, invY, invZ;
    float Sx, Sy, Sz;
    float okx, oky, okz;         // origin permuted by (kx, ky, kz)
    int32_t kx, ky, kz;
    float tStop;                 // any-hit threshold: the ray ends at an accepted hit with t <= tStop
};

// x, y or z by index, as two SELs (a ?: chain compiles to divergent branches here).
__device__ __forceinline__ float sel3(int32_t k, float x, float y, float z)
{
    float r;
    asm("{\n\t.reg .pred p0, p1;\n\t"
        "setp.eq.s32 p0, %1, 0;\n\t"
        "setp.eq.s32 p1, %1, 1;\n\t"
        "selp.f32 %0, %3, %4, p1;\n\t"
        "selp.f32 %0, %2, %0, p0;\n\t}"
        : "=&f"(r) : "r"(k), "f"(x), "f"(y), "f"(z));
    return r;
}

__device__ __forceinline__ void setRay(RayRegs &r, float ox, float oy, float oz, const RayDirRec &d)
{
    r.ox = ox; r.oy = oy; r.oz = oz;
    r.invX = d.invX; r.invY = d.invY; r.invZ = d.invZ;
    r.Sx = d.Sx; r.Sy = d.Sy; r.Sz = d.Sz;
    r.kx = d.kperm & 3; r.ky = (d.kperm >> 2) & 3; r.kz = (d.kperm >> 4) & 3;
    r.okx = sel3(r.kx, ox, oy, oz);
    r.oky = sel3(r.ky, ox, oy, oz);
    r.okz = sel3(r.kz, ox, oy, oz);
}

// rayTriangle of mp_geom.h (mesh_bvh.inl:433-554) on a FastTri; identical
// arithmetic: comp(tri - org, k) == comp(tri, k) - comp(org, k).
__device__ __forceinline__ bool rayTriangleFast(const FastTri *tri, const RayRegs &r, float t_max, float *out_t)
{
    // the permuted components come straight from shared memory by index (9 LDS.32) --
    // 3 LDS.128 + nine two-select sel3 chains cost 27 more issue slots per triangle
    const float *tk = tri->v;
    const float *px = tk + r.kx, *py = tk + r.ky, *pz = tk + r.kz;
    const float Akx = px[0] - r.okx, Aky = py[0] - r.oky, Akz = pz[0] - r.okz;
    const float Bkx = px[3] - r.okx, Bky = py[3] - r.oky, Bkz = pz[3] - r.okz;
    const float Ckx = px[6] - r.okx, Cky = py[6] - r.oky, Ckz = pz[6] - r.okz;

    const float Ax = fmaf(-r.Sx, Akz, Akx);
    const float Ay = fmaf(-r.Sy, Akz, Aky);
    const float Bx = fmaf(-r.Sx, Bkz, Bkx);
    const float By = fmaf(-r.Sy, Bkz, Bky);
    const float Cx = fmaf(-r.Sx, Ckz, Ckx);
    const float Cy = fmaf(-r.Sy, Ckz, Cky);

    float U = fmaf(Cx, By, -Cy * Bx);
    float V = fmaf(Ax, Cy, -Ay * Cx);
    float W = fmaf(Bx, Ay, -By * Ax);
    if (U < 0.0f || V < 0.0f || W < 0.0f) return false;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        double CxBy = (double)Cx * (double)By;
        double CyBx = (double)Cy * (double)Bx;
        U = (float)(CxBy - CyBx);
        double AxCy = (double)Ax * (double)Cy;
        double AyCx = (double)Ay * (double)Cx;
        V = (float)(AxCy - AyCx);
        double BxAy = (double)Bx * (double)Ay;
        double ByAx = (double)By * (double)Ax;
        W = (float)(BxAy - ByAx);
        if (U < 0.0f || V < 0.0f || W < 0.0f) return false;
    }
    const float det = U + V + W;
    if (det == 0.f) return false;
    const float Az = r.Sz * Akz, Bz = r.Sz * Bkz, Cz = r.Sz * Ckz;
    const float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
    if (T < 0.0f || T > t_max * det) return false;
    const float rcpDet = 1.0f / det;
    *out_t = T * rcpDet;
    return true;
}

// One child's slab test against the visit-time constants (mesh_bvh.inl:146-177).
// Returns t_near, or +inf when the child is absent or its box is missed.
__device__ __forceinline__ float childNear(const FastChild &c, float dqx, float dqy, float dqz,
                                           float oqx, float oqy, float oqz)
{
    const float t_near_x = c.qMinX * dqx + oqx;
    const float t_near_y = c.qMinY * dqy + oqy;
    const float t_near_z = c.qMinZ * dqz + oqz;
    const float t_far_x = c.qMaxX * dqx + oqx;
    const float t_far_y = c.qMaxY * dqy + oqy;
    const float t_far_z = c.qMaxZ * dqz + oqz;
    const float t_near = fmaxf(fminf(t_near_x, t_far_x), fmaxf(fminf(t_near_y, t_far_y),
        fmaxf(fminf(t_near_z, t_far_z), 0.f)));
    const float t_far3 = fminf(fmaxf(t_far_x, t_near_x), fminf(fmaxf(t_far_y, t_near_y),
        fmaxf(t_far_z, t_near_z)));
    return t_near <= t_far3 ? t_near : __int_as_float(0x7F800000);      // absent children carry NaN bounds
}

// Pool interface (duck-typed):
//   int  count()                                 rays in this warp's pool
//   void load(int ray, RayRegs &r)               fill the lane's registers
//   void finish(int ray, bool hit, float t)      closest mesh hit of a finished ray
//
// Must be entered by all 32 lanes (warp-synchronous: ballots decide the batched
// steps).  Two things are batched so that their code runs with many lanes
// instead of the few that happen to need it in a given iteration:
//   * leaf tests: a lane with a leaf to test parks until kLeafBatch lanes have one;
//   * ray refill: a lane whose ray is finished parks until kRefillBatch lanes are
//     idle; ray indices are handed out by ballot rank (no atomics).
// Either fires at once when no other lane can make progress.
template <int kLeafBatch, int kRefillBatch, class Pool>
__device__ __forceinline__ void traversePool(const FastBVH &bvh, Pool &pool)
{
    const unsigned lane_lt = (1u << (threadIdx.x & 31)) - 1u;
    const int32_t pool_count = pool.count();
    int32_t next_ray = 0;            // warp-uniform
    int32_t stack[32];
    int32_t sp = 0;
    int32_t ray = -1;
    RayRegs r;
    float t_max = FLT_MAX;
    bool hit = false, done = false;
    // the current node: t_near / child / triSize per slot and the slots still to consider
    float tn[4] = {0.f, 0.f, 0.f, 0.f};
    int32_t ch[4] = {0, 0, 0, 0};
    uint32_t tsz = 0;
    uint32_t rem = 0, leafmask = 0;
    int32_t leaf = -1, ntri = 0;

    for (;;) {
        const bool idle = !done && leaf < 0 && rem == 0 && sp == 0;
        const unsigned idle_m = __ballot_sync(0xFFFFFFFFu, idle);
        const unsigned adv_m = __ballot_sync(0xFFFFFFFFu, !done && !idle && leaf < 0);
        const unsigned pend_m = __ballot_sync(0xFFFFFFFFu, leaf >= 0);
        if ((idle_m | adv_m | pend_m) == 0) break;
        if (idle_m != 0 && (__popc(idle_m) >= kRefillBatch || adv_m == 0)) {
            if (idle) {
                if (ray >= 0) pool.finish(ray, hit, t_max);
                ray = next_ray + __popc(idle_m & lane_lt);
                if (ray < pool_count) {
                    pool.load(ray, r);
                    t_max = FLT_MAX;
                    hit = false;
                    stack[0] = 0;
                    sp = 1;
                } else {
                    ray = -1;
                    done = true;
                }
            }
            next_ray += __popc(idle_m);
        }
        if (!done && leaf < 0 && rem == 0 && sp > 0) {
            const int32_t node = stack[--sp];
            const FastNode *nd = bvh.nodes + node;
            const float4 h0 = reinterpret_cast<const float4 *>(nd)[0];    // minX minY minZ sX
            const float4 h1 = reinterpret_cast<const float4 *>(nd)[1];    // sY sZ leafMask triSizes
            const float dqx = h0.w * r.invX, dqy = h1.x * r.invY, dqz = h1.y * r.invZ;
            const float oqx = (h0.x - r.ox) * r.invX;
            const float oqy = (h0.y - r.oy) * r.invY;
            const float oqz = (h0.z - r.oz) * r.invZ;
            leafmask = __float_as_uint(h1.z);
            tsz = __float_as_uint(h1.w);
            rem = 0xFu;                  // absent / missed children carry t_near = +inf
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const FastChild k = nd->c[i];
                tn[i] = childNear(k, dqx, dqy, dqz, oqx, oqy, oqz);
                ch[i] = k.child;
            }
        }
        // Children in order until a leaf needs testing; internal nodes are pushed.
        // t_near <= t_max is evaluated at the child's turn, as the reference does:
        // t_max is constant during one scan (only triangle tests change it), so the
        // four comparisons are taken at once and the in-order walk becomes bit logic.
        if (leaf < 0 && rem != 0) {
            uint32_t hits = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) hits |= (tn[i] <= t_max ? 1u : 0u) << i;
            hits &= rem;
            const uint32_t leaf_hits = hits & leafmask;
            const int32_t first = leaf_hits ? (__ffs(leaf_hits) - 1) : 4;       // slot of the first leaf hit
            const uint32_t below = (1u << first) - 1u;
            const uint32_t push = hits & ~leafmask & below;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if ((push >> i) & 1u) stack[sp++] = ch[i];
            }
            if (first < 4) {
                const int32_t c = first == 0 ? ch[0] : first == 1 ? ch[1] : first == 2 ? ch[2] : ch[3];
                leaf = c & 0x7FFFFFFF;
                ntri = (int32_t)((tsz >> (8 * first)) & 0xFFu);
                rem &= ~(below | (1u << first));
            } else {
                rem = 0;
            }
        }
        const unsigned pend = __ballot_sync(0xFFFFFFFFu, leaf >= 0);
        const unsigned can_advance = __ballot_sync(0xFFFFFFFFu, !done && leaf < 0 && (rem != 0 || sp > 0));
        if (leaf >= 0 && (__popc(pend) >= kLeafBatch || can_advance == 0)) {
            // traceRayLeaf (mesh_bvh.inl:360-431): a later triangle replaces an
            // earlier one whenever T <= t_cur * det.
            for (int32_t k = 0; k < ntri; k++) {
                float t;
                if (rayTriangleFast(bvh.tris + leaf + k, r, t_max, &t)) {
                    hit = true;
                    t_max = t;
                }
            }
            leaf = -1;
            if (hit && t_max <= r.tStop) { sp = 0; rem = 0; }     // occluder found: this ray is done
        }
    }
}

// ---- agent capsules (utils.cpp:39-61 + rayZCapsule of mp_geom.h) -------------
// Everything of the capsule test that depends on (ray origin, agent) only.
struct __align__(16) CapsuleRec {
    float ox, oy, oz, oz1;       // origin relative to the capsule's lower sphere centre; oz1 = oz - h
    float cCyl, cCap0, cCap1;
    int32_t inside;              // origin inside the capsule: the test returns "no hit"
};
static_assert(sizeof(CapsuleRec) == 32, "CapsuleRec is 2 x 16 bytes");

__device__ __forceinline__ CapsuleRec makeCapsuleRec(Vec3 ray_o, Vec3 agent_pos)
{
    const float r = kAgentRadius, h = kStandHeight - 2.f * kAgentRadius;
    Vec3 capsule_origin = agent_pos;
    capsule_origin.z += kAgentRadius;
    const Vec3 o = ray_o - capsule_origin;
    const float r2 = r * r;
    CapsuleRec c;
    c.ox = o.x; c.oy = o.y; c.oz = o.z; c.oz1 = o.z - h;
    const float zc = fminf(fmaxf(o.z, 0.f), h);
    const float dz = o.z - zc;
    const float dist2 = o.x * o.x + o.y * o.y + dz * dz;
    c.inside = dist2 <= r2 ? 1 : 0;
    c.cCyl = o.x * o.x + o.y * o.y - r2;
    c.cCap0 = o.x * o.x + o.y * o.y + o.z * o.z - r2;
    c.cCap1 = o.x * o.x + o.y * o.y + c.oz1 * c.oz1 - r2;
    return c;
}

// rayZCapsule(o, d, 15, 35) with the origin-only terms taken from `c`.  The three
// discriminants are computed branch-free; the square roots / divisions run only
// for a ray that can actually touch this capsule.
__device__ __forceinline__ float rayCapsuleRec(const CapsuleRec &c, float dx, float dy, float dz,
                                               float a2, float dd)
{
    const float h = kStandHeight - 2.f * kAgentRadius;
    const float bxy = c.ox * dx + c.oy * dy;
    const float discC = bxy * bxy - a2 * c.cCyl;
    const float b0 = bxy + c.oz * dz;
    const float disc0 = b0 * b0 - dd * c.cCap0;
    const float b1 = bxy + c.oz1 * dz;
    const float disc1 = b1 * b1 - dd * c.cCap1;
    const bool side = a2 > 0.f && discC >= 0.f;
    if (c.inside || !(side || disc0 >= 0.f || disc1 >= 0.f)) return 0.f;
    float t_hit = FLT_MAX;
    if (side) {
        const float t = (-bxy - sqrtf(discC)) / a2;
        const float z = c.oz + t * dz;
        if (t > 0.f && z >= 0.f && z <= h) t_hit = t;
    }
    if (disc0 >= 0.f) {
        const float t = (-b0 - sqrtf(disc0)) / dd;
        if (t > 0.f && t < t_hit) t_hit = t;
    }
    if (disc1 >= 0.f) {
        const float t = (-b1 - sqrtf(disc1)) / dd;
        if (t > 0.f && t < t_hit) t_hit = t;
    }
    return t_hit == FLT_MAX ? 0.f : t_hit;
}

// The capsule half of traceRayWorld (utils.cpp:39-61): starts from the mesh
// result and returns the closest hit over the 12 agents of the world.
__device__ __forceinline__ bool closestWithCapsules(const CapsuleRec *caps, bool mesh_hit, float mesh_t,
                                                    float dx, float dy, float dz, float a2, float dd,
                                                    float *t_out, int32_t *agent_out)
{
    float min_hit_t = mesh_hit ? mesh_t : FLT_MAX;
    bool hit = mesh_hit;
    int32_t hit_agent = -1;
#pragma unroll
    for (int i = 0; i < kAgents; i++) {
        const float t = rayCapsuleRec(caps[i], dx, dy, dz, a2, dd);
        if (t != 0 && t < min_hit_t) {
            min_hit_t = t;
            hit = true;
            hit_agent = i;
        }
    }
    *t_out = min_hit_t;
    *agent_out = hit_agent;
    return hit;
}

}  // namespace mpenv
