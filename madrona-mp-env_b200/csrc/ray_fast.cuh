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

#include <cstddef>
#include <cuda_runtime.h>

#include "mp_geom.h"

namespace mpenv {

// FastNode (240 bytes): header, children, then the quantised child bounds as
// floats in (near, far) PAIRS, once per sign of the ray direction:
//   b[axis][order][child] = order 0: (qMin, qMax)   order 1: (qMax, qMin)
// A ray with a positive direction component enters a slab at qMin and leaves at
// qMax (negative: the other way round), so the reference's
//   fminf(t_near_x, t_far_x) / fmaxf(t_far_x, t_near_x)        (mesh_bvh.inl:168-171)
// are picked by ADDRESS (b + 32 * order) instead of computed: q * dq + oq is
// monotonic in q for a finite dq of known sign, so min/max of the two results is
// the result of the min/max operand, bit for bit.  (dq is finite and non-zero for
// every ray the engine accepts, see rayIsFast; others take the per-thread path.)
// Each pair sits in an aligned register pair after one LDS.128 (two children), which
// is the operand form of the packed FP32 instructions (FFMA2).
struct __align__(16) FastNode {
    float minX, minY, minZ;
    uint32_t meta;          // bit i: child i is a leaf; bits 4 + 7 i .. : triSize of child i (<= kFastMaxTriSize)
    float sX, sY, sZ;       // 2^exp of each axis
    uint32_t pad;
    int32_t child[4];       // kBVHSentinel, leaf (bit 31) or node index
    float2 b[3][2][4];
};
constexpr uint32_t kFastMaxTriSize = 127;
static_assert(sizeof(FastNode) == 240, "FastNode is 15 x 16 bytes");
struct __align__(16) FastTri { float v[12]; };   // a.xyz b.xyz c.xyz + pad

struct FastBVH {
    const FastNode *nodes;
    const FastTri *tris;
};

__host__ __device__ inline size_t fastBVHBytes(uint32_t num_nodes, uint32_t num_tris)
{
    return (size_t)num_nodes * sizeof(FastNode) + (size_t)(num_tris + 1) * sizeof(FastTri);
}

// Re-encodes the reference-format BVH into the FastBVH image at `dst` (FastNode[numNodes], then
// FastTri[numTris + 1]); the threads of ONE CTA.  Done once per simulator (k_encode_fast_bvh); the
// kernels copy the image into their shared memory.
__device__ __forceinline__ void encodeFastBVH(const BVHView &g, void *dst)
{
    FastNode *sn = reinterpret_cast<FastNode *>(dst);
    FastTri *st = reinterpret_cast<FastTri *>(sn + g.numNodes);
    for (uint32_t i = threadIdx.x; i < g.numNodes * 4u; i += blockDim.x) {
        const uint32_t n = i >> 2, k = i & 3u;
        const BVHNode &src = g.nodes[n];
        float lo[3] = { (float)src.qMinX[k], (float)src.qMinY[k], (float)src.qMinZ[k] };
        float hi[3] = { (float)src.qMaxX[k], (float)src.qMaxY[k], (float)src.qMaxZ[k] };
        if (src.children[k] == kBVHSentinel) {
            // absent child: NaN bounds make the slab test fail by itself (fminf / fmaxf drop the
            // NaN operand, so t_near = 0 and t_far = NaN, and 0 <= NaN is false) -- no sentinel
            // compare in the visit
            const float qnan = __int_as_float(0x7FC00000);
            for (int ax = 0; ax < 3; ax++) lo[ax] = hi[ax] = qnan;
        }
        for (int ax = 0; ax < 3; ax++) {
            sn[n].b[ax][0][k] = make_float2(lo[ax], hi[ax]);
            sn[n].b[ax][1][k] = make_float2(hi[ax], lo[ax]);
        }
        sn[n].child[k] = src.children[k];
        if (k == 0) {
            sn[n].minX = src.minX; sn[n].minY = src.minY; sn[n].minZ = src.minZ;
            sn[n].sX = u2f((uint32_t)((int32_t)src.expX + 127) << 23);
            sn[n].sY = u2f((uint32_t)((int32_t)src.expY + 127) << 23);
            sn[n].sZ = u2f((uint32_t)((int32_t)src.expZ + 127) << 23);
            uint32_t meta = 0;
            for (int q = 0; q < 4; q++) {
                if (src.children[q] != kBVHSentinel && (src.children[q] & 0x80000000)) meta |= 1u << q;
                meta |= ((uint32_t)src.triSize[q] & 0x7Fu) << (4 + 7 * q);
            }
            sn[n].meta = meta; sn[n].pad = 0;
        }
    }
    const float *gv = reinterpret_cast<const float *>(g.verts);
    for (uint32_t i = threadIdx.x; i < (g.numTris + 1u) * 12u; i += blockDim.x) {
        const uint32_t t = i / 12u, k = i - t * 12u;
        st[t].v[k] = k < 9u ? gv[t * 9u + k] : 0.f;
    }
}

__global__ void __launch_bounds__(256) k_encode_fast_bvh(BVHView g, void *dst) { encodeFastBVH(g, dst); }

// Copies the FastBVH image (global memory, 16-byte units) into `smem`.
__device__ __forceinline__ FastBVH stageFastBVH(const uint4 *image, uint32_t num_nodes, uint32_t num_tris, void *smem)
{
    const uint32_t n16 = (uint32_t)(fastBVHBytes(num_nodes, num_tris) / 16);
    uint4 *d = reinterpret_cast<uint4 *>(smem);
    for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) d[i] = __ldg(image + i);
    __syncthreads();
    FastBVH f;
    f.nodes = reinterpret_cast<const FastNode *>(smem);
    f.tris = reinterpret_cast<const FastTri *>(f.nodes + num_nodes);
    return f;
}

// The triangle-test constants of a direction (Woop shear + axis permutation,
// mesh_bvh.inl:584-741).  They stay in the pool: a lane keeps only the record's
// shared-memory offset and fetches them when it has a leaf to test.
struct __align__(16) RayTriRec {
    float Sx, Sy, Sz;
    int32_t kperm;               // kx | ky << 2 | kz << 4
};

// Per-direction constants parked in the pool (48 bytes).
struct __align__(16) RayDirRec {
    float dx, dy, dz, invX;      // invX/Y/Z: the slab-test reciprocals (mesh_bvh.inl:125-135)
    float invY, invZ;
    float a2;                    // dx*dx + dy*dy        (capsule side quadratic)
    float dd;                    // a2 + dz*dz           (capsule cap quadratic)
    RayTriRec tri;
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
    r.tri.Sx = tx.Sx; r.tri.Sy = tx.Sy; r.tri.Sz = tx.Sz;
    r.tri.kperm = tx.kx | (tx.ky << 2) | (tx.kz << 4);
    r.a2 = d.x * d.x + d.y * d.y;
    r.dd = r.a2 + d.z * d.z;
    return r;
}

// What a lane keeps in registers of the ray it is traversing: the node-visit part only.
// Everything else (triangle constants, the node's children / leaf mask) is re-read from
// shared memory at the few places that need it -- the traversal loop holds 28 warps per SM
// at 72 registers.  Shared-memory objects are named by their byte offset from the CTA's
// dynamic shared memory base (one register, and the loads stay LDS).
struct RayRegs {
    float ox, oy, oz;
    float invX, invY, invZ;
    int32_t offX, offY, offZ;    // byte offset of the ray's (near, far) order within a FastNode
    uint32_t tri;                // offset of its RayTriRec
    float tStop;                 // any-hit threshold: the ray ends at an accepted hit with t <= tStop
};

// Picking the reference's per-axis fminf / fmaxf by the sign of the direction needs finite
// reciprocals whose products with the node scales and origins stay finite (no NaN from inf - inf
// or 0 * inf, which fminf / fmaxf would drop): true for any direction component that is 0 (1e7 is
// used, mesh_bvh.inl:125-135) or not absurdly small.  A ray that fails this (none does in practice:
// directions are normalised) is marked offX < 0 and visits nodes with visitNodeRef.
__device__ __forceinline__ bool rayIsFast(float invX, float invY, float invZ)
{
    return fabsf(invX) <= 1e30f && fabsf(invY) <= 1e30f && fabsf(invZ) <= 1e30f;
}

__device__ __forceinline__ void setRay(RayRegs &r, float ox, float oy, float oz, float invX, float invY,
                                       float invZ, uint32_t tri_off)
{
    r.ox = ox; r.oy = oy; r.oz = oz;
    r.invX = invX; r.invY = invY; r.invZ = invZ;
    constexpr int kB = (int)offsetof(FastNode, b);
    r.offX = !rayIsFast(invX, invY, invZ) ? -1 : kB + (invX < 0.f ? 32 : 0);
#ifdef MPENV_FORCE_REF_VISIT
    r.offX = -1;         // test builds: every ray through visitNodeRef
#endif
    r.offY = kB + 64 + (invY < 0.f ? 32 : 0);
    r.offZ = kB + 128 + (invZ < 0.f ? 32 : 0);
    r.tri = tri_off;
}

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

// The per-leaf part of the triangle test: constants of the lane's ray, origin permuted.
struct TriConsts {
    float Sx, Sy, Sz;
    float okx, oky, okz;
    int32_t kx, ky, kz;
};
__device__ __forceinline__ TriConsts loadTriConsts(const char *smem, const RayRegs &r)
{
    const float4 v = *reinterpret_cast<const float4 *>(smem + r.tri);
    const int32_t kperm = __float_as_int(v.w);
    TriConsts c;
    c.Sx = v.x; c.Sy = v.y; c.Sz = v.z;
    c.kx = kperm & 3; c.ky = (kperm >> 2) & 3; c.kz = (kperm >> 4) & 3;
    c.okx = sel3(c.kx, r.ox, r.oy, r.oz);
    c.oky = sel3(c.ky, r.ox, r.oy, r.oz);
    c.okz = sel3(c.kz, r.ox, r.oy, r.oz);
    return c;
}

// rayTriangle of mp_geom.h (mesh_bvh.inl:433-554) on a FastTri; identical
// arithmetic: comp(tri - org, k) == comp(tri, k) - comp(org, k).
__device__ __forceinline__ bool rayTriangleFast(const FastTri *tri, const TriConsts &r, float t_max, float *out_t)
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

// ---- packed FP32 (Blackwell FFMA2: two lanes of data per instruction) ------------
// The slab test is  t = q * dq + oq  with two roundings (the reference is compiled without
// contraction).  ptxas contracts mul.f32x2 + add.f32x2 into one FFMA2 whatever
// the rounding qualifiers and -fmad say, so both steps are written as fma.rn.f32x2:
//   p = fma(q, dq, -0)  == rn(q * dq)          (x + -0 == x for every x, signed zeros included)
//   t = fma(p, ONE, oq) == rn(p + oq)          with ONE = 1.0f read from the kernel arguments,
// a value ptxas cannot see, so it cannot fold the pair back into a single rounding.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 bcast2(float x)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// One node visit (mesh_bvh.inl:142-177): t_near of the four children, +inf for a child that is
// absent or whose box is missed without regard to t_max (the t_near <= t_max part of the
// reference's test is taken at the child's turn, see traversePool).  `nd` = the node in
// shared memory.
__device__ __forceinline__ void visitNode(const char *nd, const RayRegs &r, float one, float tn[4])
{
    const float4 h0 = reinterpret_cast<const float4 *>(nd)[0];    // minX minY minZ meta
    const float4 h1 = reinterpret_cast<const float4 *>(nd)[1];    // sX sY sZ -
#ifndef MPENV_SCALAR_SLAB
    const f32x2 dqx = bcast2(h1.x * r.invX), dqy = bcast2(h1.y * r.invY), dqz = bcast2(h1.z * r.invZ);
    const f32x2 oqx = bcast2((h0.x - r.ox) * r.invX);
    const f32x2 oqy = bcast2((h0.y - r.oy) * r.invY);
    const f32x2 oqz = bcast2((h0.z - r.oz) * r.invZ);
    const f32x2 one2 = bcast2(one), nz2 = 0x8000000080000000ull;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const ulonglong2 qx = *reinterpret_cast<const ulonglong2 *>(nd + r.offX + 16 * half);
        const ulonglong2 qy = *reinterpret_cast<const ulonglong2 *>(nd + r.offY + 16 * half);
        const ulonglong2 qz = *reinterpret_cast<const ulonglong2 *>(nd + r.offZ + 16 * half);
#pragma unroll
        for (int k = 0; k < 2; k++) {
            float nx, fx, ny, fy, nz, fz;
            unpack2(fma2(fma2(k ? qx.y : qx.x, dqx, nz2), one2, oqx), nx, fx);
            unpack2(fma2(fma2(k ? qy.y : qy.x, dqy, nz2), one2, oqy), ny, fy);
            unpack2(fma2(fma2(k ? qz.y : qz.x, dqz, nz2), one2, oqz), nz, fz);
            const float t_near = fmaxf(nx, fmaxf(ny, fmaxf(nz, 0.f)));
            const float t_far3 = fminf(fx, fminf(fy, fz));
            tn[2 * half + k] = t_near <= t_far3 ? t_near : __int_as_float(0x7F800000);
        }
    }
#else
    // scalar FMUL + FADD (measured: k_lidar 1.226 ms against 1.172 ms for the packed form)
    (void)one;
    const float dqx = h1.x * r.invX, dqy = h1.y * r.invY, dqz = h1.z * r.invZ;
    const float oqx = (h0.x - r.ox) * r.invX;
    const float oqy = (h0.y - r.oy) * r.invY;
    const float oqz = (h0.z - r.oz) * r.invZ;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const float4 qx = *reinterpret_cast<const float4 *>(nd + r.offX + 16 * half);   // near far | near far
        const float4 qy = *reinterpret_cast<const float4 *>(nd + r.offY + 16 * half);
        const float4 qz = *reinterpret_cast<const float4 *>(nd + r.offZ + 16 * half);
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const float nx = (k ? qx.z : qx.x) * dqx + oqx, fx = (k ? qx.w : qx.y) * dqx + oqx;
            const float ny = (k ? qy.z : qy.x) * dqy + oqy, fy = (k ? qy.w : qy.y) * dqy + oqy;
            const float nz = (k ? qz.z : qz.x) * dqz + oqz, fz = (k ? qz.w : qz.y) * dqz + oqz;
            const float t_near = fmaxf(nx, fmaxf(ny, fmaxf(nz, 0.f)));
            const float t_far3 = fminf(fx, fminf(fy, fz));
            tn[2 * half + k] = t_near <= t_far3 ? t_near : __int_as_float(0x7F800000);
        }
    }
#endif
}

// The same for a ray that fails rayIsFast: the reference's expressions with both bounds of every
// axis computed and ordered by fminf / fmaxf (NaN operands dropped as the reference drops them).
__device__ __noinline__ float4 visitNodeRef(const char *nd, float ox, float oy, float oz, float invX, float invY,
                                           float invZ)
{
    float tn[4];
    const float4 h0 = reinterpret_cast<const float4 *>(nd)[0];
    const float4 h1 = reinterpret_cast<const float4 *>(nd)[1];
    const float dqx = h1.x * invX, dqy = h1.y * invY, dqz = h1.z * invZ;
    const float oqx = (h0.x - ox) * invX;
    const float oqy = (h0.y - oy) * invY;
    const float oqz = (h0.z - oz) * invZ;
    const float2 *b = reinterpret_cast<const float2 *>(nd + offsetof(FastNode, b));      // [axis][order][child]
    for (int i = 0; i < 4; i++) {
        const float2 qx = b[0 * 8 + i], qy = b[1 * 8 + i], qz = b[2 * 8 + i];             // order 0: (qMin, qMax)
        const float t_near_x = qx.x * dqx + oqx, t_far_x = qx.y * dqx + oqx;
        const float t_near_y = qy.x * dqy + oqy, t_far_y = qy.y * dqy + oqy;
        const float t_near_z = qz.x * dqz + oqz, t_far_z = qz.y * dqz + oqz;
        const float t_near = fmaxf(fminf(t_near_x, t_far_x), fmaxf(fminf(t_near_y, t_far_y),
            fmaxf(fminf(t_near_z, t_far_z), 0.f)));
        const float t_far3 = fminf(fmaxf(t_far_x, t_near_x), fminf(fmaxf(t_far_y, t_near_y),
            fmaxf(t_far_z, t_near_z)));
        tn[i] = t_near <= t_far3 ? t_near : __int_as_float(0x7F800000);
    }
    return make_float4(tn[0], tn[1], tn[2], tn[3]);
}

// The in-order walk over the children of the lane's current node (mesh_bvh.inl:178-200) up to
// the first leaf that needs testing; internal nodes are pushed.  t_near <= t_max is evaluated at
// the child's turn, as the reference does: t_max is constant during one scan (only triangle tests
// change it), so the four comparisons are taken at once and the walk becomes bit logic.
// rem: slots still to consider.  Returns the leaf (first triangle index | triangle count << 24) or -1.
__device__ __forceinline__ int32_t walkChildren(uint32_t meta, const int32_t child[4], const float tn[4], float t_max,
                                                uint32_t &rem, int32_t *stack, int32_t &sp)
{
    const uint32_t leafmask = meta & 0xFu;
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
        if ((push >> i) & 1u) stack[sp++] = child[i];
    }
    if (first < 4) {
        const int32_t ntri = (int32_t)((meta >> (4 + 7 * first)) & 0x7Fu);
        rem &= ~(below | (1u << first));
        return (child[first] & 0xFFFFFF) | (ntri << 24);
    }
    rem = 0;
    return -1;
}

// The same with the node's meta word and children re-read from shared memory (for a loop that has
// no registers to keep them in between the visit and a walk resumed after a leaf test).
__device__ __forceinline__ int32_t walkChildren(const char *nd, const float tn[4], float t_max, uint32_t &rem,
                                                int32_t *stack, int32_t &sp)
{
    return walkChildren(reinterpret_cast<const uint32_t *>(nd)[3], reinterpret_cast<const int32_t *>(nd + 32), tn,
                        t_max, rem, stack, sp);
}

// Pool interface (duck-typed):
//   int  count()                                 rays in this warp's pool
//   float load(int ray, RayRegs &r)              fill the lane's registers; returns the ray's initial t_max
//   void finish(int ray, float t_max)            t_max of a finished ray (closest accepted hit, or the initial value)
//
// Initial t_max.  The reference starts every ray at FLT_MAX.  Starting at a finite bound U gives the
// SAME final t whenever that t is below U: the visiting order is fixed (it does not depend on the
// ray), so both runs see the same triangles in the same order, except those whose boxes the smaller
// t_max culled or whose T <= t_max * det test it failed -- triangles at t >= U (1 - 4 delta), delta =
// the relative rounding slop of those two tests (~1e-6).  By induction over that order, the two
// runs' t_max are either equal or both >= U (1 - 4 delta); the first accepted triangle below that
// level makes them equal for good.  So: final t < U (1 - 4 delta) => identical to the reference;
// otherwise the reference's closest hit (if any) lies at >= U (1 - 4 delta).  A caller that only
// needs "is there a hit at t <= tStop" (line of sight) starts at U = tStop (1 + 3e-4) and reads
// t_max <= tStop: exact, and no box beyond the target is ever visited.
// A ray that fails rayIsFast is traversed by the same loop with visitNodeRef instead of visitNode.
//
// Must be entered by all 32 lanes (warp-synchronous: ballots decide the batched
// steps).  Two things are batched so that their code runs with many lanes
// instead of the few that happen to need it in a given iteration:
//   * leaf tests: a lane with a leaf to test parks until kLeafBatch lanes have one;
//   * ray refill: a lane whose ray is finished parks until kRefillBatch lanes are
//     idle; ray indices are handed out by ballot rank (no atomics).
// Either fires at once when no other lane can make progress.
// smem = the CTA's dynamic shared memory, which starts with the FastBVH.
template <int kLeafBatch, int kRefillBatch, class Pool>
__device__ __forceinline__ void traversePool(const char *smem, const FastBVH &bvh, Pool &pool, float one)
{
    const unsigned lane_lt = (1u << (threadIdx.x & 31)) - 1u;
    const int32_t pool_count = pool.count();
    int32_t next_ray = 0;            // warp-uniform
    int32_t stack[32];
    int32_t sp = 0;
    int32_t ray = -1;
    RayRegs r;
    float t_max = FLT_MAX;
    bool done = false;
    // the current node: meta word, children, t_near per slot and the slots still to consider (kept
    // in registers: the callers of this loop run at 80+ registers)
    uint32_t meta = 0;
    int32_t ch[4] = {0, 0, 0, 0};
    float tn[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t rem = 0;
    int32_t leaf = -1;               // leaf waiting for its triangle tests: first triangle | count << 24

    for (;;) {
        const bool idle = !done && leaf < 0 && rem == 0 && sp == 0;
        const unsigned idle_m = __ballot_sync(0xFFFFFFFFu, idle);
        const unsigned adv_m = __ballot_sync(0xFFFFFFFFu, !done && !idle && leaf < 0);
        const unsigned pend_m = __ballot_sync(0xFFFFFFFFu, leaf >= 0);
        if ((idle_m | adv_m | pend_m) == 0) break;
        if (idle_m != 0 && (__popc(idle_m) >= kRefillBatch || adv_m == 0)) {
            if (idle) {
                if (ray >= 0) pool.finish(ray, t_max);
                ray = next_ray + __popc(idle_m & lane_lt);
                if (ray < pool_count) {
                    t_max = pool.load(ray, r);
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
            const char *nd = smem + (uint32_t)stack[--sp] * (uint32_t)sizeof(FastNode);
            if (r.offX >= 0) {
                visitNode(nd, r, one, tn);
            } else {
                const float4 v = visitNodeRef(nd, r.ox, r.oy, r.oz, r.invX, r.invY, r.invZ);
                tn[0] = v.x; tn[1] = v.y; tn[2] = v.z; tn[3] = v.w;
            }
            meta = reinterpret_cast<const uint32_t *>(nd)[3];
            const int4 cc = reinterpret_cast<const int4 *>(nd)[2];
            ch[0] = cc.x; ch[1] = cc.y; ch[2] = cc.z; ch[3] = cc.w;
            rem = 0xFu;                  // absent / missed children carry t_near = +inf
        }
        if (leaf < 0 && rem != 0) leaf = walkChildren(meta, ch, tn, t_max, rem, stack, sp);
        const unsigned pend = __ballot_sync(0xFFFFFFFFu, leaf >= 0);
        const unsigned can_advance = __ballot_sync(0xFFFFFFFFu, !done && leaf < 0 && (rem != 0 || sp > 0));
        if (leaf >= 0 && (__popc(pend) >= kLeafBatch || can_advance == 0)) {
            // traceRayLeaf (mesh_bvh.inl:360-431): a later triangle replaces an
            // earlier one whenever T <= t_cur * det.
            const TriConsts tc = loadTriConsts(smem, r);
            const FastTri *tri = bvh.tris + (leaf & 0xFFFFFF);
            for (int32_t k = 0; k < (leaf >> 24); k++) {
                float t;
                if (rayTriangleFast(tri + k, tc, t_max, &t)) t_max = t;
            }
            leaf = -1;
            if (t_max <= r.tStop) { sp = 0; rem = 0; }     // occluder found: this ray is done
        }
    }
}

// One ray, one thread: the reference traversal (mesh_bvh.inl:110-208) on the FastBVH, for the few
// rays that do not come in pools or packets (the hitscan of fireSystem: 0.08 per agent-step).
// Same visiting order as bvhTraceRay -- children in slot order, leaves inline, internal children
// pushed and popped LIFO -- at a quarter of its instructions per node visit.
struct FastMeshTrace {
    const char *smem;            // the CTA's dynamic shared memory, which starts with the FastBVH
    FastBVH bvh;
    float one;
    __device__ bool operator()(Vec3 o, Vec3 d, float *t_out) const
    {
        const RayDirRec rec = makeRayDirRec(d);
        RayRegs r;
        setRay(r, o.x, o.y, o.z, rec.invX, rec.invY, rec.invZ, 0);
        TriConsts tc;
        tc.Sx = rec.tri.Sx; tc.Sy = rec.tri.Sy; tc.Sz = rec.tri.Sz;
        tc.kx = rec.tri.kperm & 3; tc.ky = (rec.tri.kperm >> 2) & 3; tc.kz = (rec.tri.kperm >> 4) & 3;
        tc.okx = sel3(tc.kx, o.x, o.y, o.z);
        tc.oky = sel3(tc.ky, o.x, o.y, o.z);
        tc.okz = sel3(tc.kz, o.x, o.y, o.z);
        int32_t stack[32];
        int32_t sp = 0;
        stack[sp++] = 0;
        float t_max = FLT_MAX;
        while (sp > 0) {
            const char *nd = smem + (uint32_t)stack[--sp] * (uint32_t)sizeof(FastNode);
            float tn[4];
            if (r.offX >= 0) {
                visitNode(nd, r, one, tn);
            } else {
                const float4 v = visitNodeRef(nd, r.ox, r.oy, r.oz, r.invX, r.invY, r.invZ);
                tn[0] = v.x; tn[1] = v.y; tn[2] = v.z; tn[3] = v.w;
            }
            const uint32_t meta = reinterpret_cast<const uint32_t *>(nd)[3];
            const int4 ch = reinterpret_cast<const int4 *>(nd)[2];
            const int32_t child[4] = { ch.x, ch.y, ch.z, ch.w };
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (!(tn[i] <= t_max)) continue;
                if ((meta >> i) & 1u) {
                    const FastTri *tri = bvh.tris + (child[i] & 0x7FFFFFFF);
                    const int ntri = (int)((meta >> (4 + 7 * i)) & 0x7Fu);
                    for (int k = 0; k < ntri; k++) {
                        float t;
                        if (rayTriangleFast(tri + k, tc, t_max, &t)) t_max = t;
                    }
                } else {
                    stack[sp++] = child[i];
                }
            }
        }
        *t_out = t_max;
        return t_max != FLT_MAX;
    }
};

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

// Conservative reject in front of rayCapsuleRec: a ray whose horizontal projection passes the
// capsule's axis at more than r + 2 cannot touch the capsule -- all three discriminants are
// negative, so rayCapsuleRec would return 0 ("no hit").  |cross| / sqrt(a2) is that distance;
// the sphere caps' 3-D distance to the ray is at least as large.  Margin: at the cull radius the
// true discriminants are <= -64 a2 (-64 dd for the caps) while their rounding error is < 12 a2
// for any origin inside a 5,700-unit world (6e-8 * 6 * |o|^2), so no touching ray is rejected.
// (ox, oy): ray origin relative to the capsule, as CapsuleRec holds it.
constexpr float kCapsuleCullR2 = (kAgentRadius + 2.f) * (kAgentRadius + 2.f);
__device__ __forceinline__ bool capsuleMayHit(float ox, float oy, float dx, float dy, float a2)
{
    const float cr = oy * dx - ox * dy;
    return !(cr * cr > kCapsuleCullR2 * a2);
}

// The capsule half of traceRayWorld (utils.cpp:39-61): starts from the mesh
// result and returns the closest hit over the 12 agents of the world.  The agents are
// visited in index order (an earlier agent keeps a tie), skipping those capsuleMayHit rules out.
__device__ __forceinline__ bool closestWithCapsules(const CapsuleRec *caps, bool mesh_hit, float mesh_t,
                                                    float dx, float dy, float dz, float a2, float dd,
                                                    float *t_out, int32_t *agent_out)
{
    float min_hit_t = mesh_hit ? mesh_t : FLT_MAX;
    bool hit = mesh_hit;
    int32_t hit_agent = -1;
    uint32_t cand = 0;
#pragma unroll
    for (int i = 0; i < kAgents; i++) {
        const float2 oxy = *reinterpret_cast<const float2 *>(&caps[i]);
        cand |= capsuleMayHit(oxy.x, oxy.y, dx, dy, a2) ? 1u << i : 0u;
    }
    while (cand) {
        const int i = __ffs(cand) - 1;
        cand &= cand - 1u;
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
