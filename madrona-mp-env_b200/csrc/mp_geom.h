// mp_geom.h -- geometry queries of the MadronaMPEnv step: closest-hit ray
// traversal and swept-sphere casts against the static 4-wide quantised mesh
// BVH, the ray-vs-agent-capsule test and the combined world query.
//
// Behaviour follows the reference (cited per function) operation for
// operation: the traversal ORDER (LIFO stack, children 0..3, leaves tested
// inline) and the exact float expressions decide which of two near-equal hits
// wins, so they are kept.  The memory layout is not the reference's: vertices
// are 12-byte positions (no UVs), and the node / vertex arrays are expected to
// sit in shared memory when the caller stages them there (BVHView just holds
// pointers).
#pragma once

#include "mp_math.h"

namespace mpenv {

// Ray census ("BVH rays/s ... counted with a device counter in an
// instrumented build, not in the timed build"): libmpenv_b200_count.so is the same
// source compiled with -DMPENV_COUNT_RAYS; the product library compiles these away.
// [0] lidar rays, [1] line-of-sight rays, [2] hitscan rays, [3] swept-sphere casts --
// counted where the reference would call traceRayAgainstWorld / sphereCast.
#if defined(MPENV_COUNT_RAYS) && defined(__CUDACC__)
__device__ unsigned long long g_rayCounts[4];
#if defined(__CUDA_ARCH__)
#define MP_COUNT_RAYS(kind, n) atomicAdd(&g_rayCounts[kind], (unsigned long long)(n))
#else
#define MP_COUNT_RAYS(kind, n) ((void)0)
#endif
#else
#define MP_COUNT_RAYS(kind, n) ((void)0)
#endif

// ---------------------------------------------------------------- ray / tri
// Woop et al. watertight ray/triangle setup (reference mesh_bvh.inl:584-741;
// only the shear constants are consumed by the triangle test).
struct RayTriTxfm { int32_t kx, ky, kz; float Sx, Sy, Sz; };

MP_HD RayTriTxfm computeRayTriTxfm(Vec3 d, Vec3 inv_d)
{
    float abs_x = fabsf(d.x), abs_y = fabsf(d.y), abs_z = fabsf(d.z);
    int32_t kz;
    if (abs_x > abs_y && abs_x > abs_z) kz = 0;
    else if (abs_y > abs_z) kz = 1;
    else kz = 2;
    int32_t kx = kz + 1; if (kx == 3) kx = 0;
    int32_t ky = kx + 1; if (ky == 3) ky = 0;
    if (comp(d, kz) < 0.f) { int32_t t = kx; kx = ky; ky = t; }
    RayTriTxfm r;
    r.kx = kx; r.ky = ky; r.kz = kz;
    r.Sx = comp(d, kx) * comp(inv_d, kz);
    r.Sy = comp(d, ky) * comp(inv_d, kz);
    r.Sz = comp(inv_d, kz);
    return r;
}

// reference mesh_bvh.inl:433-554 with MADRONA_MESHBVH_BACKFACE_CULLING defined
// (mesh_bvh.inl:3).  Returns true and the hit distance when the triangle is
// hit at T <= t_max * det.
MP_HD bool rayTriangle(Vec3 tri_a, Vec3 tri_b, Vec3 tri_c, const RayTriTxfm &tx,
                       Vec3 org, float t_max, float *out_t)
{
    const Vec3 A = tri_a - org, B = tri_b - org, C = tri_c - org;
    const float Akz = comp(A, tx.kz), Bkz = comp(B, tx.kz), Ckz = comp(C, tx.kz);
    const float Ax = fmaf(-tx.Sx, Akz, comp(A, tx.kx));
    const float Ay = fmaf(-tx.Sy, Akz, comp(A, tx.ky));
    const float Bx = fmaf(-tx.Sx, Bkz, comp(B, tx.kx));
    const float By = fmaf(-tx.Sy, Bkz, comp(B, tx.ky));
    const float Cx = fmaf(-tx.Sx, Ckz, comp(C, tx.kx));
    const float Cy = fmaf(-tx.Sy, Ckz, comp(C, tx.ky));

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

    float det = U + V + W;
    if (det == 0.f) return false;

    const float Az = tx.Sz * Akz, Bz = tx.Sz * Bkz, Cz = tx.Sz * Ckz;
    const float T = fmaf(U, Az, fmaf(V, Bz, W * Cz));
    if (T < 0.0f || T > t_max * det) return false;

    const float rcpDet = 1.0f / det;
    *out_t = T * rcpDet;
    return true;
}

// reference mesh_bvh.inl:110-208 (traceRay) + :360-431 (traceRayLeaf).
// Returns true on a hit with *t_out = closest distance.
MP_HD_NOINLINE bool bvhTraceRay(const BVHView &bvh, Vec3 ray_o, Vec3 ray_d, float *t_out)
{
    constexpr float diveps = 0.0000001f;
    Vec3 inv_d = v3(1.f / ray_d.x, 1.f / ray_d.y, 1.f / ray_d.z);
    RayTriTxfm tx = computeRayTriTxfm(ray_d, inv_d);

    const float rayXInv = copysignf(ray_d.x == 0 ? 1 / diveps : 1 / ray_d.x, ray_d.x);
    const float rayYInv = copysignf(ray_d.y == 0 ? 1 / diveps : 1 / ray_d.y, ray_d.y);
    const float rayZInv = copysignf(ray_d.z == 0 ? 1 / diveps : 1 / ray_d.z, ray_d.z);

    int32_t stack[32];
    int32_t stack_size = 0;
    stack[stack_size++] = 0;

    float t_max = FLT_MAX;
    bool ray_hit = false;

    while (stack_size > 0) {
        const BVHNode &node = bvh.nodes[stack[--stack_size]];

        float dirQuantX = u2f((uint32_t)((int32_t)node.expX + 127) << 23) * rayXInv;
        float dirQuantY = u2f((uint32_t)((int32_t)node.expY + 127) << 23) * rayYInv;
        float dirQuantZ = u2f((uint32_t)((int32_t)node.expZ + 127) << 23) * rayZInv;

        float originQuantX = (node.minX - ray_o.x) * rayXInv;
        float originQuantY = (node.minY - ray_o.y) * rayYInv;
        float originQuantZ = (node.minZ - ray_o.z) * rayZInv;

#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < 4; i++) {
            int32_t child = node.children[i];
            if (child == kBVHSentinel) continue;

            float t_near_x = (float)node.qMinX[i] * dirQuantX + originQuantX;
            float t_near_y = (float)node.qMinY[i] * dirQuantY + originQuantY;
            float t_near_z = (float)node.qMinZ[i] * dirQuantZ + originQuantZ;
            float t_far_x = (float)node.qMaxX[i] * dirQuantX + originQuantX;
            float t_far_y = (float)node.qMaxY[i] * dirQuantY + originQuantY;
            float t_far_z = (float)node.qMaxZ[i] * dirQuantZ + originQuantZ;

            float t_near = fmaxf(fminf(t_near_x, t_far_x), fmaxf(fminf(t_near_y, t_far_y),
                fmaxf(fminf(t_near_z, t_far_z), 0.f)));
            float t_far = fminf(fmaxf(t_far_x, t_near_x), fminf(fmaxf(t_far_y, t_near_y),
                fminf(fmaxf(t_far_z, t_near_z), t_max)));

            if (t_near <= t_far) {
                if (child & 0x80000000) {
                    int32_t leaf_idx = child & ~0x80000000;
                    int32_t num_tris = node.triSize[i];
                    // traceRayLeaf: a later triangle replaces an earlier one
                    // whenever T <= t_cur * det.
                    float leaf_t_max = t_max;
                    float hit_t = 0.f;
                    bool hit_tri = false;
                    for (int32_t k = 0; k < num_tris; k++) {
                        const Vec3 *tv = bvh.verts + (size_t)(leaf_idx + k) * 3;
                        float t;
                        if (rayTriangle(tv[0], tv[1], tv[2], tx, ray_o, leaf_t_max, &t)) {
                            hit_tri = true;
                            hit_t = t;
                            leaf_t_max = t;
                        }
                    }
                    if (hit_tri) {
                        ray_hit = true;
                        t_max = hit_t;
                    }
                } else {
                    stack[stack_size++] = child;
                }
            }
        }
    }

    *t_out = t_max;
    return ray_hit;
}

// ------------------------------------------------------------------ capsule
// Ray vs z-aligned capsule whose axis runs from the origin to (0,0,h), radius
// r.  Returns the entry distance, or 0 for "no hit" (miss, or origin inside).
// The reference calls madrona::geo::intersectRayZOriginCapsule
// (utils.cpp:45-47), which lives in the un-vendored Madrona engine; this is the
// project's restatement of that published contract (parity unpinned, see
// DESIGN.md) and is the same code the oracle's shim uses
// (oracle/shim/madrona/geo.hpp).
MP_HD float rayZCapsule(Vec3 o, Vec3 d, float r, float h)
{
    const float r2 = r * r;
    {
        float zc = fminf(fmaxf(o.z, 0.f), h);
        float dz = o.z - zc;
        float dist2 = o.x * o.x + o.y * o.y + dz * dz;
        if (dist2 <= r2) return 0.f;
    }
    float t_hit = FLT_MAX;
    {
        float a = d.x * d.x + d.y * d.y;
        float b = o.x * d.x + o.y * d.y;
        float c = o.x * o.x + o.y * o.y - r2;
        if (a > 0.f) {
            float disc = b * b - a * c;
            if (disc >= 0.f) {
                float t = (-b - sqrtf(disc)) / a;
                float z = o.z + t * d.z;
                if (t > 0.f && z >= 0.f && z <= h) t_hit = t;
            }
        }
    }
    float dd = d.x * d.x + d.y * d.y + d.z * d.z;
    for (int cap = 0; cap < 2; cap++) {
        float oz = cap == 0 ? o.z : o.z - h;
        float b = o.x * d.x + o.y * d.y + oz * d.z;
        float c = o.x * o.x + o.y * o.y + oz * oz - r2;
        float disc = b * b - dd * c;
        if (disc >= 0.f) {
            float t = (-b - sqrtf(disc)) / dd;
            if (t > 0.f && t < t_hit) t_hit = t;
        }
    }
    return t_hit == FLT_MAX ? 0.f : t_hit;
}

// The static-mesh part of traceRayAgainstWorld as a functor, so that a kernel can substitute
// its own traversal of the same BVH (ray_fast.cuh) without touching the logic around it.
struct RefMeshTrace {
    const BVHView &bvh;
    MP_HD bool operator()(Vec3 o, Vec3 d, float *t) const { return bvhTraceRay(bvh, o, d, t); }
};

// reference utils.cpp:10-72.  agent_pos: the world's 12 agent positions.
// Returns hit flag; *hit_agent = agent index or -1 for the static mesh.
template <class MeshTrace>
MP_HD bool traceRayWorldWith(const MeshTrace &mesh, const Vec3 *agent_pos, Vec3 o, Vec3 d,
                             float *hit_t_out, int32_t *hit_agent_out)
{
    float min_hit_t = FLT_MAX;
    float t;
    bool hit = mesh(o, d, &t);
    if (hit) min_hit_t = t;
    int32_t hit_agent = -1;
    for (int i = 0; i < kAgents; i++) {
        Vec3 capsule_origin = agent_pos[i];
        capsule_origin.z += kAgentRadius;
        Vec3 translated_o = o - capsule_origin;
        float entity_hit_t = rayZCapsule(translated_o, d, kAgentRadius,
                                         kStandHeight - 2.f * kAgentRadius);
        if (entity_hit_t != 0 && entity_hit_t < min_hit_t) {
            min_hit_t = entity_hit_t;
            hit = true;
            hit_agent = i;
        }
    }
    *hit_t_out = min_hit_t;
    *hit_agent_out = hit_agent;
    return hit;
}

MP_HD bool traceRayWorld(const BVHView &bvh, const Vec3 *agent_pos, Vec3 o, Vec3 d,
                         float *hit_t_out, int32_t *hit_agent_out)
{
    return traceRayWorldWith(RefMeshTrace{bvh}, agent_pos, o, d, hit_t_out, hit_agent_out);
}

// -------------------------------------------------------------- sphere cast
// madrona::geo::computeTriangleGeoNormal / triangleClosestPointToOrigin are
// un-vendored too (mesh_bvh.inl:929-930, 946, 1120); restated from the
// algorithms the call sites name (Jolt's shortest-edge normal; RTCD 5.1.5).
MP_HD Vec3 triangleGeoNormal(Vec3 ab, Vec3 ac, Vec3 bc)
{
    float ac_len2 = ac.x * ac.x + ac.y * ac.y + ac.z * ac.z;
    float bc_len2 = bc.x * bc.x + bc.y * bc.y + bc.z * bc.z;
    Vec3 e = bc_len2 < ac_len2 ? bc : ac;
    return cross(ab, e);
}

MP_HD Vec3 triangleClosestPointToOrigin(Vec3 a, Vec3 b, Vec3 c, Vec3 ab, Vec3 ac)
{
    Vec3 ap = -a;
    float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) return a;
    Vec3 bp = -b;
    float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) return b;
    float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
        float v = d1 / (d1 - d3);
        return a + ab * v;
    }
    Vec3 cp = -c;
    float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) return c;
    float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) {
        float w = d2 / (d2 - d6);
        return a + ac * w;
    }
    float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
        float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        return b + (c - b) * w;
    }
    float denom = 1.f / (va + vb + vc);
    float v = vb * denom;
    float w = vc * denom;
    return a + ab * v + ac * w;
}

// reference mesh_bvh.inl:1008-1071 (edge = capless cylinder, RTCD 5.3.7)
MP_HD float sphereCastEdge(Vec3 ray_d, float ray_d_len2, float sphere_r2,
                           Vec3 axis, Vec3 cylinder_base, float hit_t)
{
    constexpr float edge_epsilon = 1e-6f;
    Vec3 start = -cylinder_base;
    const float start_dot_axis = dot(start, axis);
    const float dir_dot_axis = dot(ray_d, axis);
    const float end_dot_axis = start_dot_axis + dir_dot_axis;
    if (start_dot_axis < 0.0f && end_dot_axis < 0.0f) return hit_t;
    const float axis_len2 = length2(axis);
    if (start_dot_axis > axis_len2 && end_dot_axis > axis_len2) return hit_t;
    float a = axis_len2 * ray_d_len2 - dir_dot_axis * dir_dot_axis;
    if (fabsf(a) < edge_epsilon) return hit_t;
    float b = axis_len2 * dot(start, ray_d) - dir_dot_axis * start_dot_axis;
    float c = axis_len2 * (length2(start) - sphere_r2) - start_dot_axis * start_dot_axis;
    float det = b * b - a * c;
    if (det < 0.0f) return hit_t;
    float t = -(b + sqrtf(det)) / a;
    if (t < 0.0f || t >= hit_t) return hit_t;
    if (start_dot_axis + t * dir_dot_axis < 0.0f ||
        start_dot_axis + t * dir_dot_axis > axis_len2) return hit_t;
    return t;
}

// reference mesh_bvh.inl:1073-1104.  NOTE: the reference passes the
// origin-relative vertex and still subtracts it from the absolute ray origin
// (m = ray_o - (tri_v - ray_o)); kept as is.
MP_HD float sphereCastVert(Vec3 ray_o, Vec3 ray_d, float sphere_r2, Vec3 v, float hit_t)
{
    Vec3 m = ray_o - v;
    float b = dot(m, ray_d);
    float c = dot(m, m) - sphere_r2;
    if (c > 0.0f && b > 0.0f) return hit_t;
    float discr = b * b - c;
    if (discr < 0.0f) return hit_t;
    float t = -b - sqrtf(discr);
    if (t < 0.f) return 0.f;
    if (t >= hit_t) return hit_t;
    return t;
}

// reference mesh_bvh.inl:885-1127.  *out_n is written only on a hit.
// Bounded sweeps (bvhSphereCast's t_bound) and the reference's float arithmetic: the edge test of a
// LONG edge (axis_len2 ~ 2e6 times |start|^2 ~ 2e6, differences of products near 5e12) reports contact
// when the sphere is still up to ~0.02 units farther than its radius from the edge, i.e. BEFORE the
// sphere reaches the triangle's plane: grazing a 1300-unit wall at 15.008 units the reference's own
// loop returns the diagonal edge at t = 6.88 although plane_t = 7.16.  The reference's result then
// depends on whether some EARLIER triangle of its unbounded sweep has already pulled t_max below
// plane_t -- state a bounded sweep does not have.  So a bounded sweep also DETECTS whether such a hit
// is possible (kSweepSlack is a generous bound on that error, in distance units) and reports it through
// `risky`; the caller then repeats the cast unbounded, which is the reference's loop itself.
#ifndef MPENV_SWEEP_SLACK
#define MPENV_SWEEP_SLACK 0.5f
#endif
constexpr float kSweepSlack = MPENV_SWEEP_SLACK;

// The three edge and three vertex tests of sphereCastTriangle, from the running bound hit_t.
MP_HD float sphereCastFeatures(Vec3 e01, Vec3 e02, Vec3 e12, Vec3 v0, Vec3 v1, Vec3 v2, Vec3 ray_o, Vec3 ray_d,
                               float sphere_r2, float hit_t)
{
    const float ray_d_len2 = length2(ray_d);
    hit_t = sphereCastEdge(ray_d, ray_d_len2, sphere_r2, e01, v0, hit_t);
    hit_t = sphereCastEdge(ray_d, ray_d_len2, sphere_r2, e02, v0, hit_t);
    hit_t = sphereCastEdge(ray_d, ray_d_len2, sphere_r2, e12, v1, hit_t);
    hit_t = sphereCastVert(ray_o, ray_d, sphere_r2, v0, hit_t);
    hit_t = sphereCastVert(ray_o, ray_d, sphere_r2, v1, hit_t);
    hit_t = sphereCastVert(ray_o, ray_d, sphere_r2, v2, hit_t);
    return hit_t;
}

// Detection only: would the feature tests of this triangle report contact before t_max?
MP_HD bool sphereCastFeaturesBefore(Vec3 tri_a, Vec3 tri_b, Vec3 tri_c, Vec3 ray_o, Vec3 ray_d, float t_max,
                                    float sphere_r)
{
    const Vec3 e01 = tri_b - tri_a, e02 = tri_c - tri_a, e12 = tri_c - tri_b;
    const Vec3 v0 = tri_a - ray_o, v1 = tri_b - ray_o, v2 = tri_c - ray_o;
    return sphereCastFeatures(e01, e02, e12, v0, v1, v2, ray_o, ray_d, sphere_r * sphere_r, t_max) < t_max;
}

// detect (bounded sweeps): returns kSweepRisky (< 0, no contact time is) when the plane early-out below
// hides a feature hit (see kSweepSlack).
constexpr float kSweepRisky = -1.f;
MP_HD_NOINLINE float sphereCastTriangle(Vec3 tri_a, Vec3 tri_b, Vec3 tri_c, Vec3 ray_o,
                                        Vec3 ray_d, float t_max, float sphere_r, Vec3 *out_n,
                                        bool detect = false)
{
    const Vec3 e01 = tri_b - tri_a, e02 = tri_c - tri_a, e12 = tri_c - tri_b;
    const Vec3 v0 = tri_a - ray_o, v1 = tri_b - ray_o, v2 = tri_c - ray_o;

    Vec3 n_un = triangleGeoNormal(e01, e02, e12);
    float n_len = length(n_un);
    Vec3 n = n_un / n_len;
    const float n_dot_d = dot(n, ray_d);
    const float sphere_r2 = sphere_r * sphere_r;

    if (fabsf(dot(v0, n)) <= sphere_r) {
        Vec3 q = triangleClosestPointToOrigin(v0, v1, v2, e01, e02);
        float q_len2 = length2(q);
        if (q_len2 <= sphere_r2) {
            float q_len = sqrtf(q_len2);
            *out_n = q_len > 0.0f ? q / q_len : kUp();
            return 0.f;
        }
    } else {
        float abs_n_dot_d = fabsf(n_dot_d);
        if (abs_n_dot_d > 1.0e-6f) {
            float sgn = copysignf(1.f, n_dot_d);
            Vec3 extruded_delta = (sgn * sphere_r) * n;
            Vec3 v0_ext = v0 - extruded_delta;
            float plane_t = dot(v0_ext, n) / n_dot_d;
            if (plane_t * abs_n_dot_d < -sphere_r) return t_max;
            if (plane_t >= t_max) {
                // the plane is reached at or after the bound; the features can only be (mis)reported
                // before it if the sphere is within the slack of the plane by then
                if (detect && (plane_t - t_max) * abs_n_dot_d <= kSweepSlack &&
                    sphereCastFeatures(e01, e02, e12, v0, v1, v2, ray_o, ray_d, sphere_r2, t_max) < t_max)
                    return kSweepRisky;
                return t_max;
            }
            if (plane_t >= 0.0f) {
                Vec3 e = cross(ray_d, v0_ext);
                float v = -dot(e02, e) * sgn;
                float w = dot(e01, e) * sgn;
                if (v >= 0.f && w >= 0.f && v + w <= n_len * abs_n_dot_d) {
                    *out_n = (-sgn) * n;
                    return plane_t;
                }
            }
        }
    }

    const float hit_t = sphereCastFeatures(e01, e02, e12, v0, v1, v2, ray_o, ray_d, sphere_r2, t_max);
    if (hit_t >= t_max) return t_max;

    Vec3 hit_pt = ray_d * hit_t;
    Vec3 closest = triangleClosestPointToOrigin(v0 - hit_pt, v1 - hit_pt, v2 - hit_pt, e01, e02);
    *out_n = normalize(closest);
    return hit_t;
}

// reference mesh_bvh.inl:817-855
MP_HD bool sphereCastNodeCheck(Vec3 o, Vec3 inv_d, float t_max, float sphere_r,
                               Vec3 bmin, Vec3 bmax)
{
    bmin.x -= sphere_r; bmin.y -= sphere_r; bmin.z -= sphere_r;
    bmax.x += sphere_r; bmax.y += sphere_r; bmax.z += sphere_r;
    float t_min = 0.f;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 3; i++) {
        float inv_d_i = comp(inv_d, i);
        float b_min, b_max;
        if ((f2u(inv_d_i) >> 31) == 0) { b_min = comp(bmin, i); b_max = comp(bmax, i); }
        else { b_min = comp(bmax, i); b_max = comp(bmin, i); }
        float oi = comp(o, i);
        float i_min = (b_min - oi) * inv_d_i;
        float i_max = (b_max - oi) * inv_d_i;
        t_min = i_min > t_min ? i_min : t_min;
        t_max = i_max < t_max ? i_max : t_max;
    }
    return t_min < t_max;
}

// reference mesh_bvh.inl:743-815 (sphereCast) + :857-883 (sphereCastLeaf,
// which ALWAYS tests two consecutive triangles whatever the leaf's triSize;
// the vertex array is padded by one triangle so the last leaf's over-read is
// defined).  Returns FLT_MAX on a miss; *out_n untouched on a miss.
//
// t_bound (default: none): the caller declares that it only distinguishes hits
// with t < t_bound; anything at or beyond it is reported as a miss (FLT_MAX,
// *out_n untouched).  Hits below the bound are bit-identical to the unbounded
// cast: a triangle's contact time does not depend on the running bound (the
// bound only rejects candidates >= it, in every stage of sphereCastTriangle --
// including its plane_t >= t_max early-out, since no feature of a triangle is
// touched before its plane), first-wins tie order is unchanged, and a culled
// node's expanded box is entered at t >= t_bound.  The reference sweeps to
// infinity (mesh_bvh.inl:760) and the step then clamps the result to the
// distance moved this tick; bounding the sweep skips the far side of the map.
MP_HD_NOINLINE float bvhSphereCast(const BVHView &bvh, Vec3 ray_o, Vec3 ray_d, float sphere_r,
                                   Vec3 *out_n, float t_bound = FLT_MAX, bool *risky = nullptr)
{
    MP_COUNT_RAYS(3, 1);
    Vec3 inv_d = v3(1.f / ray_d.x, 1.f / ray_d.y, 1.f / ray_d.z);
    // risky (bounded sweeps): see kSweepSlack.  Besides the plane early-out inside sphereCastTriangle,
    // the bound culls nodes the reference's unbounded sweep would enter later; a subtree whose boxes,
    // grown by the slack, are entered BEFORE the bound is walked in a detect-only mode (its triangles'
    // feature tests only).  Needed only until the first hit: from then on the running bound is the
    // reference's own.
    constexpr int32_t kDetectOnly = 0x40000000;
    const bool detect = risky != nullptr && t_bound < FLT_MAX;
    int32_t stack[32];
    int32_t stack_size = 0;
    stack[stack_size++] = 0;
    Vec3 closest_n = v3(0.f, 0.f, 0.f);
    const float t_max = t_bound;
    float hit_t = t_max;
    while (stack_size > 0) {
        const int32_t entry = stack[--stack_size];
        const bool detect_only = (entry & kDetectOnly) != 0;
        const BVHNode &node = bvh.nodes[entry & ~kDetectOnly];
        const float sx = u2f(((uint32_t)(int32_t)node.expX + 127u) << 23);
        const float sy = u2f(((uint32_t)(int32_t)node.expY + 127u) << 23);
        const float sz = u2f(((uint32_t)(int32_t)node.expZ + 127u) << 23);
        for (int i = 0; i < 4; i++) {
            int32_t child = node.children[i];
            if (child == kBVHSentinel) continue;
            Vec3 bmin = v3(node.minX + sx * (float)node.qMinX[i],
                           node.minY + sy * (float)node.qMinY[i],
                           node.minZ + sz * (float)node.qMinZ[i]);
            Vec3 bmax = v3(node.minX + sx * (float)node.qMaxX[i],
                           node.minY + sy * (float)node.qMaxY[i],
                           node.minZ + sz * (float)node.qMaxZ[i]);
            if (detect_only || !sphereCastNodeCheck(ray_o, inv_d, hit_t, sphere_r, bmin, bmax)) {
                if (!detect || *risky || hit_t < t_max) continue;
                // the reference comes here only if the sweep meets the box at all; a feature inside can be
                // (mis)reported before the bound only if the grown box is entered before the bound
                if (!sphereCastNodeCheck(ray_o, inv_d, FLT_MAX, sphere_r, bmin, bmax)) continue;
                if (!sphereCastNodeCheck(ray_o, inv_d, t_max, sphere_r + kSweepSlack, bmin, bmax)) continue;
                if (child & 0x80000000) {
                    const int32_t leaf_idx = child & ~0x80000000;
                    for (int k = 0; k < 2; k++) {
                        const Vec3 *tv = bvh.verts + (size_t)(leaf_idx + k) * 3;
                        if (sphereCastFeaturesBefore(tv[0], tv[1], tv[2], ray_o, ray_d, t_max, sphere_r)) *risky = true;
                    }
                } else {
                    stack[stack_size++] = child | kDetectOnly;
                }
                continue;
            }
            if (child & 0x80000000) {
                int32_t leaf_idx = child & ~0x80000000;
                Vec3 leaf_n = v3(0.f, 0.f, 0.f);
                float leaf_t = hit_t;
                for (int k = 0; k < 2; k++) {
                    const Vec3 *tv = bvh.verts + (size_t)(leaf_idx + k) * 3;
                    const float t = sphereCastTriangle(tv[0], tv[1], tv[2], ray_o, ray_d, leaf_t,
                                                       sphere_r, &leaf_n, detect);
                    if (t < 0.f) *risky = true;        // kSweepRisky: leaf_t stays
                    else leaf_t = t;
                }
                if (leaf_t < hit_t) { hit_t = leaf_t; closest_n = leaf_n; }
            } else {
                stack[stack_size++] = child;
            }
        }
    }
    if (hit_t < t_max) *out_n = closest_n;
    return hit_t < t_max ? hit_t : FLT_MAX;
}

// When is a BOUNDED sweep (t_bound above) exact?  sphereCastVert is handed the ORIGIN-RELATIVE vertex
// and subtracts it from the absolute ray origin (see its comment: the reference's quirk, kept), so next
// to every real vertex v a cast also tests a sphere around the MIRRORED point v - ray_o -- a point that
// has nothing to do with where v's triangle sits in the BVH.  The argument above ("a culled node's box
// is entered at t >= t_bound") covers a triangle's real features only: a mirrored vertex of a far
// triangle, which the reference's unbounded sweep does visit, can lie right in front of the sphere.
// (Found by a 256-world soak: one agent-step in 1.4 million.)  A mirrored vertex can be hit below the bound
// only if some v lies within sphere_r of the segment 2 ray_o + t ray_d, t in [0, t_bound]; this
// returns false when the vertex occupancy grid (x, y only: conservative) has no vertex near that
// segment, and the callers drop the bound otherwise.
MP_HD bool sphereSweepMayTouchMirroredVertex(const MapConsts &mc, Vec3 ray_o, Vec3 ray_d, float t_bound, float sphere_r)
{
    const float pad = sphere_r + 1.f;
    const float ax = 2.f * ray_o.x, ay = 2.f * ray_o.y;
    const float bx = ax + ray_d.x * t_bound, by = ay + ray_d.y * t_bound;
    const float x0 = fminf(ax, bx) - pad, x1 = fmaxf(ax, bx) + pad;
    const float y0 = fminf(ay, by) - pad, y1 = fmaxf(ay, by) + pad;
    const float fi0 = (x0 - mc.vertGridMinX) * mc.vertGridInvCellX, fi1 = (x1 - mc.vertGridMinX) * mc.vertGridInvCellX;
    const float fj0 = (y0 - mc.vertGridMinY) * mc.vertGridInvCellY, fj1 = (y1 - mc.vertGridMinY) * mc.vertGridInvCellY;
    if (!(fi1 >= 0.f && fj1 >= 0.f && fi0 < (float)kVertGrid && fj0 < (float)kVertGrid)) {
        // outside the grid -- or not comparable (NaN): be safe
        return !(fi1 < 0.f || fj1 < 0.f || fi0 >= (float)kVertGrid || fj0 >= (float)kVertGrid);
    }
    // vertices on the upper border are clamped into the last cell (assets.cpp)
    const int i0 = fi0 < 0.f ? 0 : (int)fi0, i1 = fi1 >= (float)kVertGrid ? kVertGrid - 1 : (int)fi1;
    const int j0 = fj0 < 0.f ? 0 : (int)fj0, j1 = fj1 >= (float)kVertGrid ? kVertGrid - 1 : (int)fj1;
    for (int j = j0; j <= j1; j++)
        for (int i = i0; i <= i1; i++)
            if ((mc.vertGrid[(j * kVertGrid + i) >> 5] >> ((j * kVertGrid + i) & 31)) & 1u) return true;
    return false;
}

// A sweep of which the caller only distinguishes hits below `bound`: bounded where that is provably the
// reference's answer, else the reference's unbounded loop.  *out_n untouched on a miss.
template <typename BVH>
MP_HD_NOINLINE float sphereSweepBelow(const BVH &bvh, const MapConsts &mc, Vec3 ray_o, Vec3 ray_d, float sphere_r, Vec3 *out_n,
                             float bound)
{
    if (!sphereSweepMayTouchMirroredVertex(mc, ray_o, ray_d, bound, sphere_r)) {
        bool risky = false;
        Vec3 n = *out_n;
        const float t = bvhSphereCast(bvh, ray_o, ray_d, sphere_r, &n, bound, &risky);
        if (!risky) { *out_n = n; return t; }
    }
    return bvhSphereCast(bvh, ray_o, ray_d, sphere_r, out_n);
}

}  // namespace mpenv
