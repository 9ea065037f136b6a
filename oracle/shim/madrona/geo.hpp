// TEST INFRASTRUCTURE (oracle shim): restatement of the three madrona::geo
// routines the reference path calls.  madrona::geo lives in the un-vendored
// Madrona engine (shacklettbp/madrona, version UNPINNED by the reference), so
// these follow the published algorithms the call sites name:
//   * utils.cpp:45-47          geo::intersectRayZOriginCapsule
//   * mesh_bvh.inl:929-930     geo::computeTriangleGeoNormal   (Jolt: shortest-edge cross)
//   * mesh_bvh.inl:946,1120    geo::triangleClosestPointToOrigin (RTCD 5.1.5)
// PARITY UNPINNED for these three: no reference test or golden vector exists.
#pragma once

#include "shim_core.hpp"

namespace madrona::geo {

// Ray vs capsule whose segment runs from the origin to (0, 0, h), radius r.
// Returns the entry distance t > 0, or 0 when the ray misses, points away, or
// starts inside the capsule (utils.cpp:49 relies on "0 == no hit").
inline float intersectRayZOriginCapsule(math::Vector3 o, math::Vector3 d,
                                        float r, float h)
{
    const float r2 = r * r;

    // Origin inside the capsule?
    {
        float zc = fminf(fmaxf(o.z, 0.f), h);
        float dz = o.z - zc;
        float dist2 = o.x * o.x + o.y * o.y + dz * dz;
        if (dist2 <= r2) {
            return 0.f;
        }
    }

    float t_hit = FLT_MAX;

    // Side of the cylinder (infinite cylinder clipped to 0 <= z <= h).
    {
        float a = d.x * d.x + d.y * d.y;
        float b = o.x * d.x + o.y * d.y;
        float c = o.x * o.x + o.y * o.y - r2;
        if (a > 0.f) {
            float disc = b * b - a * c;
            if (disc >= 0.f) {
                float t = (-b - sqrtf(disc)) / a;
                float z = o.z + t * d.z;
                if (t > 0.f && z >= 0.f && z <= h) {
                    t_hit = t;
                }
            }
        }
    }

    // End-cap spheres at z = 0 and z = h.
    float dd = d.x * d.x + d.y * d.y + d.z * d.z;
    for (int cap = 0; cap < 2; cap++) {
        float oz = cap == 0 ? o.z : o.z - h;
        float b = o.x * d.x + o.y * d.y + oz * d.z;
        float c = o.x * o.x + o.y * o.y + oz * oz - r2;
        float disc = b * b - dd * c;
        if (disc >= 0.f) {
            float t = (-b - sqrtf(disc)) / dd;
            if (t > 0.f && t < t_hit) {
                t_hit = t;
            }
        }
    }

    return t_hit == FLT_MAX ? 0.f : t_hit;
}

// Unnormalised geometric normal from the edge a->b and the shorter of the
// edges a->c / b->c (Jolt's "troublesome triangle" rule); same orientation
// as cross(ab, ac).
inline math::Vector3 computeTriangleGeoNormal(math::Vector3 ab,
                                              math::Vector3 ac,
                                              math::Vector3 bc)
{
    float ac_len2 = ac.x * ac.x + ac.y * ac.y + ac.z * ac.z;
    float bc_len2 = bc.x * bc.x + bc.y * bc.y + bc.z * bc.z;
    math::Vector3 e = bc_len2 < ac_len2 ? bc : ac;
    return math::cross(ab, e);
}

// RTCD 5.1.5 ClosestPtPointTriangle with the query point at the origin.
inline math::Vector3 triangleClosestPointToOrigin(math::Vector3 a,
                                                  math::Vector3 b,
                                                  math::Vector3 c,
                                                  math::Vector3 ab,
                                                  math::Vector3 ac)
{
    using math::dot;
    math::Vector3 ap = -a;
    float d1 = dot(ab, ap);
    float d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) return a;

    math::Vector3 bp = -b;
    float d3 = dot(ab, bp);
    float d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) return b;

    float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
        float v = d1 / (d1 - d3);
        return a + ab * v;
    }

    math::Vector3 cp = -c;
    float d5 = dot(ab, cp);
    float d6 = dot(ac, cp);
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

}
