// EXPERIMENT (not product, not test): counts node visits / triangle tests of the
// reference-order traversal when it starts from a finite t_max guess.
#include "../../tests/hostsim/hostsim.cpp"

namespace {
struct Cnt { int nodes, tris; uint64_t nmask; uint64_t tmask[4]; };
bool traceBounded(const BVHView &bvh, Vec3 ray_o, Vec3 ray_d, float t_init, float *t_out, Cnt *cnt)
{
    constexpr float diveps = 0.0000001f;
    Vec3 inv_d = v3(1.f / ray_d.x, 1.f / ray_d.y, 1.f / ray_d.z);
    RayTriTxfm tx = computeRayTriTxfm(ray_d, inv_d);
    const float rayXInv = copysignf(ray_d.x == 0 ? 1 / diveps : 1 / ray_d.x, ray_d.x);
    const float rayYInv = copysignf(ray_d.y == 0 ? 1 / diveps : 1 / ray_d.y, ray_d.y);
    const float rayZInv = copysignf(ray_d.z == 0 ? 1 / diveps : 1 / ray_d.z, ray_d.z);
    int32_t stack[32]; int32_t stack_size = 0; stack[stack_size++] = 0;
    float t_max = t_init; bool ray_hit = false;
    while (stack_size > 0) {
        const BVHNode &node = bvh.nodes[stack[--stack_size]];
        cnt->nodes++; { int ni = (int)(&node - bvh.nodes); cnt->nmask |= 1ull << ni; }
        float dirQuantX = u2f((uint32_t)((int32_t)node.expX + 127) << 23) * rayXInv;
        float dirQuantY = u2f((uint32_t)((int32_t)node.expY + 127) << 23) * rayYInv;
        float dirQuantZ = u2f((uint32_t)((int32_t)node.expZ + 127) << 23) * rayZInv;
        float originQuantX = (node.minX - ray_o.x) * rayXInv;
        float originQuantY = (node.minY - ray_o.y) * rayYInv;
        float originQuantZ = (node.minZ - ray_o.z) * rayZInv;
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
                    float leaf_t_max = t_max, hit_t = 0.f; bool hit_tri = false;
                    for (int32_t k = 0; k < num_tris; k++) {
                        const Vec3 *tv = bvh.verts + (size_t)(leaf_idx + k) * 3;
                        float t; cnt->tris++; { int ti = leaf_idx + k; cnt->tmask[ti >> 6] |= 1ull << (ti & 63); }
                        if (rayTriangle(tv[0], tv[1], tv[2], tx, ray_o, leaf_t_max, &t)) {
                            hit_tri = true; hit_t = t; leaf_t_max = t;
                        }
                    }
                    if (hit_tri) { ray_hit = true; t_max = hit_t; }
                } else stack[stack_size++] = child;
            }
        }
    }
    *t_out = t_max;
    return ray_hit;
}
}

extern "C" {
// o, d: [numAgents*80][3]
void exp_lidar_rays(void *h, float *o_out, float *d_out)
{
    HostSim &s = *(HostSim *)h;
    const Ctx &c = s.ctx; const SimPtrs &p = c.p; const MapConsts &mc = *c.mc;
    for (int w = 0; w < (int)s.numWorlds; w++) for (int a = 0; a < kAgents; a++) {
        const int idx = w * kAgents + a;
        for (int ray = 0; ray < 80; ray++) {
            const bool is_fwd = ray < 64;
            const int r = is_fwd ? ray : ray - 64;
            const int width = is_fwd ? 32 : 8;
            const int hh = r / width, col = r % width;
            const Quat q = is_fwd ? p.aim[idx].rot : p.rot[idx];
            Vec3 fwd = rotateVec(q, kFwd()); Vec3 right = rotateVec(q, kRight());
            float top_height = viewHeight(p.stand[idx].cur) + kAgentRadius;
            float range = top_height - 2.f * kAgentRadius;
            float h_off = kAgentRadius + range * (float(hh) / 1.f);
            Vec3 ro = p.pos[idx]; ro.z += h_off;
            float x = is_fwd ? mc.fwdLidarX[col] : mc.rearLidarX[col];
            float y = is_fwd ? mc.fwdLidarY[col] : mc.rearLidarY[col];
            Vec3 rd = normalize(x * right + y * fwd);
            float *oo = o_out + ((size_t)idx * 80 + ray) * 3, *dd = d_out + ((size_t)idx * 80 + ray) * 3;
            oo[0] = ro.x; oo[1] = ro.y; oo[2] = ro.z; dd[0] = rd.x; dd[1] = rd.y; dd[2] = rd.z;
        }
    }
}
void exp_trace(void *h, int64_t n, const float *o, const float *d, const float *U, float *t_out, int32_t *hit_out,
               int32_t *nodes_out, int32_t *tris_out, uint64_t *nmask_out, uint64_t *tmask_out)
{
    HostSim &s = *(HostSim *)h;
    for (int64_t i = 0; i < n; i++) {
        Cnt c{0, 0, 0, {0,0,0,0}}; float t;
        bool hit = traceBounded(s.ctx.bvh, v3(o[i*3], o[i*3+1], o[i*3+2]), v3(d[i*3], d[i*3+1], d[i*3+2]), U[i], &t, &c);
        t_out[i] = t; hit_out[i] = hit; nodes_out[i] = c.nodes; tris_out[i] = c.tris; if (nmask_out) { nmask_out[i] = c.nmask; for (int q = 0; q < 4; q++) tmask_out[i*4+q] = c.tmask[q]; }
    }
}
int exp_num_nodes(void *h) { return (int)((HostSim *)h)->ctx.bvh.numNodes; }
}

// EXPERIMENT: line-of-sight rays that an early-out after the first visible sample would skip.
extern "C" void exp_los_stats(void *h, int64_t *out)
{
    HostSim &s = *(HostSim *)h;
    const Ctx &c = s.ctx; const SimPtrs &p = c.p; const float *fr = c.mc->frustum;
    int64_t rays = 0, skip_after_first = 0, skip_if_s0 = 0, pairs_traced = 0, pairs_visible = 0, vis_rays = 0;
    for (int w = 0; w < (int)s.numWorlds; w++) {
        const Vec3 *wp = p.pos + (size_t)w * kAgents; const int base = w * kAgents;
        for (int a = 0; a < kAgents; a++) for (int j = 0; j < kTeamSize; j++) {
            const int idx = base + a; const int opp = (teamOf(a) ^ 1) * kTeamSize + j;
            if (p.alive[idx] == 0.f || p.alive[base + opp] == 0.f) continue;
            Vec3 o = wp[a]; o.z += viewHeight(p.stand[idx].cur);
            const Aim aim = p.aim[idx]; const Quat inv = quatInv(aim.rot);
            const float vh = viewHeight(p.stand[base + opp].cur);
            Vec3 dr = rotateVec(aim.rot, kRight()) * 0.9f * kAgentRadius;
            bool seen = false; int n = 0; bool vis0 = false;
            for (int sidx = 0; sidx < 4; sidx++) {
                Vec3 pt = losSamplePoint(wp[opp], vh, dr, sidx);
                Vec3 tt = pt - o; Vec3 vp = rotateVec(inv, tt);
                if (vp.y <= 0.f) continue;
                if (!(vp.y * fr[1] - fabsf(vp.x) * fr[0] > -kAgentRadius)) continue;
                if (!(vp.y * fr[3] - fabsf(vp.z) * fr[2] > -kAgentRadius)) continue;
                if (length(tt) < kAgentRadius) continue;
                rays++; n++;
                if (seen) skip_after_first++;
                if (vis0 && sidx > 0) skip_if_s0++;
                bool v = losSampleVisible(c, wp, o, inv, pt, opp);
                if (v) vis_rays++;
                if (v && sidx == 0) vis0 = true;
                if (v) seen = true;
            }
            if (n) pairs_traced++;
            if (seen) pairs_visible++;
        }
    }
    out[0] = rays; out[1] = skip_after_first; out[2] = skip_if_s0; out[3] = pairs_traced; out[4] = pairs_visible; out[5] = vis_rays;
}
