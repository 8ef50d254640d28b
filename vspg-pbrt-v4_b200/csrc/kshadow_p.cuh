// Shadow rays (SampleLd back half, vspg.cpp:1193-1252) as a persistent-lane kernel, the scheduling of kmedium_rs.cuh
// applied to the shadow chain: one ray per LANE, lanes do not wait for each other.  A lane's ray moves through
//   SH_ISECT    closest hit of the current segment (opaque hit -> blocked), then either ratio tracking through the
//               segment's medium or straight to the end of the segment
//   SH_NEED_SEG / SH_IN_SEG   SampleT_maj (media_sampleTMaj.h:49-117): DDA steps and exponential steps; every tentative
//               collision runs the ratio-tracking callback with its Russian roulette at once (the tracker's RNG feeds
//               both the steps and the roulette, so the order of draws is the reference's)
//   end of segment: transmittance gone -> blocked; no surface ahead -> done; interface -> next segment towards the light
// and as soon as enough lanes of the warp are idle they fetch new rays from the stage queue.  k_shadow (one thread per
// ray, start to end) ran at 10 of 32 lanes on C2 and 7.5 on C5: rays differ in segments, collisions and early exits.
// The arithmetic per ray is exactly shadow_trace()'s (kernels2.cuh), which stays as the per-kernel parity hook.
// MEASURED SLOWER than k_shadow (C2 shadow stage 11.0 -> 14.3 ms, C5 40.1 -> 51.9 ms): shadow rays are short, and 96 registers
// plus the votes cost more than the idle lanes.  Kept behind VSPG_SHADOW_PERSIST=1 as the evidence (profiles/r02_summary.md).
#pragma once
#include "kernels2.cuh"

namespace vspg {

enum { SH_IDLE = 0, SH_ISECT = 1, SH_NEED_SEG = 2, SH_IN_SEG = 3, SH_SEG_END = 4 };

#ifndef VSPG_SHADOWP_MINB
#define VSPG_SHADOWP_MINB 5
#endif
constexpr int kShTripsPerCheck = 4, kShSearchMax = 4;

template <bool GRAY, int KINDS = kKindsAll>
__global__ void __launch_bounds__(kBlock, VSPG_SHADOWP_MINB) k_shadow_p(const DScene sc, const Wavefront W, int *cursor, const int refillIdle) {
    typedef typename SpecT<GRAY>::type Sp;
    const int n = *W.shadow.count;
    const int lane = threadIdx.x & 31;
    TrackCounters tc; tc.queries = 0; tc.dda = 0;
    unsigned nodes = 0;
    // lane state
    int slot = 0, phase = SH_IDLE, ch = 0;
    Ray ray; ray.medium = -1;                 // current segment, as spawned (unnormalised direction)
    V3 trd;                                   // its normalised direction (media_sampleTMaj.h:55-56)
    Hit h; h.prim = -1; h.t = 0; h.u = h.v = 0;
    MajorantIterator it; it.init_empty();
    Sp segSigma, T_maj, T_ray, r_l, r_u;
    float segTMax = 0, tMin = 0, u = 0;
    Rng rng;
    bool drained = false;
    while (true) {
        __syncwarp();
        unsigned idle = __ballot_sync(kFull, phase == SH_IDLE);
        if (!drained && __popc(idle) >= refillIdle) {
            int nIdle = __popc(idle), base = 0;
            if (lane == 0) base = atomicAdd(cursor, nIdle);
            base = __shfl_sync(kFull, base, 0);
            if (base + nIdle >= n) drained = true;
            if (phase == SH_IDLE) {
                int idx = base + __popc(idle & ((1u << lane) - 1));
                if (idx < n) {
                    slot = W.shadow.items[idx];
                    float4 a = W.P.sh_o[slot], b = W.P.sh_d[slot];
                    ray.o = V3(a.x, a.y, a.z); ray.d = V3(a.w, b.x, b.y); ray.medium = __float_as_int(b.z);
                    ch = (__float_as_int(b.w) >> 8) & 3;
                    T_ray = to_spec<GRAY>(splat(1.f)); r_l = T_ray; r_u = T_ray;
                    rng.set_sequence(hash_v3(ray.o), hash_v3(ray.d));
                    phase = (ray.d == V3(0, 0, 0)) ? SH_SEG_END : SH_ISECT;     // while (lightRay.d != 0)
                    h.prim = -1;
                }
            }
            __syncwarp();
            idle = __ballot_sync(kFull, phase == SH_IDLE);
        }
        if (idle == kFull) { if (drained) break; else continue; }
#pragma unroll 1
        for (int trip = 0; trip < kShTripsPerCheck; trip++) {
            // ---- (1) closest hit of a new segment
            if (phase == SH_ISECT) {
                h = scene_intersect<KINDS>(sc, ray.o, ray.d, 1 - kShadowEpsilon, nodes);
                if (h.prim >= 0 && sc.quads[h.prim].material >= 0) phase = SH_IDLE;                 // blocked
                else if (ray.medium >= 0) {
                    float tMax = h.prim >= 0 ? h.t : (1 - kShadowEpsilon);
                    u = rng.uniform();
                    tMax *= length(ray.d);
                    trd = normalize(ray.d);
                    Ray tr; tr.o = ray.o; tr.d = trd; tr.medium = ray.medium;
                    it = sample_ray(sc.media[ray.medium], tr, tMax);
                    T_maj = to_spec<GRAY>(splat(1.f));
                    phase = SH_NEED_SEG;
                } else phase = SH_SEG_END;
            }
            // ---- (2) towards the next tentative collision: DDA steps and exponential steps, bounded per trip
            float t = 0;
            bool hit = false;
#pragma unroll 1
            for (int j = 0; j < kShSearchMax; j++) {
                if (!hit) {
                    if (phase == SH_NEED_SEG) {
                        Segment seg;
                        if (!it.next(&seg)) {                                   // residual transmittance of the segment
                            tc.dda += it.steps;
                            T_ray = T_ray * (T_maj / T_maj[ch]); r_l = r_l * (T_maj / T_maj[ch]); r_u = r_u * (T_maj / T_maj[ch]);
                            phase = SH_SEG_END;
                        } else {
                            Sp sm = to_spec<GRAY>(seg.sigma_maj);
                            if (sm[ch] == 0) {
                                float dt = seg_dt(seg.tMax, seg.tMin);
                                T_maj = T_maj * fast_exp3(-dt * sm);
                            } else { segSigma = sm; segTMax = seg.tMax; tMin = seg.tMin; phase = SH_IN_SEG; }
                        }
                    }
                    if (phase == SH_IN_SEG) {
                        t = tMin + sample_exponential(u, segSigma[ch]);
                        u = rng.uniform();
                        if (t < segTMax) hit = true;
                        else {
                            float dt = seg_dt(segTMax, tMin);
                            T_maj = T_maj * fast_exp3(-dt * segSigma);
                            phase = SH_NEED_SEG;
                        }
                    }
                }
                unsigned hits = __ballot_sync(kFull, hit);
                unsigned walking = __ballot_sync(kFull, !hit && (phase == SH_NEED_SEG || phase == SH_IN_SEG));
                if (2 * __popc(hits) >= __popc(hits | walking)) break;
            }
            // ---- (3) ratio-tracking callback at the collision (vspg.cpp:1212-1236)
            if (hit) {
                const DMedium &med = sc.media[ray.medium];
                T_maj = T_maj * fast_exp3(-(t - tMin) * segSigma);
                V3 p = ray.o + trd * t;
                MediumProps mp = sample_point(med, p);
                tc.queries++;
                Sp sigma_n = clamp_zero(segSigma - to_spec<GRAY>(mp.sigma_a) - to_spec<GRAY>(mp.sigma_s));
                float pdf = T_maj[ch] * segSigma[ch];
                T_ray = T_ray * (T_maj * sigma_n / pdf);
                r_l = r_l * (T_maj * segSigma / pdf);
                r_u = r_u * (T_maj * sigma_n / pdf);
                Sp Tr = T_ray / avg3(r_l + r_u);
                if (max_comp(Tr) < 0.05f) {
                    float q = 0.75f;
                    if (rng.uniform() < q) T_ray = Sp();
                    else T_ray = T_ray / (1 - q);
                }
                if (nonzero(T_ray)) { T_maj = to_spec<GRAY>(splat(1.f)); tMin = t; }
                else {   // the callback returned false: SampleT_maj returns (1,1,1)
                    tc.dda += it.steps;
                    T_maj = to_spec<GRAY>(splat(1.f));
                    T_ray = T_ray * (T_maj / T_maj[ch]); r_l = r_l * (T_maj / T_maj[ch]); r_u = r_u * (T_maj / T_maj[ch]);
                    phase = SH_SEG_END;
                }
            }
            // ---- (4) end of a segment
            if (phase == SH_SEG_END) {
                const bool finished = h.prim < 0 || ray.d == V3(0, 0, 0);
                if (!nonzero(T_ray)) phase = SH_IDLE;                                               // blocked
                else if (finished) {
                    ShadowRec sr; shadow_load(W.P, slot, sr);
                    V3 rl3 = to_v3(r_l) * sr.rl;
                    V3 ru3 = to_v3(r_u) * sr.ru;
                    // f_hat * T_ray * ls->L / (r_l + r_u).Average()  (vspg.cpp:1247-1250), then L += beta * Ld (:479, :833)
                    V3 v = sr.delta ? sr.f_hat * to_v3(T_ray) * sr.L / avg3(rl3) : sr.f_hat * to_v3(T_ray) * sr.L / avg3(rl3 + ru3);
                    V3 Ld = sr.beta * v;
                    float4 e = W.P.rl_L[slot];
                    e.y += Ld.x; e.z += Ld.y; e.w += Ld.z;
                    W.P.rl_L[slot] = e;
                    phase = SH_IDLE;
                } else {
                    float4 pl = W.P.sh_pl[slot], pe = W.P.sh_ple[slot];
                    const float nz = W.P.sh_ru[slot].w;
                    SurfaceIntr hs = make_intr<KINDS>(sc, h, ray.o, ray.d, ray.medium);
                    spawn_ray_to(hs.p, hs.pError, hs.n, V3(pl.x, pl.y, pl.z), V3(pe.x, pe.y, pe.z), V3(pl.w, pe.w, nz), &ray.o, &ray.d);
                    ray.medium = hs.get_medium(ray.d);
                    h.prim = -1;
                    phase = (ray.d == V3(0, 0, 0)) ? SH_SEG_END : SH_ISECT;
                }
            }
            const unsigned idleNow = __ballot_sync(kFull, phase == SH_IDLE);
            if (drained ? idleNow == kFull : __popc(idleNow) >= refillIdle) break;
        }
    }
    counter_add(W.counters, CNT_QUERIES_SH, tc.queries);
    counter_add(W.counters, CNT_DDA_SH, tc.dda);
    counter_add(W.counters, CNT_BVH, nodes);
}

}  // namespace vspg
