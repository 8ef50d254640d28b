// Scatter-vertex work: light sampling (NEE set-up), GuidedPhaseFunction / GuidedBSDF (cpu/guiding.h:57-638) and the
// shadow-ray record.  Templated on the lane-group width G: G = 8/16/32 when a guiding field is loaded -- the caller
// runs one lane GROUP per vertex (all lanes of the group execute this code on identical scalars, the vMF mixture is
// spread over the group's lanes, see dguiding.cuh); G = 0 without a field -- one THREAD per vertex and every
// mixture op compiled out.
#pragma once
#include "dshapesample.cuh"
#include "dsphere.cuh"
#include "dguiding.cuh"
#include "dpath.cuh"

namespace vspg {

struct LightSampleCtx { V3 p; };

struct LiSample { bool valid; V3 L, wi; float pdf; V3 pLight, pLightErr, nLight; bool delta; };

// lightsamplers.h:140-201 CompactLightBounds::Importance on a decoded node
VD float light_importance(const DLightNode &nd, V3 p, V3 n) {
    V3 bmin = ld3(nd.bmin), bmax = ld3(nd.bmax);
    float cosTheta_o = nd.cosTheta_o, cosTheta_e = nd.cosTheta_e;
    V3 pc = (bmin + bmax) / 2;
    float d2 = dist_sq(p, pc);
    d2 = fmaxf(d2, length(bmax - bmin) / 2);
    V3 wi = normalize(p - pc);
    float cosTheta_w = dot(ld3(nd.w), wi);
    if (nd.twoSided) cosTheta_w = fabsf(cosTheta_w);
    float sinTheta_w = safe_sqrt(1 - sqr(cosTheta_w));
    // BoundSubtendedDirections(bounds, p).cosTheta  (util/vecmath.h:1815-1828)
    float cosTheta_b;
    {
        bool inside = pc.x >= bmin.x && pc.x <= bmax.x && pc.y >= bmin.y && pc.y <= bmax.y && pc.z >= bmin.z && pc.z <= bmax.z;
        float radius = inside ? length(pc - bmax) : 0;
        if (dist_sq(p, pc) < sqr(radius)) cosTheta_b = -1;
        else cosTheta_b = safe_sqrt(1 - sqr(radius) / dist_sq(pc, p));
    }
    float sinTheta_b = safe_sqrt(1 - sqr(cosTheta_b));
    float sinTheta_o = safe_sqrt(1 - sqr(cosTheta_o));
    // cosSubClamped / sinSubClamped
    float cosTheta_x = cosTheta_w > cosTheta_o ? 1.f : cosTheta_w * cosTheta_o + sinTheta_w * sinTheta_o;
    float sinTheta_x = cosTheta_w > cosTheta_o ? 0.f : sinTheta_w * cosTheta_o - cosTheta_w * sinTheta_o;
    float cosThetap = cosTheta_x > cosTheta_b ? 1.f : cosTheta_x * cosTheta_b + sinTheta_x * sinTheta_b;
    if (cosThetap <= cosTheta_e) return 0;
    float importance = nd.phi * cosThetap / d2;
    if (n.x != 0 || n.y != 0 || n.z != 0) {
        float cosTheta_i = absdotn(wi, n);
        float sinTheta_i = safe_sqrt(1 - sqr(cosTheta_i));
        float cosThetap_i = cosTheta_i > cosTheta_b ? 1.f : cosTheta_i * cosTheta_b + sinTheta_i * sinTheta_b;
        importance *= cosThetap_i;
    }
    return fmaxf(importance, 0.f);
}
// lightsamplers.h:266-321 BVHLightSampler::Sample(ctx, u)
VD int sample_light_bvh(const DScene &sc, V3 p, V3 n, float u, float *pmfOut) {
    const int nInf = sc.n_unbounded;
    const float pInfinite = float(nInf) / float(nInf + (sc.n_lbvh == 0 ? 0 : 1));
    if (u < pInfinite) {
        u /= pInfinite;
        int index = min((int)(u * nInf), nInf - 1);
        *pmfOut = pInfinite / nInf;
        return sc.unbounded_lights[index];
    }
    if (sc.n_lbvh == 0) return -1;
    u = fminf((u - pInfinite) / (1 - pInfinite), kOneMinusEpsilon);
    int nodeIndex = 0;
    float pmf = 1 - pInfinite;
    while (true) {
        const DLightNode &node = sc.lbvh[nodeIndex];
        if (!node.isLeaf) {
            float ci0 = light_importance(sc.lbvh[nodeIndex + 1], p, n), ci1 = light_importance(sc.lbvh[node.childOrLight], p, n);
            if (ci0 == 0 && ci1 == 0) return -1;
            // SampleDiscrete({ci0, ci1}, u, &nodePMF, &u)   util/sampling.h:79-112
            float sumWeights = 0; sumWeights += ci0; sumWeights += ci1;
            float up = u * sumWeights;
            if (up == sumWeights) up = next_down(up);
            int child = 0; float sum = 0, wsel = ci0;
            if (sum + ci0 <= up) { sum += ci0; child = 1; wsel = ci1; }
            pmf *= wsel / sumWeights;
            u = fminf((up - sum) / wsel, kOneMinusEpsilon);
            nodeIndex = child == 0 ? nodeIndex + 1 : node.childOrLight;
        } else {
            if (nodeIndex > 0 || light_importance(node, p, n) > 0) { *pmfOut = pmf; return node.childOrLight; }
            return -1;
        }
    }
}
// lightsamplers.h:324-359 BVHLightSampler::PMF(ctx, light)
VD float light_pmf_bvh(const DScene &sc, V3 p, V3 n, int light) {
    const int nInf = sc.n_unbounded;
    if (!sc.light_in_tree[light]) return 1.f / (nInf + (sc.n_lbvh == 0 ? 0 : 1));
    unsigned bitTrail = sc.light_trail[light];
    const float pInfinite = float(nInf) / float(nInf + (sc.n_lbvh == 0 ? 0 : 1));
    float pmf = 1 - pInfinite;
    int nodeIndex = 0;
    while (true) {
        const DLightNode &node = sc.lbvh[nodeIndex];
        if (node.isLeaf) return pmf;
        float ci[2] = {light_importance(sc.lbvh[nodeIndex + 1], p, n), light_importance(sc.lbvh[node.childOrLight], p, n)};
        pmf *= ci[bitTrail & 1] / (ci[0] + ci[1]);
        nodeIndex = (bitTrail & 1) ? node.childOrLight : (nodeIndex + 1);
        bitTrail >>= 1;
    }
}

// lightsamplers.h:33-38 UniformLightSampler / PowerLightSampler pmf (precomputed per light on the host) / BVHLightSampler.
// p, n: LightSampleContext::p() and ::ns of the vertex (n = 0 in a medium).
VD int sample_light(const DScene &sc, float u, float *pmf, V3 p, V3 n) {
    int nl = sc.n_lights;
    if (nl == 0) return -1;
    if (sc.ip.light_sampler == VSPG_LIGHTSAMPLER_BVH) return sample_light_bvh(sc, p, n, u, pmf);
    if (sc.ip.light_sampler == VSPG_LIGHTSAMPLER_UNIFORM) {
        int li = min((int)(u * nl), nl - 1);
        *pmf = 1.f / nl;
        return li;
    }
    // AliasTable::Sample (util/sampling.cpp:620-645)
    int offset = min((int)(u * nl), nl - 1);
    float up = fminf(u * nl - offset, kOneMinusEpsilon);
    if (up < sc.lights[offset].q) { *pmf = sc.lights[offset].pmf; return offset; }
    int alias = sc.lights[offset].alias;
    *pmf = sc.lights[alias].pmf;
    return alias;
}
VD float light_pmf(const DScene &sc, int light, V3 p, V3 n) {
    if (sc.n_lights == 0) return 0;
    if (sc.ip.light_sampler == VSPG_LIGHTSAMPLER_BVH) return light_pmf_bvh(sc, p, n, light);
    if (sc.ip.light_sampler == VSPG_LIGHTSAMPLER_UNIFORM) return 1.f / sc.n_lights;
    return sc.lights[light].pmf;
}

// shapes.h:1053-1125 Triangle::Sample(ctx, u) (and :1014-1048 Sample(u) for the area-sampling fallback)
__device__ __noinline__ bool sample_triangle_light(const DQuad &q, V3 ctxP, V3 ctxNs, float u0, float u1, V3 *pOut, V3 *pErr, V3 *nOut, float *pdfOut) {
    V3 p0 = ld3(q.p00), p1 = ld3(q.p10), p2 = ld3(q.p01);
    V3 n = normalize(cross(p1 - p0, p2 - p0));
    if (q.flip_normal) n = -n;
    V3 w0 = normalize(p0 - ctxP), w1 = normalize(p1 - ctxP), w2 = normalize(p2 - ctxP);
    float solidAngle = spherical_triangle_area(w0, w1, w2);
    float b[3];
    if (solidAngle < kMinSphericalSampleAreaTri || solidAngle > kMaxSphericalSampleAreaTri) {
        if (u0 < u1) { b[0] = u0 / 2; b[1] = u1 - b[0]; } else { b[1] = u1 / 2; b[0] = u0 - b[1]; }   // SampleUniformTriangle
        b[2] = 1 - b[0] - b[1];
        V3 p;   // Interaction(Point3fi(p, pError), ...) (shapes.h:1040-1046)
        point_interval(b[0] * p0 + b[1] * p1 + b[2] * p2, gammaN(6) * (vabs(b[0] * p0) + vabs(b[1] * p1) + vabs((1 - b[0] - b[1]) * p2)), &p, pErr);
        V3 wi = p - ctxP;
        if (length_sq(wi) == 0) return false;
        wi = normalize(wi);
        float pdf = (1 / q.area) / (absdotn(n, -wi) / dist_sq(ctxP, p));
        if (isinf(pdf)) return false;
        *pOut = p; *nOut = n; *pdfOut = pdf;
        return true;
    }
    float pdf = 1;
    if (!(ctxNs.x == 0 && ctxNs.y == 0 && ctxNs.z == 0)) {
        float w[4] = {fmaxf(0.01f, absdotn(ctxNs, w1)), fmaxf(0.01f, absdotn(ctxNs, w1)), fmaxf(0.01f, absdotn(ctxNs, w0)), fmaxf(0.01f, absdotn(ctxNs, w2))};
        float a0, a1;
        sample_bilinear(u0, u1, w, &a0, &a1);
        u0 = a0; u1 = a1;
        pdf = bilinear_pdf(u0, u1, w);
    }
    float triPDF;
    if (!sample_spherical_triangle(p0, p1, p2, ctxP, u0, u1, b, &triPDF)) b[0] = b[1] = b[2] = 0;
    if (triPDF == 0) return false;
    pdf *= triPDF;
    point_interval(b[0] * p0 + b[1] * p1 + b[2] * p2, gammaN(6) * (vabs(b[0] * p0) + vabs(b[1] * p1) + vabs((1 - b[0] - b[1]) * p2)), pOut, pErr);   // shapes.h:1117-1123
    *nOut = n; *pdfOut = pdf;
    return true;
}
// shapes.h:1128-1173 Triangle::PDF(ctx, wi) evaluated at the actual hit
__device__ __noinline__ float triangle_light_pdf(const DQuad &q, V3 ctxP, V3 ctxNs, V3 wi, V3 hitP, V3 hitN) {
    V3 p0 = ld3(q.p00), p1 = ld3(q.p10), p2 = ld3(q.p01);
    V3 w0 = normalize(p0 - ctxP), w1 = normalize(p1 - ctxP), w2 = normalize(p2 - ctxP);
    float solidAngle = spherical_triangle_area(w0, w1, w2);
    if (solidAngle < kMinSphericalSampleAreaTri || solidAngle > kMaxSphericalSampleAreaTri) {
        float r = (1 / q.area) / (absdotn(hitN, -wi) / dist_sq(ctxP, hitP));
        return isinf(r) ? 0 : r;
    }
    float pdf = 1 / solidAngle;
    if (!(ctxNs.x == 0 && ctxNs.y == 0 && ctxNs.z == 0)) {
        float uu0, uu1;
        invert_spherical_triangle_sample(p0, p1, p2, ctxP, wi, &uu0, &uu1);
        float w[4] = {fmaxf(0.01f, absdotn(ctxNs, w1)), fmaxf(0.01f, absdotn(ctxNs, w1)), fmaxf(0.01f, absdotn(ctxNs, w0)), fmaxf(0.01f, absdotn(ctxNs, w2))};
        pdf *= bilinear_pdf(uu0, uu1, w);
    }
    return pdf;
}
// shapes.h:300-361 Sphere::Sample(ctx, u) (shapes.cpp:38-59 Sample(u) for reference points inside the sphere); the
// reference point is exact here, so ctx.OffsetRayOrigin(pCenter) == ctx.p()
__device__ __noinline__ bool sample_sphere_light(const DQuad &q, V3 ctxP, float u0, float u1, V3 *pOut, V3 *pErr, V3 *nOut, float *pdfOut) {
    V3 pCenter = ld3(q.p00); float radius = q.p10[0];
    if (dist_sq(ctxP, pCenter) <= sqr(radius)) {
        float z = 1 - 2 * u0, r = safe_sqrt(1 - sqr(z));
        float sp, cp; vspg_sincosf(2 * kPi * u1, &sp, &cp);   // SampleUniformSphere: phi = 2 * Pi * u[1]
        V3 pObj = radius * V3(r * cp, r * sp, z);
        pObj = pObj * (radius / length(pObj));
        V3 n = normalize(pObj);
        if (q.flip_normal) n = -n;
        V3 p, dummyN, dummyDpdu;
        sphere_interaction(pObj, pCenter, radius, false, &p, pErr, &dummyN, &dummyDpdu);   // renderFromObject(Point3fi(pObj, gamma(5)|pObj|))
        V3 wi = p - ctxP;
        if (length_sq(wi) == 0) return false;
        wi = normalize(wi);
        float pdf = (1 / q.area) / (absdotn(n, -wi) / dist_sq(ctxP, p));
        if (isinf(pdf)) return false;
        *pOut = p; *nOut = n; *pdfOut = pdf;
        return true;
    }
    float sinThetaMax = radius / sqrtf(dist_sq(ctxP, pCenter));
    float sin2ThetaMax = sqr(sinThetaMax);
    float cosThetaMax = safe_sqrt(1 - sin2ThetaMax);
    float oneMinusCosThetaMax = 1 - cosThetaMax;
    float cosTheta = (cosThetaMax - 1) * u0 + 1;
    float sin2Theta = 1 - sqr(cosTheta);
    if (sin2ThetaMax < 0.00068523f) {
        sin2Theta = sin2ThetaMax * u0;
        cosTheta = sqrtf(1 - sin2Theta);
        oneMinusCosThetaMax = sin2ThetaMax / 2;
    }
    float cosAlpha = sin2Theta / sinThetaMax + cosTheta * safe_sqrt(1 - sin2Theta / sqr(sinThetaMax));
    float sinAlpha = safe_sqrt(1 - sqr(cosAlpha));
    V3 w = spherical_direction_2pi(sinAlpha, cosAlpha, u1);
    Frame f = Frame::from_z(normalize(pCenter - ctxP));
    V3 n = f.from_local(-w);
    V3 p = pCenter + radius * n;
    if (q.flip_normal) n = -n;
    point_interval(p, gammaN(5) * vabs(p), pOut, pErr);   // shapes.h:352-360 Interaction(Point3fi(p, pError), ...)
    *nOut = n; *pdfOut = 1 / (2 * kPi * oneMinusCosThetaMax);
    return true;
}
// shapes.h:364-392 Sphere::PDF(ctx, wi)
VD float sphere_light_pdf(const DQuad &q, V3 ctxP, V3 wi, V3 hitP, V3 hitN) {
    V3 pCenter = ld3(q.p00); float radius = q.p10[0];
    if (dist_sq(ctxP, pCenter) <= sqr(radius)) {
        float r = (1 / q.area) / (absdotn(hitN, -wi) / dist_sq(ctxP, hitP));
        return isinf(r) ? 0 : r;
    }
    float sin2ThetaMax = radius * radius / dist_sq(ctxP, pCenter);
    float cosThetaMax = safe_sqrt(1 - sin2ThetaMax);
    float oneMinusCosThetaMax = 1 - cosThetaMax;
    if (sin2ThetaMax < 0.00068523f) oneMinusCosThetaMax = sin2ThetaMax / 2;
    return 1 / (2 * kPi * oneMinusCosThetaMax);
}

// lights.cpp:796-818 DiffuseAreaLight::SampleLi -> shapes.cpp:1155-1214, 1254-1280 (area-sampling branch);
// lights.h:243-250 (point), 852-866 (spot), 316-323 (distant); uniform infinite: allowIncompletePDF (lights.cpp:1022-1023)
template <int KINDS = kKindsAll>
VD LiSample sample_li(const DScene &sc, int light, V3 ctxP, V3 ctxNs, float u0, float u1) {
    LiSample ls; ls.valid = false; ls.delta = false; ls.pdf = 0;
    const DLight &l = sc.lights[light];
    if (l.type == VSPG_LIGHT_POINT) {
        V3 p = ld3(l.pos);
        ls.wi = normalize(p - ctxP);
        ls.L = ld3(l.L) / dist_sq(p, ctxP);
        ls.pdf = 1; ls.pLight = p; ls.delta = true; ls.valid = true;
        return ls;
    }
    if (l.type == VSPG_LIGHT_SPOT) {   // lights.h:852-866, lights.cpp:1448-1451 (I), util/math.h:268-274 (SmoothStep)
        V3 p = ld3(l.pos);
        V3 wi = normalize(p - ctxP);
        V3 v = -wi;
        const float *m = l.toLight;
        V3 wl = normalize(V3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z));
        float ss;
        if (l.cosEnd == l.cosStart) ss = wl.z < l.cosEnd ? 0.f : 1.f;
        else {
            float t = fminf(fmaxf((wl.z - l.cosEnd) / (l.cosStart - l.cosEnd), 0.f), 1.f);
            ss = t * t * (3 - 2 * t);
        }
        float k = ss * l.scale;
        V3 Li = V3(k * l.I[0], k * l.I[1], k * l.I[2]) / dist_sq(p, ctxP);
        if (!nonzero(Li)) return ls;
        ls.wi = wi; ls.L = Li; ls.pdf = 1; ls.pLight = p; ls.delta = true; ls.valid = true;
        return ls;
    }
    if (l.type == VSPG_LIGHT_DISTANT) {
        V3 wi = normalize(ld3(l.pos));
        ls.wi = wi; ls.L = ld3(l.L); ls.pdf = 1; ls.delta = true; ls.valid = true;
        ls.pLight = ctxP + wi * (2 * sc.scene_radius);
        return ls;
    }
    if (l.type == VSPG_LIGHT_UNIFORM_INFINITE) return ls;
    const DQuad &q = sc.quads[l.shape_index];   // global primitive id
    if ((KINDS & 6) && q.kind != PRIM_QUAD) {
        V3 p, pErr, n; float pdf = 0;
        bool ok = q.kind == PRIM_TRIANGLE ? sample_triangle_light(q, ctxP, ctxNs, u0, u1, &p, &pErr, &n, &pdf)
                                          : sample_sphere_light(q, ctxP, u0, u1, &p, &pErr, &n, &pdf);
        if (!ok || pdf == 0 || dist_sq(p, ctxP) == 0) return ls;   // lights.cpp:803
        V3 wi = normalize(p - ctxP);
        V3 Le = area_L(l, n, -wi);
        if (!nonzero(Le)) return ls;
        ls.valid = true; ls.L = Le; ls.wi = wi; ls.pdf = pdf; ls.pLight = p; ls.pLightErr = pErr; ls.nLight = n;
        return ls;
    }
    V3 p00 = ld3(q.p00), p10 = ld3(q.p10), p01 = ld3(q.p01), p11 = ld3(q.p11);
    float pdf = 1, uvx, uvy;
    if (q.is_rect) {
        // shapes.cpp:1254-1327: rectangles that subtend enough solid angle are sampled as spherical rectangles, with the
        // cos-theta bilinear warp at surface reference points
        V3 v00 = normalize(p00 - ctxP), v10 = normalize(p10 - ctxP), v01 = normalize(p01 - ctxP), v11 = normalize(p11 - ctxP);
        if (!(spherical_quad_area(v00, v10, v11, v01) <= kMinSphericalSampleAreaQuad)) {
            if (!(ctxNs.x == 0 && ctxNs.y == 0 && ctxNs.z == 0)) {
                float w[4] = {fmaxf(0.01f, absdotn(v00, ctxNs)), fmaxf(0.01f, absdotn(v10, ctxNs)), fmaxf(0.01f, absdotn(v01, ctxNs)), fmaxf(0.01f, absdotn(v11, ctxNs))};
                float w0, w1;
                sample_bilinear(u0, u1, w, &w0, &w1);
                u0 = w0; u1 = w1;
                pdf *= bilinear_pdf(u0, u1, w);
            }
            V3 eu = p10 - p00, ev = p01 - p00;
            float quadPDF;
            V3 p = sample_spherical_rectangle(ctxP, p00, eu, ev, u0, u1, &quadPDF);
            pdf *= quadPDF;
            V3 n = normalize(cross(eu, ev));
            if (q.flip_normal) n = -n;
            if (pdf == 0 || dist_sq(p, ctxP) == 0) return ls;   // lights.cpp:803
            V3 wi = normalize(p - ctxP);
            V3 Le = area_L(l, n, -wi);
            if (!nonzero(Le)) return ls;
            ls.valid = true; ls.L = Le; ls.wi = wi; ls.pdf = pdf; ls.pLight = p; ls.pLightErr = V3(); ls.nLight = n;
            return ls;
        }
    }
    if (!q.is_rect) {
        float w[4] = {length(cross(p10 - p00, p01 - p00)), length(cross(p10 - p00, p11 - p10)),
                      length(cross(p01 - p00, p11 - p01)), length(cross(p11 - p10, p11 - p01))};
        sample_bilinear(u0, u1, w, &uvx, &uvy);
        pdf = bilinear_pdf(uvx, uvy, w);
    } else { uvx = u0; uvy = u1; }
    V3 pu0 = lerpv(uvy, p00, p01), pu1 = lerpv(uvy, p10, p11);
    V3 pLerp = lerpv(uvx, pu0, pu1);
    V3 dpdu = pu1 - pu0;
    V3 dpdv = lerpv(uvx, p01, p11) - lerpv(uvx, p00, p10);
    if (length_sq(dpdu) == 0 || length_sq(dpdv) == 0) return ls;
    V3 n = normalize(cross(dpdu, dpdv));
    if (q.flip_normal) n = -n;
    V3 p, pError;   // shapes.cpp:1200-1213 Interaction(Point3fi(p, pError), ...)
    point_interval(pLerp, gammaN(6) * (vabs(p00) + vabs(p01) + vabs(p10) + vabs(p11)), &p, &pError);
    float pdfA = pdf / length(cross(dpdu, dpdv));
    V3 wi = p - ctxP;
    if (length_sq(wi) == 0) return ls;
    wi = normalize(wi);
    float pdfW = pdfA / (absdotn(n, -wi) / dist_sq(ctxP, p));
    if (isinf(pdfW)) return ls;
    if (pdfW == 0) return ls;
    V3 Le = area_L(l, n, -wi);
    if (!nonzero(Le)) return ls;
    ls.valid = true; ls.L = Le; ls.wi = wi; ls.pdf = pdfW; ls.pLight = p; ls.pLightErr = pError; ls.nLight = n;
    return ls;
}
// lights.cpp:820-824 Light::PDF_Li -> Shape::PDF(ctx, wi): every shape first intersects ITSELF with ctx.SpawnRay(wi)
// (shapes.cpp:1337-1341, shapes.h:1133-1137, 366-370) -- not the path's own hit, whose ray may have been re-spawned at
// interfaces on the way.  `spawnO` is that ray's origin: OffsetRayOrigin(ctx.pi, ctx.n, wi), i.e. where the path left
// its previous vertex (ctxP itself for a medium vertex).
template <int KINDS = kKindsAll>
VD float area_pdf_li(const DScene &sc, int light, V3 ctxP, V3 ctxNs, V3 wi, V3 spawnO) {
    const int prim = sc.lights[light].shape_index;
    const DQuad &q = sc.quads[prim];
    Hit sh; sh.prim = -1;
    {
        float t, u = 0, v = 0, b2;
        bool ok;
        if ((KINDS & 4) && q.kind == PRIM_SPHERE) { V3 pObj; ok = intersect_sphere(spawnO, wi, inff(), ld3(q.p00), q.p10[0], &t, &pObj); }
        else if ((KINDS & 2) && q.kind == PRIM_TRIANGLE) ok = intersect_triangle(spawnO, wi, inff(), ld3(q.p00), ld3(q.p10), ld3(q.p01), &u, &v, &b2, &t);
        else ok = intersect_bilinear_patch(spawnO, wi, ld3(q.p00), ld3(q.p10), ld3(q.p01), ld3(q.p11), &u, &v, &t);
        if (!ok) return 0;
        sh.t = t; sh.u = u; sh.v = v; sh.prim = prim;
    }
    const SurfaceIntr hit = make_intr<KINDS>(sc, sh, spawnO, wi, -1);
    if ((KINDS & 2) && q.kind == PRIM_TRIANGLE) return triangle_light_pdf(q, ctxP, ctxNs, wi, hit.p, hit.n);
    if ((KINDS & 4) && q.kind == PRIM_SPHERE) return sphere_light_pdf(q, ctxP, wi, hit.p, hit.n);
    V3 p00 = ld3(q.p00), p10 = ld3(q.p10), p01 = ld3(q.p01), p11 = ld3(q.p11);
    float pdf = 1;
    if (q.is_rect) {   // shapes.cpp:1352-1368 PDF for a sample in the spherical rectangle
        V3 v00 = normalize(p00 - ctxP), v10 = normalize(p10 - ctxP), v01 = normalize(p01 - ctxP), v11 = normalize(p11 - ctxP);
        float area = spherical_quad_area(v00, v10, v11, v01);
        if (!(area <= kMinSphericalSampleAreaQuad)) {
            float sp = 1 / area;
            if (!(ctxNs.x == 0 && ctxNs.y == 0 && ctxNs.z == 0)) {
                float w[4] = {fmaxf(0.01f, absdotn(v00, ctxNs)), fmaxf(0.01f, absdotn(v10, ctxNs)), fmaxf(0.01f, absdotn(v01, ctxNs)), fmaxf(0.01f, absdotn(v11, ctxNs))};
                float uu0, uu1;
                invert_spherical_rectangle_sample(ctxP, p00, p10 - p00, p01 - p00, hit.p, &uu0, &uu1);
                return bilinear_pdf(uu0, uu1, w) * sp;
            }
            return sp;
        }
    }
    if (!q.is_rect) {
        float w[4] = {length(cross(p10 - p00, p01 - p00)), length(cross(p10 - p00, p11 - p10)),
                      length(cross(p01 - p00, p11 - p01)), length(cross(p11 - p10, p11 - p01))};
        pdf = bilinear_pdf(hit.u, hit.v, w);
    }
    V3 pu0 = lerpv(hit.v, p00, p01), pu1 = lerpv(hit.v, p10, p11);
    V3 dpdu = pu1 - pu0;
    V3 dpdv = lerpv(hit.u, p01, p11) - lerpv(hit.u, p00, p10);
    float pdfA = pdf / length(cross(dpdu, dpdv));
    float r = pdfA * (dist_sq(ctxP, hit.p) / absdotn(hit.n, -wi));
    return isinf(r) ? 0 : r;
}

struct ScatterSample { bool valid; V3 f; V3 wi; float pdf, scatPdf, misPdf; };  // surface: f; volume: f.x = p

// ------------------------------------------------------------------ GuidedPhaseFunction (cpu/guiding.h:360-638)
template <int G>
struct GuidedPhase {
    static constexpr bool GUIDED = G > 0;
    bool enableGuiding, enableScatterGuiding, ris;
    bool useGuiding, useScatterGuiding;
    SamplingDist<GUIDED ? G : 32> dist;
    float g;
    int region;

    VD float p(V3 wo, V3 wi) const { return henyey_greenstein(dot(wo, wi), g); }
    VD ScatterSample phase_sample(V3 wo, float u0, float u1) const {
        ScatterSample s; float pdf;
        s.wi = sample_henyey_greenstein(wo, g, u0, u1, &pdf);
        s.valid = true; s.f = V3(pdf, 0, 0); s.pdf = s.scatPdf = s.misPdf = pdf;
        return s;
    }
    // :383-398; `wo` is ray.d at the call site (vspg.cpp:810).  `r` = region chosen by Field::getRegion (k_lookup), -1 = none
    VD bool init_region(const DScene &sc, float g_, V3 pos, V3 wo, int r) {
        g = g_; region = -1;
        bool success = false;
        if (GUIDED) {
            if (r >= 0) {
                dist.init(&sc.vol.regions[r], pos);
                region = r;
                dist.apply_hg_product(sc.hg_table, wo, g);
                success = true;
            }
        }
        useGuiding = enableGuiding ? success : false;
        useScatterGuiding = enableScatterGuiding ? success : false;
        return success;
    }
    VD bool init(const DScene &sc, float g_, V3 pos, V3 wo, float &rnd, unsigned &lookups) {
        int r = -1;
        if (GUIDED) {
            int leaf;
            r = field_get_region(sc.vol, pos, &rnd, &leaf);
            if (leaf >= 0) lookups++;
            if (r >= 0 && !__ldg(&sc.vol.regions[r].valid)) r = -1;
        }
        return init_region(sc, g_, pos, wo, r);
    }
    VD float pdf(V3 wo, V3 wi) const {  // :542-558
        float ph = p(wo, wi);
        if (GUIDED && useGuiding) {
            float gp = dist.pdf(wi);
            return (0.5f * ph) + (0.5f * gp);  // guidingProbability = 0.5 in both the MIS and the RIS branch
        }
        return ph;
    }
    VD float vsp(V3 wi) const { if (GUIDED && useScatterGuiding) return dist.vsp(wi); return -1.f; }  // :564-574
    // :404-530.  RIS pulls Get2D + Get1D from the path sampler (:472, 509)
    VD ScatterSample sample_p(V3 wo, float u0, float u1, Rng &sampler) const {
        const float gp_ = 0.5f;
        if (!ris) {
            ScatterSample ps; ps.valid = false; bool samplePhase = true;
            if (GUIDED && useGuiding) {
                if (gp_ > u0) { u0 /= gp_; samplePhase = false; }
                else { u0 -= gp_; u0 /= (1.0f - gp_); samplePhase = true; }
            }
            if (samplePhase) {
                ps = phase_sample(wo, u0, u1);
                if (GUIDED && ps.valid && useGuiding) {
                    float gp = dist.pdf(ps.wi);
                    ps.scatPdf = ps.pdf;
                    ps.pdf = ((1.0f - gp_) * ps.pdf) + (gp_ * gp);
                    ps.misPdf = ps.pdf;
                }
            } else if (GUIDED) {
                V3 wi; float gp = dist.sample_pdf(u0, u1, &wi);
                float pv = p(wo, wi);
                if (pv > 0.f) {
                    float pdf = ((1.0f - gp_) * pv) + (gp_ * gp);
                    ps.valid = true; ps.f = V3(pv, 0, 0); ps.wi = wi; ps.pdf = pdf; ps.scatPdf = pv; ps.misPdf = pdf;
                }
            }
            return ps;
        }
        if (!(GUIDED && useGuiding)) return phase_sample(wo, u0, u1);
        ScatterSample out; out.valid = false;
        if (GUIDED) {
            const float uniformLiPDF = (float)(1.0f / (4.0f * 3.14159265358979323846));
            ScatterSample c0 = phase_sample(wo, u0, u1);
            float p0 = c0.f.x, ph0 = c0.pdf, gd0 = dist.pdf(c0.wi), li0 = dist.pdf_li(c0.wi);
            float mis0 = 0.5f * (ph0 + gd0);
            float s2a = sampler.uniform(), s2b = sampler.uniform();
            V3 wi1; float gd1 = dist.sample_pdf(s2a, s2b, &wi1);
            float li1 = dist.pdf_li(wi1);
            float p1 = p(wo, wi1), ph1 = p1;
            float mis1 = 0.5f * (ph1 + gd1);
            float w0 = 0, w1 = 0, sumW = 0; int nS = 0;
            if (ph0 > 0.f) { w0 = (ph0 * ((1.0f - gp_) * uniformLiPDF + gp_ * li0)); w0 /= mis0; sumW += w0; nS++; }
            if (ph1 > 0.f) { w1 = (ph1 * ((1.0f - gp_) * uniformLiPDF + gp_ * li1)); w1 /= mis1; sumW += w1; nS++; }
            if (nS == 0 || sumW <= 0.f) return out;
            float s1 = sumW * sampler.uniform();
            int idx = 0; float sumRis = 0.f;
            sumRis += w0;
            if (!(s1 <= sumRis)) { sumRis += w1; if (s1 <= sumRis) idx = 1; }
            float wS = idx ? w1 : w0, misS = idx ? mis1 : mis0;
            out.valid = true; out.f = V3(idx ? p1 : p0, 0, 0); out.wi = idx ? wi1 : c0.wi;
            out.pdf = (wS * misS) * (float(2) / sumW); out.scatPdf = idx ? ph1 : ph0; out.misPdf = misS;
        }
        return out;
    }
};

// ------------------------------------------------------------------ GuidedBSDF over DiffuseBxDF (cpu/guiding.h:57-358)
template <int G>
struct GuidedBSDF {
    static constexpr bool GUIDED = G > 0;
    bool enableGuiding, enableScatterGuiding, ris;
    bool useGuiding, useScatterGuiding;
    SamplingDist<GUIDED ? G : 32> dist;
    Frame frame; V3 R; bool hasR;
    int region;

    VD void set_bsdf(V3 ns, V3 dpdus, V3 refl) { frame = Frame::from_xz(normalize(dpdus), ns); R = refl; hasR = nonzero(R); }
    VD bool non_specular() const { return hasR; }
    VD V3 f(V3 woR, V3 wiR) const {  // bsdf.h:41-47 + bxdfs.h:39-43
        V3 wi = frame.to_local(wiR), wo = frame.to_local(woR);
        if (wo.z == 0) return V3();
        if (!(wo.z * wi.z > 0)) return V3();
        return R * kInvPi;
    }
    VD float bsdf_pdf(V3 woR, V3 wiR) const {
        V3 wo = frame.to_local(woR), wi = frame.to_local(wiR);
        if (wo.z == 0) return 0;
        if (!(wo.z * wi.z > 0)) return 0;
        return fabsf(wi.z) * kInvPi;
    }
    VD ScatterSample bsdf_sample(V3 woR, float u0, float u1) const {  // bsdf.h:58-80 + bxdfs.h:46-59
        ScatterSample s; s.valid = false;
        V3 wo = frame.to_local(woR);
        if (wo.z == 0 || !hasR) return s;
        V3 wi = sample_cosine_hemisphere(u0, u1);
        if (wo.z < 0) wi.z *= -1;
        float pdf = fabsf(wi.z) * kInvPi;
        V3 fv = R * kInvPi;
        if (!nonzero(fv) || pdf == 0 || wi.z == 0) return s;
        s.valid = true; s.f = fv; s.wi = frame.from_local(wi); s.pdf = s.scatPdf = s.misPdf = pdf;
        return s;
    }
    // :83-109.  `r` = region chosen by Field::getRegion (k_lookup), -1 = none / specular-only BSDF
    VD bool init_region(const DScene &sc, V3 p, V3 rayD, V3 ns, int r) {
        bool success = false; region = -1;
        if (GUIDED && non_specular()) {
            if (r >= 0) {
                dist.init(&sc.surf.regions[r], p);
                region = r;
                V3 n = ns;
                if (dotn(-rayD, ns) < 0.f) n = -n;
                dist.apply_cosine_product(n);
                success = true;
            }
        }
        useGuiding = enableGuiding ? success : false;
        useScatterGuiding = enableScatterGuiding ? success : false;
        return success;
    }
    VD bool init(const DScene &sc, V3 p, V3 rayD, V3 ns, float &rnd, unsigned &lookups) {
        int r = -1;
        if (GUIDED && non_specular()) {
            int leaf;
            r = field_get_region(sc.surf, p, &rnd, &leaf);
            if (leaf >= 0) lookups++;
            if (r >= 0 && !__ldg(&sc.surf.regions[r].valid)) r = -1;
        }
        return init_region(sc, p, rayD, ns, r);
    }
    VD float pdf(V3 wo, V3 wi) const {  // :271-289
        float b = bsdf_pdf(wo, wi);
        if (GUIDED && useGuiding) {
            float gp = dist.pdf(wi);
            return (0.5f * b) + (0.5f * gp);
        }
        return b;
    }
    VD float vsp(V3 wi) const { if (GUIDED && useScatterGuiding) return dist.vsp(wi); return -1.f; }
    VD ScatterSample sample_f(V3 wo, float u, float u0, float u1, Rng &sampler) const {  // :120-269
        const float gp_ = 0.5f;
        if (!ris) {
            ScatterSample bs; bs.valid = false; bool sampleBSDF = true;
            if (GUIDED && useGuiding) {
                if (gp_ > u) { u /= gp_; sampleBSDF = false; }
                else { u -= gp_; u /= (1.0f - gp_); sampleBSDF = true; }
            }
            if (sampleBSDF) {
                bs = bsdf_sample(wo, u0, u1);
                if (GUIDED && bs.valid && useGuiding) {
                    float gp = dist.pdf(bs.wi);
                    bs.scatPdf = bs.pdf;
                    bs.pdf = ((1.0f - gp_) * bs.pdf) + (gp_ * gp);
                    bs.misPdf = bs.pdf;
                }
            } else if (GUIDED) {
                V3 wi; float gp = dist.sample_pdf(u0, u1, &wi);
                V3 fv = f(wo, wi); float bp = bsdf_pdf(wo, wi);
                if (bp > 0.f) {
                    float pdf = ((1.0f - gp_) * bp) + (gp_ * gp);
                    bs.valid = true; bs.f = fv; bs.wi = wi; bs.pdf = pdf; bs.scatPdf = bp; bs.misPdf = pdf;
                }
            }
            return bs;
        }
        if (!(GUIDED && useGuiding)) return bsdf_sample(wo, u0, u1);
        ScatterSample out; out.valid = false;
        if (GUIDED) {
            const float uniformLiPDF = (float)(1.0f / (4.0f * 3.14159265358979323846));
            ScatterSample c0 = bsdf_sample(wo, u0, u1);
            float b0 = 0, gd0 = 0, li0 = 0, mis0 = 0;
            if (c0.valid) { b0 = c0.pdf; gd0 = dist.pdf(c0.wi); li0 = dist.pdf_li(c0.wi); mis0 = 0.5f * (b0 + gd0); }
            else { c0.f = V3(); c0.wi = V3(); }
            float s2a = sampler.uniform(), s2b = sampler.uniform();
            V3 wi1; float gd1 = dist.sample_pdf(s2a, s2b, &wi1);
            float li1 = dist.pdf_li(wi1);
            V3 f1 = f(wo, wi1); float b1 = bsdf_pdf(wo, wi1);
            float mis1 = 0.5f * (b1 + gd1);
            float w0 = 0, w1 = 0, sumW = 0; int nS = 0;
            if (b0 > 0.f) { w0 = (b0 * ((1.0f - gp_) * uniformLiPDF + gp_ * li0)); w0 /= mis0; sumW += w0; nS++; }
            if (b1 > 0.f) { w1 = (b1 * ((1.0f - gp_) * uniformLiPDF + gp_ * li1)); w1 /= mis1; sumW += w1; nS++; }
            if (nS == 0 || sumW <= 0.f) return out;
            float s1 = sumW * sampler.uniform();
            int idx = 0; float sumRis = 0.f;
            sumRis += w0;
            if (!(s1 <= sumRis)) { sumRis += w1; if (s1 <= sumRis) idx = 1; }
            float wS = idx ? w1 : w0, misS = idx ? mis1 : mis0;
            out.valid = true; out.f = idx ? f1 : c0.f; out.wi = idx ? wi1 : c0.wi;
            out.pdf = (wS * misS) * (float(2) / sumW); out.scatPdf = idx ? b1 : b0; out.misPdf = misS;
        }
        return out;
    }
};

}  // namespace vspg
