// One path vertex of the reference's RayCast (src/win32_main.cpp:558-823): material + texture fetch,
// Schlick-metal Fresnel, Hammon Smith-joint G, Burley alpha^2 = r^4, GGX half-vector sampling, cosine
// hemisphere and emissive-sphere cone sampling with the reference's 50/50 estimator and pdf mixture --
// every quirk of SURVEY.md Appendix A.6 kept, because the image oracle is the reference itself.
#pragma once
#ifndef PT_TEX_INLINE
#define PT_TEX_INLINE __device__ __forceinline__
#endif
#include "pt_device.cuh"
#include "pt_trace.cuh"

// SampleTexture / BespokeSampleTexture, main.cpp:1675-1709: uv = hit.xy * size * 0.5, |.|, integer part
// modulo size, bilinear on texels = byte/255 (main.cpp:1737-1739).  Only mip 0 is ever read.
__device__ __forceinline__ float3 pt_texel(const PtTexture &t, int x, int y) {
    uchar4 p = __ldg(t.texels + (size_t)y * t.w + x);
    // byte / 255.f must be the IEEE quotient (bit-identical texels); 12 divisions per bilinear fetch were 10 % of
    // the w1 kernel's instructions, so the 256 quotients come from a table the host filled with the same division
    return f3(__ldg(t.unit255 + p.x), __ldg(t.unit255 + p.y), __ldg(t.unit255 + p.z));
}
__device__ __forceinline__ float3 lerp3(float3 a, float3 b, float p) {
    float q = 1.0f - p;
    return f3(q * a.x + p * b.x, q * a.y + p * b.y, q * a.z + p * b.z);
}
PT_TEX_INLINE float3 pt_sample_texture(const PtTexture &t, float hx, float hy) {
    float u = fabsf(hx * (float)t.w * 0.5f), v = fabsf(hy * (float)t.h * 0.5f);
    int x1 = (int)u, y1 = (int)v;
    float s = fminf(1.f, fmaxf(u - (float)x1, 0.f));
    float tt = fminf(1.f, fmaxf(v - (float)y1, 0.f));
    x1 = t.wmask ? (x1 & t.wmask) : (x1 % t.w);
    y1 = t.hmask ? (y1 & t.hmask) : (y1 % t.h);
    int x2 = x1 + 1 == t.w ? 0 : x1 + 1;
    int y2 = y1 + 1 == t.h ? 0 : y1 + 1;
    float3 top = lerp3(pt_texel(t, x1, y1), pt_texel(t, x2, y1), s);
    float3 bot = lerp3(pt_texel(t, x1, y2), pt_texel(t, x2, y2), s);
    return lerp3(top, bot, tt);
}

// BuildOrthonormalBasisFromW, main.cpp:2288-2305
__device__ __forceinline__ void pt_onb(float3 w, float3 &tx, float3 &ty, float3 &tz) {
    float3 uw = normalize3(w);
    float3 A = (fabsf(uw.x) > 0.9f) ? f3(0, 1, 0) : f3(1, 0, 0);
    float3 v = normalize3(cross3(uw, A));
    tx = cross3(uw, v);
    ty = v;
    tz = uw;
}

// SchlickMetal, main.cpp:1752-1756 (powf(1-c,5) as three multiplies)
__device__ __forceinline__ float3 pt_schlick_metal(float F0, float c, float metalness, float3 metal) {
    float q = 1.f - metalness;
    float3 vF0 = f3(q * F0 + metalness * metal.x, q * F0 + metalness * metal.y, q * F0 + metalness * metal.z);
    float m = 1.f - c;
    float m2 = m * m;
    float p5 = m2 * m2 * m;
    return f3(vF0.x + p5 * (1.f - vF0.x), vF0.y + p5 * (1.f - vF0.y), vF0.z + p5 * (1.f - vF0.z));
}

// PdfValue<TO_SPHERE_PDF>, main.cpp:324-340: 1/solid angle if the direction hits the light, else 0
__device__ __forceinline__ float pt_pdf_to_sphere(float3 dir, float4 light, float3 from) {
    float t = pt_sphere_t(from, dir, pt_dot_exact(dir), light);
    if (!(t > 0.f)) return 0.f;
    float3 rel = f3(from.x - light.x, from.y - light.y, from.z - light.z);
    float cmax = sqrtf(1.f - pt_fdiv(light.w * light.w, dot3(rel, rel)));
    return pt_fdiv(1.f, 2.f * PT_PI * (1.f - cmax));
}

struct PtSurface {            // what RayCast reads from material_t for a non-emissive hit
    float ior, metalness, roughness;
    float3 albedo, metal;
    int albedo_tex, metalness_tex, roughness_tex, normal_tex;
};

__device__ __forceinline__ void pt_load_surface(const PtDeviceScene &S, uint32_t mat, PtSurface &m) {
    const float *p = reinterpret_cast<const float *>(S.materials + mat);   // 17 words, include/pt_api.h
    const int *q = reinterpret_cast<const int *>(p);
    m.ior = __ldg(p + 1);
    m.albedo_tex = __ldg(q + 2);
    m.albedo = f3(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
    m.metalness_tex = __ldg(q + 6);
    m.metalness = __ldg(p + 7);
    m.metal = f3(__ldg(p + 8), __ldg(p + 9), __ldg(p + 10));
    m.roughness_tex = __ldg(q + 11);
    m.roughness = __ldg(p + 12);
    m.normal_tex = __ldg(q + 13);
}
__device__ __forceinline__ float3 pt_load_emit(const PtDeviceScene &S, uint32_t mat) {
    const float *p = reinterpret_cast<const float *>(S.materials + mat);
    return f3(__ldg(p + 14), __ldg(p + 15), __ldg(p + 16));
}

// Shade the vertex at o + t*d with geometric normal n.  On return true the path continues along (o, d)
// with beta multiplied by 2/px * brdfTerm (main.cpp:780-781); false = no bounce (main.cpp:648, 733, 748).
template <uint32_t F>
__device__ __forceinline__ bool pt_shade_vertex(const PtDeviceScene &S, PtRng &rng, uint32_t mat_index, float3 n, float t,
                                                float3 &o, float3 &d, float3 &beta) {
    PtSurface m;
    pt_load_surface(S, mat_index, m);

    float cos_theta = dot3(n, d);
    cos_theta = (cos_theta > 0.f) ? -cos_theta : cos_theta;                 // main.cpp:596-598
    float f = pt_fdiv(1.003f - m.ior, 1.003f + m.ior);
    const float F0 = f * f;                                                  // main.cpp:600
    const float3 P = f3(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z));
    const float3 V = -d;

    float metalness = m.metalness, roughness = m.roughness;
    float3 N = n;
    if (F & PT_F_TEX) {
        if (m.metalness_tex != 0 && (S.flags & PT_SCENE_METALNESS_MAPS)) metalness = pt_sample_texture(S.tex[m.metalness_tex - 1], P.x, P.y).x;
        if (m.roughness_tex != 0 && (S.flags & PT_SCENE_ROUGHNESS_MAPS)) roughness = pt_sample_texture(S.tex[m.roughness_tex - 1], P.x, P.y).x;
        if (m.normal_tex != 0 && (S.flags & PT_SCENE_NORMAL_MAPS)) {
            float3 c = pt_sample_texture(S.tex[m.normal_tex - 1], P.x, P.y);
            N = normalize3(f3(2.f * c.x - 1.f, 2.f * c.y - 1.f, 2.f * c.z - 1.f));   // world-space, main.cpp:637-643
        }
    }

    const float NdotV = dot3(N, V);
    if (!(NdotV > 0.f)) return false;                                        // main.cpp:647-648

    float3 tx, ty, tz;
    pt_onb(N, tx, ty, tz);
    const bool smooth = roughness < 0.01f;                                   // EffectivelySmooth, main.cpp:1783-1786
    const float4 light = S.light;
    const bool use_light = (F & PT_F_LIGHT) && !(S.flags & PT_SCENE_JUST_COSINE);   // main.cpp:654-655

    float3 L, H = f3(0, 0, 0);
    float px;
    bool specular;
    int guard = 0;
    for (;;) {
        if (++guard > 64) return false;      // the reference would spin (origin inside the light); give up
        px = 1.f;                            // both specular branches set it every iteration (main.cpp:674, 730)
        const uint4 u = rng.next_block();    // one Philox block per iteration: coin, then [coin, r1, r2] or [z1, z2]
        specular = u32_to_unit(u.x) > 0.5f;                                  // main.cpp:661
        if (specular && smooth) {
            L = d - (2.0f * cos_theta) * n;                                  // mirror about the GEOMETRIC normal, main.cpp:608
        } else if (!specular) {
            bool sample_cosine = u32_to_unit(u.y) > 0.5f;                    // main.cpp:678
            float3 r;
            float sn, cs;
            sincospif(2.f * u32_to_unit(u.z), &sn, &cs);
            const float r2 = u32_to_unit(u.w);
            if (!use_light || sample_cosine) {
                // RandomCosineDirectionHemisphere, main.cpp:2258-2268
                float sr = sqrtf(r2);
                r = f3(cs * sr, sn * sr, sqrtf(1.f - r2));
            } else {
                // RandomToSphere, main.cpp:2332-2353; the basis is re-aimed at the light and STAYS so
                float3 dirl = f3(light.x - P.x, light.y - P.y, light.z - P.z);
                float term1 = 1.f - pt_fdiv(light.w * light.w, dot3(dirl, dirl));
                if (term1 < 0.f) {
                    r = f3(0, 0, 0);
                } else {
                    float z = 1.f + r2 * (sqrtf(term1) - 1.f);
                    float st = sqrtf(1.f - z * z);
                    r = f3(cs * st, sn * st, z);
                }
                pt_onb(dirl, tx, ty, tz);
            }
            if (r.x == 0.f && r.y == 0.f && r.z == 0.f) continue;            // main.cpp:700
            L = normalize3(f3(r.x * tx.x + r.y * ty.x + r.z * tz.x, r.x * tx.y + r.y * ty.y + r.z * tz.y,
                              r.x * tx.z + r.y * ty.z + r.z * tz.z));
            H = normalize3(L + V);
            // cosine pdf of the LOCAL sample (for light samples: measured from the cone axis, main.cpp:709)
            float pc = fmaxf(0.f, r.z * pt_rsqrt(dot3(r, r)) * (1.0f / PT_PI));
            if (use_light) px = 0.5f * pc + 0.5f * pt_pdf_to_sphere(L, light, P);
            else px = pc;
            if (px == 0.f) continue;                                         // main.cpp:722
        } else {
            // RandomHalfVectorGGX, main.cpp:2271-2285: cos(theta) = sqrt((1-z2)/(1+z2(a2-1)))
            float a2 = roughness * roughness * roughness * roughness;
            float z2 = u32_to_unit(u.z);
            float ct = sqrtf(pt_fdiv(1.f - z2, 1.f + z2 * (a2 - 1.f)));
            float st = sqrtf(fmaxf(0.f, 1.f - ct * ct));
            float sn, cs;
            sincospif(2.f * u32_to_unit(u.y), &sn, &cs);
            float3 r = f3(cs * st, sn * st, ct);
            H = normalize3(f3(r.x * tx.x + r.y * ty.x + r.z * N.x, r.x * tx.y + r.y * ty.y + r.z * N.y,
                              r.x * tx.z + r.y * ty.z + r.z * N.z));
            L = (2.f * dot3(V, H)) * H - V;
        }
        break;
    }

    const float NdotL = dot3(N, L);
    if (!(NdotL > 0.f)) return false;                                        // main.cpp:733

    float3 ks;
    if (smooth) {
        ks = pt_schlick_metal(F0, NdotL, metalness, m.metal);                // main.cpp:738-739
    } else {
        float HdotL = dot3(H, L);
        if (dot3(H, V) > 0.f && HdotL > 0.f) ks = pt_schlick_metal(F0, HdotL, metalness, m.metal);
        else return false;                                                   // main.cpp:747-749
    }

    float3 brdf;
    if (specular && smooth) {
        brdf = ks;                                                           // main.cpp:761-764
    } else if (specular) {
        // ks * Hammon G * |H.L| / (|N.L| |H.N|), main.cpp:1610-1620, 1773-1781
        float a2 = roughness * roughness * roughness * roughness;
        float denom = NdotV * sqrtf(a2 + (1.f - a2) * NdotL * NdotL) + NdotL * sqrtf(a2 + (1.f - a2) * NdotV * NdotV);
        float G = pt_fdiv(2.f * NdotL * NdotV, denom);
        float wgt = G * pt_fdiv(fabsf(dot3(H, L)), fabsf(NdotL) * fabsf(dot3(H, N)));
        brdf = wgt * ks;
    } else {
        // N.L * kd * albedo / pi with kd = (1 - ks)(1 - metalness), main.cpp:751-759, 772, 1595-1608
        float3 albedo = m.albedo;
        if ((F & PT_F_TEX) && m.albedo_tex != 0) albedo = pt_sample_texture(S.tex[m.albedo_tex - 1], P.x, P.y);
        float q = (1.f - metalness) * NdotL * (1.0f / PT_PI);
        brdf = f3(q * (1.f - ks.x) * albedo.x, q * (1.f - ks.y) * albedo.y, q * (1.f - ks.z) * albedo.z);
    }
    const float wpx = 2.f * pt_fdiv(1.f, px);                             // main.cpp:780
    beta = f3(beta.x * wpx * brdf.x, beta.y * wpx * brdf.y, beta.z * wpx * brdf.z);
    o = P;
    d = L;
    return true;
}

// RayCast as a loop: L = sum_k beta_k * emit_k over at most 4 vertices (main.cpp:564, 799)
template <uint32_t F>
__device__ __forceinline__ float3 pt_trace_path(const PtDeviceScene &S, PtRng &rng, float3 o, float3 d) {
    float3 beta = f3(1, 1, 1), L = f3(0, 0, 0);
    for (int depth = 0; depth < PT_MAX_BOUNCE_COUNT; depth++) {
        PtHit h = pt_closest_hit<F>(S, o, d);
        float3 n;
        uint32_t mat;
        pt_hit_surface<F>(S, h, o, d, n, mat);
        float3 e = pt_load_emit(S, mat);
        L = L + beta * e;
        bool emissive = (e.x != 0.f) || (e.y != 0.f) || (e.z != 0.f);
        if (mat == 0 || emissive || depth == PT_MAX_BOUNCE_COUNT - 1) break;   // main.cpp:579-590
        if (!pt_shade_vertex<F>(S, rng, mat, n, h.t, o, d, beta)) break;
    }
    return L;
}
