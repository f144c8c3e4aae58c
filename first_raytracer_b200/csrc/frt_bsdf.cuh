// BSDFs and sampling densities of the reference's materials, float32, branch on a type tag
// instead of virtual calls + a heap-allocated pdf object per bounce (material.h, pdf.h,
// microfacet.h, onb.h, util.h).  Each function cites what it restates; quirks are kept
// (un-normalised onb, no face-forwarding, signed cosines) because they decide pixel values.
#pragma once
#include "frt_device.cuh"

struct Surface // the fields of hit_record (hitable.h:23-36) that shading reads
{
    float3 p, wi, n;
    float t;
};

struct Onb // onb.h:18-30 — w is NOT normalised
{
    float3 u, v, w;
    __device__ __forceinline__ explicit Onb(float3 n)
    {
        w = n;
        if (fabsf(n.x) > fabsf(n.y))
        {
            const float inv = 1.0f / sqrtf(n.x * n.x + n.z * n.z);
            v = f3(n.z * inv, 0.0f, -n.x * inv);
        }
        else
        {
            const float inv = 1.0f / sqrtf(n.y * n.y + n.z * n.z);
            v = f3(0.0f, n.z * inv, -n.y * inv);
        }
        u = cross(v, w);
    }
    __device__ __forceinline__ float3 from_local(float3 a) const { return a.x * u + a.y * v + a.z * w; }
    __device__ __forceinline__ float3 to_local(float3 a) const { return f3(dot(a, u), dot(a, v), dot(a, w)); }
};

__device__ __forceinline__ float safe_sqrt(float x) { return fmaxf(0.0f, sqrtf(x)); }          // util.h:43-46
__device__ __forceinline__ float mi_weight(float a, float b) { a *= a; b *= b; return a / (a + b); } // util.h:55-60
__device__ __forceinline__ float3 reflect_u(float3 v, float3 n) { return unit_vector(v - 2.0f * dot(v, n) * n); } // util.h:73-76

// util.h:79-84
__device__ __forceinline__ float3 refract_u(float3 wi, float3 n, float eta, float cosThetaT)
{
    if (cosThetaT < 0.0f) eta = 1.0f / eta;
    return unit_vector(n * (dot(wi, n) * eta + cosThetaT) - wi * eta);
}

// util.h:86-118
__device__ __forceinline__ float fresnel_dielectric_ext(float cosThetaI_, float &cosThetaT_, float eta)
{
    if (eta == 1.0f) { cosThetaT_ = -cosThetaI_; return 0.0f; }
    const float scale = (cosThetaI_ > 0.0f) ? 1.0f / eta : eta;
    const float cosThetaTSqr = 1.0f - (1.0f - cosThetaI_ * cosThetaI_) * (scale * scale);
    if (cosThetaTSqr <= 0.0f) { cosThetaT_ = 0.0f; return 1.0f; }
    const float cosThetaI = fabsf(cosThetaI_);
    const float cosThetaT = sqrtf(cosThetaTSqr);
    const float Rs = (cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT);
    const float Rp = (eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT);
    cosThetaT_ = (cosThetaI_ > 0.0f) ? -cosThetaT : cosThetaT;
    return 0.5f * (Rs * Rs + Rp * Rp);
}

// microfacet.h:8-31
__device__ __forceinline__ float3 fresnel_conductor_exact(float cosThetaI, float3 eta, float3 k)
{
    const float c2 = cosThetaI * cosThetaI, s2 = 1.0f - c2, s4 = s2 * s2;
    const float e[3] = {eta.x, eta.y, eta.z}, kk[3] = {k.x, k.y, k.z};
    float out[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        const float temp1 = e[i] * e[i] - kk[i] * kk[i] - s2;
        const float a2pb2 = safe_sqrt(temp1 * temp1 + kk[i] * kk[i] * e[i] * e[i] * 4.0f);
        const float a = safe_sqrt((a2pb2 + temp1) * 0.5f);
        const float term1 = a2pb2 + c2, term2 = a * (2.0f * cosThetaI);
        const float Rs2 = (term1 - term2) / (term1 + term2);
        const float term3 = a2pb2 * c2 + s4, term4 = term2 * s2;
        const float Rp2 = Rs2 * (term3 - term4) / (term3 + term4);
        out[i] = 0.5f * (Rp2 + Rs2);
    }
    return f3(out[0], out[1], out[2]);
}

// util.h:241-254
__device__ __forceinline__ float hypot2(float a, float b)
{
    float r;
    if (fabsf(a) > fabsf(b)) { r = b / a; r = fabsf(a) * sqrtf(1.0f + r * r); }
    else if (b != 0.0f) { r = a / b; r = fabsf(b) * sqrtf(1.0f + r * r); }
    else r = 0.0f;
    return r;
}

// microfacet.h:47-88 (alphaU == alphaV always: rough_conductor's ctor, material.h:251)
__device__ __forceinline__ float smith_g1(float3 v, float3 m, float3 n, float alpha, int beckmann)
{
    const float cosTheta = dot(n, v);
    if (dot(v, m) * cosTheta <= 0.0f) return 0.0f;
    const float temp = 1.0f - cosTheta * cosTheta;
    if (temp <= 0.0f) return 1.0f;
    const float tanTheta = sqrtf(temp) / cosTheta;
    if (beckmann)
    {
        const float a = 1.0f / (alpha * tanTheta);
        if (a >= 1.6f) return 1.0f;
        const float aSqr = a * a;
        return (3.535f * a + 2.181f * aSqr) / (1.0f + 2.276f * a + 2.577f * aSqr);
    }
    return 2.0f / (1.0f + hypot2(1.0f, alpha * tanTheta));
}

// microfacet.h:90-128
__device__ __forceinline__ float mf_eval(float3 m, float3 n, float alpha, int beckmann)
{
    const float3 ml = Onb(n).to_local(m);
    const float cosTheta = ml.z;
    if (cosTheta <= 0.0f) return 0.0f;
    const float cos2 = cosTheta * cosTheta;
    const float ex = ((ml.x * ml.x) / (alpha * alpha) + (ml.y * ml.y) / (alpha * alpha)) / cos2;
    float result;
    if (beckmann)
        result = expf(-ex) / (FRT_PI * alpha * alpha * cos2 * cos2);
    else
    {
        const float root = (1.0f + ex) * cos2;
        result = 1.0f / (FRT_PI * alpha * alpha * root * root);
    }
    if (result * cosTheta < 1e-20f) result = 0.0f;
    return result;
}

// util.h:185-221
__device__ __forceinline__ float erfinv_giles(float x)
{
    float w = -logf((1.0f - x) * (1.0f + x));
    float p;
    if (w < 5.0f)
    {
        w = w - 2.5f;
        p = 2.81022636e-08f;
        p = 3.43273939e-07f + p * w;
        p = -3.5233877e-06f + p * w;
        p = -4.39150654e-06f + p * w;
        p = 0.00021858087f + p * w;
        p = -0.00125372503f + p * w;
        p = -0.00417768164f + p * w;
        p = 0.246640727f + p * w;
        p = 1.50140941f + p * w;
    }
    else
    {
        w = sqrtf(w) - 3.0f;
        p = -0.000200214257f;
        p = 0.000100950558f + p * w;
        p = 0.00134934322f + p * w;
        p = -0.00367342844f + p * w;
        p = 0.00573950773f + p * w;
        p = -0.0076224613f + p * w;
        p = 0.00943887047f + p * w;
        p = 1.00167406f + p * w;
        p = 2.83297682f + p * w;
    }
    return p * x;
}

// util.h:223-239
__device__ __forceinline__ float erf_as(float x)
{
    const float sign = copysignf(1.0f, x);
    x = fabsf(x);
    const float t = 1.0f / (1.0f + 0.3275911f * x);
    const float y = 1.0f - (((((1.061405429f * t + -1.453152027f) * t) + 1.421413741f) * t + -0.284496736f) * t + 0.254829592f) * t * expf(-x * x);
    return sign * y;
}

// pdf.h:280-403 (Heitz visible-normal slopes for alpha = 1)
__device__ __forceinline__ float2 sample_visible_11(float thetaI, float sx, float sy, int beckmann)
{
    const float SQRT_PI_INV = 0.56418958354775628f;
    float2 slope;
    if (beckmann)
    {
        if (thetaI < 1e-4f)
        {
            const float r = sqrtf(-logf(1.0f - sx));
            float sp, cp;
            sincospif(2.0f * sy, &sp, &cp);
            return make_float2(r * cp, r * sp);
        }
        const float tanThetaI = tanf(thetaI), cotThetaI = 1.0f / tanThetaI;
        float a = -1.0f, c = erf_as(cotThetaI);
        const float sample_x = fmaxf(sx, 1e-6f);
        const float fit = 1.0f + thetaI * (-0.876f + thetaI * (0.4265f - 0.0594f * thetaI));
        float b = c - (1.0f + c) * powf(1.0f - sample_x, fit);
        const float normalization = 1.0f / (1.0f + c + SQRT_PI_INV * tanThetaI * expf(-cotThetaI * cotThetaI));
        int it = 0;
        while (++it < 10)
        {
            if (!(b >= a && b <= c)) b = 0.5f * (a + c);
            const float invErf = erfinv_giles(b);
            const float value = normalization * (1.0f + b + SQRT_PI_INV * tanThetaI * expf(-invErf * invErf)) - sample_x;
            const float derivative = normalization * (1.0f - invErf * tanThetaI);
            if (fabsf(value) < 1e-5f) break;
            if (value > 0.0f) c = b; else a = b;
            b -= value / derivative;
        }
        slope.x = erfinv_giles(b);
        slope.y = erfinv_giles(2.0f * fmaxf(sy, 1e-6f) - 1.0f);
        return slope;
    }
    if (thetaI < 1e-4f)
    {
        const float r = safe_sqrt(sx / (1.0f - sx));
        float sp, cp;
        sincospif(2.0f * sy, &sp, &cp);
        return make_float2(r * cp, r * sp);
    }
    const float tanThetaI = tanf(thetaI);
    const float a = 1.0f / tanThetaI;
    const float G1 = 2.0f / (1.0f + safe_sqrt(1.0f + 1.0f / (a * a)));
    float A = 2.0f * sx / G1 - 1.0f;
    if (fabsf(A) == 1.0f) A -= copysignf(1.0f, A) * FRT_EPSILON;
    const float tmp = 1.0f / (A * A - 1.0f);
    const float B = tanThetaI;
    const float D = safe_sqrt(B * B * tmp * tmp - (A * A - B * B) * tmp);
    const float slope_x_1 = B * tmp - D, slope_x_2 = B * tmp + D;
    slope.x = (A < 0.0f || slope_x_2 > 1.0f / tanThetaI) ? slope_x_1 : slope_x_2;
    float S;
    if (sy > 0.5f) { S = 1.0f; sy = 2.0f * (sy - 0.5f); }
    else { S = -1.0f; sy = 2.0f * (0.5f - sy); }
    const float z = (sy * (sy * (sy * (-0.365728915865723f) + 0.790235037209296f) - 0.424965825137544f) + 0.000152998850436920f) /
                    (sy * (sy * (sy * (sy * 0.169507819808272f - 0.397203533833404f) - 0.232500544458471f) + 1.0f) - 0.539825872510702f);
    slope.y = S * z * sqrtf(1.0f + slope.x * slope.x);
    return slope;
}

// pdf.h:405-453
__device__ __forceinline__ float3 sample_visible(float3 wi_local, float alpha, float sx, float sy, int beckmann)
{
    const float3 wi = unit_vector(f3(alpha * wi_local.x, alpha * wi_local.y, wi_local.z));
    float theta = 0.0f, phi = 0.0f;
    if (wi.z < 0.99999f) { theta = acosf(wi.z); phi = atan2f(wi.y, wi.x); }
    float sinPhi, cosPhi;
    sincosf(phi, &sinPhi, &cosPhi);
    float2 slope = sample_visible_11(theta, sx, sy, beckmann);
    if (!isfinite(slope.x)) slope.x = 0.0f;
    slope = make_float2(cosPhi * slope.x - sinPhi * slope.y, sinPhi * slope.x + cosPhi * slope.y);
    slope.x *= alpha;
    slope.y *= alpha;
    const float normalization = 1.0f / sqrtf(slope.x * slope.x + slope.y * slope.y + 1.0f);
    return f3(-slope.x * normalization, -slope.y * normalization, normalization);
}

struct ScatterRec // scatter_record (pdf.h:60-70) without the heap pdf object
{
    bool is_specular;
    float3 dir;        // specular_ray direction (specular materials only)
    float sampled_pdf; // -1 unless rough_conductor set it
};

// material::scatter for every type that scatters (diffuse_light returns false: handled by caller).
// `d_in` is r_in.direction() (un-normalised), s0/s1 the first two components of get3d().
__device__ __forceinline__ void mat_scatter(const frt_material &m, const Surface &h, float3 d_in, float s0, float s1,
                                            ScatterRec &srec)
{
    srec.sampled_pdf = -1.0f;
    srec.dir = f3(0.f, 0.f, 0.f);
    switch (m.type)
    {
    case FRT_MAT_LAMBERTIAN: // material.h:55-60
        srec.is_specular = false;
        break;
    case FRT_MAT_PHONG: // material.h:82-93 + cosine_power_pdf::generate pdf.h:116-131
    {
        srec.is_specular = true;
        const float3 R = reflect_u(-h.wi, h.n);
        const float e = m.p[6];
        const float sinAlpha = sqrtf(1.0f - powf(s1, 2.0f / (e + 1.0f)));
        const float cosAlpha = powf(s1, 1.0f / (e + 1.0f));
        float sp, cp;
        sincospif(2.0f * s0, &sp, &cp);
        srec.dir = Onb(R).from_local(f3(sinAlpha * cp, sinAlpha * sp, cosAlpha));
        break;
    }
    case FRT_MAT_DIELECTRIC: // material.h:139-146 + dielectric_pdf::generate pdf.h:163-180
    {
        srec.is_specular = true;
        float cosThetaT;
        const float F = fresnel_dielectric_ext(dot(h.wi, h.n), cosThetaT, m.p[6]);
        srec.dir = (s0 <= F) ? reflect_u(-h.wi, h.n) : refract_u(h.wi, h.n, m.p[6], cosThetaT);
        break;
    }
    case FRT_MAT_ROUGH_CONDUCTOR: // material.h:262-270 + roughconductor_pdf::generate pdf.h:464-480
    {
        srec.is_specular = true;
        const float alpha = m.p[6];
        const int beck = m.p[10] != 0.0f;
        const Onb uvw(h.n);
        const float3 wi_l = uvw.to_local(h.wi);
        const float3 ml = sample_visible(wi_l, alpha, s0, s1, beck);
        const float3 mw = uvw.from_local(ml);
        // pdfVisible, pdf.cpp:5-12
        float pdf = 0.0f;
        if (wi_l.z != 0.0f)
            pdf = smith_g1(uvw.from_local(wi_l), mw, h.n, alpha, beck) * fabsf(dot(wi_l, ml)) *
                  mf_eval(mw, h.n, alpha, beck) / fabsf(wi_l.z);
        const float3 wo = reflect_u(-h.wi, mw);
        srec.sampled_pdf = pdf / (4.0f * dot(wo, mw));
        srec.dir = unit_vector(wo);
        break;
    }
    case FRT_MAT_METAL: // material.h:115-122
    default:
        srec.is_specular = true;
        srec.dir = reflect_u(unit_vector(d_in), h.n);
        break;
    }
}

// srec.pdf_ptr->value(hrec, wo) for the pdf object each material's scatter creates
__device__ __forceinline__ float mat_pdf_value(const frt_material &m, const Surface &h, float3 wo)
{
    switch (m.type)
    {
    case FRT_MAT_LAMBERTIAN: // cosine_pdf::value pdf.h:84-89
        return fmaxf(dot(h.n, unit_vector(wo)), 0.0f) / FRT_PI;
    case FRT_MAT_PHONG: // cosine_power_pdf::value pdf.h:104-114
    {
        if (dot(h.n, wo) <= 0.0f || dot(h.n, h.wi) <= 0.0f) return 0.0f;
        const float alpha = fmaxf(0.0f, dot(reflect_u(-h.wi, h.n), wo));
        return powf(alpha, m.p[6]) * (m.p[6] + 1.0f) / (2.0f * FRT_PI);
    }
    case FRT_MAT_DIELECTRIC: // dielectric_pdf::value pdf.h:142-161
    {
        float cosThetaT;
        const float F = fresnel_dielectric_ext(dot(h.wi, h.n), cosThetaT, m.p[6]);
        if (dot(h.wi, h.n) * dot(wo, h.n) >= 0.0f)
        {
            if (fabsf(dot(reflect_u(-h.wi, h.n), wo) - 1.0f) > FRT_DELTA_EPSILON) return 0.0f;
            return F;
        }
        if (fabsf(dot(refract_u(h.wi, h.n, m.p[6], cosThetaT), wo) - 1.0f) > FRT_DELTA_EPSILON) return 0.0f;
        return 1.0f - F;
    }
    case FRT_MAT_ROUGH_CONDUCTOR: // roughconductor_pdf::value pdf.h:237-250
    {
        if (dot(h.n, wo) <= 0.0f || dot(h.n, h.wi) <= 0.0f) return 0.0f;
        const float3 H = unit_vector(wo + h.wi);
        const int beck = m.p[10] != 0.0f;
        return mf_eval(H, h.n, m.p[6], beck) * smith_g1(h.wi, H, h.n, m.p[6], beck) / (4.0f * dot(h.wi, h.n));
    }
    default: // constant_pdf(1.0) pdf.h:186-201
        return 1.0f;
    }
}

// material::eval_bsdf(r_in, hrec, wo) — r_in is not read by any of them
__device__ __forceinline__ float3 mat_eval(const frt_material &m, const Surface &h, float3 wo)
{
    switch (m.type)
    {
    case FRT_MAT_LAMBERTIAN: // material.h:62-65
        return arr3(m.p) / FRT_PI;
    case FRT_MAT_PHONG: // material.h:95-103
    {
        const float3 reflected = unit_vector(reflect_u(-h.wi, h.n));
        const float alpha = fmaxf(0.0f, dot(reflected, wo));
        const float e = m.p[6];
        const float3 result = (arr3(m.p) / FRT_PI) + arr3(m.p + 3) * ((e + 2.0f) / (2.0f * FRT_PI)) * powf(alpha, e);
        return result * dot(h.n, wo);
    }
    case FRT_MAT_DIELECTRIC: // material.h:147-172
    {
        float cosThetaT;
        const float ior = m.p[6];
        const float F = fresnel_dielectric_ext(dot(h.wi, h.n), cosThetaT, ior);
        if (dot(h.wi, h.n) * dot(wo, h.n) >= 0.0f)
        {
            if (fabsf(dot(reflect_u(-h.wi, h.n), wo) - 1.0f) > FRT_DELTA_EPSILON) return f3(0.f, 0.f, 0.f);
            return arr3(m.p) * F;
        }
        if (fabsf(dot(refract_u(h.wi, h.n, ior, cosThetaT), wo) - 1.0f) > FRT_DELTA_EPSILON) return f3(0.f, 0.f, 0.f);
        const float factor = cosThetaT < 0.0f ? (1.0f / ior) : ior;
        return arr3(m.p) * factor * factor * (1.0f - F);
    }
    case FRT_MAT_ROUGH_CONDUCTOR: // material.h:276-306
    {
        const float cosWi = dot(h.wi, h.n);
        if (cosWi <= 0.0f || dot(wo, h.n) <= 0.0f) return f3(0.f, 0.f, 0.f);
        const float3 H = unit_vector(wo + h.wi);
        const float alpha = m.p[6];
        const int beck = m.p[10] != 0.0f;
        const float D = mf_eval(H, h.n, alpha, beck);
        if (D == 0.0f) return f3(0.f, 0.f, 0.f);
        const float3 F = fresnel_conductor_exact(dot(h.wi, H), arr3(m.p + 3), arr3(m.p + 7)) * arr3(m.p);
        const float G = smith_g1(h.wi, H, h.n, alpha, beck) * smith_g1(wo, H, h.n, alpha, beck);
        return F * (D * G / (4.0f * cosWi));
    }
    case FRT_MAT_METAL: // material.h:123-126
        return arr3(m.p);
    default:
        return f3(0.f, 0.f, 0.f);
    }
}
