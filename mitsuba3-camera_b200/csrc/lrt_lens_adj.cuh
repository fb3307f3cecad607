// lensrt — reverse-mode derivative of the lens step: one interface at a time, hand-derived.
//
// What is differentiated: the ray transfer of LensInterface::compute_interaction
// (realisticlens.cpp:269-297) for SpheroidLens (:359-402), PlanoLens / ApertureStop (:444-517) and
// AsphericalLens (:574-819), with respect to the incoming ray AND the interface's own parameters
// (axial position z, curvature radius R | unitless curvature c, conic K, a0..a6, the two refractive indices).
// The reference never differentiates this code (its lens tables are constants; its notebooks differentiate
// mesh stand-ins through Dr.Jit's tape, notebooks/scripts/lens_design/lens.py:67-85,210-237). Here there is no
// tape: the forward trace stores the ray entering each interface (6 floats), and the reverse sweep recomputes
// the local geometry from it. The hit distance t is an implicit function of (o, d, params):
//     F(o + t d; params) = 0   =>   dt = -(gradF . (do + t dd) + F_param dparam) / (gradF . d)
// so neither the sphere's root selection nor the asphere's Newton iterations are differentiated.
// Masks (vignetting, TIR) are piecewise constant: samples keep the mask of the strict forward trace.
//
// Arithmetic here is relaxed float32 (FFMA allowed): gradients are compared within a tolerance, not bit-exact.
#pragma once
#include "lrt_lens.cuh"

namespace lrt {

// gradient slots of one interface, same positions as the rows16 layout of include/lensrt.h
enum : int { G_Z = 1, G_R = 3, G_C = 3, G_K = 4, G_A0 = 6, G_IORL = 13, G_IORR = 14, kGradRow = 16 };

struct AdjRay { float3 o, d; };   // adjoints of a ray's origin and direction

// Reverse of one interface. (o, d): the ray that ENTERED the interface in the forward trace (taped); p: the hit point
// the forward trace found (= the origin of the ray that left the interface, so it is on the tape as well and neither
// the sphere's root nor the asphere's Newton solve is repeated); `bar` holds the adjoint of the ray that LEFT the
// interface and is overwritten with the adjoint of the ray that entered; g[16] accumulates the parameter adjoints.
LRT_HD void lens_interact_adj(const LensIf &f, float3 o, float3 d, float3 p, AdjRay &bar, float *g) {
    // ---- local geometry at the hit point: distance t, surface gradient gF (F = 0 on the surface), unit normal n as used
    const float t = dot(p - o, d); float3 n, gF;
    float u = 0.f, w = 1.f, Sp = 0.f, Spp = 0.f, inv_m = 1.f;   // asphere / normalisation temporaries
    if (f.kind == IF_SPHERE) {
        float3 c = make_float3(0.f, 0.f, f.center_z);
        float3 m = p - c;
        inv_m = 1.f / sqrtf(dot(m, m));
        n = m * (f.convex ? inv_m : -inv_m);
        gF = m * 2.f;
    } else if (f.kind == IF_ASPHERE) {
        u = (p.x * p.x + p.y * p.y) * f.rcp_R2;
        // S(u) = P(u) u^2 + c u / (1 + w), w = sqrt(1 - k1 u);  S' = P' u^2 + 2 u P + c / (2 w)
        float P = 0.f, Pp = 0.f, Ppp = 0.f;
        for (int i = 6; i >= 0; --i) { Ppp = Ppp * u + 2.f * Pp; Pp = Pp * u + P; P = P * u + f.ai[i]; }
        w = sqrtf(fmaxf(1.f - f.k1 * u, 1e-20f));
        Sp = Pp * u * u + 2.f * u * P + f.c / (2.f * w);
        Spp = Ppp * u * u + 4.f * u * Pp + 2.f * P + f.c * f.k1 / (4.f * w * w * w);
        float k = -2.f * f.rcp_R * Sp;                 // sag_x = k x, sag_y = k y
        float3 m = make_float3(k * p.x, k * p.y, -1.f);
        inv_m = 1.f / sqrtf(dot(m, m));
        n = m * inv_m;
        gF = make_float3(-m.x, -m.y, 1.f);
    } else {                                           // plane z = f.z
        n = make_float3(0.f, 0.f, -1.f);
        gF = make_float3(0.f, 0.f, 1.f);
    }

    float3 p_bar = bar.o, d_bar = make_float3(0.f, 0.f, 0.f), n_bar = make_float3(0.f, 0.f, 0.f);
    if (f.kind == IF_APERTURE) {
        d_bar = bar.d;                                 // direction passes through unchanged (:505-517)
    } else {
        // ---- refraction: d' = n g + d eti, g = ci eti + ct, ci = -d.n, ct = -sign(ci) sqrt(1 - eti^2 (1 - ci^2))
        float ci = -dot(d, n);
        bool outside = ci >= 0.f;
        float eti = outside ? f.rcp_eta : f.eta;
        float q = 1.f - (1.f - ci * ci) * eti * eti;
        float ct = sqrtf(fmaxf(q, 1e-20f)); if (outside) ct = -ct;
        float gg = ci * eti + ct;
        float3 dp = bar.d;
        n_bar = dp * gg;
        float g_bar = dot(n, dp);
        d_bar = dp * eti;
        float eti_bar = dot(d, dp) + g_bar * ci;
        float ci_bar = g_bar * eti;
        float q_bar = g_bar / (2.f * ct);              // d ct / d q = 1 / (2 ct) (sign included)
        ci_bar += q_bar * 2.f * eti * eti * ci;
        eti_bar += q_bar * (-2.f * eti * (1.f - ci * ci));
        d_bar = d_bar - n * ci_bar;
        n_bar = n_bar - d * ci_bar;
        // eti = ior_left / ior_right (outside) or ior_right / ior_left
        if (outside) { g[G_IORL] += eti_bar / f.ior_right; g[G_IORR] -= eti_bar * eti / f.ior_right; }
        else         { g[G_IORR] += eti_bar / f.ior_left;  g[G_IORL] -= eti_bar * eti / f.ior_left; }
        // ---- normal
        if (f.kind == IF_SPHERE) {
            float s = f.convex ? inv_m : -inv_m;
            float3 m_bar = (n_bar - n * dot(n, n_bar)) * s;
            p_bar += m_bar;
            g[G_Z] -= m_bar.z; g[G_R] -= m_bar.z;      // c.z = z + R
        } else if (f.kind == IF_ASPHERE) {
            float3 m_bar = (n_bar - n * dot(n, n_bar)) * inv_m;
            // m.x = k x, m.y = k y with k = -(2/Re) S'(u), u = (x^2 + y^2) / Re^2
            float k = -2.f * f.rcp_R * Sp;
            float xy = m_bar.x * p.x + m_bar.y * p.y;
            float kk = -2.f * f.rcp_R * xy;            // adjoint of S'
            float du = kk * Spp * 2.f * f.rcp_R2;      // via u
            p_bar.x += m_bar.x * k + du * p.x;
            p_bar.y += m_bar.y * k + du * p.y;
            float w3 = w * w * w;
            g[G_C] += kk * (1.f / (2.f * w) + f.one_p_K * f.c2 * u / (2.f * w3));
            g[G_K] += kk * (f.c2 * f.c * u / (4.f * w3));
            float up = u;                              // dS'/da_i = (i + 2) u^(i+1)
            for (int i = 0; i < 7; ++i) { g[G_A0 + i] += kk * (float) (i + 2) * up; up *= u; }
        }
    }
    // ---- p = o + t d, t implicit
    float3 o_bar = p_bar;
    d_bar += p_bar * t;
    float t_bar = dot(p_bar, d);
    float s = -t_bar / dot(gF, d);
    o_bar += gF * s;
    d_bar += gF * (s * t);
    if (f.kind == IF_SPHERE) {
        float mz = gF.z;                               // 2 (p - c).z ; F = |p - c|^2 - R^2
        g[G_Z] += s * (-mz);
        g[G_R] += s * (-mz - 2.f * f.curv_radius);
    } else if (f.kind == IF_ASPHERE) {
        // F = p.z - z0 + Re S(u; c, K, a)
        float Re = f.elem_radius, opw = 1.f + w;
        g[G_Z] += s * (-1.f);
        g[G_C] += s * Re * (u / opw + f.one_p_K * f.c2 * u * u / (w * opw * opw));
        g[G_K] += s * Re * (f.c2 * f.c * u * u / (2.f * w * opw * opw));
        float up = u * u;
        for (int i = 0; i < 7; ++i) { g[G_A0 + i] += s * Re * up; up *= u; }
    } else {
        g[G_Z] += s * (-1.f);
    }
    bar.o = o_bar; bar.d = d_bar;
}

// Strict forward trace front -> rear that also tapes the ray entering each interface (tape[i] for interface i).
LRT_HD bool lens_trace_from_world_taped(const LensIf *ifs, int n, LensRay &ray, float (*tape)[6]) {
    for (int i = n - 1; i >= 0; --i) {
        tape[i][0] = ray.o.x; tape[i][1] = ray.o.y; tape[i][2] = ray.o.z;
        tape[i][3] = ray.d.x; tape[i][4] = ray.d.y; tape[i][5] = ray.d.z;
        if (!lens_interact(ifs[i], ray)) return false;
    }
    return true;
}

}  // namespace lrt
