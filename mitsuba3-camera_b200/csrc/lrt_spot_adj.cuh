// lensrt — the adjoint of ONE lens connection (a reference point seen through the lens and splatted): shared by the spot
// kernels (lrt_spot_kernel.cuh) and the particle tracer's lens adjoint (lrt_ptrace.cu). Compiled with whatever arithmetic
// flavour the including translation unit selects (lrt_math.cuh).
#pragma once
#include "lrt_internal.h"
#include "lrt_lens_adj.cuh"

namespace lrt {

constexpr int kSlots = 12;          // per-interface gradient slots kept per thread: z, R|c, K, a0..a6, ior_l, ior_r

// derivative of the reconstruction filter (same polynomial / tent as filter_eval)
LRT_HD float filter_eval_d(const FilmParams &F, float x) {
    if (F.kind == FILTER_GAUSSIAN) {
        const float *c = F.coeff;
        float x2 = x * x, P = 0.f, Pp = 0.f;
        for (int i = 9; i >= 0; --i) { Pp = Pp * x2 + P; P = P * x2 + c[i]; }
        return P > 0.f ? 2.f * x * Pp : 0.f;
    }
    if (F.kind == FILTER_TENT) {
        float ax = fabsf(x * F.inv_radius);
        return (ax < 1.f && x != 0.f) ? (x > 0.f ? -F.inv_radius : F.inv_radius) : 0.f;
    }
    return 0.f;
}

// Non-coalesced footprint (imageblock.cpp:264-412) with per-axis NORMALISED weights nx, ny (W = nx_i ny_j) and
// their derivatives with respect to the sample position. Returns false if the footprint is empty.
struct FootprintD { int x0, y0, nx, ny; float wx[kMaxTaps], wy[kMaxTaps], dwx[kMaxTaps], dwy[kMaxTaps]; };

LRT_HD bool spot_footprint(const FilmParams &F, float px, float py, FootprintD &fp, bool want_d) {
    int sx = F.size_x + 2 * F.border, sy = F.size_y + 2 * F.border;
    float pfx = px + ((float) (F.border - F.off_x) - .5f), pfy = py + ((float) (F.border - F.off_y) - .5f);
    float p0x = pfx - F.radius, p0y = pfy - F.radius;
    int cx = (int) ceilf(p0x), cy = (int) ceilf(p0y);
    int x0 = max(cx, 0), y0 = max(cy, 0);
    int x1 = min((int) floorf(pfx + F.radius), sx - 1), y1 = min((int) floorf(pfy + F.radius), sy - 1);
    if (x0 > x1 || y0 > y1) return false;
    fp.x0 = x0; fp.y0 = y0; fp.nx = min(x1 - x0 + 1, F.count_max); fp.ny = min(y1 - y0 + 1, F.count_max);
    float a = 1.f, b = 1.f, da = 0.f, db = 0.f;
    if (F.normalize) {   // :318-342: sums start at the unclipped window
        a = b = 0.f;
        float rx = (float) cx - pfx, ry = (float) cy - pfy;
        for (int i = 0; i < F.count_max; ++i) {
            a += filter_eval(F, rx); b += filter_eval(F, ry);
            if (want_d) { da -= filter_eval_d(F, rx); db -= filter_eval_d(F, ry); }
            rx += 1.f; ry += 1.f;
        }
    }
    float ia = a != 0.f ? 1.f / a : 0.f, ib = b != 0.f ? 1.f / b : 0.f;
    if (F.normalize && (a == 0.f || b == 0.f)) { ia = 0.f; ib = 0.f; }
    float rx = (float) x0 - pfx, ry = (float) y0 - pfy;
    for (int i = 0; i < F.count_max; ++i) {
        float fx = filter_eval(F, rx), fy = filter_eval(F, ry);
        fp.wx[i] = fx * ia; fp.wy[i] = fy * ib;
        if (want_d) {
            fp.dwx[i] = (-filter_eval_d(F, rx) - fp.wx[i] * da) * ia;
            fp.dwy[i] = (-filter_eval_d(F, ry) - fp.wy[i] * db) * ib;
        }
        rx += 1.f; ry += 1.f;
    }
    return true;
}

// Replays sample_direction for the camera-space reference point `src_p` and aperture sample (sa, sb) with a tape (identical
// masks to lens_sample_direction), transposes the normalised splat of  intensity * weight * scale  (RGB) and  w_value  (the
// weight channel) against grad_accum [sy + 2b, sx + 2b, 4], and sweeps the lens in reverse: ADDS d(loss)/d(interface entries) to
// g_loc[interface][slot]. Returns false when the sample is masked out (nothing added).
LRT_D bool lens_connection_adjoint(const LensSystem &L, const FilmParams &F, strict::V3 src_p, float sa, float sb, float origin_offset,
                                   float i0, float i1, float i2, float scale, float w_value, const float *__restrict__ grad_accum,
                                   float (*g_loc)[kSlots]) {
    const int n_if = L.n;
    const int sw = F.size_x + 2 * F.border;
    // ---- strict forward replay with a tape (identical masks to the forward kernel)
    using namespace strict;
    if (!(src_p.z >= L.near_clip && src_p.z <= L.far_clip)) return false;
    const LensIf &front = L.ifs[n_if - 1];
    float dx, dy; square_to_disk_concentric(sa, sb, dx, dy);
    V3 ap = v3(mul(dx, front.elem_radius), mul(dy, front.elem_radius), front.z);
    V3 dirv = sub(src_p, ap);
    float dist = norm(dirv), inv_dist = rcp(dist);
    V3 dir = mul(dirv, inv_dist);
    LensRay ray; ray.o = origin_offset > 0.f ? at(ap, dir, origin_offset) : src_p; ray.d = neg(dir);
    float tape[kMaxInterfaces][6];
    if (!lens_trace_from_world_taped(L.ifs, n_if, ray, tape)) return false;
    float t = div(-ray.o.z, ray.d.z);
    V3 pf = at(ray.o, ray.d, t);
    float sx = fmadd(L.f2s_m00, pf.x, L.f2s_m03), sy = fmadd(L.f2s_m11, pf.y, L.f2s_m13);
    if (!(sx >= 0.f && sy >= 0.f && sx <= 1.f && sy <= 1.f)) return false;
    const float ct_film = -ray.d.z;
    const float value = L.normalization / (ct_film * ct_film * ct_film * ct_film) * (L.rear_z * L.rear_z);
    const float wgt = value * inv_dist * inv_dist * dir.z;
    const float uvx = sx * L.res_x, uvy = sy * L.res_y;
    FootprintD fp;
    if (!spot_footprint(F, uvx + (float) F.off_x, uvy + (float) F.off_y, fp, true)) return false;
    // ---- transpose of the splat: adjoint of the sample weight and of the splat position
    const float wv = wgt * scale;
    const float v0 = i0 * wv, v1 = i1 * wv, v2 = i2 * wv, v3 = w_value;
    float S0 = 0.f, S1 = 0.f, S2 = 0.f, px_bar = 0.f, py_bar = 0.f;
    for (int iy = 0; iy < fp.ny; ++iy) {
        const int y = fp.y0 + iy;
        for (int ix = 0; ix < fp.nx; ++ix) {
            const int x = fp.x0 + ix;
            const float4 G = __ldg(reinterpret_cast<const float4 *>(grad_accum) + (size_t) y * sw + x);
            const float w = fp.wx[ix] * fp.wy[iy];
            S0 += G.x * w; S1 += G.y * w; S2 += G.z * w;
            const float gv = G.x * v0 + G.y * v1 + G.z * v2 + G.w * v3;
            px_bar += gv * fp.dwx[ix] * fp.wy[iy];
            py_bar += gv * fp.wx[ix] * fp.dwy[iy];
        }
    }
    const float w_bar = scale * (i0 * S0 + i1 * S1 + i2 * S2);
    // uv = s * res ; s = f2s * pf
    const float pfx_bar = px_bar * L.res_x * L.f2s_m00, pfy_bar = py_bar * L.res_y * L.f2s_m11;
    // pf = o + t d, t = -o.z / d.z
    const float3 o = f3(ray.o), d = f3(ray.d);
    AdjRay bar;
    bar.o = make_float3(pfx_bar, pfy_bar, 0.f);
    bar.d = make_float3(pfx_bar * t, pfy_bar * t, 0.f);
    const float t_bar = pfx_bar * d.x + pfy_bar * d.y;
    bar.o.z += -t_bar / d.z;
    bar.d.z += -t_bar * t / d.z;
    // wgt = value * dir.z / dist^2 ; value = norm * (-d.z)^-4 * rear_z^2
    const float value_bar = w_bar * inv_dist * inv_dist * dir.z;
    bar.d.z += value_bar * (-4.f * value / d.z);
    g_loc[0][0] += value_bar * 2.f * value / L.rear_z;            // rear_z = ifs[0].z
    float dirz_bar = w_bar * value * inv_dist * inv_dist;
    float dist_bar = w_bar * value * dir.z * (-2.f) * inv_dist * inv_dist * inv_dist;
    // ---- reverse sweep rear -> front
    for (int k = 0; k < n_if; ++k) {
        float g[kGradRow];
#pragma unroll
        for (int q = 0; q < kGradRow; ++q) g[q] = 0.f;
        // the hit point on interface k is the origin of the ray that entered interface k - 1 (or the final ray)
        const float3 hp = k > 0 ? make_float3(tape[k - 1][0], tape[k - 1][1], tape[k - 1][2]) : o;
        lens_interact_adj(L.ifs[k], make_float3(tape[k][0], tape[k][1], tape[k][2]),
                          make_float3(tape[k][3], tape[k][4], tape[k][5]), hp, bar, g);
        g_loc[k][0] += g[G_Z]; g_loc[k][1] += g[G_R]; g_loc[k][2] += g[G_K];
#pragma unroll
        for (int q = 0; q < 7; ++q) g_loc[k][3 + q] += g[G_A0 + q];
        g_loc[k][10] += g[G_IORL]; g_loc[k][11] += g[G_IORR];
    }
    // ---- the ray's start: o0 = source (fixed), d0 = -dir, dir = (src - ap) / dist, ap.z = front.z
    // (with an origin offset: o0 = ap + offset * dir, the same line through ap)
    float3 dir_bar = make_float3(-bar.d.x, -bar.d.y, -bar.d.z + dirz_bar);
    float apz_bar = 0.f;
    if (origin_offset > 0.f) { dir_bar += bar.o * origin_offset; apz_bar = bar.o.z; }
    const float3 dirf = f3(dir);
    float3 v_bar = (dir_bar - dirf * dot(dirf, dir_bar)) * inv_dist + dirf * dist_bar;
    g_loc[n_if - 1][0] += apz_bar - v_bar.z;
    return true;
}

}  // namespace lrt
