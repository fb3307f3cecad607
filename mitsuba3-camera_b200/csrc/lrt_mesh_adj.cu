// lensrt — mesh-lens parity mode: the ADJOINT of the ray trace through refractive meshes.
//
// What the notebooks obtain from Dr.Jit's reverse mode over the megakernel: a ray meets a triangle (Möller–Trumbore,
// include/mitsuba/render/mesh.h:432-452, differentiable in the vertex positions: Mesh::compute_surface_interaction,
// src/render/mesh.cpp:1370-1500 with attached positions), the shading normal is interpolated from the vertex normals
// (themselves a function of the positions: Mesh::recompute_vertex_normals, mesh.cpp:321-375), the direction refracts about it
// (fresnel / refract, include/mitsuba/render/fresnel.h:35-84,311-315; the transmission lobe of the `dispersive` / `dielectric`
// BSDF) and goes on to the next surface; d(loss) / d(vertex positions) then reaches the lens variables through the sag function
// (notebooks/scripts/lens_design/lens.py:67-85,210-237).
//
// Here: one thread per ray replays k_trace_refractive (same float32 code, same hits), keeps (o, d, triangle) of every
// interaction in a short local record, and walks it backwards. Each interaction is differentiated in closed form — in double,
// at the recorded float32 state —
//     (t, u, v)  = MT(o, d, p0, p1, p2)                   p  = o + t d
//     N          = normalize((1-u-v) n0 + u n1 + v n2)     (or the face normal when the mesh has no vertex normals)
//     ci = -d.N,  eti = ci >= 0 ? 1/eta : eta,  ct = -sign(ci) sqrt(1 - eti^2 (1 - ci^2))
//     d' = normalize(eti d + (eti ci + ct) N),  o' = p + c d'  (c: the replay's constant step off the surface),  w *= 1 - F(ci, eta)
// and scatters d(loss)/d(vertex position), d(loss)/d(vertex normal) and d(loss)/d(eta of the shape) with float atomics; the chain
// normals -> positions -> sag -> lens variables is the host's (torch autograd over MeshLens.update's own expressions).
// Discrete events are detached exactly as in the reference's AD: which triangle is hit, absorption, total internal reflection.
#include "lrt_scene.h"
#include <cmath>

namespace lrt {

namespace {

struct D3 { double x, y, z; };
__device__ inline D3 d3(double x, double y, double z) { D3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ inline D3 d3(float3 v) { return d3((double) v.x, (double) v.y, (double) v.z); }
__device__ inline D3 operator+(D3 a, D3 b) { return d3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ inline D3 operator-(D3 a, D3 b) { return d3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ inline D3 operator*(D3 a, double s) { return d3(a.x * s, a.y * s, a.z * s); }
__device__ inline double dotd(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ inline D3 crossd(D3 a, D3 b) { return d3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ inline void add3(float *g, D3 v) { atomicAdd(g, (float) v.x); atomicAdd(g + 1, (float) v.y); atomicAdd(g + 2, (float) v.z); }

constexpr int kMaxRecords = 32;

}  // namespace

__global__ void __launch_bounds__(128) k_trace_refractive_adj(SceneView sc, const int3 *__restrict__ vidx, int64_t n, const float *__restrict__ od6,
                                                               const float *__restrict__ shape_eta, int max_interactions,
                                                               const float *__restrict__ grad_out8, float *__restrict__ grad_od6,
                                                               float *__restrict__ grad_vertices, float *__restrict__ grad_normals,
                                                               float *__restrict__ grad_eta) {
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const float *r = od6 + 6 * i;
        const float3 d_in = make_float3(r[3], r[4], r[5]);
        float3 o = make_float3(r[0], r[1], r[2]), d = normalize(d_in);
        float3 rec_o[kMaxRecords], rec_d[kMaxRecords]; int rec_tri[kMaxRecords]; float rec_c[kMaxRecords];
        int K = 0; bool alive = true; float weight = 1.f;
        // ---- replay of k_trace_refractive (lrt_refit.cu), statement for statement
        for (int k = 0; k < max_interactions; ++k) {
            unsigned work = 0;
            const RayHit h = bvh_intersect(sc.bvh, o, d, kLargest, false, work);
            if (h.tri < 0) break;
            const Surface si = make_surface(sc, h, o, d);
            const float eta = __ldg(shape_eta + si.shape);
            if (!(eta > 0.f)) { alive = false; break; }
            float r_i, ct, eit, eti;
            fresnel_dielectric(si.wi.z, eta, r_i, ct, eit, eti);
            if (r_i >= 1.f) { alive = false; break; }
            weight *= 1.f - r_i;
            rec_o[K] = o; rec_d[K] = d; rec_tri[K] = h.tri;
            const float3 wo = make_float3(-eti * si.wi.x, -eti * si.wi.y, ct);
            d = normalize(si.to_world(wo));
            const float c = kRayEpsilon * 0.01f * (1.f + fmaxf(fabsf(si.p.x), fmaxf(fabsf(si.p.y), fabsf(si.p.z))));
            rec_c[K] = c; ++K;
            o = si.p + d * c;
        }
        float *go = grad_od6 + 6 * i;
        if (!alive) { for (int j = 0; j < 6; ++j) go[j] = 0.f; continue; }      // a dead ray's outputs are constants (weight 0)
        const float *g8 = grad_out8 + 8 * i;
        D3 ao = d3((double) g8[0], (double) g8[1], (double) g8[2]), ad = d3((double) g8[3], (double) g8[4], (double) g8[5]);
        const double aw_total = (double) g8[7] * (double) weight;              // d loss / d log(weight)
        // ---- reverse sweep
        for (int k = K - 1; k >= 0; --k) {
            const D3 o_k = d3(rec_o[k]), d_k = d3(rec_d[k]);
            const int tri = rec_tri[k];
            const float4 *tp = sc.bvh.tris + 3 * (size_t) tri;
            const float4 ta = __ldg(tp), tb = __ldg(tp + 1), tc = __ldg(tp + 2);
            const D3 p0 = d3((double) ta.x, (double) ta.y, (double) ta.z), p1 = d3((double) ta.w, (double) tb.x, (double) tb.y), p2 = d3((double) tb.z, (double) tb.w, (double) tc.x);
            const int flags = __float_as_int(tc.w), shape = __float_as_int(tc.y);
            const double eta = (double) __ldg(shape_eta + shape);
            // Möller–Trumbore
            const D3 e1 = p1 - p0, e2 = p2 - p0, pvec = crossd(d_k, e2);
            const double det = dotd(e1, pvec), inv = 1.0 / det;
            const D3 tvec = o_k - p0, qvec = crossd(tvec, e1);
            const double U = dotd(tvec, pvec), V = dotd(d_k, qvec), T = dotd(e2, qvec);
            const double u = U * inv, v = V * inv, t = T * inv;
            // normal
            const bool smooth = (flags & 1) && sc.tri_attr != nullptr;
            D3 n0, n1, n2, hvec;
            if (smooth) {
                const float4 *ap = sc.tri_attr + 4 * (size_t) tri;
                const float4 a0 = __ldg(ap), a1 = __ldg(ap + 1), a2 = __ldg(ap + 2);
                n0 = d3((double) a0.x, (double) a0.y, (double) a0.z); n1 = d3((double) a0.w, (double) a1.x, (double) a1.y); n2 = d3((double) a1.z, (double) a1.w, (double) a2.x);
                hvec = n0 * (1.0 - u - v) + n1 * u + n2 * v;
            } else
                hvec = crossd(e1, e2);
            const double hl = sqrt(dotd(hvec, hvec));
            const D3 N = hvec * (1.0 / hl);
            // refraction
            const double ci = -dotd(d_k, N);
            const bool outside = ci >= 0.0;
            const double eti = outside ? 1.0 / eta : eta, eit = outside ? eta : 1.0 / eta;
            const double S = sqrt(fmax(1.0 - eti * eti * (1.0 - ci * ci), 1e-30)), sgn = outside ? 1.0 : -1.0, ct = -sgn * S;
            const double m = eti * ci + ct;
            const D3 g = d_k * eti + N * m;
            const double gl = sqrt(dotd(g, g));
            const D3 dn = g * (1.0 / gl);
            // ---- adjoints: o' = p + c d'
            const double c = (double) rec_c[k];
            ad = ad + ao * c;
            const D3 ap_ = ao;                                       // adjoint of p
            // d' = normalize(g)
            const D3 ag = (ad - dn * dotd(dn, ad)) * (1.0 / gl);
            D3 ad_k = ag * eti, aN = ag * m;
            const double am = dotd(N, ag);
            double aci = eti * am, aeti = dotd(ag, d_k) + am * ci;
            const double act = am;
            aci += act * (-sgn * eti * eti * ci / S);
            aeti += act * (sgn * eti * (1.0 - ci * ci) / S);
            // weight: w_k = 1 - F, F = (a_s^2 + a_p^2) / 2 in (|ci|, S, eit)
            if (aw_total != 0.0 && eta != 1.0) {
                const double cia = fabs(ci), cta = S;
                const double ds = cia + eit * cta, dp = cta + eit * cia;
                const double a_s = (cia - eit * cta) / ds, a_p = (cta - eit * cia) / dp;
                const double F = 0.5 * (a_s * a_s + a_p * a_p);
                const double aF = -aw_total / (1.0 - F);
                const double aas = aF * a_s, aap = aF * a_p;
                // a_s = (cia - eit cta) / ds: d/dcia = (1 - a_s) / ds, d/dcta = -eit (1 + a_s) / ds, d/deit = -cta (1 + a_s) / ds
                const double acia = aas * (1.0 - a_s) / ds + aap * (-eit * (1.0 + a_p)) / dp;
                const double acta = aas * (-eit * (1.0 + a_s)) / ds + aap * (1.0 - a_p) / dp;
                const double aeit = aas * (-cta * (1.0 + a_s)) / ds + aap * (-cia * (1.0 + a_p)) / dp;
                aci += acia * (ci >= 0.0 ? 1.0 : -1.0);
                // cta = S(ci, eti)
                aci += acta * (eti * eti * ci / S);
                aeti += acta * (-eti * (1.0 - ci * ci) / S);
                // eit = 1 / eti
                aeti += aeit * (-1.0 / (eti * eti));
            }
            if (grad_eta) atomicAdd(grad_eta + shape, (float) (outside ? -aeti / (eta * eta) : aeti));
            // ci = -d.N
            ad_k = ad_k - N * aci; aN = aN - d_k * aci;
            // N = normalize(hvec)
            const D3 ah = (aN - N * dotd(N, aN)) * (1.0 / hl);
            double au = 0.0, av = 0.0;
            D3 ae1 = d3(0, 0, 0), ae2 = d3(0, 0, 0);
            if (smooth) {
                au = dotd(ah, n1 - n0); av = dotd(ah, n2 - n0);
                if (grad_normals) {
                    const int3 f = vidx[tri];
                    add3(grad_normals + 3 * (size_t) f.x, ah * (1.0 - u - v)); add3(grad_normals + 3 * (size_t) f.y, ah * u); add3(grad_normals + 3 * (size_t) f.z, ah * v);
                }
            } else {                                                 // hvec = e1 x e2
                ae1 = crossd(e2, ah); ae2 = crossd(ah, e1);
            }
            // p = o + t d
            D3 ao_k = ap_;
            ad_k = ad_k + ap_ * t;
            const double at = dotd(d_k, ap_);
            // (t, u, v) = (T, U, V) / det
            const double aU = au * inv, aV = av * inv, aT = at * inv;
            const double adet = -(au * U + av * V + at * T) * inv * inv;
            D3 atvec = pvec * aU;
            D3 apvec = tvec * aU + e1 * adet;
            ae1 = ae1 + pvec * adet;
            ad_k = ad_k + qvec * aV;
            const D3 aqvec = d_k * aV + e2 * aT;
            ae2 = ae2 + qvec * aT;
            // qvec = tvec x e1, pvec = d x e2        (c = a x b: adj a = b x adj c, adj b = adj c x a)
            atvec = atvec + crossd(e1, aqvec); ae1 = ae1 + crossd(aqvec, tvec);
            ad_k = ad_k + crossd(e2, apvec); ae2 = ae2 + crossd(apvec, d_k);
            ao_k = ao_k + atvec;
            if (grad_vertices) {
                const int3 f = vidx[tri];
                add3(grad_vertices + 3 * (size_t) f.x, d3(0, 0, 0) - atvec - ae1 - ae2);
                add3(grad_vertices + 3 * (size_t) f.y, ae1); add3(grad_vertices + 3 * (size_t) f.z, ae2);
            }
            ao = ao_k; ad = ad_k;
        }
        // d = normalize(d_in)
        const D3 din = d3(d_in);
        const double dl = sqrt(dotd(din, din));
        const D3 dn0 = din * (1.0 / dl);
        const D3 adin = (ad - dn0 * dotd(dn0, ad)) * (1.0 / dl);
        go[0] = (float) ao.x; go[1] = (float) ao.y; go[2] = (float) ao.z; go[3] = (float) adin.x; go[4] = (float) adin.y; go[5] = (float) adin.z;
    }
}

}  // namespace lrt

using namespace lrt;
extern "C" {

int lrt_scene_trace_refractive_backward(const lrt_scene *s, int64_t n, const float *od6_dev, const float *shape_eta_dev, int32_t max_interactions,
                                        const float *grad_out8_dev, float *grad_od6_dev, float *grad_vertices_dev, float *grad_normals_dev,
                                        float *grad_eta_dev, void *stream) {
    LRT_NEED_DEVICE();
    LRT_REQUIRE(s && n >= 0 && max_interactions >= 0 && (n == 0 || (od6_dev && shape_eta_dev && grad_out8_dev && grad_od6_dev)), "bad argument");
    LRT_REQUIRE(max_interactions <= kMaxRecords, "lrt_scene_trace_refractive_backward keeps at most 32 interactions per ray");
    LRT_REQUIRE(s->view.n_spheres == 0, "lrt_scene_trace_refractive_backward: triangle meshes only");
    if (n == 0 || s->n_tris == 0) {
        if (n) LRT_CUDA(cudaMemsetAsync(grad_od6_dev, 0, (size_t) n * 6 * sizeof(float), (cudaStream_t) stream));
        return LRT_OK;
    }
    int64_t g = (n + 127) / 128; if (g > 148 * 32) g = 148 * 32;
    k_trace_refractive_adj<<<(int) g, 128, 0, (cudaStream_t) stream>>>(s->view, (const int3 *) s->d_tri_vidx, n, od6_dev, shape_eta_dev, max_interactions,
                                                                      grad_out8_dev, grad_od6_dev, grad_vertices_dev, grad_normals_dev, grad_eta_dev);
    LRT_LAUNCHED();
    LRT_CUDA(cudaGetLastError());
    return LRT_OK;
}

}  // extern "C"
