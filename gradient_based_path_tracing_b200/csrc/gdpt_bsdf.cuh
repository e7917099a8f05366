// BSDF device functions: eval (BSDF x cosine), pdf_sample_bsdf (solid angle) and
// sample_bsdf for the nine lajolla material variants, dispatched by a switch on
// DMaterial::type (the reference uses std::visit over a std::variant,
// /root/reference/src/material.cpp:90-119).  Each block cites the .inl it mirrors;
// reference quirks are kept on purpose (see DESIGN.md "quirks kept"):
//   microfacet helpers     src/microfacet.h:25-177
//   Lambertian             src/materials/lambertian.inl:1-50
//   RoughPlastic           src/materials/roughplastic.inl:1-171
//   RoughDielectric        src/materials/roughdielectric.inl:1-182
//   DisneyDiffuse          src/materials/disney_diffuse.inl:1-106
//   DisneyMetal            src/materials/disney_metal.inl:1-186   (pdf does not clamp roughness)
//   DisneyGlass            src/materials/disney_glass.inl:1-238
//   DisneyClearcoat        src/materials/disney_clearcoat.inl:1-113 (pdf divides by |n.out|)
//   DisneySheen            src/materials/disney_sheen.inl:1-89
//   DisneyBSDF             src/materials/disney_bsdf.inl:1-197    (uniform 1/4 lobe pick on uv[0],
//                          weighted-mixture pdf without sheen, glass lobe gets the already
//                          inverted eta and inverts it again)
#pragma once
#include "gdpt_scene.cuh"

namespace gdpt {

template <typename R> struct BsdfSample { V3<R> dir_out; R eta, roughness; };

template <typename R> GD_D R pw(R x, R e) { return pow(x, e); }
template <typename R> GD_D R sq(R x) { return pow(x, R(2)); }   // the .inl files write pow(x,2)

template <typename R> GD_D V3<R> sample_cos_hemisphere(R u0, R u1) {   // material.cpp:4-11
    R phi = Consts<R>::TWO_PI * u0;
    R tmp = sqrt(clampr(R(1) - u1, R(0), R(1)));
    return V3<R>(cos(phi) * tmp, sin(phi) * tmp, sqrt(clampr(u1, R(0), R(1))));
}

template <typename R> GD_D R fresnel_dielectric2(R n_dot_i, R n_dot_t, R eta) {
    R rs = (n_dot_i - eta * n_dot_t) / (n_dot_i + eta * n_dot_t);
    R rp = (eta * n_dot_i - n_dot_t) / (eta * n_dot_i + n_dot_t);
    return (rs * rs + rp * rp) / 2;
}
template <typename R> GD_D R fresnel_dielectric(R n_dot_i, R eta) {
    R n_dot_t_sq = R(1) - (R(1) - n_dot_i * n_dot_i) / (eta * eta);
    if (n_dot_t_sq < R(0)) return R(1);
    R n_dot_t = sqrt(n_dot_t_sq);
    return fresnel_dielectric2(fabs(n_dot_i), n_dot_t, eta);
}
template <typename R> GD_D R GTR2(R n_dot_h, R roughness) {
    R alpha = roughness * roughness;
    R a2 = alpha * alpha;
    R t = R(1) + (a2 - R(1)) * n_dot_h * n_dot_h;
    return a2 / (Consts<R>::PI * t * t);
}
// GTR2 at the half vector h.  double: the reference's expression on n.h, literally.  float: 1 + (a2 - 1) cos^2 loses
// everything to cancellation once a2 < 2^-24 (roughness 0.01 gives a2 = 1e-8), so the same number is formed from the
// local components, sin^2 + a2 cos^2, which has no subtraction.
template <typename R> GD_D R GTR2_at(const Frame<R> &f, const V3<R> &h, R roughness) {
    if (sizeof(R) == 8) return GTR2(dot(f.n, h), roughness);
    V3<R> hl = to_local(f, h);
    R alpha = roughness * roughness;
    R a2 = alpha * alpha;
    R t = hl.x * hl.x + hl.y * hl.y + a2 * hl.z * hl.z;
    return a2 / (Consts<R>::PI * t * t);
}
template <typename R> GD_D R smith_masking_gtr2(const V3<R> &v_local, R roughness) {
    R alpha = roughness * roughness;
    R a2 = alpha * alpha;
    V3<R> v2 = v_local * v_local;
    R Lambda = (R(-1) + sqrt(R(1) + (v2.x * a2 + v2.y * a2) / v2.z)) / 2;
    return R(1) / (R(1) + Lambda);
}
// sample_visible_normals / _custom (alpha_x == alpha_y == alpha for the plain version)
template <typename R> GD_D V3<R> sample_visible_normals(V3<R> local_dir_in, R alpha_x, R alpha_y, R u0, R u1) {
    bool flip = local_dir_in.z < R(0);
    if (flip) local_dir_in = -local_dir_in;
    V3<R> hemi_dir_in = normalize(V3<R>(alpha_x * local_dir_in.x, alpha_y * local_dir_in.y, local_dir_in.z));
    R r = sqrt(u0);
    R phi = R(2) * Consts<R>::PI * u1;
    R t1 = r * cos(phi);
    R t2 = r * sin(phi);
    R s = (R(1) + hemi_dir_in.z) / 2;
    t2 = (R(1) - s) * sqrt(R(1) - t1 * t1) + s * t2;
    V3<R> disk_N(t1, t2, sqrt(rmax(R(0), R(1) - t1 * t1 - t2 * t2)));
    Frame<R> hemi_frame(hemi_dir_in);
    V3<R> hemi_N = to_world(hemi_frame, disk_N);
    V3<R> n = normalize(V3<R>(alpha_x * hemi_N.x, alpha_y * hemi_N.y, rmax(R(0), hemi_N.z)));
    return flip ? -n : n;
}
template <typename R> GD_D V3<R> sample_clearcoat(R alpha, R u0, R u1) {   // microfacet.h:162-177
    R h_azim = R(2) * Consts<R>::PI * u1;
    R sin_e = pw((pw(alpha * alpha, R(1) - u0) - alpha * alpha) / (R(1) - alpha * alpha), R(0.5));
    R cos_e = pw((R(1) - pw(alpha * alpha, R(1) - u0)) / (R(1) - alpha * alpha), R(0.5));
    return normalize(V3<R>(sin_e * cos(h_azim), sin_e * sin(h_azim), cos_e));
}

template <typename R> struct BCtx {
    const DScene<R> &sc;
    const Vertex<R> &v;
    V3<R> dir_in;
    GD_D V3<R> t3(int id) const { return eval_tex3<R>(sc, id, v.uv, v.uv_screen_size); }
    GD_D R t1(int id) const { return eval_tex1<R>(sc, id, v.uv, v.uv_screen_size); }
};

template <typename R> GD_D Frame<R> flip_frame_one_sided(const Vertex<R> &v, const V3<R> &dir_in) {
    Frame<R> f = v.sf;
    if (dot(f.n, dir_in) < R(0)) f = -f;
    return f;
}
template <typename R> GD_D Frame<R> flip_frame_two_sided(const Vertex<R> &v, const V3<R> &dir_in) {
    Frame<R> f = v.sf;
    if (dot(f.n, dir_in) * dot(v.gn, dir_in) < R(0)) f = -f;
    return f;
}

// ------------------------------------------------------------------ Lambertian / cosine lobes
template <typename R> GD_D R pdf_cosine(const BCtx<R> &c, const V3<R> &dir_out) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return R(0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    return fmax(dot(f.n, dir_out), R(0)) / Consts<R>::PI;
}
template <typename R> GD_D bool sample_cosine(const BCtx<R> &c, R u0, R u1, R roughness, BsdfSample<R> &s) {
    if (dot(c.v.gn, c.dir_in) < R(0)) return false;
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    s.dir_out = to_world(f, sample_cos_hemisphere<R>(u0, u1));
    s.eta = R(0); s.roughness = roughness;
    return true;
}
template <typename R> GD_D V3<R> eval_lambertian(const BCtx<R> &c, const V3<R> &dir_out, int tex) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return V3<R>(0, 0, 0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    return fmax(dot(f.n, dir_out), R(0)) * c.t3(tex) / Consts<R>::PI;
}

// ------------------------------------------------------------------ RoughPlastic
template <typename R> GD_D V3<R> eval_roughplastic(const BCtx<R> &c, const V3<R> &dir_out, const DMaterial &m) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return V3<R>(0, 0, 0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R n_dot_h = dot(f.n, h), n_dot_in = dot(f.n, c.dir_in), n_dot_out = dot(f.n, dir_out);
    if (n_dot_out <= R(0) || n_dot_h <= R(0)) return V3<R>(0, 0, 0);
    V3<R> Kd = c.t3(m.tex[0]), Ks = c.t3(m.tex[1]);
    R roughness = clampr(c.t1(m.tex[2]), R(0.01), R(1));
    R eta = (R)m.eta;
    R F_o = fresnel_dielectric(dot(h, dir_out), eta);
    R D = GTR2_at(f, h, roughness);
    R G = smith_masking_gtr2(to_local(f, c.dir_in), roughness) * smith_masking_gtr2(to_local(f, dir_out), roughness);
    V3<R> spec = Ks * (G * F_o * D) / (R(4) * n_dot_in * n_dot_out);
    R F_i = fresnel_dielectric(dot(h, c.dir_in), eta);
    V3<R> diff = Kd * (R(1) - F_o) * (R(1) - F_i) / Consts<R>::PI;
    return (spec + diff) * n_dot_out;
}
template <typename R> GD_D R pdf_roughplastic(const BCtx<R> &c, const V3<R> &dir_out, const DMaterial &m) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return R(0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R n_dot_in = dot(f.n, c.dir_in), n_dot_out = dot(f.n, dir_out), n_dot_h = dot(f.n, h);
    if (n_dot_out <= R(0) || n_dot_h <= R(0)) return R(0);
    V3<R> S = c.t3(m.tex[1]), Rr = c.t3(m.tex[0]);
    R lS = luminance(S), lR = luminance(Rr);
    if (lS + lR <= R(0)) return R(0);
    R roughness = clampr(c.t1(m.tex[2]), R(0.01), R(1));
    R spec_prob = lS / (lS + lR);
    R diff_prob = R(1) - spec_prob;
    R G = smith_masking_gtr2(to_local(f, c.dir_in), roughness);
    R D = GTR2_at(f, h, roughness);
    spec_prob *= (G * D) / (R(4) * n_dot_in);
    diff_prob *= n_dot_out / Consts<R>::PI;
    return spec_prob + diff_prob;
}
template <typename R> GD_D bool sample_roughplastic(const BCtx<R> &c, const DMaterial &m, R u0, R u1, R w, BsdfSample<R> &s) {
    if (dot(c.v.gn, c.dir_in) < R(0)) return false;
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> Ks = c.t3(m.tex[1]), Kd = c.t3(m.tex[0]);
    R lS = luminance(Ks), lR = luminance(Kd);
    if (lS + lR <= R(0)) return false;
    R spec_prob = lS / (lS + lR);
    if (w < spec_prob) {
        V3<R> local_in = to_local(f, c.dir_in);
        R roughness = clampr(c.t1(m.tex[2]), R(0.01), R(1));
        R alpha = roughness * roughness;
        V3<R> mn = sample_visible_normals<R>(local_in, alpha, alpha, u0, u1);
        V3<R> h = to_world(f, mn);
        s.dir_out = normalize(-c.dir_in + R(2) * dot(c.dir_in, h) * h);
        s.eta = R(0); s.roughness = roughness;
    } else {
        s.dir_out = to_world(f, sample_cos_hemisphere<R>(u0, u1));
        s.eta = R(0); s.roughness = R(1);
    }
    return true;
}

// ------------------------------------------------------------------ RoughDielectric
template <typename R> GD_D V3<R> dielectric_half(const BCtx<R> &c, const V3<R> &dir_out, const Frame<R> &f, R eta, bool reflect) {
    V3<R> h = reflect ? normalize(c.dir_in + dir_out) : normalize(c.dir_in + dir_out * eta);
    if (dot(h, f.n) < R(0)) h = -h;
    return h;
}
template <typename R> GD_D V3<R> eval_roughdielectric(const BCtx<R> &c, const V3<R> &dir_out, const DMaterial &m) {
    bool reflect = dot(c.v.gn, c.dir_in) * dot(c.v.gn, dir_out) > R(0);
    Frame<R> f = flip_frame_two_sided(c.v, c.dir_in);
    R eta = dot(c.v.gn, c.dir_in) > R(0) ? (R)m.eta : R(1) / (R)m.eta;
    V3<R> Ks = c.t3(m.tex[0]), Kt = c.t3(m.tex[1]);
    R roughness = c.t1(m.tex[2]);
    V3<R> h = dielectric_half(c, dir_out, f, eta, reflect);
    roughness = clampr(roughness, R(0.01), R(1));
    R h_dot_in = dot(h, c.dir_in);
    R F = fresnel_dielectric(h_dot_in, eta);
    R D = GTR2_at(f, h, roughness);
    R G = smith_masking_gtr2(to_local(f, c.dir_in), roughness) * smith_masking_gtr2(to_local(f, dir_out), roughness);
    if (reflect) return Ks * (F * D * G) / (R(4) * fabs(dot(f.n, c.dir_in)));
    R eta_factor = R(1) / (eta * eta);   // TransportDirection::TO_LIGHT (default argument everywhere on this path)
    R h_dot_out = dot(h, dir_out);
    R sqrt_denom = h_dot_in + eta * h_dot_out;
    return Kt * (eta_factor * (R(1) - F) * D * G * eta * eta * fabs(h_dot_out * h_dot_in)) /
           (fabs(dot(f.n, c.dir_in)) * sqrt_denom * sqrt_denom);
}
template <typename R> GD_D R pdf_roughdielectric(const BCtx<R> &c, const V3<R> &dir_out, const DMaterial &m) {
    bool reflect = dot(c.v.gn, c.dir_in) * dot(c.v.gn, dir_out) > R(0);
    Frame<R> f = flip_frame_two_sided(c.v, c.dir_in);
    R eta = dot(c.v.gn, c.dir_in) > R(0) ? (R)m.eta : R(1) / (R)m.eta;
    V3<R> h = dielectric_half(c, dir_out, f, eta, reflect);
    R roughness = clampr(c.t1(m.tex[2]), R(0.01), R(1));
    R h_dot_in = dot(h, c.dir_in);
    R F = fresnel_dielectric(h_dot_in, eta);
    R D = GTR2_at(f, h, roughness);
    R G_in = smith_masking_gtr2(to_local(f, c.dir_in), roughness);
    if (reflect) return (F * D * G_in) / (R(4) * fabs(dot(f.n, c.dir_in)));
    R h_dot_out = dot(h, dir_out);
    R sqrt_denom = h_dot_in + eta * h_dot_out;
    R dh_dout = eta * eta * h_dot_out / (sqrt_denom * sqrt_denom);
    return (R(1) - F) * D * G_in * fabs(dh_dout * h_dot_in / dot(f.n, c.dir_in));
}
// shared by RoughDielectric and DisneyGlass (identical sampling code)
template <typename R> GD_D bool sample_dielectric(const BCtx<R> &c, R mat_eta, R roughness_raw, R u0, R u1, R w, BsdfSample<R> &s) {
    R eta = dot(c.v.gn, c.dir_in) > R(0) ? mat_eta : R(1) / mat_eta;
    Frame<R> f = flip_frame_two_sided(c.v, c.dir_in);
    R roughness = clampr(roughness_raw, R(0.01), R(1));
    R alpha = roughness * roughness;
    V3<R> local_in = to_local(f, c.dir_in);
    V3<R> mn = sample_visible_normals<R>(local_in, alpha, alpha, u0, u1);
    V3<R> h = to_world(f, mn);
    if (dot(h, f.n) < R(0)) h = -h;
    R h_dot_in = dot(h, c.dir_in);
    R F = fresnel_dielectric(h_dot_in, eta);
    if (w <= F) {
        s.dir_out = normalize(-c.dir_in + R(2) * dot(c.dir_in, h) * h);
        s.eta = R(0); s.roughness = roughness;
        return true;
    }
    R h_dot_out_sq = R(1) - (R(1) - h_dot_in * h_dot_in) / (eta * eta);
    if (h_dot_out_sq <= R(0)) return false;
    if (h_dot_in < R(0)) h = -h;
    R h_dot_out = sqrt(h_dot_out_sq);
    s.dir_out = -c.dir_in / eta + (fabs(h_dot_in) / eta - h_dot_out) * h;
    s.eta = eta; s.roughness = roughness;
    return true;
}

// ------------------------------------------------------------------ Disney lobes (parameters already evaluated)
template <typename R> GD_D V3<R> eval_disney_diffuse(const BCtx<R> &c, const V3<R> &dir_out, const V3<R> &base, R roughness, R subsurface) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return V3<R>(0, 0, 0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R n_in = dot(f.n, c.dir_in), n_out = dot(f.n, dir_out), h_out = dot(h, dir_out);
    roughness = clampr(roughness, R(0.01), R(1));
    R fd90 = R(0.5) + R(2) * roughness * sq(fabs(h_out));
    R p5o = pw(R(1) - fabs(n_out), R(5)), p5i = pw(R(1) - fabs(n_in), R(5));
    R fdo = R(1.0) + (fd90 - R(1.0)) * p5o, fdi = R(1.0) + (fd90 - R(1.0)) * p5i;
    V3<R> fbase = (base * fdi * fdo * fabs(n_out)) / Consts<R>::PI;
    R fss90 = roughness * sq(fabs(h_out));
    R fssi = R(1.0) + (fss90 - R(1.0)) * p5i, fsso = R(1.0) + (fss90 - R(1.0)) * p5o;
    R k = (fssi * fsso) * ((R(1) / (fabs(n_in) + fabs(n_out))) - R(0.5)) + R(0.5);   // same value on all three channels
    V3<R> fsub = ((R(1.25) * base) / Consts<R>::PI) * V3<R>(k, k, k) * fabs(n_out);
    return (R(1) - subsurface) * fbase + subsurface * fsub;
}
template <typename R> GD_D void metal_alphas(R roughness, R anisotropic, bool use_sqrt, R &ax, R &ay) {
    R aspect = use_sqrt ? sqrt(R(1) - R(0.9) * anisotropic) : pw(R(1) - R(0.9) * anisotropic, R(0.5));
    ax = rmax(R(0.0001), sq(roughness) / aspect);
    ay = rmax(R(0.0001), sq(roughness) * aspect);
}
template <typename R> GD_D R disney_lambda(const V3<R> &l, R ax, R ay, bool use_sqrt) {
    R internal = (sq(l.x * ax) + sq(l.y * ay)) / sq(l.z);
    return ((use_sqrt ? sqrt(R(1) + internal) : pw(R(1) + internal, R(0.5))) - R(1)) / 2;
}
template <typename R> GD_D V3<R> eval_disney_metal(const BCtx<R> &c, const V3<R> &dir_out, const V3<R> &base, R roughness, R anisotropic) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return V3<R>(0, 0, 0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R n_in = dot(f.n, c.dir_in), h_out = dot(h, dir_out);
    R const_denom = R(4) * fabs(n_in);
    roughness = clampr(roughness, R(0.01), R(1));
    V3<R> fm = base + (V3<R>(1, 1, 1) - base) * pw(R(1.0) - fabs(h_out), R(5));
    V3<R> pv = to_local(f, h);
    R ax, ay; metal_alphas(roughness, anisotropic, true, ax, ay);
    R D = R(1) / (Consts<R>::PI * ax * ay * sq(sq(pv.x / ax) + sq(pv.y / ay) + sq(pv.z)));
    R din = disney_lambda(to_local(f, c.dir_in), ax, ay, true), dout = disney_lambda(to_local(f, dir_out), ax, ay, true);
    R G = (R(1) / (R(1) + din)) * (R(1) / (R(1) + dout));
    return (fm * D * G) / const_denom;
}
template <typename R> GD_D R pdf_disney_metal(const BCtx<R> &c, const V3<R> &dir_out, R roughness, R anisotropic) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return R(0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R const_denom = R(4) * fabs(dot(f.n, c.dir_in));
    V3<R> pv = to_local(f, h);
    R ax, ay; metal_alphas(roughness, anisotropic, false, ax, ay);   // no roughness clamp here (disney_metal.inl:96-111)
    R D = R(1) / (Consts<R>::PI * ax * ay * sq(sq(pv.x / ax) + sq(pv.y / ay) + sq(pv.z)));
    R din = disney_lambda(to_local(f, c.dir_in), ax, ay, true);
    R G = R(1) / (R(1) + din);
    return (G * D) / const_denom;
}
template <typename R> GD_D bool sample_disney_metal(const BCtx<R> &c, R roughness, R anisotropic, R u0, R u1, BsdfSample<R> &s) {
    if (dot(c.v.gn, c.dir_in) < R(0)) return false;
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> local_in = to_local(f, c.dir_in);
    roughness = clampr(roughness, R(0.01), R(1));
    R ax, ay; metal_alphas(roughness, anisotropic, true, ax, ay);
    V3<R> mn = sample_visible_normals<R>(local_in, ax, ay, u0, u1);
    V3<R> h = to_world(f, mn);
    s.dir_out = normalize(-c.dir_in + R(2) * dot(c.dir_in, h) * h);
    s.eta = R(0); s.roughness = roughness;
    return true;
}
template <typename R> GD_D R clearcoat_D(R alpha_g, R hz) {
    return (sq(alpha_g) - R(1)) / (Consts<R>::PI * log(sq(alpha_g)) * (R(1) + (sq(alpha_g) - R(1)) * sq(hz)));
}
template <typename R> GD_D V3<R> eval_disney_clearcoat(const BCtx<R> &c, const V3<R> &dir_out, R gloss) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return V3<R>(0, 0, 0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R n_in = dot(f.n, c.dir_in), h_out = dot(h, dir_out);
    R r0 = sq(R(1.5 - 1.0)) / sq(R(1.5 + 1.0));
    R fc = r0 + (R(1) - r0) * pw(R(1) - fabs(h_out), R(5));
    R alpha_g = (R(1) - gloss) * R(0.1) + gloss * R(0.001);
    R dc = clearcoat_D(alpha_g, to_local(f, h).z);
    V3<R> li = to_local(f, c.dir_in), lo = to_local(f, dir_out);
    R dout = (pw(R(1) + (sq(lo.x * R(0.25)) + sq(lo.y * R(0.25))) / sq(lo.z), R(0.5)) - R(1.0)) / 2;
    R din = (pw(R(1) + (sq(li.x * R(0.25)) + sq(li.y * R(0.25))) / sq(li.z), R(0.5)) - R(1.0)) / 2;
    R gc = (R(1) / (R(1) + din)) * (R(1) / (R(1) + dout));
    R val = (fc * dc * gc) / (R(4) * fabs(n_in));
    return V3<R>(val, val, val);
}
template <typename R> GD_D R pdf_disney_clearcoat(const BCtx<R> &c, const V3<R> &dir_out, R gloss) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return R(0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R n_out = dot(f.n, dir_out), n_h = dot(f.n, h);
    R alpha_g = (R(1) - gloss) * R(0.1) + gloss * R(0.001);
    R dc = clearcoat_D(alpha_g, to_local(f, h).z);
    return (dc * fabs(n_h)) / (R(4) * fabs(n_out));
}
template <typename R> GD_D bool sample_disney_clearcoat(const BCtx<R> &c, R gloss, R u0, R u1, BsdfSample<R> &s) {
    if (dot(c.v.gn, c.dir_in) < R(0)) return false;
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    R alpha_g = (R(1) - gloss) * R(0.1) + gloss * R(0.001);
    V3<R> h = to_world(f, sample_clearcoat<R>(alpha_g, u0, u1));
    s.dir_out = normalize(-c.dir_in + R(2) * dot(c.dir_in, h) * h);
    s.eta = R(0); s.roughness = alpha_g;
    return true;
}
template <typename R> GD_D V3<R> eval_disney_sheen(const BCtx<R> &c, const V3<R> &dir_out, const V3<R> &base, R sheen_tint) {
    if (dot(c.v.gn, c.dir_in) < R(0) || dot(c.v.gn, dir_out) < R(0)) return V3<R>(0, 0, 0);
    Frame<R> f = flip_frame_one_sided(c.v, c.dir_in);
    V3<R> h = normalize(c.dir_in + dir_out);
    R h_out = dot(h, dir_out), n_out = dot(f.n, dir_out);
    V3<R> c_tint(1, 1, 1);
    if (luminance(base) > R(0)) c_tint = base / luminance(base);
    V3<R> c_sheen = V3<R>(R(1.0) - sheen_tint, R(1.0) - sheen_tint, R(1.0) - sheen_tint) + sheen_tint * c_tint;
    return c_sheen * pw(R(1) - fabs(h_out), R(5)) * fabs(n_out);
}
template <typename R> GD_D void glass_terms(const BCtx<R> &c, const V3<R> &dir_out, R mat_eta, R roughness_raw, R anisotropic,
                                            bool &reflect, Frame<R> &f, R &eta, V3<R> &h, R &F, R &dm, R &din, R &dout) {
    reflect = dot(c.v.gn, c.dir_in) * dot(c.v.gn, dir_out) > R(0);
    f = flip_frame_two_sided(c.v, c.dir_in);
    eta = dot(c.v.gn, c.dir_in) > R(0) ? mat_eta : R(1) / mat_eta;
    h = dielectric_half(c, dir_out, f, eta, reflect);
    R roughness = clampr(roughness_raw, R(0.01), R(1));
    F = fresnel_dielectric(dot(h, c.dir_in), eta);
    V3<R> pv = to_local(f, h);
    R ax, ay; metal_alphas(roughness, anisotropic, false, ax, ay);
    dm = R(1) / (Consts<R>::PI * ax * ay * sq(sq(pv.x) / sq(ax) + sq(pv.y) / sq(ay) + sq(pv.z)));
    dout = disney_lambda(to_local(f, dir_out), ax, ay, false);
    din = disney_lambda(to_local(f, c.dir_in), ax, ay, false);
}
template <typename R> GD_D V3<R> eval_disney_glass(const BCtx<R> &c, const V3<R> &dir_out, const V3<R> &base, R roughness, R anisotropic, R mat_eta) {
    bool reflect; Frame<R> f; R eta, F, dm, din, dout; V3<R> h;
    glass_terms(c, dir_out, mat_eta, roughness, anisotropic, reflect, f, eta, h, F, dm, din, dout);
    R gm = (R(1) / (R(1) + din)) * (R(1) / (R(1) + dout));
    if (reflect) return base * (F * dm * gm) / (R(4) * fabs(dot(f.n, c.dir_in)));
    R h_dot_in = dot(h, c.dir_in), h_dot_out = dot(h, dir_out);
    R sqrt_denom = h_dot_in + eta * h_dot_out;
    V3<R> csq(pw(base.x, R(0.5)), pw(base.y, R(0.5)), pw(base.z, R(0.5)));
    return csq * ((R(1) - F) * dm * gm * fabs(h_dot_out * h_dot_in)) / (fabs(dot(f.n, c.dir_in)) * sqrt_denom * sqrt_denom);
}
template <typename R> GD_D R pdf_disney_glass(const BCtx<R> &c, const V3<R> &dir_out, R roughness, R anisotropic, R mat_eta) {
    bool reflect; Frame<R> f; R eta, F, dm, din, dout; V3<R> h;
    glass_terms(c, dir_out, mat_eta, roughness, anisotropic, reflect, f, eta, h, F, dm, din, dout);
    R gm = R(1) / (R(1) + din);
    if (reflect) return (F * dm * gm) / (R(4) * fabs(dot(f.n, c.dir_in)));
    R h_dot_out = dot(h, dir_out), h_dot_in = dot(h, c.dir_in);
    R prod = fabs(h_dot_out * h_dot_in);
    R sqrt_denom = h_dot_in + eta * h_dot_out;
    return ((R(1) - F) * dm * gm * fabs(prod)) / (fabs(dot(f.n, c.dir_in)) * sqrt_denom * sqrt_denom);
}

// ------------------------------------------------------------------ DisneyBSDF (uber)
template <typename R> struct DisneyParams {
    V3<R> base, c0; R spec_trans, metallic, subsurface, specular, roughness, specular_tint, anisotropic, sheen, sheen_tint, clearcoat, clearcoat_gloss, eta;
};
template <typename R> GD_D DisneyParams<R> disney_params(const BCtx<R> &c, const DMaterial &m) {
    DisneyParams<R> p;
    p.base = c.t3(m.tex[0]); p.spec_trans = c.t1(m.tex[1]); p.metallic = c.t1(m.tex[2]); p.subsurface = c.t1(m.tex[3]);
    p.specular = c.t1(m.tex[4]); p.roughness = c.t1(m.tex[5]); p.specular_tint = c.t1(m.tex[6]); p.anisotropic = c.t1(m.tex[7]);
    p.sheen = c.t1(m.tex[8]); p.sheen_tint = c.t1(m.tex[9]); p.clearcoat = c.t1(m.tex[10]); p.clearcoat_gloss = c.t1(m.tex[11]);
    V3<R> c_tint(1, 1, 1);
    if (luminance(p.base) > R(0)) c_tint = p.base / luminance(p.base);
    p.eta = dot(c.v.gn, c.dir_in) > R(0) ? (R)m.eta : R(1) / (R)m.eta;
    V3<R> Ks = V3<R>(R(1) - p.specular_tint, R(1) - p.specular_tint, R(1) - p.specular_tint) + p.specular_tint * c_tint;
    R r0 = sq(p.eta - R(1)) / sq(p.eta + R(1));
    p.c0 = p.specular * r0 * (R(1) - p.metallic) * Ks + p.metallic * p.base;
    return p;
}
template <typename R> GD_D V3<R> eval_disney_bsdf(const BCtx<R> &c, const V3<R> &dir_out, const DMaterial &m) {
    DisneyParams<R> p = disney_params(c, m);
    V3<R> glass = eval_disney_glass(c, dir_out, p.base, p.roughness, p.anisotropic, p.eta);
    R wd = (R(1) - p.spec_trans) * (R(1) - p.metallic);
    R wm = (R(1) - p.spec_trans * (R(1) - p.metallic));
    R wc = R(0.25) * p.clearcoat;
    R wg = (R(1) - p.metallic) * p.spec_trans;
    R ws = (R(1) - p.metallic) * p.sheen;
    if (dot(c.v.gn, c.dir_in) <= R(0)) return wg * glass;
    V3<R> d = eval_disney_diffuse(c, dir_out, p.base, p.roughness, p.subsurface);
    V3<R> me = eval_disney_metal(c, dir_out, p.c0, p.roughness, p.anisotropic);
    V3<R> sh = eval_disney_sheen(c, dir_out, p.base, p.sheen_tint);
    V3<R> cc = eval_disney_clearcoat(c, dir_out, p.clearcoat_gloss);
    return wd * d + wm * me + wc * cc + wg * glass + ws * sh;
}
template <typename R> GD_D R pdf_disney_bsdf(const BCtx<R> &c, const V3<R> &dir_out, const DMaterial &m) {
    DisneyParams<R> p = disney_params(c, m);
    R wd = (R(1) - p.spec_trans) * (R(1) - p.metallic);
    R wm = (R(1) - p.spec_trans * (R(1) - p.metallic));
    R wc = R(0.25) * p.clearcoat;
    R wg = (R(1) - p.metallic) * p.spec_trans;
    R net = wd + wm + wc + wg;
    if (dot(c.v.gn, c.dir_in) <= R(0)) return pdf_disney_glass(c, dir_out, p.roughness, p.anisotropic, p.eta);
    return (wd / net) * pdf_cosine(c, dir_out) + (wm / net) * pdf_disney_metal(c, dir_out, p.roughness, p.anisotropic) +
           (wc / net) * pdf_disney_clearcoat(c, dir_out, p.clearcoat_gloss) + (wg / net) * pdf_disney_glass(c, dir_out, p.roughness, p.anisotropic, p.eta);
}
template <typename R> GD_D bool sample_disney_bsdf(const BCtx<R> &c, const DMaterial &m, R u0, R u1, R w, BsdfSample<R> &s) {
    DisneyParams<R> p = disney_params(c, m);
    if (u0 < R(0.25)) return sample_cosine(c, u0, u1, clampr(p.roughness, R(0.01), R(1)), s);
    if (u0 >= R(0.25) && u0 < R(0.5)) return sample_disney_metal(c, p.roughness, p.anisotropic, u0, u1, s);
    if (u0 >= R(0.5) && u0 < R(0.75)) return sample_disney_clearcoat(c, p.clearcoat_gloss, u0, u1, s);
    return sample_dielectric(c, p.eta, p.roughness, u0, u1, w, s);
}

// ------------------------------------------------------------------ dispatch
template <typename R> GD_D V3<R> bsdf_eval_impl(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, const V3<R> &dir_out) {
    const DMaterial &m = sc.materials[v.material_id];
    BCtx<R> c{sc, v, dir_in};
    switch (m.type) {
    case MAT_LAMBERTIAN: return eval_lambertian(c, dir_out, m.tex[0]);
    case MAT_ROUGHPLASTIC: return eval_roughplastic(c, dir_out, m);
    case MAT_ROUGHDIELECTRIC: return eval_roughdielectric(c, dir_out, m);
    case MAT_DISNEY_DIFFUSE: return eval_disney_diffuse(c, dir_out, c.t3(m.tex[0]), c.t1(m.tex[1]), c.t1(m.tex[2]));
    case MAT_DISNEY_METAL: return eval_disney_metal(c, dir_out, c.t3(m.tex[0]), c.t1(m.tex[1]), c.t1(m.tex[2]));
    case MAT_DISNEY_GLASS: return eval_disney_glass(c, dir_out, c.t3(m.tex[0]), c.t1(m.tex[1]), c.t1(m.tex[2]), (R)m.eta);
    case MAT_DISNEY_CLEARCOAT: return eval_disney_clearcoat(c, dir_out, c.t1(m.tex[0]));
    case MAT_DISNEY_SHEEN: return eval_disney_sheen(c, dir_out, c.t3(m.tex[0]), c.t1(m.tex[1]));
    default: return eval_disney_bsdf(c, dir_out, m);
    }
}
template <typename R> GD_D R bsdf_pdf_impl(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, const V3<R> &dir_out) {
    const DMaterial &m = sc.materials[v.material_id];
    BCtx<R> c{sc, v, dir_in};
    switch (m.type) {
    case MAT_LAMBERTIAN: case MAT_DISNEY_DIFFUSE: case MAT_DISNEY_SHEEN: return pdf_cosine(c, dir_out);
    case MAT_ROUGHPLASTIC: return pdf_roughplastic(c, dir_out, m);
    case MAT_ROUGHDIELECTRIC: return pdf_roughdielectric(c, dir_out, m);
    case MAT_DISNEY_METAL: return pdf_disney_metal(c, dir_out, c.t1(m.tex[1]), c.t1(m.tex[2]));
    case MAT_DISNEY_GLASS: return pdf_disney_glass(c, dir_out, c.t1(m.tex[1]), c.t1(m.tex[2]), (R)m.eta);
    case MAT_DISNEY_CLEARCOAT: return pdf_disney_clearcoat(c, dir_out, c.t1(m.tex[0]));
    default: return pdf_disney_bsdf(c, dir_out, m);
    }
}
template <typename R> GD_D bool bsdf_sample_impl(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, R u0, R u1, R w, BsdfSample<R> &s) {
    const DMaterial &m = sc.materials[v.material_id];
    BCtx<R> c{sc, v, dir_in};
    switch (m.type) {
    case MAT_LAMBERTIAN: case MAT_DISNEY_SHEEN: return sample_cosine(c, u0, u1, R(1), s);
    case MAT_DISNEY_DIFFUSE: return sample_cosine(c, u0, u1, clampr(c.t1(m.tex[1]), R(0.01), R(1)), s);
    case MAT_ROUGHPLASTIC: return sample_roughplastic(c, m, u0, u1, w, s);
    case MAT_ROUGHDIELECTRIC: return sample_dielectric(c, (R)m.eta, c.t1(m.tex[2]), u0, u1, w, s);
    case MAT_DISNEY_METAL: return sample_disney_metal(c, c.t1(m.tex[1]), c.t1(m.tex[2]), u0, u1, s);
    case MAT_DISNEY_GLASS: return sample_dielectric(c, (R)m.eta, c.t1(m.tex[1]), u0, u1, w, s);
    case MAT_DISNEY_CLEARCOAT: return sample_disney_clearcoat(c, c.t1(m.tex[0]), u0, u1, s);
    default: return sample_disney_bsdf(c, m, u0, u1, w, s);
    }
}

// Out-of-line generic dispatch (any of the nine materials) ...
template <typename R> GD_BIG V3<R> bsdf_eval(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, const V3<R> &dir_out) { return bsdf_eval_impl<R>(sc, v, dir_in, dir_out); }
template <typename R> GD_BIG R bsdf_pdf(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, const V3<R> &dir_out) { return bsdf_pdf_impl<R>(sc, v, dir_in, dir_out); }
template <typename R> GD_BIG bool bsdf_sample(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, R u0, R u1, R w, BsdfSample<R> &s) { return bsdf_sample_impl<R>(sc, v, dir_in, u0, u1, w, s); }

// ... and the material-set selector used by the wavefront kernels: MS_LAMBERTIAN scenes
// (cbox, Sponza: every material is a Lambertian) get the diffuse code fully inlined, with
// the register budget of that one lobe instead of the Disney uber-shader's.
enum { MS_GENERIC = 0, MS_LAMBERTIAN = 1 };
template <typename R, int MS> GD_D V3<R> bsdf_eval_ms(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, const V3<R> &dir_out) {
    if (MS == MS_LAMBERTIAN) { BCtx<R> c{sc, v, dir_in}; return eval_lambertian(c, dir_out, sc.materials[v.material_id].tex[0]); }
    return bsdf_eval<R>(sc, v, dir_in, dir_out);
}
template <typename R, int MS> GD_D R bsdf_pdf_ms(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, const V3<R> &dir_out) {
    if (MS == MS_LAMBERTIAN) { BCtx<R> c{sc, v, dir_in}; return pdf_cosine(c, dir_out); }
    return bsdf_pdf<R>(sc, v, dir_in, dir_out);
}
template <typename R, int MS> GD_D bool bsdf_sample_ms(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &dir_in, R u0, R u1, R w, BsdfSample<R> &s) {
    if (MS == MS_LAMBERTIAN) { BCtx<R> c{sc, v, dir_in}; return sample_cosine(c, u0, u1, R(1), s); }
    return bsdf_sample<R>(sc, v, dir_in, u0, u1, w, s);
}

} // namespace gdpt
