// Light sampling on the device (the `path` integrator's next-event estimation):
//   TableDist1D sample / pmf          /root/reference/src/table_dist.cpp:3-43
//   sample_light / light_pmf          /root/reference/src/scene.cpp:61-67
//   sample_point_on_light / pdf       /root/reference/src/light.cpp:57-71, lights/diffuse_area_light.inl:5-13
//   sample_point_on_shape / pdf       /root/reference/src/shapes/sphere.inl:164-231 (cone sampling seen from outside,
//                                     uniform from inside), shapes/triangle_mesh.inl:24-58 (area-proportional triangle,
//                                     sqrt parameterisation, normal flipped towards the shading normal)
//   emission of a sampled point       /root/reference/src/lights/diffuse_area_light.inl:15-20
// The tables themselves (triangle-area CDFs of emissive meshes, the light-power CDF of scene.cpp:46-52) are built
// on the host by gdpt_create exactly like make_table_dist_1d, including its quirk that the LAST cdf entry keeps the
// un-normalised total.
//   environment map                   /root/reference/src/lights/envmap.inl:1-98, TableDist2D table_dist.cpp:45-151
#pragma once
#include "gdpt_scene.cuh"

namespace gdpt {

// std::upper_bound over cdf[0..size] then clamp(idx - 1, 0, size - 1)      table_dist.cpp:28-34
template <typename R> GD_D int table_sample(const R *cdf, int size, R u) {
    int lo = 0, n = size + 1;          // first index whose value is > u
    while (n > 0) {
        int half = n >> 1;
        if (!(u < cdf[lo + half])) { lo += half + 1; n -= half + 1; } else n = half;
    }
    int off = lo - 1;
    return off < 0 ? 0 : (off > size - 1 ? size - 1 : off);
}

template <typename R> struct LightPoint { V3<R> position, normal; };

template <typename R> GD_D LightPoint<R> sample_point_on_sphere(const double *sph, const V3<R> &ref, R u0, R u1) {
    const V3<R> center((R)sph[0], (R)sph[1], (R)sph[2]);
    const R r = (R)sph[3];
    LightPoint<R> out;
    if (distance_squared(ref, center) < r * r) {           // inside: uniform over the whole sphere
        R z = R(1) - R(2) * u0;
        R r_ = sqrt(fmax(R(0), R(1) - z * z));
        R phi = R(2) * Consts<R>::PI * u1;
        V3<R> offset(r_ * cos(phi), r_ * sin(phi), z);
        out.position = center + r * offset; out.normal = offset;
        return out;
    }
    V3<R> dir_to_center = normalize(center - ref);
    Frame<R> frame(dir_to_center);
    R sin_max_sq = r * r / distance_squared(ref, center);
    R cos_max = sqrt(rmax(R(0), R(1) - sin_max_sq));
    R azimuth = u1 * R(2) * Consts<R>::PI;
    R dc = distance(ref, center);
    R cos_alpha, sin_el_sq, cos_el;
    if (sizeof(R) == 8) {   // the reference's expressions, literally
        cos_el = (R(1) - u0) + u0 * cos_max;
        sin_el_sq = rmax(R(0), R(1) - cos_el * cos_el);
    } else {                // float: a far, small light makes 1 - cos cancel; same quantities from sin^2
        R omc = u0 * (sin_max_sq / (R(1) + cos_max));           // 1 - cos_el
        cos_el = R(1) - omc;
        sin_el_sq = omc * (R(2) - omc);
    }
    R root = sqrt(rmax(R(0), r * r - dc * dc * sin_el_sq));
    R ds = dc * cos_el - root;
    if (sizeof(R) == 8) cos_alpha = (dc * dc + r * r - ds * ds) / (R(2) * dc * r);
    else cos_alpha = ((dc - ds) * (dc + ds) + r * r) / (R(2) * dc * r);
    R sin_alpha = sqrt(rmax(R(0), R(1) - cos_alpha * cos_alpha));
    V3<R> n = -to_world(frame, V3<R>(sin_alpha * cos(azimuth), sin_alpha * sin(azimuth), cos_alpha));
    out.position = r * n + center; out.normal = n;
    return out;
}

template <typename R> GD_D R pdf_point_on_sphere(const double *sph, const LightPoint<R> &p, const V3<R> &ref) {
    const V3<R> center((R)sph[0], (R)sph[1], (R)sph[2]);
    const R r = (R)sph[3];
    if (distance_squared(ref, center) < r * r) return R(1) / (R(4) * Consts<R>::PI * r * r);
    R sin_max_sq = r * r / distance_squared(ref, center);
    R cos_max = sqrt(rmax(R(0), R(1) - sin_max_sq));
    R one_minus_cos = sizeof(R) == 8 ? R(1) - cos_max : sin_max_sq / (R(1) + cos_max);
    R pdf_solid_angle = R(1) / (R(2) * Consts<R>::PI * one_minus_cos);
    V3<R> dir = normalize(p.position - ref);
    return pdf_solid_angle * fabs(dot(p.normal, dir)) / distance_squared(ref, p.position);
}

template <typename R> GD_D LightPoint<R> sample_point_on_mesh(const DScene<R> &sc, const DShape &sh, R u0, R u1, R w) {
    int tri = table_sample<R>(sc.tri_cdf + sh.cdf_off, sh.n_tris, w);
    const int *ix = sc.indices + 3 * (size_t)(sh.tri_off + tri);
    const R *P = sc.positions + 3 * (size_t)sh.vtx_off;
    V3<R> v0(P[3 * ix[0]], P[3 * ix[0] + 1], P[3 * ix[0] + 2]), v1(P[3 * ix[1]], P[3 * ix[1] + 1], P[3 * ix[1] + 2]), v2(P[3 * ix[2]], P[3 * ix[2] + 1], P[3 * ix[2] + 2]);
    V3<R> e1 = v1 - v0, e2 = v2 - v0;
    R a = sqrt(clampr(u0, R(0), R(1)));
    R b1 = R(1) - a, b2 = a * u1;
    V3<R> gn = normalize(cross(e1, e2));
    if (sh.flags & 1) {
        const R *Nn = sc.normals + 3 * (size_t)sh.nrm_off;
        V3<R> n0(Nn[3 * ix[0]], Nn[3 * ix[0] + 1], Nn[3 * ix[0] + 2]), n1(Nn[3 * ix[1]], Nn[3 * ix[1] + 1], Nn[3 * ix[1] + 2]), n2(Nn[3 * ix[2]], Nn[3 * ix[2] + 1], Nn[3 * ix[2] + 2]);
        V3<R> sn = normalize((R(1) - b1 - b2) * n0 + b1 * n1 + b2 * n2);
        if (dot(gn, sn) < R(0)) gn = -gn;
    }
    LightPoint<R> out; out.position = v0 + (e1 * b1) + (e2 * b2); out.normal = gn;
    return out;
}

// ---------------------------------------------------------------- environment map
// sample(TableDist2D, rnd)                                                  table_dist.cpp:125-148
template <typename R> GD_D V2<R> env_table_sample(const DScene<R> &sc, R r0, R r1) {
    const int w = sc.env_w, h = sc.env_h;
    int y = table_sample<R>(sc.env_cdf_marg, h, r1);
    R dy = r1 - sc.env_cdf_marg[y];
    R wy = sc.env_cdf_marg[y + 1] - sc.env_cdf_marg[y];
    if (wy > R(0)) dy /= wy;
    const R *cdf = sc.env_cdf_rows + (size_t)y * (w + 1);
    int x = table_sample<R>(cdf, w, r0);
    R dx = r0 - cdf[x];
    R wx = cdf[x + 1] - cdf[x];
    if (wx > R(0)) dx /= wx;
    return V2<R>((R(x) + dx) / R(w), (R(y) + dy) / R(h));
}
// pdf(TableDist2D, xy)                                                      table_dist.cpp:150-161
template <typename R> GD_D R env_table_pdf(const DScene<R> &sc, const V2<R> &xy) {
    const int w = sc.env_w, h = sc.env_h;
    int x = (int)clampr(xy.x * R(w), R(0), R(w - 1)), y = (int)clampr(xy.y * R(h), R(0), R(h - 1));
    return sc.env_pdf_marg[y] * sc.env_pdf_rows[(size_t)y * w + x] * R(w) * R(h);
}
// direction -> (u, v) of the lat-long map (envmap.inl:27-37, 57-65); local_dir in the map's frame
template <typename R> GD_D V2<R> env_dir_to_uv(const V3<R> &local_dir) {
    const R inv_two_pi = R(1) / (R(2) * Consts<R>::PI), inv_pi = R(1) / Consts<R>::PI;
    V2<R> uv(atan2(local_dir.x, -local_dir.z) * inv_two_pi, acos(clampr(local_dir.y, R(-1), R(1))) * inv_pi);
    if (uv.x < R(0)) uv.x += R(1);
    return uv;
}
template <typename R> GD_D LightPoint<R> env_sample_point(const DScene<R> &sc, R u0, R u1) {
    V2<R> uv = env_table_sample<R>(sc, u0, u1);
    R azimuth = uv.x * (R(2) * Consts<R>::PI), elevation = uv.y * Consts<R>::PI;
    V3<R> local_dir(sin(azimuth) * sin(elevation), cos(elevation), -cos(azimuth) * sin(elevation));
    V3<R> world_dir = xform_vector<R>(sc.env_to_world, local_dir);
    LightPoint<R> out; out.position = V3<R>(R(0), R(0), R(0)); out.normal = -world_dir;
    return out;
}
template <typename R> GD_D R env_pdf_point(const DScene<R> &sc, const LightPoint<R> &p) {
    V3<R> local_dir = xform_vector<R>(sc.env_to_local, -p.normal);
    V2<R> uv = env_dir_to_uv<R>(local_dir);
    R cos_el = local_dir.y;
    R sin_el = sqrt(clampr(R(1) - cos_el * cos_el, R(0), R(1)));
    if (sin_el <= R(0)) return R(0);
    return env_table_pdf<R>(sc, uv) / (R(2) * Consts<R>::PI * Consts<R>::PI * sin_el);
}
// emission(Envmap, view_dir, view_footprint, ...) (envmap.inl:48-74).  As shipped the footprint is
// min(|du/dw|, dv/dw_y) with dv/dw_y NEGATIVE, so the texture is always read at mip level 0; kept literally.
template <typename R> GD_D V3<R> env_emission(const DScene<R> &sc, const V3<R> &view_dir, R view_footprint) {
    (void)view_footprint;
    V3<R> w = xform_vector<R>(sc.env_to_local, -view_dir);
    V2<R> uv = env_dir_to_uv<R>(w);
    R dudwx = -w.z / (w.x * w.x + w.z * w.z), dudwz = w.x / (w.x * w.x + w.z * w.z);
    R dvdwy = R(-1) / sqrt(rmax(R(1) - w.y * w.y, R(0)));
    R footprint = rmin(sqrt(dudwx * dudwx + dudwz * dudwz), dvdwy);
    return eval_tex3<R>(sc, sc.env_tex, uv, footprint) * sc.env_scale;
}

// ---------------------------------------------------------------- lights
template <typename R> GD_D int sample_light(const DScene<R> &sc, R u) { return table_sample<R>(sc.light_cdf, sc.n_lights, u); }
template <typename R> GD_D R light_pmf(const DScene<R> &sc, int light_id) { return sc.light_pmf[light_id]; }
template <typename R> GD_D bool is_envmap(const DScene<R> &sc, int light_id) { return sc.light_shape[light_id] < 0; }

template <typename R> GD_D LightPoint<R> sample_point_on_light(const DScene<R> &sc, int light_id, const V3<R> &ref, R u0, R u1, R w) {
    const int shape_id = sc.light_shape[light_id];
    if (shape_id < 0) return env_sample_point<R>(sc, u0, u1);
    const DShape sh = sc.shapes[shape_id];
    if (sh.kind == SHAPE_SPHERE) return sample_point_on_sphere<R>(sc.spheres + 4 * (size_t)shape_id, ref, u0, u1);
    return sample_point_on_mesh<R>(sc, sh, u0, u1, w);
}

template <typename R> GD_D R pdf_point_on_light(const DScene<R> &sc, int light_id, const LightPoint<R> &p, const V3<R> &ref) {
    const int shape_id = sc.light_shape[light_id];
    if (shape_id < 0) return env_pdf_point<R>(sc, p);
    const DShape sh = sc.shapes[shape_id];
    if (sh.kind == SHAPE_SPHERE) return pdf_point_on_sphere<R>(sc.spheres + 4 * (size_t)shape_id, p, ref);
    return R(1) / sc.shape_area[shape_id];
}

// emission(light, view_dir, footprint, point_on_light, scene): DiffuseAreaLight (diffuse_area_light.inl:15-20) or Envmap
template <typename R> GD_D V3<R> light_emission(const DScene<R> &sc, int light_id, const V3<R> &view_dir, const LightPoint<R> &p) {
    if (is_envmap(sc, light_id)) return env_emission<R>(sc, view_dir, R(0));
    if (dot(p.normal, view_dir) <= R(0)) return V3<R>(R(0), R(0), R(0));
    const R *li = sc.light_intensity + 3 * (size_t)light_id;
    return V3<R>(li[0], li[1], li[2]);
}

} // namespace gdpt
