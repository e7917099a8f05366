// Device-side scene model (POD mirror of lajolla's Scene) and the
// device functions that turn a ray hit into a shading vertex.
//   sample_primary + pixel filters   /root/reference/src/camera.cpp:23-47, src/filters/*.inl
//   PathVertex construction          /root/reference/src/intersection.cpp:38-64
//   compute_shading_info (triangle)  /root/reference/src/shapes/triangle_mesh.inl:77-169
//   compute_shading_info (sphere)    /root/reference/src/shapes/sphere.inl:243-268
//   sphere quadratic                 /root/reference/src/shapes/sphere.inl:15-38,56-106
//   textures / mipmaps               /root/reference/src/texture.h:80-159, src/mipmap.h:46-88
//   ray differentials                /root/reference/src/ray.h:35-66
//   area-light emission              /root/reference/src/lights/diffuse_area_light.inl:15-20
#pragma once
#include "gdpt_math.cuh"

namespace gdpt {

enum { TEX_CONST = 0, TEX_IMAGE = 1, TEX_CHECKER = 2 };
enum { FILTER_BOX = 0, FILTER_TENT = 1, FILTER_GAUSSIAN = 2 };
enum { SHAPE_SPHERE = 0, SHAPE_MESH = 1 };
enum { MAT_LAMBERTIAN = 0, MAT_ROUGHPLASTIC, MAT_ROUGHDIELECTRIC, MAT_DISNEY_DIFFUSE, MAT_DISNEY_METAL,
       MAT_DISNEY_GLASS, MAT_DISNEY_CLEARCOAT, MAT_DISNEY_SHEEN, MAT_DISNEY_BSDF, MAT_COUNT };

template <typename R> struct DTexture {
    int kind, image_id;
    R v0[3], v1[3];
    R uscale, vscale, uoffset, voffset;
};

struct DImage {          // one mip chain inside a texel pool
    int n_levels, channels;
    int w[8], h[8];
    long long off[8];    // texel offset of each level
};

struct DMaterial {
    int type;
    int tex[12];
    double eta;
};

struct DShape {
    int kind, material_id, area_light_id, flags;   // flags: 1 = has normals, 2 = has uvs
    int vtx_off, tri_off, nrm_off, uv_off;
    int n_tris, cdf_off;                           // triangles of the mesh; offset of its area CDF in tri_cdf (-1: not emissive)
};

// texel storage: exact mode keeps doubles (3 per RGB texel, 1 per scalar texel),
// fast mode keeps float4 (RGB padded to 16 B for one vector load) / float.
template <typename R> struct TexelTypes;
template <> struct TexelTypes<double> { typedef double T3; typedef double T1; };
template <> struct TexelTypes<float> { typedef float4 T3; typedef float T1; };

template <typename R> struct DScene {
    R s2c[16], c2w[16];
    int width, height, filter_kind;
    R filter_param;
    int n_shapes, n_materials, n_lights;
    const DShape *shapes;
    const double *spheres;        // [n_shapes][4] centre + radius, always double (sphere.inl works in double)
    const R *positions;           // [V][3]
    const int *indices;           // [T][3] mesh-local vertex indices
    const R *normals;             // [Vn][3]
    const R *uvs;                 // [Vu][2]
    const R *tri_recs;            // (may be null) [T][24] = p0 p1 p2 n0 n1 n2 (3 each) uv0 uv1 uv2 (2 each): one 96-byte (f32) / 192-byte (f64) record per triangle
    const DMaterial *materials;
    const DTexture<R> *textures;
    const DImage *img3, *img1;
    const typename TexelTypes<R>::T3 *texels3;
    const typename TexelTypes<R>::T1 *texels1;
    const R *light_intensity;     // [n_lights][3]
    // light sampling tables (the `path` integrator; gdpt_light.cuh): TableDist1D over the light powers
    // (scene.cpp:46-52), area CDFs of the emissive meshes (triangle_mesh.inl:60-75), total areas
    const R *light_cdf;           // [n_lights + 1]
    const R *light_pmf;           // [n_lights]
    const int *light_shape;       // [n_lights] shape id, -1 for an environment map
    const R *tri_cdf;             // pooled; DShape::cdf_off
    const R *shape_area;          // [n_shapes] surface_area(shape)
    // environment map (src/light.h:23-30, lights/envmap.inl): image texture, scale, transforms and the TableDist2D
    // (table_dist.cpp:45-151) over luminance * sin(elevation) of its level-0 texels; env_light < 0: none
    int env_light, env_tex, env_w, env_h;
    R env_scale;
    R env_to_world[16], env_to_local[16];
    const R *env_cdf_rows;        // [h][w + 1]
    const R *env_pdf_rows;        // [h][w]
    const R *env_cdf_marg;        // [h + 1]
    const R *env_pdf_marg;        // [h]
    R eps;                        // get_intersection_epsilon (scene.h:104-106)
    int max_depth, rr_depth;
};

template <typename R> struct Vertex {
    V3<R> position, gn;
    Frame<R> sf;
    V2<R> st, uv;
    R uv_screen_size, mean_curvature, ray_radius;
    int shape_id, prim_id, material_id;
};

struct Hit { float t, u, v; int shape_id, prim_id; };   // shape_id < 0: miss
// The hit as it travels through HBM between the trace and the shade kernels: one aligned 32-byte sector, moved as
// two 16-byte vectors (the sorted trace launch writes hits in scattered order; five 4-byte stores per ray into an
// unaligned 20-byte slot cost five L2 write requests and straddle sectors)
struct alignas(32) HitRec { float t, u, v; int shape_id; int prim_id, pad0, pad1, pad2; };
GD_D void store_hit(HitRec *dst, const Hit &h) {
    int4 *d = reinterpret_cast<int4 *>(dst);
    d[0] = make_int4(__float_as_int(h.t), __float_as_int(h.u), __float_as_int(h.v), h.shape_id);
    d[1] = make_int4(h.prim_id, 0, 0, 0);
}
GD_D Hit load_hit(const HitRec *src) {
    const int4 *s = reinterpret_cast<const int4 *>(src);
    int4 a = s[0]; int b = reinterpret_cast<const int *>(src)[4];
    Hit h; h.t = __int_as_float(a.x); h.u = __int_as_float(a.y); h.v = __int_as_float(a.z); h.shape_id = a.w; h.prim_id = b;
    return h;
}

// ---------------------------------------------------------------- camera
template <typename R> GD_D V3<R> xform_point(const R *m, const V3<R> &p) {
    R x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    R y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    R z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    R w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    R inv_w = R(1) / w;
    return V3<R>(x * inv_w, y * inv_w, z * inv_w);
}
template <typename R> GD_D V3<R> xform_vector(const R *m, const V3<R> &v) {
    return V3<R>(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

template <typename R> GD_D V2<R> sample_filter(int kind, R param, R u0, R u1) {
    if (kind == FILTER_BOX) {
        return V2<R>((R(2) * u0 - R(1)) * (param / 2), (R(2) * u1 - R(1)) * (param / 2));
    } else if (kind == FILTER_GAUSSIAN) {
        R r = param * sqrt(R(-2) * log(rmax(u0, R(1e-8))));
        return V2<R>(r * cos(R(2) * Consts<R>::PI * u1), r * sin(R(2) * Consts<R>::PI * u1));
    } else {
        R h = param / 2;
        R x = u0 < R(0.5) ? h * (sqrt(R(2) * u0) - R(1)) : h * (R(1) - sqrt(R(1) - R(2) * (u0 - R(0.5))));
        R y = u1 < R(0.5) ? h * (sqrt(R(2) * u1) - R(1)) : h * (R(1) - sqrt(R(1) - R(2) * (u1 - R(0.5))));
        return V2<R>(x, y);
    }
}

// camera.cpp:23-47. screen_pos in [0,1]^2 (may leave it for edge offsets).
template <typename R> GD_D void sample_primary(const DScene<R> &sc, R sx, R sy, V3<R> &org, V3<R> &dir) {
    R px = sx * sc.width, py = sy * sc.height;
    R fx = floor(px), fy = floor(py);
    R dx = px - fx, dy = py - fy;
    V2<R> off = sample_filter<R>(sc.filter_kind, sc.filter_param, dx, dy);
    R rx = (fx + R(0.5) + off.x) / sc.width;
    R ry = (fy + R(0.5) + off.y) / sc.height;
    V3<R> pt = xform_point<R>(sc.s2c, V3<R>(rx, ry, R(0)));
    V3<R> d = normalize(pt);
    org = xform_point<R>(sc.c2w, V3<R>(R(0), R(0), R(0)));
    dir = normalize(xform_vector<R>(sc.c2w, d));
}

// ---------------------------------------------------------------- ray differentials (ray.h)
template <typename R> GD_D R rd_reflect(R radius, R spread, R mean_curvature, R roughness) {
    R spec = spread + R(2) * mean_curvature * radius;
    R diff = R(0.2);
    return fmax(spec * (R(1) - roughness) + diff * roughness, R(0));
}
template <typename R> GD_D R rd_refract(R radius, R spread, R mean_curvature, R eta, R roughness) {
    R spec = (spread + R(2) * mean_curvature * radius) / eta;
    R diff = R(0.2);
    return fmax(spec * (R(1) - roughness) + diff * roughness, R(0));
}

// ---------------------------------------------------------------- sphere (sphere.inl, double like the reference)
GD_D bool solve_quadratic_d(double a, double b, double c, double *t0, double *t1) {
    if (a == 0) {
        if (b == 0) return false;
        *t0 = *t1 = -c / b;
        return true;
    }
    double disc = b * b - 4 * a * c;
    if (disc < 0) return false;
    double rd = sqrt(disc);
    if (b >= 0) { *t0 = (-b - rd) / (2 * a); *t1 = 2 * c / (-b - rd); }
    else { *t0 = 2 * c / (-b + rd); *t1 = (-b + rd) / (2 * a); }
    return true;
}
// Ray given as the float values Embree would see. Returns t (double) or -1.
GD_D double sphere_hit_d(const double *sph, const float *o, const float *d, float tnear, float tfar) {
    double vx = (double)o[0] - sph[0], vy = (double)o[1] - sph[1], vz = (double)o[2] - sph[2];
    double dx = d[0], dy = d[1], dz = d[2];
    double A = dx * dx + dy * dy + dz * dz;
    double B = 2 * (dx * vx + dy * vy + dz * vz);
    double C = (vx * vx + vy * vy + vz * vz) - sph[3] * sph[3];
    double t0, t1;
    if (!solve_quadratic_d(A, B, C, &t0, &t1)) return -1.0;
    if (t0 > t1) { double s = t0; t0 = t1; t1 = s; }
    double t = -1;
    if (t0 >= (double)tnear && t0 < (double)tfar) t = t0;
    if (t1 >= (double)tnear && t1 < (double)tfar && t < 0) t = t1;
    if (t >= (double)tnear && t < (double)tfar) return t;
    return -1.0;
}

// 256-bit global load (sm_100: LDG.E.ENL2.256)
// eight floats / four doubles of a 32-byte aligned address, into scalars (an array filled through a pointer stays in local memory)
// (64-bit operands: the eight-float form of this asm crashes cicc 12.9 inside make_vertex_impl)
GD_D void scene_ld256(const float *p, float4 &a, float4 &b) {
    unsigned long long x0, x1, x2, x3;
    asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(x0), "=l"(x1), "=l"(x2), "=l"(x3) : "l"(p));
    a.x = __uint_as_float((unsigned)x0); a.y = __uint_as_float((unsigned)(x0 >> 32)); a.z = __uint_as_float((unsigned)x1); a.w = __uint_as_float((unsigned)(x1 >> 32));
    b.x = __uint_as_float((unsigned)x2); b.y = __uint_as_float((unsigned)(x2 >> 32)); b.z = __uint_as_float((unsigned)x3); b.w = __uint_as_float((unsigned)(x3 >> 32));
}
GD_D void scene_ld256(const double *p, double &a, double &b, double &c, double &d) {
    unsigned long long x0, x1, x2, x3;
    asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(x0), "=l"(x1), "=l"(x2), "=l"(x3) : "l"(p));
    a = __longlong_as_double((long long)x0); b = __longlong_as_double((long long)x1); c = __longlong_as_double((long long)x2); d = __longlong_as_double((long long)x3);
}
// one triangle record -> its nine vectors
GD_D void load_tri_rec(const float *r, V3<float> &p0, V3<float> &p1, V3<float> &p2, V3<float> &n0, V3<float> &n1, V3<float> &n2, V2<float> &uv0, V2<float> &uv1, V2<float> &uv2) {
    float4 q0, q1, q2, q3, q4, q5;
    scene_ld256(r, q0, q1); scene_ld256(r + 8, q2, q3); scene_ld256(r + 16, q4, q5);
    p0 = V3<float>(q0.x, q0.y, q0.z); p1 = V3<float>(q0.w, q1.x, q1.y); p2 = V3<float>(q1.z, q1.w, q2.x);
    n0 = V3<float>(q2.y, q2.z, q2.w); n1 = V3<float>(q3.x, q3.y, q3.z); n2 = V3<float>(q3.w, q4.x, q4.y);
    uv0 = V2<float>(q4.z, q4.w); uv1 = V2<float>(q5.x, q5.y); uv2 = V2<float>(q5.z, q5.w);
}
GD_D void load_tri_rec(const double *r, V3<double> &p0, V3<double> &p1, V3<double> &p2, V3<double> &n0, V3<double> &n1, V3<double> &n2, V2<double> &uv0, V2<double> &uv1, V2<double> &uv2) {
    scene_ld256(r, p0.x, p0.y, p0.z, p1.x); scene_ld256(r + 4, p1.y, p1.z, p2.x, p2.y); scene_ld256(r + 8, p2.z, n0.x, n0.y, n0.z);
    scene_ld256(r + 12, n1.x, n1.y, n1.z, n2.x); scene_ld256(r + 16, n2.y, n2.z, uv0.x, uv0.y); scene_ld256(r + 20, uv1.x, uv1.y, uv2.x, uv2.y);
}
// One thread per triangle: the vertex data a hit needs, gathered once at scene creation into one 96-byte record, so that
// make_vertex fetches it with three (f32) or six (f64) 256-bit loads instead of 3 index loads followed by 24 scalar gathers (two dependent
// levels of long-scoreboard stalls in k_primary / k_bounce / k_tail).  Same float values as the vertex pools: same results.
template <typename T>
static __global__ void k_build_tri_recs(int n_shapes, const DShape *shapes, const int *indices, const T *positions, const T *normals, const T *uvs,
                                        long long n_tris, T *out) {
    const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (t >= n_tris) return;
    int s = -1;                                                // the mesh that owns triangle t (a warp's triangles share it: uniform loop)
    for (int i = 0; i < n_shapes; i++) { const DShape sh = shapes[i]; if (sh.kind == SHAPE_MESH && t >= sh.tri_off && t < (long long)sh.tri_off + sh.n_tris) { s = i; break; } }
    T *o = out + 24 * t;
    if (s < 0) { for (int k = 0; k < 24; k++) o[k] = T(0); return; }
    const DShape sh = shapes[s];
    const int *ix = indices + 3 * t;
    const T *P = positions + 3 * (size_t)sh.vtx_off;
    for (int v = 0; v < 3; v++) for (int k = 0; k < 3; k++) o[3 * v + k] = P[3 * ix[v] + k];
    if (sh.flags & 1) { const T *N = normals + 3 * (size_t)sh.nrm_off; for (int v = 0; v < 3; v++) for (int k = 0; k < 3; k++) o[9 + 3 * v + k] = N[3 * ix[v] + k]; }
    else for (int k = 0; k < 9; k++) o[9 + k] = T(0);
    if (sh.flags & 2) { const T *U = uvs + 2 * (size_t)sh.uv_off; for (int v = 0; v < 3; v++) for (int k = 0; k < 2; k++) o[18 + 2 * v + k] = U[2 * ix[v] + k]; }
    else for (int k = 0; k < 6; k++) o[18 + k] = T(0);
}

// ---------------------------------------------------------------- vertex construction
// org/dir: the ray in working precision (what lajolla holds in double); tnear/tfar_in: the
// values that were handed to the traversal (needed to re-derive the double sphere root).
template <typename R>
GD_D void make_vertex_impl(const DScene<R> &sc, const V3<R> &org, const V3<R> &dir, float tnear, float tfar_in,
                      const Hit &hit, R rd_radius, R rd_spread, Vertex<R> &v) {
    const DShape sh = sc.shapes[hit.shape_id];
    v.shape_id = hit.shape_id; v.prim_id = hit.prim_id; v.material_id = sh.material_id;
    v.position = org + dir * R(hit.t);
    R inv_uv_size;
    if (sh.kind == SHAPE_MESH) {
        int i0 = 0, i1 = 0, i2 = 0;
        V3<R> p0, p1, p2, n0, n1, n2;
        V2<R> uv0, uv1, uv2;
        const bool packed = sc.tri_recs != nullptr;
        if (packed) {
            load_tri_rec(sc.tri_recs + 24 * (size_t)(sh.tri_off + hit.prim_id), p0, p1, p2, n0, n1, n2, uv0, uv1, uv2);
        } else {
            const int *ix = sc.indices + 3 * (size_t)(sh.tri_off + hit.prim_id);
            i0 = ix[0]; i1 = ix[1]; i2 = ix[2];
            const R *P = sc.positions + 3 * (size_t)sh.vtx_off;
            p0 = V3<R>(P[3 * i0], P[3 * i0 + 1], P[3 * i0 + 2]); p1 = V3<R>(P[3 * i1], P[3 * i1 + 1], P[3 * i1 + 2]); p2 = V3<R>(P[3 * i2], P[3 * i2 + 1], P[3 * i2 + 2]);
        }
        // Embree's Ng: float vertices, result rounded to float (intersection.cpp:41)
        {
            V3<double> a((double)(float)p0.x, (double)(float)p0.y, (double)(float)p0.z);
            V3<double> b((double)(float)p1.x, (double)(float)p1.y, (double)(float)p1.z);
            V3<double> c((double)(float)p2.x, (double)(float)p2.y, (double)(float)p2.z);
            V3<double> ng = cross(b - a, c - a);
            v.gn = normalize(V3<R>((R)(float)ng.x, (R)(float)ng.y, (R)(float)ng.z));
        }
        v.st = V2<R>((R)hit.u, (R)hit.v);
        if (sh.flags & 2) {
            if (!packed) {
                const R *U = sc.uvs + 2 * (size_t)sh.uv_off;
                uv0 = V2<R>(U[2 * i0], U[2 * i0 + 1]); uv1 = V2<R>(U[2 * i1], U[2 * i1 + 1]); uv2 = V2<R>(U[2 * i2], U[2 * i2 + 1]);
            }
        } else { uv0 = V2<R>(0, 0); uv1 = V2<R>(1, 0); uv2 = V2<R>(1, 1); }
        R b0 = R(1) - v.st.x - v.st.y;
        v.uv = V2<R>(b0 * uv0.x + v.st.x * uv1.x + v.st.y * uv2.x, b0 * uv0.y + v.st.x * uv1.y + v.st.y * uv2.y);
        V2<R> duvds(uv2.x - uv0.x, uv2.y - uv0.y), duvdt(uv2.x - uv1.x, uv2.y - uv1.y);
        R det = duvds.x * duvdt.y - duvdt.x * duvds.y;
        R dsdu = duvdt.y / det, dtdu = -duvds.y / det, dsdv = duvdt.x / det, dtdv = -duvds.x / det;
        V3<R> dpdu, dpdv;
        if (fabs(det) > R(1e-8f)) {
            V3<R> dpds = p2 - p0, dpdt = p2 - p1;
            dpdu = dpds * dsdu + dpdt * dtdu;
            dpdv = dpds * dsdv + dpdt * dtdv;
        } else {
            coordinate_system(v.gn, dpdu, dpdv);
        }
        V3<R> sn = v.gn; R mc = R(0);
        V3<R> tangent, bitangent;
        if (sh.flags & 1) {
            if (!packed) {
                const R *N = sc.normals + 3 * (size_t)sh.nrm_off;
                n0 = V3<R>(N[3 * i0], N[3 * i0 + 1], N[3 * i0 + 2]); n1 = V3<R>(N[3 * i1], N[3 * i1 + 1], N[3 * i1 + 2]); n2 = V3<R>(N[3 * i2], N[3 * i2 + 1], N[3 * i2 + 2]);
            }
            sn = normalize(b0 * n0 + v.st.x * n1 + v.st.y * n2);
            tangent = normalize(dpdu - sn * dot(sn, dpdu));
            V3<R> dnds = n2 - n0, dndt = n2 - n1;
            V3<R> dndu = dnds * dsdu + dndt * dtdu;
            V3<R> dndv = dnds * dsdv + dndt * dtdv;
            bitangent = normalize(cross(sn, tangent));
            mc = (dot(dndu, tangent) + dot(dndv, bitangent)) / R(2);
        } else {
            tangent = normalize(dpdu - sn * dot(sn, dpdu));
            bitangent = normalize(cross(sn, tangent));
        }
        v.sf = Frame<R>(tangent, bitangent, sn);
        v.mean_curvature = mc;
        inv_uv_size = rmax(length(dpdu), length(dpdv));
    } else {
        const double *sph = sc.spheres + 4 * (size_t)hit.shape_id;
        float o[3] = {(float)org.x, (float)org.y, (float)org.z}, d[3] = {(float)dir.x, (float)dir.y, (float)dir.z};
        double t = sphere_hit_d(sph, o, d, tnear, tfar_in);
        double px = (double)o[0] + t * (double)d[0], py = (double)o[1] + t * (double)d[1], pz = (double)o[2] + t * (double)d[2];
        double gx = px - sph[0], gy = py - sph[1], gz = pz - sph[2];
        v.gn = normalize(V3<R>((R)(float)gx, (R)(float)gy, (R)(float)gz));
        double inv_r = 1.0 / sph[3];
        double cx = gx * inv_r, cy = gy * inv_r, cz = gz * inv_r;
        double elev = acos(cy < -1.0 ? -1.0 : (cy > 1.0 ? 1.0 : cy));
        double azim = atan2(cz, cx);
        v.st = V2<R>((R)(float)(azim / (2.0 * 3.14159265358979323846)), (R)(float)(elev / 3.14159265358979323846));
        R rad = (R)sph[3];
        V3<R> dpdu(-rad * sin(v.st.x) * sin(v.st.y), rad * cos(v.st.x) * sin(v.st.y), R(0));
        V3<R> dpdv(rad * cos(v.st.x) * cos(v.st.y), rad * sin(v.st.x) * cos(v.st.y), -rad * sin(v.st.y));
        V3<R> tangent = normalize(dpdu - v.gn * dot(v.gn, dpdu));
        v.sf = Frame<R>(tangent, normalize(cross(v.gn, tangent)), v.gn);
        v.uv = v.st;
        v.mean_curvature = R(1) / rad;
        inv_uv_size = (length(dpdu) + length(dpdv)) / R(2);
    }
    v.ray_radius = rd_radius + rd_spread * distance(org, v.position);
    v.uv_screen_size = v.ray_radius / inv_uv_size;
    if (dot(v.gn, v.sf.n) < R(0)) v.gn = -v.gn;
}

template <typename R>
GD_BIG void make_vertex(const DScene<R> &sc, const V3<R> &org, const V3<R> &dir, float tnear, float tfar_in,
                        const Hit &hit, R rd_radius, R rd_spread, Vertex<R> &v) {
    make_vertex_impl<R>(sc, org, dir, tnear, tfar_in, hit, rd_radius, rd_spread, v);
}
// INL != 0: inline the construction into the caller (specialised wavefront kernels)
template <typename R, int INL>
GD_D void make_vertex_ms(const DScene<R> &sc, const V3<R> &org, const V3<R> &dir, float tnear, float tfar_in,
                         const Hit &hit, R rd_radius, R rd_spread, Vertex<R> &v) {
    if (INL) make_vertex_impl<R>(sc, org, dir, tnear, tfar_in, hit, rd_radius, rd_spread, v);
    else make_vertex<R>(sc, org, dir, tnear, tfar_in, hit, rd_radius, rd_spread, v);
}

// ---------------------------------------------------------------- textures
GD_D int imodulo(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }
template <typename R> GD_D R rmodulo(R a, R b) { R r = fmod(a, b); return r < R(0) ? r + b : r; }

template <typename R> GD_D V3<R> fetch3(const DScene<R> &sc, long long idx);
template <> GD_D V3<double> fetch3<double>(const DScene<double> &sc, long long idx) {
    const double *p = sc.texels3 + 3 * idx; return V3<double>(p[0], p[1], p[2]);
}
template <> GD_D V3<float> fetch3<float>(const DScene<float> &sc, long long idx) {
    float4 t = __ldg(sc.texels3 + idx); return V3<float>(t.x, t.y, t.z);
}
template <typename R> GD_D R fetch1(const DScene<R> &sc, long long idx) { return (R)sc.texels1[idx]; }

// bilinear lookup at an integer level (mipmap.h:46-69), generic over the fetch
template <typename R, typename T, typename F>
GD_D T mip_bilinear(const DImage &im, R u, R v, int level, F fetch) {
    int w = im.w[level], h = im.h[level];
    u = u * w - R(0.5);
    v = v * h - R(0.5);
    // modulo(int(u), w) etc. of mipmap.h:52-55; the integer divisions are skipped where they cannot change the value
    const int iu = (int)u, iv = (int)v;
    const int ufi = (iu >= 0 && iu < w) ? iu : imodulo(iu, w), vfi = (iv >= 0 && iv < h) ? iv : imodulo(iv, h);
    const int uci = ufi + 1 == w ? 0 : ufi + 1, vci = vfi + 1 == h ? 0 : vfi + 1;
    R uo = u - ufi, vo = v - vfi;
    long long base = im.off[level];
    T ff = fetch(base + (long long)vfi * w + ufi), fc = fetch(base + (long long)vci * w + ufi);
    T cf = fetch(base + (long long)vfi * w + uci), cc = fetch(base + (long long)vci * w + uci);
    return ff * (R(1) - uo) * (R(1) - vo) + fc * (R(1) - uo) * vo + cf * uo * (R(1) - vo) + cc * uo * vo;
}
template <typename R, typename T, typename F>
GD_D T mip_trilinear(const DImage &im, R u, R v, R level, F fetch) {
    if (level <= R(0)) return mip_bilinear<R, T>(im, u, v, 0, fetch);
    if (level < R(im.n_levels - 1)) {
        int fl = (int)floor(level); fl = fl < 0 ? 0 : (fl > im.n_levels - 1 ? im.n_levels - 1 : fl);
        int cl = fl + 1 > im.n_levels - 1 ? im.n_levels - 1 : fl + 1;
        R lo = level - fl;
        return mip_bilinear<R, T>(im, u, v, fl, fetch) * (R(1) - lo) + mip_bilinear<R, T>(im, u, v, cl, fetch) * lo;
    }
    return mip_bilinear<R, T>(im, u, v, im.n_levels - 1, fetch);
}

template <typename R> GD_D bool checker_first(const DTexture<R> &t, R lu, R lv) {
    int x = 2 * imodulo((int)(lu * 2), 2) - 1, y = 2 * imodulo((int)(lv * 2), 2) - 1;
    return x * y == 1;
}

// texture.h:123-159
template <typename R> GD_D V3<R> eval_tex3(const DScene<R> &sc, int tex_id, const V2<R> &uv, R footprint) {
    const DTexture<R> &t = sc.textures[tex_id];
    if (t.kind == TEX_CONST) return V3<R>(t.v0[0], t.v0[1], t.v0[2]);
    R lu = rmodulo(uv.x * t.uscale + t.uoffset, R(1)), lv = rmodulo(uv.y * t.vscale + t.voffset, R(1));
    if (t.kind == TEX_IMAGE) {
        const DImage &im = sc.img3[t.image_id];
        R sf = (R)(im.w[0] > im.h[0] ? im.w[0] : im.h[0]) * rmax(t.uscale, t.vscale) * footprint;
        R level = log2(rmax(sf, R(1e-8f)));
        return mip_trilinear<R, V3<R>>(im, lu, lv, level, [&](long long i) { return fetch3<R>(sc, i); });
    }
    return checker_first(t, lu, lv) ? V3<R>(t.v0[0], t.v0[1], t.v0[2]) : V3<R>(t.v1[0], t.v1[1], t.v1[2]);
}
template <typename R> GD_D R eval_tex1(const DScene<R> &sc, int tex_id, const V2<R> &uv, R footprint) {
    const DTexture<R> &t = sc.textures[tex_id];
    if (t.kind == TEX_CONST) return t.v0[0];
    R lu = rmodulo(uv.x * t.uscale + t.uoffset, R(1)), lv = rmodulo(uv.y * t.vscale + t.voffset, R(1));
    if (t.kind == TEX_IMAGE) {
        const DImage &im = sc.img1[t.image_id];
        R sf = (R)(im.w[0] > im.h[0] ? im.w[0] : im.h[0]) * rmax(t.uscale, t.vscale) * footprint;
        R level = log2(rmax(sf, R(1e-8f)));
        return mip_trilinear<R, R>(im, lu, lv, level, [&](long long i) { return fetch1<R>(sc, i); });
    }
    return checker_first(t, lu, lv) ? t.v0[0] : t.v1[0];
}

// ---------------------------------------------------------------- emission
// intersection.cpp:87-98 + diffuse_area_light.inl:15-20 (envmaps are never reached by gradpath)
template <typename R> GD_D V3<R> emission(const DScene<R> &sc, const Vertex<R> &v, const V3<R> &view_dir) {
    int lid = sc.shapes[v.shape_id].area_light_id;
    if (lid < 0 || dot(v.gn, view_dir) <= R(0)) return V3<R>(0, 0, 0);
    const R *I = sc.light_intensity + 3 * lid;
    return V3<R>(I[0], I[1], I[2]);
}
template <typename R> GD_D bool is_light(const DScene<R> &sc, int shape_id) { return sc.shapes[shape_id].area_light_id >= 0; }

} // namespace gdpt
