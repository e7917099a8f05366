// Validation kernels behind gdpt_sample_primary / gdpt_intersect / gdpt_bsdf / gdpt_pcg_fill:
// thin wrappers that expose the renderer's own device functions so the parity tests can
// compare them one call at a time with the reference (oracle/ref.py).
#pragma once
#include "gdpt_shade.cuh"
#include "gdpt_light.cuh"

namespace gdpt {

template <typename R>
static __global__ void k_dbg_sample_primary(DScene<R> sc, int n, const double *screen_pos, double *rays) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3<R> o, d;
    sample_primary<R>(sc, (R)screen_pos[2 * i], (R)screen_pos[2 * i + 1], o, d);
    double *r = rays + 6 * i;
    r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = d.x; r[4] = d.y; r[5] = d.z;
}

template <typename R, bool EXACT>
static __global__ void __launch_bounds__(TRACE_BLOCK) k_dbg_intersect(DScene<R> sc, DBvh bvh, int n, const double *rays, const double *rdiff, int *ids, double *vtx) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *q = rays + 8 * i;
    V3<R> org((R)q[0], (R)q[1], (R)q[2]), dir((R)q[3], (R)q[4], (R)q[5]);
    float o[3] = {(float)q[0], (float)q[1], (float)q[2]}, d[3] = {(float)q[3], (float)q[4], (float)q[5]};
    float tnear = (float)q[6], tfar = (float)q[7];
    Hit h = trace_closest<EXACT, false>(bvh, o, d, tnear, tfar, stack);
    double *out = vtx + 24 * i;
    if (h.shape_id < 0) { ids[3 * i] = ids[3 * i + 1] = ids[3 * i + 2] = -1; for (int k = 0; k < 24; k++) out[k] = 0; return; }
    Vertex<R> v;
    make_vertex<R>(sc, org, dir, tnear, tfar, h, rdiff ? (R)rdiff[2 * i] : R(0), rdiff ? (R)rdiff[2 * i + 1] : R(0), v);
    ids[3 * i] = v.shape_id; ids[3 * i + 1] = v.prim_id; ids[3 * i + 2] = v.material_id;
    const V3<R> *vs[5] = {&v.position, &v.gn, &v.sf.x, &v.sf.y, &v.sf.n};
    for (int k = 0; k < 5; k++) { out[3 * k] = vs[k]->x; out[3 * k + 1] = vs[k]->y; out[3 * k + 2] = vs[k]->z; }
    out[15] = v.st.x; out[16] = v.st.y; out[17] = v.uv.x; out[18] = v.uv.y;
    out[19] = v.uv_screen_size; out[20] = v.mean_curvature; out[21] = v.ray_radius; out[22] = h.t; out[23] = 0;
}

template <typename R>
static __global__ void k_dbg_bsdf(DScene<R> sc, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *q = vtx + 15 * i;
    Vertex<R> v;
    v.position = V3<R>(0, 0, 0);
    v.gn = V3<R>((R)q[0], (R)q[1], (R)q[2]);
    v.sf = Frame<R>(V3<R>((R)q[3], (R)q[4], (R)q[5]), V3<R>((R)q[6], (R)q[7], (R)q[8]), V3<R>((R)q[9], (R)q[10], (R)q[11]));
    v.uv = V2<R>((R)q[12], (R)q[13]); v.uv_screen_size = (R)q[14]; v.st = v.uv;
    v.mean_curvature = 0; v.ray_radius = 0; v.material_id = mat_id; v.shape_id = 0; v.prim_id = 0;
    V3<R> din((R)dirs[6 * i], (R)dirs[6 * i + 1], (R)dirs[6 * i + 2]), dout((R)dirs[6 * i + 3], (R)dirs[6 * i + 4], (R)dirs[6 * i + 5]);
    double *o = out + 10 * i;
    V3<R> f = bsdf_eval<R>(sc, v, din, dout);
    o[0] = f.x; o[1] = f.y; o[2] = f.z;
    o[3] = bsdf_pdf<R>(sc, v, din, dout);
    BsdfSample<R> s;
    if (bsdf_sample<R>(sc, v, din, (R)rnd[3 * i], (R)rnd[3 * i + 1], (R)rnd[3 * i + 2], s)) {
        o[4] = 1; o[5] = s.dir_out.x; o[6] = s.dir_out.y; o[7] = s.dir_out.z; o[8] = s.eta; o[9] = s.roughness;
    } else { o[4] = 0; o[5] = o[6] = o[7] = o[8] = o[9] = 0; }
}

static __global__ void k_dbg_pcg(unsigned long long stream, unsigned long long seed, int n, double *out, unsigned int *raw) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    Pcg r = pcg_init(stream, seed);
    for (int i = 0; i < n; i++) {
        Pcg c = r;
        if (raw) raw[i] = pcg_next(c);
        out[i] = pcg_real<double>(r);
    }
}

static __global__ void k_fill_default_records(int n, double *rec) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double *o = rec + (size_t)i * 23;
    for (int k = 0; k < 18; k++) o[k] = 0.0;
    for (int k = 18; k < 23; k++) o[k] = 1.0;
}

// sample_light + sample_point_on_light + pdf_point_on_light + light_pmf (scene.cpp:61-67, light.cpp:57-71):
// rnd[n][4] = light_w, uv0, uv1, shape_w; out[n][9] = light id, position3, normal3, pdf of the point, pmf of the light
template <typename R>
static __global__ void k_dbg_sample_light(DScene<R> sc, int n, const double *ref_points, const double *rnd, double *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3<R> ref((R)ref_points[3 * i], (R)ref_points[3 * i + 1], (R)ref_points[3 * i + 2]);
    const double *q = rnd + 4 * i;
    int id = sample_light<R>(sc, (R)q[0]);
    LightPoint<R> p = sample_point_on_light<R>(sc, id, ref, (R)q[1], (R)q[2], (R)q[3]);
    double *o = out + 9 * i;
    o[0] = id; o[1] = p.position.x; o[2] = p.position.y; o[3] = p.position.z; o[4] = p.normal.x; o[5] = p.normal.y; o[6] = p.normal.z;
    o[7] = pdf_point_on_light<R>(sc, id, p, ref); o[8] = light_pmf<R>(sc, id);
}

// aux_render (/root/reference/src/render.cpp:15-72): one ray through every pixel centre, a function of the primary
// vertex per Integrator value (src/scene.h:14-23): 0 depth, 1 shading normal, 2 mean curvature, 3 ray differential
// (radius, spread, 0), 4 mip level of the material's main texture (material.cpp:121-123).  img: W*H*3 doubles.
template <typename R, bool EXACT>
static __global__ void __launch_bounds__(TRACE_BLOCK) k_aux(DScene<R> sc, DBvh bvh, int kind, double *img) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sc.width * sc.height) return;
    int x = i % sc.width, y = i / sc.width;
    V3<R> org, dir;
    sample_primary<R>(sc, (R(x) + R(0.5)) / R(sc.width), (R(y) + R(0.5)) / R(sc.height), org, dir);
    const R spread0 = R(0.25) / R(sc.width > sc.height ? sc.width : sc.height);
    float o[3] = {(float)org.x, (float)org.y, (float)org.z}, d[3] = {(float)dir.x, (float)dir.y, (float)dir.z};
    Hit h = trace_closest<EXACT, false>(bvh, o, d, 0.f, __int_as_float(0x7f800000), stack);
    V3<R> color(R(0), R(0), R(0));
    if (h.shape_id >= 0) {
        Vertex<R> v;
        make_vertex<R>(sc, org, dir, 0.f, __int_as_float(0x7f800000), h, R(0), spread0, v);
        if (kind == 0) { R dist = distance(v.position, org); color = V3<R>(dist, dist, dist); }
        else if (kind == 1) color = v.sf.n;
        else if (kind == 2) color = V3<R>(v.mean_curvature, v.mean_curvature, v.mean_curvature);
        else if (kind == 3) color = V3<R>(R(0), spread0, R(0));
        else if (kind == 4) {
            const DMaterial &m = sc.materials[v.material_id];
            int tid = m.type == MAT_DISNEY_CLEARCOAT ? -1 : m.tex[0];
            if (tid >= 0 && sc.textures[tid].kind == TEX_IMAGE) {
                const DTexture<R> &t = sc.textures[tid];
                const DImage &im = sc.img3[t.image_id];
                R sf = (R)(im.w[0] > im.h[0] ? im.w[0] : im.h[0]) * rmax(t.uscale, t.vscale) * v.uv_screen_size;
                R level = log2(rmax(sf, R(1e-8f)));
                color = V3<R>(level, level, level);
            }
        }
    }
    img[3 * (size_t)i] = color.x; img[3 * (size_t)i + 1] = color.y; img[3 * (size_t)i + 2] = color.z;
}

} // namespace gdpt
