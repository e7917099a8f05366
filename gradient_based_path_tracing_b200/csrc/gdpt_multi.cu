// gradient_path_render over several GPUs of one box inside ONE process, behind the C ABI (gdpt_multi_*): one host thread
// per device, the scene replicated, the BVH built once and copied device to device, the film shared as interleaved
// 4-row strips (strip i -> device i mod N: the load balance of the reference's dynamic tile queue, parallel.cpp:183-237),
// the strips of the five accumulation images (render.cpp:264-268) brought to device 0 by strided peer copies over
// NVLink, merge + screened-Poisson solve (render.cpp:340-353) there.  No collective library is needed: the exchange is
// a gather of disjoint rows.  (bench.py's N > 1 runs use one PROCESS per GPU and NCCL instead — multi.py — because that is
// how the driver launches them; the arithmetic underneath is the same entry points.)
#include <thread>

#include "gdpt_context.cuh"

namespace gdpt {
int poisson_device(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out);
int merge_device(gdpt_ctx *ctx, int w, int h, int ya, int yb, const double *d_cx0, const double *d_cx1, const double *d_cy0, const double *d_cy1, double *d_cx, double *d_cy);
}  // namespace gdpt

struct gdpt_multi {
    std::vector<gdpt_ctx *> ctx;       // ctx[0] owns the merged images and solves
    std::vector<int> devices;
    int width = 0, height = 0;
    std::vector<std::vector<gdpt::DevBuf>> acc;     // acc[k][0..4]: the five accumulation images on device k (k > 0)
    gdpt_stats stats;
};

using namespace gdpt;

extern "C" {

void gdpt_multi_destroy(gdpt_multi *m) {
    if (!m) return;
    for (size_t k = 0; k < m->acc.size(); k++) { if (k < m->devices.size()) cudaSetDevice(m->devices[k]); for (auto &b : m->acc[k]) b.release(); }
    for (gdpt_ctx *c : m->ctx) gdpt_destroy(c);
    delete m;
}

int gdpt_multi_create(const gdpt_scene_desc *scene, int n_devices, const int *devices, gdpt_multi **out) {
    if (out) *out = nullptr;
    if (!scene || !out || n_devices < 1) return fail(GDPT_ERR_ARG, "bad argument");
    const int ndev = gdpt_device_count();
    if (ndev < 1) return fail(GDPT_ERR_CUDA, "no CUDA device: libgdpt has no CPU fallback");
    if (n_devices > ndev) return fail(GDPT_ERR_ARG, "%d devices requested, %d present", n_devices, ndev);
    gdpt_multi *m = new gdpt_multi();
    memset(&m->stats, 0, sizeof(m->stats));
    m->width = scene->width; m->height = scene->height;
    for (int k = 0; k < n_devices; k++) m->devices.push_back(devices ? devices[k] : k);
    for (int k = 0; k < n_devices; k++) for (int j = 0; j < k; j++)
        if (m->devices[k] == m->devices[j] || m->devices[k] < 0 || m->devices[k] >= ndev) { delete m; return fail(GDPT_ERR_ARG, "bad device list"); }
    m->ctx.assign(n_devices, nullptr);
    m->acc.resize(n_devices);
    // device 0 builds the BVH (host threads); the others upload the scene meanwhile and adopt device 0's tree afterwards
    std::vector<int> rc(n_devices, 0);
    std::vector<std::string> err(n_devices);
    std::vector<std::thread> th;
    for (int k = 0; k < n_devices; k++)
        th.emplace_back([&, k] {
            rc[k] = gdpt_create_ex(scene, m->devices[k], k == 0 ? 0 : GDPT_CREATE_NO_BVH, &m->ctx[k]);
            if (rc[k]) err[k] = gdpt_last_error();
        });
    for (auto &t : th) t.join();
    for (int k = 0; k < n_devices; k++) if (rc[k]) { std::string e = err[k]; int r = rc[k]; gdpt_multi_destroy(m); return fail(r, "device %d: %s", k, e.c_str()); }
    for (int k = 1; k < n_devices; k++) {
        int can = 0;     // direct NVLink access where the topology allows it; the copies below work (staged) without it too
        cudaSetDevice(m->devices[0]);
        if (cudaDeviceCanAccessPeer(&can, m->devices[0], m->devices[k]) == cudaSuccess && can) cudaDeviceEnablePeerAccess(m->devices[k], 0);
        cudaSetDevice(m->devices[k]);
        if (cudaDeviceCanAccessPeer(&can, m->devices[k], m->devices[0]) == cudaSuccess && can) cudaDeviceEnablePeerAccess(m->devices[0], 0);
        // the buffers come from each device's default memory pool (gdpt_context.cuh), whose peer access is granted separately
        for (int a = 0; a < 2; a++) {
            const int owner = a == 0 ? m->devices[k] : m->devices[0], peer = a == 0 ? m->devices[0] : m->devices[k];
            cudaMemPool_t pool;
            if (cudaDeviceCanAccessPeer(&can, peer, owner) == cudaSuccess && can && cudaDeviceGetDefaultMemPool(&pool, owner) == cudaSuccess) {
                cudaMemAccessDesc desc; memset(&desc, 0, sizeof(desc));
                desc.location.type = cudaMemLocationTypeDevice; desc.location.id = peer; desc.flags = cudaMemAccessFlagsProtReadWrite;
                cudaMemPoolSetAccess(pool, &desc, 1);
            }
        }
        cudaGetLastError();
    }
    int64_t nn = 0, np = 0; int32_t depth = 0; void *dn = nullptr, *dp = nullptr;
    if (int r = gdpt_bvh_info(m->ctx[0], &nn, &np, &depth, &dn, &dp)) { gdpt_multi_destroy(m); return r; }
    for (int k = 1; k < n_devices; k++)
        if (int r = gdpt_bvh_adopt(m->ctx[k], dn, nn, dp, np, depth)) { std::string e = gdpt_last_error(); gdpt_multi_destroy(m); return fail(r, "device %d: %s", k, e.c_str()); }
    *out = m;
    return 0;
}

int gdpt_multi_device_count(gdpt_multi *m) { return m ? (int)m->ctx.size() : 0; }

int gdpt_multi_render(gdpt_multi *m, const gdpt_render_params *params, gdpt_images *out) {
    if (!m || !params) return fail(GDPT_ERR_ARG, "null argument");
    if (params->strip_count > 1 || params->y0 > 0 || (params->y1 > 0 && params->y1 < m->height))
        return fail(GDPT_ERR_ARG, "gdpt_multi_render shares the whole film between its devices: y0 / y1 / strip_* must be left 0");
    const int n = (int)m->ctx.size(), w = m->width, h = m->height;
    const size_t bytes = (size_t)w * h * 3 * sizeof(double);
    gdpt_ctx *c0 = m->ctx[0];
    GD_CUDA(cudaSetDevice(c0->device));
    {
        gdpt::DevBuf *bufs[] = {&c0->img, &c0->cx0, &c0->cy0, &c0->cx1, &c0->cy1, &c0->cx, &c0->cy, &c0->out};
        for (auto *b : bufs) if (int r = b->ensure(bytes)) return r;
    }
    std::vector<int> rc(n, 0);
    std::vector<std::string> err(n);
    std::vector<std::thread> th;
    for (int k = 0; k < n; k++)
        th.emplace_back([&, k] {
            gdpt_ctx *c = m->ctx[k];
            auto run = [&]() -> int {
                GD_CUDA(cudaSetDevice(c->device));
                double *img[5];
                if (k == 0) { gdpt::DevBuf *a[] = {&c->img, &c->cx0, &c->cy0, &c->cx1, &c->cy1}; for (int j = 0; j < 5; j++) img[j] = a[j]->as<double>(); }
                else {
                    m->acc[k].resize(5);
                    for (int j = 0; j < 5; j++) { if (int r = m->acc[k][j].ensure(bytes)) return r; img[j] = m->acc[k][j].as<double>(); }
                }
                for (int j = 0; j < 5; j++) GD_CUDA(cudaMemsetAsync(img[j], 0, bytes, c->stream));
                gdpt_render_params p = *params;
                p.y0 = 0; p.y1 = 0; p.strip_count = n; p.strip_phase = k;
                if (int r = gdpt_render_accumulate_device(c, &p, img[0], img[1], img[2], img[3], img[4])) return r;
                GD_CUDA(cudaStreamSynchronize(c->stream));
                return 0;
            };
            rc[k] = run();
            if (rc[k]) err[k] = gdpt_last_error();
        });
    for (auto &t : th) t.join();
    for (int k = 0; k < n; k++) if (rc[k]) return fail(rc[k], "device %d: %s", k, err[k].c_str());
    // gather: the strips device k owns (4 rows every 4n rows, starting at row 4k) -> the same rows of device 0's images
    GD_CUDA(cudaSetDevice(c0->device));
    cudaStream_t st = c0->stream;
    const size_t row = (size_t)w * 3 * sizeof(double), strip = 4 * row, pitch = strip * n;
    double *dst[5] = {c0->img.as<double>(), c0->cx0.as<double>(), c0->cy0.as<double>(), c0->cx1.as<double>(), c0->cy1.as<double>()};
    const int strips = (h + 3) / 4;
    for (int k = 1; k < n; k++) {
        const int mine = strips > k ? (strips - k + n - 1) / n : 0;       // strips k, k+n, ...
        if (mine == 0) continue;
        const int full = ((size_t)(k + (size_t)(mine - 1) * n) * 4 + 4 <= (size_t)h) ? mine : mine - 1;   // the film's last strip may be short
        for (int j = 0; j < 5; j++) {
            const char *s = (const char *)m->acc[k][j].p + (size_t)k * strip;
            char *d = (char *)dst[j] + (size_t)k * strip;
            if (full > 0) GD_CUDA(cudaMemcpy2DAsync(d, pitch, s, pitch, strip, full, cudaMemcpyDefault, st));
            if (full < mine) {
                const size_t off = (size_t)(mine - 1) * pitch, y = ((size_t)k + (size_t)(mine - 1) * n) * 4;
                GD_CUDA(cudaMemcpyAsync(d + off, s + off, (h - y) * row, cudaMemcpyDefault, st));
            }
        }
    }
    const double alpha = params->data_cost > 0 ? params->data_cost : 0.04;
    if (int r = merge_device(c0, w, h, 0, h, c0->cx0.as<double>(), c0->cx1.as<double>(), c0->cy0.as<double>(), c0->cy1.as<double>(), c0->cx.as<double>(), c0->cy.as<double>())) return r;
    if (int r = poisson_device(c0, w, h, c0->img.as<double>(), c0->cx.as<double>(), c0->cy.as<double>(), alpha, c0->out.as<double>())) return r;
    if (out) {
        if (out->primal) GD_CUDA(cudaMemcpyAsync(out->primal, c0->img.p, bytes, cudaMemcpyDeviceToHost, st));
        if (out->dx) GD_CUDA(cudaMemcpyAsync(out->dx, c0->cx.p, bytes, cudaMemcpyDeviceToHost, st));
        if (out->dy) GD_CUDA(cudaMemcpyAsync(out->dy, c0->cy.p, bytes, cudaMemcpyDeviceToHost, st));
        if (out->poisson) GD_CUDA(cudaMemcpyAsync(out->poisson, c0->out.p, bytes, cudaMemcpyDeviceToHost, st));
    }
    GD_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int gdpt_multi_get_stats(gdpt_multi *m, gdpt_stats *out) {
    if (!m || !out) return fail(GDPT_ERR_ARG, "null argument");
    memset(out, 0, sizeof(*out));
    for (gdpt_ctx *c : m->ctx) {
        gdpt_stats s;
        if (int r = gdpt_get_stats(c, &s)) return r;
        out->kernel_launches += s.kernel_launches; out->rays_traced += s.rays_traced; out->paths += s.paths;
        out->bounce_iterations += s.bounce_iterations; out->tail_paths += s.tail_paths;
    }
    return 0;
}

}  // extern "C"
