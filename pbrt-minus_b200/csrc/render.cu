// render.cu -- wavefront path integrator entry points (film, render, stage probes).
#include "common.cuh"
#include "sampler.cuh"

struct b200pt_film {
    b200pt_scene *scene = nullptr;
    int32_t width = 0, height = 0;
    float4 *d_accum = nullptr; // {sum_x, sum_y, sum_z, sum_w}
};

namespace b200pt {

struct RenderData {
    int dummy = 0;
};

void free_render(b200pt_scene *s) {
    delete s->render;
    s->render = nullptr;
}

__global__ void k_probe_sampler(int kind, int spp, int sqrt_spp, const int32_t *__restrict__ pixel,
                                const int32_t *__restrict__ sample, int npairs, const int32_t *__restrict__ pattern, int ndraw,
                                int per, float *__restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= npairs) return;
    Sampler s;
    s.start(kind, spp, sqrt_spp, pixel[k], sample[k], 0);
    float *o = out + (size_t)k * per;
    for (int i = 0; i < ndraw; i++) {
        if (pattern[i] == 1) {
            *o++ = s.get_1d();
        } else {
            const float2 v = s.get_2d();
            *o++ = v.x;
            *o++ = v.y;
        }
    }
}

} // namespace b200pt

using namespace b200pt;

int b200pt_upload_render_data(b200pt_scene *s, const b200pt_scene_desc *desc) {
    (void)s;
    (void)desc;
    return B200PT_OK;
}

B200PT_API int b200pt_probe_sampler(b200pt_scene *s, int kind, int spp, const int32_t *pixel, const int32_t *sample,
                                    int npairs, const int32_t *pattern, int ndraw, float *out) {
    if (!s || npairs < 0 || ndraw < 0 || (npairs > 0 && (!pixel || !sample || !pattern || !out))) {
        set_error("b200pt_probe_sampler: null argument");
        return B200PT_ERR_INVALID;
    }
    if (npairs == 0 || ndraw == 0) return B200PT_OK;
    CU_TRY(cudaSetDevice(s->device));
    int per = 0;
    for (int i = 0; i < ndraw; i++) per += pattern[i] == 1 ? 1 : 2;
    int32_t *d_pixel = nullptr, *d_sample = nullptr, *d_pattern = nullptr;
    float *d_out = nullptr;
    auto cleanup = [&]() {
        cudaFree(d_pixel);
        cudaFree(d_sample);
        cudaFree(d_pattern);
        cudaFree(d_out);
    };
    cudaError_t e = cudaMalloc(&d_pixel, npairs * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_sample, npairs * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_pattern, ndraw * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, (size_t)npairs * per * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pixel, pixel, npairs * 4, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_sample, sample, npairs * 4, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_pattern, pattern, ndraw * 4, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) {
        const int sqrt_spp = (int)sqrt((double)spp);
        k_probe_sampler<<<(npairs + 127) / 128, 128, 0, s->stream>>>(kind, spp, sqrt_spp, d_pixel, d_sample, npairs, d_pattern,
                                                                    ndraw, per, d_out);
        e = cudaMemcpyAsync(out, d_out, (size_t)npairs * per * 4, cudaMemcpyDeviceToHost, s->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    cleanup();
    if (e != cudaSuccess) return cuda_fail(e, "b200pt_probe_sampler", __FILE__, __LINE__);
    return B200PT_OK;
}

B200PT_API int b200pt_film_create(b200pt_scene *s, b200pt_film **out) {
    (void)s;
    (void)out;
    set_error("b200pt_film_create: scene has no film description");
    return B200PT_ERR_UNSUPPORTED;
}
B200PT_API void b200pt_film_destroy(b200pt_film *film) {
    if (!film) return;
    if (film->d_accum) cudaFree(film->d_accum);
    delete film;
}
B200PT_API int b200pt_film_clear(b200pt_film *film) { (void)film; return B200PT_ERR_UNSUPPORTED; }
B200PT_API int b200pt_render(b200pt_scene *s, b200pt_film *film, const b200pt_render_opts *opts, b200pt_render_stats *st) {
    (void)s; (void)film; (void)opts; (void)st;
    set_error("b200pt_render: not built in this revision");
    return B200PT_ERR_UNSUPPORTED;
}
B200PT_API int b200pt_film_reduce(b200pt_film *film, void *comm, int root, void *stream) {
    (void)film; (void)comm; (void)root; (void)stream;
    return B200PT_ERR_UNSUPPORTED;
}
B200PT_API void *b200pt_film_device_ptr(b200pt_film *film) { return film ? (void *)film->d_accum : nullptr; }
B200PT_API int b200pt_film_dims(b200pt_film *film, int32_t *w, int32_t *h) {
    if (!film) return B200PT_ERR_INVALID;
    if (w) *w = film->width;
    if (h) *h = film->height;
    return B200PT_OK;
}
B200PT_API int b200pt_film_read(b200pt_film *film, float *rgb_out) { (void)film; (void)rgb_out; return B200PT_ERR_UNSUPPORTED; }
B200PT_API int b200pt_film_read_raw(b200pt_film *film, float *xyzw_out) { (void)film; (void)xyzw_out; return B200PT_ERR_UNSUPPORTED; }
