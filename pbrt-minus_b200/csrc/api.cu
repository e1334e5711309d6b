// api.cu -- the C-ABI of libb200pt.so (include/b200pt.h): scene upload, accelerator build,
// batched traversal, device buffer helpers.  Errors come back as status codes with a
// thread-local message; nothing here exits, throws across the boundary or runs on the CPU.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cmath>
#include <cstring>

#include "common.cuh"
#include "triangle.cuh"

namespace b200pt {

static thread_local char g_error[1024] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? B200PT_ERR_NO_DEVICE : B200PT_ERR_CUDA;
}

template <typename T>
static int upload(T **dst, const T *src, size_t count, cudaStream_t st, int64_t *bytes) {
    *dst = nullptr;
    if (!src || count == 0) return B200PT_OK;
    CU_TRY(cudaMalloc((void **)dst, count * sizeof(T)));
    CU_TRY(cudaMemcpyAsync(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
    *bytes += int64_t(count * sizeof(T));
    return B200PT_OK;
}

__global__ void k_probe_triangle(const float *__restrict__ o, const float *__restrict__ d, const float *__restrict__ tmax,
                                 const float *__restrict__ tri, long long n, int32_t *__restrict__ hit,
                                 float *__restrict__ bt) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *v = tri + 9 * i;
    TriHit th = {0.f, 0.f, 0.f, 0.f};
    const bool h = intersect_triangle(o[3 * i], o[3 * i + 1], o[3 * i + 2], d[3 * i], d[3 * i + 1], d[3 * i + 2], tmax[i], v[0],
                                      v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8], th);
    hit[i] = h ? 1 : 0;
    bt[4 * i] = h ? th.b0 : 0.f;
    bt[4 * i + 1] = h ? th.b1 : 0.f;
    bt[4 * i + 2] = h ? th.b2 : 0.f;
    bt[4 * i + 3] = h ? th.t : 0.f;
}

struct TempDev {
    std::vector<void *> ptrs;
    ~TempDev() {
        for (void *p : ptrs) cudaFree(p);
    }
    template <typename T>
    cudaError_t alloc(T **out, size_t count) {
        void *p = nullptr;
        const cudaError_t e = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(p);
        *out = (T *)p;
        return e;
    }
};

int ensure_scratch(b200pt_scene *s, size_t bytes) {
    if (bytes <= s->scratch_bytes) return B200PT_OK;
    if (s->d_scratch) {
        CU_TRY(cudaStreamSynchronize(s->stream));
        CU_TRY(cudaFree(s->d_scratch));
        s->d_scratch = nullptr;
        s->scratch_bytes = 0;
    }
    const size_t want = bytes + bytes / 4;
    CU_TRY(cudaMalloc(&s->d_scratch, want));
    s->scratch_bytes = want;
    return B200PT_OK;
}

static int select_device(int device) {
    int count = 0;
    const cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        set_error("no CUDA device available (%s); libb200pt has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return B200PT_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= count) {
        set_error("device %d out of range (0..%d)", device, count - 1);
        return B200PT_ERR_INVALID;
    }
    CU_TRY(cudaSetDevice(device));
    return B200PT_OK;
}

} // namespace b200pt

using namespace b200pt;

B200PT_API int b200pt_abi_version(void) { return B200PT_ABI_VERSION; }
B200PT_API const char *b200pt_last_error(void) { return g_error; }
B200PT_API int b200pt_device_count(void) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return 0;
    return count;
}

int b200pt_upload_render_data(b200pt_scene *s, const b200pt_scene_desc *desc); // render.cu

// Writes a scene description to `dir` as raw arrays, one file per field (tools/desc_diff.py reads them): the way two front-ends
// -- the Python one and the reference's own through the bridge -- are compared field by field, bit by bit.  No CUDA call.
B200PT_API int b200pt_debug_dump_desc(const b200pt_scene_desc *d, const char *dir) {
    if (!d || !dir) return B200PT_ERR_INVALID;
    bool ok = true;
    auto put = [&](const char *name, const void *p, size_t bytes) {
        const std::string path = std::string(dir) + "/" + name;
        FILE *f = fopen(path.c_str(), "wb");
        if (!f) {
            ok = false;
            return;
        }
        if (p && bytes) ok = fwrite(p, 1, bytes, f) == bytes && ok;
        fclose(f);
    };
    const size_t nv = (size_t)d->n_vertices, nt = (size_t)d->n_triangles;
    put("P.f32", d->P, nv * 12);
    put("N.f32", d->N, d->N ? nv * 12 : 0);
    put("uv.f32", d->uv, d->uv ? nv * 8 : 0);
    put("indices.i32", d->indices, nt * 12);
    put("tri_material.i32", d->tri_material, d->tri_material ? nt * 4 : 0);
    put("tri_light.i32", d->tri_light, d->tri_light ? nt * 4 : 0);
    put("tri_flip.u8", d->tri_flip, d->tri_flip ? nt : 0);
    put("tri_medium.u32", d->tri_medium, d->tri_medium ? nt * 4 : 0);
    put("spheres.bin", d->spheres, (size_t)d->n_spheres * sizeof(b200pt_sphere));
    put("materials.bin", d->materials, (size_t)d->n_materials * sizeof(b200pt_material));
    put("lights.bin", d->lights, (size_t)d->n_lights * sizeof(b200pt_light));
    put("textures.bin", d->textures, (size_t)d->n_textures * sizeof(b200pt_texture));
    put("media.bin", d->media, (size_t)d->n_media * sizeof(b200pt_medium));
    put("objects.bin", d->objects, (size_t)d->n_objects * sizeof(b200pt_object));
    put("instances.bin", d->instances, (size_t)d->n_instances * sizeof(b200pt_instance));
    if (d->camera) put("camera.bin", d->camera, sizeof(b200pt_camera));
    { // spectra: {type, c[4], n} records and one pool of lambda[n] then value[n] per spectrum that has them
        std::vector<int32_t> rec;
        std::vector<float> pool;
        for (int i = 0; i < d->n_spectra; i++) {
            const b200pt_spectrum &sp = d->spectra[i];
            int32_t w[6];
            w[0] = sp.type;
            memcpy(&w[1], sp.c, 16);
            w[5] = sp.n;
            rec.insert(rec.end(), w, w + 6);
            if (sp.type == B200PT_SPEC_PIECEWISE && sp.lambda && sp.value) {
                pool.insert(pool.end(), sp.lambda, sp.lambda + sp.n);
                pool.insert(pool.end(), sp.value, sp.value + sp.n);
            } else if (sp.type == B200PT_SPEC_DENSE && sp.value) {
                pool.insert(pool.end(), sp.value, sp.value + sp.n);
            }
        }
        put("spectra.i32", rec.data(), rec.size() * 4);
        put("spectra_pool.f32", pool.data(), pool.size() * 4);
    }
    if (d->filter) {
        const float hdr[4] = {d->filter->radius_x, d->filter->radius_y, (float)d->filter->nx, (float)d->filter->ny};
        put("filter_hdr.f32", hdr, sizeof(hdr));
        put("filter_f.f32", d->filter->f, (size_t)d->filter->nx * d->filter->ny * 4);
    }
    if (d->film) {
        std::vector<float> v = {(float)d->film->width, (float)d->film->height, d->film->imaging_ratio};
        v.insert(v.end(), d->film->output_rgb_from_sensor_rgb, d->film->output_rgb_from_sensor_rgb + 9);
        for (const float *t : {d->film->cie_x, d->film->cie_y, d->film->cie_z, d->film->illuminant})
            if (t) v.insert(v.end(), t, t + 471);
        put("film.f32", v.data(), v.size() * 4);
    }
    {
        std::vector<int32_t> dims;
        std::vector<float> px;
        for (int i = 0; i < d->n_images; i++) {
            dims.push_back(d->images[i].width);
            dims.push_back(d->images[i].height);
            if (d->images[i].rgb) px.insert(px.end(), d->images[i].rgb, d->images[i].rgb + (size_t)d->images[i].width * d->images[i].height * 3);
        }
        put("images_dims.i32", dims.data(), dims.size() * 4);
        put("images_rgb.f32", px.data(), px.size() * 4);
    }
    if (!ok) {
        set_error("b200pt_debug_dump_desc: cannot write to '%s'", dir);
        return B200PT_ERR_INVALID;
    }
    return B200PT_OK;
}

B200PT_API int b200pt_scene_create(const b200pt_scene_desc *desc, int device, b200pt_scene **out) {
    if (!desc || !out) {
        set_error("b200pt_scene_create: null argument");
        return B200PT_ERR_INVALID;
    }
    *out = nullptr;
    if (const char *dump = getenv("B200PT_DESC_DUMP")) b200pt_debug_dump_desc(desc, dump);
    if (desc->n_triangles < 0 || desc->n_vertices < 0 || (desc->n_triangles > 0 && (!desc->P || !desc->indices))) {
        set_error("b200pt_scene_create: inconsistent geometry description");
        return B200PT_ERR_INVALID;
    }
    int rc = select_device(device);
    if (rc) return rc;
    b200pt_scene *s = new b200pt_scene();
    s->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) s->sm_count = prop.multiProcessorCount;
    auto fail = [&](int code) {
        b200pt_scene_destroy(s);
        return code;
    };
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) {
        set_error("b200pt_scene_create: cannot create a stream");
        return fail(B200PT_ERR_CUDA);
    }
    s->n_vertices = desc->n_vertices;
    s->n_triangles = desc->n_triangles;
    // ---- instanced objects: ranges that tile the end of the triangle array
    s->n_top_triangles = desc->n_triangles;
    if (desc->n_objects < 0 || desc->n_instances < 0 || (desc->n_objects > 0 && !desc->objects) || (desc->n_instances > 0 && !desc->instances)) {
        set_error("b200pt_scene_create: objects / instances count without an array");
        return fail(B200PT_ERR_INVALID);
    }
    if (desc->n_objects > 0) {
        int64_t next = desc->objects[0].first_triangle;
        s->n_top_triangles = next;
        for (int i = 0; i < desc->n_objects; i++) {
            const b200pt_object &o = desc->objects[i];
            if (o.first_triangle != next || o.n_triangles <= 0 || o.first_triangle < 0) {
                set_error("b200pt_scene_create: object %d: ranges must be non-empty and tile the end of the triangle array in order", i);
                return fail(B200PT_ERR_INVALID);
            }
            next += o.n_triangles;
            HostObject h;
            h.first_triangle = o.first_triangle;
            h.n_triangles = o.n_triangles;
            s->objects.push_back(h);
        }
        if (next != desc->n_triangles) {
            set_error("b200pt_scene_create: object ranges end at triangle %lld of %lld", (long long)next, (long long)desc->n_triangles);
            return fail(B200PT_ERR_INVALID);
        }
        for (int64_t t = s->n_top_triangles; t < desc->n_triangles; t++)
            if ((desc->tri_light && desc->tri_light[t] != -1) || (desc->tri_medium && desc->tri_medium[t] != 0u)) {
                set_error("b200pt_scene_create: triangle %lld of an instanced object is emissive or on a medium transition", (long long)t);
                return fail(B200PT_ERR_INVALID);
            }
    }
    for (int i = 0; i < desc->n_instances; i++) {
        if (desc->instances[i].object < 0 || desc->instances[i].object >= desc->n_objects) {
            set_error("b200pt_scene_create: instance %d names object %d", i, desc->instances[i].object);
            return fail(B200PT_ERR_INVALID);
        }
        s->instances.push_back(desc->instances[i]);
    }
    // index sanity (a bad index would fault on the device)
    for (int64_t i = 0; i < desc->n_triangles * 3; i++) {
        const int32_t v = desc->indices[i];
        if (v < 0 || v >= desc->n_vertices) {
            set_error("b200pt_scene_create: vertex index %d out of range at position %lld", v, (long long)i);
            return fail(B200PT_ERR_INVALID);
        }
    }
    if ((rc = upload(&s->d_P, desc->P, size_t(desc->n_vertices) * 3, s->stream, &s->device_bytes))) return fail(rc);
    if ((rc = upload(&s->d_N, desc->N, size_t(desc->n_vertices) * 3, s->stream, &s->device_bytes))) return fail(rc);
    if ((rc = upload(&s->d_uv, desc->uv, size_t(desc->n_vertices) * 2, s->stream, &s->device_bytes))) return fail(rc);
    if ((rc = upload(&s->d_idx, desc->indices, size_t(desc->n_triangles) * 3, s->stream, &s->device_bytes))) return fail(rc);
    if ((rc = upload(&s->d_tri_material, desc->tri_material, size_t(desc->n_triangles), s->stream, &s->device_bytes))) return fail(rc);
    if ((rc = upload(&s->d_tri_light, desc->tri_light, size_t(desc->n_triangles), s->stream, &s->device_bytes))) return fail(rc);
    if ((rc = upload(&s->d_tri_flip, desc->tri_flip, size_t(desc->n_triangles), s->stream, &s->device_bytes))) return fail(rc);
    if (cudaMalloc((void **)&s->d_trace_counters, kTraceCounterWords * sizeof(unsigned long long)) != cudaSuccess) {
        set_error("b200pt_scene_create: out of device memory");
        return fail(B200PT_ERR_CUDA);
    }
    cudaMemsetAsync(s->d_trace_counters, 0, kTraceCounterWords * sizeof(unsigned long long), s->stream);
    if ((rc = b200pt_upload_render_data(s, desc))) return fail(rc);
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) {
        set_error("b200pt_scene_create: upload failed");
        return fail(B200PT_ERR_CUDA);
    }
    *out = s;
    return B200PT_OK;
}

B200PT_API void b200pt_scene_destroy(b200pt_scene *s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    free_render(s);
    void *ptrs[] = {s->d_instances, s->d_scratch, s->d_P, s->d_N, s->d_uv, s->d_idx, s->d_tri_material, s->d_tri_light, s->d_tri_flip, s->d_spheres,
                    s->d_nodes, s->d_tris, s->d_sorted_keys, s->d_sorted_prim, s->d_trace_counters, s->build_scratch};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    for (void *p : s->user_allocs) cudaFree(p);
    if (s->stream) cudaStreamDestroy(s->stream);
    if (s->stream_h2d) cudaStreamDestroy(s->stream_h2d);
    if (s->stream_d2h) cudaStreamDestroy(s->stream_d2h);
    if (s->stream2) cudaStreamDestroy(s->stream2);
    for (cudaEvent_t e : s->pipe_events) cudaEventDestroy(e);
    delete s;
}

B200PT_API int b200pt_bvh_build(b200pt_scene *s, b200pt_build_stats *stats_out) {
    if (!s) {
        set_error("b200pt_bvh_build: null scene");
        return B200PT_ERR_INVALID;
    }
    CU_TRY(cudaSetDevice(s->device));
    return build_bvh(s, stats_out);
}

namespace b200pt {
// InstRec array: object roots are known once the build has run; the instance box is the transformed object box
// (TransformedPrimitive::bounds, primitives/transformed_primitive.cu:9-12), padded like the quadric boxes
int finish_instances(b200pt_scene *s) {
    if (s->d_instances) {
        cudaFree(s->d_instances);
        s->d_instances = nullptr;
    }
    if (s->instances.empty()) return B200PT_OK;
    std::vector<InstRec> recs(s->instances.size());
    s->inst_boxes.assign(6 * s->instances.size(), 0.f);
    for (size_t i = 0; i < recs.size(); i++) {
        const b200pt_instance &in = s->instances[i];
        const HostObject &ob = s->objects[in.object];
        InstRec &r = recs[i];
        for (int k = 0; k < 12; k++) {
            r.rfi[k] = in.render_from_instance[k];
            r.ifr[k] = in.instance_from_render[k];
        }
        r.root_node = (int32_t)ob.root_node;
        r.object = in.object;
        float ext = 0.f, mag = 0.f;
        for (int a = 0; a < 3; a++) {
            float mn = in.render_from_instance[4 * a + 3], mx = mn;
            for (int k = 0; k < 3; k++) {
                const float x = in.render_from_instance[4 * a + k] * ob.box[k], y = in.render_from_instance[4 * a + k] * ob.box[3 + k];
                mn += std::min(x, y);
                mx += std::max(x, y);
            }
            r.box_lo[a] = mn;
            r.box_hi[a] = mx;
            ext = std::max(ext, mx - mn);
            mag = std::max(mag, std::max(std::fabs(mn), std::fabs(mx)));
        }
        const float pad = 1e-3f * ext + 1e-5f * mag + 1e-30f;
        for (int a = 0; a < 3; a++) {
            r.box_lo[a] -= pad;
            r.box_hi[a] += pad;
            s->inst_boxes[6 * i + a] = r.box_lo[a];
            s->inst_boxes[6 * i + 3 + a] = r.box_hi[a];
        }
    }
    CU_TRY(cudaMalloc((void **)&s->d_instances, recs.size() * sizeof(InstRec)));
    CU_TRY(cudaMemcpyAsync(s->d_instances, recs.data(), recs.size() * sizeof(InstRec), cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}
} // namespace b200pt

B200PT_API int b200pt_bvh_validate(b200pt_scene *s, int64_t out[3]) {
    if (!s || !out || !s->built) {
        set_error("b200pt_bvh_validate: build first");
        return B200PT_ERR_STATE;
    }
    CU_TRY(cudaSetDevice(s->device));
    return validate_bvh(s, out);
}

B200PT_API int b200pt_bvh_debug_sorted(b200pt_scene *s, int32_t *prim_order_out, uint64_t *keys_out) {
    if (!s || !s->built) {
        set_error("b200pt_bvh_debug_sorted: build first");
        return B200PT_ERR_STATE;
    }
    CU_TRY(cudaSetDevice(s->device));
    if (s->n_triangles == 0) return B200PT_OK;
    if (prim_order_out)
        CU_TRY(cudaMemcpyAsync(prim_order_out, s->d_sorted_prim, s->n_triangles * 4, cudaMemcpyDeviceToHost, s->stream));
    if (keys_out) CU_TRY(cudaMemcpyAsync(keys_out, s->d_sorted_keys, s->n_triangles * 8, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}

static int check_trace_status(b200pt_scene *s) {
    unsigned long long c[4];
    CU_TRY(cudaMemcpyAsync(c, s->d_trace_counters, sizeof(c), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    s->last_counts[0] = (int64_t)c[1];
    s->last_counts[1] = (int64_t)c[2];
    if (c[3]) {
        // the overflow word is sticky on the device (any launch since the last look may have set it): reading clears it
        CU_TRY(cudaMemsetAsync(s->d_trace_counters + 3, 0, sizeof(unsigned long long), s->stream));
        CU_TRY(cudaStreamSynchronize(s->stream));
        set_error("b200pt_trace: traversal stack overflow (tree deeper than %d groups)", 96);
        return B200PT_ERR_STATE;
    }
    return B200PT_OK;
}

B200PT_API int b200pt_trace_device(b200pt_scene *s, const void *rays_dev, int64_t n, int any_hit, void *hits_dev, int sync) {
    if (!s || (n > 0 && (!rays_dev || !hits_dev))) {
        set_error("b200pt_trace_device: null argument");
        return B200PT_ERR_INVALID;
    }
    CU_TRY(cudaSetDevice(s->device));
    const int rc = launch_trace(s, rays_dev, n, any_hit, hits_dev);
    if (rc) return rc;
    s->last_counts[2] = n;
    if (sync) return check_trace_status(s);
    return B200PT_OK;
}

// Several device-resident ray sets in one call: consecutive sets go to the scene's two traversal streams in turn, each with its own
// work ticket, so the CTAs of set k+1 take the SM slots the last warps of set k have left (the tail of a persistent kernel is a few
// per cent of a launch).  Ordered like one operation on the scene's stream: it starts after what the stream held and the stream
// goes on after the last set.  Counting runs and single sets keep one stream.
B200PT_API int b200pt_trace_device_batch(b200pt_scene *s, const b200pt_device_ray_set *sets, int n_sets, int sync) {
    if (!s || n_sets < 0 || (n_sets > 0 && !sets)) {
        set_error("b200pt_trace_device_batch: null argument");
        return B200PT_ERR_INVALID;
    }
    int64_t total = 0;
    for (int i = 0; i < n_sets; i++) {
        if (sets[i].n < 0 || (sets[i].n > 0 && (!sets[i].rays_dev || !sets[i].hits_dev))) {
            set_error("b200pt_trace_device_batch: bad ray set %d", i);
            return B200PT_ERR_INVALID;
        }
        total += sets[i].n;
    }
    CU_TRY(cudaSetDevice(s->device));
    const bool two = n_sets > 1 && !s->counting;
    if (two) {
        if (!s->stream2) CU_TRY(cudaStreamCreateWithFlags(&s->stream2, cudaStreamNonBlocking));
        while (s->pipe_events.size() < 2) {
            cudaEvent_t e;
            CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            s->pipe_events.push_back(e);
        }
        CU_TRY(cudaEventRecord(s->pipe_events[0], s->stream));
        CU_TRY(cudaStreamWaitEvent(s->stream2, s->pipe_events[0], 0));
    }
    for (int i = 0; i < n_sets; i++) {
        cudaStream_t cs = (two && (i & 1)) ? s->stream2 : s->stream;
        const int rc = launch_trace(s, sets[i].rays_dev, sets[i].n, sets[i].any_hit, sets[i].hits_dev, i > 0, nullptr, two ? i : -1, cs);
        if (rc) return rc;
    }
    if (two) {
        CU_TRY(cudaEventRecord(s->pipe_events[1], s->stream2));
        CU_TRY(cudaStreamWaitEvent(s->stream, s->pipe_events[1], 0));
    }
    s->last_counts[2] = total;
    if (sync) return check_trace_status(s);
    return B200PT_OK;
}

// Host-buffer traversal.  A three-stage pipeline over chunks of ALL the sets of the call: the copy-in of chunk k+1 (H2D
// engine), the traversal of chunk k and the copy-out of chunk k-1 (D2H engine) run at the same time, so with pinned host
// buffers a call costs about max(copy-in, traversal, copy-out) + one chunk of fill and drain, instead of their sum per set.
// The trace kernel reads the caller's array layout and writes the caller's output layout directly (TraceHostIO): there is
// no repacking pass.  Consecutive chunks alternate between two traversal streams, each with its own work ticket: while the
// last warps of chunk k finish their rays, the CTAs of chunk k+1 already occupy the SM slots the others have left, so the
// per-launch tail of a persistent kernel is hidden instead of paid once per chunk.  Pageable buffers still work (the copies then block the host).  One synchronisation, at the end.
B200PT_API int b200pt_trace_batch(b200pt_scene *s, const b200pt_ray_set *sets, int n_sets) {
    if (!s || n_sets < 0 || (n_sets > 0 && !sets)) {
        set_error("b200pt_trace_batch: null argument");
        return B200PT_ERR_INVALID;
    }
    int64_t total = 0;
    for (int i = 0; i < n_sets; i++) {
        const b200pt_ray_set &rs = sets[i];
        if (rs.n < 0 || (rs.n > 0 && (!rs.o_xyz || !rs.d_xyz || !rs.tmax))) {
            set_error("b200pt_trace_batch: ray set %d: null argument", i);
            return B200PT_ERR_INVALID;
        }
        total += (rs.n + 3) & ~int64_t(3);
    }
    if (!s->built) {
        set_error("b200pt_trace: call b200pt_bvh_build first");
        return B200PT_ERR_STATE;
    }
    if (!s->instances.empty()) {
        set_error("b200pt_trace: scenes with object instances are traced by the wavefront integrator only");
        return B200PT_ERR_UNSUPPORTED;
    }
    if (total == 0) return B200PT_OK;
    CU_TRY(cudaSetDevice(s->device));
    // scratch layout (every array 16-byte aligned, sets back to back): o | d | tmax | prim | t | b12
    const size_t T = (size_t)total;
    int rc0 = ensure_scratch(s, T * (12 + 12 + 4 + 4 + 4 + 8));
    if (rc0) return rc0;
    char *base = (char *)s->d_scratch;
    float *d_o = (float *)base;
    float *d_d = (float *)(base + T * 12);
    float *d_tmax = (float *)(base + T * 24);
    int32_t *d_prim = (int32_t *)(base + T * 28);
    float *d_t = (float *)(base + T * 32);
    float *d_b = (float *)(base + T * 36);
    static const int64_t chunk_env = getenv("B200PT_TRACE_CHUNK") ? atoll(getenv("B200PT_TRACE_CHUNK")) : 0;
    const int64_t chunk = chunk_env > 0 ? chunk_env : (int64_t)1 << 19;
    int n_chunks = 0;
    for (int i = 0; i < n_sets; i++) n_chunks += (int)((sets[i].n + chunk - 1) / chunk);
    if (!s->stream_h2d) CU_TRY(cudaStreamCreateWithFlags(&s->stream_h2d, cudaStreamNonBlocking));
    if (!s->stream_d2h) CU_TRY(cudaStreamCreateWithFlags(&s->stream_d2h, cudaStreamNonBlocking));
    if (!s->stream2) CU_TRY(cudaStreamCreateWithFlags(&s->stream2, cudaStreamNonBlocking));
    static const bool one_stream = getenv("B200PT_TRACE_ONE_STREAM") != nullptr;
    const bool two = !one_stream && !s->counting; // the visit counters are per launch: counting runs keep one stream
    while ((int)s->pipe_events.size() < 2 * n_chunks + 1) {
        cudaEvent_t e;
        CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        s->pipe_events.push_back(e);
    }
    // the scratch may still be read by the copy-out of the previous call's last chunk only if that call failed half way;
    // a finished call has synchronised both streams.  The copy-in stream must not run ahead of a scratch re-allocation:
    // ensure_scratch synchronises s->stream, and the other two streams are idle between calls.
    int c = 0;
    int64_t off = 0; // ray offset of the set inside the scratch arrays
    for (int i = 0; i < n_sets; i++) {
        const b200pt_ray_set &rs = sets[i];
        for (int64_t c0 = 0; c0 < rs.n; c0 += chunk, c++) {
            const int64_t nc = std::min(chunk, rs.n - c0), g0 = off + c0;
            cudaEvent_t ev_in = s->pipe_events[2 * c], ev_done = s->pipe_events[2 * c + 1];
            CU_TRY(cudaMemcpyAsync(d_o + 3 * g0, rs.o_xyz + 3 * c0, nc * 12, cudaMemcpyHostToDevice, s->stream_h2d));
            CU_TRY(cudaMemcpyAsync(d_d + 3 * g0, rs.d_xyz + 3 * c0, nc * 12, cudaMemcpyHostToDevice, s->stream_h2d));
            CU_TRY(cudaMemcpyAsync(d_tmax + g0, rs.tmax + c0, nc * 4, cudaMemcpyHostToDevice, s->stream_h2d));
            CU_TRY(cudaEventRecord(ev_in, s->stream_h2d));
            cudaStream_t cs = (two && (c & 1)) ? s->stream2 : s->stream;
            CU_TRY(cudaStreamWaitEvent(cs, ev_in, 0));
            const TraceHostIO io = {d_o + 3 * g0, d_d + 3 * g0, d_tmax + g0, rs.prim_out ? d_prim + g0 : nullptr,
                                    rs.t_out ? d_t + g0 : nullptr, rs.b1b2_out ? d_b + 2 * g0 : nullptr};
            int rc = launch_trace(s, nullptr, nc, rs.any_hit, nullptr, c > 0, &io, two ? c : -1, cs);
            if (rc) return rc;
            CU_TRY(cudaEventRecord(ev_done, cs));
            CU_TRY(cudaStreamWaitEvent(s->stream_d2h, ev_done, 0));
            if (rs.prim_out) CU_TRY(cudaMemcpyAsync(rs.prim_out + c0, d_prim + g0, nc * 4, cudaMemcpyDeviceToHost, s->stream_d2h));
            if (rs.t_out) CU_TRY(cudaMemcpyAsync(rs.t_out + c0, d_t + g0, nc * 4, cudaMemcpyDeviceToHost, s->stream_d2h));
            if (rs.b1b2_out) CU_TRY(cudaMemcpyAsync(rs.b1b2_out + 2 * c0, d_b + 2 * g0, nc * 8, cudaMemcpyDeviceToHost, s->stream_d2h));
        }
        off += (rs.n + 3) & ~int64_t(3);
    }
    s->last_counts[2] = total;
    CU_TRY(cudaStreamSynchronize(s->stream2));
    const int rc = check_trace_status(s); // synchronises the first traversal stream
    CU_TRY(cudaStreamSynchronize(s->stream_d2h));
    return rc;
}

B200PT_API int b200pt_trace(b200pt_scene *s, const float *o_xyz, const float *d_xyz, const float *tmax, int64_t n,
                            int any_hit, int32_t *prim_out, float *t_out, float *b1b2_out) {
    if (!s || n < 0 || (n > 0 && (!o_xyz || !d_xyz || !tmax))) {
        set_error("b200pt_trace: null argument");
        return B200PT_ERR_INVALID;
    }
    const b200pt_ray_set one = {o_xyz, d_xyz, tmax, n, any_hit, 0, prim_out, t_out, b1b2_out};
    return b200pt_trace_batch(s, &one, 1);
}

B200PT_API int b200pt_trace_set_counting(b200pt_scene *s, int enable) {
    if (!s) return B200PT_ERR_INVALID;
    s->counting = enable != 0;
    return B200PT_OK;
}

B200PT_API int b200pt_trace_tune(b200pt_scene *s, int key, int value) {
    if (!s || value < 0) return B200PT_ERR_INVALID;
    switch (key) {
    case 0: s->tune_refill = value < 1 ? 1 : value; break;
    case 1: s->tune_tri_steps = value < 1 ? 1 : (value > 3 ? 3 : value); break;
    case 2: s->tune_ctas_per_sm = value; break;
    case 3: s->tune_wf_refill = value; break;
    case 4: s->tune_wf_tri_steps = value > 3 ? 3 : value; break;
    case 5: s->tune_shade_occ = value; break;
    case 6: s->tune_stash = value != 0; break;
    case 7: // MB of build scratch a scene keeps between builds; lowering it below what is held releases the scratch now
        s->tune_build_scratch_mb = value;
        if (s->build_scratch && s->build_scratch_bytes > (size_t)value << 20) {
            CU_TRY(cudaSetDevice(s->device));
            CU_TRY(cudaStreamSynchronize(s->stream));
            cudaFree(s->build_scratch);
            s->build_scratch = nullptr;
            s->build_scratch_bytes = 0;
        }
        break;
    default: set_error("b200pt_trace_tune: unknown key %d", key); return B200PT_ERR_INVALID;
    }
    return B200PT_OK;
}

B200PT_API int b200pt_trace_counters(b200pt_scene *s, int64_t out[3]) {
    if (!s || !out) return B200PT_ERR_INVALID;
    CU_TRY(cudaSetDevice(s->device));
    const int rc = check_trace_status(s);
    out[0] = s->last_counts[0];
    out[1] = s->last_counts[1];
    out[2] = s->last_counts[2];
    return rc;
}

B200PT_API void *b200pt_scene_stream(b200pt_scene *s) { return s ? (void *)s->stream : nullptr; }

B200PT_API int b200pt_scene_sync(b200pt_scene *s) {
    if (!s) return B200PT_ERR_INVALID;
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}

B200PT_API int b200pt_device_alloc(b200pt_scene *s, size_t bytes, void **out_dev) {
    if (!s || !out_dev) return B200PT_ERR_INVALID;
    CU_TRY(cudaSetDevice(s->device));
    void *p = nullptr;
    CU_TRY(cudaMalloc(&p, bytes ? bytes : 1));
    s->user_allocs.push_back(p);
    *out_dev = p;
    return B200PT_OK;
}

B200PT_API int b200pt_device_free(b200pt_scene *s, void *ptr_dev) {
    if (!s) return B200PT_ERR_INVALID;
    CU_TRY(cudaSetDevice(s->device));
    for (size_t i = 0; i < s->user_allocs.size(); i++)
        if (s->user_allocs[i] == ptr_dev) {
            s->user_allocs.erase(s->user_allocs.begin() + i);
            CU_TRY(cudaFree(ptr_dev));
            return B200PT_OK;
        }
    set_error("b200pt_device_free: pointer not owned by this scene");
    return B200PT_ERR_INVALID;
}

B200PT_API int b200pt_copy_to_device(b200pt_scene *s, void *dst_dev, const void *src, size_t bytes) {
    if (!s) return B200PT_ERR_INVALID;
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaMemcpyAsync(dst_dev, src, bytes, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}

B200PT_API int b200pt_copy_to_host(b200pt_scene *s, void *dst, const void *src_dev, size_t bytes) {
    if (!s) return B200PT_ERR_INVALID;
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaMemcpyAsync(dst, src_dev, bytes, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}

B200PT_API int b200pt_probe_triangle(b200pt_scene *s, const float *o, const float *d, const float *tmax, const float *tri,
                                     int64_t n, int32_t *hit_out, float *bt_out) {
    if (!s || n < 0 || (n > 0 && (!o || !d || !tmax || !tri || !hit_out || !bt_out))) {
        set_error("b200pt_probe_triangle: null argument");
        return B200PT_ERR_INVALID;
    }
    if (n == 0) return B200PT_OK;
    CU_TRY(cudaSetDevice(s->device));
    TempDev tmp;
    float *d_o, *d_d, *d_tmax, *d_tri, *d_bt;
    int32_t *d_hit;
    CU_TRY(tmp.alloc(&d_o, n * 3));
    CU_TRY(tmp.alloc(&d_d, n * 3));
    CU_TRY(tmp.alloc(&d_tmax, n));
    CU_TRY(tmp.alloc(&d_tri, n * 9));
    CU_TRY(tmp.alloc(&d_bt, n * 4));
    CU_TRY(tmp.alloc(&d_hit, n));
    CU_TRY(cudaMemcpyAsync(d_o, o, n * 12, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(d_d, d, n * 12, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(d_tmax, tmax, n * 4, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(d_tri, tri, n * 36, cudaMemcpyHostToDevice, s->stream));
    k_probe_triangle<<<unsigned((n + 255) / 256), 256, 0, s->stream>>>(d_o, d_d, d_tmax, d_tri, n, d_hit, d_bt);
    CU_TRY(cudaMemcpyAsync(hit_out, d_hit, n * 4, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaMemcpyAsync(bt_out, d_bt, n * 16, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    return B200PT_OK;
}
