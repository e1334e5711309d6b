// ref_trace.cu -- TEST INFRASTRUCTURE ONLY (oracle/; never on the product path).
//
// ID-recording trace harness.  Linked against the UNMODIFIED reference objects (built by
// oracle/ref_build/build_ref.py from /root/reference), it
//   1. builds the reference's own TriangleMesh -> SimplePrimitive -> HLBVH over a raw triangle
//      soup (identity transform, so vertices are bit-identical to ours),
//   2. traces ray files through the reference's traversal order (a restatement of the 57-line
//      loop of HLBVH::intersect / fast_intersect, accelerator/hlbvh.cu:407-516, calling the
//      reference's own Bounds3f::fast_intersect and Primitive::intersect) and records the
//      primitive index of the winning candidate, which the reference API does not return
//      (ShapeIntersection = interaction + t_hit, base/interaction.h:170-176),
//   3. times the trace kernel with CUDA events (the "reference arm" of bench.py).
//
// Usage: ref_trace <soup.bin> [--warmup W] [--steps K] {closest|any}:<rays.bin>:<out.bin> ...
//   soup.bin : int64 nv, int64 nt, float P[nv*3], int32 idx[nt*3]
//   rays.bin : int64 n, float o[n*3], float d[n*3], float tmax[n]
//   out.bin  : int32 prim[n] (-1 = miss; any-hit: 1/0), float t[n]
// Prints one JSON line with build seconds and per-ray-set milliseconds.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <pbrt/base/interaction.h>
#include <pbrt/base/ray.h>
#include <pbrt/base/shape.h>
#include <pbrt/euclidean_space/bounds3.h>
#include <pbrt/gpu/gpu_memory_allocator.h>
#include <pbrt/util/stack.h>

// the three arrays the loop needs are private members of HLBVH (accelerator/hlbvh.h:113-117)
#define private public
#include <pbrt/accelerator/hlbvh.h>
#undef private

#include <pbrt/base/primitive.h>
#include <pbrt/shapes/triangle_mesh.h>

static constexpr int TRACE_STACK = 128; // accelerator/hlbvh.cu STACK_SIZE

// accelerator/hlbvh.cu:459-516 with the winner's primitive_idx recorded (:492-501)
__global__ void ref_closest(const HLBVH *bvh, const float *o, const float *d, const float *tmax,
                            int64_t n, int *prim_out, float *t_out) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n) {
        return;
    }
    const Ray ray(Point3f(o[3 * i], o[3 * i + 1], o[3 * i + 2]),
                  Vector3f(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
    auto best_t = tmax[i];
    int best_prim = -1;

    auto dd = ray.d;
    auto inv_dir = Vector3f(1.0 / dd.x, 1.0 / dd.y, 1.0 / dd.z);
    int dir_is_neg[3] = {int(inv_dir.x < 0.0), int(inv_dir.y < 0.0), int(inv_dir.z < 0.0)};

    Stack<int, TRACE_STACK> nodes_to_visit;
    nodes_to_visit.push(0);
    while (!nodes_to_visit.empty()) {
        auto current_node_idx = nodes_to_visit.pop();
        const auto node = bvh->build_nodes[current_node_idx];
        if (!node.bounds.fast_intersect(ray, best_t, inv_dir, dir_is_neg)) {
            continue;
        }
        if (node.is_leaf()) {
            for (int morton_idx = node.first_primitive_idx;
                 morton_idx < node.first_primitive_idx + node.num_primitives; morton_idx++) {
                const int primitive_idx = bvh->morton_primitives[morton_idx].primitive_idx;
                auto const primitive = bvh->primitives[primitive_idx];
                auto intersection = primitive->intersect(ray, best_t);
                if (!intersection) {
                    continue;
                }
                best_t = intersection->t_hit;
                best_prim = primitive_idx;
            }
            continue;
        }
        if (dir_is_neg[node.axis] > 0) {
            nodes_to_visit.push(node.left_child_idx);
            nodes_to_visit.push(node.left_child_idx + 1);
        } else {
            nodes_to_visit.push(node.left_child_idx + 1);
            nodes_to_visit.push(node.left_child_idx);
        }
    }
    prim_out[i] = best_prim;
    t_out[i] = best_prim >= 0 ? best_t : INFINITY;
}

// accelerator/hlbvh.cu:407-456 through the reference's own entry point
__global__ void ref_any(const HLBVH *bvh, const float *o, const float *d, const float *tmax,
                        int64_t n, int *prim_out, float *t_out) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n) {
        return;
    }
    const Ray ray(Point3f(o[3 * i], o[3 * i + 1], o[3 * i + 2]),
                  Vector3f(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
    prim_out[i] = bvh->fast_intersect(ray, tmax[i]) ? 1 : 0;
    t_out[i] = 0.0f;
}

static std::vector<char> slurp(const std::string &path) {
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) {
        fprintf(stderr, "ref_trace: cannot open %s\n", path.c_str());
        exit(2);
    }
    fseek(fp, 0, SEEK_END);
    const long sz = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    std::vector<char> buf(sz);
    if (fread(buf.data(), 1, sz, fp) != size_t(sz)) {
        fprintf(stderr, "ref_trace: short read on %s\n", path.c_str());
        exit(2);
    }
    fclose(fp);
    return buf;
}

int main(int argc, char **argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: ref_trace soup.bin [--warmup W] [--steps K] mode:rays:out ...\n");
        return 2;
    }
    size_t stack_size;
    cudaDeviceGetLimit(&stack_size, cudaLimitStackSize);
    CHECK_CUDA_ERROR(cudaDeviceSetLimit(cudaLimitStackSize, std::max(stack_size, size_t(64 * 1024))));

    int warmup = 0, steps = 1;
    std::vector<std::string> jobs;
    for (int a = 2; a < argc; a++) {
        if (!strcmp(argv[a], "--warmup") && a + 1 < argc) {
            warmup = atoi(argv[++a]);
        } else if (!strcmp(argv[a], "--steps") && a + 1 < argc) {
            steps = atoi(argv[++a]);
        } else {
            jobs.push_back(argv[a]);
        }
    }

    const auto soup = slurp(argv[1]);
    int64_t nv, nt;
    memcpy(&nv, soup.data(), 8);
    memcpy(&nt, soup.data() + 8, 8);
    const float *P = reinterpret_cast<const float *>(soup.data() + 16);
    const int *idx = reinterpret_cast<const int *>(soup.data() + 16 + nv * 12);
    std::vector<Point3f> points(nv);
    for (int64_t i = 0; i < nv; i++) {
        points[i] = Point3f(P[3 * i], P[3 * i + 1], P[3 * i + 2]);
    }
    std::vector<int> indices(idx, idx + nt * 3);

    GPUMemoryAllocator allocator;
    const auto t0 = std::chrono::steady_clock::now();
    // shapes/triangle_mesh.cu:88-113, base/primitive.cu:84-102 (material = nullptr: ids only)
    const auto [shapes, num] = TriangleMesh::build_triangles(Transform::identity(), false, indices,
                                                             points, {}, {}, allocator);
    const Primitive *prims = Primitive::create_simple_primitives(shapes, nullptr, num, allocator);
    std::vector<const Primitive *> gpu_primitives(num);
    for (int i = 0; i < num; i++) {
        gpu_primitives[i] = &prims[i];
    }
    const auto t1 = std::chrono::steady_clock::now();
    const HLBVH *bvh = allocator.create<HLBVH>(gpu_primitives, "for ref_trace", allocator);
    const auto t2 = std::chrono::steady_clock::now();
    const double load_s = std::chrono::duration<double>(t1 - t0).count();
    const double build_s = std::chrono::duration<double>(t2 - t1).count();

    std::string json = "{\"impl\":\"reference\",\"triangles\":" + std::to_string(nt) +
                       ",\"load_s\":" + std::to_string(load_s) +
                       ",\"bvh_build_s\":" + std::to_string(build_s) + ",\"sets\":[";
    bool first = true;
    for (const auto &job : jobs) {
        const auto c1 = job.find(':');
        const auto c2 = job.find(':', c1 + 1);
        const std::string mode = job.substr(0, c1), rays_path = job.substr(c1 + 1, c2 - c1 - 1),
                          out_path = job.substr(c2 + 1);
        const auto rays = slurp(rays_path);
        int64_t n;
        memcpy(&n, rays.data(), 8);
        float *d_o, *d_d, *d_tmax, *d_t;
        int *d_prim;
        CHECK_CUDA_ERROR(cudaMalloc(&d_o, n * 12));
        CHECK_CUDA_ERROR(cudaMalloc(&d_d, n * 12));
        CHECK_CUDA_ERROR(cudaMalloc(&d_tmax, n * 4));
        CHECK_CUDA_ERROR(cudaMalloc(&d_t, n * 4));
        CHECK_CUDA_ERROR(cudaMalloc(&d_prim, n * 4));
        CHECK_CUDA_ERROR(cudaMemcpy(d_o, rays.data() + 8, n * 12, cudaMemcpyHostToDevice));
        CHECK_CUDA_ERROR(cudaMemcpy(d_d, rays.data() + 8 + n * 12, n * 12, cudaMemcpyHostToDevice));
        CHECK_CUDA_ERROR(cudaMemcpy(d_tmax, rays.data() + 8 + n * 24, n * 4, cudaMemcpyHostToDevice));

        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        const int threads = 256; // integrators/wavefront_path.cu:645-646
        const int blocks = int((n + threads - 1) / threads);
        float ms_total = 0;
        for (int it = 0; it < warmup + steps; it++) {
            cudaEventRecord(e0);
            if (mode == "any") {
                ref_any<<<blocks, threads>>>(bvh, d_o, d_d, d_tmax, n, d_prim, d_t);
            } else {
                ref_closest<<<blocks, threads>>>(bvh, d_o, d_d, d_tmax, n, d_prim, d_t);
            }
            cudaEventRecord(e1);
            CHECK_CUDA_ERROR(cudaGetLastError());
            CHECK_CUDA_ERROR(cudaDeviceSynchronize());
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (it >= warmup) {
                ms_total += ms;
            }
        }
        std::vector<int> prim(n);
        std::vector<float> t(n);
        CHECK_CUDA_ERROR(cudaMemcpy(prim.data(), d_prim, n * 4, cudaMemcpyDeviceToHost));
        CHECK_CUDA_ERROR(cudaMemcpy(t.data(), d_t, n * 4, cudaMemcpyDeviceToHost));
        if (out_path != "-") {
            FILE *fp = fopen(out_path.c_str(), "wb");
            fwrite(prim.data(), 4, n, fp);
            fwrite(t.data(), 4, n, fp);
            fclose(fp);
        }
        json += std::string(first ? "" : ",") + "{\"mode\":\"" + mode + "\",\"rays\":" +
                std::to_string(n) + ",\"ms\":" + std::to_string(ms_total / steps) + "}";
        first = false;
        cudaFree(d_o);
        cudaFree(d_d);
        cudaFree(d_tmax);
        cudaFree(d_t);
        cudaFree(d_prim);
    }
    json += "]}";
    printf("%s\n", json.c_str());
    fflush(stdout);
    return 0;
}
