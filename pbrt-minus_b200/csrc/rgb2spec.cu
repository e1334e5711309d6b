// rgb2spec.cu -- the RGB -> sigmoid-polynomial coefficient table, fitted ON THE DEVICE at scene creation.
//
// Replaces RGBtoSpectrumTable::RGBtoSpectrumTable (spectrum_util/rgb_to_spectrum_data.cu:214-386: a host thread pool runs a
// Gauss-Newton fit per table cell at every start of the program).  Needed on the device only when a scene has image
// textures (an "imagemap" texel becomes an RGBAlbedoSpectrum at shading time, textures/spectrum_image_texture.cu:17-32); the
// RGB parameters of materials and lights are resolved by the host front-end as before.
//
// One thread fits one (max-component l, y = j/63, x = i/63) column of the 3 x 64 x 64 x 64 table: the 64 brightness levels are
// chained outwards from k = 64/5 with warm starts exactly as `compute` does (:304-358), each a <= 15-step Gauss-Newton on the
// CIELAB residual of the 283-sample Simpson integration against the CIE * D65 curves (`eval_residual`, `eval_jacobian` with
// central differences of 1e-4, :214-262), all in fp64.  The 3 x 3 step is solved by elimination with partial pivoting.
#include <cmath>
#include <vector>

#include "common.cuh"

namespace b200pt {

constexpr int RGB2SPEC_RES = 64;
constexpr int RGB2SPEC_FINE = 283; // (95 - 1) * 3 + 1

struct FitTables {
    const double *rgb_tbl; // [3][283]  xyz_to_srgb * (cie xyz * d65 * simpson weight)
    const double *white;   // [3]       xyz whitepoint of the same integration
    const double *lam01;   // [283]     (lambda - 360) / 470
};

__device__ __forceinline__ double lab_f(double t) {
    const double d = 6.0 / 29.0;
    return t > d * d * d ? cbrt(t) : t / (d * d * 3.0) + 4.0 / 29.0;
}

__device__ void cie_lab(double p[3], const double *white) {
    const double X = 0.412453 * p[0] + 0.357580 * p[1] + 0.180423 * p[2];
    const double Y = 0.212671 * p[0] + 0.715160 * p[1] + 0.072169 * p[2];
    const double Z = 0.019334 * p[0] + 0.119193 * p[1] + 0.950227 * p[2];
    const double fx = lab_f(X / white[0]), fy = lab_f(Y / white[1]), fz = lab_f(Z / white[2]);
    p[0] = 116.0 * fy - 16.0;
    p[1] = 500.0 * (fx - fy);
    p[2] = 200.0 * (fy - fz);
}

__device__ void eval_residual(const FitTables &T, const double c[3], const double rgb[3], double r[3]) {
    double out[3] = {0.0, 0.0, 0.0};
    for (int i = 0; i < RGB2SPEC_FINE; i++) {
        const double l = T.lam01[i];
        const double x = (c[0] * l + c[1]) * l + c[2];
        const double s = 0.5 * x / sqrt(1.0 + x * x) + 0.5;
        out[0] += T.rgb_tbl[i] * s;
        out[1] += T.rgb_tbl[RGB2SPEC_FINE + i] * s;
        out[2] += T.rgb_tbl[2 * RGB2SPEC_FINE + i] * s;
    }
    cie_lab(out, T.white);
    r[0] = rgb[0];
    r[1] = rgb[1];
    r[2] = rgb[2];
    cie_lab(r, T.white);
    for (int j = 0; j < 3; j++) r[j] -= out[j];
}

// solves J x = b (3 x 3, partial pivoting); false when a pivot is below 1e-15 as LUPDecompose reports it
__device__ bool solve3(double J[3][3], const double b[3], double x[3]) {
    double a[3][4];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) a[i][j] = J[i][j];
        a[i][3] = b[i];
    }
    for (int col = 0; col < 3; col++) {
        int piv = col;
        for (int r = col + 1; r < 3; r++)
            if (fabs(a[r][col]) > fabs(a[piv][col])) piv = r;
        if (fabs(a[piv][col]) < 1e-15) return false;
        if (piv != col)
            for (int j = 0; j < 4; j++) {
                const double t = a[col][j];
                a[col][j] = a[piv][j];
                a[piv][j] = t;
            }
        for (int r = col + 1; r < 3; r++) {
            const double f = a[r][col] / a[col][col];
            for (int j = col; j < 4; j++) a[r][j] -= f * a[col][j];
        }
    }
    for (int i = 2; i >= 0; i--) {
        double s = a[i][3];
        for (int j = i + 1; j < 3; j++) s -= a[i][j] * x[j];
        x[i] = s / a[i][i];
    }
    return true;
}

__device__ bool gauss_newton(const FitTables &T, const double rgb[3], double c[3]) {
    const double eps = 1e-4;
    for (int it = 0; it < 15; it++) {
        double r[3], J[3][3];
        eval_residual(T, c, rgb, r);
        for (int i = 0; i < 3; i++) {
            double tmp[3] = {c[0], c[1], c[2]}, r0[3], r1[3];
            tmp[i] = c[i] - eps;
            eval_residual(T, tmp, rgb, r0);
            tmp[i] = c[i] + eps;
            eval_residual(T, tmp, rgb, r1);
            for (int j = 0; j < 3; j++) J[j][i] = (r1[j] - r0[j]) * 1.0 / (2 * eps);
        }
        double x[3];
        if (!solve3(J, r, x)) return false;
        double rr = 0.0;
        for (int j = 0; j < 3; j++) {
            c[j] -= x[j];
            rr += r[j] * r[j];
        }
        const double mx = fmax(fmax(c[0], c[1]), c[2]);
        if (mx > 200)
            for (int j = 0; j < 3; j++) c[j] *= 200 / mx;
        if (rr < 1e-6) break;
    }
    return true;
}

// coefficients[l][k][j][i][3] (the reference's layout, spectrum_util/rgb_to_spectrum_data.h:30)
__global__ void k_rgb2spec_fit(FitTables T, const double *z_nodes, double *coefficients, unsigned *failed) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= 3 * RGB2SPEC_RES * RGB2SPEC_RES) return;
    const int i = col % RGB2SPEC_RES, j = (col / RGB2SPEC_RES) % RGB2SPEC_RES, l = col / (RGB2SPEC_RES * RGB2SPEC_RES);
    const double x = i / double(RGB2SPEC_RES - 1), y = j / double(RGB2SPEC_RES - 1);
    const double c0 = 360.0, c1 = 1.0 / (830.0 - 360.0);
    const int start = RGB2SPEC_RES / 5;
    for (int sweep = 0; sweep < 2; sweep++) {
        double c[3] = {0.0, 0.0, 0.0};
        for (int k = start; sweep == 0 ? k < RGB2SPEC_RES : k >= 0; k += sweep == 0 ? 1 : -1) {
            const double b = z_nodes[k];
            double rgb[3];
            rgb[l] = b;
            rgb[(l + 1) % 3] = x * b;
            rgb[(l + 2) % 3] = y * b;
            if (!gauss_newton(T, rgb, c)) *failed = 1u;
            const double A = c[0], B = c[1], C = c[2];
            const size_t idx = ((size_t(l) * RGB2SPEC_RES + k) * RGB2SPEC_RES + j) * RGB2SPEC_RES + i;
            coefficients[3 * idx + 0] = A * (c1 * c1);
            coefficients[3 * idx + 1] = B * c1 - 2 * A * c0 * (c1 * c1);
            coefficients[3 * idx + 2] = C - B * c0 * c1 + A * ((c0 * c1) * (c0 * c1));
        }
    }
}

// fit: [3*283 rgb_tbl | 3 white | 283 lam01] in fp64 (the front-end's Simpson tables, spectral._fit_tables).
// Outputs live until the scene is destroyed (freed with the render allocations by the caller).
int build_rgb2spec(b200pt_scene *s, const double *fit, double **d_coeff_out, double **d_znodes_out, std::vector<void *> &allocs) {
    const size_t n_fit = 3 * RGB2SPEC_FINE + 3 + RGB2SPEC_FINE;
    const size_t n_coeff = size_t(3) * RGB2SPEC_RES * RGB2SPEC_RES * RGB2SPEC_RES * 3;
    double *d_fit = nullptr, *d_z = nullptr, *d_c = nullptr;
    unsigned *d_failed = nullptr;
    CU_TRY(cudaMalloc((void **)&d_fit, n_fit * sizeof(double)));
    allocs.push_back(d_fit);
    CU_TRY(cudaMalloc((void **)&d_z, RGB2SPEC_RES * sizeof(double)));
    allocs.push_back(d_z);
    CU_TRY(cudaMalloc((void **)&d_c, n_coeff * sizeof(double)));
    allocs.push_back(d_c);
    CU_TRY(cudaMalloc((void **)&d_failed, sizeof(unsigned)));
    allocs.push_back(d_failed);
    double z[RGB2SPEC_RES];
    auto smoothstep = [](double x) { return x * x * (3.0 - 2.0 * x); };
    for (int k = 0; k < RGB2SPEC_RES; k++) z[k] = smoothstep(smoothstep(double(k) / double(RGB2SPEC_RES - 1)));
    CU_TRY(cudaMemcpyAsync(d_fit, fit, n_fit * sizeof(double), cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(d_z, z, sizeof(z), cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemsetAsync(d_failed, 0, sizeof(unsigned), s->stream));
    FitTables T{d_fit, d_fit + 3 * RGB2SPEC_FINE, d_fit + 3 * RGB2SPEC_FINE + 3};
    const int cols = 3 * RGB2SPEC_RES * RGB2SPEC_RES;
    k_rgb2spec_fit<<<(cols + 63) / 64, 64, 0, s->stream>>>(T, d_z, d_c, d_failed);
    unsigned failed = 0;
    CU_TRY(cudaMemcpyAsync(&failed, d_failed, sizeof(unsigned), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    if (failed) {
        set_error("b200pt_scene_create: RGB-to-spectrum table fit hit a singular step");
        return B200PT_ERR_STATE;
    }
    *d_coeff_out = d_c;
    *d_znodes_out = d_z;
    return B200PT_OK;
}

} // namespace b200pt
