// Spectral upsampling on the GPU: the sRGB -> parametric-spectrum table
// builder and the per-texel RGB -> coefficient map that the atlas packer applies.
//
// The reference fits 3 x 64 x 64 x 64 sigmoid-quadratic spectra by Gauss-Newton in CIELAB under
// D65 on one host thread (src/core/spectrum.cpp:240-304,365-411; "takes minutes",
// src/scene/scene.cpp:923-925) and converts every texel of every texture through a trilinear
// table lookup when it packs the atlas (src/scene/scene.cpp:1187-1212).  Both are independent
// per column / per texel:
//
//   k_fit_columns   one WARP per table column (l, j, i): the 64 fits along the scale axis are a
//                   warm-started chain (spectrum.cpp:391-407) and stay sequential; the 471-sample
//                   observation integral of every Gauss-Newton residual/Jacobian evaluation is
//                   spread over the 32 lanes (fp64, butterfly reduction => all lanes hold the same bits).
//   k_upsample      one thread per texel, float arithmetic in the host layer's order (bit-identical
//                   to GetParametricSpectrumCoefficients with --fmad=false).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "ptb.h"

namespace ptb {
int api_fail(int code, const std::string& msg);          // ptb_api.cu: sets ptb_last_error
cudaStream_t api_stream(ptb_context* ctx);
int api_device(ptb_context* ctx);
}

namespace {

constexpr int SCALE_BINS = 64, COLOR_BINS = 64;          // spectrum.hpp:7-8
constexpr int N_SAMPLES = 471;
constexpr float LAMBDA_MIN = 360.0f, LAMBDA_MAX = 830.0f;

const float D65[N_SAMPLES] = {
#include "../host/d65_table.inc"
};

struct d3 { double x, y, z; };
__host__ __device__ inline d3 operator+(d3 a, d3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__host__ __device__ inline d3 operator-(d3 a, d3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__host__ __device__ inline d3 operator*(d3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__host__ __device__ inline d3 operator/(d3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }

// Wavelength-only factor of the D65 observation integral (spectrum.cpp:137-143,151-176,190-208):
// observer(lambda) * D65(lambda) / 10566.864005, with the sample position t in .w.  Computed on the host once.
void make_weights(double4* w) {
    for (int i = 0; i < N_SAMPLES; i++) {
        const double t = i / double(N_SAMPLES - 1);
        const double offset = t * (N_SAMPLES - 1);
        const int k = std::min(std::max((int)offset, 0), N_SAMPLES - 2);
        const double a = offset - k;
        const double d65 = (double)(float)((double)D65[k] * (1.0 - a) + (double)D65[k + 1] * a);
        const float L = (float)((double)LAMBDA_MIN * (1.0 - t) + (double)LAMBDA_MAX * t);
        float X, Y, Z;
        {
            float T1 = (L - 442.0f) * (L < 442.0f ? 0.0624f : 0.0374f);
            float T2 = (L - 599.8f) * (L < 599.8f ? 0.0264f : 0.0323f);
            float T3 = (L - 501.1f) * (L < 501.1f ? 0.0490f : 0.0382f);
            X = 0.362f * std::exp(-0.5f * T1 * T1) + 1.056f * std::exp(-0.5f * T2 * T2) - 0.065f * std::exp(-0.5f * T3 * T3);
        }
        {
            float T1 = (L - 568.8f) * (L < 568.8f ? 0.0213f : 0.0247f);
            float T2 = (L - 530.9f) * (L < 530.9f ? 0.0613f : 0.0322f);
            Y = 0.821f * std::exp(-0.5f * T1 * T1) + 0.286f * std::exp(-0.5f * T2 * T2);
        }
        {
            float T1 = (L - 437.0f) * (L < 437.0f ? 0.0845f : 0.0278f);
            float T2 = (L - 459.0f) * (L < 459.0f ? 0.0385f : 0.0725f);
            Z = 1.217f * std::exp(-0.5f * T1 * T1) + 0.681f * std::exp(-0.5f * T2 * T2);
        }
        const double s = d65 / 10566.864005;
        w[i] = make_double4((double)X * s, (double)Y * s, (double)Z * s, t);
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// spectrum.cpp:183-187
__device__ __forceinline__ double sigmoid_spectrum(d3 b, double t) {
    const double x = (b.x * t + b.y) * t + b.z;
    return 0.5 + x / (2.0 * sqrt(1.0 + x * x));
}

// spectrum.cpp:192-208 for NB coefficient sets at once, the samples dealt to the lanes.
template <int NB>
__device__ __forceinline__ void observe(const double4* __restrict__ w, const d3* beta, d3* xyz, int lane) {
    d3 acc[NB];
#pragma unroll
    for (int b = 0; b < NB; b++) acc[b] = {0.0, 0.0, 0.0};
    for (int s = lane; s < N_SAMPLES; s += 32) {
        const double4 ws = w[s];
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const double g = sigmoid_spectrum(beta[b], ws.w);
            acc[b].x += ws.x * g; acc[b].y += ws.y * g; acc[b].z += ws.z * g;
        }
    }
#pragma unroll
    for (int b = 0; b < NB; b++) xyz[b] = {warp_sum(acc[b].x), warp_sum(acc[b].y), warp_sum(acc[b].z)};
}

// spectrum.cpp:213-233 (L is driven by f(X), as in the reference)
__device__ __forceinline__ double lab_f(double t) {
    const double d = 6 / 29.0;
    return t > d * d * d ? pow(t, 1 / 3.0) : t / (3 * d * d) + 4 / 29.0;
}
__device__ __forceinline__ d3 xyz_to_lab(d3 c) {
    const double fx = lab_f(c.x / 0.950489), fy = lab_f(c.y), fz = lab_f(c.z / 1.088840);
    return {116.0 * fx - 16.0, 500.0 * (fx - fy), 200.0 * (fy - fz)};
}

// spectrum.cpp:240-304: Gauss-Newton with a central-difference Jacobian; every lane runs the same scalar code.
__device__ d3 optimize(const double4* __restrict__ w, d3 beta, d3 target_xyz, int lane) {
    const double eps = 1e-5;
    const d3 target_lab = xyz_to_lab(target_xyz);
    for (int it = 0; it < 15; it++) {
        d3 xyz0;
        observe<1>(w, &beta, &xyz0, lane);
        const d3 r = xyz_to_lab(xyz0) - target_lab;
        if (sqrt(r.x * r.x + r.y * r.y + r.z * r.z) < 1e-3) break;
        d3 probe[6], obs[6];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            probe[2 * k] = beta; probe[2 * k + 1] = beta;
            (k == 0 ? probe[2 * k].x : k == 1 ? probe[2 * k].y : probe[2 * k].z) -= eps;
            (k == 0 ? probe[2 * k + 1].x : k == 1 ? probe[2 * k + 1].y : probe[2 * k + 1].z) += eps;
        }
        observe<6>(w, probe, obs, lane);
        d3 col[3];
#pragma unroll
        for (int k = 0; k < 3; k++) col[k] = (xyz_to_lab(obs[2 * k + 1]) - xyz_to_lab(obs[2 * k])) / (2 * eps);
        const d3 a = col[0], b = col[1], c = col[2];
        const double det = a.x * (b.y * c.z - c.y * b.z) - b.x * (a.y * c.z - c.y * a.z) + c.x * (a.y * b.z - b.y * a.z);
        if (fabs(det) < 1e-15) break;
        const double id = 1.0 / det;
        const d3 i0{(b.y * c.z - c.y * b.z) * id, -(a.y * c.z - c.y * a.z) * id, (a.y * b.z - b.y * a.z) * id};
        const d3 i1{-(b.x * c.z - c.x * b.z) * id, (a.x * c.z - c.x * a.z) * id, -(a.x * b.z - b.x * a.z) * id};
        const d3 i2{(b.x * c.y - c.x * b.y) * id, -(a.x * c.y - c.x * a.y) * id, (a.x * b.y - b.x * a.y) * id};
        const d3 d = (i0 * r.x + i1 * r.y) + i2 * r.z;
        beta = beta - d;
        const double mx = fmax(fmax(beta.x, beta.y), beta.z);
        if (mx > 200.0) beta = beta * (200.0 / mx);
    }
    return beta;
}

// spectrum.cpp:306-313
__host__ __device__ inline float index_to_scale(int k) {
    const float r = k / float(SCALE_BINS - 1);
    const float s = r * r * (3.f - 2.f * r);
    return s * s * (3.f - 2.f * s);
}
// spectrum.cpp:326-334 followed by spectrum.cpp:14-19 (sRGB -> XYZ, column-major literal)
__device__ __forceinline__ d3 column_target(int i, int j, int k, int l) {
    float c[3];
    c[l] = 1.0f;
    c[(l + 1) % 3] = i / float(COLOR_BINS - 1);
    c[(l + 2) % 3] = j / float(COLOR_BINS - 1);
    const float s = index_to_scale(k);
    c[0] *= s; c[1] *= s; c[2] *= s;
    const float x = (0.4124f * c[0] + 0.3576f * c[1]) + 0.1805f * c[2];
    const float y = (0.2126f * c[0] + 0.7152f * c[1]) + 0.0722f * c[2];
    const float z = (0.0193f * c[0] + 0.1192f * c[1]) + 0.9505f * c[2];
    return {x, y, z};
}
// spectrum.cpp:370-381
__device__ __forceinline__ void store_denormalized(float* table, int l, int k, int j, int i, d3 b) {
    const float C0 = LAMBDA_MIN;
    const float C1 = 1.f / (LAMBDA_MAX - LAMBDA_MIN);
    float* o = table + 3 * ((((size_t)l * SCALE_BINS + k) * COLOR_BINS + j) * COLOR_BINS + i);
    o[0] = (float)(b.x * C1 * C1);
    o[1] = (float)(b.y * C1 - 2 * b.x * C0 * C1 * C1);
    o[2] = (float)(b.z - b.y * C0 * C1 + b.x * C0 * C0 * C1 * C1);
}

__global__ void __launch_bounds__(128) k_fit_columns(const double4* __restrict__ weights, float* __restrict__ table) {
    __shared__ double4 w[N_SAMPLES];
    for (int s = threadIdx.x; s < N_SAMPLES; s += blockDim.x) w[s] = weights[s];
    __syncthreads();
    const int column = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (column >= 3 * COLOR_BINS * COLOR_BINS) return;
    const int l = column / (COLOR_BINS * COLOR_BINS), j = (column / COLOR_BINS) % COLOR_BINS, i = column % COLOR_BINS;
    d3 beta{0, 0, 0};
    for (int k = SCALE_BINS / 5; k < SCALE_BINS; k++) {        // spectrum.cpp:391-407
        beta = optimize(w, beta, column_target(i, j, k, l), lane);
        if (lane == 0) store_denormalized(table, l, k, j, i, beta);
    }
    beta = {0, 0, 0};
    for (int k = SCALE_BINS / 5; k >= 0; k--) {
        beta = optimize(w, beta, column_target(i, j, k, l), lane);
        if (lane == 0) store_denormalized(table, l, k, j, i, beta);
    }
}

struct f3 { float x, y, z; };
__device__ __forceinline__ f3 lerp3(f3 a, f3 b, float t) {
    const float u = 1.0f - t;
    return {a.x * u + b.x * t, a.y * u + b.y * t, a.z * u + b.z * t};
}
// spectrum.cpp:336-363,439-479: trilinear lookup in the host layer's operation order.
__device__ f3 lookup(const float* __restrict__ table, float r, float g, float b) {
    const int N = COLOR_BINS, M = SCALE_BINS;
    float c[3] = {fminf(fmaxf(r, 0.0f), 1.0f), fminf(fmaxf(g, 0.0f), 1.0f), fminf(fmaxf(b, 0.0f), 1.0f)};
    int l = 0;
    for (int k = 1; k < 3; k++) if (c[k] >= c[l]) l = k;
    const float scale = fmaxf(c[l], 1e-6f);
    const float x = (N - 1) * c[(l + 1) % 3] / scale;
    const float y = (N - 1) * c[(l + 2) % 3] / scale;
    int k0 = 0, k1 = M;
    while (k1 - k0 > 1) { const int k = (k0 + k1) / 2; if (scale > index_to_scale(k)) k0 = k; else k1 = k; }
    const int i = min((int)x, N - 2), j = min((int)y, N - 2), k = min(k0, M - 2);
    const float s0 = index_to_scale(k), s1 = index_to_scale(k + 1);
    const float ax = x - i, ay = y - j, az = (scale - s0) / (s1 - s0);
    auto at = [&](int kk, int jj, int ii) {
        const float* p = table + 3 * ((((size_t)l * M + kk) * N + jj) * N + ii);
        return f3{p[0], p[1], p[2]};
    };
    const f3 b00 = lerp3(at(k, j, i), at(k, j, i + 1), ax);
    const f3 b01 = lerp3(at(k, j + 1, i), at(k, j + 1, i + 1), ax);
    const f3 b10 = lerp3(at(k + 1, j, i), at(k + 1, j, i + 1), ax);
    const f3 b11 = lerp3(at(k + 1, j + 1, i), at(k + 1, j + 1, i + 1), ax);
    return lerp3(lerp3(b00, b01, ay), lerp3(b10, b11, ay), az);
}

// Texel conversion of the atlas packer (scene.cpp:1187-1212): type 0 raw copy, 1 reflectance + alpha, 2 radiance.
__global__ void k_upsample(const float* __restrict__ table, const float4* __restrict__ in, float4* __restrict__ out, size_t n, uint32_t type) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const float4 v = in[t];
    float4 o = v;
    if (type == 1u) {
        const f3 b = lookup(table, v.x, v.y, v.z);
        o = make_float4(b.x, b.y, b.z, v.w);
    } else if (type == 2u) {
        const float I = 2 * fmaxf(fmaxf(v.x, v.y), v.z);
        if (I > 1e-6f) { const f3 b = lookup(table, v.x / I, v.y / I, v.z / I); o = make_float4(b.x, b.y, b.z, I); }
        else o = make_float4(0, 0, 0, 0);
    }
    out[t] = o;
}

#define SP(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { rc = ptb::api_fail(PTB_E_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); goto done; } } while (0)

}  // namespace

extern "C" {

int ptb_spectrum_table_build(ptb_context* ctx, float* host_table, size_t count_floats, float* device_ms) {
    const size_t need = (size_t)3 * SCALE_BINS * COLOR_BINS * COLOR_BINS * 3;
    if (!ctx || !host_table) return ptb::api_fail(PTB_E_INVALID, "ptb_spectrum_table_build: NULL argument");
    if (count_floats != need) return ptb::api_fail(PTB_E_INVALID, "ptb_spectrum_table_build: the table is float[3][64][64][64][3]");
    int rc = PTB_OK;
    cudaStream_t st = ptb::api_stream(ctx);
    double4* d_w = nullptr; float* d_table = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    std::vector<double4> w(N_SAMPLES);
    make_weights(w.data());
    SP(cudaSetDevice(ptb::api_device(ctx)));
    SP(cudaMalloc((void**)&d_w, N_SAMPLES * sizeof(double4)));
    SP(cudaMalloc((void**)&d_table, need * sizeof(float)));
    SP(cudaEventCreate(&e0)); SP(cudaEventCreate(&e1));
    SP(cudaMemcpyAsync(d_w, w.data(), N_SAMPLES * sizeof(double4), cudaMemcpyHostToDevice, st));
    SP(cudaEventRecord(e0, st));
    k_fit_columns<<<(3 * COLOR_BINS * COLOR_BINS + 3) / 4, 128, 0, st>>>(d_w, d_table);
    SP(cudaGetLastError());
    SP(cudaEventRecord(e1, st));
    SP(cudaMemcpyAsync(host_table, d_table, need * sizeof(float), cudaMemcpyDeviceToHost, st));
    SP(cudaStreamSynchronize(st));
    if (device_ms) SP(cudaEventElapsedTime(device_ms, e0, e1));
done:
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(d_w); cudaFree(d_table);
    return rc;
}

int ptb_spectrum_upsample(ptb_context* ctx, const float* host_table, size_t count_floats, const float* host_rgba_in,
                          float* host_rgba_out, size_t texels, uint32_t texture_type) {
    const size_t need = (size_t)3 * SCALE_BINS * COLOR_BINS * COLOR_BINS * 3;
    if (!ctx || !host_table || (texels && (!host_rgba_in || !host_rgba_out))) return ptb::api_fail(PTB_E_INVALID, "ptb_spectrum_upsample: NULL argument");
    if (count_floats != need) return ptb::api_fail(PTB_E_INVALID, "ptb_spectrum_upsample: the table is float[3][64][64][64][3]");
    if (texture_type > 2u) return ptb::api_fail(PTB_E_INVALID, "ptb_spectrum_upsample: unknown texture type");
    if (texels == 0) return PTB_OK;
    int rc = PTB_OK;
    cudaStream_t st = ptb::api_stream(ctx);
    float* d_table = nullptr; float4 *d_in = nullptr, *d_out = nullptr;
    SP(cudaSetDevice(ptb::api_device(ctx)));
    SP(cudaMalloc((void**)&d_table, need * sizeof(float)));
    SP(cudaMalloc((void**)&d_in, texels * sizeof(float4)));
    SP(cudaMalloc((void**)&d_out, texels * sizeof(float4)));
    SP(cudaMemcpyAsync(d_table, host_table, need * sizeof(float), cudaMemcpyHostToDevice, st));
    SP(cudaMemcpyAsync(d_in, host_rgba_in, texels * sizeof(float4), cudaMemcpyHostToDevice, st));
    k_upsample<<<(unsigned)((texels + 255) / 256), 256, 0, st>>>(d_table, d_in, d_out, texels, texture_type);
    SP(cudaGetLastError());
    SP(cudaMemcpyAsync(host_rgba_out, d_out, texels * sizeof(float4), cudaMemcpyDeviceToHost, st));
    SP(cudaStreamSynchronize(st));
done:
    cudaFree(d_table); cudaFree(d_in); cudaFree(d_out);
    return rc;
}

}  // extern "C"
