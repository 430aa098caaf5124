// See spectrum.hpp.  Follows the algorithm of /root/reference/src/core/spectrum.cpp
// (cited per function) but is organised differently: the wavelength-only factor
// observer(lambda) * D65(lambda) is tabulated once, and table columns are fitted
// in parallel (the reference is single-threaded and takes minutes, scene.cpp:923-925).
#include "spectrum.hpp"
#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>

namespace ptb { namespace host {

namespace {

constexpr float LAMBDA_MIN = 360.0f, LAMBDA_MAX = 830.0f;
constexpr int   N_SAMPLES = 471;

const float D65[N_SAMPLES] = {
#include "d65_table.inc"
};

struct d3 { double x, y, z; };
inline d3 operator+(d3 a, d3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline d3 operator-(d3 a, d3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline d3 operator*(d3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline d3 operator/(d3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }

// spectrum.cpp:137-143: lerp of the float table evaluated in double, returned as float.
double sample_d65(double t) {
    double offset = t * (N_SAMPLES - 1);
    int i = std::min(std::max((int)offset, 0), N_SAMPLES - 2);
    double a = offset - i;
    return (double)(float)((double)D65[i] * (1.0 - a) + (double)D65[i + 1] * a);
}

// spectrum.cpp:151-176: Wyman fit, evaluated in fp32 like the reference.
d3 sample_observer(double t) {
    float L = (float)((double)LAMBDA_MIN * (1.0 - t) + (double)LAMBDA_MAX * t);
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
    return {X, Y, Z};
}

// Wavelength-only factor of the D65 observation integral (spectrum.cpp:190-208).
struct weights {
    d3 ow[N_SAMPLES];
    double t[N_SAMPLES];
    weights() {
        for (int i = 0; i < N_SAMPLES; i++) {
            t[i] = i / double(N_SAMPLES - 1);
            double w = sample_d65(t[i]) / 10566.864005;
            ow[i] = sample_observer(t[i]) * w;
        }
    }
};
const weights& W() { static weights w; return w; }

// spectrum.cpp:183-187 (normalised coefficients, lambda in [0,1])
inline double sigmoid_spectrum(d3 b, double t) {
    double x = (b.x * t + b.y) * t + b.z;
    return 0.5 + x / (2.0 * std::sqrt(1.0 + x * x));
}

// spectrum.cpp:192-208; DeltaLambda = (830-360+1)/471 = 1.
d3 observe_under_d65(d3 beta) {
    const weights& w = W();
    const double delta = (double)(LAMBDA_MAX - LAMBDA_MIN + 1) / N_SAMPLES;
    d3 xyz{0, 0, 0};
    for (int i = 0; i < N_SAMPLES; i++) xyz = xyz + (w.ow[i] * sigmoid_spectrum(beta, w.t[i])) * delta;
    return xyz;
}

// spectrum.cpp:213-233 (note L is driven by f(X), as in the reference).
d3 xyz_to_lab(d3 c) {
    auto f = [](double t) {
        const double d = 6 / 29.0;
        return t > d * d * d ? std::pow(t, 1 / 3.0) : t / (3 * d * d) + 4 / 29.0;
    };
    double fx = f(c.x / 0.950489), fy = f(c.y), fz = f(c.z / 1.088840);
    return {116.0 * fx - 16.0, 500.0 * (fx - fy), 200.0 * (fy - fz)};
}

// spectrum.cpp:240-304  Gauss–Newton with a central-difference Jacobian.
d3 optimize(d3 beta, d3 target_xyz, int iterations = 15) {
    const double eps = 1e-5;
    const d3 target_lab = xyz_to_lab(target_xyz);
    for (int it = 0; it < iterations; it++) {
        d3 r = xyz_to_lab(observe_under_d65(beta)) - target_lab;
        double err = std::sqrt(r.x * r.x + r.y * r.y + r.z * r.z);
        if (err < 1e-3) break;
        d3 col[3];
        for (int k = 0; k < 3; k++) {
            d3 b0 = beta, b1 = beta;
            (&b0.x)[k] -= eps; (&b1.x)[k] += eps;
            col[k] = (xyz_to_lab(observe_under_d65(b1)) - xyz_to_lab(observe_under_d65(b0))) / (2 * eps);
        }
        // J = [col0 col1 col2]; solve J * d = r by cofactors.
        const d3 &a = col[0], &b = col[1], &c = col[2];
        double det = a.x * (b.y * c.z - c.y * b.z) - b.x * (a.y * c.z - c.y * a.z) + c.x * (a.y * b.z - b.y * a.z);
        if (std::fabs(det) < 1e-15) break;
        double id = 1.0 / det;
        d3 i0{(b.y * c.z - c.y * b.z) * id, -(a.y * c.z - c.y * a.z) * id, (a.y * b.z - b.y * a.z) * id};
        d3 i1{-(b.x * c.z - c.x * b.z) * id, (a.x * c.z - c.x * a.z) * id, -(a.x * b.z - b.x * a.z) * id};
        d3 i2{(b.x * c.y - c.x * b.y) * id, -(a.x * c.y - c.x * a.y) * id, (a.x * b.y - b.x * a.y) * id};
        d3 d = (i0 * r.x + i1 * r.y) + i2 * r.z;     // inverse columns times residual
        beta = beta - d;
        double mx = std::max(std::max(beta.x, beta.y), beta.z);
        if (mx > 200.0) beta = beta * (200.0 / mx);
    }
    return beta;
}

// spectrum.cpp:14-19,26-30 (column-major literal)
d3 srgb_to_xyz(vec3 c) {
    float x = (0.4124f * c.x + 0.3576f * c.y) + 0.1805f * c.z;
    float y = (0.2126f * c.x + 0.7152f * c.y) + 0.0722f * c.z;
    float z = (0.0193f * c.x + 0.1192f * c.y) + 0.9505f * c.z;
    return {x, y, z};
}

// spectrum.cpp:306-313
float index_to_scale(int k) {
    const int M = parametric_spectrum_table::SCALE_BINS;
    float r = k / float(M - 1);
    float s = r * r * (3.f - 2.f * r);
    return s * s * (3.f - 2.f * s);
}
// spectrum.cpp:315-324
int scale_to_index(float scale) {
    int k0 = 0, k1 = parametric_spectrum_table::SCALE_BINS;
    while (k1 - k0 > 1) {
        int k = (k0 + k1) / 2;
        if (scale > index_to_scale(k)) k0 = k; else k1 = k;
    }
    return k0;
}
// spectrum.cpp:326-334
vec3 index_to_color(int i, int j, int k, int l) {
    const int N = parametric_spectrum_table::COLOR_BINS;
    vec3 c;
    c[l] = 1.0f;
    c[(l + 1) % 3] = i / float(N - 1);
    c[(l + 2) % 3] = j / float(N - 1);
    return c * index_to_scale(k);
}
// spectrum.cpp:370-381: normalised-lambda coefficients -> nanometre coefficients.
vec3 denormalize(d3 b) {
    const float C0 = LAMBDA_MIN;
    const float C1 = 1.f / (LAMBDA_MAX - LAMBDA_MIN);
    return {(float)(b.x * C1 * C1),
            (float)(b.y * C1 - 2 * b.x * C0 * C1 * C1),
            (float)(b.z - b.y * C0 * C1 + b.x * C0 * C0 * C1 * C1)};
}

// One (l, j, i) column of the table: spectrum.cpp:391-407.
void fit_column(parametric_spectrum_table* T, int l, int j, int i) {
    const int M = parametric_spectrum_table::SCALE_BINS;
    d3 beta{0, 0, 0};
    for (int k = M / 5; k < M; k++) {
        beta = optimize(beta, srgb_to_xyz(index_to_color(i, j, k, l)));
        T->Coefficients[l][k][j][i] = denormalize(beta);
    }
    beta = {0, 0, 0};
    for (int k = M / 5; k >= 0; k--) {
        beta = optimize(beta, srgb_to_xyz(index_to_color(i, j, k, l)));
        T->Coefficients[l][k][j][i] = denormalize(beta);
    }
}

}  // namespace

// spectrum.cpp:365-411, columns distributed over host threads.
void BuildParametricSpectrumTableForSRGB(parametric_spectrum_table* Table, int Threads) {
    const int N = parametric_spectrum_table::COLOR_BINS;
    if (Threads <= 0) Threads = (int)std::max(1u, std::thread::hardware_concurrency());
    (void)W();
    std::atomic<int> next(0);
    const int total = 3 * N * N;
    std::vector<std::thread> pool;
    for (int t = 0; t < Threads; t++) {
        pool.emplace_back([&]() {
            for (;;) {
                int c = next.fetch_add(1);
                if (c >= total) break;
                fit_column(Table, c / (N * N), (c / N) % N, c % N);
            }
        });
    }
    for (auto& th : pool) th.join();
}

bool SaveParametricSpectrumTable(parametric_spectrum_table const* Table, char const* Path) {
    FILE* f = std::fopen(Path, "wb");
    if (!f) return false;
    bool ok = std::fwrite(Table->Coefficients, sizeof(Table->Coefficients), 1, f) == 1;
    std::fclose(f);
    return ok;
}

bool LoadParametricSpectrumTable(parametric_spectrum_table* Table, char const* Path) {
    FILE* f = std::fopen(Path, "rb");
    if (!f) return false;
    bool ok = std::fread(Table->Coefficients, sizeof(Table->Coefficients), 1, f) == 1;
    std::fclose(f);
    return ok;
}

// spectrum.cpp:336-363 + 439-479: trilinear lookup.
vec3 GetParametricSpectrumCoefficients(parametric_spectrum_table const* Table, vec3 const& InColor) {
    const int N = parametric_spectrum_table::COLOR_BINS, M = parametric_spectrum_table::SCALE_BINS;
    vec3 c = {std::min(std::max(InColor.x, 0.0f), 1.0f), std::min(std::max(InColor.y, 0.0f), 1.0f), std::min(std::max(InColor.z, 0.0f), 1.0f)};
    int l = 0;
    for (int k = 1; k < 3; k++) if (c[k] >= c[l]) l = k;
    float scale = std::max(c[l], 1e-6f);
    float x = (N - 1) * c[(l + 1) % 3] / scale;
    float y = (N - 1) * c[(l + 2) % 3] / scale;
    int i = std::min((int)x, N - 2), j = std::min((int)y, N - 2), k = std::min(scale_to_index(scale), M - 2);
    float s0 = index_to_scale(k), s1 = index_to_scale(k + 1);
    float ax = x - i, ay = y - j, az = (scale - s0) / (s1 - s0);
    auto lerp = [](vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; };
    auto& C = Table->Coefficients[l];
    vec3 b00 = lerp(C[k][j][i], C[k][j][i + 1], ax);
    vec3 b01 = lerp(C[k][j + 1][i], C[k][j + 1][i + 1], ax);
    vec3 b10 = lerp(C[k + 1][j][i], C[k + 1][j][i + 1], ax);
    vec3 b11 = lerp(C[k + 1][j + 1][i], C[k + 1][j + 1][i + 1], ax);
    return lerp(lerp(b00, b01, ay), lerp(b10, b11, ay), az);
}

// spectrum.cpp:481-485
float SampleParametricSpectrum(vec3 const& Beta, float Lambda) {
    float X = (Beta.x * Lambda + Beta.y) * Lambda + Beta.z;
    return 0.5f + X / (2.0f * std::sqrt(1.0f + X * X));
}

vec3 FitParametricSpectrum(vec3 const& SRGBColor) {
    return denormalize(optimize({0, 0, 0}, srgb_to_xyz(SRGBColor), 50));
}

}}  // namespace ptb::host
