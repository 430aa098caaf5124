// Host-side RGB -> parametric-spectrum coefficients ("spectral upsampling").
// Mirrors the role of /root/reference/src/core/spectrum.{hpp,cpp}: a
// 3 x 64 x 64 x 64 table of (beta0,beta1,beta2) fitted by Gauss–Newton in
// CIELAB under D65, looked up trilinearly.  The table file format is the
// reference's raw dump (spectrum.cpp:413-437): float[3][64][64][64][3].
#pragma once
#include "hmath.hpp"

namespace ptb { namespace host {

struct parametric_spectrum_table {
    static constexpr int SCALE_BINS = 64;   // spectrum.hpp:7
    static constexpr int COLOR_BINS = 64;   // spectrum.hpp:8
    vec3 Coefficients[3][SCALE_BINS][COLOR_BINS][COLOR_BINS];
};

// spectrum.hpp:13-21
void BuildParametricSpectrumTableForSRGB(parametric_spectrum_table* Table, int Threads = 0);
bool SaveParametricSpectrumTable(parametric_spectrum_table const* Table, char const* Path);
bool LoadParametricSpectrumTable(parametric_spectrum_table* Table, char const* Path);
vec3 GetParametricSpectrumCoefficients(parametric_spectrum_table const* Table, vec3 const& Color);
float SampleParametricSpectrum(vec3 const& Beta, float Lambda);

// Direct per-colour fit (no table): used to validate table entries in tests.
vec3 FitParametricSpectrum(vec3 const& SRGBColor);

}}  // namespace ptb::host
