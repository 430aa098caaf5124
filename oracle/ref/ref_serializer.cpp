// ORACLE - TEST INFRASTRUCTURE ONLY.
//
// Second translation unit of oracle/_ref/libref_host.so: the reference's src/scene/serializer.cpp (LoadScene / SaveScene:
// JSON through its vendored json.hpp, zlib side files through its vendored miniz), compiled from oracle/_ref/gen/
// serializer_host.cpp - the file as it lies under /root/reference but for one MSVC-only std::max (oracle/ref/hostgen.py).
// It defines SERIALIZER_IMPLEMENTATION before scene.hpp itself, which is why it cannot share ref_host.cpp's unit.
// tests/test_reference_host.py exchanges scene files between this and path_tracer_b200/host/serializer.cpp.
#include <cassert>
#include <cstdio>
#include <cstring>
static inline int fopen_s(FILE** f, const char* path, const char* mode) { *f = std::fopen(path, mode); return *f ? 0 : 1; }
#include "serializer_host.cpp"
