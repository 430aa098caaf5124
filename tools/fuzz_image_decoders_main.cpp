#include "scene.hpp"
#include <cstdio>
using namespace ptb::host;
int main(int argc, char** argv) {
    int ok = 0;
    for (int i = 1; i < argc; i++) { int w, h; std::vector<vec4> px; if (DecodeImageFile(argv[i], w, h, px)) ok++; }
    std::printf("decoded %d of %d\n", ok, argc - 1);
    return 0;
}
