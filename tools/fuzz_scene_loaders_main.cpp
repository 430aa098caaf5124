#include "scene.hpp"
#include <cstdio>
#include <cstring>
using namespace ptb::host;
int main(int argc, char** argv) {
    int ok = 0;
    for (int i = 2; i < argc; i++) {
        if (!std::strcmp(argv[1], "obj")) {
            scene* S = CreateScene(false);
            prefab* P = LoadModelAsPrefab(S, argv[i]);
            if (P) { CreateEntity(S, P); PackSceneData(S); ok++; }
            DestroyScene(S);
        } else {
            std::string err;
            scene* S = LoadScene(argv[i], &err);
            if (S) { PackSceneData(S); ok++; DestroyScene(S); }
        }
    }
    std::printf("loaded %d of %d\n", ok, argc - 2);
    return 0;
}
