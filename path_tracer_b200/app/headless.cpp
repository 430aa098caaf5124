// ptb_headless: the reference application's per-frame loop (src/application/application.cpp:100-124) without a
// window - scene in, image file out - written against host/integrator.hpp, i.e. against the reference's own
// integrator interface.  Swapchain, ImGui and input are out of scope (no graphics stack on this box); the frame
// RenderSampleBuffer would present is written as a PNG, the film optionally as raw RGBA32F.
//
//   ptb_headless --scene cornell|terrain[:grid]|room[:detail[:atlas]]|spherefield[:copies[:level[:camera]]]|FILE.json
//                [--size WxH] [--spp N] [--max-frames F] [--camera I] [--tone clamp|reinhard|hable|aces] [--brightness B] [--white W]
//                [--device D] [--builder sah-host|lbvh-gpu|ploc-gpu|auto] [--out frame.png] [--film film.rgba32f] [--table PATH]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../host/integrator.hpp"
#include "../host/procedural.hpp"

using namespace ptb::host;

struct application {                              // application.hpp: the members Update() touches
    device*         Device = nullptr;             // Vulkan
    device_scene*   DeviceScene = nullptr;        // VulkanScene
    scene*          Scene = nullptr;
    sample_buffer*  SampleBuffer = nullptr;
    basic_renderer* BasicRenderer = nullptr;
    resolve_parameters ResolveParameters;
    uint            CameraIndex = 0;
};

// One pass of Update() (application.cpp:60-124) for a scene that may have changed.
static void Update(application* App, bool Restart) {
    uint32_t DirtyFlags = PackSceneData(App->Scene);
    if (DirtyFlags) UpdateDeviceScene(App->Device, App->DeviceScene, App->Scene, DirtyFlags);     // (it waits for the device: scene.cpp:1704)
    if (Restart) {
        App->BasicRenderer->CameraIndex = App->CameraIndex;
        App->BasicRenderer->Scene = App->DeviceScene;
        App->BasicRenderer->RenderFlags = RENDER_FLAG_ACCUMULATE | RENDER_FLAG_SAMPLE_JITTER;
        App->BasicRenderer->PathTerminationProbability = 0.0f;
        ResetBasicRenderer(App->Device, App->BasicRenderer);
        RunBasicRenderer(App->Device, App->BasicRenderer, 2);
    } else {
        RunBasicRenderer(App->Device, App->BasicRenderer, 1);
    }
}

static std::vector<double> SplitNumbers(std::string const& S) {
    std::vector<double> V;
    size_t I = 0;
    while (I < S.size()) {
        size_t J = S.find(':', I);
        if (J == std::string::npos) J = S.size();
        V.push_back(std::atof(S.substr(I, J - I).c_str()));
        I = J + 1;
    }
    return V;
}

static scene* OpenScene(std::string const& Spec, std::string* Error) {
    std::string Name = Spec.substr(0, Spec.find(':'));
    std::vector<double> P = Spec.find(':') == std::string::npos ? std::vector<double>() : SplitNumbers(Spec.substr(Spec.find(':') + 1));
    auto Arg = [&](size_t I, double D) { return I < P.size() ? P[I] : D; };
    if (Name == "cornell") return MakeCornellScene();
    if (Name == "terrain") return MakeTerrainScene((uint32_t)Arg(0, 708));
    if (Name == "room") return MakeRoomScene((uint32_t)Arg(0, 1), (uint32_t)Arg(1, 4096));
    if (Name == "spherefield") return MakeSphereFieldScene((uint32_t)Arg(0, 488), (uint32_t)Arg(1, 5), (int)Arg(2, 2));
    return LoadScene(Spec.c_str(), Error);
}

int main(int argc, char** argv) {
    std::string SceneSpec = "cornell", Out, FilmOut, Tone = "aces", Builder, Table;
    uint Width = 640, Height = 480, Spp = 64, Camera = 0, MaxFrames = 1u << 20;
    int CudaDevice = 0;
    float Brightness = 1.0f, White = 1.0f;
    for (int I = 1; I < argc; I++) {
        std::string A = argv[I];
        auto Next = [&]() -> char const* { if (I + 1 >= argc) { std::fprintf(stderr, "ptb_headless: %s needs a value\n", A.c_str()); std::exit(2); } return argv[++I]; };
        if (A == "--scene") SceneSpec = Next();
        else if (A == "--size") { if (std::sscanf(Next(), "%ux%u", &Width, &Height) != 2) { std::fprintf(stderr, "ptb_headless: --size WxH\n"); return 2; } }
        else if (A == "--spp") Spp = (uint)std::atoi(Next());
        else if (A == "--max-frames") MaxFrames = (uint)std::atoi(Next());
        else if (A == "--camera") Camera = (uint)std::atoi(Next());
        else if (A == "--tone") Tone = Next();
        else if (A == "--brightness") Brightness = (float)std::atof(Next());
        else if (A == "--white") White = (float)std::atof(Next());
        else if (A == "--device") CudaDevice = std::atoi(Next());
        else if (A == "--builder") Builder = Next();
        else if (A == "--out") Out = Next();
        else if (A == "--film") FilmOut = Next();
        else if (A == "--table") Table = Next();
        else if (A == "--help" || A == "-h") {
            std::printf("usage: ptb_headless --scene NAME[:params]|FILE.json [--size WxH] [--spp N] [--max-frames F] [--camera I] [--tone clamp|reinhard|hable|aces]\n"
                        "                    [--brightness B] [--white W] [--device D] [--builder sah-host|lbvh-gpu|ploc-gpu|auto]\n"
                        "                    [--out frame.png] [--film film.rgba32f] [--table spectrum_table.dat]\n");
            return 0;
        } else { std::fprintf(stderr, "ptb_headless: unknown argument %s\n", A.c_str()); return 2; }
    }
    tone_mapping_mode Mode = TONE_MAPPING_MODE__COUNT;
    for (uint M = 0; M < TONE_MAPPING_MODE__COUNT; M++) {
        std::string N = ToneMappingModeName((tone_mapping_mode)M);
        for (auto& C : N) C = (char)std::tolower(C);
        if (N == Tone) Mode = (tone_mapping_mode)M;
    }
    if (Mode == TONE_MAPPING_MODE__COUNT) { std::fprintf(stderr, "ptb_headless: unknown tone mapping mode %s\n", Tone.c_str()); return 2; }
    uint32_t BuilderId = PTB_BUILDER_SAH_HOST;
    if (Builder == "lbvh-gpu") BuilderId = PTB_BUILDER_LBVH_GPU;
    else if (Builder == "ploc-gpu") BuilderId = PTB_BUILDER_PLOC_GPU;
    else if (Builder == "auto") BuilderId = PTB_BUILDER_AUTO;
    else if (!Builder.empty() && Builder != "sah-host") { std::fprintf(stderr, "ptb_headless: unknown builder %s\n", Builder.c_str()); return 2; }

    application App;
    App.Device = CreateDevice(CudaDevice);        // fails loudly without a CUDA device: there is no CPU path
    if (!App.Device) return 1;
    // The RGB -> spectrum table (scene.cpp:918-927): loaded from the file, or - where the reference spends minutes of
    // host time on first run - fitted by the CUDA kernels of ptb_spectrum.cu and saved in the same raw format.
    if (Table.empty()) Table = "sRGBSpectrumTable.dat";
    SetSpectrumTablePath(Table.c_str());
    if (FILE* F = std::fopen(Table.c_str(), "rb")) std::fclose(F);
    else {
        std::vector<float> Coefficients((size_t)3 * 64 * 64 * 64 * 3);
        float Milliseconds = 0.0f;
        if (ptb_spectrum_table_build(App.Device->Context, Coefficients.data(), Coefficients.size(), &Milliseconds) != PTB_OK) { std::fprintf(stderr, "ptb_headless: %s\n", ptb_last_error()); return 1; }
        FILE* W = std::fopen(Table.c_str(), "wb");
        if (!W || std::fwrite(Coefficients.data(), sizeof(float), Coefficients.size(), W) != Coefficients.size()) { std::fprintf(stderr, "ptb_headless: cannot write %s\n", Table.c_str()); return 1; }
        std::fclose(W);
        std::fprintf(stderr, "ptb_headless: %s fitted on the GPU in %.1f ms\n", Table.c_str(), Milliseconds);
    }
    std::string Error;
    App.Scene = OpenScene(SceneSpec, &Error);
    if (!App.Scene) { std::fprintf(stderr, "ptb_headless: cannot open scene %s: %s\n", SceneSpec.c_str(), Error.c_str()); return 1; }
    App.DeviceScene = CreateDeviceScene(App.Device);
    App.SampleBuffer = CreateSampleBuffer(App.Device, Width, Height);
    if (!App.DeviceScene || !App.SampleBuffer) return 1;
    if (!Builder.empty() && ptb_scene_set_builder(App.DeviceScene->Handle, BuilderId) != PTB_OK) { std::fprintf(stderr, "ptb_headless: %s\n", ptb_last_error()); return 1; }
    PackSceneData(App.Scene);                     // (the Make*Scene helpers and LoadScene may have packed already: nothing is dirty then)
    UpdateDeviceScene(App.Device, App.DeviceScene, App.Scene, SCENE_DIRTY_ALL);     // a new device scene needs every array
    App.BasicRenderer = CreateBasicRenderer(App.Device, App.DeviceScene, App.SampleBuffer);
    if (!App.BasicRenderer) return 1;
    App.CameraIndex = Camera;
    App.ResolveParameters.Brightness = Brightness;
    App.ResolveParameters.ToneMappingMode = Mode;
    App.ResolveParameters.ToneMappingWhiteLevel = White;

    // Frames until every pixel has `Spp` finished paths on average (the film's alpha counts them).
    auto T0 = std::chrono::steady_clock::now();
    Update(&App, true);
    ptb_stats Stats = {};
    uint Frames = 0;
    double const Target = (double)Spp * Width * Height;
    for (;;) {
        uint64_t const Before = Stats.rays_traced;
        if (ptb_renderer_stats(App.BasicRenderer->Handle, &Stats) != PTB_OK) { std::fprintf(stderr, "ptb_headless: %s\n", ptb_last_error()); return 1; }
        if ((double)Stats.paths_finished >= Target) break;
        if (Stats.rays_traced == Before || Frames >= MaxFrames) {      // a failed Run* has already said why on stderr
            std::fprintf(stderr, "ptb_headless: giving up after %u frames (%llu of %.0f paths finished)\n", Frames + 1,
                         (unsigned long long)Stats.paths_finished, Target);
            return 1;
        }
        uint Batch = Frames < 64 ? 8 : 64;
        for (uint K = 0; K < Batch; K++) Update(&App, false);
        Frames += Batch;
    }
    RenderSampleBuffer(App.Device, App.SampleBuffer, &App.ResolveParameters);
    double Seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - T0).count();
    std::printf("{\"scene\": \"%s\", \"width\": %u, \"height\": %u, \"frames\": %u, \"rays_traced\": %llu, \"paths_finished\": %llu, "
                "\"seconds\": %.4f, \"mrays_per_s\": %.1f}\n", SceneSpec.c_str(), Width, Height, Frames + 1,
                (unsigned long long)Stats.rays_traced, (unsigned long long)Stats.paths_finished, Seconds, Stats.rays_traced / Seconds / 1e6);
    int Rc = 0;
    if (!Out.empty() && !WriteImagePNG(Out.c_str(), Width, Height, App.SampleBuffer->Frame.data())) { std::fprintf(stderr, "ptb_headless: cannot write %s\n", Out.c_str()); Rc = 1; }
    if (!FilmOut.empty()) {
        std::vector<float> Film((size_t)Width * Height * 4);
        FILE* F = nullptr;
        if (ptb_film_download_accum(App.SampleBuffer->Film, Film.data(), Film.size()) != PTB_OK || !(F = std::fopen(FilmOut.c_str(), "wb")) ||
            std::fwrite(Film.data(), sizeof(float), Film.size(), F) != Film.size()) { std::fprintf(stderr, "ptb_headless: cannot write %s\n", FilmOut.c_str()); Rc = 1; }
        if (F) std::fclose(F);
    }
    DestroyBasicRenderer(App.Device, App.BasicRenderer);
    DestroySampleBuffer(App.Device, App.SampleBuffer);
    DestroyDeviceScene(App.Device, App.DeviceScene);
    DestroyScene(App.Scene);
    DestroyDevice(App.Device);
    return Rc;
}
