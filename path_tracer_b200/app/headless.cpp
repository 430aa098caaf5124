// ptb_headless: the reference application's per-frame loop (src/application/application.cpp:100-124) without a
// window - scene in, image file out - written against host/integrator.hpp, i.e. against the reference's own
// integrator interface.  Swapchain, ImGui and input are out of scope (no graphics stack on this box); the frame
// RenderSampleBuffer would present is written as a PNG, the film optionally as raw RGBA32F.
//
//   ptb_headless --scene cornell|terrain[:grid]|room[:detail[:atlas]]|spherefield[:copies[:level[:camera]]]|FILE.json
//                [--size WxH] [--spp N] [--max-frames F] [--camera I] [--tone clamp|reinhard|hable|aces] [--brightness B] [--white W]
//                [--device D | --devices D0,D1,...] [--builder sah-host|lbvh-gpu|ploc-gpu|auto] [--out frame.png] [--film film.rgba32f] [--table PATH]
// --devices renders ONE image on several GPUs of the box (sample split; host/integrator.hpp device_group): the scene is packed and
// flattened once and replicated over NVLink, every GPU renders its own seed block on its own host thread, and one fused
// peer-memory reduce + resolve delivers the combined frame.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <string>
#include <thread>
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
        size_t J = S.find_first_of(":,", I);
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

// ---- --devices: one image, several GPUs ----
struct multi_options {
    std::vector<int> CudaDevices;
    uint Width, Height, Spp, Camera, MaxFrames;
    uint32_t BuilderId; bool HasBuilder;
    resolve_parameters Resolve;
    std::string SceneSpec, Out, FilmOut;
};

static int RenderOnDevices(scene* Scene, multi_options const& O) {
    int const N = (int)O.CudaDevices.size();
    std::vector<device*> Devices;
    for (int D : O.CudaDevices) { device* Dev = CreateDevice(D); if (!Dev) return 1; Devices.push_back(Dev); }
    device_group* Group = CreateDeviceGroup(Devices);
    if (!Group) return 1;
    PackSceneData(Scene);
    struct rank_state { application App; ptb_stats Stats = {}; uint Frames = 0; double SetupSeconds = 0, RenderSeconds = 0; };
    std::vector<rank_state> Ranks(N);
    sample_buffer* Sum = nullptr;
    std::atomic<int> Failed{0};
    uint const FramesPerRank = O.MaxFrames + 64;
    double const TargetPerRank = std::ceil((double)O.Spp * O.Width * O.Height / N);
    auto T0 = std::chrono::steady_clock::now();
    auto Work = [&](int K) {
        rank_state& R = Ranks[K];
        application& App = R.App;
        App.Device = Devices[K];
        App.Scene = Scene;
        App.DeviceScene = CreateDeviceScene(App.Device);
        App.SampleBuffer = CreateSampleBuffer(App.Device, O.Width, O.Height);
        bool Ok = App.DeviceScene && App.SampleBuffer;
        if (Ok && K == 0) {
            if (O.HasBuilder && ptb_scene_set_builder(App.DeviceScene->Handle, O.BuilderId) != PTB_OK) Ok = false;
            if (Ok) UpdateDeviceScene(App.Device, App.DeviceScene, Scene, SCENE_DIRTY_ALL);
            if (!O.FilmOut.empty()) Sum = CreateSampleBuffer(App.Device, O.Width, O.Height);
        }
        // collectives are entered by every rank even after a local failure, so that nobody waits forever
        if (!App.DeviceScene || !ReplicateDeviceScene(Group, K, App.DeviceScene)) Ok = false;
        R.SetupSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - T0).count();
        if (Ok) App.BasicRenderer = CreateBasicRenderer(App.Device, App.DeviceScene, App.SampleBuffer);
        Ok = Ok && App.BasicRenderer;
        if (Ok) {
            App.CameraIndex = O.Camera;
            App.BasicRenderer->FrameIndex = FirstFrameIndex(Group, K, FramesPerRank);
            App.BasicRenderer->CameraIndex = O.Camera;
            App.BasicRenderer->Scene = App.DeviceScene;
            App.BasicRenderer->RenderFlags = RENDER_FLAG_ACCUMULATE | RENDER_FLAG_SAMPLE_JITTER;
            ResetBasicRenderer(App.Device, App.BasicRenderer);
            RunBasicRenderer(App.Device, App.BasicRenderer, 2);
            for (;;) {
                uint64_t const Before = R.Stats.rays_traced;
                if (ptb_renderer_stats(App.BasicRenderer->Handle, &R.Stats) != PTB_OK) { Ok = false; break; }
                if ((double)R.Stats.paths_finished >= TargetPerRank) break;
                if (R.Stats.rays_traced == Before || R.Frames >= O.MaxFrames) { Ok = false; break; }
                uint Batch = R.Frames < 64 ? 8 : 64;
                for (uint I = 0; I < Batch; I++) RunBasicRenderer(App.Device, App.BasicRenderer, 1);
                R.Frames += Batch;
            }
        }
        if (!Ok) Failed++;
        resolve_parameters P = O.Resolve;
        if (App.SampleBuffer && !RenderCombinedSampleBuffer(Group, K, App.SampleBuffer, &P, Sum)) Failed++;
        R.RenderSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - T0).count();
    };
    std::vector<std::thread> Threads;
    for (int K = 0; K < N; K++) Threads.emplace_back(Work, K);
    for (auto& T : Threads) T.join();
    double const Seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - T0).count();
    if (Failed) { std::fprintf(stderr, "ptb_headless: %d rank step(s) failed: %s\n", (int)Failed, ptb_last_error()); return 1; }
    unsigned long long Rays = 0, Paths = 0;
    uint Frames = 0;
    double Setup = 0;
    for (auto& R : Ranks) { Rays += R.Stats.rays_traced; Paths += R.Stats.paths_finished; Frames = std::max(Frames, R.Frames + 1); Setup = std::max(Setup, R.SetupSeconds); }
    ptb_reduce_times Reduce = {};
    ptb_group_last_reduce(Group->Handles[0], &Reduce);
    std::string ByDevice;
    for (auto& R : Ranks) ByDevice += (ByDevice.empty() ? "" : ", ") + std::to_string(R.Frames + 1);
    std::printf("{\"scene\": \"%s\", \"width\": %u, \"height\": %u, \"devices\": %d, \"frames_per_device\": %u, \"frames_by_device\": [%s], \"rays_traced\": %llu, \"paths_finished\": %llu, "
                "\"seconds\": %.4f, \"setup_seconds\": %.4f, \"exchange_ms\": %.4f, \"mrays_per_s\": %.1f}\n", O.SceneSpec.c_str(), O.Width, O.Height, N, Frames, ByDevice.c_str(),
                Rays, Paths, Seconds, Setup, Reduce.exchange_ms, Rays / std::max(1e-9, Seconds - Setup) / 1e6);
    int Rc = 0;
    sample_buffer* Root = Ranks[0].App.SampleBuffer;
    if (!O.Out.empty() && !WriteImagePNG(O.Out.c_str(), O.Width, O.Height, Root->Frame.data())) { std::fprintf(stderr, "ptb_headless: cannot write %s\n", O.Out.c_str()); Rc = 1; }
    if (!O.FilmOut.empty() && Sum) {
        std::vector<float> Film((size_t)O.Width * O.Height * 4);
        FILE* F = nullptr;
        if (ptb_film_download_accum(Sum->Film, Film.data(), Film.size()) != PTB_OK || !(F = std::fopen(O.FilmOut.c_str(), "wb")) ||
            std::fwrite(Film.data(), sizeof(float), Film.size(), F) != Film.size()) { std::fprintf(stderr, "ptb_headless: cannot write %s\n", O.FilmOut.c_str()); Rc = 1; }
        if (F) std::fclose(F);
    }
    for (auto& R : Ranks) {
        DestroyBasicRenderer(R.App.Device, R.App.BasicRenderer);
        DestroySampleBuffer(R.App.Device, R.App.SampleBuffer);
        DestroyDeviceScene(R.App.Device, R.App.DeviceScene);
    }
    if (Sum) DestroySampleBuffer(Devices[0], Sum);
    DestroyDeviceGroup(Group);
    for (device* D : Devices) DestroyDevice(D);
    return Rc;
}

int main(int argc, char** argv) {
    std::string SceneSpec = "cornell", Out, FilmOut, Tone = "aces", Builder, Table;
    uint Width = 640, Height = 480, Spp = 64, Camera = 0, MaxFrames = 1u << 20;
    int CudaDevice = 0;
    std::vector<int> CudaDevices;
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
        else if (A == "--devices") { for (double D : SplitNumbers(std::string(Next()).c_str())) CudaDevices.push_back((int)D); }
        else if (A == "--builder") Builder = Next();
        else if (A == "--out") Out = Next();
        else if (A == "--film") FilmOut = Next();
        else if (A == "--table") Table = Next();
        else if (A == "--help" || A == "-h") {
            std::printf("usage: ptb_headless --scene NAME[:params]|FILE.json [--size WxH] [--spp N] [--max-frames F] [--camera I] [--tone clamp|reinhard|hable|aces]\n"
                        "                    [--brightness B] [--white W] [--device D | --devices D0,D1,...] [--builder sah-host|lbvh-gpu|ploc-gpu|auto]\n"
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

    if (!CudaDevices.empty()) CudaDevice = CudaDevices[0];
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
    if (CudaDevices.size() > 1) {
        DestroyDevice(App.Device);                // (it was only needed to fit the spectrum table)
        multi_options O;
        O.CudaDevices = CudaDevices; O.Width = Width; O.Height = Height; O.Spp = Spp; O.Camera = Camera; O.MaxFrames = MaxFrames;
        O.BuilderId = BuilderId; O.HasBuilder = !Builder.empty();
        O.Resolve.Brightness = Brightness; O.Resolve.ToneMappingMode = Mode; O.Resolve.ToneMappingWhiteLevel = White;
        O.SceneSpec = SceneSpec; O.Out = Out; O.FilmOut = FilmOut;
        int Rc = RenderOnDevices(App.Scene, O);
        DestroyScene(App.Scene);
        return Rc;
    }
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
