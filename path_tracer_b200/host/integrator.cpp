// See integrator.hpp.  Every body is a call into the C ABI; nothing is computed here.
#include "integrator.hpp"

#include <cstdarg>
#include <cstdio>

namespace ptb { namespace host {

static void Errorf(char const* What) {            // core/vulkan.cpp:17-24: message to stderr, caller goes on
    std::fprintf(stderr, "ptb: %s: %s\n", What, ptb_last_error());
}

char const* ToneMappingModeName(tone_mapping_mode Mode) {
    switch (Mode) {
        case TONE_MAPPING_MODE_CLAMP:    return "Clamp";
        case TONE_MAPPING_MODE_REINHARD: return "Reinhard";
        case TONE_MAPPING_MODE_HABLE:    return "Hable";
        case TONE_MAPPING_MODE_ACES:     return "ACES";
        default:                         return nullptr;
    }
}

device* CreateDevice(int CudaDevice) {
    device* D = new device;
    D->CudaDevice = CudaDevice;
    if (ptb_context_create(CudaDevice, &D->Context) != PTB_OK) { Errorf("CreateDevice"); delete D; return nullptr; }
    return D;
}

void DestroyDevice(device* Device) {
    if (!Device) return;
    ptb_context_destroy(Device->Context);
    delete Device;
}

device_scene* CreateDeviceScene(device* Device) {
    device_scene* S = new device_scene;
    if (ptb_scene_create(Device->Context, &S->Handle) != PTB_OK) { Errorf("CreateDeviceScene"); delete S; return nullptr; }
    return S;
}

void UpdateDeviceScene(device*, device_scene* DeviceScene, scene* Scene, uint32_t DirtyFlags) {
    ptb_scene_view View = Scene->View();
    if (ptb_scene_update(DeviceScene->Handle, &View, DirtyFlags) != PTB_OK) Errorf("UpdateDeviceScene");
}

void DestroyDeviceScene(device*, device_scene* DeviceScene) {
    if (!DeviceScene) return;
    ptb_scene_destroy(DeviceScene->Handle);
    delete DeviceScene;
}

sample_buffer* CreateSampleBuffer(device* Device, uint Width, uint Height) {
    sample_buffer* B = new sample_buffer;
    B->Width = Width; B->Height = Height;
    if (ptb_film_create(Device->Context, Width, Height, nullptr, &B->Film) != PTB_OK) { Errorf("CreateSampleBuffer"); delete B; return nullptr; }
    B->Frame.assign((size_t)Width * Height * 4, 0);
    return B;
}

void DestroySampleBuffer(device*, sample_buffer* SampleBuffer) {
    if (!SampleBuffer) return;
    ptb_film_destroy(SampleBuffer->Film);
    delete SampleBuffer;
}

void RenderSampleBuffer(device*, sample_buffer* SampleBuffer, resolve_parameters* Parameters) {
    ptb_resolve_params P = {Parameters->Brightness, (uint32_t)Parameters->ToneMappingMode, Parameters->ToneMappingWhiteLevel};
    if (ptb_film_resolve(SampleBuffer->Film, &P, SampleBuffer->Frame.data(), nullptr) != PTB_OK) Errorf("RenderSampleBuffer");
}

basic_renderer* CreateBasicRenderer(device* Device, device_scene* Scene, sample_buffer* SampleBuffer) {
    basic_renderer* R = new basic_renderer;
    R->Scene = Scene; R->SampleBuffer = SampleBuffer;
    if (ptb_renderer_create(Device->Context, Scene->Handle, SampleBuffer->Film, &R->Handle) != PTB_OK) { Errorf("CreateBasicRenderer"); delete R; return nullptr; }
    return R;
}

void DestroyBasicRenderer(device*, basic_renderer* Renderer) {
    if (!Renderer) return;
    ptb_renderer_destroy(Renderer->Handle);
    delete Renderer;
}

// The public fields travel with every call, as the push constants do in the reference (basic.cpp:294-301, :318-325).
static bool PushParameters(basic_renderer* R) {
    ptb_render_params P = {};
    P.CameraIndex = R->CameraIndex; P.RenderFlags = R->RenderFlags; P.PathLengthLimit = R->PathLengthLimit;
    P.PathTerminationProbability = R->PathTerminationProbability;
    P.Integrator = PTB_INTEGRATOR_WAVEFRONT; P.Traversal = PTB_TRAVERSAL_WIDE; P.TieMode = PTB_TIE_CANONICAL;
    P.TraceVariant = PTB_TRACE_VARIANT_DEFAULT; P.ScatterMode = PTB_SCATTER_AUTO;
    return ptb_renderer_set_params(R->Handle, &P) == PTB_OK && ptb_renderer_set_frame_index(R->Handle, R->FrameIndex) == PTB_OK;
}

void ResetBasicRenderer(device*, basic_renderer* Renderer) {
    if (!PushParameters(Renderer) || ptb_renderer_reset(Renderer->Handle) != PTB_OK) Errorf("ResetBasicRenderer");
}

void RunBasicRenderer(device*, basic_renderer* Renderer, uint Rounds) {
    if (!PushParameters(Renderer) || ptb_renderer_run(Renderer->Handle, Rounds) != PTB_OK) { Errorf("RunBasicRenderer"); return; }
    Renderer->FrameIndex += 1;                    // basic.cpp:315
}

device_group* CreateDeviceGroup(std::vector<device*> const& Devices) {
    device_group* G = new device_group;
    G->Devices = Devices;
    G->Handles.assign(Devices.size(), nullptr);
    std::vector<ptb_context*> Contexts;
    for (device* D : Devices) Contexts.push_back(D->Context);
    if (ptb_group_create_local(Contexts.data(), (int)Contexts.size(), G->Handles.data()) != PTB_OK) { Errorf("CreateDeviceGroup"); delete G; return nullptr; }
    return G;
}

void DestroyDeviceGroup(device_group* Group) {
    if (!Group) return;
    for (ptb_group* H : Group->Handles) ptb_group_destroy(H);
    delete Group;
}

uint FirstFrameIndex(device_group* Group, int Rank, uint FramesPerRank) {
    return ptb_group_first_frame(Rank, (int)Group->Handles.size(), FramesPerRank, 0);
}

bool ReplicateDeviceScene(device_group* Group, int Rank, device_scene* DeviceScene) {
    if (ptb_scene_replicate(DeviceScene->Handle, Group->Handles[Rank], 0) != PTB_OK) { Errorf("ReplicateDeviceScene"); return false; }
    return true;
}

bool RenderCombinedSampleBuffer(device_group* Group, int Rank, sample_buffer* SampleBuffer, resolve_parameters* Parameters, sample_buffer* Sum) {
    ptb_resolve_params P = {Parameters->Brightness, (uint32_t)Parameters->ToneMappingMode, Parameters->ToneMappingWhiteLevel};
    if (ptb_film_reduce_resolve(SampleBuffer->Film, Group->Handles[Rank], 0, PTB_REDUCE_P2P_FUSED, &P, (Rank == 0 && Sum) ? Sum->Film : nullptr,
                                Rank == 0 ? SampleBuffer->Frame.data() : nullptr) != PTB_OK) { Errorf("RenderCombinedSampleBuffer"); return false; }
    return true;
}

}}  // namespace ptb::host
