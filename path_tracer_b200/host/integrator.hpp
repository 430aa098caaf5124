// Host-side integrator interface in C++ over the C ABI (include/ptb.h): the reference's seven
// integrator functions and three scene-mirror functions under their own names, argument order and
// error behaviour, so that the reference's application code (src/application/application.cpp:100-124,
// :352-362, :403-407) compiles against this header with `vulkan*` read as `device*`.
//
//   reference (file:line)                                              here
//   CreateVulkan / DestroyVulkan          core/vulkan.cpp              CreateDevice / DestroyDevice
//   CreateVulkanScene                     scene/scene.cpp:1643         CreateDeviceScene
//   UpdateVulkanScene                     scene/scene.cpp:1698-1960    UpdateDeviceScene
//   DestroyVulkanScene                    scene/scene.cpp:1963-2006    DestroyDeviceScene
//   CreateSampleBuffer                    integrator/integrator.cpp:15   same name
//   DestroySampleBuffer                   integrator/integrator.cpp:100  same name
//   RenderSampleBuffer                    integrator/integrator.cpp:110  same name (resolves into SampleBuffer->Frame)
//   CreateBasicRenderer                   integrator/basic.cpp:166     same name
//   DestroyBasicRenderer                  integrator/basic.cpp:258     same name
//   ResetBasicRenderer                    integrator/basic.cpp:285     same name
//   RunBasicRenderer                      integrator/basic.cpp:306     same name
//
// Error behaviour is the reference's: Create* return nullptr, the others return void, and the
// message goes to stderr (Errorf, core/vulkan.cpp:17-24).  There is no CPU fallback: without
// libptb.so's CUDA device CreateDevice fails.
#pragma once
#include <cstdint>
#include <vector>

#include "ptb.h"
#include "scene.hpp"

namespace ptb { namespace host {

typedef unsigned int uint;

enum render_flag : uint {                         // integrator.hpp:6-10
    RENDER_FLAG_ACCUMULATE    = 1 << 0,
    RENDER_FLAG_SAMPLE_JITTER = 1 << 1,
};

enum tone_mapping_mode : uint {                   // integrator.hpp:12-19
    TONE_MAPPING_MODE_CLAMP    = 0,
    TONE_MAPPING_MODE_REINHARD = 1,
    TONE_MAPPING_MODE_HABLE    = 2,
    TONE_MAPPING_MODE_ACES     = 3,
    TONE_MAPPING_MODE__COUNT   = 4,
};
char const* ToneMappingModeName(tone_mapping_mode Mode);

struct device {                                   // stands where `vulkan` stands in the reference's signatures
    ptb_context* Context = nullptr;
    int CudaDevice = 0;
};

struct device_scene {                             // vulkan_scene
    ptb_scene* Handle = nullptr;
};

struct sample_buffer {                            // vulkan_sample_buffer
    ptb_film* Film = nullptr;
    uint Width = 0, Height = 0;
    std::vector<uint8_t> Frame;                   // what RenderSampleBuffer presents: Width x Height sRGB8 RGBA, row 0 first
};

struct resolve_parameters {                       // integrator.hpp:43-48
    float             Brightness = 1.0f;
    tone_mapping_mode ToneMappingMode = TONE_MAPPING_MODE_CLAMP;
    float             ToneMappingWhiteLevel = 1.0f;
};

struct basic_renderer {                           // basic.hpp:6-26: the public fields the application pokes
    sample_buffer* SampleBuffer = nullptr;
    ptb_renderer*  Handle = nullptr;
    uint  FrameIndex = 0;
    uint  CameraIndex = 0;
    device_scene* Scene = nullptr;
    uint  RenderFlags = 0;
    uint  PathLengthLimit = 0;
    float PathTerminationProbability = 0.0f;
};

device* CreateDevice(int CudaDevice = 0);
void DestroyDevice(device* Device);

device_scene* CreateDeviceScene(device* Device);
void UpdateDeviceScene(device* Device, device_scene* DeviceScene, scene* Scene, uint32_t DirtyFlags);
void DestroyDeviceScene(device* Device, device_scene* DeviceScene);

sample_buffer* CreateSampleBuffer(device* Device, uint Width, uint Height);
void DestroySampleBuffer(device* Device, sample_buffer* SampleBuffer);
void RenderSampleBuffer(device* Device, sample_buffer* SampleBuffer, resolve_parameters* Parameters);

basic_renderer* CreateBasicRenderer(device* Device, device_scene* Scene, sample_buffer* SampleBuffer);
void DestroyBasicRenderer(device* Device, basic_renderer* Renderer);
void ResetBasicRenderer(device* Device, basic_renderer* Renderer);
void RunBasicRenderer(device* Device, basic_renderer* Renderer, uint Rounds);

// ---- several GPUs of one box (no reference counterpart: it binds one VkPhysicalDevice, core/vulkan.cpp:1485-1506) ----
// One process drives every device: ONE HOST THREAD PER DEVICE makes that device's calls (the two functions below are
// collective: they return when every rank has joined).  The scene is packed and flattened once, on rank 0, and sent
// device-to-device; every rank renders its own block of RandomSeed values into its own sample buffer; one exchange
// step sums the films and resolves the combined frame into rank 0's SampleBuffer->Frame.
struct device_group {
    std::vector<device*>    Devices;              // rank k renders on Devices[k]
    std::vector<ptb_group*> Handles;
};
device_group* CreateDeviceGroup(std::vector<device*> const& Devices);
void DestroyDeviceGroup(device_group* Group);
// First FrameIndex of `Rank` when every rank renders at most FramesPerRank frames after its reset (disjoint seed blocks).
uint FirstFrameIndex(device_group* Group, int Rank, uint FramesPerRank);
// Rank 0's device scene (after UpdateDeviceScene) to every other rank's freshly created one.  Collective.
bool ReplicateDeviceScene(device_group* Group, int Rank, device_scene* DeviceScene);
// RenderSampleBuffer on the SUM of all ranks' sample buffers: the frame lands in rank 0's SampleBuffer->Frame; Sum
// (rank 0 only, optional) also receives the float film.  Collective.
bool RenderCombinedSampleBuffer(device_group* Group, int Rank, sample_buffer* SampleBuffer, resolve_parameters* Parameters, sample_buffer* Sum = nullptr);

}}  // namespace ptb::host
