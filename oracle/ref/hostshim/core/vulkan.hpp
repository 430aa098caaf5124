// ORACLE - test infrastructure.  Stand-in for the reference's core/vulkan.hpp (Vulkan + GLFW, absent here): only the
// names that scene.hpp's declarations mention.  The Vulkan half of scene.cpp (from its "--- Vulkan ---" banner on) is
// not compiled (oracle/ref/hostgen.py cuts the generated copy there), so none of these is ever used.
#pragma once
#include <cassert>
#include <cstdint>
typedef struct VkDescriptorSetLayout_T* VkDescriptorSetLayout;
typedef struct VkDescriptorSet_T* VkDescriptorSet;
#ifndef VK_NULL_HANDLE
#define VK_NULL_HANDLE nullptr
#endif
struct vulkan;
struct vulkan_buffer { uint64_t Size = 0; };
struct vulkan_image {};
