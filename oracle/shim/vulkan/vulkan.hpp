// Stand-in for <vulkan/vulkan.hpp>: TEST INFRASTRUCTURE ONLY (oracle build).
// The reference's chunk_manager.hpp pulls in resources.hpp, which names a few Vulkan handle
// types in structs the voxel CPU path never touches. No Vulkan exists in this image, so this
// header supplies opaque placeholders with those names. It is our file, not a copy of any
// Vulkan header, and nothing in the product includes it.
#pragma once
#include <array>
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>
namespace vk {
using DeviceSize = uint64_t;
struct Buffer { void* h = nullptr; };
struct Image { void* h = nullptr; };
struct ImageView { void* h = nullptr; };
struct Sampler { void* h = nullptr; };
struct Semaphore { void* h = nullptr; };
struct Fence { void* h = nullptr; };
struct CommandPool { void* h = nullptr; };
struct CommandBuffer { void* h = nullptr; };
struct AccelerationStructureKHR { void* h = nullptr; };
enum class Format { eUndefined = 0 };
enum class SampleCountFlagBits { e1 = 1 };
enum class ImageLayout { eUndefined = 0 };
}  // namespace vk
