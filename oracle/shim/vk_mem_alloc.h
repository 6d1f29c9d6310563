// Stand-in for <vk_mem_alloc.h>: TEST INFRASTRUCTURE ONLY (oracle build). See shim/vulkan/vulkan.hpp.
#pragma once
typedef struct VmaAllocation_T* VmaAllocation;
typedef struct VmaAllocator_T* VmaAllocator;
