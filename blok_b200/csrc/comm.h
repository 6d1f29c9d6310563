// NCCL reached through dlopen: the library has no link-time dependency on NCCL (a single-GPU host never needs it), and a process
// that already carries an NCCL — its own, or the one bundled with torch — gets THAT one (same soname), so a caller-owned
// ncclComm_t handed to blokb_set_comm and the entry points used on it come from the same library.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <string>

namespace blokb {

using NcclComm = void*;  // ncclComm_t (opaque)
struct NcclId { char internal[128]; };  // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128)

struct NcclApi {
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int nranks, NcclId id, int rank) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*CommCount)(NcclComm, int*) = nullptr;
    int (*CommUserRank)(NcclComm, int*) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int dtype, int op, NcclComm, cudaStream_t) = nullptr;
    int (*ReduceScatter)(const void*, void*, size_t recvcount, int dtype, int op, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t sendcount, int dtype, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclFloat32 = 7, kNcclSum = 0;  // ncclDataType_t / ncclRedOp_t values (nccl.h, stable across 2.x)

// nullptr + `err` filled when no usable libnccl is found.
const NcclApi* nccl_api(std::string& err);

}  // namespace blokb
