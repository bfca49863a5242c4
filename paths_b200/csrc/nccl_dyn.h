// nccl_dyn.h -- the handful of NCCL entry points the frame combine uses, bound at run time.
//
// libpaths_b200.so does not carry a DT_NEEDED entry for libnccl: a host process that already loaded an NCCL (PyTorch ships its own
// libnccl.so.2) must keep exactly one copy alive, and a single-GPU user needs none.  The first multi-GPU call binds the library that
// is already mapped into the process (RTLD_NOLOAD), else loads libnccl.so.2 from the loader path; failure is PATHS_ERR_NCCL.
#pragma once
#include <cuda_runtime_api.h>
#include <nccl.h>

#include <string>

namespace paths {

struct NcclApi {
    ncclResult_t (*GetVersion)(int *);
    ncclResult_t (*GetUniqueId)(ncclUniqueId *);
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char *(*GetErrorString)(ncclResult_t);
    std::string origin; // which file the symbols came from
};

// null + err when no NCCL can be bound
const NcclApi *nccl_api(std::string &err);

#define PATHS_NCCL(api, call)                                                                  \
    do {                                                                                       \
        ncclResult_t r_ = (api)->call;                                                         \
        if (r_ != ncclSuccess) {                                                               \
            err = std::string("nccl" #call ": ") + (api)->GetErrorString(r_);                  \
            return PATHS_ERR_NCCL;                                                             \
        }                                                                                      \
    } while (0)

} // namespace paths
