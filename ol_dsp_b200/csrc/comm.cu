// comm.cu -- (e2) the one exchange step of the hot path: the cross-GPU voice bus.
//
// Polyvoice::Process adds every voice into one frame (modules/synthlib/Polyvoice.h:28-33; the stereo duplicate is the
// caller's, modules/ol_daisy/app/synth/main.cpp:83).  With the voices sharded over the GPUs of a box (one process per
// GPU), every rank's k_voice + k_bus_reduce leave a PARTIAL bus [frames][channels] in device memory; one in-place sum
// all-reduce over NVLink / NVSwitch turns it into the global mixdown on every rank.  The partial bus IS the send buffer
// (no staging copy).  The asynchronous form runs the reduce of block k on the communicator's own stream while the
// caller's stream already computes block k + 1 into the other bus buffer.
//
// NCCL is the transport (ncclAllReduce, fp32 sum); this file is the only place that names it.  The id is created on
// rank 0 and carried to the other ranks by whatever the host has (bench.py: torch.distributed; a C++ host: its own
// launcher) -- the library never opens a socket of its own.
//
// libnccl.so.2 is bound at the FIRST olb_comm_* call (dlopen), not at load time: a process has room for one NCCL, and a
// host application that brings its own (PyTorch ships a newer one than the system's) must get to load it first, whatever
// the order in which it and libolbank.so were loaded.
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

#include "common.cuh"

namespace olb {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

static const NcclApi &nccl() {
    static NcclApi api = [] {
        NcclApi a;
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return a;
        a.GetUniqueId = (decltype(a.GetUniqueId)) dlsym(h, "ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank)) dlsym(h, "ncclCommInitRank");
        a.CommDestroy = (decltype(a.CommDestroy)) dlsym(h, "ncclCommDestroy");
        a.AllReduce = (decltype(a.AllReduce)) dlsym(h, "ncclAllReduce");
        a.GetErrorString = (decltype(a.GetErrorString)) dlsym(h, "ncclGetErrorString");
        a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.GetErrorString;
        return a;
    }();
    return api;
}

static int need_nccl() {
    if (nccl().ok) return OLB_OK;
    set_error("libnccl.so.2 could not be loaded (%s): the cross-GPU voice bus needs NCCL", dlerror() ? dlerror() : "symbols missing");
    return OLB_E_NCCL;
}

static int fail_nccl(ncclResult_t r, const char *what, const char *file, int line) {
    set_error("NCCL error %d (%s) in %s at %s:%d", (int) r, nccl().GetErrorString(r), what, file, line);
    return OLB_E_NCCL;
}
#define OLB_NCCL(call)                                                          \
    do {                                                                        \
        ncclResult_t _r = (call);                                               \
        if (_r != ncclSuccess) return olb::fail_nccl(_r, #call, __FILE__, __LINE__); \
    } while (0)

constexpr int kSlots = 4;

}  // namespace olb

using namespace olb;

struct olb_comm {
    std::mutex mu;
    int device = 0, rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    cudaStream_t st = nullptr;            // the reduces of the asynchronous form run here
    cudaEvent_t produced = nullptr;       // "the partial bus is complete" (recorded on the producer's stream)
    cudaEvent_t done[kSlots] = {};        // "the reduce of this bus slot is complete"
    bool used[kSlots] = {};
};

static_assert(OLB_COMM_ID_BYTES >= sizeof(ncclUniqueId), "OLB_COMM_ID_BYTES must hold an ncclUniqueId");

OLB_API int olb_comm_unique_id(void *id_out) {
    OLB_ARG(id_out, "olb_comm_unique_id: id_out is NULL");
    int rc = need_nccl();
    if (rc) return rc;
    ncclUniqueId id;
    OLB_NCCL(nccl().GetUniqueId(&id));
    memset(id_out, 0, OLB_COMM_ID_BYTES);
    memcpy(id_out, &id, sizeof(id));
    return OLB_OK;
}

OLB_API int olb_comm_init(int device, int rank, int world, const void *id_bytes, olb_comm **out) {
    OLB_ARG(out, "olb_comm_init: out is NULL");
    *out = nullptr;
    OLB_ARG(id_bytes, "olb_comm_init: id is NULL");
    OLB_ARG(world >= 1 && rank >= 0 && rank < world, "olb_comm_init: rank %d of %d", rank, world);
    int rc = need_nccl();
    if (rc) return rc;
    rc = require_sm100(device);
    if (rc) return rc;
    DeviceGuard g(device);
    olb_comm *c = new olb_comm();
    c->device = device;
    c->rank = rank;
    c->world = world;
    ncclUniqueId id;
    memcpy(&id, id_bytes, sizeof(id));
    ncclResult_t r = nccl().CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) {
        delete c;
        return fail_nccl(r, "ncclCommInitRank", __FILE__, __LINE__);
    }
    cudaError_t e = cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->produced, cudaEventDisableTiming);
    for (int i = 0; i < kSlots && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&c->done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) {
        olb_comm_destroy(c);
        return fail_cuda(e, "creating the communicator's stream and events", __FILE__, __LINE__);
    }
    *out = c;
    return OLB_OK;
}

OLB_API int olb_comm_destroy(olb_comm *c) {
    if (!c) return OLB_OK;
    DeviceGuard g(c->device);
    if (c->st) cudaStreamSynchronize(c->st);
    if (c->comm) nccl().CommDestroy(c->comm);
    if (c->produced) cudaEventDestroy(c->produced);
    for (cudaEvent_t e : c->done)
        if (e) cudaEventDestroy(e);
    if (c->st) cudaStreamDestroy(c->st);
    delete c;
    return OLB_OK;
}

OLB_API int olb_comm_info(olb_comm *c, int *rank, int *world) {
    OLB_ARG(c, "olb_comm_info: NULL handle");
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    return OLB_OK;
}

// In-place sum of the partial buses of all ranks, on the caller's stream (ordered after what produced the bus).
OLB_API int olb_voice_bus_allreduce(olb_comm *c, float *bus_dev, uint64_t frames, uint32_t channels, void *stream) {
    OLB_ARG(c && bus_dev, "olb_voice_bus_allreduce: NULL argument");
    OLB_ARG(channels >= 1 && channels <= 8, "olb_voice_bus_allreduce: channels must be 1..8");
    if (frames == 0) return OLB_OK;
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    OLB_NCCL(nccl().AllReduce(bus_dev, bus_dev, (size_t) frames * channels, ncclFloat, ncclSum, c->comm, (cudaStream_t) stream));
    return OLB_OK;
}

// Same, overlapped: the reduce waits (on the device) for what `producer_stream` has queued so far and then runs on the
// communicator's stream, so the producer can go straight on with the next block into ANOTHER bus buffer.  `slot` names
// the buffer (0..3) for olb_comm_wait.
OLB_API int olb_voice_bus_allreduce_async(olb_comm *c, float *bus_dev, uint64_t frames, uint32_t channels, void *producer_stream, int slot) {
    OLB_ARG(c && bus_dev, "olb_voice_bus_allreduce_async: NULL argument");
    OLB_ARG(channels >= 1 && channels <= 8, "olb_voice_bus_allreduce_async: channels must be 1..8");
    OLB_ARG(slot >= 0 && slot < kSlots, "olb_voice_bus_allreduce_async: slot must be 0..%d", kSlots - 1);
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    OLB_CUDA(cudaEventRecord(c->produced, (cudaStream_t) producer_stream));
    OLB_CUDA(cudaStreamWaitEvent(c->st, c->produced, 0));
    if (frames) OLB_NCCL(nccl().AllReduce(bus_dev, bus_dev, (size_t) frames * channels, ncclFloat, ncclSum, c->comm, c->st));
    OLB_CUDA(cudaEventRecord(c->done[slot], c->st));
    c->used[slot] = true;
    return OLB_OK;
}

// `stream` waits (on the device) until the last reduce queued for `slot` is complete: call it before reading the
// reduced bus, and before the bank writes that buffer again.
OLB_API int olb_comm_wait(olb_comm *c, int slot, void *stream) {
    OLB_ARG(c, "olb_comm_wait: NULL handle");
    OLB_ARG(slot >= 0 && slot < kSlots, "olb_comm_wait: slot must be 0..%d", kSlots - 1);
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard g(c->device);
    if (c->used[slot]) OLB_CUDA(cudaStreamWaitEvent((cudaStream_t) stream, c->done[slot], 0));
    return OLB_OK;
}

// The host waits for every reduce queued on the communicator's stream.
OLB_API int olb_comm_sync(olb_comm *c) {
    OLB_ARG(c, "olb_comm_sync: NULL handle");
    DeviceGuard g(c->device);
    OLB_CUDA(cudaStreamSynchronize(c->st));
    return OLB_OK;
}
