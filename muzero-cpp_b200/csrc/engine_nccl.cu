// engine_nccl.cu — weight synchronisation across the GPUs of a box: one ncclBroadcast of the packed weight blob.
// Replaces the reference's disk round trip (learner.cpp:93-101: SaveCheckpoint on device 0, LoadCheckpoint on every
// other device; same at start-up, muzero.cpp:142-148).  Trees never cross GPUs, so this is the only collective.
// NCCL is resolved at run time (dlopen) so the library neither links libnccl nor needs torch's copy at build time;
// in a process that already imported torch the bundled libnccl.so.2 is the one found.
#include <dlfcn.h>

#include <cstring>

#include "engine.h"

using namespace mzb;

namespace {
typedef struct { char internal[128]; } nccl_uid;
typedef int (*fn_get_uid)(nccl_uid*);
typedef int (*fn_init_rank)(void** comm, int nranks, nccl_uid id, int rank);
typedef int (*fn_bcast)(const void* send, void* recv, size_t count, int dtype, int root, void* comm, cudaStream_t s);
typedef int (*fn_destroy)(void* comm);
typedef const char* (*fn_errstr)(int);

struct NcclApi {
    void* lib = nullptr;
    fn_get_uid get_uid = nullptr;
    fn_init_rank init_rank = nullptr;
    fn_bcast bcast = nullptr;
    fn_destroy destroy = nullptr;
    fn_errstr errstr = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
    if (g_nccl.lib) return MZB_OK;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) { set_error("cannot dlopen libnccl.so.2: %s", dlerror()); return MZB_ERR_NCCL; }
    g_nccl.get_uid = (fn_get_uid)dlsym(g_nccl.lib, "ncclGetUniqueId");
    g_nccl.init_rank = (fn_init_rank)dlsym(g_nccl.lib, "ncclCommInitRank");
    g_nccl.bcast = (fn_bcast)dlsym(g_nccl.lib, "ncclBroadcast");
    g_nccl.destroy = (fn_destroy)dlsym(g_nccl.lib, "ncclCommDestroy");
    g_nccl.errstr = (fn_errstr)dlsym(g_nccl.lib, "ncclGetErrorString");
    if (!g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.bcast) { set_error("libnccl lacks expected symbols"); return MZB_ERR_NCCL; }
    return MZB_OK;
}
int nccl_fail(const char* what, int code) {
    set_error("%s failed: %s", what, g_nccl.errstr ? g_nccl.errstr(code) : "nccl error");
    return MZB_ERR_NCCL;
}
}  // namespace

extern "C" {

int mzb_nccl_unique_id(char id_out[128]) {
    MZB_REQUIRE(id_out, "null argument");
    int rc = load_nccl();
    if (rc) return rc;
    nccl_uid id;
    int r = g_nccl.get_uid(&id);
    if (r != 0) return nccl_fail("ncclGetUniqueId", r);
    memcpy(id_out, id.internal, 128);
    return MZB_OK;
}

int mzb_nccl_init(mzb_engine* e, const char id[128], int rank, int world_size) {
    MZB_REQUIRE(e && id && rank >= 0 && rank < world_size, "bad arguments");
    int rc = load_nccl();
    if (rc) return rc;
    MZB_CUDA(cudaSetDevice(e->device));
    nccl_uid uid;
    memcpy(uid.internal, id, 128);
    int r = g_nccl.init_rank(&e->nccl_comm, world_size, uid, rank);
    if (r != 0) return nccl_fail("ncclCommInitRank", r);
    return MZB_OK;
}

// Two-phase hand-over ("into a shadow buffer, swapped between moves").  Phase 1 receives the packed blob into a SHADOW
// buffer on a stream of its own: a search in flight on the engine's stream keeps reading the live weights and is not waited for —
// the learner's broadcast overlaps self-play.  Phase 2, mzb_weights_commit, is enqueued on the SEARCH stream: it waits for the
// receive and moves shadow -> live with one device-to-device copy (25 MB for C4d: ~10 us), ordered between two searches by the
// stream itself, so no search ever sees a mixture of old and new tensors.  (The engine's kernels hold pointers into the live blob;
// a copy in stream order gives the semantics of the pointer swap without re-deriving them.)  The root only sends.
int mzb_broadcast_weights_begin(mzb_engine* e, int root, int rank) {
    MZB_REQUIRE(e && e->net && e->weights_loaded, "no weights loaded");
    MZB_REQUIRE(e->nccl_comm, "mzb_nccl_init was not called");
    MZB_CUDA(cudaSetDevice(e->device));
    if (!e->sync_stream) {
        MZB_CUDA(cudaStreamCreateWithFlags(&e->sync_stream, cudaStreamNonBlocking));
        MZB_CUDA(cudaEventCreateWithFlags(&e->ev_sync, cudaEventDisableTiming));
    }
    const size_t bytes = e->net->blob_bytes();
    const bool is_root = rank == root;
    if (!is_root && e->shadow_blob.bytes != bytes) MZB_CUDA(e->shadow_blob.alloc(bytes));
    void* recv = is_root ? e->net->blob_ptr() : e->shadow_blob.p;
    // ncclChar = 0 / ncclUint8 = 1: bytes
    int r = g_nccl.bcast(e->net->blob_ptr(), recv, bytes, 1, root, e->nccl_comm, e->sync_stream);
    if (r != 0) return nccl_fail("ncclBroadcast", r);
    MZB_CUDA(cudaEventRecord(e->ev_sync, e->sync_stream));
    e->sync_pending = is_root ? 1 : 2;
    return MZB_OK;
}

int mzb_weights_commit(mzb_engine* e) {
    MZB_REQUIRE(e && e->net, "null argument");
    if (!e->sync_pending) return MZB_OK;
    MZB_CUDA(cudaSetDevice(e->device));
    MZB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_sync, 0));
    if (e->sync_pending == 2)
        MZB_CUDA(cudaMemcpyAsync(e->net->blob_ptr(), e->shadow_blob.p, e->net->blob_bytes(), cudaMemcpyDeviceToDevice, e->stream));
    e->sync_pending = 0;
    return MZB_OK;
}

// The blocking form: receive in place on the search stream (after whatever search is queued there) and wait.
int mzb_broadcast_weights(mzb_engine* e, int root) {
    MZB_REQUIRE(e && e->net && e->weights_loaded, "no weights loaded");
    MZB_REQUIRE(e->nccl_comm, "mzb_nccl_init was not called");
    MZB_CUDA(cudaSetDevice(e->device));
    int r = g_nccl.bcast(e->net->blob_ptr(), e->net->blob_ptr(), e->net->blob_bytes(), 1, root, e->nccl_comm, e->stream);
    if (r != 0) return nccl_fail("ncclBroadcast", r);
    MZB_CUDA(cudaStreamSynchronize(e->stream));
    return MZB_OK;
}

// engine teardown: the communicator (mzb_engine_destroy)
void mzb_nccl_release(mzb_engine* e) {
    if (e && e->nccl_comm && g_nccl.destroy) { g_nccl.destroy(e->nccl_comm); e->nccl_comm = nullptr; }
}

}  // extern "C"
