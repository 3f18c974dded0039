// net_backend.h — device-side executor interface of the VPRNet towers.
// Two implementations: Fp32Backend (net_fp32.cu, check mode) and Bf16Backend (net_bf16.cu, tcgen05).
// Hidden states never leave HBM: both read the parent state from, and write the child state into, a
// pool of fixed-size slots ("node pool"), addressed by slot index arrays living on the device.
#pragma once
#include "common.h"
#include "net_plan.h"

namespace mzb {

// where the heads leave their results (device SoA, fp64 like VPRNetModel::InferenceOutputs, vprnet.h:64-69)
struct NetOutputs {
    double* value = nullptr;   // [B]
    double* reward = nullptr;  // [B]
    double* policy = nullptr;  // [B][A]
};

class NetBackend {
public:
    virtual ~NetBackend() {}
    virtual const char* name() const = 0;
    // upload folded weights; (re)callable for weight sync
    virtual int upload(const NetPlan& plan) = 0;
    // packed device weight blob (one contiguous allocation) for NCCL broadcast / checksum
    virtual void* blob_ptr() = 0;
    virtual size_t blob_bytes() const = 0;
    // node pool
    virtual int alloc_pool(size_t n_slots, int max_batch) = 0;
    virtual size_t pool_bytes() const = 0;
    virtual size_t slot_bytes() const = 0;
    // NCHW fp32 rows <-> pool slots (C-ABI compatibility path)
    virtual int import_hidden(const float* d_nchw, int B, const int* d_slot, cudaStream_t s) = 0;
    virtual int export_hidden(float* d_nchw, int B, const int* d_slot, cudaStream_t s) = 0;
    // VPRNetModel::InitialInference (vprnet.cpp:77-129): obs rows NCHW fp32 on device -> hidden into out slots
    virtual int initial(const float* d_obs, int B, const int* d_out_slot, const NetOutputs& out, cudaStream_t s) = 0;
    // VPRNetModel::RecurrentInference (vprnet.cpp:131-189): parent slots + actions -> child slots
    virtual int recurrent(int B, const int* d_in_slot, const int* d_action, const int* d_out_slot,
                          const NetOutputs& out, cudaStream_t s) = 0;
    // kernels launched by one initial()/recurrent() call (for gpu_launches accounting)
    // true when two recurrent() calls on disjoint tree ranges may be in flight on different streams at once (no shared
    // scratch): lets the search pipeline two half-cohorts so that one half's tree step hides behind the other's towers
    virtual bool recurrent_concurrent() const { return false; }
    // hint: the following recurrent() calls are parts of a pipelined loop (the GPU always has another part's CTAs to run), so
    // tiles should be as full as possible rather than balanced over the SMs of one launch
    virtual void set_pipelined(bool) {}
    // cut points t0[0..K] of a cohort of T trees into the K parts of a pipelined loop.  Default: equal parts.  A backend that
    // knows its tile geometry cuts so that every part's launch is whole waves of tiles over the SMs (net_bf16.cu)
    virtual void pipeline_cut(int T, int K, int* t0) const { for (int k = 0; k <= K; ++k) t0[k] = (int)((long)T * k / K); }
    // pipelined loop, alternating parts: the NEXT recurrent() call waits for `wait_before` (may be null) ahead of its tower
    // kernel and records `record_after` right behind it — i.e. before any follow-up kernel of the same pass (head MLPs of a
    // wide-head network), which is the work that should overlap the other part's tower
    virtual void set_tower_events(cudaEvent_t /*wait_before*/, cudaEvent_t /*record_after*/) {}
    // true when a recurrent pass launches further kernels behind its tower kernel
    virtual bool recurrent_has_followup() const { return false; }
    virtual int launches_initial() const = 0;
    virtual int launches_recurrent() const = 0;
};

NetBackend* make_fp32_backend();
NetBackend* make_bf16_backend();

}  // namespace mzb
