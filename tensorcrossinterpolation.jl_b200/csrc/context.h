// context.h — internal: handle layout, error plumbing, workspace helpers shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "../../include/tci_b200.h"

struct tci_workspace {           // grow-only device buffer
    void* ptr = nullptr;
    size_t bytes = 0;
};

struct tci_context {
    int device = 0;
    int num_sms = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;          // stream in use (own or caller's)
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::string err;
    int64_t launches = 0;
    // rrLU
    tci_workspace ws_matrix;                // working copy of A for the host-buffer entry point
    tci_workspace ws_lu;                    // L and U staging for the host-buffer entry point
    tci_workspace ws_small;                 // state, perms, partial argmax candidates, pivot row copy
    void* pinned = nullptr;                 // pinned host staging (state read-back, perms)
    size_t pinned_bytes = 0;
    double rrlu_ms[3] = {0, 0, 0};
    int64_t rrlu_steps = 0;
    int rrlu_smem_optin = 0;                // the single-CTA kernel's > 48 KB dynamic shared memory has been enabled
    int solve_smem_optin = 0;               // the small-solve kernel's > 48 KB dynamic shared memory has been enabled
    int rrlu_extracted = 0;                 // the last rrlu_run already wrote L and U (single-CTA kernel): rrlu_finish skips its launch
    int rrlu_pdl = 1;                       // programmatic dependent launch between the step kernels (TCI_RRLU_PDL=0 disables)
    int rrlu_late_state = 1;                // single-CTA kernel: read the final state once, in rrlu_finish (TCI_RRLU_LATE_STATE=0: twice)
    int rrlu_profile = 0;                   // 1: bracket every update-kernel launch with events
    long long l2_window_bytes = -1;         // persisting-L2 window of the step kernels (-1: not initialised, 0: off)
    std::vector<cudaEvent_t> ev_pool;
    std::vector<float> rrlu_launch_ms;      // per-launch update-kernel times of the last profiled rrlu
    // generic scratch for the other kernels
    tci_workspace ws_a, ws_b, ws_c;
    tci_workspace dist_parked;              // state block of the last destroyed sharded rrLU, reused by the next create
};

int tci_fail(tci_handle h, int code, const char* fmt, ...);
int tci_ws_reserve(tci_handle h, tci_workspace& ws, size_t bytes);
int tci_pinned_reserve(tci_handle h, size_t bytes);

#define TCI_CUDA(h, expr)                                                                      \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess)                                                                \
            return tci_fail((h), TCI_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                 \
                            cudaGetErrorString(e__), __FILE__, __LINE__);                      \
    } while (0)

#define TCI_LAUNCH_CHECK(h)                                                                    \
    do {                                                                                       \
        (h)->launches++;                                                                       \
        cudaError_t e__ = cudaGetLastError();                                                  \
        if (e__ != cudaSuccess)                                                                \
            return tci_fail((h), TCI_ERR_CUDA, "kernel launch failed: %s (%s:%d)",             \
                            cudaGetErrorString(e__), __FILE__, __LINE__);                      \
    } while (0)
