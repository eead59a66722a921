// host_common.h -- host-side types shared by select_host.cu and stream_host.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <functional>
#include <vector>

#include "../../include/baitstr_select.h"

namespace bstr {

struct BatchState;
struct ReadResult;
struct StreamEngine;

constexpr size_t IN_SLACK = 64;    // readable bytes kept on both sides of a device input batch

// Device (and pinned host) buffers for one in-flight batch.
struct DeviceSlot {
    uint8_t *d_in_raw, *d_in;      // d_in = d_in_raw + IN_SLACK
    uint32_t *d_nlpos, *d_tiles, *d_outlen;
    unsigned long long *d_lb;      // fused path: look-back words + ticket
    ReadResult *d_res;
    uint8_t *d_out;
    BatchState *d_state, *h_state;
    uint8_t *h_out;                // pinned
    ReadResult *h_res;             // pinned
    double rec_bytes, hdr_bytes;   // average record size / first header length of the batch in d_in (tile geometry)
    size_t cap_in, cap_nlpos, cap_tiles, cap_res, cap_outlen, cap_out, cap_hout, cap_hres, cap_lb;

    int reserve(struct ::bstr_ctx *ctx, uint64_t nbytes, uint64_t out_need);
    int reserve_generic(struct ::bstr_ctx *ctx, uint64_t nbytes);   // staged-kernel buffers, on demand
    void release();
};

constexpr int BSTR_LANES_PER_DEVICE = 3;

// One in-flight batch: its own stream, events and buffers on one device.  A batch moves
// through submit (H2D + kernels + state read-back enqueued) and collect (wait, D2H of the
// exact output size); lanes of one device overlap each other's copies and kernels.
struct Lane {
    DeviceSlot slot;
    cudaStream_t st;
    cudaEvent_t ev[8];
    int dev;                       // CUDA ordinal
    int dev_index;                 // index into bstr_params.devices
    const uint8_t *raw;            // host input of the batch in flight
    uint64_t nbytes;
    int final;
    bool busy;
    // a batch that already lives in device memory (the GPU inflater's text) borrows the slot's input
    // pointers for one submit/collect; the slot's own buffer is put back afterwards
    uint8_t *own_in_raw; size_t own_cap_in; bool borrowed;
};

}  // namespace bstr

struct bstr_ctx {
    bstr_params p;
    int dev;
    int sm_count;
    int exp_mask;                  // BSTR_EXP: phase-skipping bits for profiling experiments
    uint32_t geom_tile, geom_threads, geom_stage;   // BSTR_GEOM overrides (0 = automatic)
    bool auto_tile_off;            // BSTR_AUTOTILE=0
    int force_wmax;                // BSTR_WMAX: 2, 4, 6 or 8 (0 = from the record size)
    bool force_generic;            // BSTR_FORCE_GENERIC=1: staged kernels only (tests)
    cudaStream_t stream, stream2;  // stream2: second launch stream of bstr_bench_resident
    cudaEvent_t ev[8];
    bstr::DeviceSlot slot;         // used by bstr_scan_batch / bstr_bench_resident
    uint8_t *d_seqs; uint64_t *d_off; size_t cap_seqs, cap_off;   // bstr_match_reads
    uint32_t *d_flush; size_t cap_flush;
    std::vector<bstr::DeviceSlot> *bench_slots;   // bstr_bench_resident
    bstr::StreamEngine *engine;    // bstr_run_files (stream_host.cu)
    bstr::Lane *lanes;             // bstr_submit_batch / bstr_collect_batch, created on first use
    int n_lanes;
    void *lock;                    // std::mutex guarding stats and err (worker threads of bstr_run_files)
    bstr_stats stats;
    char err[512];
    bstr_ctx() { memset(this, 0, sizeof(*this)); }
};

namespace bstr {

void set_error(bstr_ctx *ctx, const char *fmt, ...);
int cuda_fail(bstr_ctx *ctx, cudaError_t e, const char *what, int line);
struct ScanParams scan_params(const bstr_ctx *ctx);
int launch_pipeline(bstr_ctx *ctx, DeviceSlot &s, uint64_t nbytes, int final, cudaStream_t st,
                    cudaEvent_t *ev, bool generic = false);
int decode_bad(bstr_ctx *ctx, unsigned long long first_bad, int64_t *bad_record);
void stream_release(bstr_ctx *ctx);
int lanes_create(bstr_ctx *ctx);
void lanes_release(bstr_ctx *ctx);
int lane_submit(bstr_ctx *ctx, Lane &ln, const uint8_t *raw, uint64_t nbytes, int final);
// on_state (optional): called once the batch's state is final -- `consumed` input bytes were whole
// records -- and BEFORE the output is copied back; a non-zero return aborts the collect with that code
int lane_collect(bstr_ctx *ctx, Lane &ln, bstr_batch_out *out, const std::function<int(uint64_t consumed)> *on_state = nullptr);
// The batch is d_text[0, nbytes) in the lane's device memory space: 16-byte aligned, IN_SLACK bytes
// readable before it and IN_SLACK writable bytes after it.  head = its first bytes on the host (tile
// geometry is estimated from them).  Collect with lane_collect_device.
int lane_submit_device(bstr_ctx *ctx, Lane &ln, uint8_t *d_text, uint64_t nbytes, int final, const uint8_t *head, size_t head_bytes);
int lane_collect_device(bstr_ctx *ctx, Lane &ln, bstr_batch_out *out, const std::function<int(uint64_t consumed)> *on_state);
double estimate_record_bytes(const uint8_t *raw, uint64_t nbytes, double *hdr_bytes);
int pipeline_stage_count(const bstr_ctx *ctx);
const char *pipeline_stage_name(const bstr_ctx *ctx, int k);
int pipeline_launches_per_batch(const bstr_ctx *ctx);
constexpr int BSTR_MAX_STAGES = 8;

#define CUDA_TRY(call)                                                          \
    do {                                                                        \
        cudaError_t e_ = (call);                                                \
        if (e_ != cudaSuccess) return bstr::cuda_fail(ctx, e_, #call, __LINE__); \
    } while (0)

}  // namespace bstr
