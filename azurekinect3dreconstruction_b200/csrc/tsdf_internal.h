// tsdf_internal.h -- the state behind a b2r_tsdf handle (shared by tsdf.cu and extract.cu).
#pragma once
#include "common.cuh"

// device counters (int32 words)
enum {
    B2R_C_NUM_UNITS = 0,   // units allocated so far
    B2R_C_ERR = 1,         // sticky error bits (1 = pool overflow, 2 = key out of range, 4 = hash full)
    B2R_C_TOUCHED0 = 2,    // touched-unit counts of the three rotating batch buffers (words 2,3,4)
    B2R_C_UPDATED = 6,     // u64 (words 6,7): voxels whose record changed, cumulative (N_upd)
    B2R_C_TOUCHED_TOTAL = 8,  // u64 (words 8,9): cumulative (unit, frame) pairs fused
    B2R_C_LAST_TOUCHED = 10,  // units touched by the most recent fused batch
    B2R_C_WORK0 = 11,         // K2 work-item counters of the three rotating batch buffers (words 11,12,13)
    B2R_C_TICKET = 14,        // K2: CTAs that have finished the running pass (the last one recycles the previous batch's counters)
    B2R_C_WORDS = 16
};

#define B2R_HOST_RING 6
#define B2R_MAX_BATCH 32   // frames fused per pass over the touched units (bits of B2RSlot::mask)

// what the fusion kernel needs to know about one pending frame (kernel parameter, by value)
struct B2RFrame {
    float E[12];             // (float)extrinsic rows 0..2
    float Es02, Es12, Es22;  // E(:,2) * voxel_length (f32)
    float fx, fy, cx, cy, safe_w, safe_h;
    int W, H;
    const float2* texel;     // packed frame: {depth m, rgb8 | gray f32 bits}
    const float* mult;       // ray multiplier image of the frame's intrinsics (handle cache)
};

struct b2r_tsdf {
    double voxel_length, sdf_trunc, unit_length;
    int color_type;   // 0 NoColor, 1 RGB8, 2 Gray32
    int res, stride, planes;
    int64_t cap_units;
    int rank, world, device;
    B2RHash hash;
    uint32_t nslots;
    int32_t* unit_keys;    // cap_units * 3
    float* pool;           // cap_units * planes * 4096 : planes = tsdf, weight[, r, g, b]
    int32_t* touched[3];   // cap_units each: hash slots touched by the batch using that buffer
    int32_t* touched_sorted;   // short passes: the pending batch's list ordered heaviest unit first (sort_touched)
    int sort_touched;
    int32_t* counters;     // B2R_C_WORDS
    // pending batch: frames packed + allocated (K1 done) but not yet fused (K2 runs at flush)
    int q;                 // rotating buffer index of the pending batch
    int pending;           // frames in it
    int max_batch;         // <= B2R_MAX_BATCH
    B2RFrame frames[B2R_MAX_BATCH];
    float2* texel[2][B2R_MAX_BATCH];   // overlap mode packs batch n + 1 into the other set while batch n is being fused
    size_t texel_px;
    float* mult;             // depth-to-camera-distance multiplier image of the cached intrinsics (shared by every frame)
    size_t mult_px;
    double mult_key[6];      // W, H, fx, fy, cx, cy the image was computed for
    int64_t batches;       // fused so far (alternates the sweep direction)
    int n_sm;              // SMs of the handle's device
    int fuse_grid;         // persistent grid size of the fusion kernel
    int fuse_smem;         // dynamic shared memory of the fusion kernel
    int fuse_xl;           // x-layers per K2 work item (4 or 2)
    int fuse_ring_rounds;  // K2 passes shorter than this many items per CTA claim one item at a time
    // Overlap mode (B2R_TSDF_OVERLAP, default on for sharded handles): the fusion passes run on an INTERNAL stream, so the K1 of
    // batch n + 1 (caller's stream) runs beside the K2 of batch n; K1 of batch n waits for K2 of batch n - 2 (texel set and
    // touched buffer reuse), every reader of the volume joins the internal stream first.
    int overlap;
    cudaStream_t fuse_stream;
    cudaEvent_t ev_k1, ev_k2[3];   // ev_k2[n % 3]: pass n has finished
    int unjoined;                  // a pass was launched on fuse_stream that the caller's stream has not waited for yet
    // host-input pipeline (b2r_tsdf_integrate_host)
    cudaStream_t copy_stream;
    cudaEvent_t ev_copied[B2R_HOST_RING], ev_consumed[B2R_HOST_RING];
    void* stage_depth[B2R_HOST_RING];
    void* stage_color[B2R_HOST_RING];
    size_t stage_depth_bytes, stage_color_bytes;
    int64_t host_seq;
    // extraction scratch
    void* scratch;
    size_t scratch_bytes;
    // The Open3D-shaped callers extract with a two-call protocol (sizes first, then the arrays).  generation counts
    // every change of the volume (integrate / reset / import); mesh_gen remembers for which generation the sorted
    // order, per-voxel vertex slots and scans in `scratch` are valid, so the second call only runs the fill pass.
    uint64_t generation, mesh_gen, mesh_slab;
    uint64_t unit_epoch, sort_epoch;   // unit_epoch: bumped by reset / import (anything but appends that changes the unit set)
    int sort_n;                        // unit count the sorted order in scratch was built for (-1: none)
    int mesh_n, mesh_own[2];
    int32_t mesh_v01[2], mesh_t01[2];
    // optional per-kernel timing (b2r_tsdf_profile): event pairs around every K1 / K2 launch
    int profile;
    cudaEvent_t* prof_ev;
    int* prof_kind;        // per pair: 1 = pack+alloc, 2 = fuse
    int prof_cap, prof_n;
    int64_t prof_frames;
};

struct B2RDeviceGuard {
    int prev;
    explicit B2RDeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~B2RDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int b2r_tsdf_ensure_scratch(b2r_tsdf* h, size_t bytes);
// fuses whatever is pending (K2).  Every entry point that reads or rewrites the volume calls it first.
int b2r_tsdf_flush_pending(b2r_tsdf* h, cudaStream_t stream);
extern "C" int b2r_tsdf_stats(b2r_tsdf* h, int64_t* out, cudaStream_t stream);
extern "C" int b2r_tsdf_reset(b2r_tsdf* h, cudaStream_t stream);
