// tsdf_internal.h -- the state behind a b2r_tsdf handle (shared by tsdf.cu and extract.cu).
#pragma once
#include "common.cuh"

// device counters (int32 words)
enum {
    B2R_C_NUM_UNITS = 0,   // units allocated so far
    B2R_C_ERR = 1,         // sticky error bits (1 = pool overflow, 2 = key out of range, 4 = hash full)
    B2R_C_TOUCHED0 = 2,    // touched-unit count, buffer 0
    B2R_C_TOUCHED1 = 3,    // touched-unit count, buffer 1
    B2R_C_UPDATED = 4,     // u64 (words 4,5): voxels whose record changed, cumulative (N_upd)
    B2R_C_TOUCHED_TOTAL = 6,  // u64 (words 6,7): cumulative touched units
    B2R_C_WORDS = 16
};

#define B2R_HOST_RING 3

struct b2r_tsdf {
    double voxel_length, sdf_trunc, unit_length;
    int color_type;   // 0 NoColor, 1 RGB8, 2 Gray32
    int res, stride, planes;
    int64_t cap_units;
    int rank, world, device;
    B2RHash hash;
    uint32_t nslots;
    int32_t* slot_stamp;   // nslots: frame id that last touched the slot
    int32_t* unit_keys;    // cap_units * 3
    float* pool;           // cap_units * planes * 4096 : planes = tsdf, weight[, r, g, b]
    int32_t* touched[2];   // cap_units each: hash slots touched by the frame in flight
    int32_t* counters;     // B2R_C_WORDS
    float4* texel[2];      // packed frame: {depth m, ray multiplier, rgb8|gray bits, 0}
    size_t texel_px;
    int frame_id;
    int fuse_grid;         // persistent grid size of the fusion kernel
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
    // optional per-kernel timing (b2r_tsdf_profile): 3 events per frame (before K1, between, after K2)
    int profile;
    cudaEvent_t* prof_ev;
    int prof_cap, prof_n;
};

struct B2RDeviceGuard {
    int prev;
    explicit B2RDeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~B2RDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int b2r_tsdf_ensure_scratch(b2r_tsdf* h, size_t bytes);
extern "C" int b2r_tsdf_stats(b2r_tsdf* h, int64_t* out, cudaStream_t stream);
