// common.cuh -- shared device/host helpers for libb2recon (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>

#define B2R_API extern "C" __attribute__((visibility("default")))

// error codes (include/b2recon.h)
#define B2R_OK 0
#define B2R_ERR_BAD_ARG (-1)
#define B2R_ERR_CUDA (-2)
#define B2R_ERR_CAPACITY (-3)
#define B2R_ERR_FORMAT (-4)
#define B2R_ERR_SINGULAR (-5)
#define B2R_ERR_RANGE (-6)

void b2r_set_error(const char* fmt, ...);

#define B2R_CUDA(call)                                                                               \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            b2r_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));      \
            return B2R_ERR_CUDA;                                                                     \
        }                                                                                            \
    } while (0)

#define B2R_CHECK_LAUNCH() B2R_CUDA(cudaGetLastError())

// ---- exactly-rounded, never-contracted arithmetic (parity with a non-FMA x86-64 build) ---------
__device__ __forceinline__ float xfmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xfadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xfsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float xfdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double xdmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double xdadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double xdsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double xddiv(double a, double b) { return __ddiv_rn(a, b); }

// ---- spatial hash over 16^3 volume units -----------------------------------------------------------
#define B2R_KEY_EMPTY 0xFFFFFFFFFFFFFFFFull
#define B2R_KEY_BIAS (1 << 20)

__host__ __device__ __forceinline__ uint64_t b2r_pack_key(int x, int y, int z) {
    return ((uint64_t)(uint32_t)(x + B2R_KEY_BIAS) << 42) | ((uint64_t)(uint32_t)(y + B2R_KEY_BIAS) << 21) |
           (uint64_t)(uint32_t)(z + B2R_KEY_BIAS);
}
__host__ __device__ __forceinline__ bool b2r_key_in_range(int x, int y, int z) {
    return x >= -B2R_KEY_BIAS && x < B2R_KEY_BIAS && y >= -B2R_KEY_BIAS && y < B2R_KEY_BIAS && z >= -B2R_KEY_BIAS &&
           z < B2R_KEY_BIAS;
}
__host__ __device__ __forceinline__ uint64_t b2r_mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}
// shard owner of a unit key: uses the high hash bits, the slot index uses the low ones
__host__ __device__ __forceinline__ int b2r_owner(uint64_t key, int world) {
    return world <= 1 ? 0 : (int)((b2r_mix64(key) >> 33) % (uint64_t)world);
}

struct B2RHash {
    uint64_t* keys;   // nslots, B2R_KEY_EMPTY when free
    int32_t* vals;    // nslots -> unit index (or -1 while unassigned / on overflow)
    uint32_t mask;    // nslots - 1
};

__device__ __forceinline__ int b2r_hash_find(const B2RHash& h, uint64_t key) {
    uint32_t s = (uint32_t)b2r_mix64(key) & h.mask;
    for (uint32_t probe = 0; probe <= h.mask; probe++) {
        uint64_t k = h.keys[s];
        if (k == key) return h.vals[s];
        if (k == B2R_KEY_EMPTY) return -1;
        s = (s + 1) & h.mask;
    }
    return -1;
}

#define B2R_UNIT_RES 16
#define B2R_UNIT_VOX 4096
