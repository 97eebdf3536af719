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

// One 32-byte slot = one L2 sector: a probe reads key, unit id and the three rotating per-batch frame
// masks with a single access.  mask[q] bit f = frame f of the batch using buffer q touches this unit.
struct __align__(32) B2RSlot {
    uint64_t key;        // B2R_KEY_EMPTY when free
    int32_t val;         // unit index (-1 while unassigned / on pool overflow)
    uint32_t mask[3];
    uint32_t pad[2];
};

struct B2RHash {
    B2RSlot* slots;   // nslots
    uint32_t mask;    // nslots - 1
};

__device__ __forceinline__ int b2r_hash_find(const B2RHash& h, uint64_t key) {
    uint32_t s = (uint32_t)b2r_mix64(key) & h.mask;
    for (uint32_t probe = 0; probe <= h.mask; probe++) {
        const ulonglong2 kv = *reinterpret_cast<const ulonglong2*>(&h.slots[s]);   // key + val in one 16-byte load
        if (kv.x == key) return (int)(uint32_t)kv.y;
        if (kv.x == B2R_KEY_EMPTY) return -1;
        s = (s + 1) & h.mask;
    }
    return -1;
}
__host__ __device__ __forceinline__ void b2r_unpack_key(uint64_t key, int& x, int& y, int& z) {
    x = (int)((key >> 42) & 0x1FFFFFu) - B2R_KEY_BIAS;
    y = (int)((key >> 21) & 0x1FFFFFu) - B2R_KEY_BIAS;
    z = (int)(key & 0x1FFFFFu) - B2R_KEY_BIAS;
}

// ---- IEEE-exact f32 division with a shared reciprocal ------------------------------------------------
// Several quotients by the same divisor (u and v projections by z; tsdf/r/g/b running averages by w+1)
// share ONE correctly rounded reciprocal, then Markstein's correction gives RN(a/b) in three instructions:
//   y  = RN(1/b)      : MUFU.RCP (< 1 ulp) + one FMA Newton step -- correctly rounded for every significand
//                       except all-ones (Markstein 1990), which b2r_div_domain() excludes;
//   q0 = RN(a*y); r = fma(-b, q0, a) (exact); q = fma(r, y, q0) = RN(a/b)   when nothing under/overflows.
// b2r_div_domain(b) bounds b to [2^-60, 2^60); the callers' numerators are either 0, or large enough for r to
// be exact, or so small that the quotient cannot influence the result (see tsdf.cu).  Everything outside the
// domain takes div.rn.  b2r_selftest_div() (tsdf.cu) brute-forces both claims on the GPU in the tests.
__device__ __forceinline__ float b2r_rcp_markstein(float b) {
    float y0;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y0) : "f"(b));
    const float e = __fmaf_rn(-b, y0, 1.0f);
    return __fmaf_rn(y0, e, y0);
}
__device__ __forceinline__ bool b2r_div_domain(float b) {   // 2^-60 <= b < 2^60 and significand not all ones
    const uint32_t u = __float_as_uint(b);
    return ((u >> 23) - 67u) < 120u && (u & 0x7FFFFFu) != 0x7FFFFFu;
}
__device__ __forceinline__ float b2r_div_fast(float a, float b, float y) {
    const float q0 = __fmul_rn(a, y);
    const float r = __fmaf_rn(-b, q0, a);
    return __fmaf_rn(r, y, q0);
}

#define B2R_UNIT_RES 16
#define B2R_UNIT_VOX 4096
