// tsdf.cu -- ScalableTSDFVolume on sm_100a: spatial hash of 16^3 volume units + projective fusion.
//
// Replaces, behind the C-ABI in include/b2recon.h, what the reference reaches through
//   o3d.pipelines.integration.ScalableTSDFVolume(...).integrate(rgbd, intrinsic, extrinsic)
// (/root/reference/CodeThatWorks/GoodColorReconstruction.py:111-115,270;
//  CodeThatWorks/bundlefusionattemps/bundlefusion_reconstruction.py:108-112,276,315).
// Semantics restated in SURVEY.md Appendix A.4; the f32 operation order of the per-voxel update is
// reproduced exactly (explicit round-to-nearest intrinsics, no FMA contraction, IEEE-exact division) so
// TSDF/weight/colour match the CPU oracle bit for bit and the (int)u_f / (int)v_f pixel choices can
// never diverge.
//
// Two kernels, both on the caller's stream:
//   K1 tsdf_pack_alloc_kernel (at every integrate call; one launch can take a batch of frames; the two roles are
//      dealt out evenly over the launch order so the latency-bound one runs beside the bandwidth-bound one):
//        pack role  - a CTA stages 2048 pixels of depth and colour by two TMA bulk copies, fuses the RGBDImage
//                     conversion (u16 -> f32 metres, trunc) and writes ONE 8-byte texel {d, rgb8 | gray} per pixel,
//                     256 contiguous bytes per warp store (the depth-to-camera-distance multiplier image depends on
//                     the intrinsics only: it is cached in the handle and gathered by K2 beside the texel);
//        alloc role - one thread per depth_sampling_stride-th pixel unprojects in f64 (divisions by fx, fy and the
//                     unit length through host-rounded reciprocals, exact); runs of equal key boxes are merged, the
//                     run heads write their cells' keys into a shared list and the block probes them in parallel
//                     (one 32-byte slot read per cell, all of a thread's probes in flight together); first touches
//                     of a (unit, batch) are appended to the batch's touched list (one counter atomic per warp) and
//                     the frame's bit is OR-ed into the slot's batch mask.
//   K2 tsdf_fuse_kernel (when max_batch frames are pending, or when anything reads the volume):
//        persistent CTAs draw work items (touched unit x 4 x-slabs) from a counter.  A slab (planes x 4 KiB) is
//        staged into shared memory by TMA bulk copies (cp.async.bulk + mbarrier) through a 2-deep ring, so the
//        next slab is in flight while the current one is fused.  A thread owns 4 consecutive z voxels, keeps
//        them in registers while it applies EVERY pending frame that touches the unit (in frame order,
//        so the result equals frame-by-frame integration bit for bit), and writes the record back once,
//        only if it changed.  HBM traffic per frame therefore falls with the batch size; the voxel
//        arithmetic (not DRAM) becomes the limiter, which is why the divisions share reciprocals.
#include "tsdf_internal.h"
#include "host_math.h"
#include <new>
#include <string.h>
#include <stdlib.h>

// ------------------------------------------------------------------------------------------------
// error string
// ------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
void b2r_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
B2R_API const char* b2r_last_error(void) { return g_err; }
B2R_API int b2r_version(void) { return 200; }

// ------------------------------------------------------------------------------------------------
// K1: pack + allocate
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

struct PackFrame {
    const void* depth;        // u16 or f32, W*H
    const uint8_t* color;     // u8, channels 3|4 (may be null)
    const float* gray;        // f32 intensity (Gray32), may be null
    float2* texel;
    double pose[12];          // extrinsic^-1, rows 0..2
    uint32_t bit;             // 1 << (index of the frame in the pending batch)
    uint32_t pad;
};

struct PackParams {
    int W, H;
    int depth_is_u16;
    float depth_scale, depth_trunc;
    int color_channels, color_bgr;
    float fcx, fcy, ifx, ify;      // (float)cx, (float)cy, 1.0f/(float)fx, 1.0f/(float)fy
    double fx, fy, cx, cy;         // f64 intrinsics for the allocation samples
    double trunc, unit_length;
    int stride, Ws, Hs, tiles_x;   // sample grid and its tiles of 32 x (8 * ALLOC_SPT) samples
    int n_alloc_blocks;            // allocation tiles per frame
    int n_frames, tiles_per_frame, n_pack;   // frames in this launch, pixel tiles per frame, persistent pack CTAs
    uint32_t pack_stage_bytes, pack_coff;    // ring stage size and the colour part's offset inside a stage
    const float* mult;             // cached multiplier image (pack_mult of every pixel)
    float rscale; int scale_fast;  // RN(1/depth_scale); depth_scale inside b2r_div_domain
    double rfx, rfy, runit;        // RN(1/fx), RN(1/fy), RN(1/unit_length)
    int ddiv_fast;                 // all three f64 divisors qualify for the shared-reciprocal division
    int tma_ok;                    // every tile of every image is a 16-byte aligned multiple of 16 bytes: TMA staging in the pack role
    int rank, world;
    int q;                         // batch buffer (mask word) in use
    B2RHash hash;
    int32_t* unit_keys;
    int32_t* touched;
    int32_t* counters;
    int cap_units;
    PackFrame fr[B2R_MAX_BATCH];
};

// u16 -> metres.  0x4B00vvvv is the float 8388608 + v, so the conversion is one LOP + one FADD (the I2F unit runs
// at quarter rate); the division shares RN(1/depth_scale) when the host found the divisor inside b2r_div_domain.
__device__ __forceinline__ float depth_from_u16(const PackParams& P, uint32_t v) {
    const float fv = __fadd_rn(__uint_as_float(0x4B000000u | v), -8388608.0f);
    const float d = P.scale_fast ? b2r_div_fast(fv, P.depth_scale, P.rscale) : xfdiv(fv, P.depth_scale);
    return d >= P.depth_trunc ? 0.0f : d;
}
__device__ __forceinline__ float load_depth_m(const PackParams& P, const void* depth, int idx) {
    if (P.depth_is_u16) return depth_from_u16(P, (uint32_t)__ldg((const uint16_t*)depth + idx));
    return __ldg((const float*)depth + idx);
}
// f64 division by one of the call's constants (fx, fy, unit_length): q0 = RN(a*y), r = a - b*q0 (exact), RN(q0 + r*y)
// == RN(a/b) when y = RN(1/b) (computed with an IEEE division on the host) and b's significand is not all ones
// (Markstein); b2r_selftest_ddiv brute-forces it in the tests.  Non-finite a gives NaN instead of +-inf: both fail
// the key-range test below in the same way.
__device__ __forceinline__ double ddiv_const(double a, double b, double y, int fast) {
    if (fast) {
        const double q0 = __dmul_rn(a, y);
        const double r = __fma_rn(-b, q0, a);
        return __fma_rn(r, y, q0);
    }
    return xddiv(a, b);
}

// slow path of the allocator: insert-or-find `key` (linear probing), mark it touched by frame `bit`
__device__ __noinline__ void alloc_key(const PackParams& P, uint64_t key, uint32_t bit) {
    int kx, ky, kz;
    b2r_unpack_key(key, kx, ky, kz);
    uint32_t s = (uint32_t)b2r_mix64(key) & P.hash.mask;
    for (uint32_t probe = 0; probe <= P.hash.mask; probe++) {
        B2RSlot* slot = &P.hash.slots[s];
        unsigned long long k = __ldcg(reinterpret_cast<const unsigned long long*>(&slot->key));   // L2: coherent
        if (k == B2R_KEY_EMPTY) {
            k = atomicCAS((unsigned long long*)&slot->key, (unsigned long long)B2R_KEY_EMPTY, (unsigned long long)key);
            if (k == B2R_KEY_EMPTY) {
                // we own the slot: bump-allocate a zeroed unit
                int id = atomicAdd(&P.counters[B2R_C_NUM_UNITS], 1);
                if (id >= P.cap_units) {
                    atomicOr(&P.counters[B2R_C_ERR], 1);
                    id = -1;
                } else {
                    P.unit_keys[3 * id] = kx; P.unit_keys[3 * id + 1] = ky; P.unit_keys[3 * id + 2] = kz;
                }
                slot->val = id;   // consumed by the NEXT kernel (K2), no intra-kernel reader
                k = key;
            }
        }
        if (k == key) {
            const uint32_t old = atomicOr(&slot->mask[P.q], bit);
            if (old == 0) {   // first touch of this unit in this batch -> append to the touched list
                const int pos = atomicAdd(&P.counters[B2R_C_TOUCHED0 + P.q], 1);
                if (pos < P.cap_units) P.touched[pos] = (int)s;
            }
            return;
        }
        s = (s + 1) & P.hash.mask;
    }
    atomicOr(&P.counters[B2R_C_ERR], 4);
}

#define PACK_PX 8       // pixels per thread in the pack role
#define PACK_TILE (256 * PACK_PX)   // pixels per CTA of the pack role
#ifndef PACK_STAGES
#define PACK_STAGES 3   // ring of staged tiles per persistent pack CTA (dynamic shared memory; the alloc role's key set aliases it)
#endif
#ifndef PACK_CTAS_PER_SM
#define PACK_CTAS_PER_SM 2
#endif
#ifndef ALLOC_SPT
#define ALLOC_SPT 4        // samples per thread of an allocation tile in batched launches: 32 x (8 * ALLOC_SPT) samples per CTA
                           // (launches of one or two frames use 1: the GPU is empty then, and the CTA's dependent chain -- four f64
                           // sample chains, a four times longer key list -- is what the call waits for)
#endif
#define ALLOC_LIST 2048  // compacted key list of an allocation tile (entries beyond it take the direct path)
#define ALLOC_SMEM (4096 + ALLOC_LIST * 2)
#ifndef ALLOC_IT
#define ALLOC_IT 2       // keys per thread whose hash probes are in flight together
#endif
#ifndef PACK_MIN_CTAS
#define PACK_MIN_CTAS 5
#endif

// 8-bit colours ride in the texel as one 0x00BBGGRR word
// byte n of c -> float, without the quarter-rate I2F: 0x4B0000nn is the float 8388608 + nn
template <int N>
__device__ __forceinline__ float byte_to_float(uint32_t c) {
    return __fadd_rn(__uint_as_float(__byte_perm(c, 0x4B000000u, 0x7540 | N)), -8388608.0f);
}

// CreateDepthToCameraDistanceMultiplierFloatImage: depends on the intrinsics only, so it is computed
// once per (W, H, fx, fy, cx, cy) and cached in the handle; every frame's pack reads it back (L2 resident).
__global__ void tsdf_mult_kernel(float* __restrict__ mult, int W, int H, float fcx, float fcy, float ifx, float ify) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= W * H) return;
    const int i = idx / W, j = idx - i * W;
    const float xx = xfmul(xfsub((float)j, fcx), ifx);
    const float yy = xfmul(xfsub((float)i, fcy), ify);
    mult[idx] = __fsqrt_rn(xfadd(xfadd(xfmul(xx, xx), xfmul(yy, yy)), 1.0f));
}

#ifdef B2R_K1_TRACE
__device__ unsigned long long g_k1_trace[4 * 8192];
__device__ unsigned long long g_k1_marks[8 * 8192];   // per CTA: stage timestamps of the alloc role (thread 0)
#define K1_MARK(i) do { if (threadIdx.x == 0 && blockIdx.x < 8192) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_k1_marks[8 * blockIdx.x + (i)] = t_; } } while (0)
__device__ __forceinline__ unsigned long long k1_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
struct K1Trace {
    unsigned long long t0; int role;
    __device__ K1Trace(int r) : t0(k1_now()), role(r) {}
    __device__ ~K1Trace() {
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned L = blockIdx.x;
            unsigned sm; asm("mov.u32 %0, %smid;" : "=r"(sm));
            if (L < 8192) { g_k1_trace[4 * L] = t0; g_k1_trace[4 * L + 1] = k1_now(); g_k1_trace[4 * L + 2] = sm; g_k1_trace[4 * L + 3] = role; }
        }
    }
};
B2R_API int b2r_debug_k1_marks(unsigned long long* out, int n) {
    return cudaMemcpyFromSymbol(out, g_k1_marks, sizeof(unsigned long long) * 8 * (size_t)n) == cudaSuccess ? 0 : -1;
}
B2R_API int b2r_debug_k1_trace(unsigned long long* out, int n) {
    return cudaMemcpyFromSymbol(out, g_k1_trace, sizeof(unsigned long long) * 4 * (size_t)n) == cudaSuccess ? 0 : -1;
}
#else
#define K1_MARK(i) do { } while (0)
#endif
template <int SPT>
__global__ void __launch_bounds__(256, PACK_MIN_CTAS) tsdf_pack_alloc_kernel(const __grid_constant__ PackParams P) {
    // One launch, two roles.  The first n_pack CTAs are PERSISTENT pack CTAs: each walks the launch's pixel tiles with stride
    // n_pack through a PACK_STAGES-deep ring of TMA bulk copies, so a tile's HBM latency is paid while the tiles before it
    // are converted (one tile per short-lived CTA spent two thirds of its life waiting for its own copy).  Every CTA after
    // them takes one allocation tile; those are latency-bound and run beside the bandwidth-bound pack CTAs for the whole
    // launch (CTAs start in linear block order, so the pack CTAs are resident first).
    const unsigned L = blockIdx.x;
    const bool role_pack = L < (unsigned)P.n_pack;
#ifdef B2R_K1_TRACE
    K1Trace trace_(role_pack ? 1 : 0);
#endif
    if (role_pack) {
        // ---- pack role: a1 depth conversion + colour.  A warp owns 32 * PACK_PX consecutive pixels of a tile; in step k lane l
        //      takes pixel 32 k + l, so every 8-byte texel store of the warp lands in 256 contiguous bytes (the store REQUEST
        //      rate, not the bytes, bounded the thread-contiguous layout).
        extern __shared__ __align__(128) unsigned char k1_smem[];
        __shared__ __align__(8) uint64_t tile_bar[PACK_STAGES];
        const int npx = P.W * P.H;
        const int lane = threadIdx.x & 31;
        const int total = P.n_frames * P.tiles_per_frame;
        const uint32_t dbytes = P.depth_is_u16 ? 2u : 4u;
        const uint32_t cbytes = P.fr[0].color ? (uint32_t)P.color_channels : (P.fr[0].gray ? 4u : 0u);
        const uint32_t sel = P.color_bgr ? 0x4012u : 0x4210u;   // swap byte 0 and byte 2 for BGR; byte 3 cleared
        auto issue = [&](int t, int st) {                        // thread 0: tile t -> ring stage st
            const int f = t / P.tiles_per_frame, tile0 = (t - f * P.tiles_per_frame) * PACK_TILE;
            const int tpx = min(PACK_TILE, npx - tile0);
            const PackFrame& F = P.fr[f];
            unsigned char* sb = k1_smem + (size_t)st * P.pack_stage_bytes;
            mbar_expect_tx(&tile_bar[st], (uint32_t)tpx * (dbytes + cbytes));
            tma_load_1d(sb, (const unsigned char*)F.depth + (size_t)tile0 * dbytes, (uint32_t)tpx * dbytes, &tile_bar[st]);
            if (cbytes) tma_load_1d(sb + P.pack_coff, (F.color ? (const unsigned char*)F.color : (const unsigned char*)F.gray) + (size_t)tile0 * cbytes, (uint32_t)tpx * cbytes, &tile_bar[st]);
        };
        if (P.tma_ok) {
            if (threadIdx.x == 0) {
                for (int st = 0; st < PACK_STAGES; st++) mbar_init(&tile_bar[st], 1);
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                for (int k = 0; k < PACK_STAGES - 1; k++)
                    if ((int)L + k * P.n_pack < total) issue((int)L + k * P.n_pack, k);
            }
            __syncthreads();                 // the barriers' initialisation is visible to the waiters
        }
        int k = 0;
        for (int t = (int)L; t < total; t += P.n_pack, k++) {
            const int f = t / P.tiles_per_frame, tile0 = (t - f * P.tiles_per_frame) * PACK_TILE;
            const int tpx = min(PACK_TILE, npx - tile0);
            const PackFrame& F = P.fr[f];
            const int l0 = (threadIdx.x >> 5) * (32 * PACK_PX) + lane;
            uint32_t dv[PACK_PX], cv[PACK_PX];
            if (P.tma_ok) {
                const int st = k % PACK_STAGES;
                // the stage freed by the previous iteration's closing barrier takes the tile PACK_STAGES - 1 ahead
                if (threadIdx.x == 0 && t + (PACK_STAGES - 1) * P.n_pack < total) issue(t + (PACK_STAGES - 1) * P.n_pack, (k + PACK_STAGES - 1) % PACK_STAGES);
                mbar_wait(&tile_bar[st], (uint32_t)(k / PACK_STAGES) & 1u);
                const unsigned char* s_d = k1_smem + (size_t)st * P.pack_stage_bytes;
                const unsigned char* s_c = s_d + P.pack_coff;
#pragma unroll
                for (int q = 0; q < PACK_PX; q++) {
                    const int li = min(l0 + 32 * q, tpx - 1);
                    dv[q] = P.depth_is_u16 ? (uint32_t)reinterpret_cast<const uint16_t*>(s_d)[li] : reinterpret_cast<const uint32_t*>(s_d)[li];
                    if (cbytes == 4u) cv[q] = reinterpret_cast<const uint32_t*>(s_c)[li];
                    else if (cbytes == 3u) cv[q] = (uint32_t)s_c[3 * li] | ((uint32_t)s_c[3 * li + 1] << 8) | ((uint32_t)s_c[3 * li + 2] << 16);
                    else cv[q] = 0u;
                }
            } else {
                // images that are not 16-byte aligned tile by tile: the same tile through plain loads
#pragma unroll
                for (int q = 0; q < PACK_PX; q++) {
                    const int idx = tile0 + min(l0 + 32 * q, tpx - 1);
                    dv[q] = P.depth_is_u16 ? (uint32_t)__ldg((const uint16_t*)F.depth + idx) : __float_as_uint(__ldg((const float*)F.depth + idx));
                    if (F.color) {
                        if (P.color_channels == 4) {
                            cv[q] = __ldg(reinterpret_cast<const uint32_t*>(F.color) + idx);      // 0x..BBGGRR in memory order: byte 0 = channel 0
                        } else {
                            const uint8_t* c3 = F.color + (size_t)idx * 3;
                            cv[q] = (uint32_t)__ldg(c3) | ((uint32_t)__ldg(c3 + 1) << 8) | ((uint32_t)__ldg(c3 + 2) << 16);
                        }
                    } else if (F.gray) {
                        cv[q] = __float_as_uint(__ldg(F.gray + idx));
                    } else {
                        cv[q] = 0u;
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < PACK_PX; q++) {
                const int li = l0 + 32 * q;
                const float d = P.depth_is_u16 ? depth_from_u16(P, dv[q]) : __uint_as_float(dv[q]);
                const uint32_t c = F.color ? __byte_perm(cv[q], 0, sel) : cv[q];
                if (li < tpx) F.texel[tile0 + li] = make_float2(d, __uint_as_float(c));
            }
            if (P.tma_ok) __syncthreads();   // every thread has read this stage: it may be refilled
        }
        return;
    }
    const unsigned ridx = L - (unsigned)P.n_pack;
    const int fidx = (int)(ridx / (unsigned)P.n_alloc_blocks), bidx = (int)(ridx - (unsigned)fidx * (unsigned)P.n_alloc_blocks);
    const PackFrame& F = P.fr[fidx];
    // ---- alloc role (a8): block allocation from the stride-sampled pixels (f64 chain).
    // A CTA owns a 32 x 32 TILE of samples (128 x 128 pixels at stride 4; SPT samples per thread): a compact patch of the
    // surface touches a few dozen units, against the hundreds a raster strip of as many samples crosses, and the latency of
    // the tile's depth loads and of its hash round trips is paid once per 1024 samples.  Three stages:
    //   (1) per sample the upstream f64 chain -> integer key box;
    //   (2) the boxes are OR-ed into a shared-memory bitmap of unit coordinates relative to the tile's minimum corner, which is
    //       then compacted into a list: every unit key the tile touches, exactly once;
    //   (3) the threads take the list ALLOC_IT keys at a time: hash-slot probes of all of a thread's keys in flight together
    //       (unit exists and this frame's bit is set: done), then one atomicOr per key that needs it, results consumed after
    //       all are in flight, first touches appended to the touched list with one counter atomic per warp; absent keys take
    //       the CAS probing slow path.  Only the owner rank's keys are probed (sharded handles).
    K1_MARK(0);
    const int tile_x = bidx % P.tiles_x, tile_y = bidx / P.tiles_x;
    const unsigned lane = threadIdx.x & 31u;
    const int warp = threadIdx.x >> 5;
    // thread t takes samples (row 8 k + warp, column lane) of the tile, k = 0 .. SPT - 1: a warp holds 32 neighbours of
    // one sample row.  All depth loads are issued before the first f64 chain starts.
    const int sj = tile_x * 32 + (int)lane;
    float dpt[SPT];
#pragma unroll
    for (int k = 0; k < SPT; k++) {
        const int si = tile_y * (8 * SPT) + 8 * k + warp;
        dpt[k] = (sj < P.Ws && si < P.Hs) ? load_depth_m(P, F.depth, si * P.stride * P.W + sj * P.stride) : 0.0f;
    }
    uint64_t bkey[SPT];      // packed lo corner of the sample's key box
    uint32_t ecode[SPT];     // its extents, 10 bits each; 0 = the sample contributes nothing
    int mn[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
#pragma unroll
    for (int k = 0; k < SPT; k++) {
        bkey[k] = 0; ecode[k] = 0u;
        const float d = dpt[k];
        if (d > 0.0f) {
            const int i = (tile_y * (8 * SPT) + 8 * k + warp) * P.stride, j = sj * P.stride;
            const double z = (double)d;
            const double x = ddiv_const(xdmul(xdsub((double)j, P.cx), z), P.fx, P.rfx, P.ddiv_fast);
            const double y = ddiv_const(xdmul(xdsub((double)i, P.cy), z), P.fy, P.rfy, P.ddiv_fast);
            long long ext[3];
            int lo[3];
#pragma unroll
            for (int a = 0; a < 3; a++) {
                const double p = xdadd(xdadd(xdadd(xdmul(F.pose[4 * a], x), xdmul(F.pose[4 * a + 1], y)), xdmul(F.pose[4 * a + 2], z)), F.pose[4 * a + 3]);
                const double fl = floor(ddiv_const(xdsub(p, P.trunc), P.unit_length, P.runit, P.ddiv_fast));
                const double fh = floor(ddiv_const(xdadd(p, P.trunc), P.unit_length, P.runit, P.ddiv_fast));
                const bool ok = fl >= -(double)B2R_KEY_BIAS && fh < (double)B2R_KEY_BIAS;   // also false for NaN
                lo[a] = ok ? (int)fl : 0;
                ext[a] = ok ? (long long)((int)fh - (int)fl + 1) : -1;
            }
            if (ext[0] < 0 || ext[1] < 0 || ext[2] < 0 || ext[0] * ext[1] * ext[2] > 4096) {
                atomicOr(&P.counters[B2R_C_ERR], 2);   // key out of range / degenerate box: report, don't loop
            } else {
                bkey[k] = b2r_pack_key(lo[0], lo[1], lo[2]);
                ecode[k] = (uint32_t)ext[0] | ((uint32_t)ext[1] << 10) | ((uint32_t)ext[2] << 20);
                mn[0] = min(mn[0], lo[0]); mn[1] = min(mn[1], lo[1]); mn[2] = min(mn[2], lo[2]);
            }
        }
    }
    K1_MARK(1);
    // (2) The tile's keys as a BITMAP: bit (x, y, z) of a 32 x 32 x 32 window of unit coordinates anchored at the tile's minimum
    //     corner (a 128 x 128-pixel patch spans a few units per axis; 32 units are 2 m at 4 mm voxels).  A sample whose box
    //     differs from its left neighbour's ORs one word per (x, y) of its box -- the z extent is a run of bits -- so building
    //     the tile's duplicate-free key set costs a handful of native 32-bit shared-memory atomics per box and no probing
    //     loop.  Cells outside the window (a depth discontinuity of more than 31 units inside one tile) go straight to the
    //     global table.
    extern __shared__ __align__(128) unsigned char k1_smem[];
    uint32_t* s_bm = reinterpret_cast<uint32_t*>(k1_smem);                         // 1024 words
    uint16_t* s_list = reinterpret_cast<uint16_t*>(k1_smem + 4096);                 // ALLOC_LIST compacted cell ids
    __shared__ int s_wmin[8][3];
    __shared__ int s_wcnt[8];
    {
        const int m0 = __reduce_min_sync(0xffffffffu, mn[0]);
        const int m1 = __reduce_min_sync(0xffffffffu, mn[1]);
        const int m2 = __reduce_min_sync(0xffffffffu, mn[2]);
        if (lane == 0) { s_wmin[warp][0] = m0; s_wmin[warp][1] = m1; s_wmin[warp][2] = m2; }
    }
    for (int q = threadIdx.x; q < 1024; q += blockDim.x) s_bm[q] = 0u;
    __syncthreads();
    int org[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        int m = s_wmin[0][a];
#pragma unroll
        for (int w = 1; w < 8; w++) m = min(m, s_wmin[w][a]);
        org[a] = m;
    }
#pragma unroll
    for (int k = 0; k < SPT; k++) {
        const uint64_t pk = __shfl_up_sync(0xffffffffu, bkey[k], 1);
        const uint32_t pe = __shfl_up_sync(0xffffffffu, ecode[k], 1);
        if (ecode[k] && (lane == 0 || pk != bkey[k] || pe != ecode[k])) {
            int lo0, lo1, lo2;
            b2r_unpack_key(bkey[k], lo0, lo1, lo2);
            const int e0 = (int)(ecode[k] & 1023u), e1 = (int)((ecode[k] >> 10) & 1023u), e2 = (int)(ecode[k] >> 20);
            const int rz = lo2 - org[2];                      // >= 0
            const int nz_in = max(0, min(e2, 32 - rz));      // cells of the z run inside the window
            const uint32_t zmask = nz_in >= 32 ? 0xffffffffu : (((1u << nz_in) - 1u) << (rz & 31));
            for (int a = 0; a < e0; a++)
                for (int b = 0; b < e1; b++) {
                    const int rx = lo0 + a - org[0], ry = lo1 + b - org[1];
                    const bool in_xy = rx < 32 && ry < 32;
                    if (in_xy && nz_in > 0) atomicOr(&s_bm[rx * 32 + ry], zmask);
                    for (int c = in_xy ? nz_in : 0; c < e2; c++) {          // outside the window: direct (rare)
                        const uint64_t key = b2r_pack_key(lo0 + a, lo1 + b, lo2 + c);
                        if (P.world <= 1 || b2r_owner(key, P.world) == P.rank) alloc_key(P, key, F.bit);
                    }
                }
        }
    }
    __syncthreads();
    // compaction: thread t owns words 4 t .. 4 t + 3; set bits -> consecutive entries of s_list (cell id = word * 32 + bit)
    uint32_t wv[4];
    int cnt = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) { wv[q] = s_bm[4 * threadIdx.x + q]; cnt += __popc(wv[q]); }
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if ((int)lane >= o) incl += v; }
    if (lane == 31) s_wcnt[warp] = incl;
    __syncthreads();
    int pos0 = incl - cnt, nkeys = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) { if (w < warp) pos0 += s_wcnt[w]; nkeys += s_wcnt[w]; }
#pragma unroll
    for (int q = 0; q < 4; q++)
        for (uint32_t m = wv[q]; m; m &= m - 1) {
            const int cell = (4 * (int)threadIdx.x + q) * 32 + (__ffs(m) - 1);
            if (pos0 < ALLOC_LIST) {
                s_list[pos0] = (uint16_t)cell;
            } else {                                                           // more keys than the list holds: direct (rare)
                const uint64_t key = b2r_pack_key(org[0] + (cell >> 10), org[1] + ((cell >> 5) & 31), org[2] + (cell & 31));
                if (P.world <= 1 || b2r_owner(key, P.world) == P.rank) alloc_key(P, key, F.bit);
            }
            pos0++;
        }
    __syncthreads();
    nkeys = min(nkeys, ALLOC_LIST);
    K1_MARK(2);
    // (3) global probes: every key of the tile once, ALLOC_IT per thread in flight together
    for (int c0 = 0; c0 < nkeys; c0 += 256 * ALLOC_IT) {
        uint64_t want[ALLOC_IT], got[ALLOC_IT];
        uint32_t slot_i[ALLOC_IT], mk[ALLOC_IT], old[ALLOC_IT];
#pragma unroll
        for (int it = 0; it < ALLOC_IT; it++) {
            const int c = c0 + it * 256 + (int)threadIdx.x;
            want[it] = B2R_KEY_EMPTY; got[it] = 0; mk[it] = 0; slot_i[it] = 0;
            if (c < nkeys) {
                const int cell = s_list[c];
                const uint64_t key = b2r_pack_key(org[0] + (cell >> 10), org[1] + ((cell >> 5) & 31), org[2] + (cell & 31));
                const uint64_t hsh = b2r_mix64(key);
                if (P.world <= 1 || (int)((uint32_t)(hsh >> 33) % (uint32_t)P.world) == P.rank) {
                    want[it] = key;
                    slot_i[it] = (uint32_t)hsh & P.hash.mask;
                    const B2RSlot* slot = &P.hash.slots[slot_i[it]];
                    got[it] = __ldcg(reinterpret_cast<const unsigned long long*>(&slot->key));
                    mk[it] = __ldcg(&slot->mask[P.q]);
                }
            }
        }
        unsigned first = 0;                          // bit it: this lane saw the first touch of (unit, batch)
#pragma unroll
        for (int it = 0; it < ALLOC_IT; it++) {
            old[it] = 1u;
            const bool live = want[it] != B2R_KEY_EMPTY && !(got[it] == want[it] && (mk[it] & F.bit));
            if (live) {
                if (got[it] == want[it]) { old[it] = atomicOr(&P.hash.slots[slot_i[it]].mask[P.q], F.bit); first |= 1u << it; }
                else alloc_key(P, want[it], F.bit);
            }
        }
#pragma unroll
        for (int it = 0; it < ALLOC_IT; it++) {
            const bool app = ((first >> it) & 1u) && old[it] == 0u;
            const unsigned am = __ballot_sync(0xffffffffu, app);
            if (am == 0u) continue;
            int pos = 0;
            if (lane == (unsigned)(__ffs(am) - 1)) pos = atomicAdd(&P.counters[B2R_C_TOUCHED0 + P.q], __popc(am));
            pos = __shfl_sync(0xffffffffu, pos, __ffs(am) - 1) + __popc(am & ((1u << lane) - 1u));
            if (app && pos < P.cap_units) P.touched[pos] = (int)slot_i[it];
        }
    }
    K1_MARK(4);
}

// ------------------------------------------------------------------------------------------------
// K2: fuse
// ------------------------------------------------------------------------------------------------
#ifndef FUSE_MIN_CTAS
#define FUSE_MIN_CTAS 4   // 64 registers/thread: 32 resident warps per SM (measured best; 3 and 5 are slower)
#endif

struct FuseParams {
    int nframes;
    int q;                   // batch buffer being fused; (q+2)%3 = previous batch: its masks get cleared by this pass and its
                             // counters are zeroed by the pass's last CTA (the buffer's next user is the batch after next --
                             // (q+1)%3 may already be filling: the next batch's K1 can run beside this pass)
    int ring_rounds;         // passes with at least this many items per CTA run the prefetching ring
    int reverse;             // sweep direction (alternates per batch: the tail of the last sweep is still in L2)
    float vl, hv, trunc, inv_trunc;
    double unit_length;
    const int32_t* touched;
    const int32_t* touched_prev;
    B2RSlot* slots;
    float* pool;
    int64_t unit_stride;     // floats per unit
    int32_t* counters;
    int cap_units;
    B2RFrame fr[B2R_MAX_BATCH];
};

template <int COLOR, int FUSE_STAGES, int XL>
__global__ void __launch_bounds__(64 * XL, FUSE_MIN_CTAS * 4 / XL) tsdf_fuse_kernel(const __grid_constant__ FuseParams P) {
    constexpr int PLANES = COLOR ? 5 : 2;
    // XL = x-layers per work item (4: 256 threads, 4 items per unit; 2: 128 threads, 8 items per unit -- the finer
    // quantum keeps a sharded rank's short passes balanced over the persistent grid)
    constexpr int FUSE_THREADS = 64 * XL, FUSE_SLAB = 256 * XL, IS = XL == 4 ? 2 : 3, IM = (1 << IS) - 1;
    extern __shared__ __align__(128) unsigned char fuse_smem[];
    float* stage = reinterpret_cast<float*>(fuse_smem);                        // [FUSE_STAGES][PLANES][FUSE_SLAB]
    __shared__ __align__(8) uint64_t full_bar[FUSE_STAGES];

    const int tid = threadIdx.x;
    const int T = min(P.counters[B2R_C_TOUCHED0 + P.q], P.cap_units);
    const int nwork = T << IS;

    // ---- housekeeping for the rotating batch buffers
    {
        const int qp = (P.q + 2) % 3;
        const int Tp = min(P.counters[B2R_C_TOUCHED0 + qp], P.cap_units);
        for (int i = blockIdx.x * FUSE_THREADS + tid; i < Tp; i += gridDim.x * FUSE_THREADS) P.slots[P.touched_prev[i]].mask[qp] = 0u;
        if (blockIdx.x == 0 && tid == 0) P.counters[B2R_C_LAST_TOUCHED] = T;
    }
    if (tid == 0) {
        for (int s = 0; s < FUSE_STAGES; s++) mbar_init(&full_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const int zq = tid & 3, y = (tid >> 2) & 15, xl = tid >> 6;
    const int f4 = tid;                          // this thread's float4 inside a 1024-float slab plane: 4 consecutive z voxels
    unsigned int upd = 0, pairs = 0;
    uint32_t it_seq = 0;   // items issued through the ring so far by this CTA (stage = seq % STAGES)

    // ---- dynamic work distribution: items differ in cost (1..16 frames each), so CTAs draw items from a counter as they
    // finish instead of owning a fixed stride of the list.  Long passes (>= ring_rounds items per CTA) run the TMA ring:
    // the next item is claimed and in flight while the current one is fused, and thread 0 also pre-claims the index and
    // touched-list entry of the item after that (consumed one fetch later, so a fetch waits for ONE L2 round trip) -- but
    // only while the list is far from its end, so the tail is not hoarded.  Short passes (a sharded rank sees ~2 items
    // per CTA) claim strictly one item at a time: balance matters more there than the exposed load, which the other
    // resident CTAs of the SM cover.
    __shared__ int s_us[FUSE_STAGES];            // per ring stage: (unit << IS) | slab, or -1 = the list is exhausted
    __shared__ uint32_t s_fm[FUSE_STAGES];
    __shared__ int s_k[FUSE_STAGES][3];
    int* const work_ctr = &P.counters[B2R_C_WORK0 + P.q];
    const bool ring = nwork >= P.ring_rounds * (int)gridDim.x;
    const int look_end = ring ? nwork - 3 * (int)gridDim.x : 0;   // pre-claim only items before this index
    int w1 = -1, sl1 = 0;                        // thread 0: pre-claimed item (-1 = none) and its slot index
    auto fetch = [&](uint32_t st) {              // thread 0: next live item -> stage st (info + TMA), or -1
        for (;;) {
            int cw = w1, csl = sl1;
            w1 = -1;
            if (cw < 0) {
                cw = atomicAdd(work_ctr, 1);
                if (cw < nwork) csl = __ldcg(&P.touched[(P.reverse ? nwork - 1 - cw : cw) >> IS]);
            }
            if (cw >= nwork) { s_us[st] = -1; return; }
            const B2RSlot* slot = &P.slots[csl];
            const uint4 lo = __ldcg(reinterpret_cast<const uint4*>(slot));
            const uint4 hi = __ldcg(reinterpret_cast<const uint4*>(slot) + 1);
            if (cw < look_end) {
                w1 = atomicAdd(work_ctr, 1);
                if (w1 < nwork) sl1 = __ldcg(&P.touched[(P.reverse ? nwork - 1 - w1 : w1) >> IS]);
            }
            const uint32_t m = P.q == 0 ? lo.w : (P.q == 1 ? hi.x : hi.y);
            if ((int)lo.z < 0 || m == 0u) continue;    // unit without storage (pool overflow) or nothing pending
            const int w = P.reverse ? nwork - 1 - cw : cw;
            int kx, ky, kz;
            b2r_unpack_key(((uint64_t)lo.y << 32) | lo.x, kx, ky, kz);
            s_us[st] = ((int)lo.z << IS) | (w & IM); s_fm[st] = m;
            s_k[st][0] = kx; s_k[st][1] = ky; s_k[st][2] = kz;
            const float* src = P.pool + (int64_t)(int)lo.z * P.unit_stride + (w & IM) * FUSE_SLAB;
            mbar_expect_tx(&full_bar[st], PLANES * FUSE_SLAB * 4);
#pragma unroll
            for (int pl = 0; pl < PLANES; pl++) tma_load_1d(stage + (st * PLANES + pl) * FUSE_SLAB, src + pl * B2R_UNIT_VOX, FUSE_SLAB * 4, &full_bar[st]);
            return;
        }
    };
    if (tid == 0) {
        for (int s = 0; s < (ring ? FUSE_STAGES : 1); s++) fetch(s);
    }
    __syncthreads();   // barriers initialised, first items published
    {
        for (;; it_seq++) {
            const uint32_t st = it_seq % FUSE_STAGES;
            const int us = s_us[st];
            if (us < 0) break;
            const uint32_t fmask = s_fm[st];
            {
                const int unit = us >> IS, x = (us & IM) * XL + xl;
                const int kx = s_k[st][0], ky = s_k[st][1], kz = s_k[st][2];
                // voxel-centre column base: float(hv + vl*x + origin.x) with the f32->f64 promotion of upstream
                const float px = (float)xdadd((double)xfadd(P.hv, xfmul(P.vl, (float)x)), xdmul((double)kx, P.unit_length));
                const float py = (float)xdadd((double)xfadd(P.hv, xfmul(P.vl, (float)y)), xdmul((double)ky, P.unit_length));
                const float pz = (float)xdadd((double)P.hv, xdmul((double)kz, P.unit_length));
                mbar_wait(&full_bar[st], (it_seq / FUSE_STAGES) & 1u);
                const float* sb = stage + st * PLANES * FUSE_SLAB + f4 * 4;
                float4 ts = *reinterpret_cast<const float4*>(sb);
                float4 ws = *reinterpret_cast<const float4*>(sb + FUSE_SLAB);
                float4 rs, gs, bs;
                if (COLOR) {
                    rs = *reinterpret_cast<const float4*>(sb + 2 * FUSE_SLAB);
                    gs = *reinterpret_cast<const float4*>(sb + 3 * FUSE_SLAB);
                    bs = *reinterpret_cast<const float4*>(sb + 4 * FUSE_SLAB);
                }
                float* tp = reinterpret_cast<float*>(&ts);
                float* wp = reinterpret_cast<float*>(&ws);
                float* rp = reinterpret_cast<float*>(&rs);
                float* gp = reinterpret_cast<float*>(&gs);
                float* bp = reinterpret_cast<float*>(&bs);
                const unsigned upd0 = upd;
                pairs += (tid == 0 && (us & IM) == 0) ? __popc(fmask) : 0;
                // fmask is the same in every lane; the warp reduction says so to the compiler, which then keeps the frame loop
                // and the frame's constants in the uniform datapath (LDCU into uniform registers, not one LDC per value per thread)
                for (uint32_t fm = __reduce_or_sync(0xffffffffu, fmask); fm; fm &= fm - 1) {
                    const B2RFrame& F = P.fr[__ffs(fm) - 1];
                    float pcx = xfadd(xfadd(xfadd(xfmul(F.E[0], px), xfmul(F.E[1], py)), xfmul(F.E[2], pz)), F.E[3]);
                    float pcy = xfadd(xfadd(xfadd(xfmul(F.E[4], px), xfmul(F.E[5], py)), xfmul(F.E[6], pz)), F.E[7]);
                    float pcz = xfadd(xfadd(xfadd(xfmul(F.E[8], px), xfmul(F.E[9], py)), xfmul(F.E[10], pz)), F.E[11]);
                    // sequential z-march to this thread's first voxel (same f32 accumulation as the reference
                    // loop): 4*zq steps, written as three predicated groups of four so there is no loop
#pragma unroll
                    for (int g = 0; g < 3; g++) {
                        if (zq > g) {
#pragma unroll
                            for (int s = 0; s < 4; s++) { pcx = xfadd(pcx, F.Es02); pcy = xfadd(pcy, F.Es12); pcz = xfadd(pcz, F.Es22); }
                        }
                    }
                    // pass 1: project the 4 voxels, issue the 4 texel gathers together
                    float zc[4];
                    float2 tx[4];
                    float mk[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        zc[k] = pcz;
                        tx[k] = make_float2(0.f, 0.f);
                        mk[k] = 0.f;
                        {
                            // u_f = pc.x*fx/pc.z + cx + 0.5f, v_f likewise: both quotients share RN(1/pc.z).
                            // Inside b2r_div_domain (2^-60 <= pc.z < 2^60, which implies pc.z > 0) the Markstein quotient is
                            // RN(a/b) unless |a| is so small that r underflows, and then |q| < 2^-40 cannot change
                            // RN(q + cx) -- the pixel picked is the same either way.  The fast form is computed without a
                            // branch (its value is unused outside the domain); a positive pc.z outside the domain takes
                            // div.rn on a cold path; pc.z <= 0 or NaN never projects.
                            const bool dom = b2r_div_domain(pcz);
                            const float ax = xfmul(pcx, F.fx), ay = xfmul(pcy, F.fy);
                            const float inv = b2r_rcp_markstein(pcz);
                            float qx = b2r_div_fast(ax, pcz, inv), qy = b2r_div_fast(ay, pcz, inv);
                            bool ok = dom;
                            if (__builtin_expect(!dom && pcz > 0.0f, 0)) { qx = __fdiv_rn(ax, pcz); qy = __fdiv_rn(ay, pcz); ok = true; }
                            const float u_f = xfadd(xfadd(qx, F.cx), 0.5f);
                            const float v_f = xfadd(xfadd(qy, F.cy), 0.5f);
                            const bool in = ok & (u_f >= 0.0001f) & (u_f < F.safe_w) & (v_f >= 0.0001f) & (v_f < F.safe_h);
                            if (in) {
                                const int pix = (int)v_f * F.W + (int)u_f;
                                tx[k] = __ldg(&F.texel[pix]);
                                mk[k] = __ldg(&F.mult[pix]);       // frame-independent, L1/L2 resident
                            }
                        }
                        pcx = xfadd(pcx, F.Es02); pcy = xfadd(pcy, F.Es12); pcz = xfadd(pcz, F.Es22);
                    }
                    // pass 2: running averages in registers.  w1 = weight + 1 is an integer in [1, 2^24]: always inside
                    // the division domain except 2^24 - 1 (all-ones significand); numerators are 0 or >= 2^-60.
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        {
                            const float sdf = xfmul(xfsub(tx[k].x, zc[k]), mk[k]);
                            if ((tx[k].x > 0.0f) & (sdf > -P.trunc)) {
                                const float tv = fminf(1.0f, xfmul(sdf, P.inv_trunc));
                                const float wgt = wp[k];
                                const float w1 = xfadd(wgt, 1.0f);
                                float c0 = 0.f, c1 = 0.f, c2 = 0.f;
                                if (COLOR == 1) {
                                    const uint32_t c = __float_as_uint(tx[k].y);
                                    c0 = byte_to_float<0>(c); c1 = byte_to_float<1>(c); c2 = byte_to_float<2>(c);
                                } else if (COLOR == 2) {
                                    c0 = c1 = c2 = tx[k].y;
                                }
                                const float at = xfadd(xfmul(tp[k], wgt), tv);
                                float ar = 0.f, ag = 0.f, ab = 0.f;
                                if (COLOR) { ar = xfadd(xfmul(rp[k], wgt), c0); ag = xfadd(xfmul(gp[k], wgt), c1); ab = xfadd(xfmul(bp[k], wgt), c2); }
                                {
                                    const float inv = b2r_rcp_markstein(w1);
                                    tp[k] = b2r_div_fast(at, w1, inv);
                                    if (COLOR) { rp[k] = b2r_div_fast(ar, w1, inv); gp[k] = b2r_div_fast(ag, w1, inv); bp[k] = b2r_div_fast(ab, w1, inv); }
                                }
                                if (__builtin_expect(!b2r_div_domain(w1), 0)) {     // cold: w1 = 2^24 - 1, or a non-integer weight imported into the pool
                                    tp[k] = __fdiv_rn(at, w1);
                                    if (COLOR) { rp[k] = __fdiv_rn(ar, w1); gp[k] = __fdiv_rn(ag, w1); bp[k] = __fdiv_rn(ab, w1); }
                                }
                                wp[k] = w1;
                                upd++;
                            }
                        }
                    }
                }
                if (upd != upd0) {
                    float* base = P.pool + (int64_t)unit * P.unit_stride + (us & IM) * FUSE_SLAB + f4 * 4;
                    *reinterpret_cast<float4*>(base) = ts;
                    *reinterpret_cast<float4*>(base + B2R_UNIT_VOX) = ws;
                    if (COLOR) {
                        *reinterpret_cast<float4*>(base + 2 * B2R_UNIT_VOX) = rs;
                        *reinterpret_cast<float4*>(base + 3 * B2R_UNIT_VOX) = gs;
                        *reinterpret_cast<float4*>(base + 4 * B2R_UNIT_VOX) = bs;
                    }
                }
            }
            __syncthreads();   // every thread has read its float4s and the info of this stage -> the stage is free
            if (tid == 0) fetch(ring ? st : (it_seq + 1) % FUSE_STAGES);
            if (!ring) __syncthreads();   // one item at a time: the claim just made is the next iteration's
        }
    }
    // N_upd accounting: warp shuffle -> one atomic per warp
    for (int o = 16; o > 0; o >>= 1) upd += __shfl_down_sync(0xffffffffu, upd, o);
    if ((tid & 31) == 0) {
        if (upd) atomicAdd(reinterpret_cast<unsigned long long*>(&P.counters[B2R_C_UPDATED]), (unsigned long long)upd);
        if (tid == 0 && pairs) atomicAdd(reinterpret_cast<unsigned long long*>(&P.counters[B2R_C_TOUCHED_TOTAL]), (unsigned long long)pairs);
    }
    // The last CTA out recycles the previous batch's buffer: every CTA has read its touched count (the mask-clearing loop
    // above) before it takes a ticket.
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(&P.counters[B2R_C_TICKET], 1) == (int)gridDim.x - 1) {
            const int qp = (P.q + 2) % 3;
            P.counters[B2R_C_TOUCHED0 + qp] = 0;
            P.counters[B2R_C_WORK0 + qp] = 0;      // the pass after next draws its items from a fresh counter
            P.counters[B2R_C_TICKET] = 0;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// handle lifecycle
// ------------------------------------------------------------------------------------------------
static uint32_t next_pow2(uint64_t v) {
    uint64_t p = 1;
    while (p < v) p <<= 1;
    return (uint32_t)p;
}

static int env_int(const char* name, int dflt, int lo, int hi) {
    const char* s = getenv(name);
    if (!s || !*s) return dflt;
    const int v = atoi(s);
    return v < lo ? lo : (v > hi ? hi : v);
}

static void tsdf_free(b2r_tsdf* h) {
    if (!h) return;
    cudaFree(h->hash.slots); cudaFree(h->unit_keys);
    cudaFree(h->pool); cudaFree(h->counters); cudaFree(h->scratch);
    for (int i = 0; i < 3; i++) cudaFree(h->touched[i]);
    cudaFree(h->touched_sorted);
    for (int i = 0; i < B2R_MAX_BATCH; i++) { cudaFree(h->texel[0][i]); cudaFree(h->texel[1][i]); }
    if (h->fuse_stream) cudaStreamDestroy(h->fuse_stream);
    if (h->ev_k1) cudaEventDestroy(h->ev_k1);
    for (int i = 0; i < 3; i++) if (h->ev_k2[i]) cudaEventDestroy(h->ev_k2[i]);
    cudaFree(h->mult);
    for (int i = 0; i < B2R_HOST_RING; i++) {
        if (h->ev_copied[i]) cudaEventDestroy(h->ev_copied[i]);
        if (h->ev_consumed[i]) cudaEventDestroy(h->ev_consumed[i]);
        cudaFree(h->stage_depth[i]); cudaFree(h->stage_color[i]);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (int i = 0; i < 2 * h->prof_cap; i++) cudaEventDestroy(h->prof_ev[i]);
    delete[] h->prof_ev;
    delete[] h->prof_kind;
    delete h;
}

typedef void (*fuse_fn_t)(const FuseParams);
template <int XL>
static fuse_fn_t fuse_fn_xl(int color_type, int stages) {
    if (stages >= 3) return color_type == 0 ? tsdf_fuse_kernel<0, 3, XL> : (color_type == 1 ? tsdf_fuse_kernel<1, 3, XL> : tsdf_fuse_kernel<2, 3, XL>);
    return color_type == 0 ? tsdf_fuse_kernel<0, 2, XL> : (color_type == 1 ? tsdf_fuse_kernel<1, 2, XL> : tsdf_fuse_kernel<2, 2, XL>);
}
static fuse_fn_t fuse_fn(int color_type, int stages, int xl) {
    return xl == 2 ? fuse_fn_xl<2>(color_type, stages) : fuse_fn_xl<4>(color_type, stages);
}

B2R_API b2r_tsdf* b2r_tsdf_create(double voxel_length, double sdf_trunc, int color_type, int unit_resolution,
                                  int depth_sampling_stride, int64_t capacity_units, int rank, int world) {
    if (unit_resolution != B2R_UNIT_RES) { b2r_set_error("volume_unit_resolution must be 16 (got %d)", unit_resolution); return nullptr; }
    if (!(voxel_length > 0) || !(sdf_trunc > 0) || depth_sampling_stride < 1 || capacity_units < 1 || color_type < 0 ||
        color_type > 2 || world < 1 || rank < 0 || rank >= world || capacity_units > (1 << 28)) {
        b2r_set_error("b2r_tsdf_create: bad argument");
        return nullptr;
    }
    b2r_tsdf* h = new (std::nothrow) b2r_tsdf();
    if (!h) return nullptr;
    memset(h, 0, sizeof(*h));
    h->generation = 1;   // mesh_gen == 0 never matches: the extraction cache starts invalid
    h->sort_n = -1;
    h->voxel_length = voxel_length; h->sdf_trunc = sdf_trunc;
    h->unit_length = voxel_length * unit_resolution;
    h->color_type = color_type; h->res = unit_resolution; h->stride = depth_sampling_stride;
    h->planes = color_type == 0 ? 2 : 5;
    h->cap_units = capacity_units; h->rank = rank; h->world = world;
    h->max_batch = env_int("B2R_TSDF_BATCH", B2R_MAX_BATCH, 1, B2R_MAX_BATCH);
    cudaError_t e = cudaGetDevice(&h->device);
    h->nslots = next_pow2((uint64_t)capacity_units * 2);
    h->hash.mask = h->nslots - 1;
    const size_t pool_bytes = (size_t)capacity_units * h->planes * B2R_UNIT_VOX * sizeof(float);
    auto A = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
    A((void**)&h->hash.slots, (size_t)h->nslots * sizeof(B2RSlot));
    A((void**)&h->unit_keys, (size_t)capacity_units * 12);
    for (int i = 0; i < 3; i++) A((void**)&h->touched[i], (size_t)capacity_units * 4);
    A((void**)&h->counters, B2R_C_WORDS * 4);
    A((void**)&h->pool, pool_bytes);
    if (e == cudaSuccess) e = cudaMemset(h->pool, 0, pool_bytes);
    if (e == cudaSuccess) e = cudaMemset(h->counters, 0, B2R_C_WORDS * 4);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking);
    for (int i = 0; i < B2R_HOST_RING && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&h->ev_copied[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_consumed[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) {
        const int stages = env_int("B2R_FUSE_STAGES", 2, 2, 3);
        // work quantum of K2: 2 x-layers per item (128-thread CTAs, 8 per SM).  With the heaviest-first item order the finer quantum is
        // the better one at every world size (us/frame of K2, 4 -> 2 layers: 31.9 -> 31.7 at N = 1, 8.66 -> 8.55 at 4, 4.74 -> 4.61 at 8)
        h->fuse_xl = env_int("B2R_FUSE_XL", 2, 2, 4) == 2 ? 2 : 4;
        h->fuse_ring_rounds = env_int("B2R_FUSE_RING_ROUNDS", 6, 0, 1 << 20);
        // heaviest-first ordering of a multi-frame pass's items: + 5 % on a rank of 8 (2 - 3 items per CTA), + 2 % at one GPU
        h->sort_touched = env_int("B2R_FUSE_SORT", 1, 0, 1);
        if (h->sort_touched && e == cudaSuccess) e = cudaMalloc((void**)&h->touched_sorted, (size_t)capacity_units * 4);
        h->fuse_smem = stages * h->planes * 256 * h->fuse_xl * 4;
        fuse_fn_t fn = fuse_fn(color_type, stages, h->fuse_xl);
        e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, h->fuse_smem);
        int nsm = 148, per_sm = 1;
        cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, h->device);
        h->n_sm = nsm;
        if (e == cudaSuccess) e = cudaFuncSetAttribute(tsdf_pack_alloc_kernel<ALLOC_SPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, PACK_STAGES * PACK_TILE * 8 + 256);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(tsdf_pack_alloc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, PACK_STAGES * PACK_TILE * 8 + 256);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 64 * h->fuse_xl, h->fuse_smem);
        if (per_sm < 1) per_sm = 1;
        // B2R_FUSE_CTAS_PER_SM: fewer resident fusion CTAs than fit leave registers / shared memory on every SM for what runs
        // beside the pass (the next batch's K1 on the side stream, NCCL's gather kernels); default: all that fit
        // Overlap mode is OFF by default: measured (profiles/r2_overlap_probe.json) it loses -- the pass slows by more than K1
        // costs (K1 streams 150 MB per batch through the L2 the pass gathers its texels from: 5.1 -> 6.5 us/frame on a rank of
        // 8, 32.4 -> 35.6 at one GPU) -- so K1 and K2 stay back to back on the caller's stream.
        h->overlap = env_int("B2R_TSDF_OVERLAP", 0, 0, 1);
        // overlap mode keeps one CTA slot per SM free for what runs beside a pass (K1 of the next batch)
        per_sm = env_int("B2R_FUSE_CTAS_PER_SM", h->overlap && per_sm > 1 ? per_sm - 1 : per_sm, 1, per_sm);
        h->fuse_grid = nsm * per_sm;
        if (h->overlap) {
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->fuse_stream, cudaStreamNonBlocking);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_k1, cudaEventDisableTiming);
            for (int i = 0; i < 3 && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&h->ev_k2[i], cudaEventDisableTiming);
        }
    }
    if (e == cudaSuccess && b2r_tsdf_reset(h, 0) != B2R_OK) e = cudaErrorUnknown;
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        b2r_set_error("b2r_tsdf_create: %s (pool %.1f MiB)", cudaGetErrorString(e), pool_bytes / 1048576.0);
        tsdf_free(h);
        return nullptr;
    }
    return h;
}

B2R_API void b2r_tsdf_destroy(b2r_tsdf* h) {
    if (!h) return;
    B2RDeviceGuard g(h->device);
    cudaDeviceSynchronize();
    tsdf_free(h);
}

__global__ void tsdf_clear_slots_kernel(B2RSlot* slots, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    B2RSlot s;
    s.key = B2R_KEY_EMPTY; s.val = -1; s.mask[0] = s.mask[1] = s.mask[2] = 0u; s.pad[0] = s.pad[1] = 0u;
    slots[i] = s;
}

// zeroes the units allocated so far; the count is read ON THE DEVICE, so reset() needs no host round trip
__global__ void tsdf_clear_pool_kernel(float4* pool, int64_t unit_f4, const int32_t* counters, int cap) {
    const int n = min(counters[B2R_C_NUM_UNITS], cap);
    const int64_t total = (int64_t)n * unit_f4;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) pool[i] = z;
}

// volume.reset()  (bundlefusion_reconstruction.py:309, capture_frames.py:94).  Pending frames are dropped:
// integrate-then-reset leaves an empty volume either way.  Fully asynchronous on `stream` (three launches, no copy back).
B2R_API int b2r_tsdf_reset(b2r_tsdf* h, cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int nsm = 148;
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, h->device);
    if (h->overlap && h->unjoined) {     // passes still running on the internal stream read what is about to be cleared
        B2R_CUDA(cudaStreamWaitEvent(stream, h->ev_k2[(h->batches + 2) % 3], 0));
        h->unjoined = 0;
    }
    tsdf_clear_pool_kernel<<<nsm * 8, 256, 0, stream>>>(reinterpret_cast<float4*>(h->pool), (int64_t)h->planes * (B2R_UNIT_VOX / 4), h->counters, (int)h->cap_units);
    B2R_CHECK_LAUNCH();
    tsdf_clear_slots_kernel<<<(h->nslots + 255) / 256, 256, 0, stream>>>(h->hash.slots, h->nslots);
    B2R_CHECK_LAUNCH();
    B2R_CUDA(cudaMemsetAsync(h->counters, 0, B2R_C_WORDS * 4, stream));
    h->q = 0; h->pending = 0; h->batches = 0;
    h->generation++; h->unit_epoch++;
    return B2R_OK;
}

int b2r_tsdf_ensure_scratch(b2r_tsdf* h, size_t bytes) {
    if (h->scratch_bytes >= bytes) return B2R_OK;
    if (h->scratch) cudaFree(h->scratch);
    h->scratch = nullptr; h->scratch_bytes = 0;
    B2R_CUDA(cudaMalloc(&h->scratch, bytes));
    h->scratch_bytes = bytes;
    return B2R_OK;
}

// ---- profiling hooks: an event pair around every kernel launch while enabled -----------------------
static bool prof_begin(b2r_tsdf* h, int kind, cudaStream_t stream) {
    if (!h->profile || h->prof_n >= h->prof_cap) return false;
    h->prof_kind[h->prof_n] = kind;
    cudaEventRecord(h->prof_ev[2 * h->prof_n], stream);
    return true;
}
static void prof_end(b2r_tsdf* h, bool on, cudaStream_t stream) {
    if (!on) return;
    cudaEventRecord(h->prof_ev[2 * h->prof_n + 1], stream);
    h->prof_n++;
}

// ------------------------------------------------------------------------------------------------
// K2 launch: fuse the pending batch
// ------------------------------------------------------------------------------------------------
// Short passes (a sharded rank: 2 - 3 work items per fusion CTA, each costing 1 .. 32 frames) are bounded by the CTA that
// happens to draw the heaviest items last.  One small CTA orders the touched list by the number of frames that touch the
// unit, heaviest first (counting sort on popc(mask)), so the counter hands out the long items first and the short ones fill
// the gaps (longest-processing-time-first list scheduling).
__global__ void __launch_bounds__(1024) tsdf_sort_touched_kernel(const int32_t* __restrict__ touched, int32_t* __restrict__ sorted,
                                                                 const B2RSlot* __restrict__ slots, const int32_t* __restrict__ counters, int q, int cap) {
    __shared__ int hist[33], cursor[33];
    const int T = min(counters[B2R_C_TOUCHED0 + q], cap);
    if (threadIdx.x < 33) hist[threadIdx.x] = 0;
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += blockDim.x) atomicAdd(&hist[32 - __popc(slots[touched[t]].mask[q])], 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int b = 0; b < 33; b++) { cursor[b] = acc; acc += hist[b]; }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const int sl = touched[t];
        sorted[atomicAdd(&cursor[32 - __popc(slots[sl].mask[q])], 1)] = sl;
    }
}

// Launches the pass over the pending batch.  Overlap mode: on the internal stream, after everything the caller's stream has
// queued so far (the batch's K1 launches); the caller's stream does NOT wait for it.
static int tsdf_launch_fuse(b2r_tsdf* h, cudaStream_t stream) {
    if (!h->pending) return B2R_OK;
    FuseParams F;
    memset(&F, 0, sizeof(F));
    F.nframes = h->pending; F.q = h->q; F.reverse = (int)(h->batches & 1);
    F.ring_rounds = h->fuse_ring_rounds;
    F.vl = (float)h->voxel_length; F.hv = F.vl * 0.5f;
    F.trunc = (float)h->sdf_trunc; F.inv_trunc = 1.0f / F.trunc;
    F.unit_length = h->unit_length;
    F.touched = h->touched[h->q]; F.touched_prev = h->touched[(h->q + 2) % 3];
    F.slots = h->hash.slots; F.pool = h->pool; F.unit_stride = (int64_t)h->planes * B2R_UNIT_VOX;
    F.counters = h->counters; F.cap_units = (int)h->cap_units;
    for (int i = 0; i < h->pending; i++) F.fr[i] = h->frames[i];
    const int stages = h->fuse_smem / (h->planes * 256 * h->fuse_xl * 4);
    cudaStream_t fs = stream;
    if (h->overlap) {
        fs = h->fuse_stream;
        B2R_CUDA(cudaEventRecord(h->ev_k1, stream));
        B2R_CUDA(cudaStreamWaitEvent(fs, h->ev_k1, 0));
    }
    if (h->sort_touched && h->pending > 1) {
        tsdf_sort_touched_kernel<<<1, 1024, 0, fs>>>(h->touched[h->q], h->touched_sorted, h->hash.slots, h->counters, h->q, (int)h->cap_units);
        B2R_CHECK_LAUNCH();
        F.touched = h->touched_sorted;     // the same set, heaviest units first; the mask clearing of the pass after next uses either
        F.reverse = 0;
    }
    const bool prof = prof_begin(h, 2, fs);
    fuse_fn(h->color_type, stages, h->fuse_xl)<<<h->fuse_grid, 64 * h->fuse_xl, h->fuse_smem, fs>>>(F);
    B2R_CHECK_LAUNCH();
    prof_end(h, prof, fs);
    if (h->overlap) {
        B2R_CUDA(cudaEventRecord(h->ev_k2[h->batches % 3], fs));
        h->unjoined = 1;
    }
    h->q = (h->q + 1) % 3;
    h->pending = 0;
    h->batches++;
    return B2R_OK;
}

// Fuses whatever is pending AND makes `stream` wait for every pass in flight: what every reader of the volume calls first.
int b2r_tsdf_flush_pending(b2r_tsdf* h, cudaStream_t stream) {
    int rc = tsdf_launch_fuse(h, stream);
    if (rc != B2R_OK) return rc;
    if (h->overlap && h->unjoined) {
        B2R_CUDA(cudaStreamWaitEvent(stream, h->ev_k2[(h->batches + 2) % 3], 0));   // passes run in order: the last one covers all
        h->unjoined = 0;
    }
    return B2R_OK;
}

// ------------------------------------------------------------------------------------------------
// integrate (device inputs): K1 now, K2 when the batch is full or the volume is read
// ------------------------------------------------------------------------------------------------
static bool host_fdiv_domain(float b) {    // b2r_div_domain on the host
    uint32_t u; memcpy(&u, &b, 4);
    return ((u >> 23) - 67u) < 120u && (u & 0x7FFFFFu) != 0x7FFFFFu;
}
static bool host_ddiv_domain(double b) {   // positive, 2^-60 <= b < 2^60, significand not all ones
    uint64_t u; memcpy(&u, &b, 8);
    const uint64_t e = u >> 52, m = u & 0xFFFFFFFFFFFFFull;
    return e >= 1023 - 60 && e < 1023 + 60 && m != 0xFFFFFFFFFFFFFull;
}
// Frames are either stored back to back (depth / color / gray = base pointers) or scattered (depth_ptrs / color_ptrs = HOST
// arrays of nf device pointers, e.g. the slots of an all-gather buffer in the order they are to be fused).
static int tsdf_integrate_impl(b2r_tsdf* h, int nf, const void* depth, int depth_is_u16, float depth_scale, float depth_trunc,
                               const uint8_t* color, int color_channels, int color_bgr, const float* gray, int W,
                               int H, double fx, double fy, double cx, double cy, const double* extrinsics,
                               cudaStream_t stream, const void* const* depth_ptrs = nullptr, const uint8_t* const* color_ptrs = nullptr) {
    if (depth_ptrs) { depth = depth_ptrs[0]; if (color_ptrs) color = color_ptrs[0]; }
    if (!h || !depth || !extrinsics || W < 1 || H < 1 || nf < 1) { b2r_set_error("b2r_tsdf_integrate: bad argument"); return B2R_ERR_BAD_ARG; }
    // upstream: "[ScalableTSDFVolume::Integrate] Unsupported image format."
    if (h->color_type == 1 && (!color || (color_channels != 3 && color_channels != 4))) { b2r_set_error("[ScalableTSDFVolume::Integrate] Unsupported image format."); return B2R_ERR_FORMAT; }
    if (h->color_type == 2 && !gray) { b2r_set_error("[ScalableTSDFVolume::Integrate] Unsupported image format."); return B2R_ERR_FORMAT; }
    h->generation++;
    const size_t npx = (size_t)W * H;
    if (h->texel_px < npx) {
        int rc = b2r_tsdf_flush_pending(h, stream);
        if (rc != B2R_OK) return rc;
        B2R_CUDA(cudaStreamSynchronize(stream));
        for (int t = 0; t < 2; t++)
            for (int i = 0; i < B2R_MAX_BATCH; i++) { if (h->texel[t][i]) cudaFree(h->texel[t][i]); h->texel[t][i] = nullptr; }
        h->texel_px = 0;
        for (int t = 0; t < (h->overlap ? 2 : 1); t++)
            for (int i = 0; i < h->max_batch; i++) B2R_CUDA(cudaMalloc((void**)&h->texel[t][i], npx * sizeof(float2)));
        h->texel_px = npx;
    }
    const float fcx = (float)cx, fcy = (float)cy, ifx = 1.0f / (float)fx, ify = 1.0f / (float)fy;
    const double mkey[6] = {(double)W, (double)H, fx, fy, cx, cy};
    if (!h->mult || h->mult_px < npx || memcmp(mkey, h->mult_key, sizeof(mkey)) != 0) {
        // K2 reads the multiplier image at fusion time: frames still pending were packed for the old intrinsics
        int rc = b2r_tsdf_flush_pending(h, stream);
        if (rc != B2R_OK) return rc;
        if (h->mult_px < npx) {
            B2R_CUDA(cudaStreamSynchronize(stream));
            if (h->mult) cudaFree(h->mult);
            h->mult = nullptr; h->mult_px = 0;
            B2R_CUDA(cudaMalloc((void**)&h->mult, npx * sizeof(float)));
            h->mult_px = npx;
        }
        tsdf_mult_kernel<<<(unsigned)((npx + 255) / 256), 256, 0, stream>>>(h->mult, W, H, fcx, fcy, ifx, ify);
        B2R_CHECK_LAUNCH();
        memcpy(h->mult_key, mkey, sizeof(mkey));
    }
    const size_t dstep = npx * (depth_is_u16 ? 2 : 4), cstep = npx * (size_t)color_channels;
    int done = 0;
    while (done < nf) {
        if (h->pending == h->max_batch) {
            int rc = tsdf_launch_fuse(h, stream);        // overlap mode: the caller's stream goes on to the next batch's K1
            if (rc != B2R_OK) return rc;
        }
        const int take = (nf - done < h->max_batch - h->pending) ? nf - done : h->max_batch - h->pending;
        const int tset = h->overlap ? (int)(h->batches & 1) : 0;
        if (h->overlap && h->pending == 0 && h->batches >= 2) {
            // first K1 of batch n: its texel set and its touched buffer were last used by pass n - 2
            B2R_CUDA(cudaStreamWaitEvent(stream, h->ev_k2[(h->batches + 1) % 3], 0));
        }
        PackParams P;
        memset(&P, 0, sizeof(P));
        P.W = W; P.H = H; P.depth_is_u16 = depth_is_u16;
        P.depth_scale = depth_scale; P.depth_trunc = depth_trunc;
        P.color_channels = color_channels; P.color_bgr = color_bgr;
        P.fcx = fcx; P.fcy = fcy; P.ifx = ifx; P.ify = ify;
        P.mult = h->mult;
        P.rscale = 1.0f / depth_scale; P.scale_fast = host_fdiv_domain(depth_scale) ? 1 : 0;
        P.rfx = 1.0 / fx; P.rfy = 1.0 / fy; P.runit = 1.0 / h->unit_length;
        P.ddiv_fast = (host_ddiv_domain(fx) && host_ddiv_domain(fy) && host_ddiv_domain(h->unit_length)) ? 1 : 0;
        P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
        P.trunc = h->sdf_trunc; P.unit_length = h->unit_length; P.stride = h->stride;
        P.Ws = (W + h->stride - 1) / h->stride;
        P.Hs = (H + h->stride - 1) / h->stride;
        P.tiles_x = (P.Ws + 31) / 32;
        const int spt = take <= 2 ? 1 : ALLOC_SPT;
        P.n_alloc_blocks = P.tiles_x * ((P.Hs + 8 * spt - 1) / (8 * spt));
        P.rank = h->rank; P.world = h->world; P.q = h->q;
        P.hash = h->hash; P.unit_keys = h->unit_keys; P.touched = h->touched[h->q];
        P.counters = h->counters; P.cap_units = (int)h->cap_units;
        for (int f = 0; f < take; f++) {
            const double* E = extrinsics + 16 * (size_t)(done + f);
            double pose[16];
            if (!b2r::m4_inverse(E, pose)) { b2r_set_error("b2r_tsdf_integrate: singular extrinsic"); return B2R_ERR_SINGULAR; }
            const int slot = h->pending + f;
            PackFrame& pf = P.fr[f];
            pf.depth = depth_ptrs ? depth_ptrs[done + f] : (const void*)((const char*)depth + (size_t)(done + f) * dstep);
            pf.color = (h->color_type == 1) ? (color_ptrs ? color_ptrs[done + f] : color + (size_t)(done + f) * cstep) : nullptr;
            if (!pf.depth || (h->color_type == 1 && !pf.color)) { b2r_set_error("b2r_tsdf_integrate_frames: null frame pointer"); return B2R_ERR_BAD_ARG; }
            pf.gray = (h->color_type == 2) ? gray + (size_t)(done + f) * npx : nullptr;
            pf.texel = h->texel[tset][slot];
            for (int i = 0; i < 12; i++) pf.pose[i] = pose[i];
            pf.bit = 1u << slot;
            B2RFrame& fr = h->frames[slot];
            for (int i = 0; i < 12; i++) fr.E[i] = (float)E[i];
            const float vl = (float)h->voxel_length;
            fr.Es02 = fr.E[2] * vl; fr.Es12 = fr.E[6] * vl; fr.Es22 = fr.E[10] * vl;
            fr.fx = (float)fx; fr.fy = (float)fy; fr.cx = (float)cx; fr.cy = (float)cy;
            fr.safe_w = (float)W - 0.0001f; fr.safe_h = (float)H - 0.0001f;
            fr.W = W; fr.H = H; fr.texel = h->texel[tset][slot]; fr.mult = h->mult;
        }
        {
            const void* cimg = (h->color_type == 1) ? (const void*)color : (h->color_type == 2 ? (const void*)gray : nullptr);
            const size_t cb = (h->color_type == 1) ? (size_t)color_channels : (h->color_type == 2 ? 4 : 0);
            P.tma_ok = (npx % 16 == 0) && ((uintptr_t)depth % 16 == 0) && (dstep % 16 == 0) &&
                       (cb == 0 || ((uintptr_t)cimg % 16 == 0 && (npx * cb) % 16 == 0)) ? 1 : 0;
            for (int f = 0; f < take && depth_ptrs; f++)
                if ((uintptr_t)P.fr[f].depth % 16 != 0 || (P.fr[f].color && (uintptr_t)P.fr[f].color % 16 != 0)) P.tma_ok = 0;
            const char* e = getenv("B2R_K1_TMA");      // "0": force the LDG form of the pack role (parity test of the unaligned-image path)
            if (e && atoi(e) == 0) P.tma_ok = 0;
        }
        P.n_frames = take;
        P.tiles_per_frame = (int)((npx + PACK_TILE - 1) / PACK_TILE);
        {
            const size_t cb = (h->color_type == 1) ? (size_t)color_channels : (h->color_type == 2 ? 4 : 0);
            P.pack_coff = (uint32_t)(((size_t)PACK_TILE * (depth_is_u16 ? 2 : 4) + 127) & ~(size_t)127);
            P.pack_stage_bytes = P.pack_coff + (uint32_t)(((size_t)PACK_TILE * cb + 127) & ~(size_t)127);
        }
        const long long tiles = (long long)take * P.tiles_per_frame;
        P.n_pack = (int)(tiles < (long long)h->n_sm * PACK_CTAS_PER_SM ? tiles : (long long)h->n_sm * PACK_CTAS_PER_SM);
        size_t k1_smem = (size_t)PACK_STAGES * P.pack_stage_bytes;
        if (k1_smem < (size_t)ALLOC_SMEM) k1_smem = (size_t)ALLOC_SMEM;
        const unsigned grid = (unsigned)(P.n_pack + take * P.n_alloc_blocks);
        const bool prof = prof_begin(h, 1, stream);
        if (spt == 1) tsdf_pack_alloc_kernel<1><<<grid, 256, k1_smem, stream>>>(P);
        else tsdf_pack_alloc_kernel<ALLOC_SPT><<<grid, 256, k1_smem, stream>>>(P);
        B2R_CHECK_LAUNCH();
        prof_end(h, prof, stream);
        h->pending += take;
        h->prof_frames += h->profile ? take : 0;
        done += take;
    }
    return B2R_OK;
}

// volume.integrate(rgbd, intrinsic, extrinsic) with DEVICE images (torch tensors).
//   depth: f32 metres (an RGBDImage's depth) or raw u16 with depth_scale/depth_trunc (a1 fused).
//   color: u8, 3 (RGB) or 4 channels; color_bgr=1 for the camera's BGRA order (cv2.cvtColor fused).
//   extrinsic: HOST 16 f64, row-major, world -> camera.
// The images are consumed by K1 before the call's work on `stream` completes; the voxel update itself is
// deferred until max_batch frames are pending or the volume is read (results are identical either way).
B2R_API int b2r_tsdf_integrate(b2r_tsdf* h, const void* depth, int depth_is_u16, float depth_scale, float depth_trunc,
                               const uint8_t* color, int color_channels, int color_bgr, const float* gray, int W,
                               int H, double fx, double fy, double cx, double cy, const double* extrinsic,
                               cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    return tsdf_integrate_impl(h, 1, depth, depth_is_u16, depth_scale, depth_trunc, color, color_channels, color_bgr, gray,
                               W, H, fx, fy, cx, cy, extrinsic, stream);
}

// Batch of n same-shaped frames stored back to back (re-integration loop of
// bundlefusion_reconstruction.py:309-315 / capture_frames.py:92-105). extrinsics: HOST n*16 f64.
B2R_API int b2r_tsdf_integrate_batch(b2r_tsdf* h, int n, const void* depth, int depth_is_u16, float depth_scale,
                                     float depth_trunc, const uint8_t* color, int color_channels, int color_bgr,
                                     int W, int H, double fx, double fy, double cx, double cy,
                                     const double* extrinsics, cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    if (n == 0) return B2R_OK;
    if (h->color_type == 2) { b2r_set_error("b2r_tsdf_integrate_batch: Gray32 volumes take frames through b2r_tsdf_integrate"); return B2R_ERR_FORMAT; }
    B2RDeviceGuard g(h->device);
    return tsdf_integrate_impl(h, n, depth, depth_is_u16, depth_scale, depth_trunc, color, color_channels, color_bgr, nullptr,
                               W, H, fx, fy, cx, cy, extrinsics, stream);
}

// n same-shaped frames that are NOT stored back to back: depth_ptrs / color_ptrs are HOST arrays of n DEVICE pointers, fused
// in array order (config 5: the slots of a per-tick all-gather buffer, taken tick-major so that the fusion order -- which
// decides the running averages bit for bit -- does not depend on the number of ranks).  extrinsics: HOST n*16 f64.
B2R_API int b2r_tsdf_integrate_frames(b2r_tsdf* h, int n, const void* const* depth_ptrs, int depth_is_u16, float depth_scale,
                                      float depth_trunc, const uint8_t* const* color_ptrs, int color_channels, int color_bgr,
                                      int W, int H, double fx, double fy, double cx, double cy,
                                      const double* extrinsics, cudaStream_t stream) {
    if (!h || !depth_ptrs || n < 0) return B2R_ERR_BAD_ARG;
    if (n == 0) return B2R_OK;
    if (h->color_type == 2) { b2r_set_error("b2r_tsdf_integrate_frames: Gray32 volumes take frames through b2r_tsdf_integrate"); return B2R_ERR_FORMAT; }
    if (h->color_type == 1 && !color_ptrs) { b2r_set_error("[ScalableTSDFVolume::Integrate] Unsupported image format."); return B2R_ERR_FORMAT; }
    B2RDeviceGuard g(h->device);
    return tsdf_integrate_impl(h, n, nullptr, depth_is_u16, depth_scale, depth_trunc, nullptr, color_channels, color_bgr, nullptr,
                               W, H, fx, fy, cx, cy, extrinsics, stream, depth_ptrs, color_ptrs);
}

// Same, with HOST images (pinned for true overlap): each chunk's H2D copy runs on an internal copy
// stream into a B2R_HOST_RING-deep device ring while earlier frames are packed / fused on `stream`.
// Lifetime contract: the host buffers must be fully written when the call is made (the copies do not wait for work
// queued on `stream`); they may be reused or freed as soon as the call returns (it waits for its last copy).
B2R_API int b2r_tsdf_integrate_host(b2r_tsdf* h, int n, const void* depth_host, int depth_is_u16, float depth_scale,
                                    float depth_trunc, const uint8_t* color_host, int color_channels, int color_bgr,
                                    int W, int H, double fx, double fy, double cx, double cy,
                                    const double* extrinsics, cudaStream_t stream) {
    if (!h || !depth_host || n < 0) return B2R_ERR_BAD_ARG;
    if (h->color_type == 2) { b2r_set_error("b2r_tsdf_integrate_host: Gray32 volumes take frames through b2r_tsdf_integrate"); return B2R_ERR_FORMAT; }
    B2RDeviceGuard g(h->device);
    const size_t npx = (size_t)W * H;
    const size_t dbytes = npx * (depth_is_u16 ? 2 : 4), cbytes = color_host ? npx * (size_t)color_channels : 0;
    // a ring slot stages one chunk of up to `chunk` frames: two copies and ONE K1 launch per chunk.  Chunks are shorter than a
    // fusion batch so that the first K1 starts after 8 frames' worth of PCIe time, not 32, and the ring is deep enough to
    // keep the copy engine busy through a whole K2 pass (B2R_HOST_RING x chunk frames buffered >= one batch).
    int chunk = env_int("B2R_HOST_CHUNK", 8, 1, B2R_MAX_BATCH);
    if (chunk > h->max_batch) chunk = h->max_batch;
    const size_t need_d = dbytes * chunk, need_c = cbytes ? cbytes * chunk : 16;
    if (h->stage_depth_bytes < need_d || h->stage_color_bytes < need_c) {
        B2R_CUDA(cudaStreamSynchronize(stream));
        B2R_CUDA(cudaStreamSynchronize(h->copy_stream));
        for (int i = 0; i < B2R_HOST_RING; i++) {
            cudaFree(h->stage_depth[i]); cudaFree(h->stage_color[i]);
            h->stage_depth[i] = h->stage_color[i] = nullptr;
        }
        h->stage_depth_bytes = h->stage_color_bytes = 0;
        for (int i = 0; i < B2R_HOST_RING; i++) {
            B2R_CUDA(cudaMalloc(&h->stage_depth[i], need_d));
            B2R_CUDA(cudaMalloc(&h->stage_color[i], need_c));
        }
        h->stage_depth_bytes = need_d; h->stage_color_bytes = need_c;
        h->host_seq = 0;
    }
    int f = 0, last_r = -1;
    while (f < n) {
        int room = h->max_batch - h->pending;
        if (room <= 0) room = h->max_batch;
        if (room > chunk) room = chunk;
        const int take = n - f < room ? n - f : room;
        const int r = (int)(h->host_seq % B2R_HOST_RING);
        if (h->host_seq >= B2R_HOST_RING) B2R_CUDA(cudaStreamWaitEvent(h->copy_stream, h->ev_consumed[r], 0));
        B2R_CUDA(cudaMemcpyAsync(h->stage_depth[r], (const char*)depth_host + (size_t)f * dbytes, dbytes * take, cudaMemcpyHostToDevice, h->copy_stream));
        if (cbytes) B2R_CUDA(cudaMemcpyAsync(h->stage_color[r], color_host + (size_t)f * cbytes, cbytes * take, cudaMemcpyHostToDevice, h->copy_stream));
        B2R_CUDA(cudaEventRecord(h->ev_copied[r], h->copy_stream));
        B2R_CUDA(cudaStreamWaitEvent(stream, h->ev_copied[r], 0));
        int rc = tsdf_integrate_impl(h, take, h->stage_depth[r], depth_is_u16, depth_scale, depth_trunc,
                                     cbytes ? (const uint8_t*)h->stage_color[r] : nullptr, color_channels, color_bgr,
                                     nullptr, W, H, fx, fy, cx, cy, extrinsics + 16 * (size_t)f, stream);
        if (rc != B2R_OK) return rc;
        B2R_CUDA(cudaEventRecord(h->ev_consumed[r], stream));   // K1 has packed the chunk: the ring slot is free
        h->host_seq++;
        f += take;
        last_r = r;
    }
    // The caller may free or rewrite its host buffer as soon as this returns (a pinned per-chunk temporary, say): wait -- on
    // the HOST, for the copy engine only -- until the call's last H2D copy has left it.  K1 / K2 stay asynchronous.
    if (last_r >= 0) B2R_CUDA(cudaEventSynchronize(h->ev_copied[last_r]));
    return B2R_OK;
}

// ------------------------------------------------------------------------------------------------
// stats / export / import (parity + checkpoint hooks)
// ------------------------------------------------------------------------------------------------
// out[0]=num_units out[1]=error bits out[2]=units touched by the last fused batch out[3]=cumulative updated
// voxels (N_upd) out[4]=cumulative (unit, frame) pairs fused.  Synchronises `stream`.
B2R_API int b2r_tsdf_stats(b2r_tsdf* h, int64_t* out, cudaStream_t stream) {
    if (!h || !out) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    int32_t c[B2R_C_WORDS];
    B2R_CUDA(cudaMemcpyAsync(c, h->counters, sizeof(c), cudaMemcpyDeviceToHost, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    out[0] = c[B2R_C_NUM_UNITS] > h->cap_units ? h->cap_units : c[B2R_C_NUM_UNITS];
    out[1] = c[B2R_C_ERR];
    out[2] = c[B2R_C_LAST_TOUCHED];
    memcpy(&out[3], &c[B2R_C_UPDATED], 8);
    memcpy(&out[4], &c[B2R_C_TOUCHED_TOTAL], 8);
    if (c[B2R_C_ERR] & 1) { b2r_set_error("TSDF unit pool overflow (capacity %lld units)", (long long)h->cap_units); return B2R_ERR_CAPACITY; }
    if (c[B2R_C_ERR] & 4) { b2r_set_error("TSDF hash table full"); return B2R_ERR_CAPACITY; }
    if (c[B2R_C_ERR] & 2) { b2r_set_error("TSDF unit key out of range"); return B2R_ERR_RANGE; }
    return B2R_OK;
}

B2R_API int64_t b2r_tsdf_num_units(b2r_tsdf* h, cudaStream_t stream) {
    int64_t s[5];
    int rc = b2r_tsdf_stats(h, s, stream);
    if (rc != B2R_OK) return rc;
    return s[0];
}

B2R_API double b2r_tsdf_voxel_length(b2r_tsdf* h) { return h ? h->voxel_length : 0.0; }
B2R_API double b2r_tsdf_sdf_trunc(b2r_tsdf* h) { return h ? h->sdf_trunc : 0.0; }
B2R_API int b2r_tsdf_color_type(b2r_tsdf* h) { return h ? h->color_type : -1; }
B2R_API int64_t b2r_tsdf_capacity(b2r_tsdf* h) { return h ? h->cap_units : 0; }
B2R_API int b2r_tsdf_flush(b2r_tsdf* h, cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    return b2r_tsdf_flush_pending(h, stream);
}
B2R_API int b2r_tsdf_set_batch(b2r_tsdf* h, int max_batch, cudaStream_t stream) {
    if (!h || max_batch < 1 || max_batch > B2R_MAX_BATCH) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    if (max_batch > h->max_batch && h->texel_px) {
        B2R_CUDA(cudaStreamSynchronize(stream));
        for (int t = 0; t < (h->overlap ? 2 : 1); t++)
            for (int i = h->max_batch; i < max_batch; i++)
                if (!h->texel[t][i]) B2R_CUDA(cudaMalloc((void**)&h->texel[t][i], h->texel_px * sizeof(float2)));
    }
    h->max_batch = max_batch;
    return B2R_OK;
}

// keys3: DEVICE int32[n*3] in pool order (n = num_units at call time)
B2R_API int b2r_tsdf_export_keys(b2r_tsdf* h, int32_t* keys3, int64_t n, cudaStream_t stream) {
    if (!h || !keys3) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    B2R_CUDA(cudaMemcpyAsync(keys3, h->unit_keys, (size_t)n * 12, cudaMemcpyDeviceToDevice, stream));
    return B2R_OK;
}

// Copies units [first, first+count) out as planes: tsdf[count][4096], weight[count][4096],
// rgb[count][3][4096] (DEVICE pointers; rgb may be null).
B2R_API int b2r_tsdf_export_units(b2r_tsdf* h, int64_t first, int64_t count, float* tsdf, float* weight, float* rgb,
                                  cudaStream_t stream) {
    if (!h || first < 0 || count < 0 || first + count > h->cap_units) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    if (count == 0) return B2R_OK;
    const size_t us = (size_t)h->planes * B2R_UNIT_VOX * 4, pl = B2R_UNIT_VOX * 4;
    const char* src = (const char*)h->pool + first * us;
    if (tsdf) B2R_CUDA(cudaMemcpy2DAsync(tsdf, pl, src, us, pl, count, cudaMemcpyDeviceToDevice, stream));
    if (weight) B2R_CUDA(cudaMemcpy2DAsync(weight, pl, src + pl, us, pl, count, cudaMemcpyDeviceToDevice, stream));
    if (rgb && h->planes == 5) B2R_CUDA(cudaMemcpy2DAsync(rgb, 3 * pl, src + 2 * pl, us, 3 * pl, count, cudaMemcpyDeviceToDevice, stream));
    return B2R_OK;
}

// one CTA per (selected unit, plane): a unit's planes are gathered out of the pool.  units != null: unit-major records
// [n][planes][4096] (the pool's own layout -- what a slab handle receives straight into its pool); else the export layout.
__global__ void tsdf_export_selected_kernel(const float* __restrict__ pool, int64_t unit_stride, int planes, const int32_t* __restrict__ unit_keys,
                                            const int32_t* __restrict__ ids, int32_t* keys3, float* tsdf, float* weight, float* rgb, float* units) {
    const int u = blockIdx.x / planes, pl = blockIdx.x - u * planes;
    const int id = ids[u];
    if (pl == 0 && threadIdx.x < 3 && keys3) keys3[3 * u + threadIdx.x] = unit_keys[3 * id + threadIdx.x];
    float* dst;
    if (units) dst = units + ((int64_t)u * planes + pl) * B2R_UNIT_VOX;
    else dst = pl == 0 ? tsdf + (int64_t)u * B2R_UNIT_VOX : (pl == 1 ? weight + (int64_t)u * B2R_UNIT_VOX : (rgb ? rgb + ((int64_t)u * 3 + (pl - 2)) * B2R_UNIT_VOX : nullptr));
    if (!dst) return;
    const float4* src = reinterpret_cast<const float4*>(pool + (int64_t)id * unit_stride + (int64_t)pl * B2R_UNIT_VOX);
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int v = threadIdx.x; v < B2R_UNIT_VOX / 4; v += blockDim.x) d4[v] = __ldg(src + v);
}

// Copies the units ids[0..n) (DEVICE int32 pool indices) out in export layout: keys3 [n][3], tsdf / weight [n][4096],
// rgb [n][3][4096] (rgb may be null).
B2R_API int b2r_tsdf_export_selected(b2r_tsdf* h, const int32_t* ids, int64_t n, int32_t* keys3, float* tsdf, float* weight, float* rgb,
                                     cudaStream_t stream) {
    if (!h || !ids || !tsdf || !weight || n < 0 || n > (1 << 26)) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    if (n == 0) return B2R_OK;
    tsdf_export_selected_kernel<<<(unsigned)(n * h->planes), 256, 0, stream>>>(h->pool, (int64_t)h->planes * B2R_UNIT_VOX, h->planes, h->unit_keys, ids,
                                                                             keys3, tsdf, weight, h->planes == 5 ? rgb : nullptr, nullptr);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// The same units as pool-layout records: units [n][planes * 4096] floats.  The multi-GPU extraction routes units to slab
// ranks with it; the receiver lets the collective write them straight into its pool (b2r_tsdf_pool, b2r_tsdf_adopt_units).
B2R_API int b2r_tsdf_export_records(b2r_tsdf* h, const int32_t* ids, int64_t n, int32_t* keys3, float* units, cudaStream_t stream) {
    if (!h || !ids || !units || !keys3 || n < 0 || n > (1 << 26)) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    if (n == 0) return B2R_OK;
    tsdf_export_selected_kernel<<<(unsigned)(n * h->planes), 256, 0, stream>>>(h->pool, (int64_t)h->planes * B2R_UNIT_VOX, h->planes, h->unit_keys, ids,
                                                                             keys3, nullptr, nullptr, nullptr, units);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// DEVICE address of the unit pool ([capacity][planes][4096] floats) and the floats per unit: a collective may deposit unit
// records directly behind the units the handle already holds.
B2R_API float* b2r_tsdf_pool(b2r_tsdf* h, int64_t* floats_per_unit) {
    if (!h) return nullptr;
    if (floats_per_unit) *floats_per_unit = (int64_t)h->planes * B2R_UNIT_VOX;
    return h->pool;
}

__global__ void tsdf_adopt_keys_kernel(B2RHash hash, int32_t* unit_keys, int32_t* counters, const int32_t* keys3, int n, int base) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int kx = keys3[3 * i], ky = keys3[3 * i + 1], kz = keys3[3 * i + 2];
    const int id = base + i;
    unit_keys[3 * id] = kx; unit_keys[3 * id + 1] = ky; unit_keys[3 * id + 2] = kz;
    if (i == 0) counters[B2R_C_NUM_UNITS] = base + n;
    if (!b2r_key_in_range(kx, ky, kz)) { atomicOr(&counters[B2R_C_ERR], 2); return; }
    const uint64_t key = b2r_pack_key(kx, ky, kz);
    uint32_t s = (uint32_t)b2r_mix64(key) & hash.mask;
    for (uint32_t probe = 0; probe <= hash.mask; probe++) {
        unsigned long long k = atomicCAS((unsigned long long*)&hash.slots[s].key, (unsigned long long)B2R_KEY_EMPTY, (unsigned long long)key);
        if (k == B2R_KEY_EMPTY) { hash.slots[s].val = id; return; }
        if (k == key) { atomicOr(&counters[B2R_C_ERR], 8); return; }   // duplicate
        s = (s + 1) & hash.mask;
    }
    atomicOr(&counters[B2R_C_ERR], 4);
}

// Registers n units whose records already lie in the pool at indices [first, first + n) (written there by the caller, e.g. as
// the receive buffer of an all-to-all): keys3 = DEVICE int32 [n][3].  `first` must equal the handle's current unit count.
B2R_API int b2r_tsdf_adopt_units(b2r_tsdf* h, const int32_t* keys3, int64_t first, int64_t n, cudaStream_t stream) {
    if (!h || !keys3 || n < 0 || first < 0 || first + n > h->cap_units) return B2R_ERR_BAD_ARG;
    if (n == 0) return B2R_OK;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    h->generation++; h->unit_epoch++;
    tsdf_adopt_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(h->hash, h->unit_keys, h->counters, keys3, (int)n, (int)first);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

__global__ void tsdf_import_keys_kernel(B2RHash hash, int32_t* unit_keys, int32_t* counters, const int32_t* keys3,
                                        int n, int cap, int32_t* out_ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int kx = keys3[3 * i], ky = keys3[3 * i + 1], kz = keys3[3 * i + 2];
    out_ids[i] = -1;
    if (!b2r_key_in_range(kx, ky, kz)) { atomicOr(&counters[B2R_C_ERR], 2); return; }
    const uint64_t key = b2r_pack_key(kx, ky, kz);
    uint32_t s = (uint32_t)b2r_mix64(key) & hash.mask;
    for (uint32_t probe = 0; probe <= hash.mask; probe++) {
        unsigned long long k = atomicCAS((unsigned long long*)&hash.slots[s].key, (unsigned long long)B2R_KEY_EMPTY, (unsigned long long)key);
        if (k == B2R_KEY_EMPTY) {
            int id = atomicAdd(&counters[B2R_C_NUM_UNITS], 1);
            if (id >= cap) { atomicOr(&counters[B2R_C_ERR], 1); id = -1; }
            else { unit_keys[3 * id] = kx; unit_keys[3 * id + 1] = ky; unit_keys[3 * id + 2] = kz; }
            hash.slots[s].val = id;
            out_ids[i] = id;
            return;
        }
        if (k == key) { atomicOr(&counters[B2R_C_ERR], 8); return; }   // duplicate / already present
        s = (s + 1) & hash.mask;
    }
    atomicOr(&counters[B2R_C_ERR], 4);
}

// one CTA per (unit, quarter): the unit index rides on grid.x (up to 2^31 - 1 blocks), not on grid.y (65 535)
__global__ void tsdf_import_planes_kernel(float* pool, int64_t unit_stride, int planes, const int32_t* ids, int n,
                                          const float* tsdf, const float* weight, const float* rgb) {
    const int u = blockIdx.x >> 2, part = blockIdx.x & 3;
    const int id = ids[u];
    if (id < 0) return;
    float* base = pool + (int64_t)id * unit_stride;
    for (int v = part * 1024 + threadIdx.x; v < (part + 1) * 1024; v += blockDim.x) {
        base[v] = tsdf[(int64_t)u * B2R_UNIT_VOX + v];
        base[B2R_UNIT_VOX + v] = weight[(int64_t)u * B2R_UNIT_VOX + v];
        if (planes == 5 && rgb)
            for (int c = 0; c < 3; c++) base[(2 + c) * B2R_UNIT_VOX + v] = rgb[((int64_t)u * 3 + c) * B2R_UNIT_VOX + v];
    }
}

// Adds n NEW units with given keys (DEVICE int32[n*3]) and voxel planes (same layout as export).
// Used to load analytic volumes in tests and to restore a saved volume.
B2R_API int b2r_tsdf_import_units(b2r_tsdf* h, const int32_t* keys3, int64_t n, const float* tsdf, const float* weight,
                                  const float* rgb, cudaStream_t stream) {
    if (!h || !keys3 || !tsdf || !weight || n < 0 || n > (1 << 28)) return B2R_ERR_BAD_ARG;
    if (n == 0) return B2R_OK;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    h->generation++; h->unit_epoch++;
    rc = b2r_tsdf_ensure_scratch(h, (size_t)n * 4);
    if (rc != B2R_OK) return rc;
    int32_t* ids = (int32_t*)h->scratch;
    tsdf_import_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(h->hash, h->unit_keys, h->counters, keys3, (int)n, (int)h->cap_units, ids);
    B2R_CHECK_LAUNCH();
    tsdf_import_planes_kernel<<<(unsigned)(4 * n), 256, 0, stream>>>(h->pool, (int64_t)h->planes * B2R_UNIT_VOX, h->planes, ids, (int)n, tsdf, weight, rgb);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// Per-kernel timing for the roofline line of bench.py: enable for up to max_launches kernel launches
// (0 = off).  Timing events sit on the caller's stream around every K1 / K2 launch.
B2R_API int b2r_tsdf_profile(b2r_tsdf* h, int max_launches) {
    if (!h || max_launches < 0) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    if (max_launches > h->prof_cap) {
        cudaEvent_t* ev = new cudaEvent_t[2 * max_launches];
        int* kind = new int[max_launches];
        for (int i = 0; i < 2 * h->prof_cap; i++) ev[i] = h->prof_ev[i];
        for (int i = 2 * h->prof_cap; i < 2 * max_launches; i++) B2R_CUDA(cudaEventCreate(&ev[i]));
        delete[] h->prof_ev; delete[] h->prof_kind;
        h->prof_ev = ev; h->prof_kind = kind; h->prof_cap = max_launches;
    }
    h->profile = max_launches > 0; h->prof_n = 0; h->prof_frames = 0;
    return B2R_OK;
}
// out[0] = frames packed while profiling, out[1] = total ms in K1 (pack+alloc), out[2] = total ms in K2 (fuse),
// out[3] = K1 launches, out[4] = K2 launches.  Fuses pending frames and synchronises the device.
B2R_API int b2r_tsdf_profile_read(b2r_tsdf* h, double* out5, cudaStream_t stream) {
    if (!h || !out5) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_flush_pending(h, stream);
    if (rc != B2R_OK) return rc;
    B2R_CUDA(cudaDeviceSynchronize());
    double ms[3] = {0, 0, 0}, cnt[3] = {0, 0, 0};
    for (int i = 0; i < h->prof_n; i++) {
        float a = 0.f;
        B2R_CUDA(cudaEventElapsedTime(&a, h->prof_ev[2 * i], h->prof_ev[2 * i + 1]));
        ms[h->prof_kind[i]] += a; cnt[h->prof_kind[i]] += 1;
    }
    out5[0] = (double)h->prof_frames; out5[1] = ms[1]; out5[2] = ms[2]; out5[3] = cnt[1]; out5[4] = cnt[2];
    h->prof_n = 0; h->prof_frames = 0;
    return B2R_OK;
}

// ------------------------------------------------------------------------------------------------
// self-test of the shared-reciprocal division (tests/test_gpu_parity.py): counts inputs for which
// b2r_div_fast differs bitwise from __fdiv_rn, or b2r_rcp_markstein from __frcp_rn.  mode 0: a random, b random in the projection range;
// mode 1: a random, b = small integer weights (1 .. 2^16); mode 2: b sweeps ALL 2^23 significands.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t selftest_rng(uint64_t& s) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    return (uint32_t)(s >> 32);
}
__global__ void selftest_div_kernel(int mode, uint64_t seed, int iters, unsigned long long* mismatches) {
    uint64_t s = b2r_mix64(seed + ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull);
    unsigned bad = 0;
    for (int i = 0; i < iters; i++) {
        float a, b;
        const uint32_t r0 = selftest_rng(s), r1 = selftest_rng(s);
        if (mode == 0) {
            a = __uint_as_float((r0 & 0x807FFFFFu) | ((100u + (r1 & 63u)) << 23));        // |a| in 2^-27 .. 2^36
            b = __uint_as_float((r1 >> 9) | ((117u + ((r0 >> 23) & 15u)) << 23));          // b in 2^-10 .. 2^5
        } else if (mode == 1) {
            a = __uint_as_float((r0 & 0x807FFFFFu) | ((110u + (r1 & 31u)) << 23));
            b = (float)(1u + ((r1 >> 8) & 0xFFFFu));
        } else {
            const uint32_t idx = (blockIdx.x * blockDim.x + threadIdx.x) * (uint32_t)iters + (uint32_t)i;
            b = __uint_as_float((idx & 0x7FFFFFu) | (127u << 23));
            a = __uint_as_float((r0 & 0x807FFFFFu) | ((120u + (r1 & 15u)) << 23));
        }
        const float y = b2r_rcp_markstein(b);
        const float q = b2r_div_domain(b) ? b2r_div_fast(a, b, y) : __fdiv_rn(a, b);
        const float ref = __fdiv_rn(a, b);
        bad += (__float_as_uint(q) != __float_as_uint(ref)) ? 1u : 0u;
        if (b2r_div_domain(b)) bad += (__float_as_uint(y) != __float_as_uint(__frcp_rn(b))) ? 1u : 0u;
    }
    if (bad) atomicAdd(mismatches, (unsigned long long)bad);
}
// *mismatches_host receives the count (must be 0).  n_per_mode: inputs tested per mode (rounded up).
B2R_API int b2r_selftest_div(int64_t n_per_mode, uint64_t seed, int64_t* mismatches_host, cudaStream_t stream) {
    if (!mismatches_host || n_per_mode < 1) return B2R_ERR_BAD_ARG;
    unsigned long long* d = nullptr;
    B2R_CUDA(cudaMalloc(&d, 8));
    B2R_CUDA(cudaMemsetAsync(d, 0, 8, stream));
    const int threads = 256, blocks = 1024;
    for (int mode = 0; mode < 3; mode++) {
        const int64_t n = mode == 2 ? (int64_t)1 << 23 : n_per_mode;
        const int iters = (int)((n + (int64_t)threads * blocks - 1) / ((int64_t)threads * blocks));
        selftest_div_kernel<<<blocks, threads, 0, stream>>>(mode, seed + mode, iters, d);
    }
    unsigned long long hcount = 0;
    cudaError_t e = cudaMemcpyAsync(&hcount, d, 8, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(d);
    if (e != cudaSuccess) { b2r_set_error("b2r_selftest_div: %s", cudaGetErrorString(e)); return B2R_ERR_CUDA; }
    *mismatches_host = (int64_t)hcount;
    return B2R_OK;
}

// Self-test of ddiv_const (the f64 shared-reciprocal division of the allocation samples): for the divisor b (fx, fy or
// the unit length of a call) and its host-rounded reciprocal, counts numerators whose quotient differs bitwise from
// div.rn.f64.  Numerators: random sign/significand, exponents 2^-40 .. 2^23, plus exact multiples of b and 0.
__global__ void selftest_ddiv_kernel(double b, double y, uint64_t seed, int iters, unsigned long long* mismatches) {
    uint64_t s = b2r_mix64(seed + ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull);
    unsigned bad = 0;
    for (int i = 0; i < iters; i++) {
        const uint64_t r0 = ((uint64_t)selftest_rng(s) << 32) | selftest_rng(s);
        const uint32_t r1 = selftest_rng(s);
        double a = __longlong_as_double((long long)((r0 & 0x800FFFFFFFFFFFFFull) | ((uint64_t)(1023 - 40 + (r1 & 63u)) << 52)));
        if ((r1 >> 8) % 61u == 0u) a = __dmul_rn(b, (double)((int)(r1 >> 16) - 32768));   // exact multiples, incl. 0
        const double q = ddiv_const(a, b, y, 1), ref = __ddiv_rn(a, b);
        bad += (__double_as_longlong(q) != __double_as_longlong(ref)) ? 1u : 0u;
    }
    if (bad) atomicAdd(mismatches, (unsigned long long)bad);
}
// Returns B2R_ERR_BAD_ARG when b is outside the domain the integrate call would accept for the fast path.
B2R_API int b2r_selftest_ddiv(double b, int64_t n, uint64_t seed, int64_t* mismatches_host, cudaStream_t stream) {
    if (!mismatches_host || n < 1 || !host_ddiv_domain(b)) return B2R_ERR_BAD_ARG;
    unsigned long long* d = nullptr;
    B2R_CUDA(cudaMalloc(&d, 8));
    B2R_CUDA(cudaMemsetAsync(d, 0, 8, stream));
    const int threads = 256, blocks = 1024;
    const int iters = (int)((n + (int64_t)threads * blocks - 1) / ((int64_t)threads * blocks));
    selftest_ddiv_kernel<<<blocks, threads, 0, stream>>>(b, 1.0 / b, seed, iters, d);
    unsigned long long hcount = 0;
    cudaError_t e = cudaMemcpyAsync(&hcount, d, 8, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(d);
    if (e != cudaSuccess) { b2r_set_error("b2r_selftest_ddiv: %s", cudaGetErrorString(e)); return B2R_ERR_CUDA; }
    *mismatches_host = (int64_t)hcount;
    return B2R_OK;
}
