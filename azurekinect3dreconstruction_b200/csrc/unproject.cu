// unproject.cu -- depth -> point cloud on sm_100a (north_star subsystem 1).
//
// One launch sequence fuses what the reference does in four Open3D/OpenCV calls
//   RGBDImage.create_from_color_and_depth  (u16 -> f32 metres, depth_trunc)         [a1]
//   PointCloud.create_from_rgbd_image / create_from_depth_image                       [a3/a4]
//   pcd.transform(T) (flip / camera extrinsic of the 2-camera scripts)                [a5]
//   cv2.cvtColor(BGRA->RGB) and the colour-to-depth RGB gather                        [a6]
// (/root/reference/CodeThatKindaWorks/Depth2PC.py:51-93,
//  CodeThatWorks/BetterVisualizer2CamWorksCOLOR.py:515-541,634-640; upstream SURVEY.md A.1/A.3).
//
// Output is stream-compacted in RASTER ORDER exactly like upstream's count-then-fill loops:
//   K1 count   : per 1024-pixel chunk, valid-pixel count (128-bit loads: 8 u16 depths per thread)
//   K2 scan    : one CTA per frame, exclusive scan of chunk counts -> chunk offsets + frame count
//   K3 fill    : recompute validity, warp-ballot/popc ranks, unproject (+gather +transform), stage the
//                chunk's xyz/rgb/index in shared memory, then write them out fully coalesced.
// f64 mode follows the upstream f64 chain operation by operation (bit-exact vs the oracle);
// f32 mode is the throughput path (same integer results, coordinates within a few ulp).
#include "common.cuh"
#include "host_math.h"

#ifndef UNP_THREADS
#define UNP_THREADS 128
#endif
#include <type_traits>
#include <algorithm>
#include <string.h>

static bool host_div_domain(float b) {   // same test as b2r_div_domain, on the host
    uint32_t u; memcpy(&u, &b, 4);
    return ((u >> 23) - 67u) < 120u && (u & 0x7FFFFFu) != 0x7FFFFFu;
}

struct UnprojParams {
    int W, H, B;               // depth image size, batch
    int Ws, Hs, stride;        // sampled grid
    int n_samples, n_chunks;   // per frame
    const void* depth; int depth_is_u16; float depth_scale, depth_trunc;
    const uint8_t* color; int color_channels, color_bgr; int Wc, Hc;
    const float* gray;
    int gather;                // 1: colour lives in its own camera -> project + gather
    double fx, fy, cx, cy;     // depth intrinsics (f64)
    float ax, bx, ay, by;      // normalised ray of a pixel: xn = fma(u, ax, bx), yn = fma(v, ay, by)   (a6 chain, f32 path)
    float gm[12];              // colour projection folded into the depth->colour rigid transform (rows u, v, z)
    float fWc, fHc;            // colour image size as floats (bounds test in the float domain)
    float r255hi, r255lo;      // 1/255 as an unevaluated f32 pair: RN(c/255) == fma(c, hi, RN(c*lo)) for c = 0..255
    float rscale; int scale_fast;
    int u16_count_fast;        // depth_scale > 0: the integer validity test below is equivalent
    uint32_t u16_vmax;         // largest raw u16 depth that survives the conversion (RN(v/scale) < trunc); valid <=> 1 <= v <= vmax   // RN(1/depth_scale) and whether depth_scale is inside the Markstein domain
    const double* T;           // DEVICE B*16: camera -> output transform (extrinsic^-1, optionally pre-multiplied)
    int valid_only;
    void* xyz; void* rgb; int32_t* src_index; int32_t* gather_index;
    int64_t capacity;          // points per frame in the output buffers
    int32_t* chunk_counts;     // B * n_chunks (scratch)
    int32_t* chunk_offsets;    // B * n_chunks (scratch)
    int32_t* counts;           // B
    unsigned long long* tile_state;   // B * n_chunks look-back words {flag:32 | value:32} (single-pass f32 path), zeroed per call
    int run_len, runs_per_frame;      // chunks per CTA run (single-pass f32 path)
    int fast_frame;            // stride 1, u16 depth with an in-domain scale > 0, W % 32 == 0, W >= 256, valid_only
};

// a6: nearest colour texel of depth pixel (u, v) at depth z -- the chain DEFINED in oracle.cpp "a6" (f32, correctly
// rounded FMAs over host-folded constants; RN(1/Z) by the Markstein reciprocal, div.rn outside its domain).
__device__ __forceinline__ int gather_texel_n(const UnprojParams& P, float xn, float yn, float z) {
    const float du = __fmaf_rn(P.gm[0], xn, __fmaf_rn(P.gm[1], yn, P.gm[2]));
    const float dv = __fmaf_rn(P.gm[4], xn, __fmaf_rn(P.gm[5], yn, P.gm[6]));
    const float dz = __fmaf_rn(P.gm[8], xn, __fmaf_rn(P.gm[9], yn, P.gm[10]));
    const float U = __fmaf_rn(z, du, P.gm[3]), V = __fmaf_rn(z, dv, P.gm[7]), Z = __fmaf_rn(z, dz, P.gm[11]);
    float inv;
    if (__builtin_expect(b2r_div_domain(Z), 1)) inv = b2r_rcp_markstein(Z); else inv = __frcp_rn(Z);
    const float uf = __fmul_rn(U, inv), vf = __fmul_rn(V, inv);
    // 0 <= uf < Wc (sign bit clear) as one unsigned compare of the bit patterns; NaN and -0.0 fall outside
    const bool in = Z > 0.0f && __float_as_uint(uf) < __float_as_uint(P.fWc) && __float_as_uint(vf) < __float_as_uint(P.fHc);
    return in ? __float2int_rz(vf) * P.Wc + __float2int_rz(uf) : -1;
}
__device__ __forceinline__ int gather_texel(const UnprojParams& P, float u, float v, float z) {
    return gather_texel_n(P, __fmaf_rn(u, P.ax, P.bx), __fmaf_rn(v, P.ay, P.by), z);
}

__device__ __forceinline__ float load_depth(const UnprojParams& P, int b, int pix) {
    const size_t o = (size_t)b * P.W * P.H + pix;
    if (P.depth_is_u16) {
        float d = xfdiv((float)((const uint16_t*)P.depth)[o], P.depth_scale);
        if (d >= P.depth_trunc) d = 0.0f;
        return d;
    }
    return ((const float*)P.depth)[o];
}

// depths of PPT consecutive samples starting at sample s0 (one 128-bit load on the aligned u16 path)
template <int PPT>
__device__ __forceinline__ void load_depths(const UnprojParams& P, int b, int s0, float* d, int* pix) {
    if (PPT == 8 && P.stride == 1 && P.depth_is_u16 && ((P.W * P.H) % 8 == 0) && s0 + 8 <= P.n_samples) {
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>((const uint16_t*)P.depth + (size_t)b * P.W * P.H + s0));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t v = (w[k >> 1] >> ((k & 1) * 16)) & 0xFFFFu;
            float dd = P.scale_fast ? b2r_div_fast((float)v, P.depth_scale, P.rscale) : xfdiv((float)v, P.depth_scale);
            if (dd >= P.depth_trunc) dd = 0.0f;
            d[k] = dd; pix[k] = s0 + k;
        }
        return;
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const int s = s0 + k;
        if (s < P.n_samples) {
            const int si = s / P.Ws, sj = s - si * P.Ws;
            pix[k] = si * P.stride * P.W + sj * P.stride;
            d[k] = load_depth(P, b, pix[k]);
        } else { pix[k] = -1; d[k] = 0.0f; }
    }
}

template <int PPT>
__global__ void __launch_bounds__(UNP_THREADS) unproject_count_kernel(const UnprojParams P) {
    const int b = blockIdx.y, chunk = blockIdx.x;
    const int s0 = (chunk * UNP_THREADS + threadIdx.x) * PPT;
    int c = 0;
    if (PPT == 8 && P.u16_count_fast && P.stride == 1 && P.depth_is_u16 && ((P.W * P.H) % 8 == 0) && s0 + 8 <= P.n_samples) {
        // integer-only validity: RN(v/scale) is monotone in v, so "0 < d < trunc" is "1 <= v <= vmax" (vmax from the host)
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>((const uint16_t*)P.depth + (size_t)b * P.W * P.H + s0));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t v = (w[k >> 1] >> ((k & 1) * 16)) & 0xFFFFu;
            c += (!P.valid_only || (v - 1u) < P.u16_vmax) ? 1 : 0;
        }
    } else {
        float d[PPT]; int pix[PPT];
        load_depths<PPT>(P, b, s0, d, pix);
#pragma unroll
        for (int k = 0; k < PPT; k++) c += (pix[k] >= 0 && (d[k] > 0.0f || !P.valid_only)) ? 1 : 0;
    }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_down_sync(0xffffffffu, c, o);
    __shared__ int ws[UNP_THREADS / 32];
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < UNP_THREADS / 32; i++) t += ws[i];
        P.chunk_counts[b * P.n_chunks + chunk] = t;
    }
}

__global__ void __launch_bounds__(1024) unproject_scan_kernel(const UnprojParams P) {
    const int b = blockIdx.x;
    __shared__ int wsum[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < P.n_chunks; base += 1024) {
        const int i = base + threadIdx.x;
        int v = i < P.n_chunks ? P.chunk_counts[b * P.n_chunks + i] : 0;
        int incl = v;
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += t; }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            int w = wsum[threadIdx.x], wi = w;
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, wi, o); if (threadIdx.x >= o) wi += t; }
            wsum[threadIdx.x] = wi - w;
        }
        __syncthreads();
        const int excl = carry + wsum[threadIdx.x >> 5] + incl - v;
        if (i < P.n_chunks) P.chunk_offsets[b * P.n_chunks + i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) P.counts[b] = carry;
}

template <typename OutT, int PPT>
__global__ void __launch_bounds__(UNP_THREADS) unproject_fill_kernel(const UnprojParams P) {
    constexpr int CHUNK = UNP_THREADS * PPT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    OutT* s_xyz = reinterpret_cast<OutT*>(smem_raw);
    OutT* s_rgb = s_xyz + CHUNK * 3;
    int32_t* s_idx = reinterpret_cast<int32_t*>(s_rgb + CHUNK * 3);
    int32_t* s_gidx = s_idx + CHUNK;
    __shared__ int ws[UNP_THREADS / 32];

    const int b = blockIdx.y, chunk = blockIdx.x;
    const int s0 = (chunk * UNP_THREADS + threadIdx.x) * PPT;
    float d[PPT]; int pix[PPT];
    load_depths<PPT>(P, b, s0, d, pix);
    unsigned m = 0;
#pragma unroll
    for (int k = 0; k < PPT; k++)
        if (pix[k] >= 0 && (d[k] > 0.0f || !P.valid_only)) m |= 1u << k;
    const int cnt = __popc(m);
    // exclusive rank of this thread's first output inside the chunk
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += t; }
    if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = incl;
    __syncthreads();
    int wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < UNP_THREADS / 32; i++) { if (i < (int)(threadIdx.x >> 5)) wbase += ws[i]; total += ws[i]; }
    int rank = wbase + incl - cnt;

    const double* T = P.T + (size_t)b * 16;
    const bool want_color = P.rgb != nullptr;
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        if (!(m & (1u << k))) continue;
        const int p = pix[k];
        const int i = p / P.W, j = p - i * P.W;
        OutT X, Y, Z;
        if (d[k] > 0.0f) {
            if (sizeof(OutT) == 8) {
                const double z = (double)d[k];
                const double x = xddiv(xdmul(xdsub((double)j, P.cx), z), P.fx);
                const double y = xddiv(xdmul(xdsub((double)i, P.cy), z), P.fy);
                X = (OutT)xdadd(xdadd(xdadd(xdmul(T[0], x), xdmul(T[1], y)), xdmul(T[2], z)), T[3]);
                Y = (OutT)xdadd(xdadd(xdadd(xdmul(T[4], x), xdmul(T[5], y)), xdmul(T[6], z)), T[7]);
                Z = (OutT)xdadd(xdadd(xdadd(xdmul(T[8], x), xdmul(T[9], y)), xdmul(T[10], z)), T[11]);
            } else {
                const float z = d[k];
                const float x = xfdiv(xfmul(xfsub((float)j, (float)P.cx), z), (float)P.fx);
                const float y = xfdiv(xfmul(xfsub((float)i, (float)P.cy), z), (float)P.fy);
                X = (OutT)(((float)T[0] * x + (float)T[1] * y) + (float)T[2] * z + (float)T[3]);
                Y = (OutT)(((float)T[4] * x + (float)T[5] * y) + (float)T[6] * z + (float)T[7]);
                Z = (OutT)(((float)T[8] * x + (float)T[9] * y) + (float)T[10] * z + (float)T[11]);
            }
        } else {
            const float nanv = __int_as_float(0x7fc00000);
            X = Y = Z = (OutT)nanv;
        }
        s_xyz[3 * rank] = X; s_xyz[3 * rank + 1] = Y; s_xyz[3 * rank + 2] = Z;
        s_idx[rank] = p;
        int gi = p;   // colour texel index
        if (P.gather) gi = (d[k] > 0.0f) ? gather_texel(P, (float)j, (float)i, d[k]) : -1;
        s_gidx[rank] = gi;
        if (want_color) {
            OutT r = 0, g = 0, bl = 0;
            if (P.color) {
                if (gi >= 0) {
                    const size_t cpix = (size_t)b * (P.gather ? (size_t)P.Wc * P.Hc : (size_t)P.W * P.H) + gi;
                    uint32_t c0, c1, c2;
                    if (P.color_channels == 4) {   // one 32-bit gather per BGRA/RGBA texel
                        const uint32_t t = __ldg(reinterpret_cast<const uint32_t*>(P.color) + cpix);
                        c0 = t & 255u; c1 = (t >> 8) & 255u; c2 = (t >> 16) & 255u;
                    } else {
                        const uint8_t* c = P.color + cpix * 3;
                        c0 = c[0]; c1 = c[1]; c2 = c[2];
                    }
                    if (P.color_bgr) { const uint32_t t = c0; c0 = c2; c2 = t; }
                    if (sizeof(OutT) == 8) {
                        r = (OutT)xddiv((double)c0, 255.0); g = (OutT)xddiv((double)c1, 255.0); bl = (OutT)xddiv((double)c2, 255.0);
                    } else {
                        r = (OutT)xfdiv((float)c0, 255.0f); g = (OutT)xfdiv((float)c1, 255.0f); bl = (OutT)xfdiv((float)c2, 255.0f);
                    }
                }
            } else if (P.gray) {
                r = g = bl = (OutT)P.gray[(size_t)b * P.W * P.H + p];
            }
            s_rgb[3 * rank] = r; s_rgb[3 * rank + 1] = g; s_rgb[3 * rank + 2] = bl;
        }
        rank++;
    }
    __syncthreads();
    // ---- coalesced write-out of the compacted chunk
    const int64_t off = (int64_t)P.chunk_offsets[b * P.n_chunks + chunk];
    if (off + total > P.capacity) total = (int)max((int64_t)0, P.capacity - off);
    const int64_t obase = (int64_t)b * P.capacity + off;
    OutT* gx = reinterpret_cast<OutT*>(P.xyz) + obase * 3;
    for (int t = threadIdx.x; t < total * 3; t += UNP_THREADS) gx[t] = s_xyz[t];
    if (want_color) {
        OutT* gc = reinterpret_cast<OutT*>(P.rgb) + obase * 3;
        for (int t = threadIdx.x; t < total * 3; t += UNP_THREADS) gc[t] = s_rgb[t];
    }
    if (P.src_index)
        for (int t = threadIdx.x; t < total; t += UNP_THREADS) P.src_index[obase + t] = s_idx[t];
    if (P.gather_index)
        for (int t = threadIdx.x; t < total; t += UNP_THREADS) P.gather_index[obase + t] = s_gidx[t];
}

// ---- f32 throughput path: ONE pass (count + raster-order scan + fill) ------------------------------------
// Same integer results (validity, raster order, source and gather indices) as the generic kernels; coordinates
// within a few ulp of the f64 chain; colours RN(c/255) exactly.
//  * single pass: a CTA owns a run of up to 8 consecutive 1024-sample chunks of one frame; it counts the run's valid
//    samples (integer test on the raw u16 depths, all loads in flight), publishes the count and obtains its raster
//    offset by decoupled look-back over the 64-bit {flag | value} words of the preceding runs of the frame, then
//    unprojects chunk by chunk (depths re-read through L1/L2, the next chunk's requested ahead): no count kernel,
//    no scan kernel, no chunk-offset round trip through HBM;
//  * per pixel the work is FMAs over folded constants: normalised ray xn = fma(u, ax, bx) (yn once per thread),
//    output point = z * (T * (xn, yn, 1)) + t, gather texel through the folded projection (a6 chain);
//  * u16 / u8 -> f32 conversions through PRMT into the mantissa of 2^23 (no I2F), exact division by the depth
//    scale through the shared Markstein quotient, c / 255 through a two-term constant (2 instructions, exact);
//  * all colour gathers of a thread in flight together; branch-free predicated stores into a shared-memory
//    stage laid out at the alignment phase of the global destination, written out as 128-bit stores.
// CMODE: 0 no colour, 1 f32 gray, 3 / 4 u8 channels.  GATHER: colour lives in its own camera.
#ifndef UNP_FUSED_MIN_CTAS
#define UNP_FUSED_MIN_CTAS 7          // 72 registers x 128 threads, no spills: 7 CTAs per SM (measured best of 6..9, 64..256 threads)
#endif
#define UNP_MAX_RUN 8
#define UNP_FLAG_AGG 1ull
#define UNP_FLAG_PREFIX 2ull

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Executed by one full warp.  st = look-back words of this frame; returns the number of valid samples in chunks < chunk.
__device__ __forceinline__ int tile_lookback(unsigned long long* st, int chunk, int total) {
    const int lane = threadIdx.x & 31;
    if (chunk == 0) {
        if (lane == 0) st_relaxed_u64(st, (UNP_FLAG_PREFIX << 32) | (unsigned)total);
        return 0;
    }
    if (lane == 0) st_relaxed_u64(st + chunk, (UNP_FLAG_AGG << 32) | (unsigned)total);
    int excl = 0;
    for (int idx = chunk - 1;; idx -= 32) {
        const int j = idx - lane;
        unsigned long long w = (UNP_FLAG_PREFIX << 32);
        if (j >= 0) {
            w = ld_relaxed_u64(st + j);
            while ((w >> 32) == 0) { __nanosleep(20); w = ld_relaxed_u64(st + j); }
        }
        const unsigned pm = __ballot_sync(0xffffffffu, (w >> 32) == UNP_FLAG_PREFIX);
        const int first = __ffs(pm) - 1;                  // pm != 0 at the latest when the window reaches j < 0
        int v = (pm == 0 || lane <= first) ? (int)(unsigned)w : 0;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        excl += v;
        if (pm) break;
    }
    if (lane == 0) st_relaxed_u64(st + chunk, (UNP_FLAG_PREFIX << 32) | (unsigned)(excl + total));
    return excl;
}

// Write-out of a staged chunk: scalar head up to 16-byte alignment, 128-bit streaming stores (fully unrolled), scalar tail.
template <int CMODE, bool WANT_IDX>
__device__ __forceinline__ void unproject_write_out(const UnprojParams& P, int64_t obase, int64_t off, int total, int ph3, int ph1,
                                                    const float* s_xyz, const float* s_rgb, const int32_t* s_idx, const int32_t* s_gidx) {
    constexpr int CHUNK = UNP_THREADS * 8;
    int nout = total;
    if (off + nout > P.capacity) nout = (int)max((int64_t)0, P.capacity - off);
    auto copy_out = [&](float* g, const float* s, int n, int ph, auto max_vec) {   // g[t] = s[ph + t]; (g + t) is 16B-aligned when (ph + t) % 4 == 0
        constexpr int ITERS = (decltype(max_vec)::value + UNP_THREADS - 1) / UNP_THREADS;
        const int head = min(n, (4 - ph) & 3);
        if ((int)threadIdx.x < head) g[threadIdx.x] = s[ph + threadIdx.x];
        const int nv = (n - head) >> 2;
        const float4* sv = reinterpret_cast<const float4*>(s + ph + head);
        float4* gv = reinterpret_cast<float4*>(g + head);
#pragma unroll
        for (int it = 0; it < ITERS; it++) {
            const int t = threadIdx.x + it * UNP_THREADS;
            if (t < nv) __stcs(gv + t, sv[t]);
        }
        const int t = head + (nv << 2) + threadIdx.x;
        if (t < n) g[t] = s[ph + t];
    };
    copy_out(reinterpret_cast<float*>(P.xyz) + obase * 3, s_xyz, nout * 3, ph3, std::integral_constant<int, CHUNK * 3 / 4>());
    if (CMODE != 0) copy_out(reinterpret_cast<float*>(P.rgb) + obase * 3, s_rgb, nout * 3, ph3, std::integral_constant<int, CHUNK * 3 / 4>());
    if (WANT_IDX && P.src_index) copy_out(reinterpret_cast<float*>(P.src_index) + obase, reinterpret_cast<const float*>(s_idx), nout, ph1, std::integral_constant<int, CHUNK / 4>());
    if (WANT_IDX && P.gather_index) copy_out(reinterpret_cast<float*>(P.gather_index) + obase, reinterpret_cast<const float*>(s_gidx), nout, ph1, std::integral_constant<int, CHUNK / 4>());
}

// FAST chunk body: stride 1, u16 depth with an in-domain scale, W % 32 == 0, W >= 256, valid_only, full chunk.
// Pixel mapping: a warp owns 256 consecutive samples and lane l handles samples l, l+32, ..., l+224 of them, so that in
// every step the 32 lanes hold 32 CONSECUTIVE pixels of ONE image row: the depth loads coalesce, raster ranks come
// from one ballot per step (no shuffle scan), and the staged xyz/rgb stores of a warp are 3 words apart per lane --
// bank-conflict-free (8 consecutive pixels per thread would put lanes 24 words apart: 8-way conflicts on every store).
// (row, col) = image position of the warp's first sample.
template <int CMODE, bool GATHER, bool WANT_IDX>
__device__ __forceinline__ int unproject_chunk_fast(const UnprojParams& P, int b, int seg0, int row, int col, int64_t off,
                                                    const float* T, float* s_xyz, float* s_rgb, int32_t* s_idx, int32_t* s_gidx, int* ws) {
    constexpr int PPT = 8;
    const int lane = threadIdx.x & 31;
    const size_t fbase = (size_t)b * P.W * P.H;
    const uint16_t* dp = (const uint16_t*)P.depth + fbase + seg0 + lane;
    uint32_t v[PPT];
#pragma unroll
    for (int k = 0; k < PPT; k++) v[k] = __ldg(dp + 32 * k);
    float d[PPT];
    unsigned bal[PPT];
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const float fv = __fsub_rn(__uint_as_float(v[k] | 0x4B000000u), 8388608.0f);   // (float)v exactly
        float dd = b2r_div_fast(fv, P.depth_scale, P.rscale);      // == div.rn (host checked the divisor's domain)
        dd = (dd >= P.depth_trunc) ? 0.0f : dd;
        d[k] = dd;
        bal[k] = __ballot_sync(0xffffffffu, dd > 0.0f);
        cnt += __popc(bal[k]);
    }
    if (lane == 0) ws[threadIdx.x >> 5] = cnt;                       // warp-uniform count of the warp's 256 samples

    // ---- pass 1: normalised rays + gather texel
    const float flane = (float)lane, fW = (float)P.W;
    float xn[PPT], yn[PPT];
    int gi[PPT];
    {
        float fc = __fadd_rn((float)col, flane), fr = (float)row;      // exact small integers
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            xn[k] = __fmaf_rn(fc, P.ax, P.bx);
            yn[k] = __fmaf_rn(fr, P.ay, P.by);
            fc = __fadd_rn(fc, 32.0f);
            if (fc >= fW) { fc = __fsub_rn(fc, fW); fr = __fadd_rn(fr, 1.0f); }
        }
    }
    if (GATHER) {
        const uint32_t wlim = __float_as_uint(P.fWc), hlim = __float_as_uint(P.fHc);
        unsigned bad = 0;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const float z = d[k];
            const float U = __fmaf_rn(z, __fmaf_rn(P.gm[0], xn[k], __fmaf_rn(P.gm[1], yn[k], P.gm[2])), P.gm[3]);
            const float V = __fmaf_rn(z, __fmaf_rn(P.gm[4], xn[k], __fmaf_rn(P.gm[5], yn[k], P.gm[6])), P.gm[7]);
            const float Z = __fmaf_rn(z, __fmaf_rn(P.gm[8], xn[k], __fmaf_rn(P.gm[9], yn[k], P.gm[10])), P.gm[11]);
            const float inv = b2r_rcp_markstein(Z);                  // == RN(1/Z) inside the domain; the rest is redone below
            if (!b2r_div_domain(Z)) bad |= 1u << k;
            const float uf = __fmul_rn(U, inv), vf = __fmul_rn(V, inv);
            int g = -1;
            if (Z > 0.0f) if (__float_as_uint(uf) < wlim) if (__float_as_uint(vf) < hlim) g = __float2int_rz(vf) * P.Wc + __float2int_rz(uf);
            gi[k] = g;
        }
        if (__builtin_expect(bad != 0, 0)) {                         // Z outside the Markstein domain (never for metric depths): div.rn path
#pragma unroll
            for (int k = 0; k < PPT; k++)
                if ((bad & (1u << k)) && d[k] > 0.0f) gi[k] = gather_texel_n(P, xn[k], yn[k], d[k]);
        }
    } else {
#pragma unroll
        for (int k = 0; k < PPT; k++) gi[k] = seg0 + lane + 32 * k;
    }
    // ---- pass 2: all colour gathers in flight together; 32-bit texel offsets from the batch base (the host checked
    //      that B * Wc * Hc * channels fits 32 bits before declaring the frame FAST): one IMAD.WIDE per address
    uint32_t cw[PPT];
    const uint32_t cb = (uint32_t)b * (uint32_t)(GATHER ? P.Wc * P.Hc : P.W * P.H);
    const uint32_t* c4 = reinterpret_cast<const uint32_t*>(P.color);
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        cw[k] = 0;
        const bool ld = d[k] > 0.0f && gi[k] >= 0;
        if (CMODE == 4) { if (ld) cw[k] = __ldg(c4 + (cb + (uint32_t)gi[k])); }
        else if (CMODE == 3) { if (ld) { const uint8_t* c = P.color + 3u * (cb + (uint32_t)gi[k]); cw[k] = (uint32_t)__ldg(c) | ((uint32_t)__ldg(c + 1) << 8) | ((uint32_t)__ldg(c + 2) << 16); } }
        else if (CMODE == 1) { if (ld) cw[k] = __float_as_uint(__ldg(P.gray + ((uint32_t)b * (uint32_t)(P.W * P.H) + (uint32_t)(seg0 + lane + 32 * k)))); }
    }
    __syncthreads();                                                 // ws
    int wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < UNP_THREADS / 32; i++) { if (i < (int)(threadIdx.x >> 5)) wbase += ws[i]; total += ws[i]; }
    const int64_t obase = (int64_t)b * P.capacity + off;
    const int ph3 = (int)((obase * 3) & 3), ph1 = (int)(obase & 3);     // alignment phase of the destinations (floats / ints)

    // ---- pass 3a: output points + indices into the stage (conflict-free predicated stores) while the colour gathers fly
    const unsigned lt = (1u << lane) - 1u;
    int pos[PPT];
    {
        int base = wbase;                                            // raster rank of the step's first valid pixel (warp-uniform)
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const float z = d[k];
            const float ox = __fmaf_rn(z, __fmaf_rn(T[0], xn[k], __fmaf_rn(T[1], yn[k], T[2])), T[3]);
            const float oy = __fmaf_rn(z, __fmaf_rn(T[4], xn[k], __fmaf_rn(T[5], yn[k], T[6])), T[7]);
            const float oz = __fmaf_rn(z, __fmaf_rn(T[8], xn[k], __fmaf_rn(T[9], yn[k], T[10])), T[11]);
            pos[k] = base + __popc(bal[k] & lt);
            if (z > 0.0f) {
                float* q = s_xyz + ph3 + 3 * pos[k];
                q[0] = ox; q[1] = oy; q[2] = oz;
                if (WANT_IDX) { s_idx[ph1 + pos[k]] = seg0 + lane + 32 * k; s_gidx[ph1 + pos[k]] = gi[k]; }
            }
            base += __popc(bal[k]);
        }
    }
    // ---- pass 3b: colours
    if (CMODE != 0) {
        const uint32_t selr = P.color_bgr ? 0x7542u : 0x7540u, selb = P.color_bgr ? 0x7540u : 0x7542u;   // PRMT: byte -> mantissa of 2^23
        const float rhi = P.r255hi, rlo = P.r255lo;
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            float r, g, bl;
            if (CMODE >= 3) {
                const float fr = __fsub_rn(__uint_as_float(__byte_perm(cw[k], 0x4B000000u, selr)), 8388608.0f);
                const float fg = __fsub_rn(__uint_as_float(__byte_perm(cw[k], 0x4B000000u, 0x7541u)), 8388608.0f);
                const float fb = __fsub_rn(__uint_as_float(__byte_perm(cw[k], 0x4B000000u, selb)), 8388608.0f);
                r = __fmaf_rn(fr, rhi, __fmul_rn(fr, rlo));
                g = __fmaf_rn(fg, rhi, __fmul_rn(fg, rlo));
                bl = __fmaf_rn(fb, rhi, __fmul_rn(fb, rlo));
            } else {
                r = g = bl = __uint_as_float(cw[k]);
            }
            if (d[k] > 0.0f) { float* qc = s_rgb + ph3 + 3 * pos[k]; qc[0] = r; qc[1] = g; qc[2] = bl; }
        }
    }
    __syncthreads();
    unproject_write_out<CMODE, WANT_IDX>(P, obase, off, total, ph3, ph1, s_xyz, s_rgb, s_idx, s_gidx);
    __syncthreads();                                                 // the stage and ws are reused by the next chunk
    return total;
}

// One chunk (1024 samples) of one frame: unproject, gather, stage, write out at raster offset `off`; returns the
// chunk's valid count.  FAST: stride 1, u16 depth with an in-domain scale, W % 8 == 0, valid_only, full chunk -- every
// sample index is s0 + k, a thread's 8 samples share an image row (row, col), no per-sample validity of the index.
template <int CMODE, bool GATHER, bool WANT_IDX, bool FAST>
__device__ __forceinline__ int unproject_chunk(const UnprojParams& P, int b, int chunk, int row, int col, int64_t off, const float* T,
                                               float* s_xyz, float* s_rgb, int32_t* s_idx, int32_t* s_gidx, int* ws) {
    constexpr int PPT = 8;
    const int s0 = (chunk * UNP_THREADS + threadIdx.x) * PPT;
    const size_t fbase = (size_t)b * P.W * P.H;
    float d[PPT]; int pix[PPT];
    unsigned m = 0;
    if (FAST) {
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>((const uint16_t*)P.depth + fbase + s0));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            // (float)v exactly: v in the low mantissa bits of 2^23, minus 2^23
            const float fv = __fsub_rn(__uint_as_float(__byte_perm(w[k >> 1], 0x4B000000u, (k & 1) ? 0x7532u : 0x7510u)), 8388608.0f);
            float dd = b2r_div_fast(fv, P.depth_scale, P.rscale);      // == div.rn (host checked the divisor's domain)
            if (dd >= P.depth_trunc) dd = 0.0f;
            d[k] = dd; pix[k] = s0 + k;
            m |= (dd > 0.0f) ? (1u << k) : 0u;
        }
    } else {
        load_depths<PPT>(P, b, s0, d, pix);
#pragma unroll
        for (int k = 0; k < PPT; k++) m |= (pix[k] >= 0 && (d[k] > 0.0f || !P.valid_only)) ? (1u << k) : 0u;
    }
    const int cnt = __popc(m);
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += t; }
    if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = incl;
    __syncthreads();
    int wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < UNP_THREADS / 32; i++) { if (i < (int)(threadIdx.x >> 5)) wbase += ws[i]; total += ws[i]; }
    int rank = wbase + incl - cnt;

    // ---- pass 1: normalised rays + gather texel index
    float xn[PPT], yn[PPT];
    int gi[PPT];
    if (FAST) {
        const float fj0 = (float)col, yn0 = __fmaf_rn((float)row, P.ay, P.by);
#pragma unroll
        for (int k = 0; k < PPT; k++) { xn[k] = __fmaf_rn(__fadd_rn(fj0, (float)k), P.ax, P.bx); yn[k] = yn0; }   // (float)(col + k) exactly
    } else {
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const int p = pix[k] < 0 ? 0 : pix[k];
            const int i = p / P.W, j = p - i * P.W;
            xn[k] = __fmaf_rn((float)j, P.ax, P.bx);
            yn[k] = __fmaf_rn((float)i, P.ay, P.by);
        }
    }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        gi[k] = pix[k];
        if (GATHER) gi[k] = (d[k] > 0.0f) ? gather_texel_n(P, xn[k], yn[k], d[k]) : -1;
    }
    // ---- pass 2: all colour gathers in flight together (32-bit texel offsets from a per-frame base pointer)
    uint32_t cw[PPT];
    const size_t cbase = GATHER ? (size_t)b * P.Wc * P.Hc : fbase;
    const uint32_t* c4 = reinterpret_cast<const uint32_t*>(P.color) + cbase;
    const uint8_t* c3 = P.color + cbase * 3;
    const float* cg = P.gray + fbase;
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        cw[k] = 0;
        if (CMODE != 0 && (m & (1u << k)) && gi[k] >= 0) {
            if (CMODE == 4) cw[k] = __ldg(c4 + gi[k]);
            else if (CMODE == 3) { const uint8_t* c = c3 + 3 * gi[k]; cw[k] = (uint32_t)__ldg(c) | ((uint32_t)__ldg(c + 1) << 8) | ((uint32_t)__ldg(c + 2) << 16); }
            else cw[k] = __float_as_uint(__ldg(cg + pix[k]));
        }
    }
    const int64_t obase = (int64_t)b * P.capacity + off;
    const int ph3 = (int)((obase * 3) & 3), ph1 = (int)(obase & 3);     // alignment phase of the destinations (floats / ints)

    // ---- pass 3: output point, colour conversion, predicated stores into the stage
    const uint32_t selr = P.color_bgr ? 0x7542u : 0x7540u, selb = P.color_bgr ? 0x7540u : 0x7542u;   // PRMT: byte -> mantissa of 2^23
    float* q = s_xyz + ph3 + 3 * rank;
    float* qc = s_rgb + ph3 + 3 * rank;
    int32_t* qi = s_idx + ph1 + rank;
    int32_t* qg = s_gidx + ph1 + rank;
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const float z = d[k];
        float ox = __fmaf_rn(z, __fmaf_rn(T[0], xn[k], __fmaf_rn(T[1], yn[k], T[2])), T[3]);
        float oy = __fmaf_rn(z, __fmaf_rn(T[4], xn[k], __fmaf_rn(T[5], yn[k], T[6])), T[7]);
        float oz = __fmaf_rn(z, __fmaf_rn(T[8], xn[k], __fmaf_rn(T[9], yn[k], T[10])), T[11]);
        if (!FAST && !(z > 0.0f)) ox = oy = oz = __int_as_float(0x7fc00000);      // project_valid_depth_only = False keeps NaN points
        float r = 0.f, g = 0.f, bl = 0.f;
        if (CMODE >= 3) {
            const float fr = __fsub_rn(__uint_as_float(__byte_perm(cw[k], 0x4B000000u, selr)), 8388608.0f);
            const float fg = __fsub_rn(__uint_as_float(__byte_perm(cw[k], 0x4B000000u, 0x7541u)), 8388608.0f);
            const float fb = __fsub_rn(__uint_as_float(__byte_perm(cw[k], 0x4B000000u, selb)), 8388608.0f);
            r = __fmaf_rn(fr, P.r255hi, __fmul_rn(fr, P.r255lo));
            g = __fmaf_rn(fg, P.r255hi, __fmul_rn(fg, P.r255lo));
            bl = __fmaf_rn(fb, P.r255hi, __fmul_rn(fb, P.r255lo));
        } else if (CMODE == 1) {
            r = g = bl = __uint_as_float(cw[k]);
        }
        if (m & (1u << k)) {
            q[0] = ox; q[1] = oy; q[2] = oz; q += 3;
            if (CMODE != 0) { qc[0] = r; qc[1] = g; qc[2] = bl; qc += 3; }
            if (WANT_IDX) { *qi++ = pix[k]; *qg++ = gi[k]; }
        }
    }
    __syncthreads();
    unproject_write_out<CMODE, WANT_IDX>(P, obase, off, total, ph3, ph1, s_xyz, s_rgb, s_idx, s_gidx);
    __syncthreads();                                                 // the stage and ws are reused by the next chunk
    return total;
}

template <int CMODE, bool GATHER, bool WANT_IDX>
__global__ void __launch_bounds__(UNP_THREADS, UNP_FUSED_MIN_CTAS) unproject_fused_f32_kernel(const __grid_constant__ UnprojParams P) {
    constexpr int PPT = 8, CHUNK = UNP_THREADS * PPT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s_xyz = reinterpret_cast<float*>(smem_raw);            // CHUNK*3 + 4
    float* s_rgb = s_xyz + CHUNK * 3 + 4;                          // CHUNK*3 + 4 (absent when CMODE == 0)
    int32_t* s_idx = reinterpret_cast<int32_t*>(s_rgb + (CMODE != 0 ? CHUNK * 3 + 4 : 0));   // CHUNK + 4 (WANT_IDX only)
    int32_t* s_gidx = s_idx + CHUNK + 4;                           // CHUNK + 4
    __shared__ int ws[UNP_THREADS / 32];
    __shared__ int s_excl;

    // A CTA owns a RUN of consecutive chunks of one frame: blockIdx.x = run within the frame, blockIdx.y = frame.
    // The look-back of run x waits only on runs x' < x of the same frame, which have lower linear block ids and are
    // therefore dispatched no later (the in-order dispatch cub::DeviceScan's look-back also relies on).
    const int b = blockIdx.y, rif = blockIdx.x;
    const int c0 = rif * P.run_len;
    const int nch = min(P.run_len, P.n_chunks - c0);
    const size_t fbase = (size_t)b * P.W * P.H;

    // ---- phase A: valid samples of the whole run (depth re-read from L1/L2 in phase B), then the run's raster offset
    int c = 0;
    if (P.fast_frame) {                                              // integer-only validity: 1 <= v <= vmax; all loads of the run in flight
        uint4 raw[UNP_MAX_RUN];
#pragma unroll
        for (int ch = 0; ch < UNP_MAX_RUN; ch++) {
            const int s0 = ((c0 + ch) * UNP_THREADS + threadIdx.x) * PPT;
            raw[ch] = make_uint4(0, 0, 0, 0);
            if (ch < nch && s0 + 8 <= P.n_samples) raw[ch] = __ldg(reinterpret_cast<const uint4*>((const uint16_t*)P.depth + fbase + s0));
        }
#pragma unroll
        for (int ch = 0; ch < UNP_MAX_RUN; ch++) {
            const uint32_t w[4] = {raw[ch].x, raw[ch].y, raw[ch].z, raw[ch].w};
#pragma unroll
            for (int k = 0; k < 8; k++) c += (((w[k >> 1] >> ((k & 1) * 16)) & 0xFFFFu) - 1u < P.u16_vmax) ? 1 : 0;
        }
        const int s_tail = ((c0 + nch - 1) * UNP_THREADS + threadIdx.x) * PPT;      // ragged end of the frame (only its last run)
        if (s_tail + 8 > P.n_samples)
            for (int k = 0; k < 8; k++)
                if (s_tail + k < P.n_samples) c += ((uint32_t)((const uint16_t*)P.depth)[fbase + s_tail + k] - 1u < P.u16_vmax) ? 1 : 0;
    } else {
        for (int ch = 0; ch < nch; ch++) {
            float d[PPT]; int pix[PPT];
            load_depths<PPT>(P, b, ((c0 + ch) * UNP_THREADS + threadIdx.x) * PPT, d, pix);
#pragma unroll
            for (int k = 0; k < PPT; k++) c += (pix[k] >= 0 && (d[k] > 0.0f || !P.valid_only)) ? 1 : 0;
        }
    }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x < 32) {
        int total = 0;
#pragma unroll
        for (int i = 0; i < UNP_THREADS / 32; i++) total += ws[i];
        const int excl = tile_lookback(P.tile_state + (size_t)b * P.runs_per_frame, rif, total);
        if (threadIdx.x == 0) {
            s_excl = excl;
            if (c0 + nch == P.n_chunks) P.counts[b] = excl + total;
        }
    }
    const double* Td = P.T + (size_t)b * 16;
    float T[12];
#pragma unroll
    for (int i = 0; i < 12; i++) T[i] = (float)__ldg(Td + i);
    __syncthreads();
    int64_t off = (int64_t)s_excl;

    // ---- phase B: the chunks of the run
    int row = 0, col = 0;                                            // image position of the warp's first sample (FAST frames)
    const int seg_first = c0 * CHUNK + (threadIdx.x >> 5) * 256;
    if (P.fast_frame) { row = seg_first / P.W; col = seg_first - row * P.W; }
    for (int ch = 0; ch < nch; ch++) {
        const int chunk = c0 + ch;
        if (P.fast_frame && (chunk + 1) * CHUNK <= P.n_samples)
            off += unproject_chunk_fast<CMODE, GATHER, WANT_IDX>(P, b, seg_first + ch * CHUNK, row, col, off, T, s_xyz, s_rgb, s_idx, s_gidx, ws);
        else
            off += unproject_chunk<CMODE, GATHER, WANT_IDX, false>(P, b, chunk, row, col, off, T, s_xyz, s_rgb, s_idx, s_gidx, ws);
        col += CHUNK;
        while (col >= P.W) { col -= P.W; row++; }
    }
}

template <int CMODE, bool GATHER, bool WANT_IDX>
static int launch_fused_f32_t2(const UnprojParams& P, cudaStream_t stream) {
    constexpr int CHUNK = UNP_THREADS * 8;
    const size_t smem = (size_t)(CHUNK * 3 + 4) * 4 * (CMODE != 0 ? 2 : 1) + (WANT_IDX ? (size_t)(CHUNK + 4) * 4 * 2 : 0);
    static bool attr_set = false;
    if (!attr_set) {
        B2R_CUDA(cudaFuncSetAttribute(unproject_fused_f32_kernel<CMODE, GATHER, WANT_IDX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    // look-back words were zeroed by the caller
    unproject_fused_f32_kernel<CMODE, GATHER, WANT_IDX><<<dim3(P.runs_per_frame, P.B), UNP_THREADS, smem, stream>>>(P);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

template <int CMODE, bool GATHER>
static int launch_fused_f32_t(const UnprojParams& P, cudaStream_t stream) {
    return (P.src_index || P.gather_index) ? launch_fused_f32_t2<CMODE, GATHER, true>(P, stream) : launch_fused_f32_t2<CMODE, GATHER, false>(P, stream);
}

static int launch_fused_f32(const UnprojParams& P, cudaStream_t stream) {
    const int cmode = (P.rgb && P.color) ? P.color_channels : ((P.rgb && P.gray) ? 1 : 0);
    if (cmode == 4) return P.gather ? launch_fused_f32_t<4, true>(P, stream) : launch_fused_f32_t<4, false>(P, stream);
    if (cmode == 3) return P.gather ? launch_fused_f32_t<3, true>(P, stream) : launch_fused_f32_t<3, false>(P, stream);
    if (cmode == 1) return launch_fused_f32_t<1, false>(P, stream);
    return launch_fused_f32_t<0, false>(P, stream);
}

template <typename OutT, int PPT>
static int launch_fill(const UnprojParams& P, cudaStream_t stream) {
    constexpr int CHUNK = UNP_THREADS * PPT;
    const size_t smem = (size_t)CHUNK * (6 * sizeof(OutT) + 8);
    static bool attr_set = false;
    if (!attr_set) {
        B2R_CUDA(cudaFuncSetAttribute(unproject_fill_kernel<OutT, PPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    dim3 grid(P.n_chunks, P.B);
    unproject_fill_kernel<OutT, PPT><<<grid, UNP_THREADS, smem, stream>>>(P);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// Scratch needed by b2r_unproject for a batch (int32 words, 8-byte aligned): 2 * B * n_chunks + 2
B2R_API int64_t b2r_unproject_scratch_words(int W, int H, int B, int stride, int out_f64) {
    const int Ws = (W + stride - 1) / stride, Hs = (H + stride - 1) / stride;
    const int chunk = UNP_THREADS * (out_f64 ? 4 : 8);
    const int n_chunks = (Ws * Hs + chunk - 1) / chunk;
    return (int64_t)2 * B * n_chunks + 2;
}

// See include/b2recon.h for the argument contract.
B2R_API int b2r_unproject(const void* depth, int depth_is_u16, float depth_scale, float depth_trunc, int W, int H, int B,
                          const uint8_t* color, int color_channels, int color_bgr, int Wc, int Hc, const float* gray,
                          const double* depth_K, const double* color_K, const double* d2c_host, const double* T_dev,
                          int stride, int out_f64, int valid_only, void* xyz, void* rgb, int32_t* src_index,
                          int32_t* gather_index, int64_t capacity, int32_t* counts, int32_t* scratch,
                          cudaStream_t stream) {
    if (!depth || !depth_K || !T_dev || !xyz || !counts || !scratch || W < 1 || H < 1 || B < 1 || stride < 1) {
        b2r_set_error("b2r_unproject: bad argument");
        return B2R_ERR_BAD_ARG;
    }
    if (color && color_channels != 3 && color_channels != 4) { b2r_set_error("b2r_unproject: colour must have 3 or 4 channels"); return B2R_ERR_FORMAT; }
    UnprojParams P;
    memset(&P, 0, sizeof(P));
    P.W = W; P.H = H; P.B = B; P.stride = stride;
    P.Ws = (W + stride - 1) / stride; P.Hs = (H + stride - 1) / stride;
    P.n_samples = P.Ws * P.Hs;
    const int chunk = UNP_THREADS * (out_f64 ? 4 : 8);
    P.n_chunks = (P.n_samples + chunk - 1) / chunk;
    P.depth = depth; P.depth_is_u16 = depth_is_u16; P.depth_scale = depth_scale; P.depth_trunc = depth_trunc;
    P.color = color; P.color_channels = color_channels; P.color_bgr = color_bgr; P.gray = gray;
    P.gather = (color && color_K && d2c_host) ? 1 : 0;
    P.Wc = P.gather ? Wc : W; P.Hc = P.gather ? Hc : H;
    if (color && !P.gather && (Wc != W || Hc != H) && Wc > 0) {
        // upstream RGBDImage::CreateFromColorAndDepth: LogError("Unsupported image format.")
        b2r_set_error("Unsupported image format.");
        return B2R_ERR_FORMAT;
    }
    P.fx = depth_K[0]; P.fy = depth_K[1]; P.cx = depth_K[2]; P.cy = depth_K[3];
    {   // 1/255 as an unevaluated f32 pair (hi + lo): fma(c, hi, RN(c*lo)) == RN(c/255) for every byte c (b2r_selftest_u8_div255)
        const double r = 1.0 / 255.0;
        P.r255hi = (float)r; P.r255lo = (float)(r - (double)P.r255hi);
    }
    P.rscale = 1.0f / depth_scale; P.scale_fast = host_div_domain(depth_scale) ? 1 : 0;
    P.u16_vmax = 0; P.u16_count_fast = (depth_is_u16 && depth_scale > 0.0f) ? 1 : 0;
    if (P.u16_count_fast) {   // binary search of the monotone predicate RN((float)v / scale) < trunc (IEEE float division, as on the device)
        uint32_t lo = 0, hi = 65535;
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            volatile float q = (float)mid / depth_scale;
            if (q < depth_trunc) lo = mid; else hi = mid - 1;
        }
        P.u16_vmax = lo;
    }
    {
        float gc[16];
        const double idK[4] = {1.0, 1.0, 0.0, 0.0}, idT[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
        b2r::gather_constants(depth_K, P.gather ? color_K : idK, P.gather ? d2c_host : idT, gc);
        P.ax = gc[0]; P.bx = gc[1]; P.ay = gc[2]; P.by = gc[3];
        for (int i = 0; i < 12; i++) P.gm[i] = gc[4 + i];
        P.fWc = (float)P.Wc; P.fHc = (float)P.Hc;
    }
    P.T = T_dev; P.valid_only = valid_only;
    P.xyz = xyz; P.rgb = (color || gray) ? rgb : nullptr; P.src_index = src_index; P.gather_index = gather_index;
    P.capacity = capacity;
    P.counts = counts;
    if (!out_f64) {
        // single pass: look-back words (u64 per run) live in the caller's scratch, zeroed per call
        if ((reinterpret_cast<uintptr_t>(scratch) & 7u) != 0) { b2r_set_error("b2r_unproject: scratch must be 8-byte aligned"); return B2R_ERR_BAD_ARG; }
        const size_t n_tiles = (size_t)B * P.n_chunks;
        // run length: 8 chunks per CTA once that still leaves >= 2 waves of CTAs, fewer for small batches (latency case)
        P.run_len = (int)std::min<size_t>(UNP_MAX_RUN, std::max<size_t>(1, n_tiles / 2048));
        P.runs_per_frame = (P.n_chunks + P.run_len - 1) / P.run_len;
        P.fast_frame = (P.scale_fast && P.u16_count_fast && stride == 1 && depth_is_u16 && (W % 32) == 0 && W >= 256 && valid_only &&
                        (double)B * P.Wc * P.Hc * 4.0 < 4294967296.0 && (double)B * W * H * 4.0 < 4294967296.0) ? 1 : 0;
        P.tile_state = reinterpret_cast<unsigned long long*>(scratch);
        B2R_CUDA(cudaMemsetAsync(scratch, 0, (size_t)B * P.runs_per_frame * sizeof(unsigned long long), stream));
        return launch_fused_f32(P, stream);
    }
    P.chunk_counts = scratch; P.chunk_offsets = scratch + (size_t)B * P.n_chunks; P.counts = counts;
    dim3 grid(P.n_chunks, B);
    unproject_count_kernel<4><<<grid, UNP_THREADS, 0, stream>>>(P);
    B2R_CHECK_LAUNCH();
    unproject_scan_kernel<<<B, 1024, 0, stream>>>(P);
    B2R_CHECK_LAUNCH();
    return launch_fill<double, 4>(P, stream);
}

// ---- a5: pcd.transform(T) on device arrays (f64 or f32 points, optional normals) ---------------
template <typename T>
__global__ void transform_points_kernel(T* xyz, T* nrm, int64_t n, const double* M) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x = (double)xyz[3 * i], y = (double)xyz[3 * i + 1], z = (double)xyz[3 * i + 2];
    double h[4];
#pragma unroll
    for (int r = 0; r < 4; r++) h[r] = xdadd(xdadd(xdadd(xdmul(M[4 * r], x), xdmul(M[4 * r + 1], y)), xdmul(M[4 * r + 2], z)), M[4 * r + 3]);
    xyz[3 * i] = (T)xddiv(h[0], h[3]); xyz[3 * i + 1] = (T)xddiv(h[1], h[3]); xyz[3 * i + 2] = (T)xddiv(h[2], h[3]);
    if (nrm) {
        const double a = (double)nrm[3 * i], b = (double)nrm[3 * i + 1], c = (double)nrm[3 * i + 2];
#pragma unroll
        for (int r = 0; r < 3; r++) nrm[3 * i + r] = (T)xdadd(xdadd(xdmul(M[4 * r], a), xdmul(M[4 * r + 1], b)), xdmul(M[4 * r + 2], c));
    }
}

struct M16 { double m[16]; };
template <typename T>
__global__ void transform_points_kernel_v(T* xyz, T* nrm, int64_t n, const M16 M) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x = (double)xyz[3 * i], y = (double)xyz[3 * i + 1], z = (double)xyz[3 * i + 2];
    double h[4];
#pragma unroll
    for (int r = 0; r < 4; r++) h[r] = xdadd(xdadd(xdadd(xdmul(M.m[4 * r], x), xdmul(M.m[4 * r + 1], y)), xdmul(M.m[4 * r + 2], z)), M.m[4 * r + 3]);
    xyz[3 * i] = (T)xddiv(h[0], h[3]); xyz[3 * i + 1] = (T)xddiv(h[1], h[3]); xyz[3 * i + 2] = (T)xddiv(h[2], h[3]);
    if (nrm) {
        const double a = (double)nrm[3 * i], b = (double)nrm[3 * i + 1], c = (double)nrm[3 * i + 2];
#pragma unroll
        for (int r = 0; r < 3; r++) nrm[3 * i + r] = (T)xdadd(xdadd(xdmul(M.m[4 * r], a), xdmul(M.m[4 * r + 1], b)), xdmul(M.m[4 * r + 2], c));
    }
}

// T_host: HOST 16 f64.  xyz/normals: DEVICE n*3 (f64 when is_f64 else f32); normals may be null.
B2R_API int b2r_transform_points(void* xyz, void* normals, int64_t n, int is_f64, const double* T_host, cudaStream_t stream) {
    if (!xyz || !T_host || n < 0) return B2R_ERR_BAD_ARG;
    if (n == 0) return B2R_OK;
    M16 M;
    memcpy(M.m, T_host, sizeof(M.m));
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (is_f64) transform_points_kernel_v<double><<<grid, 256, 0, stream>>>((double*)xyz, (double*)normals, n, M);
    else transform_points_kernel_v<float><<<grid, 256, 0, stream>>>((float*)xyz, (float*)normals, n, M);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// ---- a1 stand-alone conversions (normally fused into the consumers) ----------------------------
__global__ void depth_to_float_kernel(const uint16_t* in, int64_t n, float scale, float trunc, float* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float d = xfdiv((float)in[i], scale);
    if (d >= trunc) d = 0.0f;
    out[i] = d;
}
__global__ void rgb_to_intensity_kernel(const uint8_t* rgb, int64_t n, int channels, int bgr, float* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* p = rgb + i * channels;
    const float r = (float)p[bgr ? 2 : 0], g = (float)p[1], b = (float)p[bgr ? 0 : 2];
    out[i] = xfdiv(xfadd(xfadd(xfmul(0.2990f, r), xfmul(0.5870f, g)), xfmul(0.1140f, b)), 255.0f);
}
B2R_API int b2r_depth_to_float(const uint16_t* depth, int64_t n, float depth_scale, float depth_trunc, float* out, cudaStream_t stream) {
    if (!depth || !out || n < 0) return B2R_ERR_BAD_ARG;
    if (n) depth_to_float_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(depth, n, depth_scale, depth_trunc, out);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}
// ConvertDepthToFloatImage on the other 1-channel depth types upstream's CreateFloatImage accepts:
// dtype 1 = f32 (copied, then / scale), dtype 2 = u8 (/ 255.0f first), dtype 0 = u16 (same as b2r_depth_to_float).
template <typename T, bool DIV255>
__global__ void depth_any_to_float_kernel(const T* in, int64_t n, float scale, float trunc, float* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float d = (float)in[i];
    if (DIV255) d = xfdiv(d, 255.0f);
    d = xfdiv(d, scale);
    if (d >= trunc) d = 0.0f;
    out[i] = d;
}
B2R_API int b2r_depth_to_float_typed(const void* depth, int dtype, int64_t n, float depth_scale, float depth_trunc, float* out, cudaStream_t stream) {
    if (!depth || !out || n < 0 || dtype < 0 || dtype > 2) return B2R_ERR_BAD_ARG;
    const unsigned g = (unsigned)((n + 255) / 256);
    if (n && dtype == 0) depth_to_float_kernel<<<g, 256, 0, stream>>>((const uint16_t*)depth, n, depth_scale, depth_trunc, out);
    if (n && dtype == 1) depth_any_to_float_kernel<float, false><<<g, 256, 0, stream>>>((const float*)depth, n, depth_scale, depth_trunc, out);
    if (n && dtype == 2) depth_any_to_float_kernel<uint8_t, true><<<g, 256, 0, stream>>>((const uint8_t*)depth, n, depth_scale, depth_trunc, out);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}
B2R_API int b2r_rgb_to_intensity(const uint8_t* rgb, int64_t n, int channels, int bgr, float* out, cudaStream_t stream) {
    if (!rgb || !out || n < 0 || (channels != 3 && channels != 4)) return B2R_ERR_BAD_ARG;
    if (n) rgb_to_intensity_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(rgb, n, channels, bgr, out);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// Host-side exhaustive check of the two-term 1/255 constant used by the f32 unprojection path: number of bytes c for
// which fma(c, hi, RN(c*lo)) != RN(c/255).  (FFMA/FMUL on the device are the same correctly rounded operations.)
B2R_API int b2r_selftest_u8_div255(void) {
    const double r = 1.0 / 255.0;
    const float hi = (float)r, lo = (float)(r - (double)hi);
    int bad = 0;
    for (int c = 0; c < 256; c++) {
        volatile float l = (float)c * lo;
        const float q = fmaf((float)c, hi, l);
        volatile float ref = (float)c / 255.0f;
        bad += (q != ref) ? 1 : 0;
    }
    return bad;
}

// host helpers shared with the Python shim so pose algebra is done once, in f64, the same way everywhere
B2R_API int b2r_m4_inverse(const double* in, double* out) { return b2r::m4_inverse(in, out) ? B2R_OK : B2R_ERR_SINGULAR; }
B2R_API void b2r_m4_mul(const double* a, const double* b, double* out) { b2r::m4_mul(a, b, out); }
// shard owner of a volume unit: the rank whose handle allocates and fuses unit (kx,ky,kz)
B2R_API int b2r_unit_owner(int kx, int ky, int kz, int world) {
    if (world < 1 || !b2r_key_in_range(kx, ky, kz)) return -1;
    return b2r_owner(b2r_pack_key(kx, ky, kz), world);
}
