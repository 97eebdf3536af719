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

#define UNP_THREADS 128
#include <type_traits>
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
    float gfx, gfy, gcx, gcy;  // depth intrinsics (f32) for the gather chain
    float Fx, Fy, Cx, Cy;      // colour intrinsics (f32)
    float d2c[12];             // depth -> colour rigid transform (f32)
    float rgfx, rgfy, r255;    // RN(1/gfx), RN(1/gfy), RN(1/255): constant divisors go through the Markstein quotient
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
};

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
        if (P.gather) {
            gi = -1;
            const float z = d[k];
            if (z > 0.0f) {
                const float x = xfdiv(xfmul(xfsub((float)j, P.gcx), z), P.gfx);
                const float y = xfdiv(xfmul(xfsub((float)i, P.gcy), z), P.gfy);
                const float Xc = xfadd(xfadd(xfadd(xfmul(P.d2c[0], x), xfmul(P.d2c[1], y)), xfmul(P.d2c[2], z)), P.d2c[3]);
                const float Yc = xfadd(xfadd(xfadd(xfmul(P.d2c[4], x), xfmul(P.d2c[5], y)), xfmul(P.d2c[6], z)), P.d2c[7]);
                const float Zc = xfadd(xfadd(xfadd(xfmul(P.d2c[8], x), xfmul(P.d2c[9], y)), xfmul(P.d2c[10], z)), P.d2c[11]);
                if (Zc > 0.0f) {
                    const float uf = xfadd(xfadd(xfdiv(xfmul(Xc, P.Fx), Zc), P.Cx), 0.5f);
                    const float vf = xfadd(xfadd(xfdiv(xfmul(Yc, P.Fy), Zc), P.Cy), 0.5f);
                    const int uc = (int)floorf(uf), vc = (int)floorf(vf);
                    if (uc >= 0 && uc < P.Wc && vc >= 0 && vc < P.Hc) gi = vc * P.Wc + uc;
                }
            }
        }
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

// ---- f32 throughput path -------------------------------------------------------------------------------
// Same integer results (validity, raster order, gather indices) and the same IEEE-rounded camera-space
// x, y, colours as the generic kernel, but: constant divisors (fx, fy, 255) use a host-computed RN(1/b) + the
// Markstein correction (3 instructions, bit-identical to div.rn for these operand ranges), the u/v projection
// shares one reciprocal, all 8 colour gathers of a thread are in flight together, the output transform uses
// FMAs (f32 mode promises a few ulp, not bit parity), and the compacted chunk leaves shared memory as
// 128-bit stores whose alignment phase matches the global destination.
// CMODE: 0 no colour, 1 f32 gray, 3 / 4 u8 channels.  GATHER: colour lives in its own camera.
template <int CMODE, bool GATHER, bool WANT_IDX>
__global__ void __launch_bounds__(UNP_THREADS) unproject_fill_f32_kernel(const __grid_constant__ UnprojParams P) {
    constexpr int PPT = 8, CHUNK = UNP_THREADS * PPT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s_xyz = reinterpret_cast<float*>(smem_raw);            // CHUNK*3 + 4
    float* s_rgb = s_xyz + CHUNK * 3 + 4;                          // CHUNK*3 + 4 (absent when CMODE == 0)
    int32_t* s_idx = reinterpret_cast<int32_t*>(s_rgb + (CMODE != 0 ? CHUNK * 3 + 4 : 0));   // CHUNK + 4 (WANT_IDX only)
    int32_t* s_gidx = s_idx + CHUNK + 4;                           // CHUNK + 4
    __shared__ int ws[UNP_THREADS / 32];

    const int b = blockIdx.y, chunk = blockIdx.x;
    const int s0 = (chunk * UNP_THREADS + threadIdx.x) * PPT;
    float d[PPT]; int pix[PPT];
    const bool fast_u16 = P.scale_fast && P.stride == 1 && P.depth_is_u16 && ((P.W * P.H) % 8 == 0) && s0 + 8 <= P.n_samples;
    if (fast_u16) {
        // RN(v / depth_scale) through the host-computed RN(1/depth_scale) + Markstein correction (== div.rn here)
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>((const uint16_t*)P.depth + (size_t)b * P.W * P.H + s0));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t v = (w[k >> 1] >> ((k & 1) * 16)) & 0xFFFFu;
            float dd = b2r_div_fast((float)v, P.depth_scale, P.rscale);
            if (dd >= P.depth_trunc) dd = 0.0f;
            d[k] = dd; pix[k] = s0 + k;
        }
    } else {
        load_depths<PPT>(P, b, s0, d, pix);
    }
    unsigned m = 0;
#pragma unroll
    for (int k = 0; k < PPT; k++)
        if (pix[k] >= 0 && (d[k] > 0.0f || !P.valid_only)) m |= 1u << k;
    const int cnt = __popc(m);
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += t; }
    if ((threadIdx.x & 31) == 31) ws[threadIdx.x >> 5] = incl;
    const int64_t off = (int64_t)P.chunk_offsets[b * P.n_chunks + chunk];
    __syncthreads();
    int wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < UNP_THREADS / 32; i++) { if (i < (int)(threadIdx.x >> 5)) wbase += ws[i]; total += ws[i]; }
    int rank = wbase + incl - cnt;
    const int64_t obase = (int64_t)b * P.capacity + off;
    const int ph3 = (int)((obase * 3) & 3), ph1 = (int)(obase & 3);     // alignment phase of the destinations (floats / ints)

    const double* Td = P.T + (size_t)b * 16;
    float T[12];
#pragma unroll
    for (int i = 0; i < 12; i++) T[i] = (float)__ldg(Td + i);
    const float nanv = __int_as_float(0x7fc00000);
    // pass 1: camera-space point + gather index.  (i, j) of the first pixel by one division, then incremental
    // when the thread's pixels are consecutive (stride 1).
    float X[PPT], Y[PPT];
    int gi[PPT];
    // rows_ok: the thread's 8 consecutive pixels never straddle a row (stride 1, W % 8 == 0): one division per thread
    const bool rows_ok = P.stride == 1 && (P.W & 7) == 0 && pix[0] >= 0;
    int ci = 0, cj = 0;
    if (rows_ok) { ci = pix[0] / P.W; cj = pix[0] - ci * P.W; }
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        int i = ci, j = cj + k;
        X[k] = Y[k] = nanv; gi[k] = -1;
        if (!(m & (1u << k)) || !(d[k] > 0.0f)) continue;
        const int p = pix[k];
        if (!rows_ok) { i = p / P.W; j = p - i * P.W; }
        const float z = d[k];
        X[k] = b2r_div_fast(xfmul(xfsub((float)j, P.gcx), z), P.gfx, P.rgfx);
        Y[k] = b2r_div_fast(xfmul(xfsub((float)i, P.gcy), z), P.gfy, P.rgfy);
        gi[k] = p;
        if (GATHER) {
            gi[k] = -1;
            const float Xc = xfadd(xfadd(xfadd(xfmul(P.d2c[0], X[k]), xfmul(P.d2c[1], Y[k])), xfmul(P.d2c[2], z)), P.d2c[3]);
            const float Yc = xfadd(xfadd(xfadd(xfmul(P.d2c[4], X[k]), xfmul(P.d2c[5], Y[k])), xfmul(P.d2c[6], z)), P.d2c[7]);
            const float Zc = xfadd(xfadd(xfadd(xfmul(P.d2c[8], X[k]), xfmul(P.d2c[9], Y[k])), xfmul(P.d2c[10], z)), P.d2c[11]);
            if (Zc > 0.0f) {
                float qu, qv;
                const float au = xfmul(Xc, P.Fx), av = xfmul(Yc, P.Fy);
                if (__builtin_expect(b2r_div_domain(Zc), 1)) { const float inv = b2r_rcp_markstein(Zc); qu = b2r_div_fast(au, Zc, inv); qv = b2r_div_fast(av, Zc, inv); }
                else { qu = xfdiv(au, Zc); qv = xfdiv(av, Zc); }
                const float uf = xfadd(xfadd(qu, P.Cx), 0.5f), vf = xfadd(xfadd(qv, P.Cy), 0.5f);
                const int uc = (int)floorf(uf), vc = (int)floorf(vf);
                if ((unsigned)uc < (unsigned)P.Wc && (unsigned)vc < (unsigned)P.Hc) gi[k] = vc * P.Wc + uc;
            }
        }
    }
    // pass 2: all colour gathers in flight together
    uint32_t cw[PPT];
    const size_t cbase = (size_t)b * (GATHER ? (size_t)P.Wc * P.Hc : (size_t)P.W * P.H);
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        cw[k] = 0;
        if (CMODE != 0 && (m & (1u << k))) {
            if (CMODE == 4) { if (gi[k] >= 0) cw[k] = __ldg(reinterpret_cast<const uint32_t*>(P.color) + cbase + gi[k]); }
            else if (CMODE == 3) { if (gi[k] >= 0) { const uint8_t* c = P.color + (cbase + gi[k]) * 3; cw[k] = (uint32_t)__ldg(c) | ((uint32_t)__ldg(c + 1) << 8) | ((uint32_t)__ldg(c + 2) << 16); } }
            else cw[k] = __float_as_uint(__ldg(P.gray + (size_t)b * P.W * P.H + pix[k]));
        }
    }
    // pass 3: output transform, stage in shared memory at the destination's alignment phase
    const uint32_t sel = P.color_bgr ? 0x4012u : 0x4210u;   // PRMT selector: swap bytes 0 and 2 (BGR) or keep; byte 3 cleared
#pragma unroll
    for (int k = 0; k < PPT; k++) {
        if (!(m & (1u << k))) continue;
        float ox = nanv, oy = nanv, oz = nanv;
        if (d[k] > 0.0f) {
            const float z = d[k];
            ox = __fmaf_rn(T[0], X[k], __fmaf_rn(T[1], Y[k], __fmaf_rn(T[2], z, T[3])));
            oy = __fmaf_rn(T[4], X[k], __fmaf_rn(T[5], Y[k], __fmaf_rn(T[6], z, T[7])));
            oz = __fmaf_rn(T[8], X[k], __fmaf_rn(T[9], Y[k], __fmaf_rn(T[10], z, T[11])));
        }
        float* q = s_xyz + ph3 + 3 * rank;
        q[0] = ox; q[1] = oy; q[2] = oz;
        if (WANT_IDX) { s_idx[ph1 + rank] = pix[k]; s_gidx[ph1 + rank] = gi[k]; }
        if (CMODE != 0) {
            float r = 0.f, g = 0.f, bl = 0.f;
            if (CMODE >= 3) {
                if (gi[k] >= 0) {
                    const uint32_t c = __byte_perm(cw[k], 0, sel);
                    r = b2r_div_fast((float)(c & 255u), 255.0f, P.r255);
                    g = b2r_div_fast((float)((c >> 8) & 255u), 255.0f, P.r255);
                    bl = b2r_div_fast((float)(c >> 16), 255.0f, P.r255);
                }
            } else {
                r = g = bl = __uint_as_float(cw[k]);
            }
            float* qc = s_rgb + ph3 + 3 * rank;
            qc[0] = r; qc[1] = g; qc[2] = bl;
        }
        rank++;
    }
    __syncthreads();
    // ---- write-out: scalar head up to 16-byte alignment, 128-bit body (fully unrolled), scalar tail
    if (off + total > P.capacity) total = (int)max((int64_t)0, P.capacity - off);
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
            if (t < nv) gv[t] = sv[t];
        }
        const int t = head + (nv << 2) + threadIdx.x;
        if (t < n) g[t] = s[ph + t];
    };
    copy_out(reinterpret_cast<float*>(P.xyz) + obase * 3, s_xyz, total * 3, ph3, std::integral_constant<int, CHUNK * 3 / 4>());
    if (CMODE != 0) copy_out(reinterpret_cast<float*>(P.rgb) + obase * 3, s_rgb, total * 3, ph3, std::integral_constant<int, CHUNK * 3 / 4>());
    if (WANT_IDX && P.src_index) copy_out(reinterpret_cast<float*>(P.src_index) + obase, reinterpret_cast<const float*>(s_idx), total, ph1, std::integral_constant<int, CHUNK / 4>());
    if (WANT_IDX && P.gather_index) copy_out(reinterpret_cast<float*>(P.gather_index) + obase, reinterpret_cast<const float*>(s_gidx), total, ph1, std::integral_constant<int, CHUNK / 4>());
}

template <int CMODE, bool GATHER, bool WANT_IDX>
static int launch_fill_f32_t2(const UnprojParams& P, cudaStream_t stream) {
    constexpr int CHUNK = UNP_THREADS * 8;
    const size_t smem = (size_t)(CHUNK * 3 + 4) * 4 * (CMODE != 0 ? 2 : 1) + (WANT_IDX ? (size_t)(CHUNK + 4) * 4 * 2 : 0);
    static bool attr_set = false;
    if (!attr_set) {
        B2R_CUDA(cudaFuncSetAttribute(unproject_fill_f32_kernel<CMODE, GATHER, WANT_IDX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    dim3 grid(P.n_chunks, P.B);
    unproject_fill_f32_kernel<CMODE, GATHER, WANT_IDX><<<grid, UNP_THREADS, smem, stream>>>(P);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

template <int CMODE, bool GATHER>
static int launch_fill_f32_t(const UnprojParams& P, cudaStream_t stream) {
    return (P.src_index || P.gather_index) ? launch_fill_f32_t2<CMODE, GATHER, true>(P, stream) : launch_fill_f32_t2<CMODE, GATHER, false>(P, stream);
}

static int launch_fill_f32(const UnprojParams& P, cudaStream_t stream) {
    const int cmode = (P.rgb && P.color) ? P.color_channels : ((P.rgb && P.gray) ? 1 : 0);
    if (cmode == 4) return P.gather ? launch_fill_f32_t<4, true>(P, stream) : launch_fill_f32_t<4, false>(P, stream);
    if (cmode == 3) return P.gather ? launch_fill_f32_t<3, true>(P, stream) : launch_fill_f32_t<3, false>(P, stream);
    if (cmode == 1) return launch_fill_f32_t<1, false>(P, stream);
    return launch_fill_f32_t<0, false>(P, stream);
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

// Scratch needed by b2r_unproject for a batch (int32 words): 2 * B * n_chunks
B2R_API int64_t b2r_unproject_scratch_words(int W, int H, int B, int stride, int out_f64) {
    const int Ws = (W + stride - 1) / stride, Hs = (H + stride - 1) / stride;
    const int chunk = UNP_THREADS * (out_f64 ? 4 : 8);
    const int n_chunks = (Ws * Hs + chunk - 1) / chunk;
    return (int64_t)2 * B * n_chunks;
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
    P.gfx = (float)depth_K[0]; P.gfy = (float)depth_K[1]; P.gcx = (float)depth_K[2]; P.gcy = (float)depth_K[3];
    P.rgfx = 1.0f / P.gfx; P.rgfy = 1.0f / P.gfy; P.r255 = 1.0f / 255.0f;   // host IEEE division: correctly rounded reciprocals
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
    if (P.gather) {
        P.Fx = (float)color_K[0]; P.Fy = (float)color_K[1]; P.Cx = (float)color_K[2]; P.Cy = (float)color_K[3];
        for (int i = 0; i < 12; i++) P.d2c[i] = (float)d2c_host[i];
    }
    P.T = T_dev; P.valid_only = valid_only;
    P.xyz = xyz; P.rgb = (color || gray) ? rgb : nullptr; P.src_index = src_index; P.gather_index = gather_index;
    P.capacity = capacity;
    P.chunk_counts = scratch; P.chunk_offsets = scratch + (size_t)B * P.n_chunks; P.counts = counts;
    dim3 grid(P.n_chunks, B);
    if (out_f64) unproject_count_kernel<4><<<grid, UNP_THREADS, 0, stream>>>(P);
    else unproject_count_kernel<8><<<grid, UNP_THREADS, 0, stream>>>(P);
    B2R_CHECK_LAUNCH();
    unproject_scan_kernel<<<B, 1024, 0, stream>>>(P);
    B2R_CHECK_LAUNCH();
    if (out_f64) return launch_fill<double, 4>(P, stream);
    return launch_fill_f32(P, stream);
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
B2R_API int b2r_rgb_to_intensity(const uint8_t* rgb, int64_t n, int channels, int bgr, float* out, cudaStream_t stream) {
    if (!rgb || !out || n < 0 || (channels != 3 && channels != 4)) return B2R_ERR_BAD_ARG;
    if (n) rgb_to_intensity_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(rgb, n, channels, bgr, out);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// host helpers shared with the Python shim so pose algebra is done once, in f64, the same way everywhere
B2R_API int b2r_m4_inverse(const double* in, double* out) { return b2r::m4_inverse(in, out) ? B2R_OK : B2R_ERR_SINGULAR; }
B2R_API void b2r_m4_mul(const double* a, const double* b, double* out) { b2r::m4_mul(a, b, out); }
// shard owner of a volume unit: the rank whose handle allocates and fuses unit (kx,ky,kz)
B2R_API int b2r_unit_owner(int kx, int ky, int kz, int world) {
    if (world < 1 || !b2r_key_in_range(kx, ky, kz)) return -1;
    return b2r_owner(b2r_pack_key(kx, ky, kz), world);
}
