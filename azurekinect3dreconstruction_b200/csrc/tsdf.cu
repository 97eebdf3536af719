// tsdf.cu -- ScalableTSDFVolume on sm_100a: spatial hash of 16^3 volume units + projective fusion.
//
// Replaces, behind the C-ABI in include/b2recon.h, what the reference reaches through
//   o3d.pipelines.integration.ScalableTSDFVolume(...).integrate(rgbd, intrinsic, extrinsic)
// (/root/reference/CodeThatWorks/GoodColorReconstruction.py:111-115,270;
//  CodeThatWorks/bundlefusionattemps/bundlefusion_reconstruction.py:108-112,276,315).
// Semantics restated in SURVEY.md Appendix A.4; the f32 operation order of the per-voxel update is
// reproduced exactly (explicit __f*_rn intrinsics, no FMA contraction) so TSDF/weight match the CPU
// oracle bit for bit and the (int)u_f / (int)v_f pixel choices can never diverge.
//
// Per frame two kernels run on the caller's stream:
//   K1 tsdf_pack_alloc_kernel : one thread per pixel. Fuses RGBDImage conversion (u16 -> f32 metres,
//        trunc), the depth-to-camera-distance multiplier and the colour fetch into ONE 16-byte texel
//        {d, mult, rgb8, 0}; every depth_sampling_stride-th pixel also unprojects in f64 and inserts
//        the integer AABB of touched unit keys into the hash (allocation = pointer bump into a
//        zeroed pool), appending first touches to the frame's touched list.
//   K2 tsdf_fuse_kernel : persistent grid over (touched unit x 4 x-slabs); a thread owns 4
//        consecutive z voxels (one 128-bit load/store per plane), decides from the texel gather
//        alone whether any of them changes and only then touches the voxel record, so DRAM
//        traffic follows the updated voxels (sector granular), not the block.
#include "tsdf_internal.h"
#include "host_math.h"
#include <new>
#include <string.h>

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
B2R_API int b2r_version(void) { return 100; }

// ------------------------------------------------------------------------------------------------
// kernel parameters
// ------------------------------------------------------------------------------------------------
struct PackParams {
    int W, H;
    const void* depth;        // u16 or f32, W*H
    int depth_is_u16;
    float depth_scale, depth_trunc;
    const uint8_t* color;     // u8, channels 3|4 (may be null)
    int color_channels, color_bgr;
    const float* gray;        // f32 intensity (Gray32), may be null
    float4* texel;
    float fcx, fcy, ifx, ify;      // (float)cx, (float)cy, 1.0f/(float)fx, 1.0f/(float)fy
    double fx, fy, cx, cy;         // f64 intrinsics for the allocation samples
    double pose[12];               // extrinsic^-1, rows 0..2
    double trunc, unit_length;
    int stride;
    int frame_id, rank, world;
    B2RHash hash;
    int32_t* slot_stamp;
    int32_t* unit_keys;
    int32_t* touched;
    int32_t* counters;
    int touched_word;
    int cap_units;
};

__device__ __forceinline__ void alloc_key(const PackParams& P, int kx, int ky, int kz) {
    if (!b2r_key_in_range(kx, ky, kz)) { atomicOr(&P.counters[B2R_C_ERR], 2); return; }
    const uint64_t key = b2r_pack_key(kx, ky, kz);
    if (P.world > 1 && b2r_owner(key, P.world) != P.rank) return;
    uint32_t s = (uint32_t)b2r_mix64(key) & P.hash.mask;
    for (uint32_t probe = 0; probe <= P.hash.mask; probe++) {
        unsigned long long k = P.hash.keys[s];
        if (k == B2R_KEY_EMPTY) {
            k = atomicCAS((unsigned long long*)&P.hash.keys[s], (unsigned long long)B2R_KEY_EMPTY, (unsigned long long)key);
            if (k == B2R_KEY_EMPTY) {
                // we own the slot: bump-allocate a zeroed unit
                int id = atomicAdd(&P.counters[B2R_C_NUM_UNITS], 1);
                if (id >= P.cap_units) {
                    atomicOr(&P.counters[B2R_C_ERR], 1);
                    id = -1;
                } else {
                    P.unit_keys[3 * id] = kx; P.unit_keys[3 * id + 1] = ky; P.unit_keys[3 * id + 2] = kz;
                }
                P.hash.vals[s] = id;   // consumed by the NEXT kernel (K2), no intra-kernel reader
                k = key;
            }
        }
        if (k == key) {
            // first touch of this slot in this frame -> append to the touched list
            if (P.slot_stamp[s] != P.frame_id) {
                if (atomicExch(&P.slot_stamp[s], P.frame_id) != P.frame_id) {
                    int pos = atomicAdd(&P.counters[P.touched_word], 1);
                    if (pos < P.cap_units) P.touched[pos] = (int)s;
                }
            }
            return;
        }
        s = (s + 1) & P.hash.mask;
    }
    atomicOr(&P.counters[B2R_C_ERR], 4);
}

__global__ void __launch_bounds__(256) tsdf_pack_alloc_kernel(const PackParams P) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.W * P.H) return;
    const int i = idx / P.W, j = idx - i * P.W;
    // ---- a1: depth conversion (Image::ConvertDepthToFloatImage) fused
    float d;
    if (P.depth_is_u16) {
        d = xfdiv((float)((const uint16_t*)P.depth)[idx], P.depth_scale);
        if (d >= P.depth_trunc) d = 0.0f;
    } else {
        d = ((const float*)P.depth)[idx];
    }
    // ---- CreateDepthToCameraDistanceMultiplierFloatImage
    const float xx = xfmul(xfsub((float)j, P.fcx), P.ifx);
    const float yy = xfmul(xfsub((float)i, P.fcy), P.ify);
    const float mult = __fsqrt_rn(xfadd(xfadd(xfmul(xx, xx), xfmul(yy, yy)), 1.0f));
    float cbits = 0.0f;
    if (P.color) {
        const uint8_t* c = P.color + (size_t)idx * P.color_channels;
        uint32_t r = c[0], g = c[1], b = c[2];
        if (P.color_bgr) { uint32_t t = r; r = b; b = t; }
        cbits = __uint_as_float(r | (g << 8) | (b << 16));
    } else if (P.gray) {
        cbits = P.gray[idx];
    }
    P.texel[idx] = make_float4(d, mult, cbits, 0.0f);

    // ---- a8: block allocation from the stride-sampled pixels (f64 chain)
    if ((i % P.stride) == 0 && (j % P.stride) == 0 && d > 0.0f) {
        const double z = (double)d;
        const double x = xddiv(xdmul(xdsub((double)j, P.cx), z), P.fx);
        const double y = xddiv(xdmul(xdsub((double)i, P.cy), z), P.fy);
        int lo[3], hi[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const double p = xdadd(xdadd(xdadd(xdmul(P.pose[4 * k], x), xdmul(P.pose[4 * k + 1], y)), xdmul(P.pose[4 * k + 2], z)), P.pose[4 * k + 3]);
            lo[k] = (int)floor(xddiv(xdsub(p, P.trunc), P.unit_length));
            hi[k] = (int)floor(xddiv(xdadd(p, P.trunc), P.unit_length));
        }
        for (int a = lo[0]; a <= hi[0]; a++)
            for (int b = lo[1]; b <= hi[1]; b++)
                for (int c = lo[2]; c <= hi[2]; c++) alloc_key(P, a, b, c);
    }
}

struct FuseParams {
    int W, H;
    const float4* texel;
    float fx, fy, cx, cy;
    float E[12];             // (float)extrinsic rows 0..2
    float Es02, Es12, Es22;  // E(:,2) * voxel_length (f32)
    float vl, hv, trunc, inv_trunc, safe_w, safe_h;
    double unit_length;
    const int32_t* touched;
    const int32_t* hash_vals;
    const int32_t* unit_keys;
    float* pool;
    int64_t unit_stride;     // floats per unit
    int32_t* counters;
    int touched_word;
    int cap_units;
};

template <int COLOR>
__global__ void __launch_bounds__(256) tsdf_fuse_kernel(const FuseParams P) {
    const int T = min(P.counters[P.touched_word], P.cap_units);
    const int nwork = T * 4;
    const int tid = threadIdx.x;
    const int zq = tid & 3, y = (tid >> 2) & 15, xl = tid >> 6;
    unsigned int upd = 0;
    for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
        const int slot = P.touched[w >> 2];
        const int unit = P.hash_vals[slot];
        if (unit < 0) continue;
        const int x = ((w & 3) << 2) + xl;
        const int kx = P.unit_keys[3 * unit], ky = P.unit_keys[3 * unit + 1], kz = P.unit_keys[3 * unit + 2];
        // voxel-centre column base: float(hv + vl*x + origin.x) with the f32->f64 promotion of upstream
        const float px = (float)xdadd((double)xfadd(P.hv, xfmul(P.vl, (float)x)), xdmul((double)kx, P.unit_length));
        const float py = (float)xdadd((double)xfadd(P.hv, xfmul(P.vl, (float)y)), xdmul((double)ky, P.unit_length));
        const float pz = (float)xdadd((double)P.hv, xdmul((double)kz, P.unit_length));
        float pcx = xfadd(xfadd(xfadd(xfmul(P.E[0], px), xfmul(P.E[1], py)), xfmul(P.E[2], pz)), P.E[3]);
        float pcy = xfadd(xfadd(xfadd(xfmul(P.E[4], px), xfmul(P.E[5], py)), xfmul(P.E[6], pz)), P.E[7]);
        float pcz = xfadd(xfadd(xfadd(xfmul(P.E[8], px), xfmul(P.E[9], py)), xfmul(P.E[10], pz)), P.E[11]);
        // sequential z-march to this thread's first voxel (same f32 accumulation as the reference loop)
        for (int s = 0; s < 4 * zq; s++) {
            pcx = xfadd(pcx, P.Es02); pcy = xfadd(pcy, P.Es12); pcz = xfadd(pcz, P.Es22);
        }
        float tv[4];
        float cv[4];
        unsigned act = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            tv[k] = 0.f; cv[k] = 0.f;
            if (pcz > 0.0f) {
                const float u_f = xfadd(xfadd(xfdiv(xfmul(pcx, P.fx), pcz), P.cx), 0.5f);
                const float v_f = xfadd(xfadd(xfdiv(xfmul(pcy, P.fy), pcz), P.cy), 0.5f);
                if (u_f >= 0.0001f && u_f < P.safe_w && v_f >= 0.0001f && v_f < P.safe_h) {
                    const int u = (int)u_f, v = (int)v_f;
                    const float4 tx = __ldg(&P.texel[v * P.W + u]);
                    if (tx.x > 0.0f) {
                        const float sdf = xfmul(xfsub(tx.x, pcz), tx.y);
                        if (sdf > -P.trunc) {
                            tv[k] = fminf(1.0f, xfmul(sdf, P.inv_trunc));
                            cv[k] = tx.z;
                            act |= 1u << k;
                        }
                    }
                }
            }
            pcx = xfadd(pcx, P.Es02); pcy = xfadd(pcy, P.Es12); pcz = xfadd(pcz, P.Es22);
        }
        if (act) {
            float* base = P.pool + (int64_t)unit * P.unit_stride + (x * 256 + y * 16 + zq * 4);
            float4 ts = *reinterpret_cast<float4*>(base);
            float4 ws = *reinterpret_cast<float4*>(base + B2R_UNIT_VOX);
            float4 rs, gs, bs;
            if (COLOR) {
                rs = *reinterpret_cast<float4*>(base + 2 * B2R_UNIT_VOX);
                gs = *reinterpret_cast<float4*>(base + 3 * B2R_UNIT_VOX);
                bs = *reinterpret_cast<float4*>(base + 4 * B2R_UNIT_VOX);
            }
            float* tp = reinterpret_cast<float*>(&ts);
            float* wp = reinterpret_cast<float*>(&ws);
            float* rp = reinterpret_cast<float*>(&rs);
            float* gp = reinterpret_cast<float*>(&gs);
            float* bp = reinterpret_cast<float*>(&bs);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (act & (1u << k)) {
                    const float wgt = wp[k];
                    const float w1 = xfadd(wgt, 1.0f);
                    tp[k] = xfdiv(xfadd(xfmul(tp[k], wgt), tv[k]), w1);
                    if (COLOR == 1) {
                        const uint32_t c = __float_as_uint(cv[k]);
                        rp[k] = xfdiv(xfadd(xfmul(rp[k], wgt), (float)(c & 255u)), w1);
                        gp[k] = xfdiv(xfadd(xfmul(gp[k], wgt), (float)((c >> 8) & 255u)), w1);
                        bp[k] = xfdiv(xfadd(xfmul(bp[k], wgt), (float)((c >> 16) & 255u)), w1);
                    } else if (COLOR == 2) {
                        rp[k] = xfdiv(xfadd(xfmul(rp[k], wgt), cv[k]), w1);
                        gp[k] = xfdiv(xfadd(xfmul(gp[k], wgt), cv[k]), w1);
                        bp[k] = xfdiv(xfadd(xfmul(bp[k], wgt), cv[k]), w1);
                    }
                    wp[k] = w1;
                    upd++;
                }
            }
            *reinterpret_cast<float4*>(base) = ts;
            *reinterpret_cast<float4*>(base + B2R_UNIT_VOX) = ws;
            if (COLOR) {
                *reinterpret_cast<float4*>(base + 2 * B2R_UNIT_VOX) = rs;
                *reinterpret_cast<float4*>(base + 3 * B2R_UNIT_VOX) = gs;
                *reinterpret_cast<float4*>(base + 4 * B2R_UNIT_VOX) = bs;
            }
        }
    }
    // N_upd accounting: warp shuffle -> one atomic per warp
    for (int o = 16; o > 0; o >>= 1) upd += __shfl_down_sync(0xffffffffu, upd, o);
    if ((tid & 31) == 0) {
        if (upd) atomicAdd(reinterpret_cast<unsigned long long*>(&P.counters[B2R_C_UPDATED]), (unsigned long long)upd);
        if (blockIdx.x == 0 && tid == 0)
            atomicAdd(reinterpret_cast<unsigned long long*>(&P.counters[B2R_C_TOUCHED_TOTAL]), (unsigned long long)T);
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

static void tsdf_free(b2r_tsdf* h) {
    if (!h) return;
    cudaFree(h->hash.keys); cudaFree(h->hash.vals); cudaFree(h->slot_stamp); cudaFree(h->unit_keys);
    cudaFree(h->pool); cudaFree(h->touched[0]); cudaFree(h->touched[1]); cudaFree(h->counters);
    cudaFree(h->texel[0]); cudaFree(h->texel[1]); cudaFree(h->scratch);
    for (int i = 0; i < B2R_HOST_RING; i++) {
        if (h->ev_copied[i]) cudaEventDestroy(h->ev_copied[i]);
        if (h->ev_consumed[i]) cudaEventDestroy(h->ev_consumed[i]);
        cudaFree(h->stage_depth[i]); cudaFree(h->stage_color[i]);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (int i = 0; i < h->prof_cap; i++) cudaEventDestroy(h->prof_ev[i]);
    delete[] h->prof_ev;
    delete h;
}

B2R_API b2r_tsdf* b2r_tsdf_create(double voxel_length, double sdf_trunc, int color_type, int unit_resolution,
                                  int depth_sampling_stride, int64_t capacity_units, int rank, int world) {
    if (unit_resolution != B2R_UNIT_RES) { b2r_set_error("volume_unit_resolution must be 16 (got %d)", unit_resolution); return nullptr; }
    if (!(voxel_length > 0) || !(sdf_trunc > 0) || depth_sampling_stride < 1 || capacity_units < 1 || color_type < 0 ||
        color_type > 2 || world < 1 || rank < 0 || rank >= world) {
        b2r_set_error("b2r_tsdf_create: bad argument");
        return nullptr;
    }
    b2r_tsdf* h = new (std::nothrow) b2r_tsdf();
    if (!h) return nullptr;
    memset(h, 0, sizeof(*h));
    h->voxel_length = voxel_length; h->sdf_trunc = sdf_trunc;
    h->unit_length = voxel_length * unit_resolution;
    h->color_type = color_type; h->res = unit_resolution; h->stride = depth_sampling_stride;
    h->planes = color_type == 0 ? 2 : 5;
    h->cap_units = capacity_units; h->rank = rank; h->world = world;
    cudaGetDevice(&h->device);
    h->nslots = next_pow2((uint64_t)capacity_units * 2);
    h->hash.mask = h->nslots - 1;
    h->frame_id = 0;
    const size_t pool_bytes = (size_t)capacity_units * h->planes * B2R_UNIT_VOX * sizeof(float);
    cudaError_t e = cudaSuccess;
    auto A = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
    A((void**)&h->hash.keys, (size_t)h->nslots * 8);
    A((void**)&h->hash.vals, (size_t)h->nslots * 4);
    A((void**)&h->slot_stamp, (size_t)h->nslots * 4);
    A((void**)&h->unit_keys, (size_t)capacity_units * 12);
    A((void**)&h->touched[0], (size_t)capacity_units * 4);
    A((void**)&h->touched[1], (size_t)capacity_units * 4);
    A((void**)&h->counters, B2R_C_WORDS * 4);
    A((void**)&h->pool, pool_bytes);
    if (e == cudaSuccess) e = cudaMemset(h->hash.keys, 0xFF, (size_t)h->nslots * 8);
    if (e == cudaSuccess) e = cudaMemset(h->hash.vals, 0xFF, (size_t)h->nslots * 4);
    if (e == cudaSuccess) e = cudaMemset(h->slot_stamp, 0, (size_t)h->nslots * 4);
    if (e == cudaSuccess) e = cudaMemset(h->counters, 0, B2R_C_WORDS * 4);
    if (e == cudaSuccess) e = cudaMemset(h->pool, 0, pool_bytes);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking);
    for (int i = 0; i < B2R_HOST_RING && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&h->ev_copied[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_consumed[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) {
        int nsm = 148, per_sm = 8;
        cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, h->device);
        if (color_type == 0) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tsdf_fuse_kernel<0>, 256, 0);
        else if (color_type == 1) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tsdf_fuse_kernel<1>, 256, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tsdf_fuse_kernel<2>, 256, 0);
        if (per_sm < 1) per_sm = 1;
        h->fuse_grid = nsm * per_sm;
        e = cudaGetLastError();
    }
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

// volume.reset()  (bundlefusion_reconstruction.py:309, capture_frames.py:94)
B2R_API int b2r_tsdf_reset(b2r_tsdf* h, cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int32_t n = 0;
    B2R_CUDA(cudaMemcpyAsync(&n, h->counters + B2R_C_NUM_UNITS, 4, cudaMemcpyDeviceToHost, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    if (n > h->cap_units) n = (int32_t)h->cap_units;
    B2R_CUDA(cudaMemsetAsync(h->pool, 0, (size_t)n * h->planes * B2R_UNIT_VOX * sizeof(float), stream));
    B2R_CUDA(cudaMemsetAsync(h->hash.keys, 0xFF, (size_t)h->nslots * 8, stream));
    B2R_CUDA(cudaMemsetAsync(h->hash.vals, 0xFF, (size_t)h->nslots * 4, stream));
    B2R_CUDA(cudaMemsetAsync(h->slot_stamp, 0, (size_t)h->nslots * 4, stream));
    B2R_CUDA(cudaMemsetAsync(h->counters, 0, B2R_C_WORDS * 4, stream));
    h->frame_id = 0;
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

// ------------------------------------------------------------------------------------------------
// integrate (device inputs)
// ------------------------------------------------------------------------------------------------
static int tsdf_integrate_impl(b2r_tsdf* h, const void* depth, int depth_is_u16, float depth_scale, float depth_trunc,
                               const uint8_t* color, int color_channels, int color_bgr, const float* gray, int W,
                               int H, double fx, double fy, double cx, double cy, const double* extrinsic,
                               cudaStream_t stream) {
    if (!h || !depth || !extrinsic || W < 1 || H < 1) { b2r_set_error("b2r_tsdf_integrate: bad argument"); return B2R_ERR_BAD_ARG; }
    // upstream: "[ScalableTSDFVolume::Integrate] Unsupported image format."
    if (h->color_type == 1 && (!color || (color_channels != 3 && color_channels != 4))) { b2r_set_error("[ScalableTSDFVolume::Integrate] Unsupported image format."); return B2R_ERR_FORMAT; }
    if (h->color_type == 2 && !gray) { b2r_set_error("[ScalableTSDFVolume::Integrate] Unsupported image format."); return B2R_ERR_FORMAT; }
    double pose[16];
    if (!b2r::m4_inverse(extrinsic, pose)) { b2r_set_error("b2r_tsdf_integrate: singular extrinsic"); return B2R_ERR_SINGULAR; }
    const size_t npx = (size_t)W * H;
    if (h->texel_px < npx) {
        B2R_CUDA(cudaStreamSynchronize(stream));
        for (int i = 0; i < 2; i++) { if (h->texel[i]) cudaFree(h->texel[i]); h->texel[i] = nullptr; }
        h->texel_px = 0;
        for (int i = 0; i < 2; i++) B2R_CUDA(cudaMalloc((void**)&h->texel[i], npx * sizeof(float4)));
        h->texel_px = npx;
    }
    h->frame_id++;
    const int buf = h->frame_id & 1;
    const int tw = buf ? B2R_C_TOUCHED1 : B2R_C_TOUCHED0;
    B2R_CUDA(cudaMemsetAsync(h->counters + tw, 0, 4, stream));

    PackParams P;
    memset(&P, 0, sizeof(P));
    P.W = W; P.H = H; P.depth = depth; P.depth_is_u16 = depth_is_u16;
    P.depth_scale = depth_scale; P.depth_trunc = depth_trunc;
    P.color = h->color_type == 1 ? color : nullptr; P.color_channels = color_channels; P.color_bgr = color_bgr;
    P.gray = h->color_type == 2 ? gray : nullptr;
    P.texel = h->texel[buf];
    P.fcx = (float)cx; P.fcy = (float)cy; P.ifx = 1.0f / (float)fx; P.ify = 1.0f / (float)fy;
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
    for (int i = 0; i < 12; i++) P.pose[i] = pose[i];
    P.trunc = h->sdf_trunc; P.unit_length = h->unit_length; P.stride = h->stride;
    P.frame_id = h->frame_id; P.rank = h->rank; P.world = h->world;
    P.hash = h->hash; P.slot_stamp = h->slot_stamp; P.unit_keys = h->unit_keys; P.touched = h->touched[buf];
    P.counters = h->counters; P.touched_word = tw; P.cap_units = (int)h->cap_units;
    const bool prof = h->profile && h->prof_n + 3 <= h->prof_cap;
    if (prof) cudaEventRecord(h->prof_ev[h->prof_n], stream);
    tsdf_pack_alloc_kernel<<<(unsigned)((npx + 255) / 256), 256, 0, stream>>>(P);
    B2R_CHECK_LAUNCH();
    if (prof) cudaEventRecord(h->prof_ev[h->prof_n + 1], stream);

    FuseParams F;
    memset(&F, 0, sizeof(F));
    F.W = W; F.H = H; F.texel = h->texel[buf];
    F.fx = (float)fx; F.fy = (float)fy; F.cx = (float)cx; F.cy = (float)cy;
    for (int i = 0; i < 12; i++) F.E[i] = (float)extrinsic[i];
    F.vl = (float)h->voxel_length; F.hv = F.vl * 0.5f;
    F.trunc = (float)h->sdf_trunc; F.inv_trunc = 1.0f / F.trunc;
    F.Es02 = F.E[2] * F.vl; F.Es12 = F.E[6] * F.vl; F.Es22 = F.E[10] * F.vl;
    F.safe_w = (float)W - 0.0001f; F.safe_h = (float)H - 0.0001f;
    F.unit_length = h->unit_length;
    F.touched = h->touched[buf]; F.hash_vals = h->hash.vals; F.unit_keys = h->unit_keys;
    F.pool = h->pool; F.unit_stride = (int64_t)h->planes * B2R_UNIT_VOX;
    F.counters = h->counters; F.touched_word = tw; F.cap_units = (int)h->cap_units;
    if (h->color_type == 0) tsdf_fuse_kernel<0><<<h->fuse_grid, 256, 0, stream>>>(F);
    else if (h->color_type == 1) tsdf_fuse_kernel<1><<<h->fuse_grid, 256, 0, stream>>>(F);
    else tsdf_fuse_kernel<2><<<h->fuse_grid, 256, 0, stream>>>(F);
    B2R_CHECK_LAUNCH();
    if (prof) { cudaEventRecord(h->prof_ev[h->prof_n + 2], stream); h->prof_n += 3; }
    return B2R_OK;
}

// volume.integrate(rgbd, intrinsic, extrinsic) with DEVICE images (torch tensors).
//   depth: f32 metres (an RGBDImage's depth) or raw u16 with depth_scale/depth_trunc (a1 fused).
//   color: u8, 3 (RGB) or 4 channels; color_bgr=1 for the camera's BGRA order (cv2.cvtColor fused).
//   extrinsic: HOST 16 f64, row-major, world -> camera.
B2R_API int b2r_tsdf_integrate(b2r_tsdf* h, const void* depth, int depth_is_u16, float depth_scale, float depth_trunc,
                               const uint8_t* color, int color_channels, int color_bgr, const float* gray, int W,
                               int H, double fx, double fy, double cx, double cy, const double* extrinsic,
                               cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    return tsdf_integrate_impl(h, depth, depth_is_u16, depth_scale, depth_trunc, color, color_channels, color_bgr, gray,
                               W, H, fx, fy, cx, cy, extrinsic, stream);
}

// Batch of n same-shaped frames stored back to back (re-integration loop of
// bundlefusion_reconstruction.py:309-315 / capture_frames.py:92-105). extrinsics: HOST n*16 f64.
B2R_API int b2r_tsdf_integrate_batch(b2r_tsdf* h, int n, const void* depth, int depth_is_u16, float depth_scale,
                                     float depth_trunc, const uint8_t* color, int color_channels, int color_bgr,
                                     int W, int H, double fx, double fy, double cx, double cy,
                                     const double* extrinsics, cudaStream_t stream) {
    if (!h) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    const size_t npx = (size_t)W * H;
    const size_t dstep = npx * (depth_is_u16 ? 2 : 4), cstep = npx * (size_t)color_channels;
    for (int f = 0; f < n; f++) {
        int rc = tsdf_integrate_impl(h, (const char*)depth + f * dstep, depth_is_u16, depth_scale, depth_trunc,
                                     color ? color + f * cstep : nullptr, color_channels, color_bgr, nullptr, W, H, fx,
                                     fy, cx, cy, extrinsics + 16 * f, stream);
        if (rc != B2R_OK) return rc;
    }
    return B2R_OK;
}

// Same, with HOST images (pinned for true overlap): each frame's H2D copy runs on an internal copy
// stream into a 3-deep device ring while the previous frame fuses on `stream`.
B2R_API int b2r_tsdf_integrate_host(b2r_tsdf* h, int n, const void* depth_host, int depth_is_u16, float depth_scale,
                                    float depth_trunc, const uint8_t* color_host, int color_channels, int color_bgr,
                                    int W, int H, double fx, double fy, double cx, double cy,
                                    const double* extrinsics, cudaStream_t stream) {
    if (!h || !depth_host) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    const size_t npx = (size_t)W * H;
    const size_t dbytes = npx * (depth_is_u16 ? 2 : 4), cbytes = color_host ? npx * (size_t)color_channels : 0;
    if (h->stage_depth_bytes < dbytes || h->stage_color_bytes < cbytes) {
        B2R_CUDA(cudaStreamSynchronize(stream));
        B2R_CUDA(cudaStreamSynchronize(h->copy_stream));
        for (int i = 0; i < B2R_HOST_RING; i++) {
            cudaFree(h->stage_depth[i]); cudaFree(h->stage_color[i]);
            h->stage_depth[i] = h->stage_color[i] = nullptr;
        }
        h->stage_depth_bytes = h->stage_color_bytes = 0;
        for (int i = 0; i < B2R_HOST_RING; i++) {
            B2R_CUDA(cudaMalloc(&h->stage_depth[i], dbytes));
            B2R_CUDA(cudaMalloc(&h->stage_color[i], cbytes ? cbytes : 16));
        }
        h->stage_depth_bytes = dbytes; h->stage_color_bytes = cbytes ? cbytes : 16;
        h->host_seq = 0;
    }
    for (int f = 0; f < n; f++) {
        const int r = (int)(h->host_seq % B2R_HOST_RING);
        if (h->host_seq >= B2R_HOST_RING) B2R_CUDA(cudaStreamWaitEvent(h->copy_stream, h->ev_consumed[r], 0));
        B2R_CUDA(cudaMemcpyAsync(h->stage_depth[r], (const char*)depth_host + f * dbytes, dbytes, cudaMemcpyHostToDevice, h->copy_stream));
        if (cbytes) B2R_CUDA(cudaMemcpyAsync(h->stage_color[r], color_host + f * cbytes, cbytes, cudaMemcpyHostToDevice, h->copy_stream));
        B2R_CUDA(cudaEventRecord(h->ev_copied[r], h->copy_stream));
        B2R_CUDA(cudaStreamWaitEvent(stream, h->ev_copied[r], 0));
        int rc = tsdf_integrate_impl(h, h->stage_depth[r], depth_is_u16, depth_scale, depth_trunc,
                                     cbytes ? (const uint8_t*)h->stage_color[r] : nullptr, color_channels, color_bgr,
                                     nullptr, W, H, fx, fy, cx, cy, extrinsics + 16 * f, stream);
        if (rc != B2R_OK) return rc;
        B2R_CUDA(cudaEventRecord(h->ev_consumed[r], stream));
        h->host_seq++;
    }
    return B2R_OK;
}

// ------------------------------------------------------------------------------------------------
// stats / export / import (parity + checkpoint hooks)
// ------------------------------------------------------------------------------------------------
// out[0]=num_units out[1]=error bits out[2]=touched units of the last frame out[3]=cumulative updated
// voxels (N_upd) out[4]=cumulative touched units.  Synchronises `stream`.
B2R_API int b2r_tsdf_stats(b2r_tsdf* h, int64_t* out, cudaStream_t stream) {
    if (!h || !out) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int32_t c[B2R_C_WORDS];
    B2R_CUDA(cudaMemcpyAsync(c, h->counters, sizeof(c), cudaMemcpyDeviceToHost, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    out[0] = c[B2R_C_NUM_UNITS] > h->cap_units ? h->cap_units : c[B2R_C_NUM_UNITS];
    out[1] = c[B2R_C_ERR];
    out[2] = c[(h->frame_id & 1) ? B2R_C_TOUCHED1 : B2R_C_TOUCHED0];
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
    const size_t us = (size_t)h->planes * B2R_UNIT_VOX * 4, pl = B2R_UNIT_VOX * 4;
    const char* src = (const char*)h->pool + first * us;
    if (tsdf) B2R_CUDA(cudaMemcpy2DAsync(tsdf, pl, src, us, pl, count, cudaMemcpyDeviceToDevice, stream));
    if (weight) B2R_CUDA(cudaMemcpy2DAsync(weight, pl, src + pl, us, pl, count, cudaMemcpyDeviceToDevice, stream));
    if (rgb && h->planes == 5) B2R_CUDA(cudaMemcpy2DAsync(rgb, 3 * pl, src + 2 * pl, us, 3 * pl, count, cudaMemcpyDeviceToDevice, stream));
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
        unsigned long long k = atomicCAS((unsigned long long*)&hash.keys[s], (unsigned long long)B2R_KEY_EMPTY, (unsigned long long)key);
        if (k == B2R_KEY_EMPTY) {
            int id = atomicAdd(&counters[B2R_C_NUM_UNITS], 1);
            if (id >= cap) { atomicOr(&counters[B2R_C_ERR], 1); id = -1; }
            else { unit_keys[3 * id] = kx; unit_keys[3 * id + 1] = ky; unit_keys[3 * id + 2] = kz; }
            hash.vals[s] = id;
            out_ids[i] = id;
            return;
        }
        if (k == key) { atomicOr(&counters[B2R_C_ERR], 8); return; }   // duplicate / already present
        s = (s + 1) & hash.mask;
    }
    atomicOr(&counters[B2R_C_ERR], 4);
}

__global__ void tsdf_import_planes_kernel(float* pool, int64_t unit_stride, int planes, const int32_t* ids, int n,
                                          const float* tsdf, const float* weight, const float* rgb) {
    const int u = blockIdx.y;
    const int id = ids[u];
    if (id < 0) return;
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < B2R_UNIT_VOX; v += gridDim.x * blockDim.x) {
        float* base = pool + (int64_t)id * unit_stride;
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
    if (!h || !keys3 || !tsdf || !weight || n < 0) return B2R_ERR_BAD_ARG;
    if (n == 0) return B2R_OK;
    B2RDeviceGuard g(h->device);
    int rc = b2r_tsdf_ensure_scratch(h, (size_t)n * 4);
    if (rc != B2R_OK) return rc;
    int32_t* ids = (int32_t*)h->scratch;
    tsdf_import_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(h->hash, h->unit_keys, h->counters, keys3, (int)n, (int)h->cap_units, ids);
    B2R_CHECK_LAUNCH();
    dim3 grid(4, (unsigned)n);
    tsdf_import_planes_kernel<<<grid, 256, 0, stream>>>(h->pool, (int64_t)h->planes * B2R_UNIT_VOX, h->planes, ids, (int)n, tsdf, weight, rgb);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// Per-kernel timing for the roofline line of bench.py: enable (max_frames > 0) / disable (0).
B2R_API int b2r_tsdf_profile(b2r_tsdf* h, int max_frames) {
    if (!h || max_frames < 0) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    const int need = max_frames * 3;
    if (need > h->prof_cap) {
        cudaEvent_t* ev = new cudaEvent_t[need];
        for (int i = 0; i < h->prof_cap; i++) ev[i] = h->prof_ev[i];
        for (int i = h->prof_cap; i < need; i++) B2R_CUDA(cudaEventCreate(&ev[i]));
        delete[] h->prof_ev;
        h->prof_ev = ev; h->prof_cap = need;
    }
    h->profile = max_frames > 0; h->prof_n = 0;
    return B2R_OK;
}
// out[0] = frames timed, out[1] = total ms in K1 (pack+alloc), out[2] = total ms in K2 (fuse). Synchronises.
B2R_API int b2r_tsdf_profile_read(b2r_tsdf* h, double* out3) {
    if (!h || !out3) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    B2R_CUDA(cudaDeviceSynchronize());
    double k1 = 0.0, k2 = 0.0;
    for (int i = 0; i + 3 <= h->prof_n; i += 3) {
        float a = 0.f, b = 0.f;
        B2R_CUDA(cudaEventElapsedTime(&a, h->prof_ev[i], h->prof_ev[i + 1]));
        B2R_CUDA(cudaEventElapsedTime(&b, h->prof_ev[i + 1], h->prof_ev[i + 2]));
        k1 += a; k2 += b;
    }
    out3[0] = h->prof_n / 3; out3[1] = k1; out3[2] = k2;
    h->prof_n = 0;
    return B2R_OK;
}
