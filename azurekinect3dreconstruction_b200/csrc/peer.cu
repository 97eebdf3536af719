// peer.cu -- peer-memory plumbing for the multi-GPU frame exchange (one process per GPU, NVLink / NVSwitch).
//
// The reference is single-process (its many-cameras-one-volume precedent: CodeThatKindaWorks/BetterVisualizerD2camerasWORKSkinda.py:
// 533-570); BASELINE.json's north_star asks that every rank see every frame.  An NCCL all-gather does that with KERNELS, which
// cannot start while a persistent fusion pass holds every SM (measured: profiles/r2_e2e_timeline_n8.json).  Instead each rank
// keeps its uploaded frames in a buffer the other ranks map through CUDA IPC, and K1 (tsdf_pack_alloc_kernel) reads the peers'
// pixels where they lie: its TMA bulk copies pull the tiles over NVLink straight into shared memory -- the transfer IS the
// kernel's load phase.  Cross-rank ordering uses stream memory operations on 32-bit counters (cuStreamWriteValue32 /
// cuStreamWaitValue32): no kernel, no SM, no host round trip, and -- unlike interprocess events -- value based, so a waiter
// can never latch on to a stale record.
#include "common.cuh"
#include <cuda.h>
#include <string.h>

namespace {
typedef CUresult (*write32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
typedef CUresult (*wait32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
write32_fn g_write32 = nullptr;
wait32_fn g_wait32 = nullptr;

int load_memops() {
    if (g_write32 && g_wait32) return B2R_OK;
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    B2R_CUDA(cudaGetDriverEntryPoint("cuStreamWriteValue32", &f, cudaEnableDefault, &q));
    if (!f || q != cudaDriverEntryPointSuccess) { b2r_set_error("cuStreamWriteValue32 is not available"); return B2R_ERR_CUDA; }
    g_write32 = (write32_fn)f;
    f = nullptr;
    B2R_CUDA(cudaGetDriverEntryPoint("cuStreamWaitValue32", &f, cudaEnableDefault, &q));
    if (!f || q != cudaDriverEntryPointSuccess) { b2r_set_error("cuStreamWaitValue32 is not available"); return B2R_ERR_CUDA; }
    g_wait32 = (wait32_fn)f;
    return B2R_OK;
}
}  // namespace

// A zeroed device buffer other processes may map: *dptr = its address here, handle64 = what the peers pass to b2r_peer_open.
B2R_API int b2r_peer_alloc(int64_t bytes, void** dptr, unsigned char* handle64) {
    if (!dptr || !handle64 || bytes < 1) return B2R_ERR_BAD_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void* p = nullptr;
    B2R_CUDA(cudaMalloc(&p, (size_t)bytes));
    cudaError_t e = cudaMemset(p, 0, (size_t)bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); b2r_set_error("b2r_peer_alloc: %s", cudaGetErrorString(e)); return B2R_ERR_CUDA; }
    memcpy(handle64, &h, 64);
    *dptr = p;
    return B2R_OK;
}
B2R_API int b2r_peer_free(void* dptr) {
    if (dptr) B2R_CUDA(cudaFree(dptr));
    return B2R_OK;
}
// Maps a peer's buffer into this process (peer access over NVLink is enabled on first use).
B2R_API int b2r_peer_open(const unsigned char* handle64, void** dptr) {
    if (!handle64 || !dptr) return B2R_ERR_BAD_ARG;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    B2R_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B2R_OK;
}
B2R_API int b2r_peer_close(void* dptr) {
    if (dptr) B2R_CUDA(cudaIpcCloseMemHandle(dptr));
    return B2R_OK;
}
// Copy-engine copy between any two device addresses this process can see (own memory or mapped peer buffers): no SM involved.
B2R_API int b2r_peer_copy(void* dst, const void* src, int64_t bytes, cudaStream_t stream) {
    if (!dst || !src || bytes < 0) return B2R_ERR_BAD_ARG;
    if (bytes) B2R_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, stream));
    return B2R_OK;
}
// Stream-ordered 32-bit write (everything queued before it on `stream` is visible to whoever sees the value) and
// stream-ordered wait until *addr >= value (wrap-safe comparison).  addr: device memory of this GPU or a mapped peer buffer.
B2R_API int b2r_stream_write32(cudaStream_t stream, void* addr, uint32_t value) {
    int rc = load_memops();
    if (rc != B2R_OK) return rc;
    const CUresult r = g_write32((CUstream)stream, (CUdeviceptr)(uintptr_t)addr, value, CU_STREAM_WRITE_VALUE_DEFAULT);
    if (r != CUDA_SUCCESS) { b2r_set_error("cuStreamWriteValue32 failed (%d)", (int)r); return B2R_ERR_CUDA; }
    return B2R_OK;
}
// the same on n addresses (one call from the host language instead of n)
B2R_API int b2r_stream_write32_multi(cudaStream_t stream, void* const* addrs, int n, uint32_t value) {
    if (!addrs || n < 0) return B2R_ERR_BAD_ARG;
    int rc = load_memops();
    for (int i = 0; i < n && rc == B2R_OK; i++) {
        const CUresult r = g_write32((CUstream)stream, (CUdeviceptr)(uintptr_t)addrs[i], value, CU_STREAM_WRITE_VALUE_DEFAULT);
        if (r != CUDA_SUCCESS) { b2r_set_error("cuStreamWriteValue32 failed (%d)", (int)r); rc = B2R_ERR_CUDA; }
    }
    return rc;
}
B2R_API int b2r_stream_wait32_geq_multi(cudaStream_t stream, void* const* addrs, int n, uint32_t value) {
    if (!addrs || n < 0) return B2R_ERR_BAD_ARG;
    int rc = load_memops();
    for (int i = 0; i < n && rc == B2R_OK; i++) {
        const CUresult r = g_wait32((CUstream)stream, (CUdeviceptr)(uintptr_t)addrs[i], value, CU_STREAM_WAIT_VALUE_GEQ);
        if (r != CUDA_SUCCESS) { b2r_set_error("cuStreamWaitValue32 failed (%d)", (int)r); rc = B2R_ERR_CUDA; }
    }
    return rc;
}
// One peer's share of a super-chunk, pulled by the copy engine: wait until *ready >= value, copy, then *done = value.
B2R_API int b2r_peer_pull(cudaStream_t stream, void* ready, void* dst, const void* src, int64_t bytes, void* done, uint32_t value) {
    if (!ready || !dst || !src || !done || bytes < 0) return B2R_ERR_BAD_ARG;
    int rc = load_memops();
    if (rc != B2R_OK) return rc;
    CUresult r = g_wait32((CUstream)stream, (CUdeviceptr)(uintptr_t)ready, value, CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) { b2r_set_error("cuStreamWaitValue32 failed (%d)", (int)r); return B2R_ERR_CUDA; }
    if (bytes) B2R_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, stream));
    r = g_write32((CUstream)stream, (CUdeviceptr)(uintptr_t)done, value, CU_STREAM_WRITE_VALUE_DEFAULT);
    if (r != CUDA_SUCCESS) { b2r_set_error("cuStreamWriteValue32 failed (%d)", (int)r); return B2R_ERR_CUDA; }
    return B2R_OK;
}
B2R_API int b2r_stream_wait32_geq(cudaStream_t stream, void* addr, uint32_t value) {
    int rc = load_memops();
    if (rc != B2R_OK) return rc;
    const CUresult r = g_wait32((CUstream)stream, (CUdeviceptr)(uintptr_t)addr, value, CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) { b2r_set_error("cuStreamWaitValue32 failed (%d)", (int)r); return B2R_ERR_CUDA; }
    return B2R_OK;
}
