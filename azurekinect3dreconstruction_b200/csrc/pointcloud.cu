// pointcloud.cu -- point-cloud conditioning on either side of ICP, sm_100a.
//
// Replaces the Open3D calls the reference makes on every frame between unprojection (a3) and ICP (a13):
//   pcd.remove_statistical_outlier(20, 2.0)                         CodeThatWorks/GoodColorReconstruction.py:144
//   pcd.voxel_down_sample(voxel)                                    GoodColorReconstruction.py:147
//   pcd.estimate_normals(KDTreeSearchParamHybrid(radius, max_nn))   GoodColorReconstruction.py:150,
//                                                                   CodeThatWorks/BetterVisualizer2CamWorksCOLOR.py:537
//   pcd.orient_normals_towards_camera_location(origin)              GoodColorReconstruction.py:151
// (upstream: geometry/PointCloud.cpp, EstimateNormals.cpp, KDTreeFlann.cpp; restated in oracle/oracle_pcd.cpp).
//
// Design (not a port of upstream's kd-tree): EXACT nearest neighbours through a uniform grid.
//   * b2r_knn_create: bounds -> cell size (given, or chosen so a non-empty cell holds ~k/2 points) -> 63-bit cell keys
//     -> stable radix sort -> points gathered in cell order (queries then walk memory almost linearly) -> an
//     open-addressing hash of the non-empty cells {key -> [start, end)} (one 16-byte slot per probe);
//   * a query visits the 27-cell block around its cell; a k-NN query grows the block ring by ring until the k-th
//     distance is strictly inside the distance to the block's boundary (so the result is exact, not approximate);
//     a hybrid query (radius, max_nn) visits the block that covers the radius once;
//   * the k best are kept sorted by (squared distance, index) -- ties go to the lower index, the same total order the
//     oracle uses -- so the covariance cumulants and the mean distances are summed in the SAME order as upstream sums
//     them (ascending distance) and come out bit-identical in f64;
//   * one thread per query: neighbours of grid-sorted queries share cells, so a warp's loads coalesce in L1/L2.
// All arithmetic that decides an integer result (neighbour sets, voxel indices, inlier masks) is f64, as upstream.
#include "common.cuh"
#include <cub/cub.cuh>
#include <math.h>
#include <string.h>

// Temporaries come from the stream-ordered pool (cudaMallocAsync): with the release threshold raised once, repeated calls
// reuse the same blocks instead of paying cudaMalloc / cudaFree (hundreds of microseconds each) per call.
static cudaError_t pcd_alloc(void** p, size_t bytes, cudaStream_t stream) {
    static bool pool_ready = false;
    if (!pool_ready) {
        int dev = 0;
        cudaMemPool_t pool;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            uint64_t keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        pool_ready = true;
    }
    return cudaMallocAsync(p, bytes ? bytes : 1, stream);
}
template <typename T>
static cudaError_t pcd_alloc(T** p, size_t bytes, cudaStream_t stream) { return pcd_alloc(reinterpret_cast<void**>(p), bytes, stream); }
static void pcd_free(void* p, cudaStream_t stream) { if (p) cudaFreeAsync(p, stream); }

namespace {

#define PCD_KEY_EMPTY 0xFFFFFFFFFFFFFFFFull
#define PCD_CELL_BITS 21
#define PCD_CELL_MAX ((1 << PCD_CELL_BITS) - 1)

struct __align__(16) CellSlot {
    unsigned long long key;
    int32_t start, end;
};

struct GridView {
    const double* pts;       // [n][3] in cell order
    const int32_t* order;    // cell-order position -> original index
    const CellSlot* slots;
    uint32_t hmask;
    int64_t n;
    double origin[3], cell;
    int dims[3];             // cells per axis (ring search stops once the block covers the grid)
};

__host__ __device__ __forceinline__ unsigned long long pcd_pack(int x, int y, int z) {
    return ((unsigned long long)(uint32_t)x << (2 * PCD_CELL_BITS)) | ((unsigned long long)(uint32_t)y << PCD_CELL_BITS) | (unsigned long long)(uint32_t)z;
}

__device__ __forceinline__ int pcd_cell_coord(double p, double origin, double cell, int dim) {
    int c = (int)floor(xddiv(xdsub(p, origin), cell));
    return min(max(c, 0), dim - 1);
}

__device__ __forceinline__ bool grid_find(const GridView& G, unsigned long long key, int& start, int& end) {
    uint32_t s = (uint32_t)b2r_mix64(key) & G.hmask;
    for (uint32_t probe = 0; probe <= G.hmask; probe++) {
        const ulonglong2 kv = *reinterpret_cast<const ulonglong2*>(&G.slots[s]);
        if (kv.x == key) { start = (int)(uint32_t)kv.y; end = (int)(uint32_t)(kv.y >> 32); return true; }
        if (kv.x == PCD_KEY_EMPTY) return false;
        s = (s + 1) & G.hmask;
    }
    return false;
}

// ---- bounds ----------------------------------------------------------------------------------------------
__global__ void bounds_kernel(const double* xyz, int64_t n, double* mm /*6: min xyz, max xyz*/) {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
#pragma unroll
        for (int a = 0; a < 3; a++) { const double v = xyz[3 * i + a]; lo[a] = fmin(lo[a], v); hi[a] = fmax(hi[a], v); }
    __shared__ double s[6][8];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        for (int o = 16; o > 0; o >>= 1) { lo[a] = fmin(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o)); hi[a] = fmax(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o)); }
        if ((threadIdx.x & 31) == 0) { s[a][threadIdx.x >> 5] = lo[a]; s[3 + a][threadIdx.x >> 5] = hi[a]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = s[threadIdx.x][0];
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) v = threadIdx.x < 3 ? fmin(v, s[threadIdx.x][w]) : fmax(v, s[threadIdx.x][w]);
        // min / max of doubles through the ordered-integer trick (atomics on the raw bits)
        long long bits = __double_as_longlong(v);
        bits = bits >= 0 ? bits : (bits ^ 0x7FFFFFFFFFFFFFFFll);
        if (threadIdx.x < 3) atomicMin(reinterpret_cast<long long*>(mm) + threadIdx.x, bits);
        else atomicMax(reinterpret_cast<long long*>(mm) + threadIdx.x, bits);
    }
}
__global__ void bounds_init_kernel(double* mm) {
    if (threadIdx.x < 3) reinterpret_cast<long long*>(mm)[threadIdx.x] = 0x7FFFFFFFFFFFFFFFll;
    else if (threadIdx.x < 6) reinterpret_cast<long long*>(mm)[threadIdx.x] = (long long)0x8000000000000000ull;
}
__global__ void bounds_decode_kernel(double* mm) {
    if (threadIdx.x < 6) {
        long long bits = reinterpret_cast<long long*>(mm)[threadIdx.x];
        bits = bits >= 0 ? bits : (bits ^ 0x7FFFFFFFFFFFFFFFll);
        mm[threadIdx.x] = __longlong_as_double(bits);
    }
}

static int device_bounds(const double* xyz, int64_t n, double* mm_dev, cudaStream_t stream) {
    bounds_init_kernel<<<1, 32, 0, stream>>>(mm_dev);
    const int64_t want = (n + 255) / 256;
    const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
    bounds_kernel<<<blocks, 256, 0, stream>>>(xyz, n, mm_dev);
    bounds_decode_kernel<<<1, 32, 0, stream>>>(mm_dev);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// ---- grid build ------------------------------------------------------------------------------------------
struct GridParams { double origin[3], cell; int dims[3]; };

__global__ void cell_keys_kernel(const double* xyz, int64_t n, GridParams g, unsigned long long* keys, int32_t* ids) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cx = pcd_cell_coord(xyz[3 * i], g.origin[0], g.cell, g.dims[0]);
    const int cy = pcd_cell_coord(xyz[3 * i + 1], g.origin[1], g.cell, g.dims[1]);
    const int cz = pcd_cell_coord(xyz[3 * i + 2], g.origin[2], g.cell, g.dims[2]);
    keys[i] = pcd_pack(cx, cy, cz);
    ids[i] = (int32_t)i;
}

__global__ void count_heads_kernel(const unsigned long long* keys, int64_t n, int32_t* count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool head = i < n && (i == 0 || keys[i] != keys[i - 1]);
    const unsigned b = __ballot_sync(0xffffffffu, head);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, __popc(b));
}

__global__ void gather_points_kernel(const double* xyz, const int32_t* order, int64_t n, double* out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t j = order[i];
    out[3 * i] = xyz[3 * j]; out[3 * i + 1] = xyz[3 * j + 1]; out[3 * i + 2] = xyz[3 * j + 2];
}

__global__ void clear_slots_kernel(CellSlot* slots, uint32_t nslots) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nslots) { slots[i].key = PCD_KEY_EMPTY; slots[i].start = 0; slots[i].end = 0; }
}

// heads claim a slot and write start; tails (last point of a cell) find the slot and write end
__global__ void insert_heads_kernel(const unsigned long long* keys, int64_t n, CellSlot* slots, uint32_t hmask) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || (i > 0 && keys[i] == keys[i - 1])) return;
    const unsigned long long key = keys[i];
    uint32_t s = (uint32_t)b2r_mix64(key) & hmask;
    for (;;) {
        const unsigned long long prev = atomicCAS(&slots[s].key, PCD_KEY_EMPTY, key);
        if (prev == PCD_KEY_EMPTY) { slots[s].start = (int32_t)i; return; }
        s = (s + 1) & hmask;
    }
}
__global__ void insert_tails_kernel(const unsigned long long* keys, int64_t n, CellSlot* slots, uint32_t hmask) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || (i + 1 < n && keys[i] == keys[i + 1])) return;
    const unsigned long long key = keys[i];
    uint32_t s = (uint32_t)b2r_mix64(key) & hmask;
    while (slots[s].key != key) s = (s + 1) & hmask;
    slots[s].end = (int32_t)(i + 1);
}

}  // namespace

struct b2r_knn {
    int64_t n;
    double cell, origin[3];
    int dims[3];
    double* pts;
    int32_t* order;
    CellSlot* slots;
    uint32_t nslots;
    int64_t n_cells;
};

static GridView view_of(const b2r_knn* h) {
    GridView G;
    G.pts = h->pts; G.order = h->order; G.slots = h->slots; G.hmask = h->nslots - 1; G.n = h->n; G.cell = h->cell;
    for (int a = 0; a < 3; a++) { G.origin[a] = h->origin[a]; G.dims[a] = h->dims[a]; }
    return G;
}

B2R_API void b2r_knn_destroy(b2r_knn* h) {
    if (!h) return;
    cudaStream_t stream = 0;
    pcd_free(h->pts, stream); pcd_free(h->order, stream); pcd_free(h->slots, stream);
    delete h;
}

// xyz: DEVICE f64 [n][3].  cell > 0: grid cell edge (use the search radius for hybrid queries).  cell <= 0: chosen so
// that a non-empty cell holds about k_hint / 2 points (k-NN queries).  Synchronises (bounds and cell count come back to
// the host).  Returns NULL on error (b2r_last_error()).
B2R_API b2r_knn* b2r_knn_create(const double* xyz, int64_t n, double cell, int k_hint, cudaStream_t stream) {
    if (!xyz || n < 1 || n > 0x7FFFFFF0ll) { b2r_set_error("b2r_knn_create: bad argument"); return nullptr; }
    b2r_knn* h = new b2r_knn();
    memset(h, 0, sizeof(*h));
    h->n = n;
    double* mm_dev = nullptr;
    unsigned long long *k_in = nullptr, *k_out = nullptr;
    int32_t *id_in = nullptr, *count_dev = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    auto fail = [&](const char* what, cudaError_t e) -> b2r_knn* {
        if (e != cudaSuccess) b2r_set_error("b2r_knn_create: %s: %s", what, cudaGetErrorString(e));
        pcd_free(mm_dev, stream); pcd_free(k_in, stream); pcd_free(k_out, stream); pcd_free(id_in, stream); pcd_free(count_dev, stream); pcd_free(tmp, stream);
        b2r_knn_destroy(h);
        return nullptr;
    };
    cudaError_t e;
    if ((e = pcd_alloc(&mm_dev, 6 * sizeof(double), stream)) != cudaSuccess) return fail("alloc", e);
    if (device_bounds(xyz, n, mm_dev, stream) != B2R_OK) return fail("bounds", cudaGetLastError());
    double mm[6];
    if ((e = cudaMemcpyAsync(mm, mm_dev, sizeof(mm), cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return fail("bounds copy", e);
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return fail("bounds sync", e);
    for (int a = 0; a < 3; a++)
        if (!isfinite(mm[a]) || !isfinite(mm[3 + a])) { b2r_set_error("b2r_knn_create: non-finite point coordinates"); return fail("bounds", cudaSuccess); }
    const double ext[3] = {mm[3] - mm[0], mm[4] - mm[1], mm[5] - mm[2]};
    const double max_ext = fmax(ext[0], fmax(ext[1], ext[2]));
    const bool auto_cell = !(cell > 0.0);
    const double target_occ = fmax(4.0, 0.5 * (double)(k_hint > 0 ? k_hint : 16));
    if (auto_cell) {
        // first guess: the points sample a surface of about the bounding box's cross-sections
        const double area = fmax(ext[0] * ext[1] + ext[1] * ext[2] + ext[0] * ext[2], 1e-12);
        cell = sqrt(target_occ * area / (double)n);
        if (!(cell > 0.0)) cell = 1.0;
    }
    if ((e = pcd_alloc(&k_in, n * 8, stream)) != cudaSuccess || (e = pcd_alloc(&k_out, n * 8, stream)) != cudaSuccess ||
        (e = pcd_alloc(&id_in, n * 4, stream)) != cudaSuccess || (e = pcd_alloc(&h->order, n * 4, stream)) != cudaSuccess ||
        (e = pcd_alloc(&count_dev, 4, stream)) != cudaSuccess)
        return fail("alloc", e);
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, id_in, h->order, (int)n, 0, 63, stream);
    if ((e = pcd_alloc(&tmp, tmp_bytes, stream)) != cudaSuccess) return fail("alloc", e);
    const unsigned grid = (unsigned)((n + 255) / 256);
    int32_t n_cells = 0;
    for (int attempt = 0; attempt < 4; attempt++) {
        cell = fmax(cell, max_ext / (double)(PCD_CELL_MAX - 1));        // coordinates must fit 21 bits
        GridParams g;
        g.cell = cell;
        for (int a = 0; a < 3; a++) {
            g.origin[a] = mm[a];
            g.dims[a] = (int)fmin((double)PCD_CELL_MAX, floor(ext[a] / cell) + 1.0);
            h->origin[a] = g.origin[a]; h->dims[a] = g.dims[a];
        }
        h->cell = cell;
        cell_keys_kernel<<<grid, 256, 0, stream>>>(xyz, n, g, k_in, id_in);
        if ((e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, id_in, h->order, (int)n, 0, 63, stream)) != cudaSuccess) return fail("sort", e);
        cudaMemsetAsync(count_dev, 0, 4, stream);
        count_heads_kernel<<<grid, 256, 0, stream>>>(k_out, n, count_dev);
        if ((e = cudaMemcpyAsync(&n_cells, count_dev, 4, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return fail("count copy", e);
        if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return fail("count sync", e);
        if (!auto_cell) break;
        const double occ = (double)n / (double)(n_cells > 0 ? n_cells : 1);
        if (occ > target_occ / 1.6 && occ < target_occ * 1.6) break;
        cell *= sqrt(target_occ / occ);                                  // occupancy of a surface grows with cell^2
    }
    h->n_cells = n_cells;
    uint32_t nslots = 64;
    while (nslots < 2u * (uint32_t)n_cells) nslots <<= 1;
    h->nslots = nslots;
    if ((e = pcd_alloc(&h->slots, (size_t)nslots * sizeof(CellSlot), stream)) != cudaSuccess || (e = pcd_alloc(&h->pts, n * 24, stream)) != cudaSuccess) return fail("alloc", e);
    clear_slots_kernel<<<(nslots + 255) / 256, 256, 0, stream>>>(h->slots, nslots);
    insert_heads_kernel<<<grid, 256, 0, stream>>>(k_out, n, h->slots, nslots - 1);
    insert_tails_kernel<<<grid, 256, 0, stream>>>(k_out, n, h->slots, nslots - 1);
    gather_points_kernel<<<grid, 256, 0, stream>>>(xyz, h->order, n, h->pts);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail("launch", e);
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return fail("sync", e);
    pcd_free(mm_dev, stream); pcd_free(k_in, stream); pcd_free(k_out, stream); pcd_free(id_in, stream); pcd_free(count_dev, stream); pcd_free(tmp, stream);
    return h;
}

B2R_API int b2r_knn_info(const b2r_knn* h, double* cell, int64_t* n_cells) {
    if (!h) return B2R_ERR_BAD_ARG;
    if (cell) *cell = h->cell;
    if (n_cells) *n_cells = h->n_cells;
    return B2R_OK;
}

// ---- exact neighbour query ---------------------------------------------------------------------------------
namespace {

// Inserts (d2, id) into the ascending list best[0..cnt) of capacity k under the total order (d2, id).
template <int KMAX>
__device__ __forceinline__ void knn_insert(double* bd, int32_t* bi, int& cnt, int k, double d2, int32_t id) {
    if (cnt == k) {
        const double wd = bd[k - 1];
        if (d2 > wd || (d2 == wd && id > bi[k - 1])) return;
    }
    int j = cnt < k ? cnt : k - 1;
    while (j > 0 && (bd[j - 1] > d2 || (bd[j - 1] == d2 && bi[j - 1] > id))) { bd[j] = bd[j - 1]; bi[j] = bi[j - 1]; j--; }
    bd[j] = d2; bi[j] = id;
    if (cnt < k) cnt++;
}

// Visits every point of cell (cx, cy, cz).
template <int KMAX>
__device__ __forceinline__ void knn_visit_cell(const GridView& G, int cx, int cy, int cz, double qx, double qy, double qz, double r2max,
                                               double* bd, int32_t* bi, int& cnt, int k) {
    int s, e;
    if (!grid_find(G, pcd_pack(cx, cy, cz), s, e)) return;
    for (int p = s; p < e; p++) {
        const double dx = xdsub(G.pts[3 * p], qx), dy = xdsub(G.pts[3 * p + 1], qy), dz = xdsub(G.pts[3 * p + 2], qz);
        const double d2 = xdadd(xdadd(xdmul(dx, dx), xdmul(dy, dy)), xdmul(dz, dz));
        if (d2 < r2max) knn_insert<KMAX>(bd, bi, cnt, k, d2, G.order[p]);
    }
}

// k nearest neighbours of q (radius <= 0) or the at most k nearest with d2 < radius^2 (hybrid).  Returns the count;
// bd / bi = ascending (d2, original index).
template <int KMAX>
__device__ int knn_query(const GridView& G, double qx, double qy, double qz, int k, double radius, double* bd, int32_t* bi) {
    int cnt = 0;
    const int cx = pcd_cell_coord(qx, G.origin[0], G.cell, G.dims[0]);
    const int cy = pcd_cell_coord(qy, G.origin[1], G.cell, G.dims[1]);
    const int cz = pcd_cell_coord(qz, G.origin[2], G.cell, G.dims[2]);
    const bool hybrid = radius > 0.0;
    const double r2max = hybrid ? xdmul(radius, radius) : INFINITY;
    const int maxdim = max(G.dims[0], max(G.dims[1], G.dims[2]));
    int R0 = 1;
    if (hybrid) R0 = max(1, (int)ceil(radius / G.cell));           // the block that covers the radius, visited once
    int Rprev = -1;
    for (int R = R0;; R++) {
        const int x0 = max(cx - R, 0), x1 = min(cx + R, G.dims[0] - 1);
        const int y0 = max(cy - R, 0), y1 = min(cy + R, G.dims[1] - 1);
        const int z0 = max(cz - R, 0), z1 = min(cz + R, G.dims[2] - 1);
        for (int x = x0; x <= x1; x++)
            for (int y = y0; y <= y1; y++)
                for (int z = z0; z <= z1; z++) {
                    // only the shell that the previous ring did not cover
                    if (Rprev >= 0 && abs(x - cx) <= Rprev && abs(y - cy) <= Rprev && abs(z - cz) <= Rprev) continue;
                    knn_visit_cell<KMAX>(G, x, y, z, qx, qy, qz, r2max, bd, bi, cnt, k);
                }
        if (hybrid || R >= maxdim) break;
        if (cnt == k) {
            // distance from q to the nearest face of the visited block that still has cells behind it
            double bound = INFINITY;
            const double q[3] = {qx, qy, qz};
            const int c[3] = {cx, cy, cz};
#pragma unroll
            for (int a = 0; a < 3; a++) {
                if (c[a] - R > 0) bound = fmin(bound, xdsub(q[a], xdadd(G.origin[a], xdmul((double)(c[a] - R), G.cell))));
                if (c[a] + R < G.dims[a] - 1) bound = fmin(bound, xdsub(xdadd(G.origin[a], xdmul((double)(c[a] + R + 1), G.cell)), q[a]));
            }
            // cell assignment (floor of a quotient) and face positions (products) round independently: pull the bound in by
            // far more than those few ulps, so a point just behind a face can never be missed (costs at most one more ring).
            // A clamped query (outside the grid) can make bound negative: keep growing until it is safely positive.
            if (bound != INFINITY) bound = xdsub(bound, xdmul(1e-9, G.cell));
            if (bound > 0.0 && bd[k - 1] < xdmul(bound, bound)) break;
            if (bound == INFINITY) break;                             // the block already covers the whole grid
        }
        Rprev = R;
    }
    return cnt;
}

template <int KMAX>
__global__ void __launch_bounds__(128) knn_search_kernel(const GridView G, const double* query, int64_t nq, int self_query, int k, double radius,
                                                         int32_t* idx, double* d2, int32_t* counts) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq) return;
    // self queries run in cell order (neighbouring threads visit the same cells); results go to the original slot
    const int64_t i = self_query ? (int64_t)G.order[t] : t;
    const double* q = self_query ? G.pts + 3 * t : query + 3 * t;
    double bd[KMAX]; int32_t bi[KMAX];
    const int cnt = knn_query<KMAX>(G, q[0], q[1], q[2], k, radius, bd, bi);
    counts[i] = cnt;
    for (int j = 0; j < k; j++) {
        idx[i * k + j] = j < cnt ? bi[j] : -1;
        if (d2) d2[i * k + j] = j < cnt ? bd[j] : -1.0;
    }
}

}  // namespace

// query: DEVICE f64 [nq][3], or NULL for a self query (every indexed point, itself included as its first neighbour).
// radius <= 0: k nearest; radius > 0: the at most k nearest with squared distance < radius^2 (KDTreeFlann::SearchHybrid).
// idx [nq][k] original indices, ascending (distance, index), -1 padded; d2 (optional) [nq][k]; counts [nq].
B2R_API int b2r_knn_search(const b2r_knn* h, const double* query, int64_t nq, int k, double radius, int32_t* idx, double* d2,
                           int32_t* counts, cudaStream_t stream) {
    if (!h || !idx || !counts || k < 1 || k > 128) { b2r_set_error("b2r_knn_search: bad argument (1 <= k <= 128)"); return B2R_ERR_BAD_ARG; }
    const int self_query = query == nullptr;
    if (self_query) nq = h->n;
    if (nq <= 0) return B2R_OK;
    const GridView G = view_of(h);
    const unsigned grid = (unsigned)((nq + 127) / 128);
    if (k <= 32) knn_search_kernel<32><<<grid, 128, 0, stream>>>(G, query, nq, self_query, k, radius, idx, d2, counts);
    else knn_search_kernel<128><<<grid, 128, 0, stream>>>(G, query, nq, self_query, k, radius, idx, d2, counts);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// ---- estimate_normals ----------------------------------------------------------------------------------------
namespace {

struct D3 { double x, y, z; };
__device__ __forceinline__ D3 d3cross(const D3& a, const D3& b) {
    return {xdsub(xdmul(a.y, b.z), xdmul(a.z, b.y)), xdsub(xdmul(a.z, b.x), xdmul(a.x, b.z)), xdsub(xdmul(a.x, b.y), xdmul(a.y, b.x))};
}
__device__ __forceinline__ double d3dot(const D3& a, const D3& b) { return xdadd(xdadd(xdmul(a.x, b.x), xdmul(a.y, b.y)), xdmul(a.z, b.z)); }

// A = {a00, a01, a02, a11, a12, a22}; same operation order as oracle_pcd.cpp (Eberly's robust symmetric 3x3 solver)
__device__ D3 pcd_eigenvector0(const double* A, double e) {
    const D3 r0{xdsub(A[0], e), A[1], A[2]}, r1{A[1], xdsub(A[3], e), A[4]}, r2{A[2], A[4], xdsub(A[5], e)};
    const D3 c01 = d3cross(r0, r1), c02 = d3cross(r0, r2), c12 = d3cross(r1, r2);
    const double d0 = d3dot(c01, c01), d1 = d3dot(c02, c02), d2 = d3dot(c12, c12);
    double dmax = d0; int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    if (imax == 0) { const double s = sqrt(d0); return {xddiv(c01.x, s), xddiv(c01.y, s), xddiv(c01.z, s)}; }
    if (imax == 1) { const double s = sqrt(d1); return {xddiv(c02.x, s), xddiv(c02.y, s), xddiv(c02.z, s)}; }
    const double s = sqrt(d2); return {xddiv(c12.x, s), xddiv(c12.y, s), xddiv(c12.z, s)};
}

__device__ D3 pcd_eigenvector1(const double* A, const D3& e0, double eval1) {
    D3 U;
    if (fabs(e0.x) > fabs(e0.y)) { const double inv = xddiv(1.0, sqrt(xdadd(xdmul(e0.x, e0.x), xdmul(e0.z, e0.z)))); U = {xdmul(-e0.z, inv), 0.0, xdmul(e0.x, inv)}; }
    else { const double inv = xddiv(1.0, sqrt(xdadd(xdmul(e0.y, e0.y), xdmul(e0.z, e0.z)))); U = {0.0, xdmul(e0.z, inv), xdmul(-e0.y, inv)}; }
    const D3 V = d3cross(e0, U);
    const D3 AU{xdadd(xdadd(xdmul(A[0], U.x), xdmul(A[1], U.y)), xdmul(A[2], U.z)), xdadd(xdadd(xdmul(A[1], U.x), xdmul(A[3], U.y)), xdmul(A[4], U.z)),
                xdadd(xdadd(xdmul(A[2], U.x), xdmul(A[4], U.y)), xdmul(A[5], U.z))};
    const D3 AV{xdadd(xdadd(xdmul(A[0], V.x), xdmul(A[1], V.y)), xdmul(A[2], V.z)), xdadd(xdadd(xdmul(A[1], V.x), xdmul(A[3], V.y)), xdmul(A[4], V.z)),
                xdadd(xdadd(xdmul(A[2], V.x), xdmul(A[4], V.y)), xdmul(A[5], V.z))};
    double m00 = xdsub(d3dot(U, AU), eval1), m01 = d3dot(U, AV), m11 = xdsub(d3dot(V, AV), eval1);
    const double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
    if (a00 >= a11) {
        if (fmax(a00, a01) > 0.0) {
            if (a00 >= a01) { m01 = xddiv(m01, m00); m00 = xddiv(1.0, sqrt(xdadd(1.0, xdmul(m01, m01)))); m01 = xdmul(m01, m00); }
            else { m00 = xddiv(m00, m01); m01 = xddiv(1.0, sqrt(xdadd(1.0, xdmul(m00, m00)))); m00 = xdmul(m00, m01); }
            return {xdsub(xdmul(m01, U.x), xdmul(m00, V.x)), xdsub(xdmul(m01, U.y), xdmul(m00, V.y)), xdsub(xdmul(m01, U.z), xdmul(m00, V.z))};
        }
        return U;
    }
    if (fmax(a11, a01) > 0.0) {
        if (a11 >= a01) { m01 = xddiv(m01, m11); m11 = xddiv(1.0, sqrt(xdadd(1.0, xdmul(m01, m01)))); m01 = xdmul(m01, m11); }
        else { m11 = xddiv(m11, m01); m01 = xddiv(1.0, sqrt(xdadd(1.0, xdmul(m11, m11)))); m11 = xdmul(m11, m01); }
        return {xdsub(xdmul(m11, U.x), xdmul(m01, V.x)), xdsub(xdmul(m11, U.y), xdmul(m01, V.y)), xdsub(xdmul(m11, U.z), xdmul(m01, V.z))};
    }
    return U;
}

__device__ D3 pcd_fast_eigen3x3(const double* C) {
    double A[6];
    double maxc = C[0];
#pragma unroll
    for (int i = 1; i < 6; i++) maxc = fmax(maxc, C[i]);
    if (maxc == 0.0) return {0.0, 0.0, 0.0};
#pragma unroll
    for (int i = 0; i < 6; i++) A[i] = xddiv(C[i], maxc);
    const double norm = xdadd(xdadd(xdmul(A[1], A[1]), xdmul(A[2], A[2])), xdmul(A[4], A[4]));
    if (norm > 0.0) {
        const double q = xddiv(xdadd(xdadd(A[0], A[3]), A[5]), 3.0);
        const double b00 = xdsub(A[0], q), b11 = xdsub(A[3], q), b22 = xdsub(A[5], q);
        const double p = sqrt(xddiv(xdadd(xdadd(xdadd(xdmul(b00, b00), xdmul(b11, b11)), xdmul(b22, b22)), xdmul(norm, 2.0)), 6.0));
        const double c00 = xdsub(xdmul(b11, b22), xdmul(A[4], A[4]));
        const double c01 = xdsub(xdmul(A[1], b22), xdmul(A[4], A[2]));
        const double c02 = xdsub(xdmul(A[1], A[4]), xdmul(b11, A[2]));
        const double det = xddiv(xdadd(xdsub(xdmul(b00, c00), xdmul(A[1], c01)), xdmul(A[2], c02)), xdmul(xdmul(p, p), p));
        const double half_det = fmin(fmax(xdmul(det, 0.5), -1.0), 1.0);
        const double angle = xddiv(acos(half_det), 3.0);
        const double two_thirds_pi = 2.09439510239319549;
        const double beta2 = xdmul(cos(angle), 2.0);
        const double beta0 = xdmul(cos(xdadd(angle, two_thirds_pi)), 2.0);
        const double beta1 = -xdadd(beta0, beta2);
        const double e0 = xdadd(q, xdmul(p, beta0)), e1 = xdadd(q, xdmul(p, beta1)), e2 = xdadd(q, xdmul(p, beta2));
        if (half_det >= 0.0) {
            const D3 v2 = pcd_eigenvector0(A, e2);
            if (e2 < e0 && e2 < e1) return v2;
            const D3 v1 = pcd_eigenvector1(A, v2, e1);
            if (e1 < e0 && e1 < e2) return v1;
            return d3cross(v1, v2);
        }
        const D3 v0 = pcd_eigenvector0(A, e0);
        if (e0 < e1 && e0 < e2) return v0;
        const D3 v1 = pcd_eigenvector1(A, v0, e1);
        if (e1 < e0 && e1 < e2) return v1;
        return d3cross(v0, v1);
    }
    if (C[0] < C[3] && C[0] < C[5]) return {1.0, 0.0, 0.0};
    if (C[3] < C[0] && C[3] < C[5]) return {0.0, 1.0, 0.0};
    return {0.0, 0.0, 1.0};
}

// search + covariance (cumulants summed in ascending-distance order, as upstream) + eigenvector + orientation
template <int KMAX>
__global__ void __launch_bounds__(128) estimate_normals_kernel(const GridView G, const double* xyz, int k, double radius, double* normals,
                                                               int has_normals, double* cov_out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= G.n) return;
    const int64_t i = G.order[t];
    double bd[KMAX]; int32_t bi[KMAX];
    const int cnt = knn_query<KMAX>(G, G.pts[3 * t], G.pts[3 * t + 1], G.pts[3 * t + 2], k, radius, bd, bi);
    double C[6] = {1.0, 0.0, 0.0, 1.0, 0.0, 1.0};
    if (cnt >= 3) {
        double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int j = 0; j < cnt; j++) {
            const double x = xyz[3 * (int64_t)bi[j]], y = xyz[3 * (int64_t)bi[j] + 1], z = xyz[3 * (int64_t)bi[j] + 2];
            c[0] = xdadd(c[0], x); c[1] = xdadd(c[1], y); c[2] = xdadd(c[2], z);
            c[3] = xdadd(c[3], xdmul(x, x)); c[4] = xdadd(c[4], xdmul(x, y)); c[5] = xdadd(c[5], xdmul(x, z));
            c[6] = xdadd(c[6], xdmul(y, y)); c[7] = xdadd(c[7], xdmul(y, z)); c[8] = xdadd(c[8], xdmul(z, z));
        }
        const double m = (double)cnt;
#pragma unroll
        for (int j = 0; j < 9; j++) c[j] = xddiv(c[j], m);
        C[0] = xdsub(c[3], xdmul(c[0], c[0])); C[3] = xdsub(c[6], xdmul(c[1], c[1])); C[5] = xdsub(c[8], xdmul(c[2], c[2]));
        C[1] = xdsub(c[4], xdmul(c[0], c[1])); C[2] = xdsub(c[5], xdmul(c[0], c[2])); C[4] = xdsub(c[7], xdmul(c[1], c[2]));
    }
    if (cov_out)
#pragma unroll
        for (int j = 0; j < 6; j++) cov_out[6 * i + j] = C[j];
    D3 nrm = pcd_fast_eigen3x3(C);
    const double nn = sqrt(d3dot(nrm, nrm));
    const D3 old{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]};
    if (nn == 0.0) nrm = has_normals ? old : D3{0.0, 0.0, 1.0};
    if (has_normals && d3dot(nrm, old) < 0.0) { nrm.x = -nrm.x; nrm.y = -nrm.y; nrm.z = -nrm.z; }
    normals[3 * i] = nrm.x; normals[3 * i + 1] = nrm.y; normals[3 * i + 2] = nrm.z;
}

__global__ void orient_normals_kernel(const double* xyz, int64_t n, double cx, double cy, double cz, double* normals) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const D3 ref{xdsub(cx, xyz[3 * i]), xdsub(cy, xyz[3 * i + 1]), xdsub(cz, xyz[3 * i + 2])};
    D3 nm{normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]};
    const double nn = sqrt(d3dot(nm, nm));
    if (nn == 0.0) {
        const double rn = sqrt(d3dot(ref, ref));
        if (rn == 0.0) nm = {0.0, 0.0, 1.0};
        else nm = {xddiv(ref.x, rn), xddiv(ref.y, rn), xddiv(ref.z, rn)};
    } else if (d3dot(nm, ref) < 0.0) {
        nm = {-nm.x, -nm.y, -nm.z};
    }
    normals[3 * i] = nm.x; normals[3 * i + 1] = nm.y; normals[3 * i + 2] = nm.z;
}

}  // namespace

// PointCloud::EstimateNormals(search_param, fast_normal_computation = true).  h indexes xyz (the same n points, original
// order).  radius <= 0: KDTreeSearchParamKNN(k); radius > 0: KDTreeSearchParamHybrid(radius, k).  normals: DEVICE f64
// [n][3] in/out; has_normals = 1 orients each result along the existing normal.  cov_out: optional [n][6].
B2R_API int b2r_pcd_estimate_normals(const b2r_knn* h, const double* xyz, int k, double radius, double* normals, int has_normals,
                                     double* cov_out, cudaStream_t stream) {
    if (!h || !xyz || !normals || k < 1 || k > 128) { b2r_set_error("b2r_pcd_estimate_normals: bad argument (1 <= k <= 128)"); return B2R_ERR_BAD_ARG; }
    const GridView G = view_of(h);
    const unsigned grid = (unsigned)((h->n + 127) / 128);
    if (k <= 32) estimate_normals_kernel<32><<<grid, 128, 0, stream>>>(G, xyz, k, radius, normals, has_normals, cov_out);
    else estimate_normals_kernel<128><<<grid, 128, 0, stream>>>(G, xyz, k, radius, normals, has_normals, cov_out);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// PointCloud::OrientNormalsTowardsCameraLocation(camera_location)
B2R_API int b2r_pcd_orient_normals(const double* xyz, int64_t n, const double* cam3_host, double* normals, cudaStream_t stream) {
    if (!xyz || !normals || !cam3_host || n < 0) return B2R_ERR_BAD_ARG;
    if (n) orient_normals_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(xyz, n, cam3_host[0], cam3_host[1], cam3_host[2], normals);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// ---- remove_statistical_outlier --------------------------------------------------------------------------------
namespace {

template <int KMAX>
__global__ void __launch_bounds__(128) mean_distance_kernel(const GridView G, int k, double* avg) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= G.n) return;
    double bd[KMAX]; int32_t bi[KMAX];
    const int cnt = knn_query<KMAX>(G, G.pts[3 * t], G.pts[3 * t + 1], G.pts[3 * t + 2], k, -1.0, bd, bi);
    double s = 0.0;
    for (int j = 0; j < cnt; j++) s = xdadd(s, sqrt(bd[j]));       // ascending distance, the order upstream sums in
    avg[G.order[t]] = cnt > 0 ? xddiv(s, (double)cnt) : -1.0;
}

struct PositiveOrZero {
    __host__ __device__ double operator()(const double& v) const { return v > 0.0 ? v : 0.0; }
};
struct SquaredDeviation {
    const double* mean;
    __host__ __device__ double operator()(const double& v) const { const double d = v - *mean; return v > 0.0 ? d * d : 0.0; }
};

// st[0] = sum -> mean;  then st[1] = sum of squared deviations -> st[2] = threshold
__global__ void stat_mean_kernel(double* st, int64_t n) { st[0] = xddiv(st[0], (double)n); }
__global__ void stat_threshold_kernel(double* st, int64_t n, double std_ratio) {
    const double sd = sqrt(xddiv(st[1], (double)(n - 1)));
    st[2] = xdadd(st[0], xdmul(std_ratio, sd));
}
__global__ void stat_keep_kernel(const double* avg, int64_t n, const double* st, uint8_t* keep) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keep[i] = (avg[i] > 0.0 && avg[i] < st[2]) ? 1 : 0;
}

template <typename T>
__global__ void gather_rows_kernel(const T* src, const int32_t* idx, int64_t m, int width, T* dst) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m * width) return;
    const int64_t r = t / width;
    const int c = (int)(t - r * width);
    dst[t] = src[(int64_t)idx[r] * width + c];
}

}  // namespace

// PointCloud::RemoveStatisticalOutliers(nb_neighbors, std_ratio) over the points indexed by h.
// keep: DEVICE u8 [n]; inlier_idx: DEVICE int32 [n] (ascending original indices of the inliers, first *n_inliers_host valid);
// avg_out: optional DEVICE f64 [n] mean neighbour distance (the point is its own first neighbour).  Synchronises.
// The cloud mean / deviation are device-wide reductions (cub, deterministic for a given n), upstream's are sequential
// sums: the threshold can differ in its last bits, which only matters for a point within ~1e-15 of it.
B2R_API int b2r_pcd_statistical_outlier(const b2r_knn* h, int nb_neighbors, double std_ratio, uint8_t* keep, int32_t* inlier_idx,
                                        int64_t* n_inliers_host, double* avg_out, cudaStream_t stream) {
    if (!h || !keep || !inlier_idx || !n_inliers_host) return B2R_ERR_BAD_ARG;
    if (nb_neighbors < 1 || !(std_ratio > 0.0)) {      // upstream: LogError("Illegal input parameters, ...")
        b2r_set_error("[RemoveStatisticalOutliers] Illegal input parameters, number of neighbors and standard deviation ratio must be positive.");
        return B2R_ERR_BAD_ARG;
    }
    if (nb_neighbors > 128) { b2r_set_error("b2r_pcd_statistical_outlier: nb_neighbors <= 128"); return B2R_ERR_BAD_ARG; }
    const int64_t n = h->n;
    const GridView G = view_of(h);
    double *avg = avg_out, *st = nullptr;
    void* tmp = nullptr;
    int32_t* nsel = nullptr;
    B2R_CUDA(pcd_alloc(&st, 4 * sizeof(double), stream));
    B2R_CUDA(pcd_alloc(&nsel, 4, stream));
    if (!avg) B2R_CUDA(pcd_alloc(&avg, n * sizeof(double), stream));
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (nb_neighbors <= 32) mean_distance_kernel<32><<<grid, 128, 0, stream>>>(G, nb_neighbors, avg);
    else mean_distance_kernel<128><<<grid, 128, 0, stream>>>(G, nb_neighbors, avg);
    cub::TransformInputIterator<double, PositiveOrZero, const double*> it_pos(avg, PositiveOrZero());
    SquaredDeviation sq_op; sq_op.mean = st;
    cub::TransformInputIterator<double, SquaredDeviation, const double*> it_sq(avg, sq_op);
    cub::CountingInputIterator<int32_t> counting(0);
    size_t b0 = 0, b1 = 0, b2 = 0;
    cub::DeviceReduce::Sum(nullptr, b0, it_pos, st, (int)n, stream);
    cub::DeviceReduce::Sum(nullptr, b1, it_sq, st + 1, (int)n, stream);
    cub::DeviceSelect::Flagged(nullptr, b2, counting, keep, inlier_idx, nsel, (int)n, stream);
    size_t tmp_bytes = b0 > b1 ? b0 : b1;
    if (b2 > tmp_bytes) tmp_bytes = b2;
    B2R_CUDA(pcd_alloc(&tmp, tmp_bytes, stream));
    B2R_CUDA(cub::DeviceReduce::Sum(tmp, tmp_bytes, it_pos, st, (int)n, stream));
    stat_mean_kernel<<<1, 1, 0, stream>>>(st, n);
    B2R_CUDA(cub::DeviceReduce::Sum(tmp, tmp_bytes, it_sq, st + 1, (int)n, stream));
    stat_threshold_kernel<<<1, 1, 0, stream>>>(st, n, std_ratio);
    stat_keep_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(avg, n, st, keep);
    B2R_CUDA(cub::DeviceSelect::Flagged(tmp, tmp_bytes, counting, keep, inlier_idx, nsel, (int)n, stream));
    int32_t m = 0;
    B2R_CUDA(cudaMemcpyAsync(&m, nsel, 4, cudaMemcpyDeviceToHost, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    *n_inliers_host = m;
    pcd_free(st, stream); pcd_free(nsel, stream); pcd_free(tmp, stream);
    if (!avg_out) pcd_free(avg, stream);
    return B2R_OK;
}

// select_by_index on device arrays: dst[r][:] = src[idx[r]][:]   (f64 rows of `width` values, or int32 when is_i32)
B2R_API int b2r_pcd_gather_rows(const void* src, const int32_t* idx, int64_t m, int width, int is_i32, void* dst, cudaStream_t stream) {
    if (!src || !idx || !dst || m < 0 || width < 1) return B2R_ERR_BAD_ARG;
    if (m == 0) return B2R_OK;
    const unsigned grid = (unsigned)((m * width + 255) / 256);
    if (is_i32) gather_rows_kernel<int32_t><<<grid, 256, 0, stream>>>((const int32_t*)src, idx, m, width, (int32_t*)dst);
    else gather_rows_kernel<double><<<grid, 256, 0, stream>>>((const double*)src, idx, m, width, (double*)dst);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

// ---- RegistrationResult.correspondence_set: the (source, target) pairs of the matched source points, ascending by source ----
namespace {
__global__ void corr_flags_kernel(const int32_t* corr, int64_t n, int32_t* flags) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = corr[i] >= 0 ? 1 : 0;
    if (i == n) flags[i] = 0;
}
__global__ void corr_pairs_kernel(const int32_t* corr, const int32_t* scan, int64_t n, int32_t* pairs) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || corr[i] < 0) return;
    pairs[2 * (int64_t)scan[i]] = (int32_t)i;
    pairs[2 * (int64_t)scan[i] + 1] = corr[i];
}
}  // namespace
// corr: DEVICE int32[n] (target index per source point, -1 = none, as b2r_icp_* write it); pairs_out: DEVICE int32[n][2] (room for
// every point); *count_host = pairs written.  Synchronises `stream` to return the count.
B2R_API int b2r_correspondence_set(const int32_t* corr, int64_t n, int32_t* pairs_out, int64_t* count_host, cudaStream_t stream) {
    if (!corr || !pairs_out || !count_host || n < 0 || n > 0x7ffffff0) return B2R_ERR_BAD_ARG;
    *count_host = 0;
    if (n == 0) return B2R_OK;
    size_t tmp_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (int32_t*)nullptr, (int32_t*)nullptr, (int)(n + 1), stream);
    int32_t *flags = nullptr, *scan = nullptr;
    void* tmp = nullptr;
    cudaError_t e = pcd_alloc(&flags, (size_t)(n + 1) * 4, stream);
    if (e == cudaSuccess) e = pcd_alloc(&scan, (size_t)(n + 1) * 4, stream);
    if (e == cudaSuccess) e = pcd_alloc(&tmp, tmp_bytes, stream);
    int32_t total = 0;
    if (e == cudaSuccess) {
        const unsigned grid = (unsigned)((n + 1 + 255) / 256);
        corr_flags_kernel<<<grid, 256, 0, stream>>>(corr, n, flags);
        e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flags, scan, (int)(n + 1), stream);
        if (e == cudaSuccess) corr_pairs_kernel<<<grid, 256, 0, stream>>>(corr, scan, n, pairs_out);
        if (e == cudaSuccess) e = cudaMemcpyAsync(&total, scan + n, 4, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    }
    pcd_free(flags, stream); pcd_free(scan, stream); pcd_free(tmp, stream);
    if (e != cudaSuccess) { b2r_set_error("b2r_correspondence_set: %s", cudaGetErrorString(e)); return B2R_ERR_CUDA; }
    *count_host = total;
    return B2R_OK;
}

// ---- voxel_down_sample -------------------------------------------------------------------------------------------
namespace {

__global__ void voxel_keys_kernel(const double* xyz, int64_t n, const double* mm, double voxel, unsigned long long* keys, int32_t* ids, int32_t* err) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double half = xdmul(voxel, 0.5);
    int c[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const double lo = xdsub(mm[a], half);                        // voxel_min_bound = min_bound - voxel_size * 0.5
        const double f = floor(xddiv(xdsub(xyz[3 * i + a], lo), voxel));
        if (!(f >= 0.0 && f < (double)PCD_CELL_MAX)) { *err = 1; c[a] = 0; } else c[a] = (int)f;
    }
    keys[i] = pcd_pack(c[0], c[1], c[2]);
    ids[i] = (int32_t)i;
}

struct IsHead {
    const unsigned long long* keys;
    __host__ __device__ bool operator()(const int32_t& i) const { return i == 0 || keys[i] != keys[i - 1]; }
};

// one thread per voxel: sequential sums in input order (the stable sort keeps it), exactly as upstream accumulates
__global__ void voxel_average_kernel(const double* xyz, const double* col, const double* nrm, const int32_t* order, const unsigned long long* keys,
                                     const int32_t* seg_start, int32_t m, int64_t n, double* out_xyz, double* out_col, double* out_nrm, int32_t* out_key) {
    const int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= m) return;
    const int64_t b = seg_start[s], e = s + 1 < m ? (int64_t)seg_start[s + 1] : n;
    double p[3] = {0, 0, 0}, c[3] = {0, 0, 0}, q[3] = {0, 0, 0};
    for (int64_t t = b; t < e; t++) {
        const int64_t i = order[t];
#pragma unroll
        for (int a = 0; a < 3; a++) p[a] = xdadd(p[a], xyz[3 * i + a]);
        if (nrm) {
            const double nx = nrm[3 * i], ny = nrm[3 * i + 1], nz = nrm[3 * i + 2];
            if (!isnan(nx) && !isnan(ny) && !isnan(nz)) { q[0] = xdadd(q[0], nx); q[1] = xdadd(q[1], ny); q[2] = xdadd(q[2], nz); }
        }
        if (col)
#pragma unroll
            for (int a = 0; a < 3; a++) c[a] = xdadd(c[a], col[3 * i + a]);
    }
    const double d = (double)(e - b);
#pragma unroll
    for (int a = 0; a < 3; a++) {
        out_xyz[3 * (int64_t)s + a] = xddiv(p[a], d);
        if (col) out_col[3 * (int64_t)s + a] = xddiv(c[a], d);
        if (nrm) out_nrm[3 * (int64_t)s + a] = xddiv(q[a], d);
    }
    if (out_key) {
        const unsigned long long k = keys[b];
        out_key[3 * (int64_t)s] = (int32_t)(k >> (2 * PCD_CELL_BITS)); out_key[3 * (int64_t)s + 1] = (int32_t)((k >> PCD_CELL_BITS) & PCD_CELL_MAX);
        out_key[3 * (int64_t)s + 2] = (int32_t)(k & PCD_CELL_MAX);
    }
}

}  // namespace

// PointCloud::VoxelDownSample(voxel).  xyz (and optional col / nrm): DEVICE f64 [n][3].  Outputs: DEVICE, capacity rows
// (n always suffices), in ascending voxel-index order (upstream's order is that of an unordered_map); out_key optional
// int32 [capacity][3] voxel indices.  *m_host = number of voxels; B2R_ERR_CAPACITY (with *m_host set) when it exceeds
// capacity.  Synchronises.
B2R_API int b2r_pcd_voxel_down_sample(const double* xyz, const double* col, const double* nrm, int64_t n, double voxel, double* out_xyz,
                                      double* out_col, double* out_nrm, int32_t* out_key, int64_t capacity, int64_t* m_host, cudaStream_t stream) {
    if (!xyz || !out_xyz || !m_host || n < 0 || n > 0x7FFFFFF0ll || (col && !out_col) || (nrm && !out_nrm)) return B2R_ERR_BAD_ARG;
    if (!(voxel > 0.0)) { b2r_set_error("[VoxelDownSample] voxel_size <= 0."); return B2R_ERR_BAD_ARG; }
    *m_host = 0;
    if (n == 0) return B2R_OK;
    double* mm = nullptr;
    unsigned long long *k_in = nullptr, *k_out = nullptr;
    int32_t *id_in = nullptr, *order = nullptr, *seg = nullptr, *cnt = nullptr;
    void* tmp = nullptr;
    int rc = B2R_OK;
    auto cleanup = [&]() { pcd_free(mm, stream); pcd_free(k_in, stream); pcd_free(k_out, stream); pcd_free(id_in, stream); pcd_free(order, stream); pcd_free(seg, stream); pcd_free(cnt, stream); pcd_free(tmp, stream); };
#define PCD_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { b2r_set_error("b2r_pcd_voxel_down_sample: %s", cudaGetErrorString(e_)); cleanup(); return B2R_ERR_CUDA; } } while (0)
    PCD_TRY(pcd_alloc(&mm, 6 * sizeof(double), stream));
    PCD_TRY(pcd_alloc(&k_in, n * 8, stream)); PCD_TRY(pcd_alloc(&k_out, n * 8, stream));
    PCD_TRY(pcd_alloc(&id_in, n * 4, stream)); PCD_TRY(pcd_alloc(&order, n * 4, stream)); PCD_TRY(pcd_alloc(&seg, n * 4, stream));
    PCD_TRY(pcd_alloc(&cnt, 8, stream));
    PCD_TRY(cudaMemsetAsync(cnt, 0, 8, stream));
    if ((rc = device_bounds(xyz, n, mm, stream)) != B2R_OK) { cleanup(); return rc; }
    const unsigned grid = (unsigned)((n + 255) / 256);
    voxel_keys_kernel<<<grid, 256, 0, stream>>>(xyz, n, mm, voxel, k_in, id_in, cnt + 1);
    size_t b0 = 0, b1 = 0;
    cub::CountingInputIterator<int32_t> counting(0);
    IsHead is_head; is_head.keys = k_out;
    cub::DeviceRadixSort::SortPairs(nullptr, b0, k_in, k_out, id_in, order, (int)n, 0, 63, stream);
    cub::DeviceSelect::If(nullptr, b1, counting, seg, cnt, (int)n, is_head, stream);
    size_t tmp_bytes = b0 > b1 ? b0 : b1;
    PCD_TRY(pcd_alloc(&tmp, tmp_bytes, stream));
    PCD_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, id_in, order, (int)n, 0, 63, stream));
    PCD_TRY(cub::DeviceSelect::If(tmp, tmp_bytes, counting, seg, cnt, (int)n, is_head, stream));
    int32_t hc[2] = {0, 0};
    PCD_TRY(cudaMemcpyAsync(hc, cnt, 8, cudaMemcpyDeviceToHost, stream));
    PCD_TRY(cudaStreamSynchronize(stream));
    if (hc[1]) { b2r_set_error("[VoxelDownSample] voxel_size is too small."); cleanup(); return B2R_ERR_RANGE; }
    *m_host = hc[0];
    if (hc[0] > capacity) { b2r_set_error("b2r_pcd_voxel_down_sample: %d voxels exceed the output capacity", hc[0]); cleanup(); return B2R_ERR_CAPACITY; }
    voxel_average_kernel<<<(unsigned)((hc[0] + 127) / 128), 128, 0, stream>>>(xyz, col, nrm, order, k_out, seg, hc[0], n, out_xyz, out_col, out_nrm, out_key);
    PCD_TRY(cudaGetLastError());
    PCD_TRY(cudaStreamSynchronize(stream));
#undef PCD_TRY
    cleanup();
    return B2R_OK;
}

// axis-aligned bounds of a device point array: out6_host = {min xyz, max xyz}.  Synchronises.
B2R_API int b2r_pcd_bounds(const double* xyz, int64_t n, double* out6_host, cudaStream_t stream) {
    if (!xyz || !out6_host || n < 1) return B2R_ERR_BAD_ARG;
    double* mm = nullptr;
    B2R_CUDA(pcd_alloc(&mm, 6 * sizeof(double), stream));
    int rc = device_bounds(xyz, n, mm, stream);
    if (rc == B2R_OK) {
        B2R_CUDA(cudaMemcpyAsync(out6_host, mm, 6 * sizeof(double), cudaMemcpyDeviceToHost, stream));
        B2R_CUDA(cudaStreamSynchronize(stream));
    }
    pcd_free(mm, stream);
    return rc;
}

// ---- f4: GetRegistrationResultAndCorrespondences for H candidate transforms at once -----------------------------------------
// (the "expensive validation" of registration_ransac_based_on_*: every source point transformed, nearest target within
// max_dist).  One CTA per transform; counts and squared-distance sums reduced in a fixed tree (deterministic).
namespace {
__global__ void __launch_bounds__(256) eval_transforms_kernel(const GridView G, const double* __restrict__ src, int64_t N,
                                                              const double* __restrict__ T, double radius, int64_t* __restrict__ out_count,
                                                              double* __restrict__ out_sum) {
    __shared__ double sT[16];
    if (threadIdx.x < 16) sT[threadIdx.x] = T[(int64_t)blockIdx.x * 16 + threadIdx.x];
    __syncthreads();
    long long cnt = 0;
    double sum = 0.0;
    for (int64_t i = threadIdx.x; i < N; i += blockDim.x) {
        const double x = src[3 * i], y = src[3 * i + 1], z = src[3 * i + 2];
        double h[4];
#pragma unroll
        for (int r = 0; r < 4; r++) h[r] = xdadd(xdadd(xdadd(xdmul(sT[4 * r], x), xdmul(sT[4 * r + 1], y)), xdmul(sT[4 * r + 2], z)), sT[4 * r + 3]);
        double bd[1]; int32_t bi[1];
        if (knn_query<1>(G, xddiv(h[0], h[3]), xddiv(h[1], h[3]), xddiv(h[2], h[3]), 1, radius, bd, bi) > 0) { cnt++; sum = xdadd(sum, bd[0]); }
    }
    __shared__ long long s_c[8];
    __shared__ double s_s[8];
    for (int o = 16; o > 0; o >>= 1) { cnt += __shfl_down_sync(0xffffffffu, cnt, o); sum = xdadd(sum, __shfl_down_sync(0xffffffffu, sum, o)); }
    if ((threadIdx.x & 31) == 0) { s_c[threadIdx.x >> 5] = cnt; s_s[threadIdx.x >> 5] = sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        long long c = 0; double s = 0.0;
        for (int w = 0; w < 8; w++) { c += s_c[w]; s = xdadd(s, s_s[w]); }
        out_count[blockIdx.x] = c; out_sum[blockIdx.x] = s;
    }
}
}  // namespace

// tree indexes the TARGET; src DEVICE [N][3]; T DEVICE [H][16] row-major.  count[h] = source points with a target point at
// squared distance < max_dist^2 after T[h] (KDTreeFlann::SearchHybrid(point, max_dist, 1) as upstream calls it), sumd2[h] = the
// sum of those squared distances: fitness = count / N, inlier_rmse = sqrt(sumd2 / count).
B2R_API int b2r_knn_evaluate_transforms(const b2r_knn* h, const double* src_xyz, int64_t N, const double* T, int64_t H, double max_dist,
                                        int64_t* count, double* sumd2, cudaStream_t stream) {
    if (!h || !src_xyz || !T || !count || !sumd2 || N < 0 || H < 0 || !(max_dist > 0.0)) { b2r_set_error("b2r_knn_evaluate_transforms: bad argument"); return B2R_ERR_BAD_ARG; }
    if (H == 0) return B2R_OK;
    eval_transforms_kernel<<<(unsigned)H, 256, 0, stream>>>(view_of(h), src_xyz, N, T, max_dist, count, sumd2);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}
