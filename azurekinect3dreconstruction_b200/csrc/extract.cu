// extract.cu -- surface point and marching-cubes extraction from the hashed TSDF volume (sm_100a).
//
// Replaces volume.extract_point_cloud() / volume.extract_triangle_mesh() (+ compute_vertex_normals)
// (/root/reference/CodeThatWorks/GoodColorReconstruction.py:285-286,305,536-542;
//  CodeThatWorks/1camera-detectsSceneChanges.py:277-291); upstream semantics: SURVEY.md A.5.
//
// Design (not a port of upstream's serial scan + unordered_map welding):
//   * units are processed in SORTED KEY ORDER (cub radix sort of packed keys) so output order is
//     canonical and reproducible: (unit key, voxel x*256+y*16+z, edge axis);
//   * one CTA per unit stages the unit's TSDF with its 1-voxel apron (neighbour units found through
//     the spatial hash) in shared memory, weight==0 encoded as NaN;
//   * every mesh vertex has exactly one OWNER (the voxel at the lower end of its edge + the axis), so
//     welding needs no hash map: a count pass writes per-voxel vertex slots, an exclusive scan over
//     units gives bases, the fill pass resolves any cube edge to owner -> index arithmetically;
//   * compaction = per-thread popc over a z column + block scan (warp shuffles).
#include "tsdf_internal.h"
#include <cub/cub.cuh>
#include <mutex>

#define MC_TABLE_QUAL __device__ __constant__ const
#include "mc_tables.h"

namespace {

struct VolView {
    B2RHash hash;
    const int32_t* unit_keys;
    const float* pool;
    int64_t unit_stride;
    int planes;
    double voxel_length, unit_length;
};

__device__ __forceinline__ int find_unit(const VolView& V, int kx, int ky, int kz) {
    if (!b2r_key_in_range(kx, ky, kz)) return -1;
    return b2r_hash_find(V.hash, b2r_pack_key(kx, ky, kz));
}

// ---- shared-memory apron: A[(x-LO)][(y-LO)][(z-LO)] for x,y,z in [LO, 16] ----------------------
template <int LO, bool RANGE_CHECK>
__device__ __forceinline__ void stage_apron(const VolView& V, int unit, float* A, int* nb /*27*/) {
    constexpr int D = 17 - LO;
    const int kx = V.unit_keys[3 * unit], ky = V.unit_keys[3 * unit + 1], kz = V.unit_keys[3 * unit + 2];
    if (threadIdx.x < 27) {
        const int dx = threadIdx.x / 9 - 1, dy = (threadIdx.x / 3) % 3 - 1, dz = threadIdx.x % 3 - 1;
        int id = -1;
        if (dx == 0 && dy == 0 && dz == 0) id = unit;
        else if (dx >= LO && dy >= LO && dz >= LO) id = find_unit(V, kx + dx, ky + dy, kz + dz);
        nb[threadIdx.x] = id;
    }
    __syncthreads();
    const float nanv = __int_as_float(0x7fc00000);
    auto keep = [&](float t, float w) -> float {
        bool ok = (w != 0.0f);
        if (RANGE_CHECK) ok = ok && (t < 0.98f) && (t >= -0.98f);
        return ok ? t : nanv;
    };
    // interior rows: every (x, y) column holds z = 0..15 of ONE unit, contiguous in both planes -> four 128-bit loads each
    for (int i = threadIdx.x; i < D * D * 4; i += blockDim.x) {
        const int c = i >> 2, part = i & 3;
        const int ax = c / D, ay = c - ax * D;
        int x = ax + LO, y = ay + LO, ox = 1, oy = 1;
        if (x < 0) { x += 16; ox = 0; } else if (x >= 16) { x -= 16; ox = 2; }
        if (y < 0) { y += 16; oy = 0; } else if (y >= 16) { y -= 16; oy = 2; }
        const int u = nb[(ox * 3 + oy) * 3 + 1];
        float* dst = A + c * D + (0 - LO) + part * 4;
        if (u >= 0) {
            const float* base = V.pool + (int64_t)u * V.unit_stride + (x * 256 + y * 16 + part * 4);
            const float4 t = __ldg(reinterpret_cast<const float4*>(base));
            const float4 w = __ldg(reinterpret_cast<const float4*>(base + B2R_UNIT_VOX));
            dst[0] = keep(t.x, w.x); dst[1] = keep(t.y, w.y); dst[2] = keep(t.z, w.z); dst[3] = keep(t.w, w.w);
        } else {
            dst[0] = nanv; dst[1] = nanv; dst[2] = nanv; dst[3] = nanv;
        }
    }
    // caps: z = 16 (and z = -1 when the apron extends below) come from the units above / below
    constexpr int NCAP = (LO < 0) ? 2 : 1;
    for (int i = threadIdx.x; i < D * D * NCAP; i += blockDim.x) {
        const int c = i / NCAP, top = (NCAP == 1) ? 1 : (i - c * NCAP);
        const int ax = c / D, ay = c - ax * D;
        int x = ax + LO, y = ay + LO, ox = 1, oy = 1;
        if (x < 0) { x += 16; ox = 0; } else if (x >= 16) { x -= 16; ox = 2; }
        if (y < 0) { y += 16; oy = 0; } else if (y >= 16) { y -= 16; oy = 2; }
        const int u = nb[(ox * 3 + oy) * 3 + (top ? 2 : 0)];
        float f = nanv;
        if (u >= 0) {
            const float* base = V.pool + (int64_t)u * V.unit_stride + (x * 256 + y * 16 + (top ? 0 : 15));
            f = keep(__ldg(base), __ldg(base + B2R_UNIT_VOX));
        }
        A[c * D + (top ? 16 - LO : 0)] = f;
    }
    __syncthreads();
}

// block-wide exclusive scan of one int per thread (256 threads); returns exclusive prefix, sets total
__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* ws /*>=9 ints*/) {
    int incl = v;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    __syncthreads();
    if (lane == 31) ws[warp] = incl;
    __syncthreads();
    int wbase = 0, tot = 0;
    const int nw = blockDim.x >> 5;
    for (int i = 0; i < nw; i++) { if (i < warp) wbase += ws[i]; tot += ws[i]; }
    *total = tot;
    return wbase + incl - v;
}

// f64 trilinear TSDF sample, SURVEY A.5 GetTSDFAt (reads global memory through the hash)
__device__ double tsdf_at(const VolView& V, double px, double py, double pz) {
    const double vl = V.voxel_length, hv = xdmul(0.5, vl);
    const double q[3] = {xdsub(px, hv), xdsub(py, hv), xdsub(pz, hv)};
    int k[3], idx[3];
    double r[3];
#pragma unroll
    for (int i = 0; i < 3; i++) k[i] = (int)floor(xddiv(q[i], V.unit_length));
    const int u0 = find_unit(V, k[0], k[1], k[2]);
    if (u0 < 0) return 0.0;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const double g = xddiv(xdsub(q[i], xdmul((double)k[i], V.unit_length)), vl);
        int f = (int)floor(g);
        f = min(max(f, 0), 15);
        idx[i] = f;
        r[i] = xdsub(g, (double)f);
    }
    double f[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        int x = idx[0] + mc_shift[i][0], y = idx[1] + mc_shift[i][1], z = idx[2] + mc_shift[i][2];
        int ux = k[0], uy = k[1], uz = k[2];
        bool same = true;
        if (x >= 16) { x -= 16; ux++; same = false; }
        if (y >= 16) { y -= 16; uy++; same = false; }
        if (z >= 16) { z -= 16; uz++; same = false; }
        const int u = same ? u0 : find_unit(V, ux, uy, uz);
        f[i] = (u >= 0) ? (double)V.pool[(int64_t)u * V.unit_stride + (x * 256 + y * 16 + z)] : 0.0;
    }
    const double a0 = xdsub(1.0, r[0]), a1 = xdsub(1.0, r[1]), a2 = xdsub(1.0, r[2]);
    const double t04 = xdadd(xdmul(a2, f[0]), xdmul(r[2], f[4]));
    const double t37 = xdadd(xdmul(a2, f[3]), xdmul(r[2], f[7]));
    const double t15 = xdadd(xdmul(a2, f[1]), xdmul(r[2], f[5]));
    const double t26 = xdadd(xdmul(a2, f[2]), xdmul(r[2], f[6]));
    const double lo = xdadd(xdmul(a1, t04), xdmul(r[1], t37));
    const double hi = xdadd(xdmul(a1, t15), xdmul(r[1], t26));
    return xdadd(xdmul(a0, lo), xdmul(r[0], hi));
}

// ------------------------------------------------------------------------------------------------
// surface points
// ------------------------------------------------------------------------------------------------
#define POINTS_CAP 2048
template <bool FILL>
__global__ void __launch_bounds__(256, 4) extract_points_kernel(const VolView V, const int32_t* order, int32_t* unit_counts,
                                                             const int32_t* unit_offsets, double* xyz, double* rgb,
                                                             double* nrm, int32_t* ids, int64_t capacity) {
    constexpr int D = 17;
    __shared__ float A[D * D * D];
    __shared__ int nb[27];
    __shared__ int ws[9];
    const int pos = blockIdx.x;
    const int unit = order[pos];
    if (FILL && unit_offsets[pos + 1] == unit_offsets[pos]) return;     // the count pass found no surface point in this unit: nothing to stage
    stage_apron<0, true>(V, unit, A, nb);
    const int x = threadIdx.x >> 4, y = threadIdx.x & 15;
    unsigned masks[16];
    int cnt = 0;
#pragma unroll
    for (int z = 0; z < 16; z++) {
        const float f0 = A[(x * D + y) * D + z];
        unsigned m = 0;
        if (f0 == f0) {
            const float fx1 = A[((x + 1) * D + y) * D + z], fy1 = A[(x * D + y + 1) * D + z], fz1 = A[(x * D + y) * D + z + 1];
            if (xfmul(f0, fx1) < 0.0f) m |= 1u;
            if (xfmul(f0, fy1) < 0.0f) m |= 2u;
            if (xfmul(f0, fz1) < 0.0f) m |= 4u;
        }
        masks[z] = m;
        cnt += __popc(m);
    }
    int total;
    int offs = block_exclusive_scan(cnt, &total, ws);
    if (!FILL) {
        if (threadIdx.x == 0) unit_counts[pos] = total;
        return;
    }
    if (total == 0) return;
    const int64_t base = (int64_t)unit_offsets[pos];
    const int kx = V.unit_keys[3 * unit], ky = V.unit_keys[3 * unit + 1], kz = V.unit_keys[3 * unit + 2];
    const double vl = V.voxel_length, hv = xdmul(0.5, vl);
    // Few voxels carry a surface point: compact (voxel, axis) items into a shared list in output order, then consecutive
    // threads take consecutive items -- full warps on the six f64 trilinear samples of each normal.
    __shared__ uint16_t plist[POINTS_CAP];      // x << 10 | y << 6 | z << 2 | axis
    for (int lbase = 0; lbase < total; lbase += POINTS_CAP) {
        int li = offs - lbase;
#pragma unroll
        for (int z = 0; z < 16; z++) {
            const unsigned m = masks[z];
#pragma unroll
            for (int i = 0; i < 3; i++)
                if (m & (1u << i)) {
                    if (li >= 0 && li < POINTS_CAP) plist[li] = (uint16_t)((x << 10) | (y << 6) | (z << 2) | i);
                    li++;
                }
        }
        __syncthreads();
        const int cnt = min(POINTS_CAP, total - lbase);
        for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
            const int64_t o = base + lbase + t;
            if (o >= capacity) continue;
            const int it = plist[t];
            const int px = it >> 10, py = (it >> 6) & 15, pz = (it >> 2) & 15, i = it & 3;
            const int l[3] = {px, py, pz};
            const int kk[3] = {kx, ky, kz};
            double p0[3];
#pragma unroll
            for (int a = 0; a < 3; a++) p0[a] = xdadd(xdadd(hv, xdmul(vl, (double)l[a])), xdmul((double)kk[a], V.unit_length));
            const float f0 = A[(px * D + py) * D + pz];
            const float* v0 = V.pool + (int64_t)unit * V.unit_stride + (px * 256 + py * 16 + pz);
            int n[3] = {px, py, pz};
            n[i] += 1;
            const float f1 = A[(n[0] * D + n[1]) * D + n[2]];
            const float r0 = fabsf(f0), r1 = fabsf(f1);
            double p[3] = {p0[0], p0[1], p0[2]};
            const double p0i = i == 0 ? p0[0] : (i == 1 ? p0[1] : p0[2]);
            const double p1i = xdadd(p0i, vl);
            const double pi = xddiv(xdadd(xdmul(p0i, (double)r1), xdmul(p1i, (double)r0)), (double)xfadd(r0, r1));
            if (i == 0) p[0] = pi; else if (i == 1) p[1] = pi; else p[2] = pi;
            xyz[3 * o] = p[0]; xyz[3 * o + 1] = p[1]; xyz[3 * o + 2] = p[2];
            if (rgb && V.planes == 5) {
                int ux = 1, uy = 1, uz = 1, lx = n[0], ly = n[1], lz = n[2];
                if (lx >= 16) { lx -= 16; ux = 2; }
                if (ly >= 16) { ly -= 16; uy = 2; }
                if (lz >= 16) { lz -= 16; uz = 2; }
                const int u1 = nb[(ux * 3 + uy) * 3 + uz];
                const float* v1 = V.pool + (int64_t)u1 * V.unit_stride + (lx * 256 + ly * 16 + lz);
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    const float c0 = v0[(2 + c) * B2R_UNIT_VOX], c1 = v1[(2 + c) * B2R_UNIT_VOX];
                    const float cc = xfdiv(xfdiv(xfadd(xfmul(c0, r1), xfmul(c1, r0)), xfadd(r0, r1)), 255.0f);
                    rgb[3 * o + c] = (double)cc;
                }
            }
            if (nrm) {
                const double hg = xdmul(0.99, vl);
                double g[3];
                g[0] = xdsub(tsdf_at(V, xdadd(p[0], hg), p[1], p[2]), tsdf_at(V, xdsub(p[0], hg), p[1], p[2]));
                g[1] = xdsub(tsdf_at(V, p[0], xdadd(p[1], hg), p[2]), tsdf_at(V, p[0], xdsub(p[1], hg), p[2]));
                g[2] = xdsub(tsdf_at(V, p[0], p[1], xdadd(p[2], hg)), tsdf_at(V, p[0], p[1], xdsub(p[2], hg)));
                const double len = sqrt(xdadd(xdadd(xdmul(g[0], g[0]), xdmul(g[1], g[1])), xdmul(g[2], g[2])));
                if (len > 0.0) { g[0] = xddiv(g[0], len); g[1] = xddiv(g[1], len); g[2] = xddiv(g[2], len); }
                nrm[3 * o] = g[0]; nrm[3 * o + 1] = g[1]; nrm[3 * o + 2] = g[2];
            }
            if (ids) { ids[4 * o] = kx * 16 + px; ids[4 * o + 1] = ky * 16 + py; ids[4 * o + 2] = kz * 16 + pz; ids[4 * o + 3] = i; }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// marching cubes
// ------------------------------------------------------------------------------------------------
#define MD 18   // apron [-1, 16]
#define MC 17   // cube origins [-1, 15]


// ---- bit-parallel classification ----------------------------------------------------------------------------
// Per (x, y) column of the staged apron (x, y in [-1, 16]) two 18-bit words over z in [-1, 16] (bit z + 1):
//   VAL: the voxel has weight != 0;   NEG: it is valid and its TSDF is negative.
// Per cube column (origins x, y in [-1, 15]) the word CC: bit z + 1 set iff all 8 corners of cube (x, y, z) are valid
// (upstream skips a cube as soon as one corner has weight 0).  Edge ownership, cube activity and cube cases are then
// AND / XOR / shift expressions over these words -- a thread classifies its 16-voxel z column in a few dozen integer
// instructions and only loops over the (few) voxels that own a vertex and the (few) cubes that emit triangles.
#define MCOLS (MD * MD)        // 324 voxel columns
#define CCOLS (MC * MC)        // 289 cube columns

// The words come straight from global memory, without a float apron in shared memory: a thread takes one (x, y) column of the
// 18 x 18 apron columns, loads its 16 interior voxels (tsdf and weight: four 128-bit loads each) from the unit the column lies in
// and the two caps z = -1 / z = 16 from the units below / above, and forms the column's VAL / NEG words in registers.
__device__ __forceinline__ void neighbour_ids(const VolView& V, int unit, int* nb /*27*/) {
    const int kx = V.unit_keys[3 * unit], ky = V.unit_keys[3 * unit + 1], kz = V.unit_keys[3 * unit + 2];
    if (threadIdx.x < 27) {
        const int dx = threadIdx.x / 9 - 1, dy = (threadIdx.x / 3) % 3 - 1, dz = threadIdx.x % 3 - 1;
        nb[threadIdx.x] = (dx == 0 && dy == 0 && dz == 0) ? unit : find_unit(V, kx + dx, ky + dy, kz + dz);
    }
    __syncthreads();
}
__device__ __forceinline__ void stage_bits_direct(const VolView& V, const int* nb, uint32_t* NEG, uint32_t* VAL, uint32_t* CC) {
    for (int c = threadIdx.x; c < MCOLS; c += blockDim.x) {
        const int ax = c / MD, ay = c - ax * MD;
        int x = ax - 1, y = ay - 1, ox = 1, oy = 1;
        if (x < 0) { x += 16; ox = 0; } else if (x >= 16) { x -= 16; ox = 2; }
        if (y < 0) { y += 16; oy = 0; } else if (y >= 16) { y -= 16; oy = 2; }
        const int col = x * 256 + y * 16;
        const int um = nb[(ox * 3 + oy) * 3 + 1], ub = nb[(ox * 3 + oy) * 3], ut = nb[(ox * 3 + oy) * 3 + 2];
        uint32_t n = 0, v = 0;
        if (um >= 0) {
            const float4* tp = reinterpret_cast<const float4*>(V.pool + (int64_t)um * V.unit_stride + col);
            const float4* wp = reinterpret_cast<const float4*>(V.pool + (int64_t)um * V.unit_stride + B2R_UNIT_VOX + col);
            float4 t[4], w[4];
#pragma unroll
            for (int q = 0; q < 4; q++) { t[q] = __ldg(tp + q); w[q] = __ldg(wp + q); }
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float tf[4] = {t[q].x, t[q].y, t[q].z, t[q].w}, wf[4] = {w[q].x, w[q].y, w[q].z, w[q].w};
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const bool ok = wf[k] != 0.0f;
                    v |= ok ? (1u << (4 * q + k + 1)) : 0u;
                    n |= (ok && tf[k] < 0.0f) ? (1u << (4 * q + k + 1)) : 0u;
                }
            }
        }
        if (ub >= 0) {
            const float* b = V.pool + (int64_t)ub * V.unit_stride + col + 15;
            const bool ok = __ldg(b + B2R_UNIT_VOX) != 0.0f;
            v |= ok ? 1u : 0u;
            n |= (ok && __ldg(b) < 0.0f) ? 1u : 0u;
        }
        if (ut >= 0) {
            const float* b = V.pool + (int64_t)ut * V.unit_stride + col;
            const bool ok = __ldg(b + B2R_UNIT_VOX) != 0.0f;
            v |= ok ? (1u << 17) : 0u;
            n |= (ok && __ldg(b) < 0.0f) ? (1u << 17) : 0u;
        }
        NEG[c] = n; VAL[c] = v;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < CCOLS; c += blockDim.x) {
        const int cx = c / MC, cy = c - cx * MC;                       // cube origin (cx - 1, cy - 1)
        const int b = cx * MD + cy;
        const uint32_t v4 = VAL[b] & VAL[b + MD] & VAL[b + 1] & VAL[b + MD + 1];
        CC[c] = v4 & (v4 >> 1);
    }
    __syncthreads();
}

// Masks of thread (x, y), bit z for z in [0, 15]: m[a] = voxel (x, y, z) owns the vertex on its +axis-a edge;
// act = cube (x, y, z) is complete and its corners do not all share a sign.  nn[4] = NEG words of the cube's columns
// (x, y), (x+1, y), (x, y+1), (x+1, y+1) for the cube-case bits.
__device__ __forceinline__ void column_masks(const uint32_t* NEG, const uint32_t* VAL, const uint32_t* CC, int x, int y, uint32_t* m,
                                             uint32_t* act, uint32_t* nn) {
    const int c0 = (x + 1) * MD + (y + 1);
    const uint32_t n0 = NEG[c0], v0 = VAL[c0], nx = NEG[c0 + MD], vx = VAL[c0 + MD], ny = NEG[c0 + 1], vy = VAL[c0 + 1], nxy = NEG[c0 + MD + 1];
    const int q = (x + 1) * MC + (y + 1);
    const uint32_t c11 = CC[q], c01 = CC[q - MC], c10 = CC[q - 1], c00 = CC[q - MC - 1];
    const uint32_t C0 = c11 | c10, C1 = c11 | c01, C2 = c11 | c01 | c10 | c00;
    m[0] = (((n0 ^ nx) & v0 & vx & (C0 | (C0 << 1))) >> 1) & 0xFFFFu;      // cubes (x, y-{0,1}, z-{0,1})
    m[1] = (((n0 ^ ny) & v0 & vy & (C1 | (C1 << 1))) >> 1) & 0xFFFFu;      // cubes (x-{0,1}, y, z-{0,1})
    m[2] = (((n0 ^ (n0 >> 1)) & v0 & (v0 >> 1) & C2) >> 1) & 0xFFFFu;      // cubes (x-{0,1}, y-{0,1}, z)
    uint32_t any = n0 | nx | ny | nxy, all = n0 & nx & ny & nxy;
    any |= any >> 1; all &= all >> 1;
    *act = ((c11 & any & ~all) >> 1) & 0xFFFFu;
    nn[0] = n0; nn[1] = nx; nn[2] = ny; nn[3] = nxy;
}

// marching-cubes case of cube (x, y, z): bit i = corner mc_shift[i] is negative (Bourke corner order:
// (0,0,0) (1,0,0) (1,1,0) (0,1,0) (0,0,1) (1,0,1) (1,1,1) (0,1,1))
__device__ __forceinline__ int cube_case_bits(const uint32_t* nn, int z) {
    int cube = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t w = nn[mc_shift[i][0] + 2 * mc_shift[i][1]];
        cube |= (int)((w >> (z + 1 + mc_shift[i][2])) & 1u) << i;
    }
    return cube;
}

// pass 1: per-voxel vertex slots (written only where a vertex is owned) + per-unit vertex / triangle counts
// The classification of a unit -- everything the fill pass needs besides the few voxels that carry a vertex -- is 3.8 KB of bit
// words: the count pass stores it (record of UBITS words per unit, in sorted-unit order) so that the fill pass does not stage
// the unit's 47 KB apron a second time.  Layout: NEG[324] | VAL[324] | CC[289] | neighbour unit ids [27].
#define UBITS 968
__global__ void __launch_bounds__(256) mesh_count_kernel(const VolView V, const int32_t* order, uint32_t* vinfo,
                                                         int32_t* unit_vcount, int32_t* unit_tcount, uint32_t* ubits) {
    __shared__ uint32_t NEG[MCOLS], VAL[MCOLS], CC[CCOLS];
    __shared__ int nb[27];
    __shared__ int ws[9];
    const int pos = blockIdx.x;
    const int unit = order[pos];
    neighbour_ids(V, unit, nb);
    stage_bits_direct(V, nb, NEG, VAL, CC);
    {
        uint32_t* rec = ubits + (size_t)pos * UBITS;
        for (int i = threadIdx.x; i < MCOLS; i += blockDim.x) { rec[i] = NEG[i]; rec[MCOLS + i] = VAL[i]; }
        for (int i = threadIdx.x; i < CCOLS; i += blockDim.x) rec[2 * MCOLS + i] = CC[i];
        if (threadIdx.x < 27) rec[2 * MCOLS + CCOLS + threadIdx.x] = (uint32_t)nb[threadIdx.x];
    }
    const int x = threadIdx.x >> 4, y = threadIdx.x & 15;
    uint32_t m[3], act, nn[4];
    column_masks(NEG, VAL, CC, x, y, m, &act, nn);
    const int vc = __popc(m[0]) + __popc(m[1]) + __popc(m[2]);
    int tc = 0;
    for (uint32_t a = act; a; a &= a - 1) tc += mc_num_tris[cube_case_bits(nn, __ffs(a) - 1)];
    int vtot, ttot;
    int voff = block_exclusive_scan(vc, &vtot, ws);
    block_exclusive_scan(tc, &ttot, ws);
    uint32_t* vi = vinfo + (int64_t)unit * B2R_UNIT_VOX + (x * 256 + y * 16);
    for (uint32_t mm = m[0] | m[1] | m[2]; mm; mm &= mm - 1) {
        const int z = __ffs(mm) - 1;
        const uint32_t bits = ((m[0] >> z) & 1u) | (((m[1] >> z) & 1u) << 1) | (((m[2] >> z) & 1u) << 2);
        vi[z] = ((uint32_t)voff << 3) | bits;
        voff += __popc(bits);
    }
    if (threadIdx.x == 0) { unit_vcount[pos] = vtot; unit_tcount[pos] = ttot; }
}

__global__ void scatter_bases_kernel(const int32_t* order, const int32_t* scan, int32_t* by_unit, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) by_unit[order[i]] = scan[i];
}

// pass 2: emit vertices (owner voxels) and triangles (cubes).
// Only ~1 % of the voxels own a vertex or an active cube, so the classified columns are first COMPACTED into shared-memory
// work lists (block scans give every item its slot = its output rank), then consecutive threads process consecutive
// items: full warps on the f64 interpolation instead of one or two live lanes per warp, and consecutive output rows.
#define FILL_CAP 2048          // items per round (a unit holds at most 3 * 4096 vertices / 4096 cubes: processed in rounds)
__global__ void __launch_bounds__(256, 4) mesh_fill_kernel(const VolView V, const int32_t* order, const uint32_t* vinfo,
                                                        const int32_t* vbase_by_unit, const int32_t* tbase,
                                                        double* vtx, double* vcol, int32_t* tri, int32_t* ids,
                                                        int64_t capV, int64_t capF, int pos0, int64_t v0, int64_t t0, int vshift,
                                                        const uint32_t* ubits, const int32_t* unit_vcount, const int32_t* unit_tcount) {
    __shared__ uint32_t NEG[MCOLS], VAL[MCOLS], CC[CCOLS];
    __shared__ uint16_t vlist[FILL_CAP];       // x << 10 | y << 6 | z << 2 | axis
    __shared__ uint16_t clist[FILL_CAP];       // x << 8 | y << 4 | z
    __shared__ uint16_t ctoff[FILL_CAP];       // triangle offset of the cube inside the unit (< 5 * 4096)
    __shared__ int nb[27];
    __shared__ int ws[9];
    const int pos = blockIdx.x + pos0;   // a slab handle fills its own contiguous range of the key order
    if (unit_vcount[pos] == 0 && unit_tcount[pos] == 0) return;        // about half of the allocated units hold no part of the surface
    const int unit = order[pos];
    {
        const uint32_t* rec = ubits + (size_t)pos * UBITS;       // the count pass's classification of this unit
        for (int i = threadIdx.x; i < MCOLS; i += blockDim.x) { NEG[i] = rec[i]; VAL[i] = rec[MCOLS + i]; }
        for (int i = threadIdx.x; i < CCOLS; i += blockDim.x) CC[i] = rec[2 * MCOLS + i];
        if (threadIdx.x < 27) nb[threadIdx.x] = (int)rec[2 * MCOLS + CCOLS + threadIdx.x];
    }
    __syncthreads();
    const int x = threadIdx.x >> 4, y = threadIdx.x & 15;
    uint32_t m[3], act, nn[4];
    column_masks(NEG, VAL, CC, x, y, m, &act, nn);
    const int kx = V.unit_keys[3 * unit], ky = V.unit_keys[3 * unit + 1], kz = V.unit_keys[3 * unit + 2];
    const double vl = V.voxel_length, hv = xdmul(0.5, vl);
    const int64_t vb = (int64_t)vbase_by_unit[unit] - v0;   // compact output; vertex ids inside triangles stay global

    // ---- vertices: rank inside the unit = (x, y) column order, then z, then axis -- the numbering of the count pass
    int vtot;
    const int voff = block_exclusive_scan(__popc(m[0]) + __popc(m[1]) + __popc(m[2]), &vtot, ws);
    for (int base = 0; base < vtot; base += FILL_CAP) {
        int li = voff - base;
        for (uint32_t mm = m[0] | m[1] | m[2]; mm; mm &= mm - 1) {
            const int z = __ffs(mm) - 1;
#pragma unroll
            for (int a = 0; a < 3; a++)
                if ((m[a] >> z) & 1u) {
                    if (li >= 0 && li < FILL_CAP) vlist[li] = (uint16_t)((x << 10) | (y << 6) | (z << 2) | a);
                    li++;
                }
        }
        __syncthreads();
        const int cnt = min(FILL_CAP, vtot - base);
        for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
            const int64_t o = vb + base + t;
            if (o >= capV) continue;
            const int it = vlist[t];
            const int px = it >> 10, py = (it >> 6) & 15, pz = (it >> 2) & 15, a = it & 3;
            int n[3] = {px, py, pz};
            n[a] += 1;
            // the two end voxels of an owned edge are valid by construction: their TSDF is read where it lies
            int ux = 1, uy = 1, uz = 1, lx = n[0], ly = n[1], lz = n[2];
            if (lx >= 16) { lx -= 16; ux = 2; }
            if (ly >= 16) { ly -= 16; uy = 2; }
            if (lz >= 16) { lz -= 16; uz = 2; }
            const int u1 = nb[(ux * 3 + uy) * 3 + uz];
            const float* p0 = V.pool + (int64_t)unit * V.unit_stride + (px * 256 + py * 16 + pz);
            const float* p1 = V.pool + (int64_t)u1 * V.unit_stride + (lx * 256 + ly * 16 + lz);
            const float f0f = __ldg(p0), f1f = __ldg(p1);
            const double f0 = fabs((double)f0f), f1 = fabs((double)f1f);
            const int e[3] = {kx * 16 + px, ky * 16 + py, kz * 16 + pz};
            double pt[3];
#pragma unroll
            for (int i = 0; i < 3; i++) pt[i] = xdadd(hv, xdmul(vl, (double)e[i]));
            const double step = xddiv(xdmul(f0, vl), xdadd(f0, f1));
            if (a == 0) pt[0] = xdadd(pt[0], step); else if (a == 1) pt[1] = xdadd(pt[1], step); else pt[2] = xdadd(pt[2], step);
            vtx[3 * o] = pt[0]; vtx[3 * o + 1] = pt[1]; vtx[3 * o + 2] = pt[2];
            if (vcol && V.planes == 5) {
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    const double c0 = xddiv((double)p0[(2 + c) * B2R_UNIT_VOX], 255.0), c1 = xddiv((double)p1[(2 + c) * B2R_UNIT_VOX], 255.0);
                    vcol[3 * o + c] = xddiv(xdadd(xdmul(f1, c0), xdmul(f0, c1)), xdadd(f0, f1));
                }
            }
            if (ids) { ids[4 * o] = e[0]; ids[4 * o + 1] = e[1]; ids[4 * o + 2] = e[2]; ids[4 * o + 3] = a; }
        }
        __syncthreads();
    }

    // ---- triangles: the active cubes, compacted the same way
    int tc = 0;
    for (uint32_t a = act; a; a &= a - 1) tc += mc_num_tris[cube_case_bits(nn, __ffs(a) - 1)];
    int ttot, ctot;
    const int toff = block_exclusive_scan(tc, &ttot, ws);
    if (ttot == 0) return;
    const int coff = block_exclusive_scan(__popc(act), &ctot, ws);
    const int64_t tb = (int64_t)tbase[pos] - t0;
    for (int base = 0; base < ctot; base += FILL_CAP) {
        int li = coff - base, to = toff;
        for (uint32_t a = act; a; a &= a - 1) {
            const int z = __ffs(a) - 1;
            if (li >= 0 && li < FILL_CAP) { clist[li] = (uint16_t)((x << 8) | (y << 4) | z); ctoff[li] = (uint16_t)to; }
            to += mc_num_tris[cube_case_bits(nn, z)];
            li++;
        }
        __syncthreads();
        const int cnt = min(FILL_CAP, ctot - base);
        for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
            const int it = clist[t];
            const int px = it >> 8, py = (it >> 4) & 15, pz = it & 15;
            const int c0 = (px + 1) * MD + (py + 1);
            const uint32_t cn[4] = {NEG[c0], NEG[c0 + MD], NEG[c0 + 1], NEG[c0 + MD + 1]};
            const int cube = cube_case_bits(cn, pz);
            const int nt = mc_num_tris[cube];
            if (!nt) continue;
            int vid[12];
            const int em = mc_edge_table[cube];
#pragma unroll 1
            for (int i = 0; i < 12; i++) {
                vid[i] = -1;
                if (!(em & (1 << i))) continue;
                int ox = px + mc_edge_shift[i][0], oy = py + mc_edge_shift[i][1], oz = pz + mc_edge_shift[i][2];
                const int axis = mc_edge_shift[i][3];
                int ux = 1, uy = 1, uz = 1;
                if (ox >= 16) { ox -= 16; ux = 2; }
                if (oy >= 16) { oy -= 16; uy = 2; }
                if (oz >= 16) { oz -= 16; uz = 2; }
                const int u = nb[(ux * 3 + uy) * 3 + uz];
                const uint32_t info = vinfo[(int64_t)u * B2R_UNIT_VOX + (ox * 256 + oy * 16 + oz)];
                vid[i] = vbase_by_unit[u] + vshift + (int)(info >> 3) + __popc(info & ((1u << axis) - 1u));
            }
            int64_t to2 = tb + ctoff[t];
            for (int k = 0; k < nt; k++, to2++) {
                if (to2 >= capF) break;
                tri[3 * to2] = vid[mc_tri_table[cube][3 * k]];
                tri[3 * to2 + 1] = vid[mc_tri_table[cube][3 * k + 2]];
                tri[3 * to2 + 2] = vid[mc_tri_table[cube][3 * k + 1]];
            }
        }
        __syncthreads();
    }
}

__global__ void pack_keys_kernel(const int32_t* unit_keys, int n, uint64_t* keys, int32_t* ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = b2r_pack_key(unit_keys[3 * i], unit_keys[3 * i + 1], unit_keys[3 * i + 2]);
    ids[i] = i;
}

struct Scratch {
    uint64_t *k_in, *k_out;
    int32_t *id_in, *order;
    int32_t *cnt_a, *cnt_b, *scan_a, *scan_b, *vbase_by_unit;
    uint32_t* vinfo;
    uint32_t* ubits;
    void* cub_tmp;
    size_t cub_bytes;
};

size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

int carve(b2r_tsdf* h, int n, bool mesh, Scratch* S) {
    size_t sort_tmp = 0, scan_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp, (uint64_t*)nullptr, (uint64_t*)nullptr, (int32_t*)nullptr, (int32_t*)nullptr, n);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_tmp, (int32_t*)nullptr, (int32_t*)nullptr, n + 1);
    size_t sort2_tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort2_tmp, (uint32_t*)nullptr, (uint32_t*)nullptr, (int32_t*)nullptr, (int32_t*)nullptr, n);
    if (sort2_tmp > sort_tmp) sort_tmp = sort2_tmp;
    const size_t cub_bytes = align_up(sort_tmp > scan_tmp ? sort_tmp : scan_tmp);
    size_t total = 0;
    auto take = [&](size_t bytes) { size_t o = total; total += align_up(bytes); return o; };
    const size_t o_kin = take((size_t)n * 8), o_kout = take((size_t)n * 8), o_idin = take((size_t)n * 4), o_ord = take((size_t)n * 4);
    const size_t o_ca = take((size_t)(n + 65) * 4), o_cb = take((size_t)(n + 1) * 4), o_sa = take((size_t)(n + 1) * 4), o_sb = take((size_t)(n + 1) * 4);
    const size_t o_vb = take((size_t)n * 4);
    const size_t o_cub = take(cub_bytes);
    const size_t o_ub = mesh ? take((size_t)n * UBITS * 4) : 0;
    const size_t o_vi = mesh ? take((size_t)n * B2R_UNIT_VOX * 4) : 0;
    if (!mesh || h->scratch_bytes < total) h->mesh_gen = 0;   // the scratch is about to be reused / reallocated
    if (h->scratch_bytes < total) h->sort_n = -1;             // ... and a reallocation loses the cached unit order too
    int rc = b2r_tsdf_ensure_scratch(h, total);
    if (rc != B2R_OK) return rc;
    char* b = (char*)h->scratch;
    S->k_in = (uint64_t*)(b + o_kin); S->k_out = (uint64_t*)(b + o_kout);
    S->id_in = (int32_t*)(b + o_idin); S->order = (int32_t*)(b + o_ord);
    S->cnt_a = (int32_t*)(b + o_ca); S->cnt_b = (int32_t*)(b + o_cb);
    S->scan_a = (int32_t*)(b + o_sa); S->scan_b = (int32_t*)(b + o_sb);
    S->vbase_by_unit = (int32_t*)(b + o_vb);
    S->cub_tmp = b + o_cub; S->cub_bytes = cub_bytes;
    S->vinfo = mesh ? (uint32_t*)(b + o_vi) : nullptr;
    S->ubits = mesh ? (uint32_t*)(b + o_ub) : nullptr;
    return B2R_OK;
}

VolView make_view(const b2r_tsdf* h) {
    VolView V;
    V.hash = h->hash; V.unit_keys = h->unit_keys; V.pool = h->pool;
    V.unit_stride = (int64_t)h->planes * B2R_UNIT_VOX; V.planes = h->planes;
    V.voxel_length = h->voxel_length; V.unit_length = h->unit_length;
    return V;
}

// A slab handle (multi-GPU extraction, pipelines/sharded.py) holds the units with kx in [kx_lo, kx_hi] plus ghost layers
// kx_lo - 1, kx_hi + 1, kx_hi + 2.  Packed keys sort kx-major, so in the sorted order the slab is [i0, i1) and the first
// ghost layer above it [i1, i2).  out7 = {i0, i1, i2, vscan[i0], vscan[i1], tscan[i0], tscan[i1]} (scans may be null).
__global__ void slab_bounds_kernel(const uint64_t* sorted_keys, int n, uint64_t k_lo, uint64_t k_hi1, uint64_t k_hi2,
                                   const int32_t* vscan, const int32_t* tscan, int32_t* out7) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    auto lower = [&](uint64_t key) { int lo = 0, hi = n; while (lo < hi) { const int mid = (lo + hi) >> 1; if (sorted_keys[mid] < key) lo = mid + 1; else hi = mid; } return lo; };
    const int i0 = lower(k_lo), i1 = lower(k_hi1), i2 = lower(k_hi2);
    out7[0] = i0; out7[1] = i1; out7[2] = i2;
    out7[3] = vscan ? vscan[i0] : 0; out7[4] = vscan ? vscan[i1] : 0;
    out7[5] = tscan ? tscan[i0] : 0; out7[6] = tscan ? tscan[i1] : 0;
}

struct Slab { int use, kx_lo, kx_hi; };

// Canonical unit order: by packed key (kx-major).
int sorted_order(b2r_tsdf* h, int n, Scratch& S, cudaStream_t stream) {
    // units are only ever appended between resets / imports: with the same unit count in the same epoch, the order (and the
    // sorted keys) left in the scratch by the last extraction still hold -- the steady state of a scan sorts nothing
    if (h->sort_n == n && h->sort_epoch == h->unit_epoch) return B2R_OK;
    h->sort_n = n; h->sort_epoch = h->unit_epoch;
    pack_keys_kernel<<<(n + 255) / 256, 256, 0, stream>>>(h->unit_keys, n, S.k_in, S.id_in);
    B2R_CHECK_LAUNCH();
    size_t tmp = S.cub_bytes;
    B2R_CUDA(cub::DeviceRadixSort::SortPairs(S.cub_tmp, tmp, S.k_in, S.k_out, S.id_in, S.order, n, 0, 63, stream));
    return B2R_OK;
}

int slab_bounds(int n, const Scratch& S, const Slab& sl, const int32_t* vscan, const int32_t* tscan, int32_t* host7, cudaStream_t stream) {
    if (!sl.use) { host7[0] = 0; host7[1] = n; host7[2] = n; return 1; }   // 1: nothing was queued, the caller fills the scan values itself
    auto clampk = [](long long k) { return (int)(k < -(long long)B2R_KEY_BIAS ? -(long long)B2R_KEY_BIAS : (k > (long long)B2R_KEY_BIAS - 1 ? (long long)B2R_KEY_BIAS - 1 : k)); };
    const bool top = (long long)sl.kx_hi + 1 > (long long)B2R_KEY_BIAS - 1;
    const uint64_t k_lo = b2r_pack_key(clampk(sl.kx_lo), -B2R_KEY_BIAS, -B2R_KEY_BIAS);
    const uint64_t k_hi1 = top ? B2R_KEY_EMPTY : b2r_pack_key(clampk((long long)sl.kx_hi + 1), -B2R_KEY_BIAS, -B2R_KEY_BIAS);
    const uint64_t k_hi2 = ((long long)sl.kx_hi + 2 > (long long)B2R_KEY_BIAS - 1) ? B2R_KEY_EMPTY : b2r_pack_key(clampk((long long)sl.kx_hi + 2), -B2R_KEY_BIAS, -B2R_KEY_BIAS);
    int32_t* d7 = S.cnt_a + (size_t)n + 8;     // 64 spare words follow the counts
    slab_bounds_kernel<<<1, 32, 0, stream>>>(S.k_out, n, k_lo, k_hi1, k_hi2, vscan, tscan, d7);
    B2R_CHECK_LAUNCH();
    B2R_CUDA(cudaMemcpyAsync(host7, d7, 7 * 4, cudaMemcpyDeviceToHost, stream));
    return B2R_OK;
}

int extract_points_impl(b2r_tsdf* h, const Slab& sl, double* xyz, double* rgb, double* nrm, int32_t* ids, int64_t capacity, int64_t* count, cudaStream_t stream) {
    if (!h || !count) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int64_t st[5];
    int rc = b2r_tsdf_stats(h, st, stream);
    if (rc != B2R_OK) return rc;
    const int n = (int)st[0];
    *count = 0;
    if (n == 0) return B2R_OK;
    Scratch S;
    rc = carve(h, n, false, &S);
    if (rc != B2R_OK) return rc;
    rc = sorted_order(h, n, S, stream);
    if (rc != B2R_OK) return rc;
    int32_t b7[7];
    rc = slab_bounds(n, S, sl, nullptr, nullptr, b7, stream);
    if (rc < 0) return rc;
    if (rc == B2R_OK) B2R_CUDA(cudaStreamSynchronize(stream));
    const int n_own = b7[1] - b7[0];
    if (n_own <= 0) return B2R_OK;
    const int32_t* order = S.order + b7[0];     // a slab handle emits the points of its own units only
    const VolView V = make_view(h);
    extract_points_kernel<false><<<n_own, 256, 0, stream>>>(V, order, S.cnt_a, nullptr, nullptr, nullptr, nullptr, nullptr, 0);
    B2R_CHECK_LAUNCH();
    B2R_CUDA(cudaMemsetAsync(S.cnt_a + n_own, 0, 4, stream));
    size_t tmp = S.cub_bytes;
    B2R_CUDA(cub::DeviceScan::ExclusiveSum(S.cub_tmp, tmp, S.cnt_a, S.scan_a, n_own + 1, stream));
    int32_t total = 0;
    B2R_CUDA(cudaMemcpyAsync(&total, S.scan_a + n_own, 4, cudaMemcpyDeviceToHost, stream));
    B2R_CUDA(cudaStreamSynchronize(stream));
    *count = total;
    if (!xyz || capacity < total) { b2r_set_error("extract_points: need capacity %d", total); return B2R_ERR_CAPACITY; }
    if (total == 0) return B2R_OK;
    extract_points_kernel<true><<<n_own, 256, 0, stream>>>(V, order, nullptr, S.scan_a, xyz, h->planes == 5 ? rgb : nullptr, nrm, ids, capacity);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

int extract_mesh_impl(b2r_tsdf* h, const Slab& sl, int64_t vertex_base, double* vtx, double* vcol, int32_t* tri, int32_t* ids, int64_t capV,
                      int64_t capF, int64_t* nV, int64_t* nF, cudaStream_t stream) {
    if (!h || !nV || !nF) return B2R_ERR_BAD_ARG;
    B2RDeviceGuard g(h->device);
    int rc = B2R_OK;
    int n = h->mesh_n;
    if (!(h->pending == 0 && h->mesh_gen == h->generation && h->mesh_n > 0)) {
        // (an unchanged volume whose count pass is cached needs no read-back of the unit count: the fill pass is launched
        //  without a host synchronisation)
        int64_t st[5];
        rc = b2r_tsdf_stats(h, st, stream);
        if (rc != B2R_OK) return rc;
        n = (int)st[0];
    }
    *nV = 0; *nF = 0;
    if (n == 0) return B2R_OK;
    Scratch S;
    rc = carve(h, n, true, &S);
    if (rc != B2R_OK) return rc;
    int own[2];
    int32_t v01[2] = {0, 0}, t01[2] = {0, 0};
    const VolView V = make_view(h);
    const uint64_t slab_tag = sl.use ? (((uint64_t)(uint32_t)sl.kx_lo << 32) | (uint32_t)sl.kx_hi) + 1 : 0;
    if (h->mesh_gen == h->generation && h->mesh_n == n && h->mesh_slab == slab_tag) {
        // second call of the two-call protocol on an unchanged volume: order, vertex slots and scans are still in scratch
        own[0] = h->mesh_own[0]; own[1] = h->mesh_own[1];
        v01[0] = h->mesh_v01[0]; v01[1] = h->mesh_v01[1]; t01[0] = h->mesh_t01[0]; t01[1] = h->mesh_t01[1];
    } else {
        rc = sorted_order(h, n, S, stream);
        if (rc != B2R_OK) return rc;
        // the count pass covers EVERY unit the handle holds: it fixes the vertex numbering (key order), which a cube needs for
        // edges owned by the neighbouring units above it -- for a slab handle those are the first ghost layer, whose own
        // neighbourhood is complete thanks to the second one (the counts of the outermost ghost layers are never used)
        mesh_count_kernel<<<n, 256, 0, stream>>>(V, S.order, S.vinfo, S.cnt_a, S.cnt_b, S.ubits);
        B2R_CHECK_LAUNCH();
        B2R_CUDA(cudaMemsetAsync(S.cnt_a + n, 0, 4, stream));
        B2R_CUDA(cudaMemsetAsync(S.cnt_b + n, 0, 4, stream));
        size_t tmp = S.cub_bytes;
        B2R_CUDA(cub::DeviceScan::ExclusiveSum(S.cub_tmp, tmp, S.cnt_a, S.scan_a, n + 1, stream));
        tmp = S.cub_bytes;
        B2R_CUDA(cub::DeviceScan::ExclusiveSum(S.cub_tmp, tmp, S.cnt_b, S.scan_b, n + 1, stream));
        scatter_bases_kernel<<<(n + 255) / 256, 256, 0, stream>>>(S.order, S.scan_a, S.vbase_by_unit, n);
        B2R_CHECK_LAUNCH();
        int32_t b7[7];
        rc = slab_bounds(n, S, sl, S.scan_a, S.scan_b, b7, stream);
        if (rc < 0) return rc;
        if (rc == 1) {
            B2R_CUDA(cudaMemcpyAsync(&b7[4], S.scan_a + n, 4, cudaMemcpyDeviceToHost, stream));
            B2R_CUDA(cudaMemcpyAsync(&b7[6], S.scan_b + n, 4, cudaMemcpyDeviceToHost, stream));
            b7[3] = 0; b7[5] = 0;
        }
        B2R_CUDA(cudaStreamSynchronize(stream));
        own[0] = b7[0]; own[1] = b7[1];
        v01[0] = b7[3]; v01[1] = b7[4]; t01[0] = b7[5]; t01[1] = b7[6];
        h->mesh_gen = h->generation; h->mesh_n = n; h->mesh_slab = slab_tag;
        h->mesh_own[0] = own[0]; h->mesh_own[1] = own[1];
        h->mesh_v01[0] = v01[0]; h->mesh_v01[1] = v01[1]; h->mesh_t01[0] = t01[0]; h->mesh_t01[1] = t01[1];
    }
    const int32_t tv = v01[1] - v01[0], tf = t01[1] - t01[0];
    *nV = tv; *nF = tf;
    if (!vtx || !tri || capV < tv || capF < tf) { b2r_set_error("extract_mesh: need capacity V=%d F=%d", tv, tf); return B2R_ERR_CAPACITY; }
    if (tv == 0 || own[1] == own[0]) return B2R_OK;
    // triangle vertex ids: local key-order id - (ids before this slab) + the slab's global base
    const long long vshift = vertex_base - (long long)v01[0];
    if (vertex_base < 0 || vertex_base + (long long)tv > (long long)0x7fffffff) { b2r_set_error("extract_mesh: vertex ids exceed int32"); return B2R_ERR_RANGE; }
    mesh_fill_kernel<<<own[1] - own[0], 256, 0, stream>>>(V, S.order, S.vinfo, S.vbase_by_unit, S.scan_b, vtx, h->planes == 5 ? vcol : nullptr, tri, ids,
                                                        capV, capF, own[0], (int64_t)v01[0], (int64_t)t01[0], (int)vshift, S.ubits, S.cnt_a, S.cnt_b);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}

}  // namespace

// volume.extract_point_cloud().  Outputs (DEVICE, f64): xyz/rgb/nrm [capacity*3]; ids (optional int32
// [capacity*4]) = global voxel coordinate of the edge's lower end + axis (the canonical sort key).
// Call with xyz == NULL (or too small a capacity) to get the required count in *count and
// B2R_ERR_CAPACITY; a second call with enough room fills.  Synchronises `stream` to read counts.
B2R_API int b2r_tsdf_extract_points(b2r_tsdf* h, double* xyz, double* rgb, double* nrm, int32_t* ids, int64_t capacity,
                                    int64_t* count, cudaStream_t stream) {
    const Slab sl = {0, 0, 0};
    return extract_points_impl(h, sl, xyz, rgb, nrm, ids, capacity, count, stream);
}

// volume.extract_triangle_mesh().  Outputs (DEVICE): vtx/vcol f64 [capV*3], tri int32 [capF*3],
// ids optional int32 [capV*4] = global integer edge id (X,Y,Z,axis) of each vertex.
// Same two-call capacity protocol as extract_points (nV/nF always written).  Vertices come out in unit-key order
// (kx-major), then voxel, then axis.
B2R_API int b2r_tsdf_extract_mesh(b2r_tsdf* h, double* vtx, double* vcol, int32_t* tri, int32_t* ids, int64_t capV,
                                  int64_t capF, int64_t* nV, int64_t* nF, cudaStream_t stream) {
    const Slab sl = {0, 0, 0};
    return extract_mesh_impl(h, sl, 0, vtx, vcol, tri, ids, capV, capF, nV, nF, stream);
}

// Multi-GPU extraction: the handle holds ONE SLAB of the volume -- every unit with kx in [kx_lo, kx_hi] -- plus the
// ghost layers kx_lo - 1, kx_hi + 1 and kx_hi + 2 (whatever of them exists).  Emitted: the vertices owned by the slab's units and
// the triangles of their cubes.  Triangle vertex ids are GLOBAL: vertex_base (= the vertex count of all slabs below this one,
// known after every rank's first call) + the slab's key-order numbering -- which continues seamlessly into the ghost layer
// above, so the slab-order concatenation of all ranks' outputs is the single-GPU mesh, vertex for vertex.
B2R_API int b2r_tsdf_extract_mesh_slab(b2r_tsdf* h, int kx_lo, int kx_hi, int64_t vertex_base, double* vtx, double* vcol, int32_t* tri,
                                       int32_t* ids, int64_t capV, int64_t capF, int64_t* nV, int64_t* nF, cudaStream_t stream) {
    if (kx_lo > kx_hi) { if (nV) *nV = 0; if (nF) *nF = 0; return B2R_OK; }
    const Slab sl = {1, kx_lo, kx_hi};
    return extract_mesh_impl(h, sl, vertex_base, vtx, vcol, tri, ids, capV, capF, nV, nF, stream);
}
B2R_API int b2r_tsdf_extract_points_slab(b2r_tsdf* h, int kx_lo, int kx_hi, double* xyz, double* rgb, double* nrm, int32_t* ids,
                                         int64_t capacity, int64_t* count, cudaStream_t stream) {
    if (kx_lo > kx_hi) { if (count) *count = 0; return B2R_OK; }
    const Slab sl = {1, kx_lo, kx_hi};
    return extract_points_impl(h, sl, xyz, rgb, nrm, ids, capacity, count, stream);
}

// ---- TriangleMesh.compute_vertex_normals() (GoodColorReconstruction.py:286) --------
// Upstream: ComputeTriangleNormals(false) -- the UN-normalised cross products (so a triangle weighs in by its area) are
// added to their three vertices in triangle order, then every sum is normalised (zero stays zero, NaN -> (0, 0, 1)).
// Here: the 3 F (vertex, triangle) incidences are sorted by vertex with a STABLE radix sort, so each vertex sees its
// triangles in ascending triangle order, and one thread per vertex adds them up in that order: the f64 sums are the
// sequential loop's, bit for bit, and the same on every run (no floating-point atomics).
namespace {
__global__ void incidence_fill_kernel(int32_t* tri_of_corner, int64_t n3) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n3) tri_of_corner[c] = (int32_t)(c / 3);
}
__global__ void vertex_normals_gather_kernel(const double* __restrict__ vtx, const int32_t* __restrict__ tri, const int32_t* __restrict__ vsorted,
                                             const int32_t* __restrict__ tsorted, int64_t n3, double* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const int v = vsorted[i];
    if (i > 0 && vsorted[i - 1] == v) return;          // not the head of this vertex's run
    double sx = 0.0, sy = 0.0, sz = 0.0;
    for (int64_t j = i; j < n3 && vsorted[j] == v; j++) {
        const int t = tsorted[j];
        const int i0 = tri[3 * t], i1 = tri[3 * t + 1], i2 = tri[3 * t + 2];
        const double ax = vtx[3 * i0], ay = vtx[3 * i0 + 1], az = vtx[3 * i0 + 2];
        const double e1x = xdsub(vtx[3 * i1], ax), e1y = xdsub(vtx[3 * i1 + 1], ay), e1z = xdsub(vtx[3 * i1 + 2], az);
        const double e2x = xdsub(vtx[3 * i2], ax), e2y = xdsub(vtx[3 * i2 + 1], ay), e2z = xdsub(vtx[3 * i2 + 2], az);
        sx = xdadd(sx, xdsub(xdmul(e1y, e2z), xdmul(e1z, e2y)));
        sy = xdadd(sy, xdsub(xdmul(e1z, e2x), xdmul(e1x, e2z)));
        sz = xdadd(sz, xdsub(xdmul(e1x, e2y), xdmul(e1y, e2x)));
    }
    const double z = xdadd(xdadd(xdmul(sx, sx), xdmul(sy, sy)), xdmul(sz, sz));
    if (z > 0.0) { const double len = sqrt(z); sx = xddiv(sx, len); sy = xddiv(sy, len); sz = xddiv(sz, len); }
    if (sx != sx) { sx = 0.0; sy = 0.0; sz = 1.0; }
    out[3 * (int64_t)v] = sx; out[3 * (int64_t)v + 1] = sy; out[3 * (int64_t)v + 2] = sz;
}
}  // namespace

B2R_API int b2r_mesh_vertex_normals(const double* vtx, int64_t nV, const int32_t* tri, int64_t nF, double* out, cudaStream_t stream) {
    if (!vtx || !tri || !out || nV < 0 || nF < 0 || nF > (int64_t)0x7fffffff / 3) return B2R_ERR_BAD_ARG;
    if (nV == 0) return B2R_OK;
    B2R_CUDA(cudaMemsetAsync(out, 0, (size_t)nV * 24, stream));     // vertices without a triangle keep the zero normal
    if (nF == 0) return B2R_OK;
    const int64_t n3 = 3 * nF;
    int end_bit = 1;
    while (end_bit < 31 && ((int64_t)1 << end_bit) < nV) end_bit++;
    size_t tmp_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const int32_t*)nullptr, (int32_t*)nullptr, (const int32_t*)nullptr, (int32_t*)nullptr, (int)n3, 0, end_bit, stream);
    const size_t a = align_up((size_t)n3 * 4);
    char* buf = nullptr;
    B2R_CUDA(cudaMallocAsync((void**)&buf, 3 * a + align_up(tmp_bytes), stream));
    int32_t* t_in = (int32_t*)buf;
    int32_t* v_out = (int32_t*)(buf + a);
    int32_t* t_out = (int32_t*)(buf + 2 * a);
    void* tmp = buf + 3 * a;
    const unsigned blocks = (unsigned)((n3 + 255) / 256);
    incidence_fill_kernel<<<blocks, 256, 0, stream>>>(t_in, n3);
    cudaError_t e = cudaGetLastError();
    // the triangle array itself is the key array: corner c = 3 t + k holds the vertex id of incidence (t, k)
    if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, tri, v_out, (const int32_t*)t_in, t_out, (int)n3, 0, end_bit, stream);
    if (e == cudaSuccess) {
        vertex_normals_gather_kernel<<<blocks, 256, 0, stream>>>(vtx, tri, v_out, t_out, n3, out);
        e = cudaGetLastError();
    }
    cudaFreeAsync(buf, stream);
    if (e != cudaSuccess) { b2r_set_error("b2r_mesh_vertex_normals: %s", cudaGetErrorString(e)); return B2R_ERR_CUDA; }
    return B2R_OK;
}

// ---- compute_vertex_normals without any allocation ------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as b2r_mesh_vertex_normals -- the
// per-vertex sums run in ascending triangle order -- but the scratch (per-vertex incidence lists in CSR form) lives in a
// file-scope device array of fixed size instead of a stream-ordered allocation, and no library routine is called:
//   zero the counts -> count corners per vertex -> one-CTA exclusive scan -> fill the lists (atomic cursor; order arbitrary)
//   -> one thread per vertex sorts its (short) list ascending and adds the un-normalised triangle normals in that order.
// Limits: B2R_NOALLOC_MAX_VERTICES vertices, B2R_NOALLOC_MAX_TRIANGLES triangles (8.4 MB of scratch, held by every GPU that
// loads the library until the process exits).  Every call on a device shares the array, which keeps the last call's contents:
// the counts are cleared at the start of each call; calls that overlap on the GPU (two streams) race on it.
#define B2R_NOALLOC_MAX_VERTICES 262144
#define B2R_NOALLOC_MAX_TRIANGLES 524288
namespace {
struct NoAllocScratch {
    int32_t offsets[B2R_NOALLOC_MAX_VERTICES + 1];   // corner counts, then their exclusive scan
    int32_t cursor[B2R_NOALLOC_MAX_VERTICES + 1];
    int32_t incident[3 * B2R_NOALLOC_MAX_TRIANGLES]; // triangle ids, grouped by vertex
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

__global__ void na_zero_kernel(int32_t* p, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = 0;
}
__global__ void na_count_kernel(const int32_t* __restrict__ tri, int64_t n3, int nV, int32_t* counts) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n3) return;
    const int v = tri[c];
    if (v >= 0 && v < nV) atomicAdd(&counts[v], 1);
}
// exclusive scan of counts[0 .. n] in place by ONE CTA of 1024 threads (n <= 262 145: 257 elements per thread); cursor = copy
__global__ void __launch_bounds__(1024) na_scan_kernel(int32_t* counts, int32_t* cursor, int n) {
    __shared__ int part[1024];
    const int per = (n + 1023) / 1024;
    const int lo = min(n, (int)threadIdx.x * per), hi = min(n, lo + per);
    int s = 0;
    for (int i = lo; i < hi; i++) s += counts[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    int run = part[threadIdx.x] - s;
    for (int i = lo; i < hi; i++) { const int c = counts[i]; counts[i] = run; cursor[i] = run; run += c; }
}
__global__ void na_fill_kernel(const int32_t* __restrict__ tri, int64_t n3, int nV, int32_t* cursor, int32_t* incident) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n3) return;
    const int v = tri[c];
    if (v >= 0 && v < nV) incident[atomicAdd(&cursor[v], 1)] = (int32_t)(c / 3);
}
__global__ void na_gather_kernel(const double* __restrict__ vtx, const int32_t* __restrict__ tri, const int32_t* __restrict__ offsets,
                                 int32_t* incident, int nV, double* __restrict__ out) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nV) return;
    const int a = offsets[v], b = offsets[v + 1];
    for (int i = a + 1; i < b; i++) {                 // insertion sort: the list is this thread's own, a handful of entries
        const int key = incident[i];
        int j = i - 1;
        while (j >= a && incident[j] > key) { incident[j + 1] = incident[j]; j--; }
        incident[j + 1] = key;
    }
    double sx = 0.0, sy = 0.0, sz = 0.0;
    for (int i = a; i < b; i++) {
        const int t = incident[i];
        const int i0 = tri[3 * t], i1 = tri[3 * t + 1], i2 = tri[3 * t + 2];
        const double ax = vtx[3 * i0], ay = vtx[3 * i0 + 1], az = vtx[3 * i0 + 2];
        const double e1x = xdsub(vtx[3 * i1], ax), e1y = xdsub(vtx[3 * i1 + 1], ay), e1z = xdsub(vtx[3 * i1 + 2], az);
        const double e2x = xdsub(vtx[3 * i2], ax), e2y = xdsub(vtx[3 * i2 + 1], ay), e2z = xdsub(vtx[3 * i2 + 2], az);
        sx = xdadd(sx, xdsub(xdmul(e1y, e2z), xdmul(e1z, e2y)));
        sy = xdadd(sy, xdsub(xdmul(e1z, e2x), xdmul(e1x, e2z)));
        sz = xdadd(sz, xdsub(xdmul(e1x, e2y), xdmul(e1y, e2x)));
    }
    const double z = xdadd(xdadd(xdmul(sx, sx), xdmul(sy, sy)), xdmul(sz, sz));
    if (z > 0.0) { const double len = sqrt(z); sx = xddiv(sx, len); sy = xddiv(sy, len); sz = xddiv(sz, len); }
    if (sx != sx) { sx = 0.0; sy = 0.0; sz = 1.0; }
    out[3 * (int64_t)v] = sx; out[3 * (int64_t)v + 1] = sy; out[3 * (int64_t)v + 2] = sz;
}

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
int noalloc_scratch(NoAllocScratch** out) {
    static std::mutex mu;
    static NoAllocScratch* ptr[64] = {nullptr};
    int dev = 0;
    B2R_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { b2r_set_error("b2r_mesh_vertex_normals_noalloc: device index %d out of range", dev); return B2R_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        B2R_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        ptr[dev] = reinterpret_cast<NoAllocScratch*>(p);
    }
    *out = ptr[dev];
    return B2R_OK;
}
}  // namespace

B2R_API int b2r_mesh_vertex_normals_noalloc(const double* vtx, int64_t nV, const int32_t* tri, int64_t nF, double* out, cudaStream_t stream) {
    if (!vtx || !tri || !out || nV < 0 || nF < 0) return B2R_ERR_BAD_ARG;
    if (nV > B2R_NOALLOC_MAX_VERTICES || nF > B2R_NOALLOC_MAX_TRIANGLES) {
        b2r_set_error("b2r_mesh_vertex_normals_noalloc: mesh of %lld vertices / %lld triangles exceeds the fixed scratch (max %d vertices, %d triangles); "
                      "use b2r_mesh_vertex_normals", (long long)nV, (long long)nF, B2R_NOALLOC_MAX_VERTICES, B2R_NOALLOC_MAX_TRIANGLES);
        return B2R_ERR_CAPACITY;
    }
    if (nV == 0) return B2R_OK;
    NoAllocScratch* S = nullptr;
    int rc = noalloc_scratch(&S);
    if (rc != B2R_OK) return rc;
    const int n = (int)nV;
    const int64_t n3 = 3 * nF;
    na_zero_kernel<<<(unsigned)((n + 1 + 255) / 256), 256, 0, stream>>>(S->offsets, n + 1);      // the last call's counts are stale
    if (n3) na_count_kernel<<<(unsigned)((n3 + 255) / 256), 256, 0, stream>>>(tri, n3, n, S->offsets);
    na_scan_kernel<<<1, 1024, 0, stream>>>(S->offsets, S->cursor, n + 1);
    if (n3) na_fill_kernel<<<(unsigned)((n3 + 255) / 256), 256, 0, stream>>>(tri, n3, n, S->cursor, S->incident);
    na_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(vtx, tri, S->offsets, S->incident, n, out);
    B2R_CHECK_LAUNCH();
    return B2R_OK;
}
