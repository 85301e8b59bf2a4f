// costmap.cu — the ego-frame costmap warp of map_engine (sm_100a) behind costmap_warp_batch().
//
// Replaces the per-cell loop of LocalCostmap::odomCallback (CILQR/src/map_engine/src/local_costmap.cpp:242-264),
// which walks a grid_map::GridMapIterator, rotates each cell centre by the ego pose and reads the global map
// through GridMap::atPosition.  The payload is a 4-byte copy, so the kernel is meant to be HBM-bound: 4 B read +
// 4 B written per cell (+ 4 B when the index map is emitted).
//
// Bit-exact index maps at copy speed
//   The reference's index is a chain of separately rounded FP64 operations (grid_map_core/src/GridMapMath.cpp:114-157,
//   GridMap.cpp:45-62) followed by a truncation.  Evaluating that chain per cell costs ~80 instructions and made the
//   first version of this kernel issue-bound (profiles/r01a_full.txt).  Now: the map (i, j) -> source coordinate is
//   affine up to those roundings, so each tile evaluates it in 7.25 fixed point relative to the tile's source bounding
//   box (32-bit integer adds: the integer part IS the index, the fraction says how far the cell centre is from a
//   source-cell border).  Whenever the fraction is within 2^-17 cell of a border — where the reference's own roundings could decide the truncation — the cell is
//   re-evaluated with the exact chain (__dmul_rn/__dadd_rn/__dsub_rn/__ddiv_rn, no FMA contraction, both subtraction
//   orders of GridMapMath.cpp:139 vs :150).  The predictor's total error is bounded by 2^-18.8 cell (derivation at
//   tile_setup), so both paths give the reference's integer; tests compare every cell of 10^7..10^8-cell maps.
//
// Two kernels, same index code:
//   warp_direct_kernel : 64x64-cell target tile per CTA, gathers straight from global/L2 through the read-only path.
//   warp_tma_kernel    : one elected thread issues a TMA tile load (cp.async.bulk.tensor.2d + mbarrier) of the
//                        bounding box of the tile's rotated footprint in the source map into shared memory
//                        (out-of-map elements are NaN-filled by the TMA unit); cells are gathered from smem with a
//                        lane mapping / row pitch chosen for few bank conflicts at every rotation angle.
// Both write 128-bit vectors along index 0 (the contiguous direction of grid_map's column-major layers).
#include "common.cuh"
#include <cuda.h>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace cilqr
{
constexpr int kTile = 64;         // target tile edge (cells)
constexpr int kBoxW = 108;        // TMA box, index-0 extent (floats): >= 91 + 3 (16-byte alignment); 108 = 12 mod 32 keeps
                                  // the rotated gather at ~2.7 shared-memory wavefronts per warp load (96 gave 3.4+, worst 16)
constexpr int kBoxH = 92;         // TMA box, index-1 extent: ceil(63*sqrt(2)) + 2
constexpr int kWarpThreads = 256;
constexpr int kFrac = 25;                        // 7.25 unsigned fixed point: box-local source coordinates are < 108 < 2^7
constexpr unsigned kFracMask = (1u << kFrac) - 1;
constexpr unsigned kDelta = 1u << (kFrac - 17);  // 2^-17 cell: fractions closer than this to a source-cell border go exact
constexpr double kMaxCells = 33554432.0;         // 2^25: coordinate magnitude (in source cells) up to which the predictor bound holds

struct WarpGeom {
    int rows_t, cols_t, rows_s, cols_s;
    int tiles_i, tiles_j;
    double base_x, base_y, r_t;                 // target: position + (0.5*length - 0.5*resolution), resolution
    double hLx, hLy, Psx, Psy, r_s, rinv_s;     // source: 0.5*length, position, resolution, 1/resolution
    double Lx, Ly;                              // source length (size*resolution)
    double ratio;                               // r_t / r_s
    double mag;                                 // (|base| + target extent + |0.5 L_s| + |P_s|) / r_s, pose offset added per CTA
    int ahead;                                  // warp_tma_kernel: prefetch the box of the tile `ahead` CTAs later into L2 (0 = off)
};

// getPositionFromIndex (GridMapMath.cpp:114-127): position = (mapPosition + offset) + resolution * (-index).
__device__ __forceinline__ double cell_centre(double base, double r, int index)
{
    return __dadd_rn(base, __dmul_rn(r, -(double)index));
}

// One coordinate of getIndexFromPosition (GridMapMath.cpp:129-142): (int)( -(((p - 0.5*L) - P) / r) ), C++ truncation,
// evaluated exactly as the reference does (IEEE division).
__device__ __noinline__ int source_index_exact(double p, double hL, double Ps, double r)
{
    const double t = -__ddiv_rn(__dsub_rn(__dsub_rn(p, hL), Ps), r);
    if (!(fabs(t) < 2.0e9))
        return t < 0.0 ? INT_MIN : INT_MAX; // NaN / out of int range: invalid either way
    return __double2int_rz(t);
}

// checkIfPositionWithinMap (GridMapMath.cpp:144-157) && checkIfIndexInRange (:191-198).  Note the different
// subtraction order ((p - P) - 0.5*L) from the index computation.
__device__ __forceinline__ bool source_valid_exact(const WarpGeom &G, double px, double py, int g0, int g1)
{
    const double qx = -__dsub_rn(__dsub_rn(px, G.Psx), G.hLx), qy = -__dsub_rn(__dsub_rn(py, G.Psy), G.hLy);
    return qx >= 0.0 && qy >= 0.0 && qx < G.Lx && qy < G.Ly && g0 >= 0 && g1 >= 0 && g0 < G.rows_s && g1 < G.cols_s;
}

struct TileShared {
    double rowC[kTile], rowS[kTile], colS[kTile], colC[kTile]; // Cx*cos, Cx*sin, Cy*sin, Cy*cos: one rounded multiply each
    unsigned Tx0, Ty0;    // 7.25 fixed-point source coordinate of tile cell (0,0), relative to (org0, org1)
    int org0, org1;       // origin of the tile's source bounding box (org0 is a multiple of 4)
    int exact_only;       // predictor bound does not hold for this pose/geometry: every cell takes the exact chain
    int inside;           // the whole bounding box lies inside the source map: no validity tests needed
};

__device__ __forceinline__ void cell_position(const TileShared &T, int li, int lj, double Vx, double Vy, double &px, double &py)
{
    px = __dadd_rn(__dsub_rn(T.rowC[li], T.colS[lj]), Vx); // (Cx*cos - Cy*sin) + Vx     local_costmap.cpp:249
    py = __dadd_rn(__dadd_rn(T.rowS[li], T.colC[lj]), Vy); // (Cx*sin + Cy*cos) + Vy     local_costmap.cpp:250
}

// Per-tile setup.  Threads 0..127 fill the shared row/column products (identical bits to the reference: each is a single
// rounded multiply there as well); thread 0 then derives the bounding-box origin from the exact indices of the four tile
// corners (floor of the reference's coordinate is monotone per axis, so the corner minimum is the exact lower bound and the tile
// spans at most ceil(63*sqrt(2)) + 1 = 91 source cells per axis) and the fixed-point start value.
//
// Predictor error budget (source cells).  The reference's t(i,j) differs from the exact affine map by at most
// ~8 roundings of magnitude <= M = (|coordinates|)/r_s, i.e. 8 * 2^-53 * M.  The predictor starts from the reference's own
// t at tile cell (0,0) (same bound, floored to 2^-25), and adds li*A + lj*B with A, B rounded to 2^-26 each:
// <= 126 * 2^-26 + 2^-25 < 2^-18.9.  For M < 2^25 cells the total is < 2^-18.9 + 2^-24 < 2^-18.8, under 30 % of the 2^-17
// guard band; about 3e-5 of all cells fall in the band and are recomputed exactly.
__device__ __forceinline__ void tile_setup(const WarpGeom &G, TileShared &T, int i0, int j0, double Vx, double Vy, double sn, double cn, int tid)
{
    // Warps 2..5 fill the row/column products; warp 0 finds the bounding box meanwhile — one lane per tile corner and one for
    // cell (0,0), each forming the products it needs itself (the same single rounded multiplies, so the same bits as the
    // table entries) — and one barrier closes the setup.  (First version: thread 0 walked the four corners one after the
    // other behind a barrier of its own; `barrier` was the kernel's top stall reason, profiles/r01h_warp_full.txt.)
    if (tid >= 2 * 32 && tid < 2 * 32 + kTile) {
        const int q = tid - 2 * 32;
        const double Cx = cell_centre(G.base_x, G.r_t, i0 + q);
        T.rowC[q] = __dmul_rn(Cx, cn);
        T.rowS[q] = __dmul_rn(Cx, sn);
    } else if (tid >= 2 * 32 + kTile && tid < 2 * 32 + 2 * kTile) {
        const int q = tid - 2 * 32 - kTile;
        const double Cy = cell_centre(G.base_y, G.r_t, j0 + q);
        T.colS[q] = __dmul_rn(Cy, sn);
        T.colC[q] = __dmul_rn(Cy, cn);
    } else if (tid < 32) {
        const int li1 = min(kTile, G.rows_t - i0) - 1, lj1 = min(kTile, G.cols_t - j0) - 1;
        // lanes 0..3: corners; lane 4 (and the idle lanes, harmlessly): cell (0,0)
        const int li = (tid < 4 && (tid & 1)) ? li1 : 0, lj = (tid < 4 && (tid & 2)) ? lj1 : 0;
        const double Cx = cell_centre(G.base_x, G.r_t, i0 + li), Cy = cell_centre(G.base_y, G.r_t, j0 + lj);
        const double px = __dadd_rn(__dsub_rn(__dmul_rn(Cx, cn), __dmul_rn(Cy, sn)), Vx); // cell_position, local_costmap.cpp:249-250
        const double py = __dadd_rn(__dadd_rn(__dmul_rn(Cx, sn), __dmul_rn(Cy, cn)), Vy);
        const double tx = -__ddiv_rn(__dsub_rn(__dsub_rn(px, G.hLx), G.Psx), G.r_s), ty = -__ddiv_rn(__dsub_rn(__dsub_rn(py, G.hLy), G.Psy), G.r_s);
        // FLOOR of the reference's coordinate (its index truncates toward zero, which is not monotone below zero)
        const int g0 = !(fabs(tx) < 2.0e9) ? (tx < 0.0 ? INT_MIN : INT_MAX) : __double2int_rd(tx);
        const int g1 = !(fabs(ty) < 2.0e9) ? (ty < 0.0 ? INT_MIN : INT_MAX) : __double2int_rd(ty);
        int lo0 = g0, hi0 = g0, lo1 = g1, hi1 = g1; // lanes >= 4 repeat corner 0: neutral in the min / max
#pragma unroll
        for (int d = 1; d < 4; d <<= 1) {
            lo0 = min(lo0, __shfl_xor_sync(0xffffffffu, lo0, d)); hi0 = max(hi0, __shfl_xor_sync(0xffffffffu, hi0, d));
            lo1 = min(lo1, __shfl_xor_sync(0xffffffffu, lo1, d)); hi1 = max(hi1, __shfl_xor_sync(0xffffffffu, hi1, d));
        }
        if (tid == 0) {
            const double mag = G.mag + (fabs(Vx) + fabs(Vy)) * G.rinv_s;
            bool exact_only = !(mag < kMaxCells) || !(G.ratio <= 1.0 + 1e-9);
            // TMA needs the innermost start coordinate on a 16-byte boundary (a misaligned one raises "illegal instruction",
            // measured on B200): round index 0 down to a multiple of 4 floats (two's complement: right for negatives too).
            if (!exact_only) {
                lo0 &= ~3;
                exact_only = (long long)hi0 - lo0 >= kBoxW || (long long)hi1 - lo1 >= kBoxH;
            }
            if (exact_only) { lo0 = 0; lo1 = 0; }
            T.Tx0 = exact_only ? 0u : (unsigned)__double2ll_rd((tx - (double)lo0) * (double)(1u << kFrac));
            T.Ty0 = exact_only ? 0u : (unsigned)__double2ll_rd((ty - (double)lo1) * (double)(1u << kFrac));
            T.org0 = lo0;
            T.org1 = lo1;
            T.exact_only = exact_only ? 1 : 0;
            T.inside = (!exact_only && lo0 >= 0 && lo1 >= 0 && lo0 + kBoxW <= G.rows_s && lo1 + kBoxH <= G.cols_s) ? 1 : 0;
        }
    }
    __syncthreads();
}

// The reference's own arithmetic for one cell (used for the few cells within 2^-20 cell of a source-cell border, and for
// every cell when the predictor bound does not hold).
struct CellResult {
    float v;
    int gidx;
};
__device__ __forceinline__ CellResult exact_cell(const WarpGeom &G, const TileShared &T, const float *__restrict__ src, int li, int lj, double Vx, double Vy)
{
    float v;
    int gidx;
    double px, py;
    cell_position(T, li, lj, Vx, Vy, px, py);
    const int g0 = source_index_exact(px, G.hLx, G.Psx, G.r_s), g1 = source_index_exact(py, G.hLy, G.Psy, G.r_s);
    const bool ok = source_valid_exact(G, px, py, g0, g1);
    gidx = ok ? g0 + G.rows_s * g1 : -1;
    v = ok ? __ldg(src + gidx) : __int_as_float(0x7fc00000); // the reference throws for out-of-map cells (GridMap.cpp:160-166)
    return CellResult{v, gidx};
}

// Overwrites the cells of one quad flagged in `unc` with exactly evaluated ones (same thread wrote the fast-path values just
// before, so program order makes these stores win).
__device__ __noinline__ void fixup_exact(const WarpGeom &G, const TileShared &T, const float *__restrict__ src, unsigned unc, int li0, int lj, int i,
                                         size_t l, double Vx, double Vy, const float *__restrict__ bbox, float *__restrict__ out,
                                         int32_t *__restrict__ idx)
{
    for (int e = 0; e < 4; ++e) {
        if (!((unc >> e) & 1u) || i + e >= G.rows_t)
            continue;
        const CellResult r = exact_cell(G, T, src, li0 + e, lj, Vx, Vy);
        float val = r.v;
        if (bbox) {
            const float b = bbox[l + e];
            if (b > 90.f) val = b;
        }
        out[l + e] = val;
        if (idx) idx[l + e] = r.gidx;
    }
}

template <bool TMA, bool INSIDE>
__device__ __forceinline__ void warp_tile_body(const WarpGeom &G, const TileShared &T, const float *__restrict__ src, const float *sbox, double Vx,
                                               double Vy, double sn, double cn, int i0, int j0, const float *__restrict__ bbox,
                                               float *__restrict__ out, int32_t *__restrict__ idx, int tid)
{
    const bool vec_ok = (G.rows_t & 3) == 0; // 16-byte aligned quads along index 0
    const float qnan = __int_as_float(0x7fc00000);
    // d(source coordinate)/d(target index) in 7.25 fixed point (wrapping unsigned arithmetic): Cx = base - r_t*i,
    // Cy = base - r_t*j, p = R(theta) C + V, t = -(p - ..)/r_s
    const double sc = G.ratio * (double)(1u << kFrac);
    const unsigned Axi = (unsigned)__double2int_rn(sc * cn), Axj = (unsigned)__double2int_rn(-sc * sn);
    const unsigned Ayi = (unsigned)__double2int_rn(sc * sn), Ayj = Axi;
    const int warp = tid >> 5, lane = tid & 31;
    const int org0 = T.org0, org1 = T.org1;
    const int gbase = org1 * G.rows_s + org0;
    const bool exact_only = T.exact_only != 0;
    unsigned unc_all = 0;
    // lane mapping: 4 quads (16 cells along index 0) x 8 rows per warp instruction — compact in the source for every rotation
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
        const int wi = warp + 8 * k;
        const int lq = 4 * (wi & 3) + (lane & 3), lj = 8 * (wi >> 2) + (lane >> 2);
        const int i = i0 + 4 * lq, j = j0 + lj;
        if (j >= G.cols_t || i >= G.rows_t)
            continue;
        float v[4];
        int gidx[4];
        unsigned unc = exact_only ? 0xFu : 0u;
        if (!exact_only) {
            unsigned tx = T.Tx0 + (unsigned)(4 * lq) * Axi + (unsigned)lj * Axj;
            unsigned ty = T.Ty0 + (unsigned)(4 * lq) * Ayi + (unsigned)lj * Ayj;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int ix = (int)(tx >> kFrac), iy = (int)(ty >> kFrac);
                if (((tx + kDelta) & kFracMask) < 2 * kDelta || ((ty + kDelta) & kFracMask) < 2 * kDelta) // within 2^-17 cell of a border
                    unc |= 1u << e;
                bool ok = true;
                if (!INSIDE)
                    ok = (unsigned)(ix + org0) < (unsigned)G.rows_s && (unsigned)(iy + org1) < (unsigned)G.cols_s;
                gidx[e] = ok ? iy * G.rows_s + ix + gbase : -1;
                if (TMA) {
                    const unsigned a = (unsigned)(iy * kBoxW + ix);
                    // out-of-map cells: the TMA unit NaN-fills them, but with its own NaN bit pattern; the ABI promises 0x7fc00000
                    v[e] = (ok && (unsigned)ix < (unsigned)kBoxW && a < (unsigned)(kBoxW * kBoxH)) ? sbox[a] : qnan;
                } else {
                    v[e] = ok ? __ldg(src + gidx[e]) : qnan;
                }
                tx += Axi;
                ty += Ayi;
            }
        }
        const size_t l = (size_t)j * G.rows_t + i;
        if (vec_ok) {
            if (bbox) { // local_costmap.cpp:260-263: bounding-box layer wins when > 90
                const float4 bb = __ldcs(reinterpret_cast<const float4 *>(bbox + l));
                if (bb.x > 90.f) v[0] = bb.x;
                if (bb.y > 90.f) v[1] = bb.y;
                if (bb.z > 90.f) v[2] = bb.z;
                if (bb.w > 90.f) v[3] = bb.w;
            }
            __stcs(reinterpret_cast<float4 *>(out + l), make_float4(v[0], v[1], v[2], v[3]));
            if (idx)
                __stcs(reinterpret_cast<int4 *>(idx + l), make_int4(gidx[0], gidx[1], gidx[2], gidx[3]));
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (i + e < G.rows_t) {
                    float val = v[e];
                    if (bbox && bbox[l + e] > 90.f) val = bbox[l + e];
                    out[l + e] = val;
                    if (idx) idx[l + e] = gidx[e];
                }
        }
        unc_all |= unc << (4 * k);
    }
    // Rare second pass, kept out of the loop above so that the call does not cost the fast path any registers: cells on (or
    // within 2^-17 cell of) a source-cell border are redone with the reference's own chain.
    if (unc_all) {
#pragma unroll 1
        for (int k = 0; k < 4; ++k) {
            const unsigned unc = (unc_all >> (4 * k)) & 0xFu;
            if (!unc)
                continue;
            const int wi = warp + 8 * k;
            const int lq = 4 * (wi & 3) + (lane & 3), lj = 8 * (wi >> 2) + (lane >> 2);
            const int i = i0 + 4 * lq, j = j0 + lj;
            fixup_exact(G, T, src, unc, 4 * lq, lj, i, (size_t)j * G.rows_t + i, Vx, Vy, bbox, out, idx);
        }
    }
}

__global__ void __launch_bounds__(kWarpThreads, 4) warp_direct_kernel(const __grid_constant__ WarpGeom G, const float *__restrict__ src,
                                                                   const double *__restrict__ poses, const float *__restrict__ bbox,
                                                                   float *__restrict__ out, int32_t *__restrict__ idx)
{
    __shared__ TileShared T;
    const int tid = threadIdx.x;
    const int tile = blockIdx.x, pose = blockIdx.y;
    const int i0 = (tile % G.tiles_i) * kTile, j0 = (tile / G.tiles_i) * kTile;
    const double Vx = poses[4 * pose], Vy = poses[4 * pose + 1], sn = poses[4 * pose + 2], cn = poses[4 * pose + 3];
    tile_setup(G, T, i0, j0, Vx, Vy, sn, cn, tid);
    const size_t cells = (size_t)G.rows_t * G.cols_t;
    const float *bb = bbox ? bbox + pose * cells : nullptr;
    int32_t *ix = idx ? idx + pose * cells : nullptr;
    if (T.inside)
        warp_tile_body<false, true>(G, T, src, nullptr, Vx, Vy, sn, cn, i0, j0, bb, out + pose * cells, ix, tid);
    else
        warp_tile_body<false, false>(G, T, src, nullptr, Vx, Vy, sn, cn, i0, j0, bb, out + pose * cells, ix, tid);
}

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(kWarpThreads, 5) warp_tma_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ WarpGeom G,
                                                                const float *__restrict__ src, const double *__restrict__ poses,
                                                                const float *__restrict__ bbox, float *__restrict__ out,
                                                                int32_t *__restrict__ idx)
{
    __shared__ __align__(128) float sbox[kBoxW * kBoxH];
    __shared__ TileShared T;
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x;
    const int tile = blockIdx.x, pose = blockIdx.y;
    const int i0 = (tile % G.tiles_i) * kTile, j0 = (tile / G.tiles_i) * kTile;
    const double Vx = poses[4 * pose], Vy = poses[4 * pose + 1], sn = poses[4 * pose + 2], cn = poses[4 * pose + 3];
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Warp 1 has nothing to do during the setup: it asks the TMA unit to bring the source box of a tile that starts later
    // (G.ahead CTAs further on in launch order) into L2.  The box origin is
    // approximated — a prefetch only has to cover the box — with the same corner rule as tile_setup.  When the source does not fit
    // L2 (4096^2 target from an 8192^2 source) a tile's own TMA load then finds its rows in L2 instead of in DRAM: 0.84 -> 0.88 of
    // the HBM copy rate with value + index map for distances of 148 .. 592 CTAs (740: 0.86, 1110: 0.84, 1480: 0.81 — the boxes
    // leave L2 again before they are used; on an L2-resident source the extra work costs 5 %, so the host switches it on by
    // source size).
    if (G.ahead > 0 && tid >= 32 && tid < 64) {
        const long long next = (long long)pose * gridDim.x + tile + G.ahead;
        const int npose = (int)(next / gridDim.x), ntile = (int)(next - (long long)npose * gridDim.x);
        if (npose < (int)gridDim.y) {
            const int lane = tid - 32;
            const int ni0 = (ntile % G.tiles_i) * kTile, nj0 = (ntile / G.tiles_i) * kTile;
            const double nVx = poses[4 * npose], nVy = poses[4 * npose + 1], nsn = poses[4 * npose + 2], ncn = poses[4 * npose + 3];
            const int li = (lane & 1) ? min(kTile, G.rows_t - ni0) - 1 : 0, lj = (lane & 2) ? min(kTile, G.cols_t - nj0) - 1 : 0;
            const double Cx = G.base_x - G.r_t * (ni0 + li), Cy = G.base_y - G.r_t * (nj0 + lj);
            const double tx = -((Cx * ncn - Cy * nsn + nVx) - G.hLx - G.Psx) * G.rinv_s, ty = -((Cx * nsn + Cy * ncn + nVy) - G.hLy - G.Psy) * G.rinv_s;
            int lo0 = fabs(tx) < 1.0e9 ? __double2int_rd(tx) : 0, lo1 = fabs(ty) < 1.0e9 ? __double2int_rd(ty) : 0;
#pragma unroll
            for (int d = 1; d < 4; d <<= 1) {
                lo0 = min(lo0, __shfl_xor_sync(0xffffffffu, lo0, d));
                lo1 = min(lo1, __shfl_xor_sync(0xffffffffu, lo1, d));
            }
            if (lane == 0)
                asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(&tmap), "r"(lo0 & ~3), "r"(lo1) : "memory");
        }
    }
    tile_setup(G, T, i0, j0, Vx, Vy, sn, cn, tid); // ends with __syncthreads(): barrier init and box origin are visible
    const bool staged = !T.exact_only;
    if (staged && tid == 0) {
        const unsigned bytes = kBoxW * kBoxH * sizeof(float);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(sbox)), "l"(&tmap), "r"(T.org0), "r"(T.org1), "r"(smem_u32(&bar))
                     : "memory");
    }
    if (staged) { // wait for the TMA bytes (phase 0)
        unsigned done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(smem_u32(&bar))
                         : "memory");
        }
    }
    const size_t cells = (size_t)G.rows_t * G.cols_t;
    const float *bb = bbox ? bbox + pose * cells : nullptr;
    int32_t *ix = idx ? idx + pose * cells : nullptr;
    if (T.inside)
        warp_tile_body<true, true>(G, T, src, sbox, Vx, Vy, sn, cn, i0, j0, bb, out + pose * cells, ix, tid);
    else
        warp_tile_body<true, false>(G, T, src, sbox, Vx, Vy, sn, cn, i0, j0, bb, out + pose * cells, ix, tid);
}

// ---- the warp on the wire format (row N2 fused into W1) ---------------------------------------------------------------
// The reference's frame is fromOccupancyGrid (int8 -> float layer, reversed order, -1 -> NaN; GridMapRosConverter.cpp:259-266,
// call site local_costmap.cpp:169) -> warp loop (local_costmap.cpp:242-264) -> toOccupancyGrid (float -> int8 0..100, NaN ->
// -1, reversed; :291-306, call site local_costmap.cpp:298).  The two conversions are a reversal and a per-cell map, so they
// fold into the gather's load and store: this kernel reads the global map as int8 occupancy data and writes the ego-frame map
// as int8 occupancy data — 1 B read + 1 B written per cell instead of 4 + 4 (plus the two conversion passes) — with the same
// index arithmetic (tile_setup, fixed-point predictor, exact chain on border cells), so the index map stays bit-identical.
// Thread = 8 consecutive cells along index 0: eight byte gathers through the read-only path (a 4096^2 source is 16.7 MB of
// int8: L2-resident), one 64-bit store of the reversed bytes, two 128-bit stores of the optional index map.
struct OccTile {
    const float *bb;
    int32_t *ix_out;
    int8_t *out;
    const int8_t *occ_src;
    const uint8_t *sbox;
    const int8_t *lut;
    int a0, a1;
    float dmin, dmax;
};
constexpr int kOccBoxWc = 144, kOccBoxHc = kBoxH; // (= kOccBoxW / kOccBoxH below)

// Thread = 8 consecutive cells along index 0, two rows of the 64 x 64 tile per thread.  FAST: the tile's whole source box lies
// inside the map and is staged — index prediction, one shared-memory byte, one table look-up per cell.  The cells within 2^-17
// cell of a source-cell border (about 3e-5 of all) are redone afterwards with the reference's own chain, outside the hot loop.
template <bool FAST>
__device__ __forceinline__ void occupancy_tile_body(const WarpGeom &G, const TileShared &T, const OccTile &O, double Vx, double Vy, double sn, double cn,
                                                    int i0, int j0, int tid, bool staged = true)
{
    const long long cells = (long long)G.rows_t * G.cols_t, ns = (long long)G.rows_s * G.cols_s;
    const bool vec_ok = (G.rows_t & 7) == 0 && (reinterpret_cast<uintptr_t>(O.out) & 7) == 0;
    const bool bb_vec = O.bb && vec_ok && (reinterpret_cast<uintptr_t>(O.bb) & 15) == 0;
    const bool ix_vec = O.ix_out && vec_ok && (reinterpret_cast<uintptr_t>(O.ix_out) & 15) == 0;
    const double sc = G.ratio * (double)(1u << kFrac);
    const unsigned Axi = (unsigned)__double2int_rn(sc * cn), Axj = (unsigned)__double2int_rn(-sc * sn);
    const unsigned Ayi = (unsigned)__double2int_rn(sc * sn), Ayj = Axi;
    const int org0 = T.org0, org1 = T.org1, gbase = org1 * G.rows_s + org0;
    const bool exact_only = !FAST && T.exact_only != 0;
    const int8_t unknown = O.lut[255]; // the conversion of an unknown (-1) / out-of-map cell
    const int lq = tid & 7;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const int lj = (tid >> 3) + 32 * pass;
        const int i = i0 + 8 * lq, j = j0 + lj;
        if (j >= G.cols_t || i >= G.rows_t)
            continue;
        unsigned tx = T.Tx0 + (unsigned)(8 * lq) * Axi + (unsigned)lj * Axj;
        unsigned ty = T.Ty0 + (unsigned)(8 * lq) * Ayi + (unsigned)lj * Ayj;
        int gidx[8];
        int8_t o[8];
        unsigned unc = exact_only ? 0xffu : 0u;
        const size_t l = (size_t)j * G.rows_t + i;
        if (!exact_only) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int ixs = (int)(tx >> kFrac), iys = (int)(ty >> kFrac);
                if (((tx + kDelta) & kFracMask) < 2 * kDelta || ((ty + kDelta) & kFracMask) < 2 * kDelta) // within 2^-17 cell of a border
                    unc |= 1u << e;
                const int g = iys * G.rows_s + ixs + gbase;
                if (FAST) {
                    gidx[e] = g;
                    // (a predicted coordinate a hair below the box origin wraps around: such a cell is flagged and redone below, the
                    // clamp only keeps this read inside the staged box)
                    o[e] = O.lut[O.sbox[(O.a1 - min(iys, kOccBoxHc - 1)) * kOccBoxWc + (O.a0 - min(ixs, kBoxW - 1))]];
                } else {
                    const bool ok = (unsigned)(ixs + org0) < (unsigned)G.rows_s && (unsigned)(iys + org1) < (unsigned)G.cols_s;
                    gidx[e] = ok ? g : -1;
                    uint8_t byte = 255;
                    if (ok) {
                        if (staged && (unsigned)ixs < (unsigned)kBoxW && (unsigned)iys < (unsigned)kOccBoxHc)
                            byte = O.sbox[(O.a1 - iys) * kOccBoxWc + (O.a0 - ixs)];
                        else
                            byte = (uint8_t)__ldg(O.occ_src + (ns - 1 - (long long)g));
                    }
                    o[e] = ok ? O.lut[byte] : unknown; // the reference throws for out-of-map cells (GridMap.cpp:160-166): unknown here
                }
                tx += Axi;
                ty += Ayi;
            }
        }
        if (unc) { // the reference's own chain for the flagged cells
#pragma unroll 1
            for (int e = 0; e < 8; ++e) {
                if (!((unc >> e) & 1u))
                    continue;
                double px, py;
                cell_position(T, 8 * lq + e, lj, Vx, Vy, px, py);
                const int g0 = source_index_exact(px, G.hLx, G.Psx, G.r_s), g1 = source_index_exact(py, G.hLy, G.Psy, G.r_s);
                const bool ok = source_valid_exact(G, px, py, g0, g1);
                const int g = g0 + G.rows_s * g1;
                const int8_t v = ok ? O.lut[(uint8_t)__ldg(O.occ_src + (ns - 1 - (long long)g))] : unknown;
#pragma unroll
                for (int q = 0; q < 8; ++q) // (constant indices keep the arrays in registers)
                    if (q == e) { gidx[q] = ok ? g : -1; o[q] = v; }
            }
        }
        if (O.bb) { // local_costmap.cpp:260-263: bounding-box layer wins when > 90
            float bbv[8];
            if (bb_vec) {
                const float4 b0 = __ldcs(reinterpret_cast<const float4 *>(O.bb + l)), b1 = __ldcs(reinterpret_cast<const float4 *>(O.bb + l + 4));
                bbv[0] = b0.x; bbv[1] = b0.y; bbv[2] = b0.z; bbv[3] = b0.w; bbv[4] = b1.x; bbv[5] = b1.y; bbv[6] = b1.z; bbv[7] = b1.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    bbv[e] = i + e < G.rows_t ? O.bb[l + e] : 0.0f;
            }
#pragma unroll
            for (int e = 0; e < 8; ++e)
                if (bbv[e] > 90.f) o[e] = occupancy_cell(bbv[e], O.dmin, O.dmax);
        }
        if (vec_ok) { // occ[n - 1 - (l + e)]: the eight bytes in reverse, one 64-bit store
            const unsigned lo = (unsigned)(uint8_t)o[7] | ((unsigned)(uint8_t)o[6] << 8) | ((unsigned)(uint8_t)o[5] << 16) | ((unsigned)(uint8_t)o[4] << 24);
            const unsigned hi = (unsigned)(uint8_t)o[3] | ((unsigned)(uint8_t)o[2] << 8) | ((unsigned)(uint8_t)o[1] << 16) | ((unsigned)(uint8_t)o[0] << 24);
            __stcs(reinterpret_cast<uint2 *>(O.out + (cells - 8 - (long long)l)), make_uint2(lo, hi));
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e)
                if (i + e < G.rows_t) O.out[cells - 1 - (long long)(l + e)] = o[e];
        }
        if (O.ix_out) {
            if (ix_vec) {
                __stcs(reinterpret_cast<int4 *>(O.ix_out + l), make_int4(gidx[0], gidx[1], gidx[2], gidx[3]));
                __stcs(reinterpret_cast<int4 *>(O.ix_out + l + 4), make_int4(gidx[4], gidx[5], gidx[6], gidx[7]));
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (i + e < G.rows_t) O.ix_out[l + e] = gidx[e];
            }
        }
    }
}

constexpr int kOccBoxW = 144; // staged box, index-0 extent in bytes: the 108 source cells a tile can span + up to 15 of alignment (TMA
                              // needs the innermost start coordinate on a 16-byte boundary), a multiple of 16; 144 B = 36 words keeps
                              // consecutive box rows 4 banks apart (128 would put every row on the same banks)
constexpr int kOccBoxH = kBoxH;
static_assert(kOccBoxW == kOccBoxWc && kOccBoxH == kOccBoxHc, "one box size");

// STAGED: the tile's source box arrives through the TMA unit (uint8 tensor map over the occupancy data, whose order is the
// layer's reversed: layer cell (i, j) is occ[(rows - 1 - i) + rows (cols - 1 - j)]), and the int8 -> float -> int8 conversion
// of a cell that no bounding box overrides is one look-up in a 256-entry table that the CTA fills with occupancy_cell itself.
// !STAGED (source rows not a multiple of 16 bytes, or an unaligned buffer): byte gathers through the read-only path.
template <bool STAGED>
__global__ void __launch_bounds__(kWarpThreads, 3) warp_occupancy_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ WarpGeom G,
                                                                       const int8_t *__restrict__ occ_src, const double *__restrict__ poses,
                                                                       const float *__restrict__ bbox, float dmin, float dmax,
                                                                       int8_t *__restrict__ occ_out, int32_t *__restrict__ idx)
{
    __shared__ __align__(128) uint8_t sbox[STAGED ? kOccBoxW * kOccBoxH : 16];
    __shared__ int8_t lut[256];
    __shared__ TileShared T;
    __shared__ __align__(8) unsigned long long bar;
    const int tid = threadIdx.x;
    const int tile = blockIdx.x, pose = blockIdx.y;
    const int i0 = (tile % G.tiles_i) * kTile, j0 = (tile / G.tiles_i) * kTile;
    const double Vx = poses[4 * pose], Vy = poses[4 * pose + 1], sn = poses[4 * pose + 2], cn = poses[4 * pose + 3];
    if (STAGED && tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    { // toOccupancyGrid(fromOccupancyGrid(b)) for every byte value (GridMapRosConverter.cpp:263, 297-302)
        const int8_t b = (int8_t)tid;
        lut[tid & 255] = occupancy_cell(b != -1 ? (float)b : __int_as_float(0x7fc00000), dmin, dmax);
    }
    tile_setup(G, T, i0, j0, Vx, Vy, sn, cn, tid); // ends with __syncthreads()
    const bool staged = STAGED && !T.exact_only;
    // occ coordinates of the box: index 0 runs backwards, so layer column org0 + ix sits at c0 + (a0 - ix)
    const int c0 = (G.rows_s - T.org0 - kBoxW) & ~15, a0 = (G.rows_s - 1 - T.org0) - c0;
    const int c1 = G.cols_s - T.org1 - kOccBoxH, a1 = (G.cols_s - 1 - T.org1) - c1;
    if (STAGED) {
        if (staged && tid == 0) {
            const unsigned bytes = kOccBoxW * kOccBoxH;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(smem_u32(sbox)), "l"(&tmap), "r"(c0), "r"(c1), "r"(smem_u32(&bar))
                         : "memory");
        }
        if (staged) {
            unsigned done = 0;
            while (!done) {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done)
                             : "r"(smem_u32(&bar))
                             : "memory");
            }
        }
    }
    const long long cells = (long long)G.rows_t * G.cols_t;
    OccTile O;
    O.bb = bbox ? bbox + pose * cells : nullptr;
    O.ix_out = idx ? idx + pose * cells : nullptr;
    O.out = occ_out + pose * cells;
    O.occ_src = occ_src;
    O.sbox = sbox;
    O.lut = lut;
    O.a0 = a0;
    O.a1 = a1;
    O.dmin = dmin;
    O.dmax = dmax;
    // whole box inside the map and staged: no validity tests, every gather comes from shared memory
    if (staged && T.inside)
        occupancy_tile_body<true>(G, T, O, Vx, Vy, sn, cn, i0, j0, tid);
    else
        occupancy_tile_body<false>(G, T, O, Vx, Vy, sn, cn, i0, j0, tid, staged);
}

// ------------------------------------------------------------------------------------------------- host helpers
static int geom_size_host(const costmap_geom *g, int *rows, int *cols)
{
    if (!g || !(g->resolution > 0.0) || !(g->length_x > 0.0) || !(g->length_y > 0.0))
        return fail(CILQR_E_ARG, "bad costmap geometry");
    *rows = (int)std::round(g->length_x / g->resolution); // GridMap.cpp:50-52
    *cols = (int)std::round(g->length_y / g->resolution);
    if (*rows < 1 || *cols < 1)
        return fail(CILQR_E_ARG, "costmap geometry has no cells");
    return 0;
}

// setGeometry (GridMap.cpp:45-62) of both maps and the constants of the index chain.
static int make_warp_geom(const costmap_geom *src_geom, const costmap_geom *tgt_geom, WarpGeom &G)
{
    if (int rc = geom_size_host(src_geom, &G.rows_s, &G.cols_s)) return rc;
    if (int rc = geom_size_host(tgt_geom, &G.rows_t, &G.cols_t)) return rc;
    if ((long long)G.rows_s * G.cols_s > 2147483647LL) return fail(CILQR_E_ARG, "source map has more than 2^31-1 cells");
    // setGeometry (GridMap.cpp:57-58): length_ = size * resolution; getVectorToFirstCell (GridMapMath.cpp:41-52)
    const double Ltx = G.rows_t * tgt_geom->resolution, Lty = G.cols_t * tgt_geom->resolution;
    G.r_t = tgt_geom->resolution;
    G.base_x = tgt_geom->pos_x + (0.5 * Ltx - 0.5 * G.r_t);
    G.base_y = tgt_geom->pos_y + (0.5 * Lty - 0.5 * G.r_t);
    G.r_s = src_geom->resolution;
    G.rinv_s = 1.0 / G.r_s;
    G.Lx = G.rows_s * G.r_s;
    G.Ly = G.cols_s * G.r_s;
    G.hLx = 0.5 * G.Lx;
    G.hLy = 0.5 * G.Ly;
    G.Psx = src_geom->pos_x;
    G.Psy = src_geom->pos_y;
    G.ratio = G.r_t / G.r_s;
    G.ahead = 0;
    G.mag = (std::fabs(G.base_x) + std::fabs(G.base_y) + Ltx + Lty + G.hLx + G.hLy + std::fabs(G.Psx) + std::fabs(G.Psy)) / G.r_s;
    G.tiles_i = (G.rows_t + kTile - 1) / kTile;
    G.tiles_j = (G.cols_t + kTile - 1) / kTile;
    return 0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
        else
            cudaGetLastError();
    }
    return fn;
}
} // namespace cilqr

using namespace cilqr;

namespace
{
struct WarpWorkspace { // staging for CILQR_MEM_HOST callers (process-wide, grown on demand)
    DevBuf src, poses, bbox, out, idx;
};
WarpWorkspace g_ws;
// Host-buffer calls share g_ws: one at a time, from the first reserve() to the synchronize that ends the call (a second
// thread's reserve() would free buffers the first thread's kernel is still using).  Device-buffer calls never touch it.
std::mutex g_ws_mutex;
} // namespace

extern "C" {

int costmap_geom_size(const costmap_geom *g, int32_t *rows, int32_t *cols)
{
    if (!rows || !cols) return fail(CILQR_E_ARG, "null argument");
    int r, c;
    if (int rc = geom_size_host(g, &r, &c)) return rc;
    *rows = r;
    *cols = c;
    return 0;
}

int costmap_warp_batch(const float *src, const costmap_geom *src_geom, const costmap_geom *tgt_geom, const double *poses, int P,
                       const float *bbox, float *out, int32_t *idx, int variant, int mem, void *stream)
{
    if (!src || !poses || !out) return fail(CILQR_E_ARG, "null argument");
    if (P < 0 || P > 65535) return fail(CILQR_E_ARG, "P = %d outside [0, 65535]", P);
    if (variant < 0 || variant > 2) return fail(CILQR_E_ARG, "unknown variant %d", variant);
    WarpGeom G;
    if (int rc = make_warp_geom(src_geom, tgt_geom, G)) return rc;
    if (P == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }

    cudaStream_t st = (cudaStream_t)stream;
    const size_t cells = (size_t)G.rows_t * G.cols_t, scells = (size_t)G.rows_s * G.cols_s;
    const float *d_src = src, *d_bbox = bbox;
    const double *d_poses = poses;
    float *d_out = out;
    int32_t *d_idx = idx;
    std::unique_lock<std::mutex> ws_lock(g_ws_mutex, std::defer_lock);
    if (mem == CILQR_MEM_HOST) {
        ws_lock.lock();
        if (int rc = g_ws.src.reserve(scells * sizeof(float))) return rc;
        if (int rc = g_ws.poses.reserve((size_t)P * 4 * sizeof(double))) return rc;
        if (int rc = g_ws.out.reserve(P * cells * sizeof(float))) return rc;
        CILQR_CUDA(cudaMemcpyAsync(g_ws.src.p, src, scells * sizeof(float), cudaMemcpyHostToDevice, st));
        CILQR_CUDA(cudaMemcpyAsync(g_ws.poses.p, poses, (size_t)P * 4 * sizeof(double), cudaMemcpyHostToDevice, st));
        d_src = g_ws.src.as<float>();
        d_poses = g_ws.poses.as<double>();
        d_out = g_ws.out.as<float>();
        if (bbox) {
            if (int rc = g_ws.bbox.reserve(P * cells * sizeof(float))) return rc;
            CILQR_CUDA(cudaMemcpyAsync(g_ws.bbox.p, bbox, P * cells * sizeof(float), cudaMemcpyHostToDevice, st));
            d_bbox = g_ws.bbox.as<float>();
        }
        if (idx) {
            if (int rc = g_ws.idx.reserve(P * cells * sizeof(int32_t))) return rc;
            d_idx = g_ws.idx.as<int32_t>();
        }
    }
    // vector stores need 16-byte aligned output rows
    if ((G.rows_t & 3) == 0 && ((((uintptr_t)d_out) & 15) || (d_idx && (((uintptr_t)d_idx) & 15)) || (d_bbox && (((uintptr_t)d_bbox) & 15))))
        return fail(CILQR_E_ARG, "out / idx / bbox must be 16-byte aligned");

    // TMA needs a 16-byte aligned base, a row pitch that is a multiple of 16 bytes and a footprint that fits the box
    const bool tma_possible = (G.rows_s % 4 == 0) && ((((uintptr_t)d_src) & 15) == 0) && (G.r_t <= G.r_s * (1.0 + 1e-9)) && encode_tiled_fn();
    if (variant == 2 && !tma_possible)
        return fail(CILQR_E_ARG, "TMA variant needs rows_src %% 4 == 0, a 16-byte aligned source, target resolution <= source resolution");
    const bool use_tma = variant == 2 || (variant == 0 && tma_possible);
    dim3 grid(G.tiles_i * G.tiles_j, P);
    if (use_tma) {
        CUtensorMap tmap;
        const cuuint64_t gdim[2] = {(cuuint64_t)G.rows_s, (cuuint64_t)G.cols_s};
        const cuuint64_t gstride[1] = {(cuuint64_t)G.rows_s * sizeof(float)};
        const cuuint32_t box[2] = {kBoxW, kBoxH};
        const cuuint32_t estr[2] = {1, 1};
        CUresult cr = encode_tiled_fn()(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(d_src), gdim, gstride, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA);
        if (cr != CUDA_SUCCESS) return fail(CILQR_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)cr);
        { // a source larger than L2: boxes are prefetched three CTAs per SM ahead — 60 % of the resident CTAs (see warp_tma_kernel)
            int dev = 0, l2 = 0, sms = 0;
            CILQR_CUDA(cudaGetDevice(&dev));
            CILQR_CUDA(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev));
            CILQR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
            G.ahead = scells * sizeof(float) > (size_t)l2 ? sms * 3 : 0;
#ifdef CILQR_PROBE_ENV
            if (const char *e = getenv("CILQR_PROBE_WARP_AHEAD")) G.ahead = atoi(e);
#endif
        }
        warp_tma_kernel<<<grid, kWarpThreads, 0, st>>>(tmap, G, d_src, d_poses, d_bbox, d_out, d_idx);
    } else {
        warp_direct_kernel<<<grid, kWarpThreads, 0, st>>>(G, d_src, d_poses, d_bbox, d_out, d_idx);
    }
    CILQR_LAUNCH_CHECK();
    if (mem == CILQR_MEM_HOST) {
        CILQR_CUDA(cudaMemcpyAsync(out, d_out, P * cells * sizeof(float), cudaMemcpyDeviceToHost, st));
        if (idx) CILQR_CUDA(cudaMemcpyAsync(idx, d_idx, P * cells * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CILQR_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

int costmap_warp_occupancy_batch(const int8_t *occ_src, const costmap_geom *src_geom, const costmap_geom *tgt_geom, const double *poses, int P,
                                 const float *bbox, float data_min, float data_max, int8_t *occ_out, int32_t *idx, int mem, void *stream)
{
    if (!occ_src || !poses || !occ_out) return fail(CILQR_E_ARG, "null argument");
    if (P < 0 || P > 65535) return fail(CILQR_E_ARG, "P = %d outside [0, 65535]", P);
    WarpGeom G;
    if (int rc = make_warp_geom(src_geom, tgt_geom, G)) return rc;
    if (P == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cells = (size_t)G.rows_t * G.cols_t, scells = (size_t)G.rows_s * G.cols_s;
    const int8_t *d_src = occ_src;
    const float *d_bbox = bbox;
    const double *d_poses = poses;
    int8_t *d_out = occ_out;
    int32_t *d_idx = idx;
    DevBuf t_src, t_poses, t_bbox, t_out, t_idx; // host buffers: staging of this call only (no shared state between calls)
    if (mem == CILQR_MEM_HOST) {
        if (int rc = t_src.reserve(scells)) return rc;
        if (int rc = t_poses.reserve((size_t)P * 4 * sizeof(double))) { t_src.release(); return rc; }
        if (int rc = t_out.reserve(P * cells)) { t_src.release(); t_poses.release(); return rc; }
        cudaMemcpyAsync(t_src.p, occ_src, scells, cudaMemcpyHostToDevice, st);
        cudaMemcpyAsync(t_poses.p, poses, (size_t)P * 4 * sizeof(double), cudaMemcpyHostToDevice, st);
        d_src = t_src.as<int8_t>();
        d_poses = t_poses.as<double>();
        d_out = t_out.as<int8_t>();
        if (bbox && t_bbox.reserve(P * cells * sizeof(float)) == 0) {
            cudaMemcpyAsync(t_bbox.p, bbox, P * cells * sizeof(float), cudaMemcpyHostToDevice, st);
            d_bbox = t_bbox.as<float>();
        }
        if (idx && t_idx.reserve(P * cells * sizeof(int32_t)) == 0) d_idx = t_idx.as<int32_t>();
    }
    int rc = 0;
    if ((bbox && d_bbox == bbox && mem == CILQR_MEM_HOST) || (idx && d_idx == idx && mem == CILQR_MEM_HOST)) {
        rc = fail(CILQR_E_ALLOC, "out of device memory for the staging buffers");
    } else {
        dim3 grid(G.tiles_i * G.tiles_j, P);
        // the TMA unit needs a 16-byte aligned base and a row pitch that is a multiple of 16 bytes (and the footprint of a tile has to
        // fit the box: target resolution <= source resolution)
        const bool staged = (G.rows_s % 16 == 0) && ((((uintptr_t)d_src) & 15) == 0) && (G.r_t <= G.r_s * (1.0 + 1e-9)) && encode_tiled_fn();
        CUtensorMap tmap;
        std::memset(&tmap, 0, sizeof(tmap));
        CUresult cr = CUDA_SUCCESS;
        if (staged) {
            const cuuint64_t gdim[2] = {(cuuint64_t)G.rows_s, (cuuint64_t)G.cols_s};
            const cuuint64_t gstride[1] = {(cuuint64_t)G.rows_s};
            const cuuint32_t box[2] = {kOccBoxW, kOccBoxH};
            const cuuint32_t estr[2] = {1, 1};
            cr = encode_tiled_fn()(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<int8_t *>(d_src), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        }
        if (staged && cr == CUDA_SUCCESS)
            warp_occupancy_kernel<true><<<grid, kWarpThreads, 0, st>>>(tmap, G, d_src, d_poses, d_bbox, data_min, data_max, d_out, d_idx);
        else
            warp_occupancy_kernel<false><<<grid, kWarpThreads, 0, st>>>(tmap, G, d_src, d_poses, d_bbox, data_min, data_max, d_out, d_idx);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess && mem == CILQR_MEM_HOST) {
            cudaMemcpyAsync(occ_out, d_out, P * cells, cudaMemcpyDeviceToHost, st);
            if (idx) cudaMemcpyAsync(idx, d_idx, P * cells * sizeof(int32_t), cudaMemcpyDeviceToHost, st);
            e = cudaStreamSynchronize(st);
        }
        if (e != cudaSuccess) rc = fail(CILQR_E_CUDA, "costmap_warp_occupancy_batch: %s", cudaGetErrorString(e));
    }
    t_src.release(); t_poses.release(); t_bbox.release(); t_out.release(); t_idx.release();
    return rc;
}

} // extern "C"
