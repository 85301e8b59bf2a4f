// costmap.cu — the ego-frame costmap warp of map_engine (sm_100a) behind costmap_warp_batch().
//
// Replaces the per-cell loop of LocalCostmap::odomCallback (CILQR/src/map_engine/src/local_costmap.cpp:242-264),
// which walks a grid_map::GridMapIterator, rotates each cell centre by the ego pose and reads the global map
// through GridMap::atPosition.  The index arithmetic of grid_map_core (src/GridMapMath.cpp:114-157, GridMap.cpp:45-62)
// is reproduced operation by operation with explicitly rounded FP64 intrinsics (no FMA contraction), so the
// source-index map is bit-identical to the reference's; the payload is a 4-byte copy, which makes the kernel
// HBM-bound: 4 B read + 4 B written per cell (+ 4 B when the index map is emitted).
//
// Two kernels, same arithmetic:
//   warp_direct_kernel : 64x64-cell target tile per CTA, gathers straight from global/L2 through the read-only path.
//   warp_tma_kernel    : one elected thread issues a TMA tile load (cp.async.bulk.tensor.2d + mbarrier) of the
//                        bounding box of the tile's rotated footprint in the source map into shared memory
//                        (out-of-map elements are NaN-filled by the TMA unit); cells are gathered from smem.
// Both write 128-bit vectors along index 0 (the contiguous direction of grid_map's column-major layers).
#include "common.cuh"
#include <cuda.h>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace cilqr
{
constexpr int kTile = 64;        // target tile edge (cells)
constexpr int kBox = 96;         // TMA box edge (source cells): ceil(64*sqrt(2)) + margins, multiple of 4
constexpr int kWarpThreads = 256;

struct WarpGeom {
    int rows_t, cols_t, rows_s, cols_s;
    int tiles_i, tiles_j;
    double base_x, base_y, r_t;                 // target: position + (0.5*length - 0.5*resolution), resolution
    double hLx, hLy, Psx, Psy, r_s, rinv_s;     // source: 0.5*length, position, resolution, 1/resolution
    double Lx, Ly;                              // source length (size*resolution)
};

// getPositionFromIndex (GridMapMath.cpp:114-127): position = (mapPosition + offset) + resolution * (-index).
__device__ __forceinline__ double cell_centre(double base, double r, int index)
{
    return __dadd_rn(base, __dmul_rn(r, -(double)index));
}

// One coordinate of getIndexFromPosition (GridMapMath.cpp:129-142): (int)( -(((p - 0.5*L) - P) / r) ), C++ truncation.
// Fast path: multiply by 1/r and take the integer part with the 2^52 trick; whenever the product is within 2^-23 of
// an integer (where the last-ulp difference between n*(1/r) and n/r could change the truncation) or negative, the
// exact IEEE division is evaluated instead, so the result always equals the reference's.
__device__ __forceinline__ int source_index(double p, double hL, double Ps, double r, double rinv)
{
    const double n = __dsub_rn(__dsub_rn(p, hL), Ps);
    const double t = -__dmul_rn(n, rinv);
    const double magic = 6755399441055744.0; // 2^52 + 2^51
    const double tm = __dadd_rn(t, magic);
    const double d = __dsub_rn(t, __dsub_rn(tm, magic)); // t - nearest integer, exact
    const int hi = __double2hiint(d);
    const bool near_int = (hi & 0x7fffffff) < 0x3E800000; // |d| < 2^-23 (also d == 0)
    if (near_int || t < 0.0 || !(t < 2.0e9))
        return __double2int_rz(-__ddiv_rn(n, r));
    return __double2loint(tm) - (hi < 0 ? 1 : 0); // floor(t) for t >= 0
}

// checkIfPositionWithinMap (GridMapMath.cpp:144-157) && checkIfIndexInRange (:191-198).  Note the different
// subtraction order ((p - P) - 0.5*L) from the index computation.
__device__ __forceinline__ bool source_valid(const WarpGeom &G, double px, double py, int g0, int g1)
{
    if (g0 >= 1 && g1 >= 1 && g0 <= G.rows_s - 2 && g1 <= G.cols_s - 2)
        return true; // a whole cell away from every border: the position test cannot fail
    const double qx = -__dsub_rn(__dsub_rn(px, G.Psx), G.hLx), qy = -__dsub_rn(__dsub_rn(py, G.Psy), G.hLy);
    return qx >= 0.0 && qy >= 0.0 && qx < G.Lx && qy < G.Ly && g0 >= 0 && g1 >= 0 && g0 < G.rows_s && g1 < G.cols_s;
}

struct TileTerms {
    double rowC[kTile], rowS[kTile], colS[kTile], colC[kTile];
};

// Cx*cos, Cx*sin per tile row and Cy*sin, Cy*cos per tile column (local_costmap.cpp:249-250); each is one rounded
// multiply in the reference as well, so sharing them across the tile does not change a single bit.
__device__ __forceinline__ void fill_tile_terms(const WarpGeom &G, TileTerms &T, int i0, int j0, double sn, double cn, int tid)
{
    if (tid < kTile) {
        const double Cx = cell_centre(G.base_x, G.r_t, i0 + tid);
        T.rowC[tid] = __dmul_rn(Cx, cn);
        T.rowS[tid] = __dmul_rn(Cx, sn);
    } else if (tid < 2 * kTile) {
        const int q = tid - kTile;
        const double Cy = cell_centre(G.base_y, G.r_t, j0 + q);
        T.colS[q] = __dmul_rn(Cy, sn);
        T.colC[q] = __dmul_rn(Cy, cn);
    }
}

__device__ __forceinline__ void cell_position(const TileTerms &T, int li, int lj, double Vx, double Vy, double &px, double &py)
{
    px = __dadd_rn(__dsub_rn(T.rowC[li], T.colS[lj]), Vx); // (Cx*cos - Cy*sin) + Vx
    py = __dadd_rn(__dadd_rn(T.rowS[li], T.colC[lj]), Vy); // (Cx*sin + Cy*cos) + Vy
}

template <bool TMA>
__device__ __forceinline__ void warp_tile_body(const WarpGeom &G, const TileTerms &T, const float *__restrict__ src, const float *sbox,
                                               int box0, int box1, double Vx, double Vy, int i0, int j0, const float *__restrict__ bbox,
                                               float *__restrict__ out, int32_t *__restrict__ idx, int tid)
{
    const bool vec_ok = (G.rows_t & 3) == 0; // 16-byte aligned quads along index 0
    const float qnan = __int_as_float(0x7fc00000);
    for (int item = tid; item < (kTile / 4) * kTile; item += kWarpThreads) {
        const int lq = item % (kTile / 4), lj = item / (kTile / 4);
        const int i = i0 + 4 * lq, j = j0 + lj;
        if (j >= G.cols_t || i >= G.rows_t)
            continue;
        float v[4];
        int gidx[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            double px, py;
            cell_position(T, 4 * lq + e, lj, Vx, Vy, px, py);
            const int g0 = source_index(px, G.hLx, G.Psx, G.r_s, G.rinv_s);
            const int g1 = source_index(py, G.hLy, G.Psy, G.r_s, G.rinv_s);
            const bool ok = (i + e < G.rows_t) && source_valid(G, px, py, g0, g1);
            gidx[e] = ok ? g0 + G.rows_s * g1 : -1;
            if (TMA) {
                const int l0 = g0 - box0, l1 = g1 - box1;
                if (ok && (unsigned)l0 < (unsigned)kBox && (unsigned)l1 < (unsigned)kBox)
                    v[e] = sbox[l1 * kBox + l0];
                else
                    v[e] = ok ? __ldg(src + gidx[e]) : qnan;
            } else {
                v[e] = ok ? __ldg(src + gidx[e]) : qnan; // the reference throws for out-of-map cells (GridMap.cpp:160-166)
            }
        }
        const size_t l = (size_t)j * G.rows_t + i;
        if (vec_ok) {
            if (bbox) { // local_costmap.cpp:260-263: bounding-box layer wins when > 90
                const float4 bb = *reinterpret_cast<const float4 *>(bbox + l);
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
    }
}

__global__ void __launch_bounds__(kWarpThreads) warp_direct_kernel(const WarpGeom G, const float *__restrict__ src,
                                                                   const double *__restrict__ poses, const float *__restrict__ bbox,
                                                                   float *__restrict__ out, int32_t *__restrict__ idx)
{
    __shared__ TileTerms T;
    const int tid = threadIdx.x;
    const int tile = blockIdx.x, pose = blockIdx.y;
    const int i0 = (tile % G.tiles_i) * kTile, j0 = (tile / G.tiles_i) * kTile;
    const double Vx = poses[4 * pose], Vy = poses[4 * pose + 1], sn = poses[4 * pose + 2], cn = poses[4 * pose + 3];
    fill_tile_terms(G, T, i0, j0, sn, cn, tid);
    __syncthreads();
    const size_t cells = (size_t)G.rows_t * G.cols_t;
    warp_tile_body<false>(G, T, src, nullptr, 0, 0, Vx, Vy, i0, j0, bbox ? bbox + pose * cells : nullptr, out + pose * cells,
                          idx ? idx + pose * cells : nullptr, tid);
}

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(kWarpThreads) warp_tma_kernel(const __grid_constant__ CUtensorMap tmap, const WarpGeom G,
                                                                const float *__restrict__ src, const double *__restrict__ poses,
                                                                const float *__restrict__ bbox, float *__restrict__ out,
                                                                int32_t *__restrict__ idx)
{
    __shared__ __align__(128) float sbox[kBox * kBox];
    __shared__ TileTerms T;
    __shared__ __align__(8) unsigned long long bar;
    __shared__ int s_box[2];
    const int tid = threadIdx.x;
    const int tile = blockIdx.x, pose = blockIdx.y;
    const int i0 = (tile % G.tiles_i) * kTile, j0 = (tile / G.tiles_i) * kTile;
    const double Vx = poses[4 * pose], Vy = poses[4 * pose + 1], sn = poses[4 * pose + 2], cn = poses[4 * pose + 3];
    fill_tile_terms(G, T, i0, j0, sn, cn, tid);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        // Bounding box of the tile's footprint: the map (i,j) -> source index is affine up to truncation, so its
        // extremes over the tile are attained at the four corners; one cell of margin absorbs the truncation.
        const int li1 = min(kTile, G.rows_t - i0) - 1, lj1 = min(kTile, G.cols_t - j0) - 1;
        int lo0 = INT_MAX, lo1 = INT_MAX;
#pragma unroll
        for (int cnr = 0; cnr < 4; ++cnr) {
            double px, py;
            cell_position(T, (cnr & 1) ? li1 : 0, (cnr & 2) ? lj1 : 0, Vx, Vy, px, py);
            lo0 = min(lo0, source_index(px, G.hLx, G.Psx, G.r_s, G.rinv_s));
            lo1 = min(lo1, source_index(py, G.hLy, G.Psy, G.r_s, G.rinv_s));
        }
        // The truncating index map is monotone per axis, so the corner minimum is the exact lower bound and the tile
        // spans at most ceil(63*sqrt(2)) + 1 = 91 source cells per axis.  TMA needs the innermost start coordinate on a
        // 16-byte boundary (a misaligned one raises "illegal instruction", measured on B200): round index 0 down to a
        // multiple of 4 floats (two's complement: also correct for negative, out-of-map coordinates) -> <= 94 <= kBox.
        lo0 &= ~3;
        s_box[0] = lo0;
        s_box[1] = lo1;
        const unsigned bytes = kBox * kBox * sizeof(float);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem_u32(sbox)), "l"(&tmap), "r"(lo0), "r"(lo1), "r"(smem_u32(&bar))
                     : "memory");
    }
    __syncthreads();
    { // wait for the TMA bytes (phase 0)
        unsigned done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done)
                         : "r"(smem_u32(&bar))
                         : "memory");
        }
    }
    const size_t cells = (size_t)G.rows_t * G.cols_t;
    warp_tile_body<true>(G, T, src, sbox, s_box[0], s_box[1], Vx, Vy, i0, j0, bbox ? bbox + pose * cells : nullptr, out + pose * cells,
                         idx ? idx + pose * cells : nullptr, tid);
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
    if (int rc = geom_size_host(src_geom, &G.rows_s, &G.cols_s)) return rc;
    if (int rc = geom_size_host(tgt_geom, &G.rows_t, &G.cols_t)) return rc;
    if ((long long)G.rows_s * G.cols_s > 2147483647LL) return fail(CILQR_E_ARG, "source map has more than 2^31-1 cells");
    if (P == 0) return 0;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }
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
    G.tiles_i = (G.rows_t + kTile - 1) / kTile;
    G.tiles_j = (G.cols_t + kTile - 1) / kTile;

    cudaStream_t st = (cudaStream_t)stream;
    const size_t cells = (size_t)G.rows_t * G.cols_t, scells = (size_t)G.rows_s * G.cols_s;
    const float *d_src = src, *d_bbox = bbox;
    const double *d_poses = poses;
    float *d_out = out;
    int32_t *d_idx = idx;
    if (mem == CILQR_MEM_HOST) {
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
        const cuuint32_t box[2] = {kBox, kBox};
        const cuuint32_t estr[2] = {1, 1};
        const char *dbg = getenv("CILQR_TMA_DEBUG"); // bring-up knob: 1 = zero OOB fill, 2 = no L2 promotion
        const int dbgv = dbg ? atoi(dbg) : 0;
        CUresult cr = encode_tiled_fn()(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(d_src), gdim, gstride, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                        (dbgv & 2) ? CU_TENSOR_MAP_L2_PROMOTION_NONE : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                        (dbgv & 1) ? CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE : CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA);
        if (cr != CUDA_SUCCESS) return fail(CILQR_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)cr);
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

int costmap_propagate_uncertainty(const float *, const costmap_geom *, double, double, double, double, double, int, int32_t *, double *,
                                  int, void *)
{
    return fail(CILQR_E_STATE, "costmap_propagate_uncertainty: not built yet");
}

} // extern "C"
