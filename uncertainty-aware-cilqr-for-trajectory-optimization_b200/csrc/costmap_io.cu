// costmap_io.cu — the data formats either side of the costmap warp, sm_100a.
//
//   costmap_from_occupancy / costmap_to_occupancy   nav_msgs/OccupancyGrid int8 wire format <-> grid_map float layer
//       (grid_map_ros/src/GridMapRosConverter.cpp:225-269 and :271-307; call sites local_costmap.cpp:169, :298).
//       Byte streams, HBM-bound: 16 cells per thread, 128-bit accesses on the int8 side, the reversal done in registers.
//   costmap_rasterize_boxes   LocalCostmap::bondingBoxHandle (local_costmap.cpp:860-922): rotated rectangles of the
//       surrounding vehicles filled with 100 in the "bounding_box_map" layer, the overlay input of the warp.  The reference
//       walks a grid_map::PolygonIterator (bounding box of the polygon, Polygon::isInside per cell, Polygon.cpp:32-44); here
//       one thread per cell of the box evaluates the same crossing test with separately rounded operations.
//       The rectangle corners are computed on the HOST with libm, expression by expression like the reference, because a cell
//       centre exactly on an edge is decided by the last bit of a corner coordinate.
#include "common.cuh"
#include <cmath>
#include <vector>

namespace cilqr
{
// ---- N2 -------------------------------------------------------------------------------------------------------
// data(i) = occ[n-1-i] != -1 ? occ[n-1-i] : NaN   (GridMapRosConverter.cpp:259-266)
__global__ void __launch_bounds__(256) from_occupancy_kernel(const int8_t *__restrict__ occ, long long n, float *__restrict__ out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c * 16 < n; c += stride) {
        const long long i0 = c * 16, src_hi = n - 1 - i0; // out[i0 + k] = occ[src_hi - k]
        const float nanv = __int_as_float(0x7fc00000);
        if (i0 + 16 <= n && ((src_hi - 15) & 15) == 0 && (reinterpret_cast<uintptr_t>(occ) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
            const int4 v = *reinterpret_cast<const int4 *>(occ + src_hi - 15);
            const int w[4] = {v.x, v.y, v.z, v.w};
            float f[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int8_t b = (int8_t)((w[(15 - k) >> 2] >> (8 * ((15 - k) & 3))) & 0xff);
                f[k] = b != -1 ? (float)b : nanv;
            }
#pragma unroll
            for (int k = 0; k < 16; k += 4)
                *reinterpret_cast<float4 *>(out + i0 + k) = make_float4(f[k], f[k + 1], f[k + 2], f[k + 3]);
        } else {
            for (int k = 0; k < 16 && i0 + k < n; ++k) {
                const int8_t b = occ[src_hi - k];
                out[i0 + k] = b != -1 ? (float)b : nanv;
            }
        }
    }
}
// occ[n-1-index] = occupancy_cell(data[index])   (GridMapRosConverter.cpp:291-306; occupancy_cell in common.cuh)
__global__ void __launch_bounds__(256) to_occupancy_kernel(const float *__restrict__ data, long long n, float dmin, float dmax, int8_t *__restrict__ occ)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c * 16 < n; c += stride) {
        const long long i0 = c * 16, dst_hi = n - 1 - i0; // occ[dst_hi - k] = f(data[i0 + k])
        if (i0 + 16 <= n && ((dst_hi - 15) & 15) == 0 && (reinterpret_cast<uintptr_t>(occ) & 15) == 0 && (reinterpret_cast<uintptr_t>(data) & 15) == 0) {
            float f[16];
#pragma unroll
            for (int k = 0; k < 16; k += 4) {
                const float4 q = *reinterpret_cast<const float4 *>(data + i0 + k);
                f[k] = q.x; f[k + 1] = q.y; f[k + 2] = q.z; f[k + 3] = q.w;
            }
            int w[4] = {0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 16; ++k)
                w[(15 - k) >> 2] |= ((int)(uint8_t)occupancy_cell(f[k], dmin, dmax)) << (8 * ((15 - k) & 3));
            *reinterpret_cast<int4 *>(occ + dst_hi - 15) = make_int4(w[0], w[1], w[2], w[3]);
        } else {
            for (int k = 0; k < 16 && i0 + k < n; ++k)
                occ[dst_hi - k] = occupancy_cell(data[i0 + k], dmin, dmax);
        }
    }
}

// ---- N3 -------------------------------------------------------------------------------------------------------
struct RasterBox {
    double vx[5], vy[5];    // closed polygon: the four corners + the first one again (local_costmap.cpp:906-911)
    int i_lo, i_hi, j_lo, j_hi; // inclusive index range of cells that can be inside
    int frame;
};
// Polygon::isInside (Polygon.cpp:32-44): crossing number over the 5 stored vertices, same expression, each operation rounded.
__device__ __forceinline__ bool polygon_inside(const RasterBox &b, double px, double py)
{
    int cross = 0;
#pragma unroll
    for (int i = 0, j = 4; i < 5; j = i++) {
        if ((b.vy[i] > py) != (b.vy[j] > py)) {
            const double rhs = __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn(b.vx[j], b.vx[i]), __dsub_rn(py, b.vy[i])), __dsub_rn(b.vy[j], b.vy[i])), b.vx[i]);
            if (px < rhs)
                ++cross;
        }
    }
    return (cross & 1) != 0;
}
// One CTA per box; threads sweep the box's index range.  Cell centre = getPositionFromIndex (GridMapMath.cpp:114-127).
__global__ void __launch_bounds__(256) rasterize_boxes_kernel(const RasterBox *__restrict__ boxes, int rows, int cols, double base_x, double base_y,
                                                              double r, float *__restrict__ layer)
{
    __shared__ RasterBox b;
    if (threadIdx.x == 0)
        b = boxes[blockIdx.x];
    __syncthreads();
    const int ni = b.i_hi - b.i_lo + 1, nj = b.j_hi - b.j_lo + 1;
    if (ni <= 0 || nj <= 0)
        return;
    float *out = layer + (size_t)b.frame * rows * cols;
    for (int c = threadIdx.x; c < ni * nj; c += blockDim.x) {
        const int i = b.i_lo + c % ni, j = b.j_lo + c / ni;
        const double px = __dadd_rn(base_x, __dmul_rn(r, -(double)i)), py = __dadd_rn(base_y, __dmul_rn(r, -(double)j));
        if (polygon_inside(b, px, py))
            out[(size_t)j * rows + i] = 100.0f;
    }
}
} // namespace cilqr

using namespace cilqr;

static int io_grid(long long chunks)
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return (int)std::min<long long>((chunks + 255) / 256, (long long)sms * 8);
}
static int need_device()
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(CILQR_E_CUDA, "no CUDA device: this library has no CPU path");
    }
    return 0;
}

extern "C" {

int costmap_from_occupancy(const int8_t *occ, int64_t n, float *data, int mem, void *stream)
{
    if (!occ || !data || n < 0) return fail(CILQR_E_ARG, "bad argument");
    if (n == 0) return 0;
    if (int rc = need_device()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DevBuf tmp;
    const int8_t *d_in = occ;
    float *d_out = data;
    if (mem == CILQR_MEM_HOST) {
        const size_t off = ((size_t)n + 255) / 256 * 256;
        if (int rc = tmp.reserve(off + (size_t)n * sizeof(float))) return rc;
        CILQR_CUDA(cudaMemcpyAsync(tmp.p, occ, (size_t)n, cudaMemcpyHostToDevice, st));
        d_in = tmp.as<int8_t>();
        d_out = reinterpret_cast<float *>(tmp.as<char>() + off);
    }
    from_occupancy_kernel<<<io_grid((n + 15) / 16), 256, 0, st>>>(d_in, n, d_out);
    CILQR_LAUNCH_CHECK();
    if (mem == CILQR_MEM_HOST) {
        CILQR_CUDA(cudaMemcpyAsync(data, d_out, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, st));
        CILQR_CUDA(cudaStreamSynchronize(st));
        tmp.release();
    }
    return 0;
}

int costmap_to_occupancy(const float *data, int64_t n, float data_min, float data_max, int8_t *occ, int mem, void *stream)
{
    if (!occ || !data || n < 0) return fail(CILQR_E_ARG, "bad argument");
    if (n == 0) return 0;
    if (int rc = need_device()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    DevBuf tmp;
    const float *d_in = data;
    int8_t *d_out = occ;
    if (mem == CILQR_MEM_HOST) {
        const size_t off = (size_t)n * sizeof(float);
        if (int rc = tmp.reserve(off + (size_t)n)) return rc;
        CILQR_CUDA(cudaMemcpyAsync(tmp.p, data, off, cudaMemcpyHostToDevice, st));
        d_in = tmp.as<float>();
        d_out = reinterpret_cast<int8_t *>(tmp.as<char>() + off);
    }
    to_occupancy_kernel<<<io_grid((n + 15) / 16), 256, 0, st>>>(d_in, n, data_min, data_max, d_out);
    CILQR_LAUNCH_CHECK();
    if (mem == CILQR_MEM_HOST) {
        CILQR_CUDA(cudaMemcpyAsync(occ, d_out, (size_t)n, cudaMemcpyDeviceToHost, st));
        CILQR_CUDA(cudaStreamSynchronize(st));
        tmp.release();
    }
    return 0;
}

// grid_map's index of a position along one axis, clamped into the map (boundPositionToRange + getIndexFromPosition,
// GridMapMath.cpp:241-264, :129-142), widened by one cell: a conservative superset of the PolygonIterator's submap.
static int clamp_index(double p, double pos, double len, double r, int size, int widen)
{
    double t = -(((p - 0.5 * len) - pos) / r);
    if (!(t == t)) return widen > 0 ? size - 1 : 0;
    t += widen;
    if (t < 0.0) return 0;
    if (t > size - 1) return size - 1;
    return (int)t;
}

int costmap_rasterize_boxes(const costmap_geom *geom, int P, const double *own_poses, int n_boxes, const double *boxes, float *layer, int mem,
                            void *stream)
{
    if (!geom || !own_poses || !layer || P < 0 || n_boxes < 0 || (n_boxes > 0 && !boxes)) return fail(CILQR_E_ARG, "bad argument");
    if (!(geom->resolution > 0.0) || !(geom->length_x > 0.0) || !(geom->length_y > 0.0)) return fail(CILQR_E_ARG, "bad costmap geometry");
    const int rows = (int)std::round(geom->length_x / geom->resolution), cols = (int)std::round(geom->length_y / geom->resolution); // GridMap.cpp:50-52
    if (rows < 1 || cols < 1) return fail(CILQR_E_ARG, "costmap geometry has no cells");
    if (P == 0 || n_boxes == 0) return 0;
    if (int rc = need_device()) return rc;
    const double r = geom->resolution, Lx = rows * r, Ly = cols * r;
    std::vector<RasterBox> hb;
    hb.reserve((size_t)P * n_boxes);
    for (int f = 0; f < P; ++f) {
        const double own_pos_x = own_poses[3 * f], own_pos_y = own_poses[3 * f + 1], own_yaw = own_poses[3 * f + 2];
        for (int v = 0; v < n_boxes; ++v) {
            const double *bx = boxes + ((size_t)f * n_boxes + v) * 5;
            // local_costmap.cpp:866-911, expression by expression (host libm, no FMA contraction in this translation unit's host code)
            const double posX = bx[0];
            const double posY = -bx[1];
            const double distance = std::sqrt(std::pow(posX - own_pos_x, 2) + std::pow(posY - own_pos_y, 2));
            if (!(distance <= 100.0))
                continue;
            const double sizeX = bx[3] + 0.2, sizeY = bx[4] + 0.2, yaw = -bx[2];
            const double half_sizeX = sizeX / 2.0, half_sizeY = sizeY / 2.0;
            const double cx[4] = {half_sizeX, half_sizeX, -half_sizeX, -half_sizeX}, cy[4] = {half_sizeY, -half_sizeY, -half_sizeY, half_sizeY};
            RasterBox b;
            double xmax = -INFINITY, xmin = INFINITY, ymax = -INFINITY, ymin = INFINITY;
            for (int k = 0; k < 4; ++k) {
                const double global_x = cos(yaw) * cx[k] - sin(yaw) * cy[k] + posX;
                const double global_y = sin(yaw) * cx[k] + cos(yaw) * cy[k] + posY;
                const double local_x = cos(own_yaw) * (global_x - own_pos_x) + sin(own_yaw) * (global_y - own_pos_y);
                const double local_y = -sin(own_yaw) * (global_x - own_pos_x) + cos(own_yaw) * (global_y - own_pos_y);
                b.vx[k] = local_x;
                b.vy[k] = local_y;
                xmax = std::fmax(xmax, local_x); xmin = std::fmin(xmin, local_x);
                ymax = std::fmax(ymax, local_y); ymin = std::fmin(ymin, local_y);
            }
            b.vx[4] = b.vx[0];
            b.vy[4] = b.vy[0];
            // larger position = smaller index (grid_map's axes point "up/left")
            b.i_lo = clamp_index(xmax, geom->pos_x, Lx, r, rows, -1);
            b.i_hi = clamp_index(xmin, geom->pos_x, Lx, r, rows, +1);
            b.j_lo = clamp_index(ymax, geom->pos_y, Ly, r, cols, -1);
            b.j_hi = clamp_index(ymin, geom->pos_y, Ly, r, cols, +1);
            b.frame = f;
            hb.push_back(b);
        }
    }
    if (hb.empty()) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cells = (size_t)P * rows * cols;
    DevBuf dboxes, dlayer;
    if (int rc = dboxes.reserve(hb.size() * sizeof(RasterBox))) return rc;
    CILQR_CUDA(cudaMemcpyAsync(dboxes.p, hb.data(), hb.size() * sizeof(RasterBox), cudaMemcpyHostToDevice, st));
    float *d_layer = layer;
    if (mem == CILQR_MEM_HOST) {
        if (int rc = dlayer.reserve(cells * sizeof(float))) { dboxes.release(); return rc; }
        CILQR_CUDA(cudaMemcpyAsync(dlayer.p, layer, cells * sizeof(float), cudaMemcpyHostToDevice, st));
        d_layer = dlayer.as<float>();
    }
    const double base_x = geom->pos_x + (0.5 * Lx - 0.5 * r), base_y = geom->pos_y + (0.5 * Ly - 0.5 * r);
    rasterize_boxes_kernel<<<(int)hb.size(), 256, 0, st>>>(dboxes.as<RasterBox>(), rows, cols, base_x, base_y, r, d_layer);
    CILQR_LAUNCH_CHECK();
    if (mem == CILQR_MEM_HOST)
        CILQR_CUDA(cudaMemcpyAsync(layer, d_layer, cells * sizeof(float), cudaMemcpyDeviceToHost, st));
    CILQR_CUDA(cudaStreamSynchronize(st)); // hb / dboxes are released below
    dboxes.release();
    dlayer.release();
    return 0;
}

} // extern "C"
