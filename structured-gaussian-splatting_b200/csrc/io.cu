// io.cu - device side of the PLY wire format.
//
// The reference stores a model as one binary little-endian PLY vertex element of 62 float properties per Gaussian at SH-3
// (x y z nx ny nz f_dc_0..2 f_rest_0..44 opacity scale_0..2 rot_0..3; scene/gaussian_model.py:187-218) and reads it back
// column by column through plyfile into numpy arrays, transposing the SH block twice (:224-266).  Here the payload goes to the
// GPU as it lies in the file (one pinned H2D copy) and ONE gather kernel per parameter tensor pulls its columns out of the
// 248-byte records - and the other way round for saving.  out[r, j] = in[r, col[j]]: rows are contiguous, so a warp's reads
// stay inside a few neighbouring records; the kernel runs at HBM speed (8 B per element moved).
#include "gsr_common.cuh"

namespace gsr {

constexpr int IO_MAX_COLS = 64;
struct ColumnMap { int col[IO_MAX_COLS]; };

__global__ void __launch_bounds__(256)
gather_columns_kernel(const float* __restrict__ in, int in_cols, float* __restrict__ out, int out_cols, long long rows,
                      const __grid_constant__ ColumnMap map)
{
    const long long total = rows * out_cols;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long r = i / out_cols;
        const int j = (int)(i - r * out_cols);
        const int c = map.col[j];
        out[i] = c >= 0 ? in[r * in_cols + c] : 0.f;          // negative column: constant 0 (the nx ny nz normals)
    }
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_gather_columns(const float* in, int in_cols, float* out, int out_cols, const int* col_of_out, long long rows,
                                  gsr_stream_t stream_)
{
    cudaStream_t s = (cudaStream_t)stream_;
    if (rows < 0 || in_cols < 0 || out_cols < 0 || out_cols > IO_MAX_COLS) return GSR_EINVAL;
    if (rows == 0 || out_cols == 0) return GSR_OK;
    if (!in || !out || !col_of_out) return GSR_EINVAL;
    ColumnMap map;
    for (int j = 0; j < out_cols; j++) {
        if (col_of_out[j] >= in_cols) return GSR_EINVAL;
        map.col[j] = col_of_out[j];
    }
    for (int j = out_cols; j < IO_MAX_COLS; j++) map.col[j] = -1;
    const long long total = rows * out_cols;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    GSR_LAUNCH(gather_columns_kernel, (unsigned)blocks, 256, 0, s, in, in_cols, out, out_cols, rows, map);
    return GSR_OK;
}
