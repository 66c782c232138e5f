// Densify / prune surgery on the per-splat tensors and their Adam state.
// Replaces the per-tensor Python loops of reference gslam/pruning.py:10-53 (`value[keep_mask]` for
// 6 parameters + 12 Adam states) and gslam/insertion.py:27-62 (`torch.cat` with the new rows /
// zeros): ONE launch moves the kept rows of every tensor to their compacted position and appends
// the new rows, so the flat [15 N] parameter / gradient / Adam-state layout survives the resize.
#include "common.cuh"

#define SURGERY_MAX_TENSORS 24

struct RowTensor {
    const float* src;     // [n_old, width]
    float* dst;           // [n_keep + n_add, width]
    const float* add;     // [n_add, width] appended rows, NULL = zeros (fresh Adam state)
    int width;
};
struct RowPack {
    RowTensor t[SURGERY_MAX_TENSORS];
};

// grid.y = tensor; x covers max(n_old, n_add) rows
__global__ void __launch_bounds__(256)
rows_resize_kernel(const RowPack pk, int64_t n_old, const int64_t* __restrict__ cum, int64_t n_keep,
                   int64_t n_add) {
    const RowTensor rt = pk.t[blockIdx.y];
    const int w = rt.width;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_old || i < n_add;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (i < n_old) {
            // cum = inclusive scan of the keep mask: row i is kept iff cum steps at i
            const int64_t c = cum ? cum[i] : i + 1;
            const bool keep = cum ? (c != (i ? cum[i - 1] : 0)) : true;
            if (keep)
                for (int j = 0; j < w; ++j) rt.dst[(c - 1) * w + j] = rt.src[i * w + j];
        }
        if (i < n_add)
            for (int j = 0; j < w; ++j) rt.dst[(n_keep + i) * w + j] = rt.add ? rt.add[i * w + j] : 0.f;
    }
}

GSB_API int gsb_rows_resize(const gsb_row_tensor* tensors_host, int n_tensors, int64_t n_old,
                            const int64_t* keep_cum, int64_t n_keep, int64_t n_add, void* stream) {
    GSB_CHECK_ARG(n_tensors >= 0 && n_tensors <= SURGERY_MAX_TENSORS && n_old >= 0 && n_keep >= 0 && n_add >= 0);
    GSB_CHECK_ARG(keep_cum || n_keep == n_old);
    if (n_tensors == 0 || (n_old == 0 && n_add == 0)) return 0;
    GSB_CHECK_ARG(tensors_host);
    RowPack pk;
    for (int k = 0; k < n_tensors; ++k) {
        const gsb_row_tensor& h = tensors_host[k];
        GSB_CHECK_ARG(h.width > 0 && h.dst && (h.src || n_old == 0));
        pk.t[k].src = h.src; pk.t[k].dst = h.dst; pk.t[k].add = h.add; pk.t[k].width = h.width;
    }
    const int64_t rows = n_old > n_add ? n_old : n_add;
    int64_t bx = (rows + 255) / 256;
    if (bx > 148 * 16) bx = 148 * 16;
    dim3 grid((unsigned)bx, (unsigned)n_tensors);
    rows_resize_kernel<<<grid, 256, 0, gsb_stream(stream)>>>(pk, n_old, keep_cum, n_keep, n_add);
    GSB_LAUNCH_CHECK();
    return 0;
}
