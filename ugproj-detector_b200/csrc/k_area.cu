// Scan-size decimation: cv::resize(orig, scan.frame_size, INTER_AREA) at reference
// src/tracker/face_tracker.cc:112-114, the step right before detectFaces.
// Bit-exact with cv2 4.13 (oracle: orc_resize_area).  HBM-bound: cn bytes read per source pixel,
// cn bytes written per destination pixel; the taps of one output pixel are adjacent, so the reads
// of a warp are contiguous row segments.
#include <cmath>
#include <vector>

#include "kernels.cuh"

namespace ugd {

namespace {
void axis_taps(int ssize, int dsize, double scale, std::vector<int32_t>& first, std::vector<AreaTap>& taps) {
    first.assign((size_t)dsize + 1, 0);
    taps.clear();
    for (int dx = 0; dx < dsize; ++dx) {
        first[dx] = (int32_t)taps.size();
        const double fsx1 = dx * scale, fsx2 = fsx1 + scale;
        const double cell = std::min(scale, ssize - fsx1);
        int sx1 = (int)ceil(fsx1), sx2 = (int)floor(fsx2);
        sx2 = std::min(sx2, ssize - 1);
        sx1 = std::min(sx1, sx2);
        if (sx1 - fsx1 > 1e-3) taps.push_back({sx1 - 1, (float)((sx1 - fsx1) / cell)});
        for (int sx = sx1; sx < sx2; ++sx) taps.push_back({sx, (float)(1.0 / cell)});
        if (fsx2 - sx2 > 1e-3) taps.push_back({sx2, (float)(std::min(std::min(fsx2 - sx2, 1.0), cell) / cell)});
    }
    first[dsize] = (int32_t)taps.size();
}
}  // namespace

bool build_area_plan(int sw, int sh, int dw, int dh, AreaPlanHost& p) {
    if (dw <= 0 || dh <= 0 || dw > sw || dh > sh) return false;
    p.sw = sw; p.sh = sh; p.dw = dw; p.dh = dh;
    const double scx = (double)sw / dw, scy = (double)sh / dh;
    const int ix = (int)nearbyint(scx), iy = (int)nearbyint(scy);
    p.fast = fabs(scx - ix) < 2.220446049250313e-16 && fabs(scy - iy) < 2.220446049250313e-16;
    p.ix = ix; p.iy = iy;
    axis_taps(sw, dw, scx, p.xfirst, p.xtaps);
    axis_taps(sh, dh, scy, p.yfirst, p.ytaps);
    return true;
}

template <int CN>
__global__ void __launch_bounds__(256) k_area(AreaArgs a) {
    const int x = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y, frame = blockIdx.z;
    if (x >= a.dw) return;
    const uint8_t* src = a.src + (size_t)frame * a.src_frame_stride;
    uint8_t* dst = a.dst + (size_t)frame * a.dst_frame_stride + ((size_t)y * a.dw + x) * CN;
    if (a.fast) {
        int sum[CN];
#pragma unroll
        for (int c = 0; c < CN; ++c) sum[c] = 0;
        for (int v = 0; v < a.iy; ++v) {
            const uint8_t* row = src + (size_t)(y * a.iy + v) * a.src_row_stride + (size_t)x * a.ix * CN;
            for (int u = 0; u < a.ix; ++u)
#pragma unroll
                for (int c = 0; c < CN; ++c) sum[c] += row[u * CN + c];
        }
        const float inv = __fdiv_rn(1.f, (float)(a.ix * a.iy));
#pragma unroll
        for (int c = 0; c < CN; ++c) {
            const int r = (a.ix == 2 && a.iy == 2) ? (sum[c] + 2) >> 2 : __float2int_rn(__fmul_rn((float)sum[c], inv));
            dst[c] = (uint8_t)min(r, 255);
        }
        return;
    }
    const int x0 = a.xfirst[x], x1 = a.xfirst[x + 1], y0 = a.yfirst[y], y1 = a.yfirst[y + 1];
    float sum[CN];
#pragma unroll
    for (int c = 0; c < CN; ++c) sum[c] = 0.f;
    for (int j = y0; j < y1; ++j) {
        const AreaTap ty = a.ytaps[j];
        const uint8_t* row = src + (size_t)ty.si * a.src_row_stride;
        float buf[CN];
#pragma unroll
        for (int c = 0; c < CN; ++c) buf[c] = 0.f;
        for (int k = x0; k < x1; ++k) {
            const AreaTap tx = a.xtaps[k];
#pragma unroll
            for (int c = 0; c < CN; ++c) buf[c] = __fadd_rn(buf[c], __fmul_rn((float)row[tx.si * CN + c], tx.a));
        }
#pragma unroll
        for (int c = 0; c < CN; ++c) sum[c] = j == y0 ? __fmul_rn(ty.a, buf[c]) : __fadd_rn(sum[c], __fmul_rn(ty.a, buf[c]));
    }
#pragma unroll
    for (int c = 0; c < CN; ++c) dst[c] = (uint8_t)min(max(__float2int_rn(sum[c]), 0), 255);
}

// Face patches (second half): what `cv::Mat img_face(curr_frame, *it)` + the copy in Face's copy
// constructor produce on the host (reference face_tracker.cc:296-302, structure.h:137-146), cut out of device-resident
// frames.  One block per patch; a row of a patch is contiguous in the frame, so the copy is row-wise and coalesced.
__global__ void __launch_bounds__(256) k_crop(const uint8_t* __restrict__ frames, size_t frame_stride, size_t row_stride, int channels,
                                              const CropItem* __restrict__ items, uint8_t* __restrict__ out) {
    const CropItem it = items[blockIdx.x];
    const uint8_t* src = frames + (size_t)it.frame * frame_stride + (size_t)it.y * row_stride + (size_t)it.x * channels;
    uint8_t* dst = out + it.offset;
    const int row_bytes = it.w * channels;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int y = warp; y < it.h; y += 8) {
        const uint8_t* s = src + (size_t)y * row_stride;
        uint8_t* d = dst + (size_t)y * row_bytes;
        for (int i = lane; i < row_bytes; i += 32) d[i] = s[i];
    }
}

void launch_crop(const uint8_t* frames, size_t frame_stride, size_t row_stride, int channels, const CropItem* items, int nitems,
                 uint8_t* out, cudaStream_t st) {
    if (nitems) k_crop<<<nitems, 256, 0, st>>>(frames, frame_stride, row_stride, channels, items, out);
}

void launch_area(const AreaArgs& a, int channels, int nframes, cudaStream_t st) {
    const dim3 grid((a.dw + 255) / 256, a.dh, nframes);
    if (channels == 3) k_area<3><<<grid, 256, 0, st>>>(a);
    else k_area<1><<<grid, 256, 0, st>>>(a);
}

}  // namespace ugd
