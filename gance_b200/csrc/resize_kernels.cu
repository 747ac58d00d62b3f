// Bicubic resize of uint8 NHWC frames (the 1024^2 -> 2160^2 output scaling of the reference's
// `scale_square_source_duplicate`, /root/reference/gance/image_sources/video_common.py:399-455, which calls
// cv2.resize(..., interpolation=cv2.INTER_CUBIC)). Bit-exact restatement of OpenCV's own 8-bit path
// (modules/imgproc/src/resize.cpp; the algorithm is spelled out in oracle/resize_ref.py): 11-bit fixed-point
// coefficient tables (built by the caller: gance_b200/image_sources/video_common.py:cubic_tables), integer
// horizontal pass, vertical pass in float with separately rounded multiply/add for the first floor(n/8)*8 elements of
// a destination row and in integers for the tail.
//
// Algorithmic bytes: 3.1 MB read + 14 MB written per 1024^2 -> 2160^2 frame. One CTA per 16 x 128-pixel destination
// tile: the source window (replicated border applied while loading; whole 32-bit words on interior tiles) goes to shared
// memory once as float, the horizontal pass writes exact float rows to shared memory, the vertical pass reads 4 of them
// per output element; each thread produces 4 consecutive bytes and stores them as one 32-bit word when the row pitch
// allows it. Measured 26 us per frame (0.66 TB/s): the kernel is instruction-issue-bound, not HBM-bound - reproducing
// OpenCV's float rounding costs ~12 instructions per output byte.
#include "common.cuh"

namespace gance {

constexpr int kRzTileW = 128, kRzTileH = 16, kRzRowElems = kRzTileW * 3, kRzThreads = kRzRowElems;

struct ResizeParams {
    const uint8_t* src;  // [B, sh, sw, 3]
    uint8_t* dst;        // [B, dh, dw, 3]
    const int* xofs;     // [dw] source index of tap 1
    const short* xcoef;  // [dw, 4]
    const int* yofs;     // [dh]
    const short* ycoef;  // [dh, 4]
    int batch, sh, sw, dh, dw;
    int win_w, win_h;    // capacity of the shared-memory source window (pixels, rows)
};

// 384 threads = one per element (pixel x channel) of a 128-pixel destination row segment.
//  1. source window -> shared memory as FLOAT (each byte converted once; border replicated while loading)
//  2. horizontal pass: thread e owns destination element e of every window row: 4 LDS + 4 FP32 FMAs. Exact: products are
//     below 2^23 and every partial sum below 2^24 (sum |coef| < 2700), so the float equals OpenCV's int32 row value.
//  3. vertical pass: thread (g, q) owns 4 consecutive elements of rows q, q+4, q+8, q+12: 4 LDS.128 per row, OpenCV's
//     separately rounded multiply/add chain, round-half-even, saturate, one 32-bit store.
__global__ void __launch_bounds__(kRzThreads)
resize_cubic_u8_kernel(const ResizeParams p) {
    extern __shared__ float smem_f[];
    const int x0 = blockIdx.x * kRzTileW, y0 = blockIdx.y * kRzTileH, b = blockIdx.z;
    const int tw = min(kRzTileW, p.dw - x0), th = min(kRzTileH, p.dh - y0);
    const int sx0 = __ldg(p.xofs + x0) - 1, sx1 = __ldg(p.xofs + x0 + tw - 1) + 2;
    const int sy0 = __ldg(p.yofs + y0) - 1, sy1 = __ldg(p.yofs + y0 + th - 1) + 2;
    const int ww3 = (sx1 - sx0 + 1) * 3, wh = sy1 - sy0 + 1;  // <= win_w * 3, win_h (the host sizes the window over all tiles)
    const int row_stride = p.win_w * 3 + 1;
    float* hbuf = smem_f;                                      // [win_h][kRzRowElems]
    float* window = smem_f + (size_t)p.win_h * kRzRowElems;    // [win_h][row_stride]
    const uint8_t* src_b = p.src + (size_t)b * p.sh * p.sw * 3;
    const int t = threadIdx.x;

    // Interior tiles (no border replication, 4-byte aligned source) read whole 32-bit words: one LDG per 4 source bytes
    // instead of one per byte (the per-byte loop was 31 % of the kernel's instructions). Border tiles take the byte path.
    const bool words_ok = sx0 >= 0 && sx1 < p.sw && sy0 >= 0 && sy1 < p.sh &&
                          ((reinterpret_cast<uintptr_t>(p.src) | ((size_t)p.batch * p.sh * p.sw * 3)) & 3) == 0;
    if (words_ok) {
        const int words = (ww3 + 6) >> 2;  // words a row can touch, whatever its alignment
        const uint32_t magic = 0xFFFFFFFFu / (uint32_t)words + 1u;
        for (int idx = t; idx < wh * words; idx += kRzThreads) {
            const int r = (int)__umulhi((uint32_t)idx, magic), j = idx - r * words;
            const uint8_t* row = src_b + ((size_t)(sy0 + r) * p.sw + sx0) * 3;
            const int misalign = (int)(reinterpret_cast<uintptr_t>(row) & 3);
            const int c0 = 4 * j - misalign;  // window column of the word's first byte
            if (c0 < ww3) {
                const uint32_t word = __ldg(reinterpret_cast<const uint32_t*>(row - misalign) + j);
                float* dst = window + r * row_stride + c0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (c0 + k >= 0 && c0 + k < ww3) dst[k] = (float)((word >> (8 * k)) & 0xffu);
            }
        }
    } else {
        const uint32_t magic = 0xFFFFFFFFu / (uint32_t)ww3 + 1u;   // idx / ww3 == umulhi(idx, magic) for idx < 2^16
        for (int idx = t; idx < wh * ww3; idx += kRzThreads) {
            const int r = (int)__umulhi((uint32_t)idx, magic), c = idx - r * ww3;
            const int px = (c * 43691) >> 17, ch = c - px * 3;     // c / 3 for c < 98304
            const int sy = min(max(sy0 + r, 0), p.sh - 1), sx = min(max(sx0 + px, 0), p.sw - 1);
            window[r * row_stride + c] = (float)__ldg(src_b + ((size_t)sy * p.sw + sx) * 3 + ch);
        }
    }
    __syncthreads();
    {
        const int dxl = (t * 43691) >> 17;
        if (dxl < tw) {
            const int dx = x0 + dxl;
            const float* w = window + (__ldg(p.xofs + dx) - 1 - sx0) * 3 + (t - dxl * 3);
            const short4 a = __ldg(reinterpret_cast<const short4*>(p.xcoef) + dx);
            const float a0 = (float)a.x, a1 = (float)a.y, a2 = (float)a.z, a3 = (float)a.w;
            for (int r = 0; r < wh; ++r, w += row_stride)
                hbuf[r * kRzRowElems + t] = fmaf(w[0], a0, fmaf(w[3], a1, fmaf(w[6], a2, w[9] * a3)));
        }
    }
    __syncthreads();
    const int n = p.dw * 3;                 // elements of a destination row
    const int body = (n >> 3) << 3;         // handled by the float vector body in OpenCV
    const bool word_store = (n & 3) == 0;   // x0 * 3 is a multiple of 4 (kRzTileW = 128)
    const int g = t % (kRzRowElems / 4), q = t / (kRzRowElems / 4);
    const int e0 = g * 4;
    if (e0 < tw * 3) {
        for (int ry = q; ry < th; ry += 4) {
            const int dy = y0 + ry;
            const float* col = hbuf + (__ldg(p.yofs + dy) - 1 - sy0) * kRzRowElems + e0;
            const short4 bq = __ldg(reinterpret_cast<const short4*>(p.ycoef) + dy);
            const float scale = 1.f / (2048.f * 2048.f);
            const float b0 = (float)bq.x * scale, b1 = (float)bq.y * scale, b2 = (float)bq.z * scale, b3 = (float)bq.w * scale;
            const float4 s0 = *reinterpret_cast<const float4*>(col), s1 = *reinterpret_cast<const float4*>(col + kRzRowElems),
                         s2 = *reinterpret_cast<const float4*>(col + 2 * kRzRowElems),
                         s3 = *reinterpret_cast<const float4*>(col + 3 * kRzRowElems);
            const float v0[4] = {s0.x, s0.y, s0.z, s0.w}, v1[4] = {s1.x, s1.y, s1.z, s1.w}, v2[4] = {s2.x, s2.y, s2.z, s2.w},
                        v3[4] = {s3.x, s3.y, s3.z, s3.w};
            uint32_t bytes[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (x0 * 3 + e0 + k < body) {
                    float x = __fmul_rn(v3[k], b3);
                    x = __fadd_rn(__fmul_rn(v2[k], b2), x);
                    x = __fadd_rn(__fmul_rn(v1[k], b1), x);
                    x = __fadd_rn(__fmul_rn(v0[k], b0), x);
                    // round-half-even + saturate to [0,255] in one instruction (was F2I + two IMNMX)
                    asm("cvt.rni.sat.u8.f32 %0, %1;" : "=r"(bytes[k]) : "f"(x));
                } else {  // OpenCV's scalar tail: integers
                    const int value = ((int)v0[k] * (int)bq.x + (int)v1[k] * (int)bq.y + (int)v2[k] * (int)bq.z +
                                       (int)v3[k] * (int)bq.w + (1 << 21)) >> 22;
                    bytes[k] = (uint32_t)min(max(value, 0), 255);
                }
            }
            uint8_t* out = p.dst + ((size_t)b * p.dh + dy) * n + x0 * 3 + e0;
            if (word_store && e0 + 3 < tw * 3) {
                *reinterpret_cast<uint32_t*>(out) = bytes[0] | (bytes[1] << 8) | (bytes[2] << 16) | (bytes[3] << 24);
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (e0 + k < tw * 3) out[k] = (uint8_t)bytes[k];
            }
        }
    }
}

}  // namespace gance

using namespace gance;

extern "C" int gance_resize_cubic_u8(const uint8_t* src, int batch, int src_h, int src_w, uint8_t* dst, int dst_h, int dst_w,
                                     const int* xofs, const short* xcoef, const int* yofs, const short* ycoef, int win_w,
                                     int win_h, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(src && dst && xofs && xcoef && yofs && ycoef, GANCE_E_NULL, "gance_resize_cubic_u8: NULL pointer");
    GANCE_REQUIRE(batch > 0 && src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0 && batch <= 65535, GANCE_E_SHAPE,
                  "gance_resize_cubic_u8: shape B=%d %dx%d -> %dx%d", batch, src_h, src_w, dst_h, dst_w);
    GANCE_REQUIRE((reinterpret_cast<uintptr_t>(xcoef) & 7u) == 0 && (reinterpret_cast<uintptr_t>(ycoef) & 7u) == 0 &&
                      (reinterpret_cast<uintptr_t>(dst) & 3u) == 0,
                  GANCE_E_ALIGN, "gance_resize_cubic_u8: coefficient tables must be 8-byte aligned, dst 4-byte aligned");
    const size_t smem = ((size_t)win_h * kRzRowElems + (size_t)win_h * (win_w * 3 + 1)) * sizeof(float);
    GANCE_REQUIRE(win_w >= 4 && win_h >= 4 && smem <= 200 * 1024 && (size_t)win_h * win_w * 3 < 65536, GANCE_E_SHAPE,
                  "gance_resize_cubic_u8: source window %d x %d needs %zu B of shared memory (limit 200 KB): the scale "
                  "factor is too small for the 16 x 128 destination tile", win_w, win_h, smem);
    static std::atomic<size_t> configured{0};
    if (smem > 48 * 1024 && smem > configured.load()) {
        cudaError_t err = cudaFuncSetAttribute(resize_cubic_u8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return fail(-(int)err, "cudaFuncSetAttribute(resize_cubic_u8_kernel): %s", cudaGetErrorString(err));
        configured.store(smem);
    }
    ResizeParams p;
    p.src = src; p.dst = dst; p.xofs = xofs; p.xcoef = xcoef; p.yofs = yofs; p.ycoef = ycoef;
    p.batch = batch; p.sh = src_h; p.sw = src_w; p.dh = dst_h; p.dw = dst_w; p.win_w = win_w; p.win_h = win_h;
    dim3 grid((dst_w + kRzTileW - 1) / kRzTileW, (dst_h + kRzTileH - 1) / kRzTileH, batch);
    resize_cubic_u8_kernel<<<grid, kRzThreads, smem, (cudaStream_t)stream>>>(p);
    return check_launch("resize_cubic_u8_kernel");
}
