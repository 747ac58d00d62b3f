// Up-convolution at its ALGORITHMIC cost: transposed 3x3 convolution as a 4-phase implicit GEMM on tcgen05
// (9 * Cin * Cout MACs per input pixel - the folded form of modconv_tc.cu executes 36), followed by one
// bandwidth-bound pass that applies the [1,3,3,1] FIR + noise + bias + leaky-ReLU + next-layer style.
//
// Reference: upsample_conv_2d (dnnlib/tflib/ops/upfirdn_2d.py) = conv2d_transpose(stride 2, VALID) -> upfirdn_2d, as
// restated in oracle/stylegan2_ref.py:109-123. With t = conv2d_transpose(x, flip(w)) of size (2H+1)^2:
//     t[2m+a, 2n+b] = sum over (dy,dx) in {0,1}^2 of  x[m-dy, n-dx] * w[iy(dy,a), ix(dx,b)],
//     iy(0,0) = 2, iy(0,1) = 1, iy(1,0) = 0, (1,1) absent            (x[-1] = x[H] = 0)
// i.e. the input shifted by (dy,dx) feeds 4 / 2 / 2 / 1 of the four output phases: exactly 9 (tap, phase) pairs.
// Pass 1 (this kernel) computes all phases of "cell" (m,n), m in [0,H], n in [0,W], as ONE accumulator row:
//     columns [P01 | P00 | P10 | P11] x 64 channels  (this order makes the phase sets of every shift contiguous:
//     shift (0,0) -> blocks 0-3, (1,0) -> 0-1, (0,1) -> 1-2, (1,1) -> 1)
// with one tcgen05.mma of N = 256 / 128 / 128 / 64 per shift and K step, multiplies by the demodulation
// coefficient and stores t as 16-bit NHWC [B, 2H+2, 2W+2, Cout] (row/column 2H+1 are the exact zeros the FIR's
// right padding needs). Pass 2 (upfir_act_kernel): out[y,x] = lrelu(sum_ab f[a] f[b] t[y-1+a, x-1+b] + noise +
// bias) * sqrt(2) * s_next, f = [1,3,3,1]/4.
#include <mutex>

#include "tc_ptx.cuh"

namespace gance {
namespace tc {

constexpr int kPhaseCC = 64;       // output channels per accumulator tile (x 4 phases = 256 columns)
constexpr int kPhaseK = 64;        // input channels per pipeline stage (one 128-byte swizzled row)
constexpr int kPhaseStages = 4;
constexpr int kPhaseABytes = kTileM * kPhaseK * 2;        // 16 KB
constexpr int kPhaseBBlock = kPhaseCC * kPhaseK * 2;      // 8 KB: one phase block of the B operand
constexpr int kPhaseStageBytes = kPhaseABytes + 4 * kPhaseBBlock;
constexpr int kPhaseBarrierOffset = kPhaseStages * kPhaseStageBytes;
constexpr int kPhaseCacheOffset = kPhaseBarrierOffset + (2 * kPhaseStages + 4) * 8 + 16;
constexpr int kPhaseSmemBytes = kPhaseCacheOffset + kEpilogueGroups * kPhaseCC * 4 + 1024;

struct PhaseParams {
    int batch, h_cells, w_cells;    // cells = input pixels + 1 per axis
    int c_in, c_out, k_chunks, n_tiles;
    int tile_w, tile_h, tiles_w, tiles_h;  // tile_w * tile_h <= 128 cells, any shape (chosen to minimise the tile count)
    const float* demod;  // [B, c_out]
    void* t;             // [B, 2*h_cells, 2*w_cells, c_out]
};

__device__ __forceinline__ void phase_decode(const PhaseParams& p, int tile, int& nt, int& b, int& h0, int& w0) {
    nt = tile % p.n_tiles;
    int m = tile / p.n_tiles;
    w0 = (m % p.tiles_w) * p.tile_w;
    m /= p.tiles_w;
    h0 = (m % p.tiles_h) * p.tile_h;
    b = m / p.tiles_h;
}

template <typename T>
__global__ void __launch_bounds__(kThreads, 1)
upconv_phase_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                    const PhaseParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kPhaseBarrierOffset);
    uint64_t* empty_bar = full_bar + kPhaseStages;
    uint64_t* tmem_full_bar = empty_bar + kPhaseStages;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
    float* demod_cache = reinterpret_cast<float*>(smem + kPhaseCacheOffset);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < kPhaseStages; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_base_slot, 512);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);
    const int total_tiles = p.batch * p.tiles_h * p.tiles_w * p.n_tiles;

    if (warp == 0) {
        // ===== TMA producer: per (shift, K chunk) one activation box + 4/2/2/1 weight blocks =====
        const bool leader = elect_one_sync();
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            int nt, b, h0, w0;
            phase_decode(p, tile, nt, b, h0, w0);
#pragma unroll
            for (int sh = 0; sh < 4; ++sh) {
                const int dy = sh & 1, dx = sh >> 1;
                const int nblk = sh == 0 ? 4 : (sh == 3 ? 1 : 2);
                // rows of the phase weight matrix: shift bases 0, 4C, 6C, 8C; inside a shift (chunk, block, channel)
                const int row0 = (sh == 0 ? 0 : sh == 1 ? 4 : sh == 2 ? 6 : 8) * p.c_out + nt * nblk * kPhaseCC;
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    mbar_wait(empty_bar + stage, phase ^ 1u);
                    uint8_t* a_dst = smem + stage * kPhaseStageBytes;
                    if (leader) {
                        mbar_expect_tx(full_bar + stage, (uint32_t)(p.tile_w * p.tile_h * kPhaseK * 2 + nblk * kPhaseBBlock));
                        tma_load_4d(a_dst, &tmap_x, full_bar + stage, kc * kPhaseK, w0 - dx, h0 - dy, b);
                        for (int j = 0; j < nblk; ++j)
                            tma_load_2d(a_dst + kPhaseABytes + j * kPhaseBBlock, &tmap_w, full_bar + stage, kc * kPhaseK,
                                        row0 + j * kPhaseCC);
                    }
                    __syncwarp();
                    if (++stage == kPhaseStages) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const bool leader = elect_one_sync();
        constexpr bool kHalf = !std::is_same<T, __nv_bfloat16>::value;
        constexpr uint32_t idesc256 = make_instr_desc(kTileM, 256, kHalf), idesc128 = make_instr_desc(kTileM, 128, kHalf),
                           idesc64 = make_instr_desc(kTileM, 64, kHalf);
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        const uint32_t smem_base = smem_u32(smem);
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
            tcgen05_fence_after();
#pragma unroll
            for (int sh = 0; sh < 4; ++sh) {
                const uint32_t idesc = sh == 0 ? idesc256 : (sh == 3 ? idesc64 : idesc128);
                const int first_block = sh == 0 ? 0 : (sh == 1 ? 0 : 1);
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256 + first_block * kPhaseCC);
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    mbar_wait(full_bar + stage, phase);
                    tcgen05_fence_after();
                    const uint64_t a_desc = make_smem_desc<128>(smem_base + (uint32_t)(stage * kPhaseStageBytes));
                    const uint64_t b_desc = a_desc + (uint64_t)(kPhaseABytes >> 4);
                    if (leader) {
#pragma unroll
                        for (int k = 0; k < kPhaseK / kUmmaK; ++k)
                            umma_bf16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc,
                                      (sh | kc | k) != 0 ? 1u : 0u);  // shift (0,0) touches every column first
                        umma_commit(empty_bar + stage);
                        if (sh == 3 && kc == p.k_chunks - 1) umma_commit(tmem_full_bar + acc);
                    }
                    __syncwarp();
                    if (++stage == kPhaseStages) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1u;
        }
    } else {
        // ===== epilogue: t[2m+a, 2n+b, co] = acc * demod[b, co], one cell per thread =====
        const int group = (warp - 2) >> 2, quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const int wl = row % p.tile_w, hl = row / p.tile_w;
        const int t_h = 2 * p.h_cells, t_w = 2 * p.w_cells;
        float* dcache = demod_cache + group * kPhaseCC;
        const uint32_t dcache_addr = smem_u32(dcache);
        const int tid = quarter * 32 + lane;
        int cached_key = -1;
        uint32_t full_phase = 0;
        for (int tile = blockIdx.x + group * gridDim.x; tile < total_tiles; tile += kEpilogueGroups * gridDim.x) {
            int nt, b, h0, w0;
            phase_decode(p, tile, nt, b, h0, w0);
            const int m = h0 + hl, n = w0 + wl;
            const bool valid = hl < p.tile_h && m < p.h_cells && n < p.w_cells;
            const int key = b * p.n_tiles + nt;
            if (key != cached_key) {  // group-uniform
                asm volatile("bar.sync %0, 128;" ::"r"(group + 1) : "memory");
                if (tid < kPhaseCC) dcache[tid] = __ldg(p.demod + (size_t)b * p.c_out + nt * kPhaseCC + tid);
                asm volatile("bar.sync %0, 128;" ::"r"(group + 1) : "memory");
                cached_key = key;
            }
            mbar_wait(tmem_full_bar + group, full_phase);
            full_phase ^= 1u;
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(group * 256);
#pragma unroll 1
            for (int c0 = 0; c0 < 256; c0 += 32) {
                uint32_t r[32];
                tmem_ld_32x32(taddr + (uint32_t)c0, r);
                const int block = c0 >> 6, ch0 = c0 & 63;
                const int a = block >> 1, bb = (block == 0 || block == 3) ? 1 : 0;  // blocks: P01, P00, P10, P11
                T* dst = reinterpret_cast<T*>(p.t) +
                         (((size_t)b * t_h + (2 * m + a)) * t_w + (2 * n + bb)) * p.c_out + nt * kPhaseCC + ch0;
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        float v[8];
#pragma unroll
                        for (int q4 = 0; q4 < 2; ++q4) {
                            float4 d;
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                         : "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w)
                                         : "r"(dcache_addr + (uint32_t)(ch0 + g * 8 + q4 * 4) * 4u));
                            v[q4 * 4 + 0] = __uint_as_float(r[g * 8 + q4 * 4 + 0]) * d.x;
                            v[q4 * 4 + 1] = __uint_as_float(r[g * 8 + q4 * 4 + 1]) * d.y;
                            v[q4 * 4 + 2] = __uint_as_float(r[g * 8 + q4 * 4 + 2]) * d.z;
                            v[q4 * 4 + 3] = __uint_as_float(r[g * 8 + q4 * 4 + 3]) * d.w;
                        }
                        Vec<T> packed;
                        packed.from_float(v);
                        packed.store(dst + g * 8);
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty_bar + group);
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---- pass 2: [1,3,3,1] FIR over t + noise + bias + leaky-ReLU + next-layer style ---------------------------------------
// HBM-bound by construction: t is read once, out written once. Persistent CTAs (two per SM); one producer thread streams
// (8+3) x (32+3)-pixel x 64-channel tiles of t into a 2-stage shared-memory ring with TMA (coordinate -1 = the FIR's zero
// padding, filled by the TMA unit). Each of the 256 consumer threads owns 4 channels x 2 adjacent output columns and
// walks DOWN the tile: one t row = 5 LDS.64 + one horizontal filter (both x phases), a 4-row sliding window of those
// gives one output row per step - every t row is converted and filtered horizontally exactly once per thread
// (16 instructions per output element; the first version recomputed it per 2x2 block: 27).
constexpr int kFirTileW = 32, kFirTileH = 8, kFirC = 64, kFirStages = 2, kFirConsumers = 256;
constexpr int kFirTW = kFirTileW + 3, kFirTH = kFirTileH + 3;
constexpr int kFirStageBytes = ((kFirTW * kFirTH * kFirC * 2) + 1023) / 1024 * 1024;
constexpr int kFirSmemBytes = kFirStages * kFirStageBytes + 2 * kFirStages * 8 + 1024;

template <typename T>
__device__ __forceinline__ void unpack4(uint2 raw, float (&v)[4]);
template <>
__device__ __forceinline__ void unpack4<__half>(uint2 raw, float (&v)[4]) {
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
template <>
__device__ __forceinline__ void unpack4<__nv_bfloat16>(uint2 raw, float (&v)[4]) {
    v[0] = __uint_as_float(raw.x << 16); v[1] = __uint_as_float(raw.x & 0xffff0000u);
    v[2] = __uint_as_float(raw.y << 16); v[3] = __uint_as_float(raw.y & 0xffff0000u);
}
__device__ __forceinline__ uint2 pack4(const float (&v)[4], __half*) {
    const __half2 a = __floats2half2_rn(v[0], v[1]), b = __floats2half2_rn(v[2], v[3]);
    return make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
__device__ __forceinline__ uint2 pack4(const float (&v)[4], __nv_bfloat16*) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    return make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}

template <typename T>
__global__ void __launch_bounds__(kFirConsumers + 32, 2)
upfir_act_kernel(const __grid_constant__ CUtensorMap tmap_t, int batch, int h_out, int w_out, int channels,
                 const float* __restrict__ noise, long long noise_batch_stride, float noise_strength,
                 const float* __restrict__ bias, const float* __restrict__ next_styles, T* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kFirStages * kFirStageBytes);
    uint64_t* empty_bar = full_bar + kFirStages;
    const int tiles_w = (w_out + kFirTileW - 1) / kFirTileW, tiles_h = (h_out + kFirTileH - 1) / kFirTileH;
    const int chunks = channels / kFirC;
    const int total_tiles = batch * tiles_h * tiles_w * chunks;
    if (threadIdx.x == 0) {
        prefetch_tmap(&tmap_t);
        for (int i = 0; i < kFirStages; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, kFirConsumers / 32);
        }
        fence_barrier_init();
    }
    __syncthreads();
    auto decode = [&](int tile, int& cc, int& b, int& y0, int& x0) {
        cc = tile % chunks;
        int m = tile / chunks;
        x0 = (m % tiles_w) * kFirTileW;
        m /= tiles_w;
        y0 = (m % tiles_h) * kFirTileH;
        b = m / tiles_h;
    };
    if (threadIdx.x >= kFirConsumers) {
        if (threadIdx.x == kFirConsumers) {  // producer
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                int cc, b, y0, x0;
                decode(tile, cc, b, y0, x0);
                mbar_wait(empty_bar + stage, phase ^ 1u);
                mbar_expect_tx(full_bar + stage, kFirTW * kFirTH * kFirC * 2);
                tma_load_4d(smem + stage * kFirStageBytes, &tmap_t, full_bar + stage, cc * kFirC, x0 - 1, y0 - 1, b);
                if (++stage == kFirStages) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
        }
        return;
    }
    const int tid = threadIdx.x, lane = tid & 31;
    const int quad = tid & 15, bj = tid >> 4;  // 4 channels, output columns 2*bj and 2*bj + 1 of the tile
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int cc, b, y0, x0;
        decode(tile, cc, b, y0, x0);
        const int c0 = cc * kFirC + quad * 4;
        const int x = x0 + 2 * bj;
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + c0));
        float4 s4 = make_float4(1.f, 1.f, 1.f, 1.f);
        if (next_styles) s4 = __ldg(reinterpret_cast<const float4*>(next_styles + (size_t)b * channels + c0));
        const float bs[4] = {b4.x, b4.y, b4.z, b4.w};
        const float st[4] = {s4.x * kLreluGain, s4.y * kLreluGain, s4.z * kLreluGain, s4.w * kLreluGain};
        const float* noise_b = noise ? noise + (size_t)b * noise_batch_stride : nullptr;
        T* out_b = out + ((size_t)b * h_out * w_out) * channels + c0;
        mbar_wait(full_bar + stage, phase);
        const uint32_t column_addr = smem_u32(smem + stage * kFirStageBytes) + (uint32_t)((2 * bj * kFirC + quad * 4) * 2);
        float window[4][2][4];  // horizontally filtered t rows r-3 .. r (slot = r & 3), both x phases, 4 channels
#pragma unroll
        for (int r = 0; r < kFirTH; ++r) {
            float v[5][4];
#pragma unroll
            for (int c = 0; c < 5; ++c) {
                uint2 raw;
                asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];"
                             : "=r"(raw.x), "=r"(raw.y)
                             : "r"(column_addr + (uint32_t)((r * kFirTW + c) * kFirC * 2)));
                unpack4<T>(raw, v[c]);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                window[r & 3][0][q] = 0.25f * (v[0][q] + v[3][q]) + 0.75f * (v[1][q] + v[2][q]);
                window[r & 3][1][q] = 0.25f * (v[1][q] + v[4][q]) + 0.75f * (v[2][q] + v[3][q]);
            }
            if (r >= 3) {  // output row y0 + r - 3 = t rows r-3 .. r with [1,3,3,1]/4
                const int y = y0 + r - 3;
                if (y < h_out && x < w_out) {
                    float2 nz = make_float2(0.f, 0.f);
                    if (noise_b) nz = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)y * w_out + x));
#pragma unroll
                    for (int px = 0; px < 2; ++px) {
                        const float nzv = (px ? nz.y : nz.x) * noise_strength;
                        float o[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float value = 0.25f * (window[(r - 3) & 3][px][q] + window[r & 3][px][q]) +
                                          0.75f * (window[(r - 2) & 3][px][q] + window[(r - 1) & 3][px][q]);
                            value += nzv + bs[q];
                            value = fmaxf(value, value * kLreluAlpha);
                            o[q] = value * st[q];
                        }
                        *reinterpret_cast<uint2*>(out_b + ((size_t)y * w_out + x + px) * channels) = pack4(o, (T*)nullptr);
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_bar + stage);
        if (++stage == kFirStages) {
            stage = 0;
            phase ^= 1u;
        }
    }
}

int encode_tensor_map(CUtensorMap* map, const void* base, bool half_operands, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes, const cuuint32_t* box, int swizzle_bytes);
int sm_count();

template <typename T>
static int launch_phase(const void* x, int batch, int h_in, int w_in, int c_in, int c_out, const void* weight_phase,
                        const float* demod, const float* noise, long long noise_batch_stride, float noise_strength,
                        const float* bias, const float* next_styles, void* scratch, void* out, cudaStream_t stream) {
    constexpr bool kHalf = std::is_same<T, __half>::value;
    PhaseParams p;
    p.batch = batch; p.h_cells = h_in + 1; p.w_cells = w_in + 1; p.c_in = c_in; p.c_out = c_out;
    p.k_chunks = c_in / kPhaseK; p.n_tiles = c_out / kPhaseCC;
    // H+1 cells per axis is never a multiple of a power of two: pick the tile shape (<= 128 cells) with the fewest tiles
    long long best = -1;
    for (int tw = 8; tw <= 16; ++tw) {
        const int th = kTileM / tw;
        const long long tiles = (long long)((p.w_cells + tw - 1) / tw) * ((p.h_cells + th - 1) / th);
        if (best < 0 || tiles < best) {
            best = tiles; p.tile_w = tw; p.tile_h = th;
        }
    }
    p.tiles_w = (p.w_cells + p.tile_w - 1) / p.tile_w; p.tiles_h = (p.h_cells + p.tile_h - 1) / p.tile_h;
    p.demod = demod; p.t = scratch;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)c_in, (cuuint64_t)w_in, (cuuint64_t)h_in, (cuuint64_t)batch};
        cuuint64_t strides[3] = {(cuuint64_t)c_in * 2, (cuuint64_t)w_in * c_in * 2, (cuuint64_t)h_in * w_in * c_in * 2};
        cuuint32_t box[4] = {(cuuint32_t)kPhaseK, (cuuint32_t)p.tile_w, (cuuint32_t)p.tile_h, 1u};
        int rc = encode_tensor_map(&tmap_x, x, kHalf, 4, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)c_in, (cuuint64_t)9 * c_out};
        cuuint64_t strides[1] = {(cuuint64_t)c_in * 2};
        cuuint32_t box[2] = {(cuuint32_t)kPhaseK, (cuuint32_t)kPhaseCC};
        int rc = encode_tensor_map(&tmap_w, weight_phase, kHalf, 2, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    const int total_tiles = batch * p.tiles_h * p.tiles_w * p.n_tiles;
    const int grid = total_tiles < sm_count() ? total_tiles : sm_count();
    upconv_phase_kernel<T><<<grid, kThreads, kPhaseSmemBytes, stream>>>(tmap_x, tmap_w, p);
    int rc = check_launch("upconv_phase_kernel");
    if (rc != GANCE_OK) return rc;
    const int h_out = 2 * h_in, w_out = 2 * w_in;
    CUtensorMap tmap_t;
    {
        cuuint64_t dims[4] = {(cuuint64_t)c_out, (cuuint64_t)(w_out + 2), (cuuint64_t)(h_out + 2), (cuuint64_t)batch};
        cuuint64_t strides[3] = {(cuuint64_t)c_out * 2, (cuuint64_t)(w_out + 2) * c_out * 2,
                                 (cuuint64_t)(h_out + 2) * (w_out + 2) * c_out * 2};
        cuuint32_t box[4] = {(cuuint32_t)kFirC, (cuuint32_t)kFirTW, (cuuint32_t)kFirTH, 1u};
        rc = encode_tensor_map(&tmap_t, scratch, kHalf, 4, dims, strides, box, 0);
        if (rc != GANCE_OK) return rc;
    }
    const int fir_tiles = batch * ((h_out + kFirTileH - 1) / kFirTileH) * ((w_out + kFirTileW - 1) / kFirTileW) * (c_out / kFirC);
    const int fir_grid = fir_tiles < 2 * sm_count() ? fir_tiles : 2 * sm_count();
    upfir_act_kernel<T><<<fir_grid, kFirConsumers + 32, kFirSmemBytes, stream>>>(tmap_t, batch, h_out, w_out, c_out, noise,
                                                                             noise_batch_stride, noise_strength, bias,
                                                                             next_styles, (T*)out);
    return check_launch("upfir_act_kernel");
}

int configure_phase_kernels() {
    cudaError_t err = cudaFuncSetAttribute(upconv_phase_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPhaseSmemBytes);
    if (err == cudaSuccess)
        err = cudaFuncSetAttribute(upconv_phase_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPhaseSmemBytes);
    if (err != cudaSuccess)
        return fail(-(int)err, "cudaFuncSetAttribute(upconv_phase_kernel, %d B): %s", kPhaseSmemBytes, cudaGetErrorString(err));
    err = cudaFuncSetAttribute(upfir_act_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirSmemBytes);
    if (err == cudaSuccess)
        err = cudaFuncSetAttribute(upfir_act_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFirSmemBytes);
    if (err != cudaSuccess)
        return fail(-(int)err, "cudaFuncSetAttribute(upfir_act_kernel, %d B): %s", kFirSmemBytes, cudaGetErrorString(err));
    return GANCE_OK;
}

}  // namespace tc
}  // namespace gance

using namespace gance;

extern "C" int gance_modconv3x3_up2_phase(const void* x, int dtype, int batch, int h_in, int w_in, int c_in, int c_out,
                                          const void* weight_phase, const float* demod, const float* noise,
                                          long long noise_batch_stride, float noise_strength, const float* bias,
                                          const float* next_styles, void* scratch, void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight_phase && demod && bias && scratch && out, GANCE_E_NULL, "gance_modconv3x3_up2_phase: NULL pointer");
    GANCE_REQUIRE(dtype == GANCE_BF16 || dtype == GANCE_F16, GANCE_E_DTYPE, "gance_modconv3x3_up2_phase: dtype %d", dtype);
    GANCE_REQUIRE(batch > 0 && h_in >= 8 && w_in >= 16 && c_in % 64 == 0 && c_out % 64 == 0, GANCE_E_SHAPE,
                  "gance_modconv3x3_up2_phase: shape B=%d H=%d W=%d Cin=%d Cout=%d (needs H >= 8, W >= 16, channels %% 64 == 0)",
                  batch, h_in, w_in, c_in, c_out);
    GANCE_REQUIRE(aligned16(x) && aligned16(weight_phase) && aligned16(scratch) && aligned16(out), GANCE_E_ALIGN,
                  "gance_modconv3x3_up2_phase: pointers must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == GANCE_F16)
        return tc::launch_phase<__half>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise, noise_batch_stride,
                                        noise_strength, bias, next_styles, scratch, out, s);
    return tc::launch_phase<__nv_bfloat16>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise, noise_batch_stride,
                                           noise_strength, bias, next_styles, scratch, out, s);
}

// ---- the same two launches for callers that cannot allocate: t lives in a file-scope device array ---------------------------
// No counterpart in the reference; added for callers that may not call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Every GPU that loads the library holds
// these kStaticScratchBytes until exit. Both kernels are the ones above, unchanged; only where t is kept differs.
//
// In host code `g_phase_static_scratch` is the host-side stand-in of the symbol, NOT a device address: the address the
// kernels get is resolved with cudaGetSymbolAddress, once per device, with that device current (phase_static_scratch()).
//
// Every call on a device shares the array and finds the previous call's t in it. Nothing has to be cleared: pass 1
// writes every element of [B, 2H+2, 2W+2, Cout] (the zero border included) before pass 2, next on the same stream,
// reads exactly that range. Calls that overlap on the GPU (two streams, two host threads) race on the array.
namespace gance {
namespace tc {

constexpr size_t kStaticScratchBytes = GANCE_PHASE_STATIC_SCRATCH_BYTES;   // 4 MiB: e.g. B=1, 32^2 -> 64^2, Cout=256 is 2.2 MB
static __device__ __align__(256) unsigned char g_phase_static_scratch[kStaticScratchBytes];

constexpr int kStaticScratchMaxDevices = 64;
static std::mutex g_static_scratch_mutex;
static void* g_static_scratch_address[kStaticScratchMaxDevices] = {};

// device address of the array on the CURRENT device; resolved on that device's first call, no allocation
static int phase_static_scratch(void** address) {
    int device = -1;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return fail(-(int)err, "gance_modconv3x3_up2_phase_static: cudaGetDevice: %s", cudaGetErrorString(err));
    if (device < 0 || device >= kStaticScratchMaxDevices)
        return fail(GANCE_E_SHAPE, "gance_modconv3x3_up2_phase_static: device %d is beyond the %d this entry tracks", device,
                    kStaticScratchMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (g_static_scratch_address[device] == nullptr) {
        void* resolved = nullptr;
        err = cudaGetSymbolAddress(&resolved, g_phase_static_scratch);
        if (err != cudaSuccess || resolved == nullptr)
            return fail(err != cudaSuccess ? -(int)err : GANCE_E_DRIVER,
                        "gance_modconv3x3_up2_phase_static: cudaGetSymbolAddress on device %d: %s", device,
                        cudaGetErrorString(err));
        g_static_scratch_address[device] = resolved;
    }
    *address = g_static_scratch_address[device];
    return GANCE_OK;
}

}  // namespace tc
}  // namespace gance

extern "C" int gance_modconv3x3_up2_phase_static(const void* x, int dtype, int batch, int h_in, int w_in, int c_in, int c_out,
                                                 const void* weight_phase, const float* demod, const float* noise,
                                                 long long noise_batch_stride, float noise_strength, const float* bias,
                                                 const float* next_styles, void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight_phase && demod && bias && out, GANCE_E_NULL, "gance_modconv3x3_up2_phase_static: NULL pointer");
    GANCE_REQUIRE(dtype == GANCE_BF16 || dtype == GANCE_F16, GANCE_E_DTYPE, "gance_modconv3x3_up2_phase_static: dtype %d", dtype);
    GANCE_REQUIRE(batch > 0 && h_in >= 8 && w_in >= 16 && c_in > 0 && c_out > 0 && c_in % 64 == 0 && c_out % 64 == 0, GANCE_E_SHAPE,
                  "gance_modconv3x3_up2_phase_static: shape B=%d H=%d W=%d Cin=%d Cout=%d (needs H >= 8, W >= 16, channels %% 64 == 0)",
                  batch, h_in, w_in, c_in, c_out);
    const unsigned long long needed = 2ull * (unsigned long long)batch * (2ull * h_in + 2) * (2ull * w_in + 2) * (unsigned long long)c_out;
    GANCE_REQUIRE(needed <= (unsigned long long)tc::kStaticScratchBytes, GANCE_E_SHAPE,
                  "gance_modconv3x3_up2_phase_static: t = [%d,%d,%d,%d] x 2 B needs %llu bytes, the static scratch holds %llu; "
                  "use gance_modconv3x3_up2_phase with a caller-owned scratch",
                  batch, 2 * h_in + 2, 2 * w_in + 2, c_out, needed, (unsigned long long)tc::kStaticScratchBytes);
    GANCE_REQUIRE(aligned16(x) && aligned16(weight_phase) && aligned16(out), GANCE_E_ALIGN,
                  "gance_modconv3x3_up2_phase_static: pointers must be 16-byte aligned");
    void* scratch = nullptr;
    const int rc = tc::phase_static_scratch(&scratch);
    if (rc != GANCE_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == GANCE_F16)
        return tc::launch_phase<__half>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise, noise_batch_stride,
                                        noise_strength, bias, next_styles, scratch, out, s);
    return tc::launch_phase<__nv_bfloat16>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise, noise_batch_stride,
                                           noise_strength, bias, next_styles, scratch, out, s);
}
