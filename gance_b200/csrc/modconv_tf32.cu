// fp32-grade validation mode ON THE TENSOR CORE: the 3x3 modulated convolution as a tcgen05 implicit GEMM with
// kind::tf32 operands and error compensation ("3xTF32").
//
// north_star asks for "bf16 with fp32 accumulate, plus an fp32/TF32 validation mode". The fp32 arbiter of this library
// is a CUDA-core FMA kernel (synthesis_kernels.cu, modconv3x3_f32_kernel): exact fp32 products, but ~40x slower than
// the 16-bit tensor path, so parity at 1024^2 costs seconds. A single tf32 pass (10-bit significand, inputs truncated by
// the tensor core) is too coarse for the 1e-3 max-abs criterion over 18 layers, so every operand is split:
//     x = x_hi + x_lo,   x_hi = x with its low 13 significand bits cleared (exactly a tf32 value), x_lo = x - x_hi (exact)
// and D = A_lo*B_hi + A_hi*B_lo + A_hi*B_hi with fp32 accumulation in TMEM: the dropped term A_lo*B_lo and the
// truncation of the lo parts are both ~2^-22 relative. Activations are split by gance_split_tf32 (which can also apply
// the per-sample style), weights once at upload.
//
// What limits the result is the accumulation itself: the tensor core TRUNCATES when it adds a K = 8 block of products to
// the accumulator, up to one ulp of the accumulator per instruction and always in the same direction, so a chain of n
// instructions drifts by ~n * 2^-24 (measured with everything in one accumulator: 3.4e-5 relative at K = 4608, 3e-6 at
// K = 288). Two thirds of those instructions are the tiny cross terms, which cost a full ulp of the LARGE sum each. So a
// tile keeps three accumulators - cross terms (magnitude 2^-11 of the sum: their ulps are negligible), hi*hi of the even
// K steps and hi*hi of the odd K steps - and the epilogue adds them in fp32 with round-to-nearest.
//
// Same structure as modconv_tc.cu: persistent CTAs, warp 0 = TMA producer (four boxes per stage: A_hi, A_lo, B_hi, B_lo,
// 128-byte swizzled rows of 32 floats), warp 1 = MMA issuer (M = 128 pixels, N = N_TILE channels, K = 8 per instruction),
// warps 2-9 = two epilogue groups draining the two TMEM accumulators:
//     out = lrelu(acc * demod[b,co] + noise * strength + bias[co]) * sqrt(2) * next_styles[b,co]   -> fp32 NHWC
// Plain convolutions only (pad 0 or 1): the up-convolutions of this mode run FIR-first (gance_upfir2x_pad) like the
// CUDA-core mode, ToRGB stays on gance_torgb_skip.
#include "tc_ptx.cuh"

namespace gance {
namespace tc {

int encode_tensor_map_f32(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
                          const cuuint64_t* strides_bytes, const cuuint32_t* box);  // modconv_tc.cu
int sm_count();

constexpr int kTf32BlockK = 32;  // floats per 128-byte swizzled row
constexpr int kTf32UmmaK = 8;

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// Instruction descriptor as in tc_ptx.cuh, with a_format = b_format = TF32 (2).
__host__ __device__ constexpr uint32_t make_instr_desc_tf32(int m, int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

template <int N_TILE, int STAGES>
struct Tf32Smem {
    static constexpr int kABytes = kTileM * kTf32BlockK * 4;   // one of A_hi / A_lo
    static constexpr int kBBytes = N_TILE * kTf32BlockK * 4;   // one of B_hi / B_lo
    static constexpr int kStageBytes = 2 * kABytes + 2 * kBBytes;
    static constexpr int kBarrierOffset = STAGES * kStageBytes;
    static constexpr int kBytes = kBarrierOffset + (2 * STAGES + 4) * 8 + 16 + 1024;
};

struct Tf32Params {
    int batch, h_out, w_out, c_in, c_out, pad;
    int tile_w, tile_h, tile_w_log2;
    int tiles_w, tiles_h;
    int n_tiles, k_chunks, a_tx_bytes;
    const float* demod;
    const float* noise;
    long long noise_batch_stride;
    float noise_strength;
    const float* bias;
    const float* next_styles;
    float* out;
};

template <int N_TILE, int STAGES>
__global__ void __launch_bounds__(kThreads, 1)
modconv3x3_tf32x3_kernel(const __grid_constant__ CUtensorMap tmap_x_hi, const __grid_constant__ CUtensorMap tmap_x_lo,
                         const __grid_constant__ CUtensorMap tmap_w_hi, const __grid_constant__ CUtensorMap tmap_w_lo,
                         const Tf32Params p) {
    using L = Tf32Smem<N_TILE, STAGES>;
    // two tile buffers x three accumulators (hi*hi even K steps, hi*hi odd K steps, cross terms) of N_TILE columns
    constexpr uint32_t kAccCols = 3 * N_TILE;
    constexpr uint32_t kTmemCols = 2 * kAccCols <= 256 ? 256 : 512;
    static_assert(2 * kAccCols <= 512, "TMEM holds 512 columns");
    static_assert(L::kABytes % 1024 == 0 && L::kBBytes % 1024 == 0, "operand buffers must stay 1024-byte aligned");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x_hi); prefetch_tmap(&tmap_x_lo); prefetch_tmap(&tmap_w_hi); prefetch_tmap(&tmap_w_lo);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 4);  // one arrival per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_base_slot, kTmemCols);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);

    const int m_tiles = p.batch * p.tiles_h * p.tiles_w;
    const int total_tiles = m_tiles * p.n_tiles;
    const int k_iters = 9 * p.k_chunks;

    auto decode = [&](int tile, int& n0, int& b0, int& h0, int& w0) {
        const int nt = tile % p.n_tiles;
        int m = tile / p.n_tiles;
        const int tw = m % p.tiles_w;
        m /= p.tiles_w;
        const int th = m % p.tiles_h;
        b0 = m / p.tiles_h;
        n0 = nt * N_TILE;
        w0 = tw * p.tile_w;
        h0 = th * p.tile_h;
    };

    if (warp == 0) {
        // ===== TMA producer =====
        const bool leader = elect_one_sync();
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            int n0, b0, h0, w0;
            decode(tile, n0, b0, h0, w0);
            for (int tap = 0; tap < 9; ++tap) {
                const int dy = tap / 3 - p.pad, dx = tap % 3 - p.pad;
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    mbar_wait(empty_bar + stage, phase ^ 1u);
                    uint8_t* a_hi = smem + stage * L::kStageBytes;
                    uint8_t* a_lo = a_hi + L::kABytes;
                    uint8_t* b_hi = a_lo + L::kABytes;
                    uint8_t* b_lo = b_hi + L::kBBytes;
                    if (leader) {
                        mbar_expect_tx(full_bar + stage, 2u * (uint32_t)p.a_tx_bytes + 2u * L::kBBytes);
                        tma_load_4d(a_hi, &tmap_x_hi, full_bar + stage, kc * kTf32BlockK, w0 + dx, h0 + dy, b0);
                        tma_load_4d(a_lo, &tmap_x_lo, full_bar + stage, kc * kTf32BlockK, w0 + dx, h0 + dy, b0);
                        tma_load_2d(b_hi, &tmap_w_hi, full_bar + stage, tap * p.c_in + kc * kTf32BlockK, n0);
                        tma_load_2d(b_lo, &tmap_w_lo, full_bar + stage, tap * p.c_in + kc * kTf32BlockK, n0);
                    }
                    __syncwarp();
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const bool leader = elect_one_sync();
        constexpr uint32_t idesc = make_instr_desc_tf32(kTileM, N_TILE);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        const uint32_t smem_base = smem_u32(smem);
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
            tcgen05_fence_after();
            const uint32_t d_hi0 = tmem_base + (uint32_t)acc * kAccCols;
            const uint32_t d_hi1 = d_hi0 + (uint32_t)N_TILE, d_lo = d_hi0 + 2u * (uint32_t)N_TILE;
            for (int it = 0; it < k_iters; ++it) {
                mbar_wait(full_bar + stage, phase);
                tcgen05_fence_after();
                const uint32_t a_addr = smem_base + (uint32_t)(stage * L::kStageBytes);
                const uint64_t a_hi = make_smem_desc<128>(a_addr);
                const uint64_t a_lo = a_hi + (uint64_t)(L::kABytes >> 4);
                const uint64_t b_hi = a_lo + (uint64_t)(L::kABytes >> 4);
                const uint64_t b_lo = b_hi + (uint64_t)(L::kBBytes >> 4);
                if (leader) {
#pragma unroll
                    for (int k = 0; k < kTf32BlockK / kTf32UmmaK; ++k) {
                        const uint64_t step = (uint64_t)(k * 2);  // +32 bytes along K inside the swizzle atom
                        umma_tf32(d_lo, a_lo + step, b_hi + step, idesc, (it | k) != 0 ? 1u : 0u);
                        umma_tf32(d_lo, a_hi + step, b_lo + step, idesc, 1u);
                        umma_tf32((k & 1) ? d_hi1 : d_hi0, a_hi + step, b_hi + step, idesc, (it | (k >> 1)) != 0 ? 1u : 0u);
                    }
                    umma_commit(empty_bar + stage);
                    if (it == k_iters - 1) umma_commit(tmem_full_bar + acc);
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1u;
        }
    } else {
        // ===== epilogue: group g drains accumulator g (tiles g, g+2, ... of this CTA) =====
        const int group = (warp - 2) >> 2;
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const int wl = row & (p.tile_w - 1);
        const int hl = row >> p.tile_w_log2;
        uint32_t full_phase = 0;
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)group * kAccCols;
        for (int tile = blockIdx.x + group * gridDim.x; tile < total_tiles; tile += 2 * gridDim.x) {
            int n0, b, h0, w0;
            decode(tile, n0, b, h0, w0);
            const int h = h0 + hl, w = w0 + wl;
            const bool valid = hl < p.tile_h && h < p.h_out && w < p.w_out;
            float nz = 0.f;
            if (valid && p.noise != nullptr)
                nz = __ldg(p.noise + (size_t)b * p.noise_batch_stride + (size_t)h * p.w_out + w) * p.noise_strength;
            mbar_wait(tmem_full_bar + group, full_phase);
            full_phase ^= 1u;
            tcgen05_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < N_TILE; c0 += 32) {
                uint32_t r[32], r1[32], r2[32];
                tmem_ld_32x32(taddr + (uint32_t)c0, r);
                tmem_ld_32x32(taddr + (uint32_t)(N_TILE + c0), r1);
                tmem_ld_32x32(taddr + (uint32_t)(2 * N_TILE + c0), r2);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i)  // (hi*hi even + hi*hi odd) + cross terms, rounded to nearest
                    r[i] = __float_as_uint(__fadd_rn(__fadd_rn(__uint_as_float(r[i]), __uint_as_float(r1[i])), __uint_as_float(r2[i])));
                if (valid) {
                    float* dst = p.out + (((size_t)b * p.h_out + h) * p.w_out + w) * p.c_out + n0 + c0;
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        float v[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const int co = n0 + c0 + g * 4 + q;
                            const size_t bc = (size_t)b * p.c_out + co;
                            float o = fmaf(__uint_as_float(r[g * 4 + q]), __ldg(p.demod + bc), nz + __ldg(p.bias + co));
                            o = (o > 0.f ? o : o * kLreluAlpha) * kLreluGain;
                            if (p.next_styles != nullptr) o *= __ldg(p.next_styles + bc);
                            v[q] = o;
                        }
                        *reinterpret_cast<float4*>(dst + g * 4) = make_float4(v[0], v[1], v[2], v[3]);
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
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// x (optionally times the per-sample style) -> hi (low 13 significand bits cleared) + lo (the exact remainder)
__global__ void split_tf32_kernel(const float4* __restrict__ x, size_t vectors, const float* __restrict__ styles,
                                  int channels, size_t pixels_per_sample, float4* __restrict__ hi,
                                  float4* __restrict__ lo) {
    const int vec_per_pixel = channels >> 2;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < vectors; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = x[i];
        if (styles != nullptr) {
            const size_t pixel = i / vec_per_pixel;
            const int c = (int)(i - pixel * vec_per_pixel) << 2;
            const float4 s = __ldg(reinterpret_cast<const float4*>(styles + (pixel / pixels_per_sample) * channels + c));
            v.x *= s.x; v.y *= s.y; v.z *= s.z; v.w *= s.w;
        }
        float4 h;
        h.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u);
        h.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u);
        h.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u);
        h.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u);
        hi[i] = h;
        lo[i] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
    }
}

template <int N_TILE, int STAGES>
static int launch_tf32(const Tf32Params& p, const float* x_hi, const float* x_lo, const float* w_hi, const float* w_lo,
                       int h_in, int w_in, cudaStream_t stream) {
    using L = Tf32Smem<N_TILE, STAGES>;
    static bool configured = false;
    if (!configured) {
        cudaError_t err = cudaFuncSetAttribute(modconv3x3_tf32x3_kernel<N_TILE, STAGES>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes);
        if (err != cudaSuccess) return fail(-(int)err, "cudaFuncSetAttribute(modconv3x3_tf32x3, %d B): %s", L::kBytes, cudaGetErrorString(err));
        configured = true;
    }
    CUtensorMap maps[4];
    const float* bases[4] = {x_hi, x_lo, w_hi, w_lo};
    for (int i = 0; i < 2; ++i) {
        cuuint64_t dims[4] = {(cuuint64_t)p.c_in, (cuuint64_t)w_in, (cuuint64_t)h_in, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {(cuuint64_t)p.c_in * 4, (cuuint64_t)w_in * p.c_in * 4, (cuuint64_t)h_in * w_in * p.c_in * 4};
        cuuint32_t box[4] = {(cuuint32_t)kTf32BlockK, (cuuint32_t)p.tile_w, (cuuint32_t)p.tile_h, 1u};
        int rc = encode_tensor_map_f32(&maps[i], bases[i], 4, dims, strides, box);
        if (rc != GANCE_OK) return rc;
    }
    for (int i = 2; i < 4; ++i) {
        cuuint64_t dims[2] = {(cuuint64_t)9 * p.c_in, (cuuint64_t)p.c_out};
        cuuint64_t strides[1] = {(cuuint64_t)9 * p.c_in * 4};
        cuuint32_t box[2] = {(cuuint32_t)kTf32BlockK, (cuuint32_t)N_TILE};
        int rc = encode_tensor_map_f32(&maps[i], bases[i], 2, dims, strides, box);
        if (rc != GANCE_OK) return rc;
    }
    const int total_tiles = p.batch * p.tiles_h * p.tiles_w * p.n_tiles;
    const int grid = total_tiles < sm_count() ? total_tiles : sm_count();
    modconv3x3_tf32x3_kernel<N_TILE, STAGES><<<grid, kThreads, L::kBytes, stream>>>(maps[0], maps[1], maps[2], maps[3], p);
    return check_launch("modconv3x3_tf32x3_kernel");
}

}  // namespace tc
}  // namespace gance

using namespace gance;

extern "C" int gance_split_tf32(const float* x, const float* styles, int batch, int h, int w, int channels, float* hi,
                                float* lo, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && hi && lo, GANCE_E_NULL, "gance_split_tf32: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h > 0 && w > 0 && channels > 0 && channels % 4 == 0, GANCE_E_SHAPE,
                  "gance_split_tf32: shape B=%d H=%d W=%d C=%d (C %% 4 == 0)", batch, h, w, channels);
    GANCE_REQUIRE(aligned16(x) && aligned16(hi) && aligned16(lo) && aligned16(styles), GANCE_E_ALIGN,
                  "gance_split_tf32: pointers must be 16-byte aligned");
    const size_t vectors = (size_t)batch * h * w * (channels / 4);
    size_t blocks = (vectors + 255) / 256;
    const size_t cap = (size_t)tc::sm_count() * 16;
    if (blocks > cap) blocks = cap;
    tc::split_tf32_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(x), vectors, styles, channels, (size_t)h * w, reinterpret_cast<float4*>(hi),
        reinterpret_cast<float4*>(lo));
    return check_launch("split_tf32_kernel");
}

extern "C" int gance_modconv3x3_tf32x3(const float* x_hi, const float* x_lo, int batch, int h_out, int w_out, int c_in,
                                       int c_out, int pad, const float* w_hi, const float* w_lo, const float* demod,
                                       const float* noise, long long noise_batch_stride, float noise_strength,
                                       const float* bias, const float* next_styles, float* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x_hi && x_lo && w_hi && w_lo && demod && bias && out, GANCE_E_NULL, "gance_modconv3x3_tf32x3: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h_out > 0 && w_out > 0 && (pad == 0 || pad == 1), GANCE_E_SHAPE,
                  "gance_modconv3x3_tf32x3: shape B=%d H=%d W=%d pad=%d", batch, h_out, w_out, pad);
    GANCE_REQUIRE(c_in % 32 == 0 && c_out % 32 == 0 && c_in > 0 && c_out > 0, GANCE_E_SHAPE,
                  "gance_modconv3x3_tf32x3: Cin %% 32 == 0 and Cout %% 32 == 0 required (got %d, %d)", c_in, c_out);
    GANCE_REQUIRE(aligned16(x_hi) && aligned16(x_lo) && aligned16(w_hi) && aligned16(w_lo) && aligned16(out), GANCE_E_ALIGN,
                  "gance_modconv3x3_tf32x3: pointers must be 16-byte aligned");
    auto is_pow2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
    tc::Tf32Params p;
    p.batch = batch; p.h_out = h_out; p.w_out = w_out; p.c_in = c_in; p.c_out = c_out; p.pad = pad;
    p.tile_w = w_out >= 16 ? 16 : w_out;
    GANCE_REQUIRE(is_pow2(p.tile_w), GANCE_E_SHAPE, "gance_modconv3x3_tf32x3: W=%d (< 16) must be a power of two", w_out);
    const int rows = tc::kTileM / p.tile_w;
    p.tile_h = h_out >= rows ? rows : h_out;
    GANCE_REQUIRE(is_pow2(p.tile_h), GANCE_E_SHAPE, "gance_modconv3x3_tf32x3: H=%d (< %d) must be a power of two", h_out, rows);
    p.tile_w_log2 = 0;
    while ((1 << p.tile_w_log2) < p.tile_w) ++p.tile_w_log2;
    p.tiles_w = (w_out + p.tile_w - 1) / p.tile_w;
    p.tiles_h = (h_out + p.tile_h - 1) / p.tile_h;
    p.k_chunks = c_in / tc::kTf32BlockK;
    p.a_tx_bytes = p.tile_w * p.tile_h * tc::kTf32BlockK * 4;
    p.demod = demod; p.noise = noise; p.noise_batch_stride = noise_batch_stride; p.noise_strength = noise_strength;
    p.bias = bias; p.next_styles = next_styles; p.out = out;
    const int h_in = h_out + 2 - 2 * pad, w_in = w_out + 2 - 2 * pad;
    if (c_out % 64 == 0) {
        p.n_tiles = c_out / 64;
        return tc::launch_tf32<64, 4>(p, x_hi, x_lo, w_hi, w_lo, h_in, w_in, (cudaStream_t)stream);
    }
    p.n_tiles = c_out / 32;
    return tc::launch_tf32<32, 4>(p, x_hi, x_lo, w_hi, w_lo, h_in, w_in, (cudaStream_t)stream);
}
