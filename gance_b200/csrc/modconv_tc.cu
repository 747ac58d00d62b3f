// 3x3 modulated convolution as a tcgen05 / TMEM implicit GEMM fed by TMA (sm_100a).
//
// Math (SURVEY.md Appendix A.4 identity, proven in tests/test_oracle_stylegan2.py):
//   modulated_conv2d_layer(x, w)  ==  conv(x * s, W) * d
// so ONE weight matrix serves the whole batch and the batch folds into the GEMM M dimension:
//   D[M = B*H*W pixels, N = Cout] = A[M, K = 9*Cin] * B[N, K]^T,   K ordered (ky, kx, ci)
// A is never materialised: for tap (ky,kx) and channel chunk c0 the A tile is the NHWC box
//   x[b0:b0+tb, h0+ky-pad : +th, w0+kx-pad : +tw, c0 : c0+BLOCK_K]
// which one 4-D TMA load drops into shared memory in exactly the K-major SWIZZLE_128B (or 64B)
// layout the UMMA shared-memory descriptor wants; out-of-range coordinates are zero-filled by the
// TMA unit, which is the SAME padding. B tiles are 2-D TMA boxes of the [Cout][9*Cin] weight.
//
// CTA = 6 warps, persistent over (m-tile, n-tile) work items:
//   warp 0   : TMA producer (one elected lane)           smem ring: full[] / empty[] mbarriers
//   warp 1   : tcgen05.mma issuer (one elected lane), owns the TMEM allocation
//   warps 2-5: epilogue, one TMEM lane quarter each: tcgen05.ld -> demod * acc + noise + bias ->
//              leaky-ReLU * sqrt(2) -> (* next layer's style) -> bf16 -> 16-byte NHWC stores
// Accumulators are double-buffered in TMEM (2 x N_TILE columns) so the epilogue of tile i
// overlaps the MMAs of tile i+1.
#include <cstdlib>

#include "tc_ptx.cuh"

namespace gance {

namespace tc {

// Per-(sample, channel) epilogue coefficients, staged in shared memory once per sample and per epilogue
// group as float4 {demod, bias, sqrt(2) * next style, -} and, for the fused ToRGB,
// float4 {sqrt(2) * rgb_style * rgb_weight[.,0..2], -}: one broadcast LDS.128 per channel (and one more for RGB).
// Up to 512 channels without ToRGB, 256 with it: 8 KB per group.
constexpr int kCacheFloat4PerGroup = 512;
constexpr int kCacheBytes = kEpilogueGroups * kCacheFloat4PerGroup * 16;

template <int N_TILE, int BLOCK_K, int STAGES>
struct SmemLayout {
    static constexpr int kABytes = kTileM * BLOCK_K * 2;
    static constexpr int kBBytes = N_TILE * BLOCK_K * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kBarrierOffset = STAGES * kStageBytes;
    static constexpr int kCacheOffset = kBarrierOffset + (2 * STAGES + 4) * 8 + 16;
    static constexpr int kBytes = kCacheOffset + kCacheBytes + 1024;  // + alignment slack
};

struct TileCoord {
    int n0, b0, h0, w0;
};

// m-tile index -> (sample, first row, first column) of the tile
__device__ __forceinline__ TileCoord decode_m_tile(const ConvParams& p, int m, int nt, int n_tile_size) {
    TileCoord t;
    int tw, th;
    if (p.tiles_w_log2 >= 0) {  // every 2^k-sized image: three runtime divisions (~60 instructions per warp and tile) become shifts
        tw = m & (p.tiles_w - 1);
        m >>= p.tiles_w_log2;
        th = m & (p.tiles_h - 1);
        t.b0 = m >> p.tiles_h_log2;
    } else {
        tw = m % p.tiles_w;
        m /= p.tiles_w;
        th = m % p.tiles_h;
        t.b0 = m / p.tiles_h;  // tile_b == 1
    }
    t.n0 = nt * n_tile_size;
    t.w0 = tw << p.tile_w_log2;
    t.h0 = th * p.tile_h;
    return t;
}

__device__ __forceinline__ TileCoord decode_tile(const ConvParams& p, int tile, int n_tile_size) {
    int nt = 0, m = tile;
    if (p.n_tiles != 1) {
        nt = tile % p.n_tiles;
        m = tile / p.n_tiles;
    }
    return decode_m_tile(p, m, nt, n_tile_size);
}

// CTA-pair kernels: work item = (pair of consecutive m-tiles, n-tile); CTA `rank` of the pair owns m-tile 2 * pair + rank.
// With an odd number of m-tiles the last pair's second tile is a GHOST: its sample index is the batch size, so the TMA
// unit zero-fills its A operand (the byte counts of the pipeline stay the same) and the epilogue stores nothing.
__device__ __forceinline__ TileCoord decode_pair_item(const ConvParams& p, int item, int rank, int n_tile_size, bool& ghost) {
    int nt = 0, mp = item;
    if (p.n_tiles != 1) {
        nt = item % p.n_tiles;
        mp = item / p.n_tiles;
    }
    const int m = 2 * mp + rank;
    ghost = m >= p.tiles_b * p.tiles_h * p.tiles_w;
    return decode_m_tile(p, m, nt, n_tile_size);
}

// Epilogue shared by both kernels. Two groups of four warps; group g drains TMEM accumulator buffer g, i.e. the
// CTA's tiles g, g+2, g+4, ... (the MMA issuer alternates buffers), so the epilogue of one tile overlaps both the
// MMAs and the epilogue of the next. Inside a group warp w owns TMEM lane quarter w % 4 and every thread one
// accumulator row = one output pixel of the tile:
//   v = lrelu(acc * demod[b,co] + noise * strength + bias[co])              (gain sqrt(2) folded into the consumers)
//   fused ToRGB (RGB): rgb += v * sqrt(2) * rgb_styles[b,co] * rgb_weight[co,:] in fp32, then
//       y = upsample_2d(y_prev) + rgb + rgb_bias -> y_out (fp32) and/or uint8 frames
//   x_out = v * sqrt(2) * next_styles[b,co] (the next convolution's modulation) -> 16-bit NHWC store (skipped if
//   out == NULL)
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ void epilogue_bar_sync(int group) {
    asm volatile("bar.sync %0, 128;" ::"r"(group + 1) : "memory");
}

// APG = TMEM accumulator buffers per epilogue group (1 or 2). The n-th tile of group g lives in buffer g + GROUPS * (n % APG).
// With one buffer per group the MMAs of tile i + GROUPS cannot start before the epilogue of tile i has drained its
// buffer: a tile's epilogue latency must fit inside (GROUPS - 1) tiles of MMA time. With two, only its THROUGHPUT has to.
template <typename T, int N_TILE, bool UP, bool RGB, int GROUPS, int CACHE_F4, bool PAIR = false, int APG = 1>
__device__ __forceinline__ void conv_epilogue(const ConvParams& p, int group, int warp, int lane, int total_tiles,
                                              uint32_t tmem_base, uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                              float4* cache_all) {
    // PAIR: `total_tiles` counts pair items, walked cluster by cluster; the drained accumulator is reported to the
    // LEADER's barrier (the leader's MMA warp issues for both CTAs)
    const int pair_rank = PAIR ? (int)cluster_ctarank() : 0;
    const int walk_first = PAIR ? (int)cluster_id_x() : (int)blockIdx.x;
    const int walk_stride = PAIR ? (int)cluster_count_x() : (int)gridDim.x;
    const uint32_t leader_empty_base = PAIR ? cluster_map(smem_u32(tmem_empty_bar), 0) : 0u;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int wl = row & (p.tile_w - 1);
    const int hl = row >> p.tile_w_log2;  // >= tile_h for the unused rows of a small-resolution tile
    const int C = p.c_real;
    const int out_w = UP ? 2 * p.w_out : p.w_out;
    float4* coef_a = cache_all + group * CACHE_F4;
    float4* coef_b = coef_a + C;
    const uint32_t coef_a_addr = smem_u32(coef_a);
    const int tid = quarter * 32 + lane;
    int cached_b = -1;
    int visit = 0;  // tiles this group has handled
    float3 rgb_bias = make_float3(0.f, 0.f, 0.f);
    if (RGB) rgb_bias = make_float3(__ldg(p.rgb_bias), __ldg(p.rgb_bias + 1), __ldg(p.rgb_bias + 2));

    for (int tile = walk_first + group * walk_stride; tile < total_tiles; tile += GROUPS * walk_stride, ++visit) {
        const int acc = group + GROUPS * (visit % APG);
        const uint32_t full_phase = (uint32_t)(visit / APG) & 1u;
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * N_TILE);
        bool ghost = false;
        const TileCoord t = PAIR ? decode_pair_item(p, tile, pair_rank, N_TILE, ghost) : decode_tile(p, tile, N_TILE);
        if (PAIR && ghost) {  // CTA-uniform: keep the accumulator hand-shake going, touch nothing else
            mbar_wait(tmem_full_bar + acc, full_phase);
            tcgen05_fence_after();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
            continue;
        }
        const int b = t.b0, h = t.h0 + hl, w = t.w0 + wl;
        const bool valid = hl < p.tile_h && h < p.h_out && w < p.w_out;
        const int hv = valid ? h : 0, wv = valid ? w : 0;
        if (b != cached_b) {      // group-uniform
            epilogue_bar_sync(group);  // everyone is done reading the previous sample's coefficients
            for (int i = tid; i < C; i += kEpilogueThreads) {
                const size_t bc = (size_t)b * C + i;
                const float st = p.next_styles ? __ldg(p.next_styles + bc) : 1.0f;
                coef_a[i] = make_float4(__ldg(p.demod + bc), __ldg(p.bias + i), st * kLreluGain, 0.f);
                if (RGB) {
                    const float s = __ldg(p.rgb_styles + bc) * kLreluGain;
                    coef_b[i] = make_float4(s * __ldg(p.rgb_weight + 3 * i), s * __ldg(p.rgb_weight + 3 * i + 1),
                                            s * __ldg(p.rgb_weight + 3 * i + 2), 0.f);
                }
            }
            epilogue_bar_sync(group);
            cached_b = b;
        }
        const float* noise_b = p.noise ? p.noise + (size_t)b * p.noise_batch_stride : nullptr;
        // everything that does not depend on the accumulator is fetched while the MMAs run
        float4 skip = make_float4(0.f, 0.f, 0.f, 0.f);
        if (RGB && valid && p.y_prev != nullptr) skip = upsample_tap(p.y_prev, b, p.h_out >> 1, p.w_out >> 1, hv, wv);
        float nz4[4] = {0.f, 0.f, 0.f, 0.f};
        if (noise_b != nullptr) {
            if (UP) {
                const float2 top = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)(2 * hv) * out_w + 2 * wv));
                const float2 bot = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)(2 * hv + 1) * out_w + 2 * wv));
                nz4[0] = __fmul_rn(top.x, p.noise_strength); nz4[1] = __fmul_rn(top.y, p.noise_strength);
                nz4[2] = __fmul_rn(bot.x, p.noise_strength); nz4[3] = __fmul_rn(bot.y, p.noise_strength);
            } else {
                nz4[0] = __fmul_rn(__ldg(noise_b + (size_t)hv * out_w + wv), p.noise_strength);
            }
        }
        float red = 0.f, green = 0.f, blue = 0.f;

        mbar_wait(tmem_full_bar + acc, full_phase);
        tcgen05_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < N_TILE; c0 += 32) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + (uint32_t)c0, r);
            // column chunk -> (phase, first channel): a 32-column chunk never straddles a phase (c_real % 32 == 0)
            const int n = t.n0 + c0;
            const int phase = UP ? n >> p.c_log2 : 0;
            const int co0 = UP ? n & (C - 1) : n;
            const int oh = UP ? 2 * hv + (phase >> 1) : hv, ow = UP ? 2 * wv + (phase & 1) : wv;
            const float nz = UP ? nz4[phase] : nz4[0];
            T* dst = nullptr;
            T* dst_plain = nullptr;
            const size_t out_offset = (((size_t)b * (UP ? 2 * p.h_out : p.h_out) + oh) * out_w + ow) * C + co0;
            if (p.out != nullptr) dst = reinterpret_cast<T*>(p.out) + out_offset;
            if (!UP && !RGB && p.out_plain != nullptr) dst_plain = reinterpret_cast<T*>(p.out_plain) + out_offset;
            const uint32_t ca = coef_a_addr + (uint32_t)co0 * 16u, cb = ca + (uint32_t)C * 16u;
            tmem_ld_wait();
            if (valid) {
#pragma unroll
                for (int g = 0; g < 4; ++g) {  // 8 channels -> one 16-byte store
                    float v[8], plain[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float4 a = lds128(ca + (uint32_t)(g * 8 + q) * 16u);
                        float o = fmaf(__uint_as_float(r[g * 8 + q]), a.x, nz + a.y);
                        o = fmaxf(o, o * kLreluAlpha);
                        if (RGB) {
                            const float4 wv4 = lds128(cb + (uint32_t)(g * 8 + q) * 16u);
                            red = fmaf(o, wv4.x, red); green = fmaf(o, wv4.y, green); blue = fmaf(o, wv4.z, blue);
                        }
                        v[q] = o * a.z;
                        plain[q] = o * kLreluGain;
                    }
                    if (dst != nullptr) {
                        Vec<T> packed;
                        packed.from_float(v);
                        packed.store(dst + g * 8);
                    }
                    if (!UP && !RGB && dst_plain != nullptr) {
                        Vec<T> packed;
                        packed.from_float(plain);
                        packed.store(dst_plain + g * 8);
                    }
                }
            }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) {  // accumulator drained: the MMAs of the tile that takes this buffer next may start
            if (PAIR) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
            else mbar_arrive(tmem_empty_bar + acc);
        }
        if (RGB && valid) {
            red += rgb_bias.x + skip.x; green += rgb_bias.y + skip.y; blue += rgb_bias.z + skip.z;
            const size_t pixel = ((size_t)b * p.h_out + hv) * p.w_out + wv;
            if (p.y_out != nullptr) reinterpret_cast<float4*>(p.y_out)[pixel] = make_float4(red, green, blue, 0.f);
            if (p.frames != nullptr) {
                uint8_t* f = p.frames + pixel * 3;
                f[0] = to_u8(red); f[1] = to_u8(green); f[2] = to_u8(blue);
                flag_nonfinite(p.nonfinite, red, green, blue);
            }
        }
    }
}

__device__ __forceinline__ uint32_t pack_pair(float a, float b, __half*) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t pack_pair(float a, float b, __nv_bfloat16*) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// ---- epilogue of the halo kernels: accumulators read in the mma-fragment layout ------------------------------------
// With one accumulator row per thread (conv_epilogue above) every coefficient is a warp-wide broadcast load from
// shared memory; at 32-64 channels a tile is only 18-36 tensor-core instructions long and those broadcasts (4 LSU
// wavefronts per LDS.128) outlast the MMAs several times over (ncu, round 1: 1017 wavefronts per 288-cycle tile).
// tcgen05.ld.16x256b hands each thread 4 pixels x C/4 channels instead, so a thread needs the coefficients of only
// C/4 <= 16 channels and keeps them in REGISTERS for as long as the CTA stays on one sample: the inner loop has no
// memory instruction at all. Thread t of warp quarter Q owns
//   rows   Q*32 + 16*h2 + 8*s + t/4          (h2, s in {0,1}; "rr" = 2*h2 + s)
//   channels 32*chunk + 8*j + 2*(t%4) + e    (j in 0..3, e in {0,1})
// The ToRGB sum over channels is finished with a two-step reduce-scatter over the four lanes that share a row
// (9 shuffles), which leaves row rr on lane t%4 == rr: every lane then adds the skip connection for ONE pixel.
// STAGED (folded up-convolution with C = 32 on 8-wide tiles only): the warp's 8 x 16 output pixels are written to its own
// 8 KB shared-memory tile and leave with ONE TMA store per tile instead of 32 sector-sized global stores per thread (the
// L1 store path takes one sector per clock; ablation, round 2: 0.36 of that layer's 0.99 ms were its global stores).
// The tile is addressed as 8 rows x 8 pixel PAIRS of 128 bytes under the 128-byte swizzle - `tmap_out` views the
// [B,H,W,32] tensor as [B,H,W/2,64] - which spreads a store instruction's 32 x 8 bytes over all banks twice (optimal).
// C > 64 (the 128 -> 128-channel layer): a thread's coefficients no longer fit its registers for the whole tile, so they are
// staged per sample in the group's shared-memory cache (`cache_all`, 2 * C float4 per group, as conv_epilogue lays them out)
// and the 8 channels of each 32-column chunk are loaded into registers at the top of the chunk: 16 LDS.128 per chunk with
// four distinct addresses per warp (one wavefront) - against one four-wavefront BROADCAST LDS.128 per channel and per
// accumulator element in the row-per-thread epilogue, which kept that layer's shared-memory pipe at 90 % (ncu, round 2).
// The same staging also serves the 64-channel fused-ToRGB flavour (COEF_SMEM): 48 instead of 96 coefficient registers let it
// keep the next tile's operands in registers and run a third epilogue group.
template <typename T, int N_TILE, int C, bool UP, bool RGB, int GROUPS, int TILE_W_LOG2 = -1, bool PAIR = false, int APG = 1,
          bool STAGED = false, bool COEF_SMEM = (C > 64), bool ROWS2 = false>
__device__ __forceinline__ void conv_epilogue_frag(const ConvParams& p, int group, int warp, int lane, int total_tiles,
                                                   uint32_t tmem_base, uint64_t* tmem_full_bar,
                                                   uint64_t* tmem_empty_bar, uint8_t* stage_tile = nullptr,
                                                   const CUtensorMap* tmap_out = nullptr, float4* cache_all = nullptr) {
    constexpr bool kCoefSmem = COEF_SMEM;
    static_assert(C <= 64 || COEF_SMEM, "more than 64 channels need the shared-memory coefficient cache");
    static_assert(!STAGED || (UP && C == 32 && TILE_W_LOG2 == 3 && !RGB), "staged output: folded up-convolution, 32 channels");
    const uint32_t stage_addr = STAGED ? smem_u32(stage_tile) : 0u;
    // PAIR: `total_tiles` counts pair items (the host guarantees an even m-tile count: no ghost tiles here), walked
    // cluster by cluster; the drained accumulator is reported to the leader's barrier
    const int pair_rank = PAIR ? (int)cluster_ctarank() : 0;
    const int walk_first = PAIR ? (int)cluster_id_x() : (int)blockIdx.x;
    const int walk_stride = PAIR ? (int)cluster_count_x() : (int)gridDim.x;
    const uint32_t leader_empty_base = PAIR ? cluster_map(smem_u32(tmem_empty_bar), 0) : 0u;
    auto decode = [&](int tile) {
        bool ghost;
        return PAIR ? decode_pair_item(p, tile, pair_rank, N_TILE, ghost) : decode_tile(p, tile, N_TILE);
    };
    static_assert(C % 32 == 0 && C <= 128, "register-resident coefficients: at most 16 channels per thread, or per chunk from shared memory");
    static_assert(N_TILE == (UP ? 4 * C : (ROWS2 ? 2 * C : C)), "GEMM columns = channels (x 4 phases for the folded up-convolution, x 2 image rows for ROWS2)");
    static_assert(!ROWS2 || (!UP && !STAGED && !PAIR && TILE_W_LOG2 >= 0), "ROWS2: plain convolution of the single-CTA halo kernel");
    static_assert(!kCoefSmem || !UP, "the shared-memory coefficient cache serves plain convolutions");
    constexpr int CPT = kCoefSmem ? 8 : C / 4;  // coefficient registers: all of the thread's channels, or one chunk's
    // ROWS2 (modconv3x3_halo2_kernel): accumulator row (hl, wl) holds the pixels (2 hl, wl) in columns [0, C) and (2 hl + 1, wl)
    // in columns [C, 2C) - the tile is drained as SUB = 2 sub-tiles, each exactly like a tile of the one-row kernel
    constexpr int SUB = ROWS2 ? 2 : 1;
    constexpr int CHUNKS = N_TILE / 32 / SUB;  // 32-column accumulator chunks per (sub-)tile
    auto image_row = [&](int h_local, int py) { return ROWS2 ? 2 * h_local + py : h_local; };
    const int quarter = warp & 3;
    const int q4 = lane & 3, r8 = lane >> 2;
    const int out_w = UP ? 2 * p.w_out : p.w_out, out_h = UP ? 2 * p.h_out : p.h_out;
#ifdef GANCE_ABL_NOSTORE
    const bool has_out = p.out != nullptr && p.batch < 0;  // ablation: no activation stores
#else
    const bool has_out = p.out != nullptr;
#endif

    float cd[CPT], cb[CPT], cs[CPT], cw0[CPT], cw1[CPT], cw2[CPT];
    float3 rgb_bias = make_float3(0.f, 0.f, 0.f);
    if (RGB) rgb_bias = make_float3(__ldg(p.rgb_bias), __ldg(p.rgb_bias + 1), __ldg(p.rgb_bias + 2));
    // tile-local pixel of each of this thread's rows (the tile is tile_w x tile_h pixels, w fastest)
    int wl[4], hl[4];
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
        const int row = quarter * 32 + 16 * (rr >> 1) + 8 * (rr & 1) + r8;
        if (TILE_W_LOG2 >= 0) {  // compile-time tile width (halo kernels: 8): the four rows of a thread become constant offsets
            wl[rr] = row & ((1 << TILE_W_LOG2) - 1);
            hl[rr] = row >> TILE_W_LOG2;
        } else {
            wl[rr] = row & (p.tile_w - 1);
            hl[rr] = row >> p.tile_w_log2;
        }
    }
    int cached_b = -1;
    int visit = 0;  // tiles this group has handled (accumulator buffer and barrier parity follow from it, see conv_epilogue)

    // Noise and skip-connection values of a tile do not depend on its accumulator: they are requested one tile AHEAD
    // (while the current tile is being computed) so their global-memory latency never sits on the tile's critical path.
    constexpr int NP = UP ? 4 : 1;  // noise values per input pixel
    constexpr bool kAheadInRegisters = !(RGB && C == 64 && !kCoefSmem);  // see the comment above `prefetch` below
    struct Ahead {
        TileCoord t;
        float nz[4][NP];
        float4 p00, p01, p10, p11;  // y_prev taps of this lane's own pixel (RGB)
        float w01, w10, w11;
        float3 skip;                // prefetch flavour: the four taps already combined (3 registers instead of 19)
        int own_h, own_w;
    };
    auto request = [&](int tile, int py) {
        Ahead n;
        n.t = decode(tile);
#ifdef GANCE_ABL_NOLOADS
        if (p.batch > 0) {  // ablation: no noise / skip fetches
#pragma unroll
            for (int rr = 0; rr < 4; ++rr)
#pragma unroll
                for (int ph = 0; ph < NP; ++ph) n.nz[rr][ph] = 0.f;
            n.p00 = n.p01 = n.p10 = n.p11 = make_float4(0.f, 0.f, 0.f, 0.f);
            n.w01 = n.w10 = n.w11 = 0.f;
            n.skip = make_float3(0.f, 0.f, 0.f);
            n.own_h = n.t.h0 + image_row(q4 == 0 ? hl[0] : q4 == 1 ? hl[1] : q4 == 2 ? hl[2] : hl[3], py);
            n.own_w = n.t.w0 + (q4 == 0 ? wl[0] : q4 == 1 ? wl[1] : q4 == 2 ? wl[2] : wl[3]);
            return n;
        }
#endif
        const float* noise_b = p.noise ? p.noise + (size_t)n.t.b0 * p.noise_batch_stride : nullptr;
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) {
            const int h = n.t.h0 + image_row(hl[rr], py), w = n.t.w0 + wl[rr];
            if (UP) {
                float2 top = make_float2(0.f, 0.f), bot = top;
                if (noise_b != nullptr) {
                    top = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)(2 * h) * out_w + 2 * w));
                    bot = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)(2 * h + 1) * out_w + 2 * w));
                }
                n.nz[rr][0] = top.x; n.nz[rr][1 % NP] = top.y; n.nz[rr][2 % NP] = bot.x; n.nz[rr][3 % NP] = bot.y;
            } else {
                // unconditional load from a valid address (the bias vector when there is no noise, multiplied by a zero
                // strength below): `noise_b ? __ldg(..) : 0` became LDG + FSEL, and the select waited out the whole fetch
                // latency right here instead of one tile later (ncu, round 2: 23 % of this kernel's stall samples)
                n.nz[rr][0] = __ldg(noise_b ? noise_b + (size_t)h * out_w + w : p.bias);
            }
        }
        n.p00 = n.p01 = n.p10 = n.p11 = make_float4(0.f, 0.f, 0.f, 0.f);
        n.w01 = n.w10 = n.w11 = 0.f;
        n.skip = make_float3(0.f, 0.f, 0.f);
        n.own_h = n.own_w = 0;
        if (RGB) {  // this lane finishes the pixel of row rr == q4
            n.own_h = n.t.h0 + image_row(q4 == 0 ? hl[0] : q4 == 1 ? hl[1] : q4 == 2 ? hl[2] : hl[3], py);
            n.own_w = n.t.w0 + (q4 == 0 ? wl[0] : q4 == 1 ? wl[1] : q4 == 2 ? wl[2] : wl[3]);
            if (p.y_prev != nullptr) {
                // upsample_2d(k=[1,3,3,1]): y[2a] = .25 p[a-1] + .75 p[a]; y[2a+1] = .75 p[a] + .25 p[a+1]; taps outside
                // the image contribute zero - fetched anyway (clamped) so the four loads are unconditional
                const int hp = p.h_out >> 1, wp = p.w_out >> 1;
                const int a = n.own_h >> 1, c = n.own_w >> 1;
                const int a2 = (n.own_h & 1) ? a + 1 : a - 1, c2 = (n.own_w & 1) ? c + 1 : c - 1;
                const bool a_in = a2 >= 0 && a2 < hp, c_in = c2 >= 0 && c2 < wp;
                const int a2c = a_in ? a2 : a, c2c = c_in ? c2 : c;
                const float4* base = reinterpret_cast<const float4*>(p.y_prev) + (size_t)n.t.b0 * hp * wp;
                // Three live components per tap (LDG.64 + LDG.32), not one float4: the unused fourth component left a dead
                // destination register that ptxas re-used a few instructions later - a write-after-write hazard that waited out
                // the whole fetch latency right here instead of one (sub-)tile later (ncu: 7-12 % of the stall samples of the
                // 1024^2 Conv1 on one LDC). A/B in one run: 1.085 -> 1.034 ms for that layer, 0.496 -> 0.489 ms for 256^2 Conv1.
                auto tap3 = [&](size_t index) {
                    const float2 xy = __ldg(reinterpret_cast<const float2*>(base + index));
                    const float z = __ldg(reinterpret_cast<const float*>(base + index) + 2);
                    return make_float4(xy.x, xy.y, z, 0.f);
                };
                n.p00 = tap3((size_t)a * wp + c);
                n.p01 = tap3((size_t)a * wp + c2c);
                n.p10 = tap3((size_t)a2c * wp + c);
                n.p11 = tap3((size_t)a2c * wp + c2c);
                n.w01 = c_in ? 0.1875f : 0.f;
                n.w10 = a_in ? 0.1875f : 0.f;
                n.w11 = (a_in && c_in) ? 0.0625f : 0.f;
                if (!kAheadInRegisters) {  // same expression, same order as the register flavour evaluates at the end
                    n.skip.x = 0.5625f * n.p00.x + n.w01 * n.p01.x + n.w10 * n.p10.x + n.w11 * n.p11.x;
                    n.skip.y = 0.5625f * n.p00.y + n.w01 * n.p01.y + n.w10 * n.p10.y + n.w11 * n.p11.y;
                    n.skip.z = 0.5625f * n.p00.z + n.w01 * n.p01.z + n.w10 * n.p10.z + n.w11 * n.p11.z;
                }
            }
        }
        return n;
    };

    // With 64 channels and the fused ToRGB (96 coefficient registers per thread) the values fetched one tile ahead do not
    // fit the register file: ptxas spilled them, and a spill STORE waits for the load it spills - the whole fetch latency
    // landed back on the critical path (ncu: 14 % of all stall samples on three STL instructions). That flavour therefore
    // only PREFETCHES the next tile's operands into L2 and loads them at the top of its own iteration, where the L2-hit
    // latency hides behind the wait for the accumulator.
    auto prefetch = [&](int next) {
        const TileCoord n = decode(next);
        if (p.noise != nullptr) {
            const float* noise_b = p.noise + (size_t)n.b0 * p.noise_batch_stride;
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const int h = n.h0 + hl[rr], w = n.w0 + wl[rr];
                const float* address = UP ? noise_b + (size_t)(2 * h) * out_w + 2 * w : noise_b + (size_t)h * out_w + w;
                asm volatile("prefetch.global.L2 [%0];" ::"l"(address));
                if (UP) asm volatile("prefetch.global.L2 [%0];" ::"l"(address + out_w));
            }
        }
        if (RGB && p.y_prev != nullptr) {
            const int own_h = n.h0 + (q4 == 0 ? hl[0] : q4 == 1 ? hl[1] : q4 == 2 ? hl[2] : hl[3]);
            const int own_w = n.w0 + (q4 == 0 ? wl[0] : q4 == 1 ? wl[1] : q4 == 2 ? wl[2] : wl[3]);
            const int hp = p.h_out >> 1, wp = p.w_out >> 1;
            const int a = own_h >> 1, c = own_w >> 1;
            const int a2 = min(max((own_h & 1) ? a + 1 : a - 1, 0), hp - 1);
            const float4* base = reinterpret_cast<const float4*>(p.y_prev) + (size_t)n.b0 * hp * wp;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (size_t)a * wp + c));   // the +-1 column shares the 32-byte sector
            asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (size_t)a2 * wp + c));  // half of the time, the line always
        }
    };

#ifdef GANCE_PIPE_STATS
    long long stat_wait = 0;
    const long long stat_begin = clock64();
#endif
    int tile = walk_first + group * walk_stride;
    Ahead ahead;
    static_assert(!ROWS2 || kAheadInRegisters, "ROWS2 keeps the next sub-tile's operands in registers");
    if (kAheadInRegisters && tile < total_tiles) ahead = request(tile, 0);
    for (; tile < total_tiles; tile += GROUPS * walk_stride, ++visit) {
        const int acc = group + GROUPS * (visit % APG);
        const uint32_t full_phase = (uint32_t)(visit / APG) & 1u;
        const bool more = tile + GROUPS * walk_stride < total_tiles;
#pragma unroll
      for (int py = 0; py < SUB; ++py) {
        if (!kAheadInRegisters) {
            ahead = request(tile, py);
            if (more) prefetch(tile + GROUPS * walk_stride);
        }
        const Ahead cur = ahead;
        const TileCoord t = cur.t;
        const int b = t.b0;
        if (kAheadInRegisters) {
            if (py + 1 < SUB) ahead = request(tile, py + 1);
            else if (more) ahead = request(tile + GROUPS * walk_stride, 0);
        }
        if (kCoefSmem && b != cached_b) {  // group-uniform: per-sample coefficients of every channel -> the group's cache
            float4* coef_a = cache_all + group * 2 * C;
            float4* coef_b = coef_a + C;
            epilogue_bar_sync(group);  // everyone is done reading the previous sample's
            for (int i = (warp & 3) * 32 + lane; i < C; i += kEpilogueThreads) {
                const size_t bc = (size_t)b * C + i;
                coef_a[i] = make_float4(__ldg(p.demod + bc), __ldg(p.bias + i),
                                        (p.next_styles ? __ldg(p.next_styles + bc) : 1.0f) * kLreluGain, 0.f);
                if (RGB) {
                    const float rs = __ldg(p.rgb_styles + bc) * kLreluGain;
                    coef_b[i] = make_float4(rs * __ldg(p.rgb_weight + 3 * i), rs * __ldg(p.rgb_weight + 3 * i + 1),
                                            rs * __ldg(p.rgb_weight + 3 * i + 2), 0.f);
                }
            }
            epilogue_bar_sync(group);
            cached_b = b;
        }
        if (!kCoefSmem && b != cached_b) {  // per-sample coefficients of this thread's channels -> registers
#pragma unroll
            for (int k = 0; k < CPT; ++k) {
                const int ch = (k >> 3) * 32 + ((k >> 1) & 3) * 8 + 2 * q4 + (k & 1);
                const size_t bc = (size_t)b * C + ch;
                cd[k] = __ldg(p.demod + bc);
                cb[k] = __ldg(p.bias + ch);
                cs[k] = (p.next_styles ? __ldg(p.next_styles + bc) : 1.0f) * kLreluGain;
                if (RGB) {
                    const float s = __ldg(p.rgb_styles + bc) * kLreluGain;
                    cw0[k] = s * __ldg(p.rgb_weight + 3 * ch);
                    cw1[k] = s * __ldg(p.rgb_weight + 3 * ch + 1);
                    cw2[k] = s * __ldg(p.rgb_weight + 3 * ch + 2);
                }
            }
            cached_b = b;
        }
        float nz[4][NP];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
#pragma unroll
            for (int ph = 0; ph < NP; ++ph)  // rounded product (never contracted into the bias add: every flavour of the epilogue
                nz[rr][ph] = __fmul_rn(cur.nz[rr][ph], p.noise ? p.noise_strength : 0.f);  // must round the same way)
        const int own_h = cur.own_h, own_wp = cur.own_w;
        float rgb[4][3];
#pragma unroll
        for (int rr = 0; rr < 4; ++rr) rgb[rr][0] = rgb[rr][1] = rgb[rr][2] = 0.f;

        if (STAGED) {  // the previous tile's TMA store has read the staging tile
            if (lane == 0) tma_store_wait_read();
            __syncwarp();
        }
#ifdef GANCE_PIPE_STATS
        const long long stat_w0 = clock64();
#endif
        if (py == 0) mbar_wait(tmem_full_bar + acc, full_phase);
#ifdef GANCE_PIPE_STATS
        stat_wait += clock64() - stat_w0;
#endif
        tcgen05_fence_after();
#ifdef GANCE_ABL_NOEPI
        if (p.batch > 0) {  // ablation: the epilogue only hands the accumulator back
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0 && py == SUB - 1) {
                if (PAIR) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
                else mbar_arrive(tmem_empty_bar + acc);
            }
            continue;
        }
#endif
#pragma unroll
        for (int chunk = 0; chunk < CHUNKS; ++chunk) {
            const int phase = UP ? (chunk * 32) / C : 0;
            const int kbase = kCoefSmem ? 0 : ((chunk * 32) % C) / 4;  // first of this thread's channels inside the chunk
            if (kCoefSmem) {  // this chunk's 8 channels: 32 * chunk + 8 j + 2 q4 + e
                const uint32_t base = smem_u32(cache_all + group * 2 * C);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int ch = chunk * 32 + (k >> 1) * 8 + 2 * q4 + (k & 1);
                    const float4 a = lds128(base + (uint32_t)ch * 16u);
                    cd[k] = a.x; cb[k] = a.y; cs[k] = a.z;
                    if (RGB) {
                        const float4 w = lds128(base + (uint32_t)(C + ch) * 16u);
                        cw0[k] = w.x; cw1[k] = w.y; cw2[k] = w.z;
                    }
                }
            }
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
                uint32_t r[16];
                tmem_ld_16x256b_x4(tmem_base + ((uint32_t)(quarter * 32 + 16 * h2) << 16) +
                                       (uint32_t)(acc * N_TILE + py * C * (ROWS2 ? 1 : 0) + chunk * 32), r);
                tmem_ld_wait();
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    const int rr = 2 * h2 + s;
                    const float nzv = nz[rr][UP ? phase : 0];
                    float v[8];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int k = kbase + 2 * j + e;
                            float o = fmaf(__uint_as_float(r[4 * j + 2 * s + e]), cd[k], nzv + cb[k]);
                            o = fmaxf(o, o * kLreluAlpha);
                            if (RGB) {
                                rgb[rr][0] = fmaf(o, cw0[k], rgb[rr][0]);
                                rgb[rr][1] = fmaf(o, cw1[k], rgb[rr][1]);
                                rgb[rr][2] = fmaf(o, cw2[k], rgb[rr][2]);
                            }
                            v[2 * j + e] = o * cs[k];
                        }
                    }
                    if (STAGED) {
                        // tile-local output pixel (2 * rr + py, 2 * r8 + px): row of the staging tile = 8 * (2 rr + py) + r8,
                        // 16-byte chunk = 4 * px + channel group, XOR-swizzled by the row's low three bits (= r8)
                        const bool odd = (q4 & 1) != 0;
                        const uint32_t row_addr = stage_addr + (uint32_t)((8 * (2 * rr + (phase >> 1)) + r8) * 128 + 4 * (q4 & 2));
#pragma unroll
                        for (int j = 0; j < 4; j += 2) {
                            const uint32_t w0 = pack_pair(v[2 * j], v[2 * j + 1], (T*)nullptr);
                            const uint32_t w1 = pack_pair(v[2 * j + 2], v[2 * j + 3], (T*)nullptr);
                            const uint32_t got = __shfl_xor_sync(0xffffffffu, odd ? w0 : w1, 1);
                            const uint32_t c16 = (uint32_t)((4 * (phase & 1) + j + (odd ? 1 : 0)) ^ r8);
                            asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(row_addr + (c16 << 4)), "r"(odd ? got : w0),
                                         "r"(odd ? w1 : got)
                                         : "memory");
                        }
                    } else if (has_out) {
                        const int h = t.h0 + image_row(hl[rr], py), w = t.w0 + wl[rr];
                        const int oh = UP ? 2 * h + (phase >> 1) : h, ow = UP ? 2 * w + (phase & 1) : w;
                        // The fragment layout leaves a lane with one channel PAIR (4 bytes) of each of four 8-channel groups.
                        // Stored as they are, a warp store covers 16 bytes in each of 8 pixels - half a 32-byte sector at a
                        // time, and the L1 store path handles one sector per clock whether it is full or not (ablation,
                        // round 2: the 4-byte stores were 0.49 of the 1.12 ms of the 1024^2 up-convolution). Lane pairs
                        // swap one word so that the even lane holds 8 contiguous bytes of group j, the odd lane of
                        // group j + 1: every store instruction then writes whole sectors.
                        const bool odd = (q4 & 1) != 0;
                        T* dst = reinterpret_cast<T*>(p.out) + (((size_t)b * out_h + oh) * out_w + ow) * C +
                                 ((chunk * 32) % C) + 2 * (q4 & 2);
#pragma unroll
                        for (int j = 0; j < 4; j += 2) {
                            const uint32_t w0 = pack_pair(v[2 * j], v[2 * j + 1], (T*)nullptr);
                            const uint32_t w1 = pack_pair(v[2 * j + 2], v[2 * j + 3], (T*)nullptr);
                            const uint32_t got = __shfl_xor_sync(0xffffffffu, odd ? w0 : w1, 1);
                            const uint2 both = odd ? make_uint2(got, w1) : make_uint2(w0, got);
                            *reinterpret_cast<uint2*>(dst + 8 * (j + (odd ? 1 : 0))) = both;
                        }
                    }
                }
            }
        }
        tcgen05_fence_before();
        if (STAGED) fence_proxy_async_smem();  // this thread's st.shared -> visible to the TMA unit
        __syncwarp();
        if (lane == 0 && py == SUB - 1) {  // accumulator drained
            if (PAIR) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
            else mbar_arrive(tmem_empty_bar + acc);
            if (STAGED) {  // the warp's 8 output rows x 16 pixels x 32 channels: rows 2 * (h0 + 4 * quarter) .., pixel pair w0 ..
                tma_store_4d(tmap_out, stage_tile, 0, t.w0, 2 * (t.h0 + 4 * quarter), b);
                tma_store_commit();
            }
        }
        if (RGB) {
            // reduce-scatter over the 4 lanes of a row group: xor 1 settles s, xor 2 settles h2
            float keep[2][3], mine[3];
            const bool s_own = (q4 & 1) != 0, h_own = (q4 & 2) != 0;
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float send = s_own ? rgb[2 * h2][c] : rgb[2 * h2 + 1][c];
                    keep[h2][c] = (s_own ? rgb[2 * h2 + 1][c] : rgb[2 * h2][c]) + __shfl_xor_sync(0xffffffffu, send, 1);
                }
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float send = h_own ? keep[0][c] : keep[1][c];
                mine[c] = (h_own ? keep[1][c] : keep[0][c]) + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            float red, green, blue;
            if (kAheadInRegisters) {
                red = mine[0] + rgb_bias.x + 0.5625f * cur.p00.x + cur.w01 * cur.p01.x + cur.w10 * cur.p10.x + cur.w11 * cur.p11.x;
                green = mine[1] + rgb_bias.y + 0.5625f * cur.p00.y + cur.w01 * cur.p01.y + cur.w10 * cur.p10.y + cur.w11 * cur.p11.y;
                blue = mine[2] + rgb_bias.z + 0.5625f * cur.p00.z + cur.w01 * cur.p01.z + cur.w10 * cur.p10.z + cur.w11 * cur.p11.z;
            } else {
                red = mine[0] + rgb_bias.x + cur.skip.x;
                green = mine[1] + rgb_bias.y + cur.skip.y;
                blue = mine[2] + rgb_bias.z + cur.skip.z;
            }
            const size_t pixel = ((size_t)b * p.h_out + own_h) * p.w_out + own_wp;
            if (p.y_out != nullptr) reinterpret_cast<float4*>(p.y_out)[pixel] = make_float4(red, green, blue, 0.f);
#ifdef GANCE_ABL_NOFRAMES
            if (p.frames != nullptr && red == 12345.678f) {  // ablation: (practically) no frame stores
#else
            if (p.frames != nullptr) {
#endif
                uint8_t* f = p.frames + pixel * 3;
                f[0] = to_u8(red); f[1] = to_u8(green); f[2] = to_u8(blue);
                flag_nonfinite(p.nonfinite, red, green, blue);
            }
        }
      }  // py
    }
    if (STAGED && lane == 0) tma_store_wait_read();  // shared memory must outlive the last store's read
#ifdef GANCE_PIPE_STATS
    if (!PAIR && p.pipe_stats != nullptr && (warp & 3) == 0 && lane == 0 && group < 2) {
        p.pipe_stats[blockIdx.x * 8 + 5 + group] = stat_wait;  // epilogue group: waiting for a finished accumulator
        if (group == 0) p.pipe_stats[blockIdx.x * 8 + 7] = clock64() - stat_begin;  // ... whole epilogue loop
    }
#endif
}

template <typename T, int N_TILE, int BLOCK_K, int STAGES, bool UP, bool RGB>
__global__ void __launch_bounds__(kThreads, 1)
modconv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                     const ConvParams p) {
    using L = SmemLayout<N_TILE, BLOCK_K, STAGES>;
    constexpr int kSwizzle = BLOCK_K * 2;
    constexpr uint32_t kTmemCols = 2 * N_TILE < 32 ? 32 : 2 * N_TILE;
    static_assert((kTmemCols & (kTmemCols - 1)) == 0 && kTmemCols <= 512, "TMEM columns must be a power of two");
    static_assert(L::kABytes % 1024 == 0 && L::kBBytes % 1024 == 0, "stage buffers must stay 1024-byte aligned");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // shfl result = provably warp-uniform
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
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

    const int total_tiles = p.tiles_b * p.tiles_h * p.tiles_w * p.n_tiles;
    const int k_iters = 9 * p.k_chunks;

    // Producer and MMA warps run their loops with all 32 lanes converged and elect one lane only around the
    // asynchronous instructions: every operand is then warp-uniform for the compiler (uniform registers, no
    // per-instruction R2UR "waterfall" loop - which cost ~130 cycles per tcgen05.mma in the first version).
    if (warp == 0) {
        const bool leader = elect_one_sync();
        {
            // ===== TMA producer =====
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const TileCoord t = decode_tile(p, tile, N_TILE);
                // K is walked channel chunk by channel chunk, the nine taps inside: the order of every 3x3 kernel of this
                // library (generic, CTA-pair and halo flavours accumulate in the same order, so a frame's bits do not
                // depend on which flavour the batch size selects)
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    for (int tap = 0; tap < 9; ++tap) {
                        const int dy = tap / 3 - p.pad, dx = tap % 3 - p.pad;
                        mbar_wait(empty_bar + stage, phase ^ 1u);
                        uint8_t* a_dst = smem + stage * L::kStageBytes;
                        uint8_t* b_dst = a_dst + L::kABytes;
                        if (leader) {
                            mbar_expect_tx(full_bar + stage, (uint32_t)p.a_tx_bytes + L::kBBytes);
                            tma_load_4d(a_dst, &tmap_x, full_bar + stage, kc * BLOCK_K, t.w0 + dx, t.h0 + dy, t.b0);
                            tma_load_2d(b_dst, &tmap_w, full_bar + stage, tap * p.c_in + kc * BLOCK_K, t.n0);
                        }
                        __syncwarp();
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1u;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        const bool leader = elect_one_sync();
        {
            // ===== MMA issuer =====
            constexpr uint32_t idesc = make_instr_desc(kTileM, N_TILE, sizeof(T) == 2 && !std::is_same<T, __nv_bfloat16>::value);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            const uint32_t smem_base = smem_u32(smem);
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);  // epilogue has drained this accumulator
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * N_TILE);
                for (int it = 0; it < k_iters; ++it) {
                    mbar_wait(full_bar + stage, phase);  // TMA bytes have landed
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_base + (uint32_t)(stage * L::kStageBytes);
                    const uint64_t a_desc = make_smem_desc<kSwizzle>(a_addr);
                    const uint64_t b_desc = a_desc + (uint64_t)(L::kABytes >> 4);
                    if (leader) {
#pragma unroll
                        for (int k = 0; k < BLOCK_K / kUmmaK; ++k) {
                            // advance along K inside the swizzle atom: +32 bytes = +2 in the >>4 address field
                            umma_bf16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc,
                                      (it | k) != 0 ? 1u : 0u);
                        }
                        umma_commit(empty_bar + stage);  // frees the smem slot when these MMAs retire
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
        }
    } else {
        conv_epilogue<T, N_TILE, UP, RGB, kEpilogueGroups, kCacheFloat4PerGroup>(p, (warp - 2) >> 2, warp, lane, total_tiles, tmem_base, tmem_full_bar,
                                          tmem_empty_bar, reinterpret_cast<float4*>(smem + L::kCacheOffset));
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---- CTA-pair flavour of the kernel above (cta_group::2) -----------------------------------------------------------------
// At N = 128 one 128 x N x 16 MMA reads 4 KB of A + 4 KB of B from shared memory (64 clk at 128 B/clk) for 64 clk of math:
// the operand port ties with the tensor pipe and neither is ever saturated (ncu, round 2: tensor pipe 54 % at N = 128,
// 76 % at N = 256). Two CTAs on the two SMs of a TPC share ONE M = 256 MMA: each supplies its own 128 pixels of A and
// HALF of the weight tile, so every SM reads A + B/2 (48 clk at N = 128) and fetches half the weight bytes through L2.
// Same roles as above in both CTAs; the leader's (rank 0) MMA warp issues for the pair, the peer's MMA warp only owns its
// CTA's TMEM allocation. Barriers: full[] lives in the leader (both CTAs' TMA bytes are counted there), empty[] and
// tmem_full[] are signalled in both CTAs by multicast commits, tmem_empty[] lives in the leader and collects the
// epilogue warps of both CTAs.
template <int N_TILE, int BLOCK_K, int STAGES>
struct PairSmemLayout {
    static constexpr int kABytes = kTileM * BLOCK_K * 2;
    static constexpr int kBBytes = (N_TILE / 2) * BLOCK_K * 2;  // this CTA's half of the weight tile
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kBarrierOffset = STAGES * kStageBytes;
    static constexpr int kCacheOffset = kBarrierOffset + (2 * STAGES + 4) * 8 + 16;
    static constexpr int kBytes = kCacheOffset + kCacheBytes + 1024;
};

template <typename T, int N_TILE, int BLOCK_K, int STAGES, bool RGB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
modconv3x3_pair_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                       const ConvParams p) {
    using L = PairSmemLayout<N_TILE, BLOCK_K, STAGES>;
    constexpr int kSwizzle = BLOCK_K * 2;
    constexpr uint32_t kTmemCols = 2 * N_TILE;
    static_assert((kTmemCols & (kTmemCols - 1)) == 0 && kTmemCols <= 512 && kTmemCols >= 32, "TMEM columns must be a power of two");
    static_assert(L::kABytes % 1024 == 0 && L::kBBytes % 1024 == 0, "stage buffers must stay 1024-byte aligned");
    static_assert(N_TILE % 32 == 0 && N_TILE <= 256, "cta_group::2: N in steps of 16 up to 256");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(full_bar + i, 1);   // the leader's expect_tx arrival; the bytes of both CTAs
            mbar_init(empty_bar + i, 1);  // one multicast commit
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 8);  // four epilogue warps of each CTA (used in the leader only)
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_base_slot, kTmemCols);
    tcgen05_fence_before();
    cluster_sync_all();  // barriers of BOTH CTAs are initialised before any remote arrival or TMA byte count
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);

    const int m_tiles = p.tiles_b * p.tiles_h * p.tiles_w;
    const int total_items = ((m_tiles + 1) >> 1) * p.n_tiles;
    const int k_iters = 9 * p.k_chunks;
    const int first_item = (int)cluster_id_x(), item_stride = (int)cluster_count_x();

    if (warp == 0) {
        // ===== TMA producer (both CTAs): own A tile + own half of the weight tile, bytes counted on the leader's barrier =====
        const bool leader_lane = elect_one_sync();
        int stage = 0;
        uint32_t phase = 0;
        const uint32_t full_leader = cluster_map(smem_u32(full_bar), 0);
        for (int item = first_item; item < total_items; item += item_stride) {
            bool ghost;
            const TileCoord t = decode_pair_item(p, item, rank, N_TILE, ghost);
            for (int kc = 0; kc < p.k_chunks; ++kc) {
                for (int tap = 0; tap < 9; ++tap) {
                    const int dy = tap / 3 - p.pad, dx = tap % 3 - p.pad;
                    mbar_wait(empty_bar + stage, phase ^ 1u);
                    uint8_t* a_dst = smem + stage * L::kStageBytes;
                    uint8_t* b_dst = a_dst + L::kABytes;
                    if (leader_lane) {
                        if (rank == 0) mbar_expect_tx(full_bar + stage, 2u * ((uint32_t)p.a_tx_bytes + L::kBBytes));
                        const uint32_t bar = full_leader + (uint32_t)(stage * 8);
                        tma_load_4d_pair(a_dst, &tmap_x, bar, kc * BLOCK_K, t.w0 + dx, t.h0 + dy, t.b0);
                        tma_load_2d_pair(b_dst, &tmap_w, bar, tap * p.c_in + kc * BLOCK_K, t.n0 + rank * (N_TILE / 2));
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
        if (rank == 0) {
            // ===== MMA issuer (leader CTA only) =====
            const bool leader_lane = elect_one_sync();
            constexpr uint32_t idesc = make_instr_desc(2 * kTileM, N_TILE, sizeof(T) == 2 && !std::is_same<T, __nv_bfloat16>::value);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            const uint32_t smem_base = smem_u32(smem);
            for (int item = first_item; item < total_items; item += item_stride) {
                mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);  // both CTAs' epilogues have drained this accumulator
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * N_TILE);
                for (int it = 0; it < k_iters; ++it) {
                    mbar_wait(full_bar + stage, phase);  // the TMA bytes of both CTAs have landed
                    tcgen05_fence_after();
                    const uint32_t a_addr = smem_base + (uint32_t)(stage * L::kStageBytes);
                    const uint64_t a_desc = make_smem_desc<kSwizzle>(a_addr);
                    const uint64_t b_desc = a_desc + (uint64_t)(L::kABytes >> 4);
                    if (leader_lane) {
#pragma unroll
                        for (int k = 0; k < BLOCK_K / kUmmaK; ++k)
                            umma_bf16_pair(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc,
                                           (it | k) != 0 ? 1u : 0u);
                        umma_commit_pair(empty_bar + stage);  // frees this stage in both CTAs
                        if (it == k_iters - 1) umma_commit_pair(tmem_full_bar + acc);
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
        }
    } else {
        conv_epilogue<T, N_TILE, false, RGB, kEpilogueGroups, kCacheFloat4PerGroup, true>(
            p, (warp - 2) >> 2, warp, lane, total_items, tmem_base, tmem_full_bar, tmem_empty_bar,
            reinterpret_cast<float4*>(smem + L::kCacheOffset));
    }

    // Neither CTA may leave (or free its TMEM) while its partner can still read its shared memory, write its TMEM or
    // signal its barriers.
    tcgen05_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// ====================================================================================================
// Halo variant for the narrow layers (Cin = 32 or 64, SAME padding): the layers of 512^2 and 1024^2.
//
// The kernel above re-fetches the activation tile once per filter tap (9x the input through L2) and the
// weight tile once per work item; with 32-64 channels that makes those layers L2-bound at ~1/8 of the
// tensor rate. Here
//   * the WHOLE weight matrix [9 taps][N][Cin] is loaded into shared memory once per CTA, and
//   * ONE TMA box per tile brings the (16+2) x (8+2)-pixel halo of a 16-row x 8-column output tile;
//     the A operand of tap (ky,kx) is the same buffer read through a shifted descriptor:
//     start += (ky*10 + kx) pixels, 8-row groups (= 8 consecutive pixels of one image row) spaced by the
//     halo pitch (SBO = 10 pixels). The swizzle XOR is a function of the shared-memory address bits for
//     both the TMA write and the UMMA read, so a start shifted by whole pixel rows stays consistent.
// L2 -> SM traffic per 128 output pixels drops from 9*128 to 180 pixel rows (6.4x) + no weight re-reads.
// ====================================================================================================
constexpr int kHaloTileW = 8, kHaloTileH = 16;
constexpr int kHaloW = kHaloTileW + 2, kHaloH = kHaloTileH + 2;
// A tile of these layers is 18-36 tensor-core instructions: four epilogue groups (= four TMEM accumulator buffers,
// four tiles in flight) hide the global-memory latency of the noise / skip-connection fetches behind one another.
// Measured with cycle counters in the warp roles (tools/pipe_stats.py, round 2): in all three instances the epilogue
// groups are busy > 90 % of the time while the MMA issuer waits 20-25 % for a drained accumulator - the epilogue warps are
// latency-bound (one instruction every ~5 cycles per warp), so the remedy is more of them where registers allow:
// the folded up-convolution (N = 128, 112 registers) runs four groups; the fused-ToRGB flavours need 128-168 registers.
// The 64-channel fused-ToRGB flavour was also tried with shared-memory coefficients (conv_epilogue_frag's COEF_SMEM: 48
// instead of 96 coefficient registers): with a third group it spills at the 128-register cap (775 vs 700 us for the 512^2
// Conv1), with two groups the per-chunk reloads cost more than the registers they free (733 us). Registers stay.
__host__ __device__ constexpr int halo_groups(int n_gemm) { return n_gemm == 128 ? 4 : (n_gemm == 32 ? 3 : 2); }  // bounded by registers (65536 / threads)
constexpr bool kHalo64CoefSmem = false;
__host__ __device__ constexpr int halo_threads(int n_gemm) { return 64 + halo_groups(n_gemm) * kEpilogueThreads; }
constexpr int kHaloApg = 2;  // TMEM accumulator buffers per epilogue group (see conv_epilogue); 1 where 512 columns are full
// The 32-channel layer (1024^2 Conv1) runs four buffers per group (12 x 32 of the 512 TMEM columns) and an eight-deep halo
// ring: its MMA pipeline alone and its epilogue alone each need ~0.82 of the kernel's time (ablations, round 2), so what is
// left to gain is their overlap - measured 8-14 % (three A/B pairs in one run). The 64-channel layer measured 4 % SLOWER
// with four buffers and six stages and keeps two and three.
__host__ __device__ constexpr int halo_apg(int n_gemm) { return n_gemm == 128 ? 1 : (n_gemm == 32 ? 4 : kHaloApg); }
constexpr int kHaloStages64 = 3, kHaloStages32 = 8;
__host__ __device__ constexpr int halo_accs(int n_gemm) { return halo_apg(n_gemm) * halo_groups(n_gemm); }

template <int N_GEMM, int CIN, int STAGES>
struct HaloSmem {
    static constexpr int kRowBytes = CIN * 2;
    static constexpr int kWeightBytes = 9 * N_GEMM * kRowBytes;
    static constexpr int kHaloBytes = kHaloH * kHaloW * kRowBytes;
    static constexpr int kHaloStride = (kHaloBytes + 1023) / 1024 * 1024;
    static constexpr int kBarrierOffset = kWeightBytes + STAGES * kHaloStride;
    static constexpr int kCacheOffset = kBarrierOffset + ((2 * STAGES + 2 * halo_accs(N_GEMM) + 1) * 8 + 16 + 15) / 16 * 16;
    static constexpr int kCacheBytes = N_GEMM == 64 ? halo_groups(N_GEMM) * 2 * N_GEMM * 16 : 0;  // coefficient cache of the 64-channel fused-ToRGB flavour
    static constexpr int kBytes = kCacheOffset + kCacheBytes + 1024;
};

template <int SWIZZLE_BYTES>
__device__ __forceinline__ uint64_t make_smem_desc_sbo(uint32_t smem_addr, uint32_t sbo_bytes) {
    constexpr uint64_t layout = SWIZZLE_BYTES == 128 ? 2ull : (SWIZZLE_BYTES == 64 ? 4ull : 6ull);
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) |
           (layout << 61);
}

template <typename T, int N_GEMM, int CIN, int STAGES, bool UP, bool RGB>
__global__ void __launch_bounds__(halo_threads(N_GEMM), 1)
modconv3x3_halo_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                       const ConvParams p) {
    using L = HaloSmem<N_GEMM, CIN, STAGES>;
    constexpr int kSwizzle = CIN * 2;  // 128 or 64 bytes per pixel row
    constexpr int kHaloGroups = halo_groups(N_GEMM);
    constexpr int kAccs = halo_accs(N_GEMM);
    constexpr uint32_t kTmemCols = kAccs * N_GEMM <= 128 ? 128 : (kAccs * N_GEMM <= 256 ? 256 : 512);
    static_assert(kAccs * N_GEMM <= 512, "TMEM columns");
    static_assert((kTmemCols & (kTmemCols - 1)) == 0 && kTmemCols <= 512, "TMEM columns must be a power of two");
    static_assert(CIN == 32 || CIN == 64, "one swizzled row per pixel");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_w = smem;
    uint8_t* smem_halo = smem + L::kWeightBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + kAccs;
    uint64_t* weight_bar = tmem_empty_bar + kAccs;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(weight_bar + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // shfl result = provably warp-uniform
    const int lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < kAccs; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 4);
        }
        mbar_init(weight_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_base_slot, kTmemCols);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);
    const int total_tiles = p.tiles_b * p.tiles_h * p.tiles_w;  // n_tiles == 1: every CTA computes all N_GEMM columns

    if (warp == 0) {
        const bool leader = elect_one_sync();
        {
            if (leader) {
                mbar_expect_tx(weight_bar, L::kWeightBytes);
                for (int tap = 0; tap < 9; ++tap)
                    tma_load_2d(smem_w + tap * N_GEMM * L::kRowBytes, &tmap_w, weight_bar, tap * CIN, 0);
            }
            __syncwarp();
            int stage = 0;
            uint32_t phase = 0;
#ifdef GANCE_PIPE_STATS
            long long waited = 0;
#endif
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const TileCoord t = decode_tile(p, tile, N_GEMM);
#ifdef GANCE_PIPE_STATS
                const long long w0 = clock64();
#endif
                mbar_wait(empty_bar + stage, phase ^ 1u);
#ifdef GANCE_PIPE_STATS
                waited += clock64() - w0;
#endif
                if (leader) {
                    mbar_expect_tx(full_bar + stage, L::kHaloBytes);
                    tma_load_4d(smem_halo + stage * L::kHaloStride, &tmap_x, full_bar + stage, 0, t.w0 - 1, t.h0 - 1, t.b0);
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
#ifdef GANCE_PIPE_STATS
            if (p.pipe_stats != nullptr && lane == 0) p.pipe_stats[blockIdx.x * 8 + 4] = waited;
#endif
        }
    } else if (warp == 1) {
        const bool leader = elect_one_sync();
        {
            constexpr uint32_t idesc = make_instr_desc(kTileM, N_GEMM, sizeof(T) == 2 && !std::is_same<T, __nv_bfloat16>::value);
            mbar_wait(weight_bar, 0);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            const uint32_t w_addr = smem_u32(smem_w);
            const uint64_t b_desc0 = make_smem_desc<kSwizzle>(w_addr);
#ifdef GANCE_PIPE_STATS
            long long wait_acc = 0, wait_data = 0;
            const long long begin = clock64();
#endif
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
#ifdef GANCE_PIPE_STATS
                const long long w0 = clock64();
#endif
                mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
#ifdef GANCE_PIPE_STATS
                const long long w1 = clock64();
#endif
                mbar_wait(full_bar + stage, phase);
#ifdef GANCE_PIPE_STATS
                wait_acc += w1 - w0;
                wait_data += clock64() - w1;
#endif
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * N_GEMM);
                // descriptors differ from these two only by constants: tap shift / K step, in 16-byte units
                const uint64_t a_desc0 = make_smem_desc_sbo<kSwizzle>(w_addr + (uint32_t)(L::kWeightBytes + stage * L::kHaloStride),
                                                                      kHaloW * L::kRowBytes);
                if (leader) {
#ifdef GANCE_ABL_NOMMA
                    if (p.batch < 0)  // ablation: no tensor-core work, the epilogue runs on whatever the accumulator holds
#endif
#pragma unroll
                    for (int tap = 0; tap < 9; ++tap) {
#ifdef GANCE_ABL_NOSWZ
                        // experiment (timing only, the data is NOT in this layout): A read as un-swizzled 8 x 16-byte core
                        // matrices, [K chunk][h][w][8 channels]: LBO = one K chunk (18 x 10 x 16 B), SBO = one halo row
                        const uint32_t halo_addr = w_addr + (uint32_t)(L::kWeightBytes + stage * L::kHaloStride);
                        const uint64_t a_desc = (uint64_t)(((halo_addr + (uint32_t)(((tap / 3) * kHaloW + (tap % 3)) * 16)) & 0x3FFFFu) >> 4) |
                                                ((uint64_t)((kHaloH * kHaloW * 16) >> 4) << 16) | ((uint64_t)((kHaloW * 16) >> 4) << 32) | (1ull << 46);
                        const uint64_t b_desc = b_desc0 + (uint64_t)((tap * N_GEMM * L::kRowBytes) >> 4);
#pragma unroll
                        for (int k = 0; k < CIN / kUmmaK; ++k)
                            umma_bf16(d_tmem, a_desc + (uint64_t)((k * 2 * kHaloH * kHaloW * 16) >> 4), b_desc + (uint64_t)(k * 2), idesc,
                                      (tap | k) != 0 ? 1u : 0u);
#else
                        const uint64_t a_desc = a_desc0 + (uint64_t)((((tap / 3) * kHaloW + (tap % 3)) * L::kRowBytes) >> 4);
                        const uint64_t b_desc = b_desc0 + (uint64_t)((tap * N_GEMM * L::kRowBytes) >> 4);
#pragma unroll
                        for (int k = 0; k < CIN / kUmmaK; ++k)
                            umma_bf16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc,
                                      (tap | k) != 0 ? 1u : 0u);
#endif
                    }
                    umma_commit(empty_bar + stage);
                    umma_commit(tmem_full_bar + acc);
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1u;
                }
                if (++acc == kAccs) {
                    acc = 0;
                    acc_phase ^= 1u;
                }
            }
#ifdef GANCE_PIPE_STATS
            if (p.pipe_stats != nullptr && lane == 0) {
                p.pipe_stats[blockIdx.x * 8 + 0] = clock64() - begin;  // MMA warp: whole loop (issue side; the tensor pipe runs behind)
                p.pipe_stats[blockIdx.x * 8 + 1] = wait_acc;           // ... of which waiting for a drained accumulator
                p.pipe_stats[blockIdx.x * 8 + 2] = wait_data;          // ... of which waiting for TMA data
            }
#endif
        }
    } else {
        conv_epilogue_frag<T, N_GEMM, (UP ? N_GEMM / 4 : N_GEMM), UP, RGB, kHaloGroups, 3, false, halo_apg(N_GEMM), false,
                           (kHalo64CoefSmem && RGB && N_GEMM == 64)>(
            p, (warp - 2) >> 2, warp, lane, total_tiles, tmem_base, tmem_full_bar, tmem_empty_bar, nullptr, nullptr,
            reinterpret_cast<float4*>(smem + L::kCacheOffset));
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---- two image rows per accumulator row: the 32 -> 32-channel layer (1024^2 Conv1) -------------------------------------------
// The one-row kernel above issues 18 MMAs of N = 32 per 128 pixels, and a 128 x 32 x 16 tcgen05.mma costs ~59 cycles whatever
// its N (the A operand is read from shared memory at the same rate; N = 64 measured 52, profiles/r02_ablation.txt): with no
// epilogue at all that layer needed 0.97 ms. Here accumulator row (g, w) of a 16 x 8 grid holds TWO output pixels, image rows
// 2g and 2g+1 of a 32 x 8-pixel tile, in columns [0,32) and [32,64). Input row 2g-1+r (r = 0..3) feeds output row 2g through
// filter row ky = r and output row 2g+1 through ky = r-1, so per (r, kx) ONE MMA multiplies the shifted halo view (8-row groups
// now spaced by TWO halo rows) with the 64 x 32 weight block [W(ky=r,kx) ; W(ky=r-1,kx)]:
//   r = 0: N = 32 into columns [0,32)        (only the even row sees input row 2g-1)
//   r = 1, 2: N = 64                         (weights are kept in shared memory as [kx][ky = 2,1,0] so the two blocks are adjacent)
//   r = 3: N = 32 into columns [32,64)
// = 12 MMAs of N = 64 + 12 of N = 32 (+1: the very first r = 1 MMA is split so that columns [32,64) start from zero) per
// 256 pixels instead of 36 of N = 32, with NO wasted MACs. Every output element accumulates (ky, kx, 16-channel step) in the
// order of the one-row kernel, so the results are bit-identical to it. The epilogue drains an accumulator as two sub-tiles
// (conv_epilogue_frag, ROWS2).
constexpr int kHalo2TileH = 32, kHalo2H = kHalo2TileH + 2;
constexpr int kHalo2Groups = 3, kHalo2Apg = 2, kHalo2Accs = kHalo2Groups * kHalo2Apg;  // 6 x 64 of the 512 TMEM columns
constexpr int kHalo2Threads = 64 + kHalo2Groups * kEpilogueThreads;
constexpr int kHalo2Stages = 4;                                                        // 4 x 256 pixels in flight (8 x 128 before)

template <int STAGES>
struct Halo2Smem {
    static constexpr int kRowBytes = 64;                        // 32 channels
    static constexpr int kBlockBytes = 32 * kRowBytes;          // one tap: 32 output channels x 32 input channels
    static constexpr int kWeightBytes = 9 * kBlockBytes;
    static constexpr int kHaloBytes = kHalo2H * kHaloW * kRowBytes;
    static constexpr int kHaloStride = (kHaloBytes + 1023) / 1024 * 1024;
    static constexpr int kBarrierOffset = (kWeightBytes + 1023) / 1024 * 1024 + STAGES * kHaloStride;
    static constexpr int kBytes = kBarrierOffset + (2 * STAGES + 2 * kHalo2Accs + 1) * 8 + 16 + 1024;
};

template <typename T, int STAGES, bool RGB>
__global__ void __launch_bounds__(kHalo2Threads, 1)
modconv3x3_halo2_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                        const ConvParams p) {
    using L = Halo2Smem<STAGES>;
    constexpr int CIN = 32, C = 32;
    constexpr uint32_t kTmemCols = 512;
    static_assert(kHalo2Accs * 2 * C <= 512, "TMEM columns");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_w = smem;
    uint8_t* smem_halo = smem + (L::kWeightBytes + 1023) / 1024 * 1024;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + kHalo2Accs;
    uint64_t* weight_bar = tmem_empty_bar + kHalo2Accs;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(weight_bar + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < kHalo2Accs; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 4);
        }
        mbar_init(weight_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_base_slot, kTmemCols);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);
    const int total_tiles = p.tiles_b * p.tiles_h * p.tiles_w;  // tiles of 32 rows x 8 columns

    if (warp == 0) {
        const bool leader = elect_one_sync();
        if (leader) {
            mbar_expect_tx(weight_bar, L::kWeightBytes);
            for (int tap = 0; tap < 9; ++tap)  // tap (ky, kx) -> block kx * 3 + (2 - ky)
                tma_load_2d(smem_w + ((tap % 3) * 3 + (2 - tap / 3)) * L::kBlockBytes, &tmap_w, weight_bar, tap * CIN, 0);
        }
        __syncwarp();
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const TileCoord t = decode_tile(p, tile, 2 * C);
            mbar_wait(empty_bar + stage, phase ^ 1u);
            if (leader) {
                mbar_expect_tx(full_bar + stage, L::kHaloBytes);
                tma_load_4d(smem_halo + stage * L::kHaloStride, &tmap_x, full_bar + stage, 0, t.w0 - 1, t.h0 - 1, t.b0);
            }
            __syncwarp();
            if (++stage == STAGES) {
                stage = 0;
                phase ^= 1u;
            }
        }
    } else if (warp == 1) {
        const bool leader = elect_one_sync();
        constexpr bool kHalfOperands = sizeof(T) == 2 && !std::is_same<T, __nv_bfloat16>::value;
        constexpr uint32_t idesc32 = make_instr_desc(kTileM, C, kHalfOperands);
        constexpr uint32_t idesc64 = make_instr_desc(kTileM, 2 * C, kHalfOperands);
        mbar_wait(weight_bar, 0);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        const uint32_t w_addr = smem_u32(smem_w);
        const uint32_t halo_addr0 = smem_u32(smem_halo);
        const uint64_t b_desc0 = make_smem_desc<64>(w_addr);
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
            mbar_wait(full_bar + stage, phase);
            tcgen05_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 2 * C);
            // 8-row groups of the A operand = 8 pixels of one halo row; consecutive groups are TWO halo rows apart
            const uint64_t a_desc0 = make_smem_desc_sbo<64>(halo_addr0 + (uint32_t)(stage * L::kHaloStride), 2 * kHaloW * L::kRowBytes);
            if (leader) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const uint64_t a_desc = a_desc0 + (uint64_t)(((r * kHaloW + kx) * L::kRowBytes) >> 4);
                        // weight blocks of this kx: [W(ky=2) | W(ky=1) | W(ky=0)]
                        const uint64_t b_kx = b_desc0 + (uint64_t)((kx * 3 * L::kBlockBytes) >> 4);
#pragma unroll
                        for (int k = 0; k < CIN / kUmmaK; ++k) {
                            const uint64_t a = a_desc + (uint64_t)(k * 2);
                            if (r == 0) {         // W(ky=0) -> even row
                                umma_bf16(d_tmem, a, b_kx + (uint64_t)((2 * L::kBlockBytes) >> 4) + (uint64_t)(k * 2), idesc32, (kx | k) != 0 ? 1u : 0u);
                            } else if (r == 3) {  // W(ky=2) -> odd row
                                umma_bf16(d_tmem + (uint32_t)C, a, b_kx + (uint64_t)(k * 2), idesc32, 1u);
                            } else if (r == 1 && kx == 0 && k == 0) {  // the odd row's first product: its columns start from zero
                                umma_bf16(d_tmem, a, b_kx + (uint64_t)((1 * L::kBlockBytes) >> 4), idesc32, 1u);
                                umma_bf16(d_tmem + (uint32_t)C, a, b_kx + (uint64_t)((2 * L::kBlockBytes) >> 4), idesc32, 0u);
                            } else {              // [W(ky=r) ; W(ky=r-1)] -> both rows
                                umma_bf16(d_tmem, a, b_kx + (uint64_t)(((2 - r) * L::kBlockBytes) >> 4) + (uint64_t)(k * 2), idesc64, 1u);
                            }
                        }
                    }
                }
                umma_commit(empty_bar + stage);
                umma_commit(tmem_full_bar + acc);
            }
            __syncwarp();
            if (++stage == STAGES) {
                stage = 0;
                phase ^= 1u;
            }
            if (++acc == kHalo2Accs) {
                acc = 0;
                acc_phase ^= 1u;
            }
        }
    } else {
        conv_epilogue_frag<T, 2 * C, C, false, RGB, kHalo2Groups, 3, false, kHalo2Apg, false, false, true>(
            p, (warp - 2) >> 2, warp, lane, total_tiles, tmem_base, tmem_full_bar, tmem_empty_bar);
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---- CTA-pair flavour of the halo kernel (cta_group::2), Cin = 64 or 128 ----------------------------------------------------
// Each CTA of the pair keeps HALF of the weight matrix resident ([K chunk][tap][N/2 rows][64 channels]), which
//   * halves the B-operand shared-memory reads per SM (N = 128: A 32 + B 32 clk per MMA -> 32 + 16 against 64 clk of math:
//     the folded 1024^2 up-convolution leaves the operand-port tie), and
//   * makes the 128 -> 128 channel layer (256^2 Conv1: 295 KB of weights) fit: 147 KB per CTA. That layer ran on the generic
//     kernel, which re-fetches every activation tile once per filter tap through L2 and sat at the L2 -> SM limit (ncu,
//     round 2: 7.3 GB through L2 in 640 us); here a tile's activations cross L2 once (two 18 x 10-pixel halo boxes).
// Halo boxes are per 64-channel K chunk (one 128-byte swizzled row per pixel); the ring is counted in K chunks.
template <int N_GEMM, int CIN, int STAGES>
struct HaloPairSmem {
    static constexpr int kChunks = CIN / 64;
    static constexpr int kRowBytes = 128;
    static constexpr int kTapBytes = (N_GEMM / 2) * kRowBytes;     // one tap, one K chunk, this CTA's half of the rows
    static constexpr int kWeightBytes = kChunks * 9 * kTapBytes;
    static constexpr int kHaloBytes = kHaloH * kHaloW * kRowBytes;
    static constexpr int kHaloStride = (kHaloBytes + 1023) / 1024 * 1024;
    static constexpr int kBarrierOffset = kWeightBytes + STAGES * kHaloStride;
    static constexpr int kAccs = kHaloApg * kEpilogueGroups;        // TMEM accumulator buffers
    static constexpr int kCacheOffset = kBarrierOffset + ((2 * STAGES + 2 * kAccs + 1) * 8 + 16 + 15) / 16 * 16;
    static constexpr int kCacheFloat4 = 2 * N_GEMM;                 // per epilogue group (row-per-thread epilogue only)
    // after the barriers: the coefficient cache of the row-per-thread epilogue (UP = false, 128 channels) or the staging
    // tiles of the TMA-stored output (folded up-convolution: 8 KB per epilogue warp, 1024-byte aligned)
    static constexpr int kStageTileBytes = 8 * 8 * 128;
    static constexpr int kStageOffset = (kCacheOffset + 1023) / 1024 * 1024;
    static constexpr int kBytesCache = kCacheOffset + kEpilogueGroups * kCacheFloat4 * 16 + 1024;
    static constexpr int kBytesStaged = kStageOffset + kEpilogueGroups * 4 * kStageTileBytes + 1024;
};

template <typename T, int N_GEMM, int CIN, int STAGES, bool UP, bool RGB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
modconv3x3_halo_pair_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                            const __grid_constant__ CUtensorMap tmap_out, const ConvParams p) {
    using L = HaloPairSmem<N_GEMM, CIN, STAGES>;
    constexpr int C = UP ? N_GEMM / 4 : N_GEMM;  // channels of the output tensor
    constexpr bool kFrag = C <= 64 || RGB;       // fragment-layout epilogue (the plain 128-channel flavour keeps the row-per-thread one: it can write the second, unscaled output)
    constexpr uint32_t kTmemCols = L::kAccs * N_GEMM;
    static_assert(CIN == 64 || CIN == 128, "64-channel K chunks");
    static_assert((kTmemCols & (kTmemCols - 1)) == 0 && kTmemCols <= 512 && kTmemCols >= 32, "TMEM columns must be a power of two");
    static_assert(L::kTapBytes % 1024 == 0, "weight blocks must stay 1024-byte aligned");
    constexpr bool kStaged = UP && C == 32;      // output through shared memory + TMA store (see conv_epilogue_frag)
    static_assert((kStaged ? L::kBytesStaged : L::kBytesCache) <= 227 * 1024, "shared memory");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_w = smem;
    uint8_t* smem_halo = smem + L::kWeightBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + L::kAccs;
    uint64_t* weight_bar = tmem_empty_bar + L::kAccs;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(weight_bar + 1);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        if (kStaged) prefetch_tmap(&tmap_out);
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < L::kAccs; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 8);  // four epilogue warps of each CTA (used in the leader only)
        }
        mbar_init(weight_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_base_slot, kTmemCols);
    tcgen05_fence_before();
    cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);
    const int total_items = (p.tiles_b * p.tiles_h * p.tiles_w) >> 1;  // even m-tile count (host)
    const int first_item = (int)cluster_id_x(), item_stride = (int)cluster_count_x();

    if (warp == 0) {
        const bool leader_lane = elect_one_sync();
        if (leader_lane) {  // this CTA's half of the weights; both halves are counted on the leader's barrier
            if (rank == 0) mbar_expect_tx(weight_bar, 2u * L::kWeightBytes);
            const uint32_t bar = cluster_map(smem_u32(weight_bar), 0);
            for (int kc = 0; kc < L::kChunks; ++kc)
                for (int tap = 0; tap < 9; ++tap)
                    tma_load_2d_pair(smem_w + (kc * 9 + tap) * L::kTapBytes, &tmap_w, bar, tap * CIN + kc * 64,
                                     rank * (N_GEMM / 2));
        }
        __syncwarp();
        int stage = 0;
        uint32_t phase = 0;
        const uint32_t full_leader = cluster_map(smem_u32(full_bar), 0);
        for (int item = first_item; item < total_items; item += item_stride) {
            bool ghost;
            const TileCoord t = decode_pair_item(p, item, rank, N_GEMM, ghost);
#pragma unroll
            for (int kc = 0; kc < L::kChunks; ++kc) {
                mbar_wait(empty_bar + stage, phase ^ 1u);
                if (leader_lane) {
                    if (rank == 0) mbar_expect_tx(full_bar + stage, 2u * L::kHaloBytes);
                    tma_load_4d_pair(smem_halo + stage * L::kHaloStride, &tmap_x, full_leader + (uint32_t)(stage * 8), kc * 64,
                                     t.w0 - 1, t.h0 - 1, t.b0);
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            const bool leader_lane = elect_one_sync();
            constexpr uint32_t idesc = make_instr_desc(2 * kTileM, N_GEMM, sizeof(T) == 2 && !std::is_same<T, __nv_bfloat16>::value);
            mbar_wait(weight_bar, 0);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            const uint32_t w_addr = smem_u32(smem_w);
            const uint64_t b_desc0 = make_smem_desc<128>(w_addr);
            for (int item = first_item; item < total_items; item += item_stride) {
                mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * N_GEMM);
#pragma unroll
                for (int kc = 0; kc < L::kChunks; ++kc) {
                    mbar_wait(full_bar + stage, phase);
                    tcgen05_fence_after();
                    const uint64_t a_desc0 = make_smem_desc_sbo<128>(w_addr + (uint32_t)(L::kWeightBytes + stage * L::kHaloStride),
                                                                     kHaloW * L::kRowBytes);
                    if (leader_lane) {
#pragma unroll
                        for (int tap = 0; tap < 9; ++tap) {
                            const uint64_t a_desc = a_desc0 + (uint64_t)((((tap / 3) * kHaloW + (tap % 3)) * L::kRowBytes) >> 4);
                            const uint64_t b_desc = b_desc0 + (uint64_t)(((kc * 9 + tap) * L::kTapBytes) >> 4);
#pragma unroll
                            for (int k = 0; k < 64 / kUmmaK; ++k)
                                umma_bf16_pair(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc,
                                               (kc | tap | k) != 0 ? 1u : 0u);
                        }
                        umma_commit_pair(empty_bar + stage);
                        if (kc == L::kChunks - 1) umma_commit_pair(tmem_full_bar + acc);
                    }
                    __syncwarp();
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
                if (++acc == L::kAccs) {
                    acc = 0;
                    acc_phase ^= 1u;
                }
            }
        }
    } else {
        if constexpr (kFrag)
            conv_epilogue_frag<T, N_GEMM, C, UP, RGB, kEpilogueGroups, 3, true, kHaloApg, kStaged>(
                p, (warp - 2) >> 2, warp, lane, total_items, tmem_base, tmem_full_bar, tmem_empty_bar,
                kStaged ? smem + L::kStageOffset + (warp - 2) * L::kStageTileBytes : nullptr, &tmap_out,
                reinterpret_cast<float4*>(smem + L::kCacheOffset));
        else
            conv_epilogue<T, N_GEMM, UP, RGB, kEpilogueGroups, L::kCacheFloat4, true, kHaloApg>(
                p, (warp - 2) >> 2, warp, lane, total_items, tmem_base, tmem_full_bar, tmem_empty_bar,
                reinterpret_cast<float4*>(smem + L::kCacheOffset));
    }

    tcgen05_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc_pair(tmem_base, kTmemCols);
    }
}

// ---- host side -------------------------------------------------------------------------------------
#ifdef GANCE_PIPE_STATS
static long long* g_pipe_stats = nullptr;  // [launch slot][CTA][8], advanced by every single-CTA halo launch
static int g_pipe_slot = 0;
#endif
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode_tiled = nullptr;
static int g_num_sms = 148;

template <typename K>
static int raise_smem_limit(K kernel, int bytes, const char* what) {
    cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (err != cudaSuccess)
        return fail(-(int)err, "cudaFuncSetAttribute(%s, %d B): %s", what, bytes, cudaGetErrorString(err));
    return GANCE_OK;
}

// Every (operand type, epilogue flavour) instance of one tile configuration: plain, folded up-convolution, fused ToRGB.
template <int N_TILE, int BLOCK_K, int STAGES>
static int configure() {
    using L = SmemLayout<N_TILE, BLOCK_K, STAGES>;
    int rc;
    if ((rc = raise_smem_limit(modconv3x3_tc_kernel<__nv_bfloat16, N_TILE, BLOCK_K, STAGES, false, false>, L::kBytes, "modconv3x3_tc")) != GANCE_OK) return rc;
    if ((rc = raise_smem_limit(modconv3x3_tc_kernel<__half, N_TILE, BLOCK_K, STAGES, false, false>, L::kBytes, "modconv3x3_tc")) != GANCE_OK) return rc;
    if ((rc = raise_smem_limit(modconv3x3_tc_kernel<__nv_bfloat16, N_TILE, BLOCK_K, STAGES, true, false>, L::kBytes, "modconv3x3_tc<up>")) != GANCE_OK) return rc;
    if ((rc = raise_smem_limit(modconv3x3_tc_kernel<__half, N_TILE, BLOCK_K, STAGES, true, false>, L::kBytes, "modconv3x3_tc<up>")) != GANCE_OK) return rc;
    if ((rc = raise_smem_limit(modconv3x3_tc_kernel<__nv_bfloat16, N_TILE, BLOCK_K, STAGES, false, true>, L::kBytes, "modconv3x3_tc<rgb>")) != GANCE_OK) return rc;
    return raise_smem_limit(modconv3x3_tc_kernel<__half, N_TILE, BLOCK_K, STAGES, false, true>, L::kBytes, "modconv3x3_tc<rgb>");
}

static int make_tmap(CUtensorMap* map, const void* base, bool half_operands, int rank, const cuuint64_t* dims,
                     const cuuint64_t* strides_bytes, const cuuint32_t* box, int swizzle_bytes) {
    cuuint32_t elem_strides[5] = {1, 1, 1, 1, 1};
    CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                                 : (swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                    : (swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                                           : CU_TENSOR_MAP_SWIZZLE_NONE));
    CUresult rc = g_encode_tiled(map, half_operands ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                                 (cuuint32_t)rank, const_cast<void*>(base),
                                 dims, strides_bytes, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) return fail(GANCE_E_DRIVER, "cuTensorMapEncodeTiled failed with CUresult %d", (int)rc);
    return GANCE_OK;
}

template <typename T, int N_TILE, int BLOCK_K, int STAGES>
static int launch_typed(const ConvParams& p, const void* x, const void* weight, int h_in, int w_in,
                        cudaStream_t stream) {
    using L = SmemLayout<N_TILE, BLOCK_K, STAGES>;
    constexpr bool kHalf = std::is_same<T, __half>::value;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)p.c_in, (cuuint64_t)w_in, (cuuint64_t)h_in, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {(cuuint64_t)p.c_in * 2, (cuuint64_t)w_in * p.c_in * 2,
                                 (cuuint64_t)h_in * w_in * p.c_in * 2};
        cuuint32_t box[4] = {(cuuint32_t)BLOCK_K, (cuuint32_t)p.tile_w, (cuuint32_t)p.tile_h, 1u};
        int rc = make_tmap(&tmap_x, x, kHalf, 4, dims, strides, box, BLOCK_K * 2);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)9 * p.c_in, (cuuint64_t)p.c_out};
        cuuint64_t strides[1] = {(cuuint64_t)9 * p.c_in * 2};
        cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, (cuuint32_t)N_TILE};
        int rc = make_tmap(&tmap_w, weight, kHalf, 2, dims, strides, box, BLOCK_K * 2);
        if (rc != GANCE_OK) return rc;
    }
    ConvParams q = p;
    q.a_tx_bytes = p.tile_w * p.tile_h * BLOCK_K * 2;
    const int total_tiles = p.tiles_b * p.tiles_h * p.tiles_w * p.n_tiles;
    const int grid = total_tiles < g_num_sms ? total_tiles : g_num_sms;
    if (p.rgb_weight != nullptr)
        modconv3x3_tc_kernel<T, N_TILE, BLOCK_K, STAGES, false, true><<<grid, kThreads, L::kBytes, stream>>>(tmap_x, tmap_w, q);
    else if (p.up)
        modconv3x3_tc_kernel<T, N_TILE, BLOCK_K, STAGES, true, false><<<grid, kThreads, L::kBytes, stream>>>(tmap_x, tmap_w, q);
    else
        modconv3x3_tc_kernel<T, N_TILE, BLOCK_K, STAGES, false, false><<<grid, kThreads, L::kBytes, stream>>>(tmap_x, tmap_w, q);
    return check_launch("modconv3x3_tc_kernel");
}

template <int N_TILE, int BLOCK_K, int STAGES>
static int launch(bool half_operands, const ConvParams& p, const void* x, const void* weight, int h_in, int w_in,
                  cudaStream_t stream) {
    if (half_operands) return launch_typed<__half, N_TILE, BLOCK_K, STAGES>(p, x, weight, h_in, w_in, stream);
    return launch_typed<__nv_bfloat16, N_TILE, BLOCK_K, STAGES>(p, x, weight, h_in, w_in, stream);
}

// ---- CTA-pair launches ----
static int g_pair_mode = 1;         // gance_set_conv_pairing(): 0 never, 1 when a layer fills a wave of CTA pairs, 2 whenever eligible
static int g_pair_clusters[2] = {0, 0};  // co-resident 2-CTA clusters for the <256,64,6> / <128,64,8> instances (0 = not usable)

template <typename T, int N_TILE, int BLOCK_K, int STAGES, bool RGB>
static int configure_pair_one(int* clusters) {
    using L = PairSmemLayout<N_TILE, BLOCK_K, STAGES>;
    auto kernel = modconv3x3_pair_kernel<T, N_TILE, BLOCK_K, STAGES, RGB>;
    int rc = raise_smem_limit(kernel, L::kBytes, "modconv3x3_pair");
    if (rc != GANCE_OK) return rc;
    cudaLaunchConfig_t config = {};
    config.gridDim = dim3(2 * g_num_sms, 1, 1);
    config.blockDim = dim3(kThreads, 1, 1);
    config.dynamicSmemBytes = L::kBytes;
    cudaLaunchAttribute attribute;
    attribute.id = cudaLaunchAttributeClusterDimension;
    attribute.val.clusterDim.x = 2; attribute.val.clusterDim.y = 1; attribute.val.clusterDim.z = 1;
    config.attrs = &attribute;
    config.numAttrs = 1;
    int n = 0;
    cudaError_t err = cudaOccupancyMaxActiveClusters(&n, kernel, &config);
    if (err != cudaSuccess) {
        cudaGetLastError();
        n = 0;  // pairing stays off for this instance; the single-CTA kernel serves it
    }
    if (*clusters == 0 || n < *clusters) *clusters = n;
    return GANCE_OK;
}

template <int N_TILE, int BLOCK_K, int STAGES>
static int configure_pair(int* clusters) {
    int rc, n = 1 << 30;
    if ((rc = configure_pair_one<__half, N_TILE, BLOCK_K, STAGES, false>(&n)) != GANCE_OK) return rc;
    if ((rc = configure_pair_one<__half, N_TILE, BLOCK_K, STAGES, true>(&n)) != GANCE_OK) return rc;
    if ((rc = configure_pair_one<__nv_bfloat16, N_TILE, BLOCK_K, STAGES, false>(&n)) != GANCE_OK) return rc;
    if ((rc = configure_pair_one<__nv_bfloat16, N_TILE, BLOCK_K, STAGES, true>(&n)) != GANCE_OK) return rc;
    *clusters = n == (1 << 30) ? 0 : n;
    return GANCE_OK;
}

template <typename T, int N_TILE, int BLOCK_K, int STAGES>
static int launch_pair_typed(const ConvParams& p, const void* x, const void* weight, int h_in, int w_in, int max_clusters,
                             cudaStream_t stream) {
    using L = PairSmemLayout<N_TILE, BLOCK_K, STAGES>;
    constexpr bool kHalf = std::is_same<T, __half>::value;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)p.c_in, (cuuint64_t)w_in, (cuuint64_t)h_in, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {(cuuint64_t)p.c_in * 2, (cuuint64_t)w_in * p.c_in * 2,
                                 (cuuint64_t)h_in * w_in * p.c_in * 2};
        cuuint32_t box[4] = {(cuuint32_t)BLOCK_K, (cuuint32_t)p.tile_w, (cuuint32_t)p.tile_h, 1u};
        int rc = make_tmap(&tmap_x, x, kHalf, 4, dims, strides, box, BLOCK_K * 2);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)9 * p.c_in, (cuuint64_t)p.c_out};
        cuuint64_t strides[1] = {(cuuint64_t)9 * p.c_in * 2};
        cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, (cuuint32_t)(N_TILE / 2)};
        int rc = make_tmap(&tmap_w, weight, kHalf, 2, dims, strides, box, BLOCK_K * 2);
        if (rc != GANCE_OK) return rc;
    }
    ConvParams q = p;
    q.a_tx_bytes = p.tile_w * p.tile_h * BLOCK_K * 2;
    const int m_tiles = p.tiles_b * p.tiles_h * p.tiles_w;
    const int items = ((m_tiles + 1) / 2) * p.n_tiles;
    const int clusters = items < max_clusters ? items : max_clusters;
    if (p.rgb_weight != nullptr)
        modconv3x3_pair_kernel<T, N_TILE, BLOCK_K, STAGES, true><<<2 * clusters, kThreads, L::kBytes, stream>>>(tmap_x, tmap_w, q);
    else
        modconv3x3_pair_kernel<T, N_TILE, BLOCK_K, STAGES, false><<<2 * clusters, kThreads, L::kBytes, stream>>>(tmap_x, tmap_w, q);
    return check_launch("modconv3x3_pair_kernel");
}

template <int N_TILE, int BLOCK_K, int STAGES>
static int launch_pair(bool half_operands, const ConvParams& p, const void* x, const void* weight, int h_in, int w_in,
                       int max_clusters, cudaStream_t stream) {
    if (half_operands) return launch_pair_typed<__half, N_TILE, BLOCK_K, STAGES>(p, x, weight, h_in, w_in, max_clusters, stream);
    return launch_pair_typed<__nv_bfloat16, N_TILE, BLOCK_K, STAGES>(p, x, weight, h_in, w_in, max_clusters, stream);
}

// The halo instances that exist: <N_GEMM, CIN> = <128,64> is the folded 1024^2 up-convolution (and plain 64->128),
// <64,64> the 512^2 Conv1 (plain / fused ToRGB), <32,32> the 1024^2 Conv1 (plain / fused ToRGB).
template <typename T, int N_GEMM, int CIN, int STAGES>
static int launch_halo_typed(const ConvParams& p, const void* x, const void* weight, cudaStream_t stream) {
    using L = HaloSmem<N_GEMM, CIN, STAGES>;
    constexpr bool kHalf = std::is_same<T, __half>::value;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)CIN, (cuuint64_t)p.w_out, (cuuint64_t)p.h_out, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {(cuuint64_t)CIN * 2, (cuuint64_t)p.w_out * CIN * 2, (cuuint64_t)p.h_out * p.w_out * CIN * 2};
        cuuint32_t box[4] = {(cuuint32_t)CIN, (cuuint32_t)kHaloW, (cuuint32_t)kHaloH, 1u};
        int rc = make_tmap(&tmap_x, x, kHalf, 4, dims, strides, box, CIN * 2);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)9 * CIN, (cuuint64_t)N_GEMM};
        cuuint64_t strides[1] = {(cuuint64_t)9 * CIN * 2};
        cuuint32_t box[2] = {(cuuint32_t)CIN, (cuuint32_t)N_GEMM};
        int rc = make_tmap(&tmap_w, weight, kHalf, 2, dims, strides, box, CIN * 2);
        if (rc != GANCE_OK) return rc;
    }
    const int total_tiles = p.tiles_b * p.tiles_h * p.tiles_w;
    const int grid = total_tiles < g_num_sms ? total_tiles : g_num_sms;
#ifdef GANCE_PIPE_STATS
    ConvParams q = p;
    q.pipe_stats = g_pipe_stats ? g_pipe_stats + (size_t)(g_pipe_slot++) * 256 * 8 : nullptr;
    const ConvParams& launch_params = q;
#else
    const ConvParams& launch_params = p;
#endif
    if constexpr (N_GEMM == 128) {  // only the folded up-convolution 64 -> 4 x 32
        modconv3x3_halo_kernel<T, N_GEMM, CIN, STAGES, true, false><<<grid, halo_threads(N_GEMM), L::kBytes, stream>>>(tmap_x, tmap_w, launch_params);
    } else if (p.rgb_weight != nullptr) {
        modconv3x3_halo_kernel<T, N_GEMM, CIN, STAGES, false, true><<<grid, halo_threads(N_GEMM), L::kBytes, stream>>>(tmap_x, tmap_w, launch_params);
    } else {
        modconv3x3_halo_kernel<T, N_GEMM, CIN, STAGES, false, false><<<grid, halo_threads(N_GEMM), L::kBytes, stream>>>(tmap_x, tmap_w, launch_params);
    }
    return check_launch("modconv3x3_halo_kernel");
}

template <int N_GEMM, int CIN, int STAGES>
static int launch_halo(bool half_operands, const ConvParams& p, const void* x, const void* weight, cudaStream_t stream) {
    if (half_operands) return launch_halo_typed<__half, N_GEMM, CIN, STAGES>(p, x, weight, stream);
    return launch_halo_typed<__nv_bfloat16, N_GEMM, CIN, STAGES>(p, x, weight, stream);
}

template <int N_GEMM, int CIN, int STAGES>
static int configure_halo() {
    using L = HaloSmem<N_GEMM, CIN, STAGES>;
    int rc;
    if constexpr (N_GEMM == 128) {
        if ((rc = raise_smem_limit(modconv3x3_halo_kernel<__nv_bfloat16, N_GEMM, CIN, STAGES, true, false>, L::kBytes, "modconv3x3_halo<up>")) != GANCE_OK) return rc;
        return raise_smem_limit(modconv3x3_halo_kernel<__half, N_GEMM, CIN, STAGES, true, false>, L::kBytes, "modconv3x3_halo<up>");
    } else {
        if ((rc = raise_smem_limit(modconv3x3_halo_kernel<__nv_bfloat16, N_GEMM, CIN, STAGES, false, false>, L::kBytes, "modconv3x3_halo")) != GANCE_OK) return rc;
        if ((rc = raise_smem_limit(modconv3x3_halo_kernel<__half, N_GEMM, CIN, STAGES, false, false>, L::kBytes, "modconv3x3_halo")) != GANCE_OK) return rc;
        if ((rc = raise_smem_limit(modconv3x3_halo_kernel<__nv_bfloat16, N_GEMM, CIN, STAGES, false, true>, L::kBytes, "modconv3x3_halo<rgb>")) != GANCE_OK) return rc;
        return raise_smem_limit(modconv3x3_halo_kernel<__half, N_GEMM, CIN, STAGES, false, true>, L::kBytes, "modconv3x3_halo<rgb>");
    }
}

// ---- two-row halo kernel (32 -> 32 channels): tiles of 32 rows x 8 columns ---------------------------------------------------
static int g_two_rows = 1;  // gance_set_conv_two_rows: 0 sends the 32 -> 32 layer to the one-row kernel (A/B runs, bit-identity tests)

template <typename T>
static int launch_halo2_typed(const ConvParams& p, const void* x, const void* weight, cudaStream_t stream) {
    using L = Halo2Smem<kHalo2Stages>;
    constexpr bool kHalf = std::is_same<T, __half>::value;
    constexpr int CIN = 32, C = 32;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)CIN, (cuuint64_t)p.w_out, (cuuint64_t)p.h_out, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {(cuuint64_t)CIN * 2, (cuuint64_t)p.w_out * CIN * 2, (cuuint64_t)p.h_out * p.w_out * CIN * 2};
        cuuint32_t box[4] = {(cuuint32_t)CIN, (cuuint32_t)kHaloW, (cuuint32_t)kHalo2H, 1u};
        int rc = make_tmap(&tmap_x, x, kHalf, 4, dims, strides, box, CIN * 2);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)9 * CIN, (cuuint64_t)C};
        cuuint64_t strides[1] = {(cuuint64_t)9 * CIN * 2};
        cuuint32_t box[2] = {(cuuint32_t)CIN, (cuuint32_t)C};
        int rc = make_tmap(&tmap_w, weight, kHalf, 2, dims, strides, box, CIN * 2);
        if (rc != GANCE_OK) return rc;
    }
    const int total_tiles = p.tiles_b * p.tiles_h * p.tiles_w;
    const int grid = total_tiles < g_num_sms ? total_tiles : g_num_sms;
    if (p.rgb_weight != nullptr)
        modconv3x3_halo2_kernel<T, kHalo2Stages, true><<<grid, kHalo2Threads, L::kBytes, stream>>>(tmap_x, tmap_w, p);
    else
        modconv3x3_halo2_kernel<T, kHalo2Stages, false><<<grid, kHalo2Threads, L::kBytes, stream>>>(tmap_x, tmap_w, p);
    return check_launch("modconv3x3_halo2_kernel");
}

static int launch_halo2(bool half_operands, const ConvParams& p, const void* x, const void* weight, cudaStream_t stream) {
    if (half_operands) return launch_halo2_typed<__half>(p, x, weight, stream);
    return launch_halo2_typed<__nv_bfloat16>(p, x, weight, stream);
}

static int configure_halo2() {
    using L = Halo2Smem<kHalo2Stages>;
    int rc;
    if ((rc = raise_smem_limit(modconv3x3_halo2_kernel<__nv_bfloat16, kHalo2Stages, false>, L::kBytes, "modconv3x3_halo2")) != GANCE_OK) return rc;
    if ((rc = raise_smem_limit(modconv3x3_halo2_kernel<__half, kHalo2Stages, false>, L::kBytes, "modconv3x3_halo2")) != GANCE_OK) return rc;
    if ((rc = raise_smem_limit(modconv3x3_halo2_kernel<__nv_bfloat16, kHalo2Stages, true>, L::kBytes, "modconv3x3_halo2<rgb>")) != GANCE_OK) return rc;
    return raise_smem_limit(modconv3x3_halo2_kernel<__half, kHalo2Stages, true>, L::kBytes, "modconv3x3_halo2<rgb>");
}

// ---- CTA-pair halo launches: <N_GEMM, CIN> = <128,64> folded 1024^2 up-convolution, <64,64> 512^2 Conv1, <128,128> 256^2 Conv1
static int g_halo_pair_clusters[3] = {0, 0, 0};

template <typename K>
static int pair_cluster_capacity(K kernel, int smem_bytes) {
    cudaLaunchConfig_t config = {};
    config.gridDim = dim3(2 * g_num_sms, 1, 1);
    config.blockDim = dim3(kThreads, 1, 1);
    config.dynamicSmemBytes = smem_bytes;
    cudaLaunchAttribute attribute;
    attribute.id = cudaLaunchAttributeClusterDimension;
    attribute.val.clusterDim.x = 2; attribute.val.clusterDim.y = 1; attribute.val.clusterDim.z = 1;
    config.attrs = &attribute;
    config.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kernel, &config) != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    return n;
}

template <int N_GEMM, int CIN, int STAGES, bool UP>
constexpr int halo_pair_smem_bytes() {
    using L = HaloPairSmem<N_GEMM, CIN, STAGES>;
    return (UP && N_GEMM == 128) ? L::kBytesStaged : L::kBytesCache;
}

template <typename T, int N_GEMM, int CIN, int STAGES, bool UP, bool RGB>
static int configure_halo_pair_one(int* clusters) {
    constexpr int kBytes = halo_pair_smem_bytes<N_GEMM, CIN, STAGES, UP>();
    auto kernel = modconv3x3_halo_pair_kernel<T, N_GEMM, CIN, STAGES, UP, RGB>;
    int rc = raise_smem_limit(kernel, kBytes, "modconv3x3_halo_pair");
    if (rc != GANCE_OK) return rc;
    const int n = pair_cluster_capacity(kernel, kBytes);
    if (n < *clusters) *clusters = n;
    return GANCE_OK;
}

template <int N_GEMM, int CIN, int STAGES, bool UP>
static int configure_halo_pair(int* clusters) {
    int rc, n = 1 << 30;
    if ((rc = configure_halo_pair_one<__half, N_GEMM, CIN, STAGES, UP, false>(&n)) != GANCE_OK) return rc;
    if ((rc = configure_halo_pair_one<__nv_bfloat16, N_GEMM, CIN, STAGES, UP, false>(&n)) != GANCE_OK) return rc;
    if constexpr (!UP) {
        if ((rc = configure_halo_pair_one<__half, N_GEMM, CIN, STAGES, UP, true>(&n)) != GANCE_OK) return rc;
        if ((rc = configure_halo_pair_one<__nv_bfloat16, N_GEMM, CIN, STAGES, UP, true>(&n)) != GANCE_OK) return rc;
    }
    *clusters = n == (1 << 30) ? 0 : n;
    return GANCE_OK;
}

template <typename T, int N_GEMM, int CIN, int STAGES, bool UP>
static int launch_halo_pair_typed(const ConvParams& p, const void* x, const void* weight, int max_clusters, cudaStream_t stream) {
    constexpr int kBytes = halo_pair_smem_bytes<N_GEMM, CIN, STAGES, UP>();
    constexpr bool kHalf = std::is_same<T, __half>::value;
    CUtensorMap tmap_x, tmap_w, tmap_out;
    {
        cuuint64_t dims[4] = {(cuuint64_t)CIN, (cuuint64_t)p.w_out, (cuuint64_t)p.h_out, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {(cuuint64_t)CIN * 2, (cuuint64_t)p.w_out * CIN * 2, (cuuint64_t)p.h_out * p.w_out * CIN * 2};
        cuuint32_t box[4] = {64u, (cuuint32_t)kHaloW, (cuuint32_t)kHaloH, 1u};
        int rc = make_tmap(&tmap_x, x, kHalf, 4, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    tmap_out = tmap_x;  // unused unless the output is staged
    if constexpr (UP && N_GEMM == 128) {
        // output [B, 2H, 2W, 32] seen as [B, 2H, W, 64]: one 128-byte row = two neighbouring pixels; a warp's tile is
        // 8 rows x 8 pixel pairs
        const cuuint64_t oh = 2ull * p.h_out, ow_pairs = (cuuint64_t)p.w_out;
        cuuint64_t dims[4] = {64ull, ow_pairs, oh, (cuuint64_t)p.batch};
        cuuint64_t strides[3] = {128ull, ow_pairs * 128ull, oh * ow_pairs * 128ull};
        cuuint32_t box[4] = {64u, 8u, 8u, 1u};
        int rc = make_tmap(&tmap_out, p.out, kHalf, 4, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)9 * CIN, (cuuint64_t)N_GEMM};
        cuuint64_t strides[1] = {(cuuint64_t)9 * CIN * 2};
        cuuint32_t box[2] = {64u, (cuuint32_t)(N_GEMM / 2)};
        int rc = make_tmap(&tmap_w, weight, kHalf, 2, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    const int items = (p.tiles_b * p.tiles_h * p.tiles_w) / 2;
    const int clusters = items < max_clusters ? items : max_clusters;
    if constexpr (UP) {
        modconv3x3_halo_pair_kernel<T, N_GEMM, CIN, STAGES, true, false><<<2 * clusters, kThreads, kBytes, stream>>>(tmap_x, tmap_w, tmap_out, p);
    } else {
        if (p.rgb_weight != nullptr)
            modconv3x3_halo_pair_kernel<T, N_GEMM, CIN, STAGES, false, true><<<2 * clusters, kThreads, kBytes, stream>>>(tmap_x, tmap_w, tmap_out, p);
        else
            modconv3x3_halo_pair_kernel<T, N_GEMM, CIN, STAGES, false, false><<<2 * clusters, kThreads, kBytes, stream>>>(tmap_x, tmap_w, tmap_out, p);
    }
    return check_launch("modconv3x3_halo_pair_kernel");
}

template <int N_GEMM, int CIN, int STAGES, bool UP>
static int launch_halo_pair(bool half_operands, const ConvParams& p, const void* x, const void* weight, int max_clusters,
                            cudaStream_t stream) {
    if (half_operands) return launch_halo_pair_typed<__half, N_GEMM, CIN, STAGES, UP>(p, x, weight, max_clusters, stream);
    return launch_halo_pair_typed<__nv_bfloat16, N_GEMM, CIN, STAGES, UP>(p, x, weight, max_clusters, stream);
}

static const bool g_use_halo = true;

// used by upconv_phase.cu
int encode_tensor_map(CUtensorMap* map, const void* base, bool half_operands, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes, const cuuint32_t* box, int swizzle_bytes) {
    return make_tmap(map, base, half_operands, rank, dims, strides_bytes, box, swizzle_bytes);
}
// used by modconv_tf32.cu: fp32 elements, 128-byte swizzled rows
int encode_tensor_map_f32(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
                          const cuuint64_t* strides_bytes, const cuuint32_t* box) {
    cuuint32_t elem_strides[5] = {1, 1, 1, 1, 1};
    CUresult rc = g_encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), dims,
                                 strides_bytes, box, elem_strides, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) return fail(GANCE_E_DRIVER, "cuTensorMapEncodeTiled (fp32) failed with CUresult %d", (int)rc);
    return GANCE_OK;
}
int sm_count() { return g_num_sms; }
int configure_phase_kernels();
int configure_fused_kernels();
int configure_fused_pair_kernels();
int conv_pair_mode() { return g_pair_mode; }

}  // namespace tc

int init_tensor_core_path() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (err != cudaSuccess || qres != cudaDriverEntryPointSuccess || fn == nullptr)
        return fail(GANCE_E_DRIVER, "cuTensorMapEncodeTiled is not exported by this driver (%s)",
                    err == cudaSuccess ? "symbol not found" : cudaGetErrorString(err));
    tc::g_encode_tiled = reinterpret_cast<tc::EncodeTiledFn>(fn);
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&tc::g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    int rc;
    if ((rc = tc::configure<256, 64, 4>()) != GANCE_OK) return rc;
    if ((rc = tc::configure<128, 64, 6>()) != GANCE_OK) return rc;
    if ((rc = tc::configure<64, 64, 8>()) != GANCE_OK) return rc;
    if ((rc = tc::configure<32, 64, 8>()) != GANCE_OK) return rc;
    if ((rc = tc::configure<32, 32, 8>()) != GANCE_OK) return rc;
    if ((rc = tc::configure_pair<256, 64, 6>(&tc::g_pair_clusters[0])) != GANCE_OK) return rc;
    if ((rc = tc::configure_pair<128, 64, 8>(&tc::g_pair_clusters[1])) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo_pair<128, 64, 3, true>(&tc::g_halo_pair_clusters[0])) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo_pair<64, 64, 4, false>(&tc::g_halo_pair_clusters[1])) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo_pair<128, 128, 3, false>(&tc::g_halo_pair_clusters[2])) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo<128, 64, 2>()) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo<64, 64, tc::kHaloStages64>()) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo<32, 32, tc::kHaloStages32>()) != GANCE_OK) return rc;
    if ((rc = tc::configure_halo2()) != GANCE_OK) return rc;
    if ((rc = tc::configure_phase_kernels()) != GANCE_OK) return rc;
    if ((rc = tc::configure_fused_kernels()) != GANCE_OK) return rc;
    if ((rc = tc::configure_fused_pair_kernels()) != GANCE_OK) return rc;
    return GANCE_OK;
}

int launch_modconv3x3_f32(const float* x, int batch, int h_out, int w_out, int c_in, int c_out, int pad,
                          const float* weight, const float* demod, const float* noise,
                          long long noise_batch_stride, float noise_strength, const float* bias,
                          const float* next_styles, float* out, cudaStream_t stream);

}  // namespace gance

using namespace gance;

struct RgbFusion {
    const float* styles;
    const float* weight;
    const float* bias;
    const float* y_prev;
    float* y_out;
    uint8_t* frames;
};

// Shared launcher of the tcgen05 path. `c_out` is the GEMM N (4 * c_real for a folded up-convolution).
static int launch_tensor_core_conv(const char* who, const void* x, int dtype, int batch, int h_out, int w_out, int c_in,
                                   int c_out, int pad, int up, int c_real, const void* weight, const float* demod,
                                   const float* noise, long long noise_batch_stride, float noise_strength,
                                   const float* bias, const float* next_styles, void* out, cudaStream_t s,
                                   const RgbFusion* rgb = nullptr, void* out_plain = nullptr) {
    GANCE_REQUIRE(dtype == GANCE_BF16 || dtype == GANCE_F16, GANCE_E_DTYPE, "%s: dtype %d", who, dtype);
    const bool half_operands = dtype == GANCE_F16;
    GANCE_REQUIRE(c_in % 32 == 0 && c_real % 32 == 0, GANCE_E_SHAPE,
                  "%s: the tcgen05 path needs Cin %% 32 == 0 and Cout %% 32 == 0 (got %d, %d)", who, c_in, c_real);
    GANCE_REQUIRE(aligned16(x) && aligned16(weight) && aligned16(out) && aligned16(demod) && aligned16(bias) &&
                      aligned16(next_styles),
                  GANCE_E_ALIGN, "%s: pointers must be 16-byte aligned", who);

    tc::ConvParams p;
    p.batch = batch; p.h_out = h_out; p.w_out = w_out; p.c_in = c_in; p.c_out = c_out; p.pad = pad;
    p.up = up; p.c_real = c_real;
    // One tile = up to 128 output pixels of ONE sample: as wide as the image allows (<= 16), then rows. Resolutions
    // below 16x8 leave accumulator rows unused (4x4: 16 of 128) - those layers are a rounding error of the frame time,
    // and one sample per tile is what lets the epilogue stage its coefficients in shared memory.
    auto is_pow2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
    auto ilog2 = [](int v) { int l = 0; while ((1 << l) < v) ++l; return l; };
    p.tile_w = w_out >= 16 ? 16 : w_out;
    GANCE_REQUIRE(is_pow2(p.tile_w), GANCE_E_SHAPE, "%s: W=%d (< 16) must be a power of two", who, w_out);
    const int rows = tc::kTileM / p.tile_w;
    p.tile_h = h_out >= rows ? rows : h_out;
    GANCE_REQUIRE(is_pow2(p.tile_h), GANCE_E_SHAPE, "%s: H=%d (< %d) must be a power of two", who, h_out, rows);
    GANCE_REQUIRE(is_pow2(c_real) && c_real <= tc::kCacheFloat4PerGroup / (rgb ? 2 : 1), GANCE_E_SHAPE,
                  "%s: Cout=%d must be a power of two <= %d", who, c_real, tc::kCacheFloat4PerGroup / (rgb ? 2 : 1));
    p.tile_b = 1;
    p.tiles_w = (w_out + p.tile_w - 1) / p.tile_w;
    p.tiles_h = (h_out + p.tile_h - 1) / p.tile_h;
    p.tiles_b = batch;
    p.tile_w_log2 = ilog2(p.tile_w); p.tile_hw_log2 = ilog2(p.tile_w * p.tile_h); p.c_log2 = ilog2(c_real);
    p.a_tx_bytes = 0;  // set per instantiation (depends on BLOCK_K)
    p.demod = demod; p.noise = noise; p.noise_batch_stride = noise_batch_stride; p.noise_strength = noise_strength;
    p.bias = bias; p.next_styles = next_styles; p.out = out; p.out_plain = out_plain;
    GANCE_REQUIRE(out_plain == nullptr || (!up && rgb == nullptr && c_in % 64 == 0 && c_in >= 128 && aligned16(out_plain)),
                  GANCE_E_SHAPE, "%s: the unscaled second output needs a plain convolution with Cin %% 64 == 0, Cin >= 128", who);
    p.rgb_styles = rgb ? rgb->styles : nullptr; p.rgb_weight = rgb ? rgb->weight : nullptr;
    p.rgb_bias = rgb ? rgb->bias : nullptr; p.y_prev = rgb ? rgb->y_prev : nullptr;
    p.y_out = rgb ? rgb->y_out : nullptr; p.frames = rgb ? rgb->frames : nullptr;
    p.nonfinite = nonfinite_flag();
    p.pipe_stats = nullptr;
    const bool one_n_tile = rgb != nullptr;  // the fused ToRGB needs every channel of a pixel in one tile
    GANCE_REQUIRE(!one_n_tile || (up == 0 && (c_out == 256 || c_out == 128 || c_out == 64 || c_out == 32)), GANCE_E_SHAPE,
                  "%s: fused ToRGB needs Cout in {32,64,128,256} (got %d)", who, c_out);
    const int h_in = h_out + 2 - 2 * pad, w_in = w_out + 2 - 2 * pad;
    const int m_tiles = p.tiles_w * p.tiles_h * p.tiles_b;
    auto set_tile_logs = [&]() {
        const bool pow2 = is_pow2(p.tiles_w) && is_pow2(p.tiles_h);
        p.tiles_w_log2 = pow2 ? ilog2(p.tiles_w) : -1;
        p.tiles_h_log2 = pow2 ? ilog2(p.tiles_h) : -1;
    };
    set_tile_logs();

    // Narrow layers (512^2 / 1024^2): halo kernel, whole weight resident, one activation fetch per tile. As CTA pairs
    // (half of the weights per CTA) also the 128 -> 128 channel layer (256^2 Conv1).
    const bool halo_shape = tc::g_use_halo && pad == 1 && w_out % tc::kHaloTileW == 0 && h_out % tc::kHaloTileH == 0;
    const int halo_m_tiles = halo_shape ? batch * (w_out / tc::kHaloTileW) * (h_out / tc::kHaloTileH) : 0;
    // CTA pairs for the halo layers: even m-tile count, and (unless forced) one work item per co-resident pair. Measured
    // (round 2, batch 32): the 128 -> 128 layer gains 14 % (660 -> 580 us: it leaves the L2-bound generic kernel); the folded
    // 1024^2 up-convolution, whose pair flavour has the shared memory to stage its output for TMA stores, 3 % (907 -> 877 us:
    // that layer now moves its 3.3 GB at 3.8 TB/s); the 512^2 Conv1 LOSES (680 -> 800 us: its tiles are 36 short MMAs and
    // every accumulator hand-over crosses the SM pair) and pairs only when forced (mode 2).
    auto halo_pair_ok = [&](int which) {
        const int clusters = tc::g_halo_pair_clusters[which];
        if (tc::g_pair_mode == 0 || clusters <= 0 || halo_m_tiles % 2 != 0) return false;
        if (tc::g_pair_mode == 2) return true;
        return which != 1 && halo_m_tiles / 2 >= clusters;
    };
    const bool wide_halo = halo_shape && c_in == 128 && c_out == 128 && !up && halo_pair_ok(2);
    if (wide_halo || (halo_shape &&
        ((c_in == 64 && ((c_out == 128 && up && c_real == 32) || (c_out == 64 && !up))) || (c_in == 32 && c_out == 32 && !up)))) {
        p.tile_w = tc::kHaloTileW; p.tile_h = tc::kHaloTileH; p.tile_b = 1;
        p.tile_w_log2 = ilog2(p.tile_w); p.tile_hw_log2 = ilog2(p.tile_w * p.tile_h);
        p.tiles_w = w_out / p.tile_w; p.tiles_h = h_out / p.tile_h; p.tiles_b = batch;
        set_tile_logs();
        p.n_tiles = 1; p.k_chunks = 1;
        if (wide_halo) return tc::launch_halo_pair<128, 128, 3, false>(half_operands, p, x, weight, tc::g_halo_pair_clusters[2], s);
        if (c_in == 64 && c_out == 128 && halo_pair_ok(0))
            return tc::launch_halo_pair<128, 64, 3, true>(half_operands, p, x, weight, tc::g_halo_pair_clusters[0], s);
        if (c_in == 64 && c_out == 64 && halo_pair_ok(1))
            return tc::launch_halo_pair<64, 64, 4, false>(half_operands, p, x, weight, tc::g_halo_pair_clusters[1], s);
        if (c_in == 32 && tc::g_two_rows != 0 && h_out % tc::kHalo2TileH == 0) {  // two image rows per accumulator row
            p.tile_h = tc::kHalo2TileH; p.tile_hw_log2 = ilog2(p.tile_w * p.tile_h);
            p.tiles_h = h_out / p.tile_h;
            set_tile_logs();
            return tc::launch_halo2(half_operands, p, x, weight, s);
        }
        if (c_in == 32) return tc::launch_halo<32, 32, tc::kHaloStages32>(half_operands, p, x, weight, s);
        if (c_out == 128) return tc::launch_halo<128, 64, 2>(half_operands, p, x, weight, s);
        return tc::launch_halo<64, 64, tc::kHaloStages64>(half_operands, p, x, weight, s);
    }
    if (c_in % 64 != 0) {  // Cin = 32 (1024^2 Conv1 in config-f): 64-byte rows, SWIZZLE_64B
        p.n_tiles = c_out / 32; p.k_chunks = c_in / 32;
        return tc::launch<32, 32, 8>(half_operands, p, x, weight, h_in, w_in, s);
    }
    p.k_chunks = c_in / 64;
    // CTA pairs (cta_group::2) for the plain / fused-ToRGB layers with at least one full wave of pair items
    const bool pairable = tc::g_pair_mode != 0 && !up;
    const int pair_items_needed = tc::g_pair_mode == 2 ? 1 : 0;  // 0: at least one item per co-resident cluster
    if (c_out % 256 == 0 && (m_tiles >= 148 || one_n_tile)) {
        p.n_tiles = c_out / 256;
        if (pairable && tc::g_pair_clusters[0] > 0 && ((m_tiles + 1) / 2) * p.n_tiles >= (pair_items_needed ? pair_items_needed : tc::g_pair_clusters[0]))
            return tc::launch_pair<256, 64, 6>(half_operands, p, x, weight, h_in, w_in, tc::g_pair_clusters[0], s);
        return tc::launch<256, 64, 4>(half_operands, p, x, weight, h_in, w_in, s);
    }
    if (c_out % 128 == 0 && !(one_n_tile && c_out != 128)) {
        p.n_tiles = c_out / 128;
        if (pairable && tc::g_pair_clusters[1] > 0 && ((m_tiles + 1) / 2) * p.n_tiles >= (pair_items_needed ? pair_items_needed : tc::g_pair_clusters[1]))
            return tc::launch_pair<128, 64, 8>(half_operands, p, x, weight, h_in, w_in, tc::g_pair_clusters[1], s);
        return tc::launch<128, 64, 6>(half_operands, p, x, weight, h_in, w_in, s);
    }
    if (c_out % 64 == 0 && !(one_n_tile && c_out != 64)) {
        p.n_tiles = c_out / 64;
        return tc::launch<64, 64, 8>(half_operands, p, x, weight, h_in, w_in, s);
    }
    p.n_tiles = c_out / 32;
    return tc::launch<32, 64, 8>(half_operands, p, x, weight, h_in, w_in, s);
}

#ifdef GANCE_PIPE_STATS
// tools/pipe_stats.py only: device buffer [slots][256][8] of int64, filled by the next single-CTA halo launches
extern "C" int gance_debug_pipe_stats(long long* buffer) {
    tc::g_pipe_stats = buffer;
    tc::g_pipe_slot = 0;
    return GANCE_OK;
}
#endif

extern "C" int gance_set_conv_pairing(int mode) {
    const int before = tc::g_pair_mode;
    tc::g_pair_mode = mode < 0 ? 0 : (mode > 2 ? 2 : mode);
    return before;
}

extern "C" int gance_set_conv_two_rows(int enabled) {
    const int before = tc::g_two_rows;
    tc::g_two_rows = enabled != 0 ? 1 : 0;
    return before;
}

extern "C" int gance_modconv3x3(const void* x, int dtype, int batch, int h_out, int w_out, int c_in, int c_out,
                                int pad, const void* weight, const float* demod, const float* noise,
                                long long noise_batch_stride, float noise_strength, const float* bias,
                                const float* next_styles, void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight && demod && bias && out, GANCE_E_NULL, "gance_modconv3x3: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h_out > 0 && w_out > 0 && c_in > 0 && c_out > 0 && (pad == 0 || pad == 1),
                  GANCE_E_SHAPE, "gance_modconv3x3: shape B=%d H=%d W=%d Cin=%d Cout=%d pad=%d", batch, h_out, w_out,
                  c_in, c_out, pad);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == GANCE_F32)
        return launch_modconv3x3_f32((const float*)x, batch, h_out, w_out, c_in, c_out, pad, (const float*)weight,
                                     demod, noise, noise_batch_stride, noise_strength, bias, next_styles,
                                     (float*)out, s);
    return launch_tensor_core_conv("gance_modconv3x3", x, dtype, batch, h_out, w_out, c_in, c_out, pad, /*up=*/0, c_out,
                                   weight, demod, noise, noise_batch_stride, noise_strength, bias, next_styles, out, s);
}

extern "C" int gance_modconv3x3_dual(const void* x, int dtype, int batch, int h_out, int w_out, int c_in, int c_out,
                                     int pad, const void* weight, const float* demod, const float* noise,
                                     long long noise_batch_stride, float noise_strength, const float* bias,
                                     const float* next_styles, void* out_scaled, void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight && demod && bias && out && out_scaled && next_styles, GANCE_E_NULL,
                  "gance_modconv3x3_dual: NULL pointer");
    GANCE_REQUIRE(out != out_scaled, GANCE_E_NULL, "gance_modconv3x3_dual: out and out_scaled must be distinct");
    GANCE_REQUIRE(batch > 0 && h_out > 0 && w_out > 0 && c_in > 0 && c_out > 0 && (pad == 0 || pad == 1),
                  GANCE_E_SHAPE, "gance_modconv3x3_dual: shape B=%d H=%d W=%d Cin=%d Cout=%d pad=%d", batch, h_out, w_out,
                  c_in, c_out, pad);
    return launch_tensor_core_conv("gance_modconv3x3_dual", x, dtype, batch, h_out, w_out, c_in, c_out, pad, /*up=*/0,
                                   c_out, weight, demod, noise, noise_batch_stride, noise_strength, bias, next_styles,
                                   out_scaled, (cudaStream_t)stream, nullptr, out);
}

extern "C" int gance_modconv3x3_torgb(const void* x, int dtype, int batch, int h, int w, int c_in, int c_out,
                                      const void* weight, const float* demod, const float* noise,
                                      long long noise_batch_stride, float noise_strength, const float* bias,
                                      const float* next_styles, void* out, const float* rgb_styles,
                                      const float* rgb_weight, const float* rgb_bias, const float* y_prev, float* y_out,
                                      uint8_t* frames, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight && demod && bias && rgb_styles && rgb_weight && rgb_bias, GANCE_E_NULL,
                  "gance_modconv3x3_torgb: NULL pointer");
    GANCE_REQUIRE(y_out || frames, GANCE_E_NULL, "gance_modconv3x3_torgb: neither y_out nor frames given");
    GANCE_REQUIRE(batch > 0 && h > 0 && w > 0 && c_in > 0 && c_out > 0, GANCE_E_SHAPE,
                  "gance_modconv3x3_torgb: shape B=%d H=%d W=%d Cin=%d Cout=%d", batch, h, w, c_in, c_out);
    GANCE_REQUIRE(y_prev == nullptr || (h % 2 == 0 && w % 2 == 0), GANCE_E_SHAPE,
                  "gance_modconv3x3_torgb: skip input needs even h,w");
    GANCE_REQUIRE(aligned16(y_prev) && aligned16(y_out), GANCE_E_ALIGN, "gance_modconv3x3_torgb: y pointers must be 16-byte aligned");
    const RgbFusion rgb{rgb_styles, rgb_weight, rgb_bias, y_prev, y_out, frames};
    return launch_tensor_core_conv("gance_modconv3x3_torgb", x, dtype, batch, h, w, c_in, c_out, /*pad=*/1, /*up=*/0, c_out,
                                   weight, demod, noise, noise_batch_stride, noise_strength, bias, next_styles, out,
                                   (cudaStream_t)stream, &rgb);
}

extern "C" int gance_modconv3x3_up2_folded(const void* x, int dtype, int batch, int h_in, int w_in, int c_in, int c_out,
                                           const void* weight_folded, const float* demod, const float* noise,
                                           long long noise_batch_stride, float noise_strength, const float* bias,
                                           const float* next_styles, void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight_folded && demod && bias && out, GANCE_E_NULL, "gance_modconv3x3_up2_folded: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h_in > 0 && w_in > 0 && c_in > 0 && c_out > 0, GANCE_E_SHAPE,
                  "gance_modconv3x3_up2_folded: shape B=%d H=%d W=%d Cin=%d Cout=%d", batch, h_in, w_in, c_in, c_out);
    return launch_tensor_core_conv("gance_modconv3x3_up2_folded", x, dtype, batch, h_in, w_in, c_in, 4 * c_out, /*pad=*/1,
                                   /*up=*/1, c_out, weight_folded, demod, noise, noise_batch_stride, noise_strength,
                                   bias, next_styles, out, (cudaStream_t)stream);
}

extern "C" int gance_modconv3x3_up2_fir(const void* x, int dtype, int batch, int h_in, int w_in, int c_in,
                                        int c_out, const float* styles, const void* weight, const float* demod,
                                        const float* noise, long long noise_batch_stride, float noise_strength,
                                        const float* bias, const float* next_styles, void* scratch, void* out,
                                        void* stream) {
    GANCE_REQUIRE(scratch != nullptr, GANCE_E_NULL, "gance_modconv3x3_up2_fir: scratch is NULL");
    int rc = gance_upfir2x_pad(x, dtype, batch, h_in, w_in, c_in, styles, scratch, stream);
    if (rc != GANCE_OK) return rc;
    return gance_modconv3x3(scratch, dtype, batch, 2 * h_in, 2 * w_in, c_in, c_out, /*pad=*/0, weight, demod, noise,
                            noise_batch_stride, noise_strength, bias, next_styles, out, stream);
}
