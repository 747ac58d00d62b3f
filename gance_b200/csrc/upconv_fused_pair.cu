// Fused up-convolution for the wide layers (Cin >= 256), second form: CTA pairs (tcgen05 cta_group::2) + ONE activation halo
// box per K chunk. Same mathematics and the same epilogue as upconv_fused.cu (4-phase transposed-convolution GEMM whose
// result t stays in shared memory, [1,3,3,1] FIR + noise + bias + leaky-ReLU + next style applied from there).
//
// Why. Ablation of upconv_fused_kernel (round 2, tools/ablate.sh): with the whole epilogue removed the Cin >= 256 layers
// keep 83-93 % of their time - they are bound by what the TMA producer pulls through L2 per K chunk of 64 channels:
// 4 shifted 16 KB activation boxes + 72 KB of weights = 136 KB for 1152 cycles of tensor math (118 B/clk/SM against the
// ~65 B/clk/SM the L2 delivered). Here, per CTA and K chunk:
//   * activations: ONE (8+1) x (16+1)-cell halo box, 19.1 KB; the A operand of shift (dx,dy) is the same buffer read
//     through a shifted UMMA descriptor (start += ((1-dy) * 9 + (1-dx)) cells, 8-row groups spaced by the 9-cell pitch).
//     That needs 8-cell-wide tiles (an 8-row core-matrix group = 8 consecutive cells of one row), so the tile is 8 x 16
//     cells = a 16 x 32-pixel patch of t and 12 x 28 output pixels - upconv_fused.cu's geometry transposed;
//   * weights: the two CTAs of a pair (two m-tiles, same channel block) each fetch HALF of every shift's weight rows and
//     share one M = 256 MMA: 36 KB.
// 55 KB instead of 136 KB per 1152 cycles.
//
// Rings, both in units of one K chunk: activations (2 halo boxes) and weights (3 x this CTA's half of all four shift tiles,
// 36 KB) - the first version's weight ring held five single-shift tiles, ~1400 cycles of MMA work ahead of the tensor core,
// and the layers stayed latency-bound on the L2 fetch; three K chunks are ~3400 cycles. Both rings are counted on the
// leader's full barriers and released in both CTAs by multicast commits. Warp roles as in upconv_fused.cu; the leader's MMA
// warp issues for the pair.
#include "tc_ptx.cuh"

namespace gance {
namespace tc {

constexpr int kFpCellsW = 8, kFpCellsH = 16;                    // cells per tile (= kTileM accumulator rows)
constexpr int kFpOutW = 2 * (kFpCellsW - 2), kFpOutH = 2 * (kFpCellsH - 2);  // 12 x 28 output pixels per tile
constexpr int kFpTW = 2 * kFpCellsW, kFpTH = 2 * kFpCellsH;     // 16 x 32 pixel patch of t
constexpr int kFpK = 64;                                        // input channels per K chunk
constexpr int kFpCC = 64;                                       // output channels per work item
constexpr int kFpHaloW = kFpCellsW + 1, kFpHaloH = kFpCellsH + 1;
constexpr int kFpHaloBytes = kFpHaloW * kFpHaloH * kFpK * 2;    // 19 584
constexpr int kFpHaloStride = (kFpHaloBytes + 1023) / 1024 * 1024;
constexpr int kFpAStages = 2, kFpBStages = 3;
constexpr int kFpBRowsPerLoad = 32;                             // weight rows per TMA box
constexpr int kFpBHalfRows = 9 * kFpCC / 2;                     // this CTA's weight rows of one K chunk: (4 + 2 + 2 + 1) x 32
constexpr int kFpBStageBytes = kFpBHalfRows * kFpK * 2;         // 36 KB
constexpr int kFpEpilogueThreads = 256;
constexpr int kFpRowSplit = 2;                                  // the 28 output rows are split over two thread groups
constexpr int kFpRowsPerThread = kFpOutH / kFpRowSplit;         // 14

struct FusedPairLayout {
    static constexpr int kBOffset = kFpAStages * kFpHaloStride;
    static constexpr int kPatchOffset = kBOffset + kFpBStages * kFpBStageBytes;
    static constexpr int kPixelBytes = kFpCC * 2;
    static constexpr int kPatchBytes = kFpTW * kFpTH * kPixelBytes;   // 64 KB
    static constexpr int kBarrierOffset = kPatchOffset + kPatchBytes;
    static constexpr int kSmemBytes = kBarrierOffset + (2 * kFpAStages + 2 * kFpBStages + 4) * 8 + 16 + 1024;
};
static_assert(FusedPairLayout::kSmemBytes <= 227 * 1024, "shared memory");

struct FusedPairParams {
    int batch, h_in, w_in, c_in, c_out, k_chunks, n_tiles, tiles_w, tiles_h;
    const float* demod;        // [B, c_out]
    const float* noise;        // [2H,2W] or [B,2H,2W] or NULL
    long long noise_batch_stride;
    float noise_strength;
    const float* bias;         // [c_out]
    const float* next_styles;  // [B, c_out] or NULL
    void* out;                 // [B, 2H, 2W, c_out]
};

// work item -> (channel block, sample, tile row, tile column) of THIS CTA's tile; ghost = the odd tile out
__device__ __forceinline__ bool fused_pair_decode(const FusedPairParams& p, int item, int rank, int& nt, int& b, int& ty, int& tx) {
    nt = item % p.n_tiles;
    int m = 2 * (item / p.n_tiles) + rank;
    tx = m % p.tiles_w;
    m /= p.tiles_w;
    ty = m % p.tiles_h;
    b = m / p.tiles_h;
    return b >= p.batch;
}

template <typename T>
__device__ __forceinline__ void fp_unpack4(uint2 raw, float (&v)[4]);
template <>
__device__ __forceinline__ void fp_unpack4<__half>(uint2 raw, float (&v)[4]) {
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
template <>
__device__ __forceinline__ void fp_unpack4<__nv_bfloat16>(uint2 raw, float (&v)[4]) {
    v[0] = __uint_as_float(raw.x << 16); v[1] = __uint_as_float(raw.x & 0xffff0000u);
    v[2] = __uint_as_float(raw.y << 16); v[3] = __uint_as_float(raw.y & 0xffff0000u);
}
__device__ __forceinline__ uint2 fp_pack4(const float (&v)[4], __half*) {
    const __half2 a = __floats2half2_rn(v[0], v[1]), b = __floats2half2_rn(v[2], v[3]);
    return make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
__device__ __forceinline__ uint2 fp_pack4(const float (&v)[4], __nv_bfloat16*) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    return make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}

__device__ __forceinline__ void fp_epilogue_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(kFpEpilogueThreads) : "memory");
}

template <int SWIZZLE_BYTES>
__device__ __forceinline__ uint64_t fp_smem_desc_sbo(uint32_t smem_addr, uint32_t sbo_bytes) {
    constexpr uint64_t layout = SWIZZLE_BYTES == 128 ? 2ull : (SWIZZLE_BYTES == 64 ? 4ull : 6ull);
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) |
           (layout << 61);
}

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
upconv_fused_pair_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                         const FusedPairParams p) {
    using L = FusedPairLayout;
    constexpr int CC = kFpCC;
    constexpr int kChunks = CC / 8;        // 16-byte chunks per patch pixel
    constexpr int kQuads = CC / 4;         // 4-channel groups per pixel
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* a_empty = a_full + kFpAStages;
    uint64_t* b_full = a_empty + kFpAStages;
    uint64_t* b_empty = b_full + kFpBStages;
    uint64_t* tmem_full_bar = b_empty + kFpBStages;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < kFpAStages; ++i) {
            mbar_init(a_full + i, 1);
            mbar_init(a_empty + i, 1);
        }
        for (int i = 0; i < kFpBStages; ++i) {
            mbar_init(b_full + i, 1);
            mbar_init(b_empty + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, 2 * kFpEpilogueThreads / 32);  // the epilogue warps of both CTAs (leader's copy)
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_base_slot, 512);
    tcgen05_fence_before();
    cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);
    const int m_tiles = p.batch * p.tiles_h * p.tiles_w;
    const int total_items = ((m_tiles + 1) >> 1) * p.n_tiles;
    const int first_item = (int)cluster_id_x(), item_stride = (int)cluster_count_x();

    if (warp == 0) {
        // ===== TMA producer (both CTAs): per K chunk one halo box, then per shift this CTA's half of the weight rows =====
        const bool leader_lane = elect_one_sync();
        int a_stage = 0, b_stage = 0;
        uint32_t a_phase = 0, b_phase = 0;
        const uint32_t a_full_leader = cluster_map(smem_u32(a_full), 0), b_full_leader = cluster_map(smem_u32(b_full), 0);
        for (int item = first_item; item < total_items; item += item_stride) {
            int nt, b, ty, tx;
            fused_pair_decode(p, item, rank, nt, b, ty, tx);  // a ghost tile has b == batch: the TMA unit zero-fills its box
            const int h0 = ty * (kFpCellsH - 2) - 1, w0 = tx * (kFpCellsW - 2) - 1;
            for (int kc = 0; kc < p.k_chunks; ++kc) {
                // weights first: their ring is one K chunk deeper than the activations', so it must not queue behind them
                mbar_wait(b_empty + b_stage, b_phase ^ 1u);
                if (leader_lane) {
                    if (rank == 0) mbar_expect_tx(b_full + b_stage, 2u * kFpBStageBytes);
                    uint8_t* dst = smem + L::kBOffset + b_stage * kFpBStageBytes;
                    const uint32_t bar = b_full_leader + (uint32_t)(b_stage * 8);
#pragma unroll
                    for (int sh = 0; sh < 4; ++sh) {
                        const int nblk = sh == 0 ? 4 : (sh == 3 ? 1 : 2);
                        const int row0 = (sh == 0 ? 0 : sh == 1 ? 4 : sh == 2 ? 6 : 8) * p.c_out + nt * nblk * CC;
                        const int half_rows = nblk * CC / 2;                                   // 128 / 64 / 64 / 32
                        const int first = (sh == 0 ? 0 : sh == 1 ? 4 : sh == 2 ? 6 : 8) * CC / 2;  // row of this tile inside the stage
                        for (int j = 0; j < half_rows / kFpBRowsPerLoad; ++j)
                            tma_load_2d_pair(dst + (first + j * kFpBRowsPerLoad) * kFpK * 2, &tmap_w, bar, kc * kFpK,
                                             row0 + rank * half_rows + j * kFpBRowsPerLoad);
                    }
                }
                __syncwarp();
                if (++b_stage == kFpBStages) {
                    b_stage = 0;
                    b_phase ^= 1u;
                }
                mbar_wait(a_empty + a_stage, a_phase ^ 1u);
                if (leader_lane) {
                    if (rank == 0) mbar_expect_tx(a_full + a_stage, 2u * kFpHaloBytes);
                    tma_load_4d_pair(smem + a_stage * kFpHaloStride, &tmap_x, a_full_leader + (uint32_t)(a_stage * 8), kc * kFpK,
                                     w0 - 1, h0 - 1, b);
                }
                __syncwarp();
                if (++a_stage == kFpAStages) {
                    a_stage = 0;
                    a_phase ^= 1u;
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            // ===== MMA issuer (leader CTA) =====
            const bool leader_lane = elect_one_sync();
            constexpr bool kHalf = !std::is_same<T, __nv_bfloat16>::value;
            constexpr uint32_t idesc4 = make_instr_desc(2 * kTileM, 4 * CC, kHalf), idesc2 = make_instr_desc(2 * kTileM, 2 * CC, kHalf),
                               idesc1 = make_instr_desc(2 * kTileM, CC, kHalf);
            int a_stage = 0, b_stage = 0, acc = 0;
            uint32_t a_phase = 0, b_phase = 0, acc_phase = 0;
            const uint32_t smem_base = smem_u32(smem);
            for (int item = first_item; item < total_items; item += item_stride) {
                mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
                tcgen05_fence_after();
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    mbar_wait(a_full + a_stage, a_phase);
                    mbar_wait(b_full + b_stage, b_phase);
                    tcgen05_fence_after();
                    const uint32_t halo = smem_base + (uint32_t)(a_stage * kFpHaloStride);
                    const uint32_t weights = smem_base + (uint32_t)(L::kBOffset + b_stage * kFpBStageBytes);
#pragma unroll
                    for (int sh = 0; sh < 4; ++sh) {
                        const int dy = sh & 1, dx = sh >> 1;
                        const uint32_t idesc = sh == 0 ? idesc4 : (sh == 3 ? idesc1 : idesc2);
                        const int first_block = sh == 0 ? 0 : (sh == 1 ? 0 : 1);
                        const int first = (sh == 0 ? 0 : sh == 1 ? 4 : sh == 2 ? 6 : 8) * CC / 2;
                        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 4 * CC + first_block * CC);
                        const uint64_t a_desc = fp_smem_desc_sbo<128>(halo + (uint32_t)(((1 - dy) * kFpHaloW + (1 - dx)) * kFpK * 2),
                                                                      kFpHaloW * kFpK * 2);
                        const uint64_t b_desc = make_smem_desc<128>(weights + (uint32_t)(first * kFpK * 2));
                        if (leader_lane) {
#pragma unroll
                            for (int k = 0; k < kFpK / kUmmaK; ++k)
                                umma_bf16_pair(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc,
                                               (sh | kc | k) != 0 ? 1u : 0u);  // shift (0,0) touches every column first
                        }
                    }
                    if (leader_lane) {
                        umma_commit_pair(a_empty + a_stage);
                        umma_commit_pair(b_empty + b_stage);
                        if (kc == p.k_chunks - 1) umma_commit_pair(tmem_full_bar + acc);
                    }
                    __syncwarp();
                    if (++a_stage == kFpAStages) {
                        a_stage = 0;
                        a_phase ^= 1u;
                    }
                    if (++b_stage == kFpBStages) {
                        b_stage = 0;
                        b_phase ^= 1u;
                    }
                }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1u;
            }
        }
    } else {
        // ===== epilogue (each CTA its own tile) =====
        const int ew = warp - 2, quarter = warp & 3, half = ew >> 2;
        const int etid = ew * 32 + lane;
        const int row = quarter * 32 + lane;            // accumulator row = cell
        const int wl = row & (kFpCellsW - 1), hl = row / kFpCellsW;
        const uint32_t patch = smem_u32(smem + L::kPatchOffset);
        const int h_out = 2 * p.h_in, w_out = 2 * p.w_in;
        // filter role: quad = 4 channels, bj = output column pair (6 of 8 slots used), rs = row group
        const int quad = etid % kQuads;
        const int bj = (etid / kQuads) % 8, rs = etid / (kQuads * 8);
        const bool filters = bj < kFpOutW / 2;
        const int chunk = quad >> 1, hb = quad & 1;
        // patch row of output row i of this thread, tap 0 = rs * kFpRowsPerThread + i + 1; patch column of tap 0 = 2 bj + 1.
        // col[c] = byte offset (from the row start) of this thread's 8 bytes of patch column 2 bj + 1 + c, swizzle included
        uint32_t col[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            const int column = (2 * bj + 1 + c) & (kFpTW - 1);  // the idle slots (bj >= 6) stay inside the row
            col[c] = (uint32_t)(column * L::kPixelBytes + (((chunk ^ ((column >> 1) & (kChunks - 1))) << 4) + hb * 8));
        }
        const uint32_t first_row = patch + (uint32_t)((rs * kFpRowsPerThread + 1) * kFpTW * L::kPixelBytes);
        constexpr uint32_t kRowBytes = kFpTW * L::kPixelBytes;
        const size_t out_row_stride = (size_t)w_out * p.c_out;
        const uint32_t leader_empty_base = cluster_map(smem_u32(tmem_empty_bar), 0);
        int acc = 0;
        uint32_t full_phase = 0;
        for (int item = first_item; item < total_items; item += item_stride) {
            int nt, b, ty, tx;
            const bool ghost = fused_pair_decode(p, item, rank, nt, b, ty, tx);
            if (ghost) {  // CTA-uniform: only the accumulator hand-shake
                mbar_wait(tmem_full_bar + acc, full_phase);
                tcgen05_fence_after();
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
                acc ^= 1;
                if (acc == 0) full_phase ^= 1u;
                continue;
            }
            // per-tile global operands of the filter step, requested BEFORE the TMEM wait so their latency hides behind
            // the accumulator drain
            const int c0 = nt * CC + quad * 4;
            const int x = tx * kFpOutW + 2 * bj;
            const int y0 = ty * kFpOutH + rs * kFpRowsPerThread;
            const int rows_valid = (filters && x < w_out) ? h_out - y0 : 0;  // output rows i < rows_valid are stored
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f), s4 = make_float4(1.f, 1.f, 1.f, 1.f), d4 = s4;
            float2 nz[kFpRowsPerThread];
#pragma unroll
            for (int i = 0; i < kFpRowsPerThread; ++i) nz[i] = make_float2(0.f, 0.f);
            if (rows_valid > 0) {
                b4 = __ldg(reinterpret_cast<const float4*>(p.bias + c0));
                d4 = __ldg(reinterpret_cast<const float4*>(p.demod + (size_t)b * p.c_out + c0));
                if (p.next_styles) s4 = __ldg(reinterpret_cast<const float4*>(p.next_styles + (size_t)b * p.c_out + c0));
                if (p.noise) {
                    const float* noise_b = p.noise + (size_t)b * p.noise_batch_stride + (size_t)y0 * w_out + x;
#pragma unroll
                    for (int i = 0; i < kFpRowsPerThread; ++i)
                        if (i < rows_valid) nz[i] = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)i * w_out));
                }
            }
            // ---- (a) TMEM -> t patch: the RAW accumulator as 16-bit ----
            mbar_wait(tmem_full_bar + acc, full_phase);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * 4 * CC + half * 2 * CC);
#pragma unroll
            for (int blk = 0; blk < 2; ++blk) {
                const int block = half * 2 + blk;                      // P01, P00 | P10, P11
                const int a = block >> 1, bb = (block == 0 || block == 3) ? 1 : 0;
                const uint32_t pixel = patch + (uint32_t)(((2 * hl + a) * kFpTW + 2 * wl + bb) * L::kPixelBytes);
#pragma unroll
                for (int cc = 0; cc < CC; cc += 32) {
                    uint32_t r[32];
                    tmem_ld_32x32(taddr + (uint32_t)(blk * CC + cc), r);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[g * 8 + e]);
                        Vec<T> packed;
                        packed.from_float(v);
                        const int j = (cc >> 3) + g;
                        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(
                                         pixel + (uint32_t)((j ^ (wl & (kChunks - 1))) << 4)),
                                     "r"(packed.raw.x), "r"(packed.raw.y), "r"(packed.raw.z), "r"(packed.raw.w)
                                     : "memory");
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
            fp_epilogue_barrier();
            // ---- (c) FIR + demod + noise + bias + lrelu + next style from the patch ----
            if (rows_valid > 0) {
                const float bs[4] = {b4.x, b4.y, b4.z, b4.w};
                const float coef[4] = {d4.x * 0.0625f, d4.y * 0.0625f, d4.z * 0.0625f, d4.w * 0.0625f};
                // lrelu(v) * g = v * 0.6 g + |v| * 0.4 g  (alpha = 0.2), g = sqrt(2) * next style
                const float pos[4] = {s4.x * (0.6f * kLreluGain), s4.y * (0.6f * kLreluGain), s4.z * (0.6f * kLreluGain),
                                      s4.w * (0.6f * kLreluGain)};
                const float neg[4] = {s4.x * (0.4f * kLreluGain), s4.y * (0.4f * kLreluGain), s4.z * (0.4f * kLreluGain),
                                      s4.w * (0.4f * kLreluGain)};
                T* out_ptr = reinterpret_cast<T*>(p.out) + (((size_t)b * h_out + y0) * w_out + x) * p.c_out + c0;
                float window[4][2][4];  // horizontally filtered patch rows (x4 scale), both x phases, 4 channels
                auto filter_row = [&](uint32_t row_addr, float (&dst)[2][4]) {
                    float v[5][4];
#pragma unroll
                    for (int c = 0; c < 5; ++c) {
                        uint2 raw;
                        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(raw.x), "=r"(raw.y) : "r"(row_addr + col[c]));
                        fp_unpack4<T>(raw, v[c]);
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        dst[0][q] = fmaf(3.f, v[1][q] + v[2][q], v[0][q] + v[3][q]);
                        dst[1][q] = fmaf(3.f, v[2][q] + v[3][q], v[1][q] + v[4][q]);
                    }
                };
#pragma unroll
                for (int r = 0; r < 3; ++r) filter_row(first_row + r * kRowBytes, window[r]);
                const uint32_t row_addr = first_row + 3 * kRowBytes;
#pragma unroll
                for (int j = 0; j < kFpRowsPerThread; ++j) {
                    filter_row(row_addr + j * kRowBytes, window[(j + 3) & 3]);
                    if (j < rows_valid) {
#pragma unroll
                        for (int px = 0; px < 2; ++px) {
                            const float nzv = (px ? nz[j].y : nz[j].x) * p.noise_strength;
                            float o[4];
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float sum = fmaf(3.f, window[(j + 1) & 3][px][q] + window[(j + 2) & 3][px][q],
                                                       window[j & 3][px][q] + window[(j + 3) & 3][px][q]);
                                const float value = fmaf(sum, coef[q], nzv + bs[q]);
                                o[q] = fmaf(fabsf(value), neg[q], value * pos[q]);
                            }
                            *reinterpret_cast<uint2*>(out_ptr + px * p.c_out) = fp_pack4(o, (T*)nullptr);
                        }
                    }
                    out_ptr += out_row_stride;
                }
            }
            fp_epilogue_barrier();  // the patch is free for the next tile
            acc ^= 1;
            if (acc == 0) full_phase ^= 1u;
        }
    }

    // neither CTA leaves while its partner can still read its shared memory, write its TMEM or signal its barriers
    tcgen05_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc_pair(tmem_base, 512);
    }
}

int encode_tensor_map(CUtensorMap* map, const void* base, bool half_operands, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes, const cuuint32_t* box, int swizzle_bytes);
int sm_count();

static int g_fused_pair_clusters = 0;

template <typename T>
int launch_fused_pair(const void* x, int batch, int h_in, int w_in, int c_in, int c_out, const void* weight_phase,
                      const float* demod, const float* noise, long long noise_batch_stride, float noise_strength,
                      const float* bias, const float* next_styles, void* out, cudaStream_t stream) {
    constexpr bool kHalf = std::is_same<T, __half>::value;
    using L = FusedPairLayout;
    FusedPairParams p;
    p.batch = batch; p.h_in = h_in; p.w_in = w_in; p.c_in = c_in; p.c_out = c_out;
    p.k_chunks = c_in / kFpK; p.n_tiles = c_out / kFpCC;
    p.tiles_w = (2 * w_in + kFpOutW - 1) / kFpOutW; p.tiles_h = (2 * h_in + kFpOutH - 1) / kFpOutH;
    p.demod = demod; p.noise = noise; p.noise_batch_stride = noise_batch_stride; p.noise_strength = noise_strength;
    p.bias = bias; p.next_styles = next_styles; p.out = out;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)c_in, (cuuint64_t)w_in, (cuuint64_t)h_in, (cuuint64_t)batch};
        cuuint64_t strides[3] = {(cuuint64_t)c_in * 2, (cuuint64_t)w_in * c_in * 2, (cuuint64_t)h_in * w_in * c_in * 2};
        cuuint32_t box[4] = {(cuuint32_t)kFpK, (cuuint32_t)kFpHaloW, (cuuint32_t)kFpHaloH, 1u};
        int rc = encode_tensor_map(&tmap_x, x, kHalf, 4, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)c_in, (cuuint64_t)9 * c_out};
        cuuint64_t strides[1] = {(cuuint64_t)c_in * 2};
        cuuint32_t box[2] = {(cuuint32_t)kFpK, (cuuint32_t)kFpBRowsPerLoad};
        int rc = encode_tensor_map(&tmap_w, weight_phase, kHalf, 2, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    const long long m_tiles = (long long)batch * p.tiles_h * p.tiles_w;
    const long long items = ((m_tiles + 1) / 2) * p.n_tiles;
    const int clusters = items < g_fused_pair_clusters ? (int)items : g_fused_pair_clusters;
    upconv_fused_pair_kernel<T><<<2 * clusters, kThreads, L::kSmemBytes, stream>>>(tmap_x, tmap_w, p);
    return check_launch("upconv_fused_pair_kernel");
}

template int launch_fused_pair<__half>(const void*, int, int, int, int, int, const void*, const float*, const float*, long long,
                                       float, const float*, const float*, void*, cudaStream_t);
template int launch_fused_pair<__nv_bfloat16>(const void*, int, int, int, int, int, const void*, const float*, const float*,
                                              long long, float, const float*, const float*, void*, cudaStream_t);

template <typename T>
static int configure_fused_pair_one(int* clusters) {
    auto kernel = upconv_fused_pair_kernel<T>;
    cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FusedPairLayout::kSmemBytes);
    if (err != cudaSuccess)
        return fail(-(int)err, "cudaFuncSetAttribute(upconv_fused_pair_kernel): %s", cudaGetErrorString(err));
    cudaLaunchConfig_t config = {};
    config.gridDim = dim3(2 * sm_count(), 1, 1);
    config.blockDim = dim3(kThreads, 1, 1);
    config.dynamicSmemBytes = FusedPairLayout::kSmemBytes;
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
    if (n < *clusters) *clusters = n;
    return GANCE_OK;
}

int configure_fused_pair_kernels() {
    int n = 1 << 30, rc;
    if ((rc = configure_fused_pair_one<__half>(&n)) != GANCE_OK) return rc;
    if ((rc = configure_fused_pair_one<__nv_bfloat16>(&n)) != GANCE_OK) return rc;
    g_fused_pair_clusters = n == (1 << 30) ? 0 : n;
    return GANCE_OK;
}

int fused_pair_clusters() { return g_fused_pair_clusters; }

}  // namespace tc
}  // namespace gance
