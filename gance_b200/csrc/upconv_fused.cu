// Up-convolution in ONE kernel at (close to) its algorithmic cost: the 4-phase transposed-convolution GEMM of
// upconv_phase.cu, whose epilogue keeps the transposed-conv result t ON CHIP (a 16-bit tile in shared memory) and
// applies the [1,3,3,1] FIR + noise + bias + leaky-ReLU + next-layer style from there. t never reaches HBM: the layer
// reads x once and writes its output once (upconv_phase.cu writes and re-reads the (2H+2)^2 x Cout intermediate).
//
// Reference: upsample_conv_2d (dnnlib/tflib/ops/upfirdn_2d.py) = conv2d_transpose(stride 2, VALID) -> upfirdn_2d(pad 1,1),
// restated in oracle/stylegan2_ref.py:109-123. Cell (m,n) of the GEMM holds t[2m+a, 2n+b] for the four phases (a,b) as one
// accumulator row [P01 | P00 | P10 | P11] x CC channels (see upconv_phase.cu for the shift / phase algebra).
//
// Tiling. One tile = 16 x 8 cells (128 accumulator rows) = a 32 x 16-pixel patch of t. out[y,x] needs t[y-1..y+2, x-1..x+2],
// so the patch yields 28 x 12 output pixels; tiles overlap by 2 cells per axis (the halo is RECOMPUTED by the tensor core:
// 128 / 84 = 1.52x the algorithmic MACs instead of the folded form's 4x, and no HBM round trip). Cell coordinates -1 and
// > H address input pixels outside the image, which the TMA unit zero-fills: exactly the FIR's zero padding.
//
// Warp roles (320 threads, one persistent CTA per SM): warp 0 TMA producer, warp 1 MMA issuer (both as in
// upconv_phase_kernel), warps 2-9 epilogue. Per tile the 8 epilogue warps (a) drain the TMEM accumulator (each warp its
// lane quarter x half of the columns) and store it as 16-bit t into the shared patch
// (16-byte chunks XOR-swizzled by the cell column so the row-per-thread stores are conflict-free), release the TMEM
// buffer to the MMA warp, (b) barrier, (c) filter: 224 threads = 14 column pairs x 16 channel quads walk DOWN the patch
// with a 4-row sliding window of horizontally filtered rows (every t element is converted and filtered horizontally once
// per thread), apply demodulation (after the filter: it is linear per channel) + noise + bias, lrelu, scale by
// sqrt(2) * next style, and store 8-byte NHWC vectors, (d) barrier.
// The MMA of tile i+1 (other TMEM buffer) overlaps the epilogue of tile i.
#include "tc_ptx.cuh"

namespace gance {
namespace tc {

constexpr int kFuCellsW = 16, kFuCellsH = 8;                   // cells per tile (= kTileM accumulator rows)
constexpr int kFuOutW = 2 * (kFuCellsW - 2), kFuOutH = 2 * (kFuCellsH - 2);  // 28 x 12 output pixels per tile
constexpr int kFuTW = 2 * kFuCellsW, kFuTH = 2 * kFuCellsH;    // 32 x 16 pixel patch of t
constexpr int kFuK = 64;                                       // input channels per pipeline stage
constexpr int kFuABytes = kTileM * kFuK * 2;                   // 16 KB

// PAIR: two CTAs (cta_group::2) share the MMAs of two tiles of one channel block; each holds HALF of a shift's weight rows,
// so a stage is 16 + <= 16 KB instead of 16 + <= 32 KB and five of them fit beside the patch.
template <int CC, bool PAIR = false>
struct FusedLayout {
    static constexpr int kStages = PAIR ? 5 : 3;  // measured at batch 32 (single CTA): 2 stages cost +6 % (Cin 128) to +31 % (Cin 512); a 4th (patch halved, two epilogue passes) gained < 4 % - profiles/README.md
    static constexpr int kBBlock = CC * kFuK * 2;              // one phase block of the B operand
    static constexpr int kStageBytes = kFuABytes + (PAIR ? 2 : 4) * kBBlock;
    static constexpr int kPatchOffset = kStages * kStageBytes;
    static constexpr int kPixelBytes = CC * 2;
    static constexpr int kPatchBytes = kFuTW * kFuTH * kPixelBytes;
    static constexpr int kBarrierOffset = kPatchOffset + kPatchBytes;
    static constexpr int kSmemBytes = kBarrierOffset + (2 * kStages + 4) * 8 + 16 + 1024;
};

struct FusedParams {
    int batch, h_in, w_in, c_in, c_out, k_chunks, n_tiles, tiles_w, tiles_h;
    const float* demod;        // [B, c_out]
    const float* noise;        // [2H,2W] or [B,2H,2W] or NULL
    long long noise_batch_stride;
    float noise_strength;
    const float* bias;         // [c_out]
    const float* next_styles;  // [B, c_out] or NULL
    void* out;                 // [B, 2H, 2W, c_out]
};

__device__ __forceinline__ void fused_decode(const FusedParams& p, int tile, int& nt, int& b, int& ty, int& tx) {
    nt = tile % p.n_tiles;
    int m = tile / p.n_tiles;
    tx = m % p.tiles_w;
    m /= p.tiles_w;
    ty = m % p.tiles_h;
    b = m / p.tiles_h;
}

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

// Eight epilogue warps. Sixteen (each thread filtering half of the rows; -DGANCE_ABL_FU_EW16 builds it) were measured 5 %
// SLOWER on the layer the filter step bounds (256^2 -> 512^2: 750 vs 715 us): the step is not short of warps but of
// shared-memory bandwidth - every row group re-reads three warm-up rows of the patch (ncu: LSU + tensor-core wavefronts
// together ~60 % of the shared-memory pipe, next to the TMA writes).
#ifdef GANCE_ABL_FU_EW16
__host__ __device__ constexpr int fused_epilogue_warps(int cc) { return cc == 64 ? 16 : 8; }
#else
__host__ __device__ constexpr int fused_epilogue_warps(int) { return 8; }
#endif

template <int THREADS>
__device__ __forceinline__ void epilogue_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(THREADS) : "memory");
}

// EW = epilogue warps (8; 16 is an A/B build, see fused_epilogue_warps).
// PAIR (launched as 2-CTA clusters): the two CTAs take two tiles of the same channel block; TMA bytes of both are counted
// on the leader's full barriers, the leader's MMA warp issues M = 256 MMAs for both, commits are multicast, the epilogue
// warps of both CTAs report to the leader. Same K order, same filter arithmetic: same bits as the single-CTA flavour.
template <typename T, int CC, int EW, bool PAIR>
__global__ void __launch_bounds__(64 + EW * 32, 1)
upconv_fused_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                    const FusedParams p) {
    using L = FusedLayout<CC, PAIR>;
    static_assert(!PAIR || CC == 64, "the CTA-pair flavour exists for 64-channel blocks");
    constexpr int kEpilogueThreads = EW * 32;
    constexpr int kChunks = CC / 8;        // 16-byte chunks per patch pixel
    constexpr int kQuads = CC / 4;         // 4-channel groups per pixel
    constexpr int kRowSplit = kEpilogueThreads / (kQuads * 16);  // row groups: 1 (CC = 64, 8 warps) or 2
    constexpr int kDrainBlocks = 16 / EW;  // 64-/32-channel phase blocks each warp copies out of TMEM (its lane quarter)
    static_assert(EW == 8 || EW == 16, "8 or 16 epilogue warps");
    constexpr int kRowsPerThread = kFuOutH / kRowSplit;
    constexpr int kInner = kRowsPerThread % 4 == 0 ? 4 : kRowsPerThread;  // rows per rolled iteration (window slots are static)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    uint64_t* empty_bar = full_bar + L::kStages;
    uint64_t* tmem_full_bar = empty_bar + L::kStages;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
    const int lane = threadIdx.x & 31;
    const int rank = PAIR ? (int)cluster_ctarank() : 0;
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_x);
        prefetch_tmap(&tmap_w);
        for (int i = 0; i < L::kStages; ++i) {
            mbar_init(full_bar + i, 1);
            mbar_init(empty_bar + i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(tmem_full_bar + i, 1);
            mbar_init(tmem_empty_bar + i, PAIR ? 2 * EW : EW);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        if (PAIR) tmem_alloc_pair(tmem_base_slot, 512);
        else tmem_alloc(tmem_base_slot, 512);
    }
    tcgen05_fence_before();
    if (PAIR) cluster_sync_all();
    else __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_base_slot, 0);
    // work items: tiles (single CTA) or pairs of tiles of one channel block (PAIR), walked CTA by CTA / cluster by cluster
    const int m_tiles = p.batch * p.tiles_h * p.tiles_w;
    const int total_tiles = PAIR ? ((m_tiles + 1) >> 1) * p.n_tiles : m_tiles * p.n_tiles;
    const int walk_first = PAIR ? (int)cluster_id_x() : (int)blockIdx.x;
    const int walk_stride = PAIR ? (int)cluster_count_x() : (int)gridDim.x;
    // item -> this CTA's tile; a PAIR's odd tile out is a ghost (sample index = batch: TMA zero-fills, nothing is stored)
    auto decode = [&](int item, int& nt, int& b, int& ty, int& tx) {
        if (!PAIR) {
            fused_decode(p, item, nt, b, ty, tx);
            return false;
        }
        nt = item % p.n_tiles;
        int m = 2 * (item / p.n_tiles) + rank;
        tx = m % p.tiles_w;
        m /= p.tiles_w;
        ty = m % p.tiles_h;
        b = m / p.tiles_h;
        return b >= p.batch;
    };

    if (warp == 0) {
        // ===== TMA producer: per (shift, K chunk) one activation box + 4/2/2/1 weight blocks =====
        const bool leader = elect_one_sync();
        int stage = 0;
        uint32_t phase = 0;
        const uint32_t full_leader = PAIR ? cluster_map(smem_u32(full_bar), 0) : 0u;
        for (int tile = walk_first; tile < total_tiles; tile += walk_stride) {
            int nt, b, ty, tx;
            decode(tile, nt, b, ty, tx);
            const int h0 = ty * (kFuCellsH - 2) - 1, w0 = tx * (kFuCellsW - 2) - 1;
            // K chunk by K chunk, the four shifts inside: the accumulation order of upconv_fused_pair.cu, so the two forms
            // produce the same bits
            for (int kc = 0; kc < p.k_chunks; ++kc) {
#pragma unroll
                for (int sh = 0; sh < 4; ++sh) {
                    const int dy = sh & 1, dx = sh >> 1;
                    const int nblk = sh == 0 ? 4 : (sh == 3 ? 1 : 2);
                    const int row0 = (sh == 0 ? 0 : sh == 1 ? 4 : sh == 2 ? 6 : 8) * p.c_out + nt * nblk * CC;
                    mbar_wait(empty_bar + stage, phase ^ 1u);
                    uint8_t* a_dst = smem + stage * L::kStageBytes;
                    if (leader) {
                        if (PAIR) {  // own activation box + own half of the shift's weight rows (in boxes of CC / 2 rows)
                            if (rank == 0) mbar_expect_tx(full_bar + stage, 2u * (uint32_t)(kFuABytes + nblk * L::kBBlock / 2));
                            const uint32_t bar = full_leader + (uint32_t)(stage * 8);
                            tma_load_4d_pair(a_dst, &tmap_x, bar, kc * kFuK, w0 - dx, h0 - dy, b);
                            for (int j = 0; j < nblk; ++j)
                                tma_load_2d_pair(a_dst + kFuABytes + j * (L::kBBlock / 2), &tmap_w, bar, kc * kFuK,
                                                 row0 + rank * (nblk * CC / 2) + j * (CC / 2));
                        } else {
                            mbar_expect_tx(full_bar + stage, (uint32_t)(kFuABytes + nblk * L::kBBlock));
                            tma_load_4d(a_dst, &tmap_x, full_bar + stage, kc * kFuK, w0 - dx, h0 - dy, b);
                            for (int j = 0; j < nblk; ++j)
                                tma_load_2d(a_dst + kFuABytes + j * L::kBBlock, &tmap_w, full_bar + stage, kc * kFuK,
                                            row0 + j * CC);
                        }
                    }
                    __syncwarp();
                    if (++stage == L::kStages) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1 && rank == 0) {
        // ===== MMA issuer (PAIR: the leader CTA's, for both) =====
        const bool leader = elect_one_sync();
        constexpr bool kHalf = !std::is_same<T, __nv_bfloat16>::value;
        constexpr int kM = PAIR ? 2 * kTileM : kTileM;
        constexpr uint32_t idesc4 = make_instr_desc(kM, 4 * CC, kHalf), idesc2 = make_instr_desc(kM, 2 * CC, kHalf),
                           idesc1 = make_instr_desc(kM, CC, kHalf);
        int stage = 0, acc = 0;
        uint32_t phase = 0, acc_phase = 0;
        const uint32_t smem_base = smem_u32(smem);
        for (int tile = walk_first; tile < total_tiles; tile += walk_stride) {
            mbar_wait(tmem_empty_bar + acc, acc_phase ^ 1u);
            tcgen05_fence_after();
            for (int kc = 0; kc < p.k_chunks; ++kc) {
#pragma unroll
                for (int sh = 0; sh < 4; ++sh) {
                    const uint32_t idesc = sh == 0 ? idesc4 : (sh == 3 ? idesc1 : idesc2);
                    const int first_block = sh == 0 ? 0 : (sh == 1 ? 0 : 1);
                    const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 4 * CC + first_block * CC);
                    mbar_wait(full_bar + stage, phase);
                    tcgen05_fence_after();
                    const uint64_t a_desc = make_smem_desc<128>(smem_base + (uint32_t)(stage * L::kStageBytes));
                    const uint64_t b_desc = a_desc + (uint64_t)(kFuABytes >> 4);
                    if (leader) {
#pragma unroll
                        for (int k = 0; k < kFuK / kUmmaK; ++k) {  // shift (0,0) touches every column first
                            if (PAIR) umma_bf16_pair(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (sh | kc | k) != 0 ? 1u : 0u);
                            else umma_bf16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (sh | kc | k) != 0 ? 1u : 0u);
                        }
                        if (PAIR) {
                            umma_commit_pair(empty_bar + stage);
                            if (sh == 3 && kc == p.k_chunks - 1) umma_commit_pair(tmem_full_bar + acc);
                        } else {
                            umma_commit(empty_bar + stage);
                            if (sh == 3 && kc == p.k_chunks - 1) umma_commit(tmem_full_bar + acc);
                        }
                    }
                    __syncwarp();
                    if (++stage == L::kStages) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1u;
        }
    } else if (warp >= 2) {
        // ===== epilogue =====
        const int ew = warp - 2, quarter = warp & 3, part = ew >> 2;  // part: which of the EW / 4 column ranges this warp drains
        const int etid = ew * 32 + lane;
        const int row = quarter * 32 + lane;            // accumulator row = cell
        const int wl = row & (kFuCellsW - 1), hl = row / kFuCellsW;
        const uint32_t patch = smem_u32(smem + L::kPatchOffset);
        const int h_out = 2 * p.h_in, w_out = 2 * p.w_in;
        // filter role: quad = 4 channels, bj = output column pair, rs = row group (CC = 32 only)
        const int quad = etid % kQuads;
        const int bj = (etid / kQuads) % 16, rs = etid / (kQuads * 16);
        const bool filters = bj < kFuOutW / 2 && rs < kRowSplit;
        const int chunk = quad >> 1, hb = quad & 1;
        // patch row of output row i of this thread, tap 0 = rs * kRowsPerThread + i + 1; patch column of tap 0 = 2 bj + 1.
        // col[c] = byte offset (from the row start) of this thread's 8 bytes of patch column 2 bj + 1 + c, swizzle included
        uint32_t col[5];
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            const int column = 2 * bj + 1 + c;
            col[c] = (uint32_t)(column * L::kPixelBytes + (((chunk ^ ((column >> 1) & (kChunks - 1))) << 4) + hb * 8));
        }
        const uint32_t first_row = patch + (uint32_t)((rs * kRowsPerThread + 1) * kFuTW * L::kPixelBytes);
        constexpr uint32_t kRowBytes = kFuTW * L::kPixelBytes;
        const size_t out_row_stride = (size_t)w_out * p.c_out;
        int acc = 0;
        uint32_t full_phase = 0;
        const uint32_t leader_empty_base = PAIR ? cluster_map(smem_u32(tmem_empty_bar), 0) : 0u;
        for (int tile = walk_first; tile < total_tiles; tile += walk_stride) {
            int nt, b, ty, tx;
            if (decode(tile, nt, b, ty, tx)) {  // ghost tile (PAIR, CTA-uniform): only the accumulator hand-shake
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
            // the accumulator drain (issued inside the row loop they cost ~600 exposed cycles per output row)
            const int c0 = nt * CC + quad * 4;
            const int x = tx * kFuOutW + 2 * bj;
            const int y0 = ty * kFuOutH + rs * kRowsPerThread;
            const int rows_valid = (filters && x < w_out) ? h_out - y0 : 0;  // output rows i < rows_valid are stored
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f), s4 = make_float4(1.f, 1.f, 1.f, 1.f), d4 = s4;
            float2 nz[kRowsPerThread];
#pragma unroll
            for (int i = 0; i < kRowsPerThread; ++i) nz[i] = make_float2(0.f, 0.f);
            if (rows_valid > 0) {
                b4 = __ldg(reinterpret_cast<const float4*>(p.bias + c0));
                d4 = __ldg(reinterpret_cast<const float4*>(p.demod + (size_t)b * p.c_out + c0));
                if (p.next_styles) s4 = __ldg(reinterpret_cast<const float4*>(p.next_styles + (size_t)b * p.c_out + c0));
                if (p.noise) {
                    const float* noise_b = p.noise + (size_t)b * p.noise_batch_stride + (size_t)y0 * w_out + x;
#pragma unroll
                    for (int i = 0; i < kRowsPerThread; ++i)
                        if (i < rows_valid) nz[i] = __ldg(reinterpret_cast<const float2*>(noise_b + (size_t)i * w_out));
                }
            }
            // ---- (a) TMEM -> t patch: the RAW accumulator as 16-bit (the FIR is linear per channel, so the demodulation
            //      coefficient is applied after it, folded into the filter's 1/16 normalisation) ----
            mbar_wait(tmem_full_bar + acc, full_phase);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * 4 * CC + part * kDrainBlocks * CC);
#ifdef GANCE_ABL_FU_NODRAIN
            if (p.batch < 0)  // ablation: the accumulator is not copied to the patch
#endif
#pragma unroll
            for (int blk = 0; blk < kDrainBlocks; ++blk) {
                const int block = part * kDrainBlocks + blk;           // P01, P00 | P10, P11
                const int a = block >> 1, bb = (block == 0 || block == 3) ? 1 : 0;
                const uint32_t pixel = patch + (uint32_t)(((2 * hl + a) * kFuTW + 2 * wl + bb) * L::kPixelBytes);
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
            if (lane == 0) {
                if (PAIR) mbar_arrive_cluster(leader_empty_base + (uint32_t)(acc * 8));
                else mbar_arrive(tmem_empty_bar + acc);
            }
            epilogue_barrier<kEpilogueThreads>();
            // ---- (c) FIR + demod + noise + bias + lrelu + next style from the patch ----
#ifdef GANCE_ABL_FU_NOFILTER
            if (rows_valid > 0 && p.batch < 0) {  // ablation: drain only
#else
            if (rows_valid > 0) {
#endif
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
                        unpack4<T>(raw, v[c]);
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        dst[0][q] = fmaf(3.f, v[1][q] + v[2][q], v[0][q] + v[3][q]);
                        dst[1][q] = fmaf(3.f, v[2][q] + v[3][q], v[1][q] + v[4][q]);
                    }
                };
#pragma unroll
                for (int r = 0; r < 3; ++r) filter_row(first_row + r * kRowBytes, window[r]);
                uint32_t row_addr = first_row + 3 * kRowBytes;
                int remaining = rows_valid;
#pragma unroll 1
                for (int it = 0; it < kRowsPerThread / kInner; ++it) {
#pragma unroll
                    for (int j = 0; j < kInner; ++j) {
                        filter_row(row_addr + j * kRowBytes, window[(j + 3) & 3]);
                        if (j < remaining) {
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
#ifdef GANCE_ABL_FU_NOSTORE
                                if (o[0] == 12345.678f)  // ablation: (practically) no global stores
#endif
                                *reinterpret_cast<uint2*>(out_ptr + px * p.c_out) = pack4(o, (T*)nullptr);
                            }
                        }
                        out_ptr += out_row_stride;
                    }
                    row_addr += kInner * kRowBytes;
                    remaining -= kInner;
                    if (kRowsPerThread > kInner) {
#pragma unroll
                        for (int i = 0; i + kInner < kRowsPerThread; ++i) nz[i] = nz[i + kInner];
                    }
                }
            }
            epilogue_barrier<kEpilogueThreads>();  // the patch is free for the next tile
            acc ^= 1;
            if (acc == 0) full_phase ^= 1u;
        }
    }

    tcgen05_fence_before();
    if (PAIR) cluster_sync_all();  // neither CTA leaves while its partner can still read its shared memory or signal its barriers
    else __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        if (PAIR) tmem_dealloc_pair(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

int encode_tensor_map(CUtensorMap* map, const void* base, bool half_operands, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes, const cuuint32_t* box, int swizzle_bytes);
int sm_count();

static int g_fused_wide_pair_clusters = 0;  // co-resident 2-CTA clusters of the PAIR flavour (0 = unusable)

template <typename T, int CC, bool PAIR = false>
static int launch_fused(const void* x, int batch, int h_in, int w_in, int c_in, int c_out, const void* weight_phase,
                        const float* demod, const float* noise, long long noise_batch_stride, float noise_strength,
                        const float* bias, const float* next_styles, void* out, cudaStream_t stream) {
    constexpr bool kHalf = std::is_same<T, __half>::value;
    using L = FusedLayout<CC, PAIR>;
    FusedParams p;
    p.batch = batch; p.h_in = h_in; p.w_in = w_in; p.c_in = c_in; p.c_out = c_out;
    p.k_chunks = c_in / kFuK; p.n_tiles = c_out / CC;
    p.tiles_w = (2 * w_in + kFuOutW - 1) / kFuOutW; p.tiles_h = (2 * h_in + kFuOutH - 1) / kFuOutH;
    p.demod = demod; p.noise = noise; p.noise_batch_stride = noise_batch_stride; p.noise_strength = noise_strength;
    p.bias = bias; p.next_styles = next_styles; p.out = out;
    CUtensorMap tmap_x, tmap_w;
    {
        cuuint64_t dims[4] = {(cuuint64_t)c_in, (cuuint64_t)w_in, (cuuint64_t)h_in, (cuuint64_t)batch};
        cuuint64_t strides[3] = {(cuuint64_t)c_in * 2, (cuuint64_t)w_in * c_in * 2, (cuuint64_t)h_in * w_in * c_in * 2};
        cuuint32_t box[4] = {(cuuint32_t)kFuK, (cuuint32_t)kFuCellsW, (cuuint32_t)kFuCellsH, 1u};
        int rc = encode_tensor_map(&tmap_x, x, kHalf, 4, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)c_in, (cuuint64_t)9 * c_out};
        cuuint64_t strides[1] = {(cuuint64_t)c_in * 2};
        cuuint32_t box[2] = {(cuuint32_t)kFuK, (cuuint32_t)(PAIR ? CC / 2 : CC)};  // PAIR: each CTA loads half of a block's rows
        int rc = encode_tensor_map(&tmap_w, weight_phase, kHalf, 2, dims, strides, box, 128);
        if (rc != GANCE_OK) return rc;
    }
    constexpr int kBlock = 64 + fused_epilogue_warps(CC) * 32;
    if constexpr (PAIR) {
        const long long m_tiles = (long long)batch * p.tiles_h * p.tiles_w;
        const long long items = ((m_tiles + 1) / 2) * p.n_tiles;
        const int clusters = items < g_fused_wide_pair_clusters ? (int)items : g_fused_wide_pair_clusters;
        cudaLaunchConfig_t config = {};
        config.gridDim = dim3(2 * clusters, 1, 1);
        config.blockDim = dim3(kBlock, 1, 1);
        config.dynamicSmemBytes = L::kSmemBytes;
        config.stream = stream;
        cudaLaunchAttribute attribute;
        attribute.id = cudaLaunchAttributeClusterDimension;
        attribute.val.clusterDim.x = 2; attribute.val.clusterDim.y = 1; attribute.val.clusterDim.z = 1;
        config.attrs = &attribute;
        config.numAttrs = 1;
        cudaError_t err = cudaLaunchKernelEx(&config, upconv_fused_kernel<T, CC, fused_epilogue_warps(CC), true>, tmap_x, tmap_w, p);
        if (err != cudaSuccess) return fail(-(int)err, "cudaLaunchKernelEx(upconv_fused_kernel<pair>): %s", cudaGetErrorString(err));
    } else {
        const long long total_tiles = (long long)batch * p.tiles_h * p.tiles_w * p.n_tiles;
        const int grid = total_tiles < sm_count() ? (int)total_tiles : sm_count();
        upconv_fused_kernel<T, CC, fused_epilogue_warps(CC), false><<<grid, kBlock, L::kSmemBytes, stream>>>(tmap_x, tmap_w, p);
    }
    return check_launch("upconv_fused_kernel");
}

template <typename T, int CC>
static cudaError_t configure_one() {
    return cudaFuncSetAttribute(upconv_fused_kernel<T, CC, fused_epilogue_warps(CC), false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                FusedLayout<CC>::kSmemBytes);
}

template <typename T>
static cudaError_t configure_wide_pair(int* clusters) {
    constexpr int kBlock = 64 + fused_epilogue_warps(64) * 32;
    auto kernel = upconv_fused_kernel<T, 64, fused_epilogue_warps(64), true>;
    cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FusedLayout<64, true>::kSmemBytes);
    if (err != cudaSuccess) return err;
    cudaLaunchConfig_t config = {};
    config.gridDim = dim3(2 * sm_count(), 1, 1);
    config.blockDim = dim3(kBlock, 1, 1);
    config.dynamicSmemBytes = FusedLayout<64, true>::kSmemBytes;
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
    return cudaSuccess;
}

int configure_fused_kernels() {
    cudaError_t err = configure_one<__half, 64>();
    if (err == cudaSuccess) err = configure_one<__nv_bfloat16, 64>();
    if (err == cudaSuccess) err = configure_one<__half, 32>();
    if (err == cudaSuccess) err = configure_one<__nv_bfloat16, 32>();
    int clusters = 1 << 30;
    if (err == cudaSuccess) err = configure_wide_pair<__half>(&clusters);
    if (err == cudaSuccess) err = configure_wide_pair<__nv_bfloat16>(&clusters);
    g_fused_wide_pair_clusters = clusters == (1 << 30) ? 0 : clusters;
    if (err != cudaSuccess)
        return fail(-(int)err, "cudaFuncSetAttribute(upconv_fused_kernel): %s", cudaGetErrorString(err));
    return GANCE_OK;
}

// upconv_fused_pair.cu: CTA pairs + one activation halo box per K chunk, for the wide layers
template <typename T>
int launch_fused_pair(const void* x, int batch, int h_in, int w_in, int c_in, int c_out, const void* weight_phase,
                      const float* demod, const float* noise, long long noise_batch_stride, float noise_strength,
                      const float* bias, const float* next_styles, void* out, cudaStream_t stream);
int fused_pair_clusters();
int conv_pair_mode();

}  // namespace tc
}  // namespace gance

using namespace gance;

extern "C" int gance_modconv3x3_up2_fused(const void* x, int dtype, int batch, int h_in, int w_in, int c_in, int c_out,
                                          const void* weight_phase, const float* demod, const float* noise,
                                          long long noise_batch_stride, float noise_strength, const float* bias,
                                          const float* next_styles, void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && weight_phase && demod && bias && out, GANCE_E_NULL, "gance_modconv3x3_up2_fused: NULL pointer");
    GANCE_REQUIRE(dtype == GANCE_BF16 || dtype == GANCE_F16, GANCE_E_DTYPE, "gance_modconv3x3_up2_fused: dtype %d", dtype);
    GANCE_REQUIRE(batch > 0 && h_in >= 4 && w_in >= 4 && c_in % 64 == 0 && (c_out % 64 == 0 || c_out == 32), GANCE_E_SHAPE,
                  "gance_modconv3x3_up2_fused: shape B=%d H=%d W=%d Cin=%d Cout=%d (needs H >= 4, W >= 4, Cin %% 64 == 0, "
                  "Cout %% 64 == 0 or Cout == 32)", batch, h_in, w_in, c_in, c_out);
    GANCE_REQUIRE(aligned16(x) && aligned16(weight_phase) && aligned16(out), GANCE_E_ALIGN,
                  "gance_modconv3x3_up2_fused: pointers must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    // Wide layers (Cin >= 256) are bound by what the producer pulls through L2: CTA pairs + halo boxes (upconv_fused_pair.cu)
    // when a layer fills the co-resident pairs (or whenever eligible in pairing mode 2)
#ifdef GANCE_ABL_FP_CIN128
    constexpr int kPairMinCin = 128;  // A/B build
#else
    constexpr int kPairMinCin = 256;
#endif
#ifdef GANCE_ABL_FP_TALL512
    constexpr int kTallMinCin = 512;  // A/B build: the Cin = 256 layer on the 16 x 8-cell pair kernel instead
#else
    constexpr int kTallMinCin = kPairMinCin;
#endif
    if (c_in >= kTallMinCin && c_out % 64 == 0 && h_in >= 32 && w_in >= 32 && tc::conv_pair_mode() != 0 && tc::fused_pair_clusters() > 0) {
        const long long m_tiles = (long long)batch * ((2 * h_in + 27) / 28) * ((2 * w_in + 11) / 12);
        if (tc::conv_pair_mode() == 2 || ((m_tiles + 1) / 2) * (c_out / 64) >= tc::fused_pair_clusters()) {
            if (dtype == GANCE_F16)
                return tc::launch_fused_pair<__half>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                                     noise_batch_stride, noise_strength, bias, next_styles, out, s);
            return tc::launch_fused_pair<__nv_bfloat16>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                                        noise_batch_stride, noise_strength, bias, next_styles, out, s);
        }
    }
    // Wide layers too small for the 8 x 16-cell tiles (16^2 -> 32^2): the 16 x 8-cell kernel itself as CTA pairs - half of the
    // weight bytes per SM through L2 and shared memory, five pipeline stages instead of three: 102 -> 79 us. Measured on
    // the Cin = 128 layer (256^2 -> 512^2) the pair flavour LOSES (745 vs 698 us: that layer is bound by its filter step, and
    // the accumulator hand-over across the SM pair adds to it), so narrow layers pair only when forced (mode 2, tests).
    const bool wide_pair_wanted = tc::conv_pair_mode() == 2 || (tc::conv_pair_mode() == 1 && c_in >= kPairMinCin);
    if (wide_pair_wanted && c_out % 64 == 0 && tc::g_fused_wide_pair_clusters > 0) {
        const long long m_tiles = (long long)batch * ((2 * h_in + 11) / 12) * ((2 * w_in + 27) / 28);
        if (tc::conv_pair_mode() == 2 || ((m_tiles + 1) / 2) * (c_out / 64) >= tc::g_fused_wide_pair_clusters) {
            if (dtype == GANCE_F16)
                return tc::launch_fused<__half, 64, true>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                                          noise_batch_stride, noise_strength, bias, next_styles, out, s);
            return tc::launch_fused<__nv_bfloat16, 64, true>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                                             noise_batch_stride, noise_strength, bias, next_styles, out, s);
        }
    }
    if (c_out == 32) {
        if (dtype == GANCE_F16)
            return tc::launch_fused<__half, 32>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                                noise_batch_stride, noise_strength, bias, next_styles, out, s);
        return tc::launch_fused<__nv_bfloat16, 32>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                                   noise_batch_stride, noise_strength, bias, next_styles, out, s);
    }
    if (dtype == GANCE_F16)
        return tc::launch_fused<__half, 64>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                            noise_batch_stride, noise_strength, bias, next_styles, out, s);
    return tc::launch_fused<__nv_bfloat16, 64>(x, batch, h_in, w_in, c_in, c_out, weight_phase, demod, noise,
                                               noise_batch_stride, noise_strength, bias, next_styles, out, s);
}
