// Bandwidth-bound and small kernels of G_mapping / G_synthesis, plus the fp32 validation-mode
// convolution. Everything is NHWC; T is float (validation mode) or __nv_bfloat16 (fast mode).
// Reference semantics: SURVEY.md Appendix A.4 (NVlabs/stylegan2 networks_stylegan2.py,
// upfirdn_2d.py, fused_bias_act.py, tfutil.convert_images_to_uint8).
#include "common.cuh"

namespace gance {

// ------------------------------------------------------------------------------------------------
// G_mapping + truncation: 8 dense 512x512 layers over the batch, ONE LAUNCH PER LAYER so that every layer's 1 MB of
// weights is pulled through all 148 SMs instead of one SM per sample (the single-launch, CTA-per-sample version spent
// 264 us on a batch of 32 - 3 % of a step - waiting on one SM's L2 path; this one takes 8 x ~4 us).
// CTA = 8 output features x 8 samples; 1024 threads = 64 outputs x 16 slices of K, reduced through shared memory.
// W is [layer][in][out]: the 8 features of a CTA are one 32-byte sector per input row.
// The layers ping-pong between rows 0 and 1 of each sample's w+ block (no scratch argument needed); the last layer
// applies the truncation lerp and mapping_broadcast_kernel copies the result to the remaining rows.
// ------------------------------------------------------------------------------------------------
constexpr int kLatent = 512;
// A layer is latency-bound (1 MB of weights, 32 x 512 x 512 MACs): the 512-long dot product of one (sample, feature) is cut
// into 16 slices of 32 so that every thread issues all of its 64 loads at once - one round trip to L2 per layer instead of
// the 32 dependent ones of the first version (4 slices, 4 loads in flight: 14-19 us per layer).
constexpr int kMapFeatures = 8, kMapSamples = 8, kMapSlices = 16;

template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(kMapFeatures * kMapSamples * kMapSlices)
mapping_layer_kernel(const float* __restrict__ x, long long x_stride, int batch, const float* __restrict__ weight,
                     const float* __restrict__ bias, const float* __restrict__ dlatent_avg, float psi,
                     float* __restrict__ out, long long out_stride) {
    __shared__ float partial[kMapSlices][kMapFeatures * kMapSamples];
    __shared__ float scale[kMapSamples];
    const int tid = threadIdx.x;
    const int f = tid % kMapFeatures, s_local = (tid / kMapFeatures) % kMapSamples, slice = tid / (kMapFeatures * kMapSamples);
    const int j = blockIdx.x * kMapFeatures + f;
    const int s0 = blockIdx.y * kMapSamples;
    const int s = min(s0 + s_local, batch - 1);
    if (FIRST) {  // normalize_2nd_moment of z: x * rsqrt(mean(x^2) + 1e-8); one warp per sample
        const int warp = tid >> 5, lane = tid & 31;
        if (warp < kMapSamples) {
            const float* row = x + (size_t)min(s0 + warp, batch - 1) * x_stride;
            float sq = 0.f;
            for (int k = lane; k < kLatent; k += 32) sq = fmaf(row[k], row[k], sq);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
            if (lane == 0) scale[warp] = rsqrtf(sq * (1.0f / kLatent) + 1e-8f);
        }
        __syncthreads();
    }
    const float* xr = x + (size_t)s * x_stride + slice * (kLatent / kMapSlices);
    const float* w = weight + (size_t)slice * (kLatent / kMapSlices) * kLatent + j;
    constexpr int kPerSlice = kLatent / kMapSlices;  // 32, in two halves of 16 (1024 threads per CTA = 64 registers each)
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
#pragma unroll
    for (int half = 0; half < kPerSlice; half += 16) {
        float xv[16], wv[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            wv[k] = __ldg(w + (size_t)(half + k) * kLatent);
            xv[k] = __ldg(xr + half + k);
        }
#pragma unroll
        for (int k = 0; k < 16; k += 4) {
            acc0 = fmaf(xv[k], wv[k], acc0);
            acc1 = fmaf(xv[k + 1], wv[k + 1], acc1);
            acc2 = fmaf(xv[k + 2], wv[k + 2], acc2);
            acc3 = fmaf(xv[k + 3], wv[k + 3], acc3);
        }
    }
    partial[slice][tid % (kMapFeatures * kMapSamples)] = (acc0 + acc1) + (acc2 + acc3);
    __syncthreads();
    if (slice == 0 && s0 + s_local < batch) {
        const int o = tid;
        float total = 0.f;
#pragma unroll
        for (int q = 0; q < kMapSlices; q += 4)
            total += (partial[q][o] + partial[q + 1][o]) + (partial[q + 2][o] + partial[q + 3][o]);
        if (FIRST) total *= scale[s_local];  // the normalisation is a per-sample scalar: apply it to the dot product
        float value = bias_lrelu(total + __ldg(bias + j));
        if (LAST && psi != 1.0f) {  // G_main 'Truncation': lerp(dlatent_avg, w, psi)
            const float avg = __ldg(dlatent_avg + j);
            value = avg + (value - avg) * psi;
        }
        out[(size_t)s * out_stride + j] = value;
    }
}

__global__ void mapping_broadcast_kernel(float* __restrict__ wplus, int batch, int num_ws, int source_row) {
    const size_t total = (size_t)batch * num_ws * kLatent;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % kLatent);
        const int row = (int)((i / kLatent) % num_ws);
        const size_t b = i / ((size_t)kLatent * num_ws);
        if (row != source_row) wplus[i] = wplus[(b * num_ws + source_row) * kLatent + k];
    }
}

// ------------------------------------------------------------------------------------------------
// Style affine + demodulation coefficients. One CTA per sample.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) style_demod_kernel(const float* __restrict__ wplus, int num_ws,
                                                          int w_index,
                                                          const float* __restrict__ mod_weight,
                                                          const float* __restrict__ mod_bias,
                                                          const float* __restrict__ wsq, int c_in,
                                                          int c_out, float* __restrict__ styles,
                                                          float* __restrict__ demod) {
    __shared__ float w[kLatent];
    __shared__ float s2[kLatent];
    const int b = blockIdx.x;
    const float* src = wplus + ((size_t)b * num_ws + w_index) * kLatent;
    for (int k = threadIdx.x; k < kLatent; k += blockDim.x) w[k] = src[k];
    __syncthreads();
    for (int ci = threadIdx.x; ci < c_in; ci += blockDim.x) {
        float acc0 = 0.f, acc1 = 0.f;
        for (int k = 0; k < kLatent; k += 2) {
            acc0 = fmaf(w[k], __ldg(mod_weight + (size_t)k * c_in + ci), acc0);
            acc1 = fmaf(w[k + 1], __ldg(mod_weight + (size_t)(k + 1) * c_in + ci), acc1);
        }
        float s = acc0 + acc1 + mod_bias[ci] + 1.0f;
        styles[(size_t)b * c_in + ci] = s;
        s2[ci] = s * s;
    }
    if (demod == nullptr) return;
    __syncthreads();
    for (int co = threadIdx.x; co < c_out; co += blockDim.x) {
        float acc0 = 0.f, acc1 = 0.f;
        for (int ci = 0; ci < c_in; ci += 2) {
            acc0 = fmaf(s2[ci], __ldg(wsq + (size_t)ci * c_out + co), acc0);
            acc1 = fmaf(s2[ci + 1], __ldg(wsq + (size_t)(ci + 1) * c_out + co), acc1);
        }
        demod[(size_t)b * c_out + co] = rsqrtf(acc0 + acc1 + 1e-8f);
    }
}

// ------------------------------------------------------------------------------------------------
// Styles and demodulation coefficients of EVERY layer of a generator in two launches.
// The per-layer affine weights are concatenated column-wise into mod_weight_cat [512, total_cols];
// a chunk = 32 consecutive columns of one layer. CTA = 32 columns x 16 slices of the 512-long (or
// Cin-long) reduction, for 8 samples at a time (the w vectors / squared styles sit in shared memory).
// Outputs are layer-major so each layer's [B, C] block is contiguous: block of layer L starts at
// B * col0(L).
// ------------------------------------------------------------------------------------------------
constexpr int kStyleSamples = 8;

constexpr int kStyleSlices = 16;  // slices of the reduction: every thread has its 32 (or Cin / 16) weight loads in flight at once

__global__ void __launch_bounds__(32 * kStyleSlices) styles_all_kernel(const float* __restrict__ wplus, int batch, int num_ws,
                                                                       const float* __restrict__ mod_weight_cat, int total_cols,
                                                                       const float* __restrict__ mod_bias_cat,
                                                                       const int4* __restrict__ chunks,
                                                                       float* __restrict__ styles_out) {
    __shared__ float w[kStyleSamples][kLatent];
    __shared__ float part[kStyleSlices][kStyleSamples][32];
    const int4 chunk = chunks[blockIdx.x];  // x: w_index, y: first column of the layer, z: Cin, w: first column of the chunk inside the layer
    const int b0 = blockIdx.y * kStyleSamples;
    for (int i = threadIdx.x; i < kStyleSamples * kLatent; i += blockDim.x) {
        const int sidx = i / kLatent, k = i % kLatent;
        w[sidx][k] = b0 + sidx < batch ? wplus[((size_t)(b0 + sidx) * num_ws + chunk.x) * kLatent + k] : 0.f;
    }
    const int col = threadIdx.x & 31, slice = threadIdx.x >> 5;
    const int gcol = chunk.y + chunk.w + col;
    constexpr int kPerSlice = kLatent / kStyleSlices;  // 32
    const float* m = mod_weight_cat + (size_t)(slice * kPerSlice) * total_cols + gcol;
    float mv[kPerSlice];
#pragma unroll
    for (int k = 0; k < kPerSlice; ++k) mv[k] = __ldg(m + (size_t)k * total_cols);
    __syncthreads();
    float acc[kStyleSamples];
#pragma unroll
    for (int i = 0; i < kStyleSamples; ++i) acc[i] = 0.f;
#pragma unroll
    for (int k = 0; k < kPerSlice; ++k) {
#pragma unroll
        for (int i = 0; i < kStyleSamples; ++i) acc[i] = fmaf(w[i][slice * kPerSlice + k], mv[k], acc[i]);
    }
#pragma unroll
    for (int i = 0; i < kStyleSamples; ++i) part[slice][i][col] = acc[i];
    __syncthreads();
    for (int i = threadIdx.x; i < kStyleSamples * 32; i += blockDim.x) {
        const int sidx = i >> 5, c = i & 31;
        if (b0 + sidx >= batch) continue;
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < kStyleSlices; q += 4)
            v += (part[q][sidx][c] + part[q + 1][sidx][c]) + (part[q + 2][sidx][c] + part[q + 3][sidx][c]);
        styles_out[(size_t)batch * chunk.y + (size_t)(b0 + sidx) * chunk.z + chunk.w + c] =
            v + mod_bias_cat[chunk.y + chunk.w + c] + 1.0f;
    }
}

struct DemodChunk {
    int wsq_offset;   // float offset of this layer's [Cin, Cout] sum-of-squared-weights matrix in wsq_cat
    int c_in, c_out;
    int style_col0;   // first column of the layer in the concatenated style space
    int co0;          // first output channel of this chunk
    int demod_col0;   // first column of the layer in the concatenated demod space
};

__global__ void __launch_bounds__(32 * kStyleSlices) demod_all_kernel(const float* __restrict__ styles_all, int batch,
                                                                      const float* __restrict__ wsq_cat,
                                                                      const DemodChunk* __restrict__ chunks,
                                                                      float* __restrict__ demod_out) {
    __shared__ float s2[kStyleSamples][kLatent];
    __shared__ float part[kStyleSlices][kStyleSamples][32];
    const DemodChunk chunk = chunks[blockIdx.x];
    const int b0 = blockIdx.y * kStyleSamples;
    const int col = threadIdx.x & 31, slice = threadIdx.x >> 5;
    const int per_slice = (chunk.c_in + kStyleSlices - 1) / kStyleSlices;  // 1 ... 32 rows per slice
    const int rows = min(per_slice, chunk.c_in - slice * per_slice);       // <= 0 for the slices past a short Cin
    const float* m = wsq_cat + chunk.wsq_offset + (size_t)(slice * per_slice) * chunk.c_out + chunk.co0 + col;
    constexpr int kMaxPerSlice = kLatent / kStyleSlices;
    float mv[kMaxPerSlice];
#pragma unroll
    for (int k = 0; k < kMaxPerSlice; ++k) mv[k] = k < rows ? __ldg(m + (size_t)k * chunk.c_out) : 0.f;
    for (int i = threadIdx.x; i < kStyleSamples * chunk.c_in; i += blockDim.x) {
        const int sidx = i / chunk.c_in, ci = i % chunk.c_in;
        float v = 0.f;
        if (b0 + sidx < batch) v = styles_all[(size_t)batch * chunk.style_col0 + (size_t)(b0 + sidx) * chunk.c_in + ci];
        s2[sidx][ci] = v * v;
    }
    __syncthreads();
    float acc[kStyleSamples];
#pragma unroll
    for (int i = 0; i < kStyleSamples; ++i) acc[i] = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxPerSlice; ++k) {
        if (k < rows) {
#pragma unroll
            for (int i = 0; i < kStyleSamples; ++i) acc[i] = fmaf(s2[i][slice * per_slice + k], mv[k], acc[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < kStyleSamples; ++i) part[slice][i][col] = acc[i];
    __syncthreads();
    for (int i = threadIdx.x; i < kStyleSamples * 32; i += blockDim.x) {
        const int sidx = i >> 5, c = i & 31;
        if (b0 + sidx >= batch) continue;
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < kStyleSlices; q += 4)
            v += (part[q][sidx][c] + part[q + 1][sidx][c]) + (part[q + 2][sidx][c] + part[q + 3][sidx][c]);
        demod_out[(size_t)batch * chunk.demod_col0 + (size_t)(b0 + sidx) * chunk.c_out + chunk.co0 + c] =
            rsqrtf(v + 1e-8f);
    }
}

// ------------------------------------------------------------------------------------------------
// const * style
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void const_input_kernel(const float* __restrict__ konst, const float* __restrict__ styles,
                                   int batch, int channels, T* __restrict__ out) {
    constexpr int V = Vec<T>::N;
    const int vec_per_pixel = channels / V;
    const size_t total = (size_t)batch * 16 * vec_per_pixel;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        int cv = (int)(i % vec_per_pixel);
        size_t rest = i / vec_per_pixel;
        int p = (int)(rest % 16);
        int b = (int)(rest / 16);
        float f[V];
#pragma unroll
        for (int k = 0; k < V; ++k)
            f[k] = konst[p * channels + cv * V + k] * styles[(size_t)b * channels + cv * V + k];
        Vec<T> v;
        v.from_float(f);
        v.store(out + ((size_t)b * 16 + p) * channels + cv * V);
    }
}

// ------------------------------------------------------------------------------------------------
// 2x FIR upsample with one sample of apron: u [B,2H+2,2W+2,C].
// One thread = one 16-byte channel vector of a 2x2 output quad, built from the 2x2 input
// neighbourhood x[a-1..a, b-1..b]; per axis even taps (.75,.25), odd taps (.25,.75).
// Reads are coalesced along C and each input vector is re-read by 4 quads out of L1/L2.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) upfir2x_pad_kernel(const T* __restrict__ x, int batch, int h, int w,
                                                          int channels, const float* __restrict__ styles,
                                                          T* __restrict__ u) {
    constexpr int V = Vec<T>::N;
    const int vec_per_pixel = channels / V;
    const int qh = h + 1, qw = w + 1;
    const int uh = 2 * h + 2, uw = 2 * w + 2;
    const size_t total = (size_t)batch * qh * qw * vec_per_pixel;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        int cv = (int)(i % vec_per_pixel);
        size_t rest = i / vec_per_pixel;
        int bq = (int)(rest % qw);
        rest /= qw;
        int aq = (int)(rest % qh);
        int b = (int)(rest / qh);
        float in[2][2][V];
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                int yy = aq - 1 + dy, xx = bq - 1 + dx;
                Vec<T> v;
                if (yy >= 0 && yy < h && xx >= 0 && xx < w)
                    v.load(x + (((size_t)b * h + yy) * w + xx) * channels + cv * V);
                else
                    v.zero();
                v.to_float(in[dy][dx]);
            }
        float s[V];
#pragma unroll
        for (int k = 0; k < V; ++k) s[k] = styles ? styles[(size_t)b * channels + cv * V + k] : 1.0f;
#pragma unroll
        for (int py = 0; py < 2; ++py) {
            const float wy0 = py == 0 ? 0.75f : 0.25f, wy1 = 1.0f - wy0;
#pragma unroll
            for (int px = 0; px < 2; ++px) {
                const float wx0 = px == 0 ? 0.75f : 0.25f, wx1 = 1.0f - wx0;
                float out[V];
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    float top = in[0][0][k] * wx0 + in[0][1][k] * wx1;
                    float bot = in[1][0][k] * wx0 + in[1][1][k] * wx1;
                    out[k] = (top * wy0 + bot * wy1) * s[k];
                }
                Vec<T> v;
                v.from_float(out);
                v.store(u + (((size_t)b * uh + 2 * aq + py) * uw + 2 * bq + px) * channels + cv * V);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// ToRGB (1x1 modulated conv, no demod) + bias + skip upsample-add (+ uint8 conversion).
// A group of LANES (<= 8) lanes owns kTorgbPixels pixels per pass and splits their channels in 16-byte vectors (a group
// reads 128 consecutive bytes). wmod[c] = styles[b,c] * weight[c,0..2] sits in shared memory (one sample per CTA row of
// the grid).
// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// x[b,h,w,c] * styles[b,c] -> out (the pre-scaling of a modulated convolution's input when the
// producing kernel could not apply it in its epilogue). 16-byte vectors, grid-stride.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) scale_channels_kernel(const T* __restrict__ x, size_t pixels_per_sample,
                                                             int channels, size_t total_vecs,
                                                             const float* __restrict__ styles, T* __restrict__ out) {
    constexpr int V = Vec<T>::N;
    const int vec_per_pixel = channels / V;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total_vecs;
         i += (size_t)gridDim.x * blockDim.x) {
        const int cv = (int)(i % vec_per_pixel);
        const size_t b = i / ((size_t)vec_per_pixel * pixels_per_sample);
        Vec<T> v;
        v.load(x + i * V);
        float f[V];
        v.to_float(f);
        const float* s = styles + b * channels + cv * V;
#pragma unroll
        for (int k = 0; k < V; ++k) f[k] *= __ldg(s + k);
        v.from_float(f);
        v.store(out + i * V);
    }
}

// ToRGB launch shape (ncu of the first version, 64^2 x 512 channels: 69 % issue-slot use against 30 % DRAM - the layer was
// bound by its own bookkeeping: a whole warp per pixel = 15 shuffle + add pairs per colour and pixel, three scalar shared
// loads per channel, one divergent tail per pixel). Now a group of at most 8 lanes owns kTorgbPixels pixels per pass:
// 3 shuffle steps, one float4 weight read per channel shared by the group's pixels, and the lanes of a group finish one
// pixel each side by side; the four groups of a warp read four 128-byte runs.
constexpr int kTorgbPixels = 4;
constexpr int kTorgbMaxLanes = 8;

template <typename T, int LANES>
__global__ void __launch_bounds__(256) torgb_skip_kernel(const T* __restrict__ x, int h, int w, int channels,
                                                         const float* __restrict__ styles,
                                                         const float* __restrict__ weight,
                                                         const float* __restrict__ bias,
                                                         const float* __restrict__ y_prev,
                                                         float* __restrict__ y_out,
                                                         uint8_t* __restrict__ frames, int* nonfinite) {
    constexpr int V = Vec<T>::N;
    // [V][channels / V] float4 (r, g, b, -): channel c = cv * V + k sits at slot k * (channels / V) + cv, so the lanes of
    // a group (consecutive cv, same k) read consecutive 16-byte words and the groups of a warp broadcast
    extern __shared__ float4 wmod[];
    const int b = blockIdx.y;
    const int vec_per_pixel = channels / V;
    for (int c = threadIdx.x; c < channels; c += blockDim.x) {
        const float s = styles[(size_t)b * channels + c];
        wmod[(c % V) * vec_per_pixel + c / V] = make_float4(s * weight[3 * c], s * weight[3 * c + 1], s * weight[3 * c + 2], 0.f);
    }
    __syncthreads();
    const int pixels = h * w;
    const int group = threadIdx.x / LANES, lane = threadIdx.x % LANES;
    const int groups_per_block = blockDim.x / LANES;
    // the loop bound is CTA-uniform so the group shuffles below always see a full warp
    for (int base = blockIdx.x * groups_per_block * kTorgbPixels; base < pixels; base += gridDim.x * groups_per_block * kTorgbPixels) {
        const int p0 = base + group * kTorgbPixels;
        const T* px[kTorgbPixels];
#pragma unroll
        for (int i = 0; i < kTorgbPixels; ++i) px[i] = x + ((size_t)b * pixels + (p0 + i < pixels ? p0 + i : 0)) * channels;
        float acc[kTorgbPixels][3];
#pragma unroll
        for (int i = 0; i < kTorgbPixels; ++i) acc[i][0] = acc[i][1] = acc[i][2] = 0.f;
#pragma unroll 2
        for (int cv = lane; cv < vec_per_pixel; cv += LANES) {
            Vec<T> v[kTorgbPixels];
#pragma unroll
            for (int i = 0; i < kTorgbPixels; ++i) v[i].load(px[i] + cv * V);
            float f[kTorgbPixels][V];
#pragma unroll
            for (int i = 0; i < kTorgbPixels; ++i) v[i].to_float(f[i]);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const float4 wv = wmod[k * vec_per_pixel + cv];
#pragma unroll
                for (int i = 0; i < kTorgbPixels; ++i) {
                    acc[i][0] = fmaf(f[i][k], wv.x, acc[i][0]);
                    acc[i][1] = fmaf(f[i][k], wv.y, acc[i][1]);
                    acc[i][2] = fmaf(f[i][k], wv.z, acc[i][2]);
                }
            }
        }
#pragma unroll
        for (int o = LANES / 2; o > 0; o >>= 1) {
#pragma unroll
            for (int i = 0; i < kTorgbPixels; ++i) {
                acc[i][0] += __shfl_xor_sync(0xffffffffu, acc[i][0], o);
                acc[i][1] += __shfl_xor_sync(0xffffffffu, acc[i][1], o);
                acc[i][2] += __shfl_xor_sync(0xffffffffu, acc[i][2], o);
            }
        }
        // every lane of the group holds all sums: lane j finishes pixels j, j + LANES, ...
#pragma unroll
        for (int t = 0; t < (kTorgbPixels + LANES - 1) / LANES; ++t) {
            const int i = t * LANES + lane;
            float r = 0.f, g = 0.f, bl = 0.f;
#pragma unroll
            for (int j = 0; j < kTorgbPixels; ++j) {
                if (j == i) { r = acc[j][0]; g = acc[j][1]; bl = acc[j][2]; }
            }
            const int p = p0 + i;
            if (i < kTorgbPixels && p < pixels) {
                r += bias[0]; g += bias[1]; bl += bias[2];
                const int hh = p / w, ww = p % w;
                if (y_prev != nullptr) {
                    float4 up = upsample_tap(y_prev, b, h >> 1, w >> 1, hh, ww);
                    r += up.x; g += up.y; bl += up.z;
                }
                if (y_out != nullptr)
                    reinterpret_cast<float4*>(y_out)[(size_t)b * pixels + p] = make_float4(r, g, bl, 0.f);
                if (frames != nullptr) {
                    uint8_t* f = frames + ((size_t)b * pixels + p) * 3;
                    f[0] = to_u8(r); f[1] = to_u8(g); f[2] = to_u8(bl);
                    flag_nonfinite(nonfinite, r, g, bl);
                }
            }
        }
    }
}

__global__ void to_uint8_kernel(const float4* __restrict__ y, size_t pixels, uint8_t* __restrict__ frames, int* nonfinite) {
    for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < pixels;
         p += (size_t)gridDim.x * blockDim.x) {
        float4 v = y[p];
        frames[p * 3 + 0] = to_u8(v.x);
        frames[p * 3 + 1] = to_u8(v.y);
        frames[p * 3 + 2] = to_u8(v.z);
        flag_nonfinite(nonfinite, v.x, v.y, v.z);
    }
}

// ------------------------------------------------------------------------------------------------
// Fresh N(0,1) noise for every noise input of a batch in ONE launch (the vector path: Gs.run leaves
// randomize_noise=True, network_functions.py:151-158, so tf.random_normal draws [B,1,H,W] per layer).
// Counter-based: element group g (4 consecutive floats) = BoxMuller(Philox4x32-10(counter = g + offset,
// key = seed)), so the stream is reproducible from (seed, offset) and independent of the launch shape.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
#pragma unroll
    for (int round = 0; round < 10; ++round) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}

__device__ __forceinline__ float2 box_muller(uint32_t a, uint32_t b) {
    const float u1 = ((float)a + 1.0f) * 2.3283064365386963e-10f;  // (0, 1]
    const float u2 = (float)b * 2.3283064365386963e-10f;           // [0, 1)
    const float radius = sqrtf(-2.0f * __logf(u1));
    float sn, cs;
    __sincosf(6.283185307179586f * u2, &sn, &cs);
    return make_float2(radius * cs, radius * sn);
}

__global__ void __launch_bounds__(256) normal_noise_kernel(float4* __restrict__ out, size_t groups, uint2 key,
                                                           unsigned long long offset,
                                                           const unsigned long long* __restrict__ counter) {
    if (counter != nullptr) offset += *counter;  // device-resident stream position (CUDA-graph replays must not repeat noise)
    for (size_t g = blockIdx.x * (size_t)blockDim.x + threadIdx.x; g < groups; g += (size_t)gridDim.x * blockDim.x) {
        const unsigned long long c = g + offset;
        const uint4 r = philox4x32_10(make_uint4((uint32_t)c, (uint32_t)(c >> 32), 0u, 0u), key);
        const float2 a = box_muller(r.x, r.y), b = box_muller(r.z, r.w);
        out[g] = make_float4(a.x, a.y, b.x, b.y);
    }
}

__global__ void advance_counter_kernel(unsigned long long* counter, unsigned long long by) { *counter += by; }

// ------------------------------------------------------------------------------------------------
// fp32 validation-mode 3x3 convolution on CUDA cores (the numerical yard-stick for the tcgen05
// path). CTA = 8x8 output pixels x 64 output channels; thread = 4 pixels x 4 channels;
// input channels are staged through shared memory 8 at a time. Weights are HWIO (TF layout).
// ------------------------------------------------------------------------------------------------
constexpr int kF32TileP = 8;    // 8x8 pixels
constexpr int kF32TileC = 64;   // output channels per CTA
constexpr int kF32ChunkK = 8;   // input channels per smem stage

__global__ void __launch_bounds__(256) modconv3x3_f32_kernel(
    const float* __restrict__ x, int batch, int h_out, int w_out, int c_in, int c_out, int pad,
    const float* __restrict__ weight, const float* __restrict__ demod, const float* __restrict__ noise,
    long long noise_batch_stride, float noise_strength, const float* __restrict__ bias,
    const float* __restrict__ next_styles, float* __restrict__ out) {
    __shared__ float xs[kF32TileP + 2][kF32TileP + 2][kF32ChunkK];
    __shared__ __align__(16) float ws[9][kF32ChunkK][kF32TileC];
    const int h_in = h_out + 2 - 2 * pad, w_in = w_out + 2 - 2 * pad;
    const int tiles_w = (w_out + kF32TileP - 1) / kF32TileP;
    const int tile_h0 = (blockIdx.x / tiles_w) * kF32TileP, tile_w0 = (blockIdx.x % tiles_w) * kF32TileP;
    const int co0 = blockIdx.y * kF32TileC;
    const int b = blockIdx.z;
    const int cg = threadIdx.x % 16;  // 4 output channels: co0 + cg*4 ..
    const int pg = threadIdx.x / 16;  // 4 pixels: row (pg/2), cols (pg%2)*4 ..
    const int prow = pg / 2, pcol0 = (pg % 2) * 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int c0 = 0; c0 < c_in; c0 += kF32ChunkK) {
        for (int i = threadIdx.x; i < (kF32TileP + 2) * (kF32TileP + 2) * kF32ChunkK; i += blockDim.x) {
            int k = i % kF32ChunkK;
            int xx = (i / kF32ChunkK) % (kF32TileP + 2);
            int yy = i / (kF32ChunkK * (kF32TileP + 2));
            int gy = tile_h0 + yy - pad, gx = tile_w0 + xx - pad;
            float v = 0.f;
            if (gy >= 0 && gy < h_in && gx >= 0 && gx < w_in && c0 + k < c_in)
                v = x[(((size_t)b * h_in + gy) * w_in + gx) * c_in + c0 + k];
            xs[yy][xx][k] = v;
        }
        for (int i = threadIdx.x; i < 9 * kF32ChunkK * kF32TileC; i += blockDim.x) {
            int co = i % kF32TileC;
            int k = (i / kF32TileC) % kF32ChunkK;
            int tap = i / (kF32TileC * kF32ChunkK);
            float v = 0.f;
            if (c0 + k < c_in && co0 + co < c_out)
                v = weight[((size_t)tap * c_in + c0 + k) * c_out + co0 + co];
            ws[tap][k][co] = v;
        }
        __syncthreads();
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int ky = tap / 3, kx = tap % 3;
#pragma unroll
            for (int k = 0; k < kF32ChunkK; ++k) {
                const float4 wv = *reinterpret_cast<const float4*>(&ws[tap][k][cg * 4]);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float xv = xs[prow + ky][pcol0 + i + kx][k];
                    acc[i][0] = fmaf(xv, wv.x, acc[i][0]);
                    acc[i][1] = fmaf(xv, wv.y, acc[i][1]);
                    acc[i][2] = fmaf(xv, wv.z, acc[i][2]);
                    acc[i][3] = fmaf(xv, wv.w, acc[i][3]);
                }
            }
        }
        __syncthreads();
    }
    const int hh = tile_h0 + prow;
    if (hh >= h_out) return;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ww = tile_w0 + pcol0 + i;
        if (ww >= w_out) continue;
        float nz = 0.f;
        if (noise != nullptr)
            nz = noise[(size_t)b * noise_batch_stride + (size_t)hh * w_out + ww] * noise_strength;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int co = co0 + cg * 4 + j;
            if (co >= c_out) continue;
            float v = acc[i][j] * demod[(size_t)b * c_out + co] + nz + bias[co];
            v = bias_lrelu(v);
            if (next_styles != nullptr) v *= next_styles[(size_t)b * c_out + co];
            out[(((size_t)b * h_out + hh) * w_out + ww) * c_out + co] = v;
        }
    }
}

int launch_modconv3x3_f32(const float* x, int batch, int h_out, int w_out, int c_in, int c_out, int pad,
                          const float* weight, const float* demod, const float* noise,
                          long long noise_batch_stride, float noise_strength, const float* bias,
                          const float* next_styles, float* out, cudaStream_t stream) {
    dim3 grid(((h_out + kF32TileP - 1) / kF32TileP) * ((w_out + kF32TileP - 1) / kF32TileP),
              (c_out + kF32TileC - 1) / kF32TileC, batch);
    modconv3x3_f32_kernel<<<grid, 256, 0, stream>>>(x, batch, h_out, w_out, c_in, c_out, pad, weight, demod,
                                                    noise, noise_batch_stride, noise_strength, bias,
                                                    next_styles, out);
    return check_launch("modconv3x3_f32_kernel");
}

static int grid_for(size_t work_items, int block) {
    size_t blocks = (work_items + block - 1) / block;
    const size_t cap = 148u * 16u;  // 148 SMs x enough resident CTAs to cover latency
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace gance

using namespace gance;

template <typename T>
static int launch_torgb(const void* x, int batch, int h, int w, int channels, const float* styles,
                        const float* weight, const float* bias, const float* y_prev, float* y_out,
                        uint8_t* frames, cudaStream_t s) {
    const int vec_per_pixel = channels / Vec<T>::N;
    const size_t smem = (size_t)channels * sizeof(float4);
    int lanes = 1;
    while (lanes < kTorgbMaxLanes && lanes < vec_per_pixel) lanes <<= 1;
    const int groups = 256 / lanes;
    int gx = (h * w + groups * kTorgbPixels - 1) / (groups * kTorgbPixels);
    const int cap = (148 * 8 + batch - 1) / batch;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    dim3 grid(gx, batch);
#define GANCE_TORGB(L)                                                                                   \
    torgb_skip_kernel<T, L><<<grid, 256, smem, s>>>((const T*)x, h, w, channels, styles, weight, bias, \
                                                    y_prev, y_out, frames, nonfinite_flag())
    switch (lanes) {
        case 1: GANCE_TORGB(1); break;
        case 2: GANCE_TORGB(2); break;
        case 4: GANCE_TORGB(4); break;
        default: GANCE_TORGB(8); break;
    }
#undef GANCE_TORGB
    return check_launch("torgb_skip_kernel");
}

extern "C" {

int gance_mapping(const float* z, int batch, const float* weights, const float* biases,
                  const float* dlatent_avg, float truncation_psi, int num_ws, float* wplus, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(z && weights && biases && dlatent_avg && wplus, GANCE_E_NULL, "gance_mapping: NULL pointer");
    GANCE_REQUIRE(batch > 0 && num_ws > 0, GANCE_E_SHAPE, "gance_mapping: batch=%d num_ws=%d", batch, num_ws);
    cudaStream_t s = (cudaStream_t)stream;
    const dim3 grid(kLatent / kMapFeatures, (batch + kMapSamples - 1) / kMapSamples);
    const int threads = kMapFeatures * kMapSamples * kMapSlices;
    const long long row_stride = (long long)num_ws * kLatent;
    float* row0 = wplus;
    float* row1 = num_ws > 1 ? wplus + kLatent : nullptr;
    GANCE_REQUIRE(num_ws >= 2, GANCE_E_SHAPE, "gance_mapping: num_ws=%d (the layers ping-pong between w+ rows 0 and 1)", num_ws);
    for (int layer = 0; layer < 8; ++layer) {
        const float* w = weights + (size_t)layer * kLatent * kLatent;
        const float* bl = biases + (size_t)layer * kLatent;
        float* dst = (layer & 1) ? row1 : row0;
        const float* src = (layer & 1) ? row0 : row1;
        if (layer == 0)
            mapping_layer_kernel<true, false><<<grid, threads, 0, s>>>(z, kLatent, batch, w, bl, dlatent_avg, truncation_psi, dst, row_stride);
        else if (layer == 7)
            mapping_layer_kernel<false, true><<<grid, threads, 0, s>>>(src, row_stride, batch, w, bl, dlatent_avg, truncation_psi, dst, row_stride);
        else
            mapping_layer_kernel<false, false><<<grid, threads, 0, s>>>(src, row_stride, batch, w, bl, dlatent_avg, truncation_psi, dst, row_stride);
        int rc = check_launch("mapping_layer_kernel");
        if (rc != GANCE_OK) return rc;
    }
    mapping_broadcast_kernel<<<grid_for((size_t)batch * num_ws * kLatent, 256), 256, 0, s>>>(wplus, batch, num_ws, 1);
    return check_launch("mapping_broadcast_kernel");
}

int gance_style_demod(const float* wplus, int batch, int num_ws, int w_index, const float* mod_weight,
                      const float* mod_bias, const float* wsq, int c_in, int c_out, float* styles,
                      float* demod, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(wplus && mod_weight && mod_bias && styles, GANCE_E_NULL, "gance_style_demod: NULL pointer");
    GANCE_REQUIRE((wsq == nullptr) == (demod == nullptr), GANCE_E_NULL,
                  "gance_style_demod: wsq and demod must both be given or both be NULL");
    GANCE_REQUIRE(batch > 0 && w_index >= 0 && w_index < num_ws, GANCE_E_SHAPE,
                  "gance_style_demod: w_index %d outside [0,%d)", w_index, num_ws);
    GANCE_REQUIRE(c_in > 0 && c_in <= kLatent && c_in % 2 == 0 && c_out > 0, GANCE_E_SHAPE,
                  "gance_style_demod: c_in=%d c_out=%d unsupported", c_in, c_out);
    style_demod_kernel<<<batch, 256, 0, (cudaStream_t)stream>>>(wplus, num_ws, w_index, mod_weight, mod_bias, wsq,
                                                                c_in, c_out, styles, demod);
    return check_launch("style_demod_kernel");
}

int gance_styles_all(const float* wplus, int batch, int num_ws, const float* mod_weight_cat, int total_cols,
                     const float* mod_bias_cat, const int* style_chunks, int n_style_chunks, float* styles_out,
                     const float* wsq_cat, const int* demod_chunks, int n_demod_chunks, float* demod_out,
                     void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(wplus && mod_weight_cat && mod_bias_cat && style_chunks && styles_out, GANCE_E_NULL,
                  "gance_styles_all: NULL pointer");
    GANCE_REQUIRE(batch > 0 && num_ws > 0 && total_cols > 0 && n_style_chunks > 0 && n_demod_chunks >= 0, GANCE_E_SHAPE,
                  "gance_styles_all: batch=%d num_ws=%d total_cols=%d chunks=%d/%d", batch, num_ws, total_cols,
                  n_style_chunks, n_demod_chunks);
    GANCE_REQUIRE(aligned16(style_chunks), GANCE_E_ALIGN, "gance_styles_all: style_chunks must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    const int sample_groups = (batch + kStyleSamples - 1) / kStyleSamples;
    styles_all_kernel<<<dim3(n_style_chunks, sample_groups), 32 * kStyleSlices, 0, s>>>(
        wplus, batch, num_ws, mod_weight_cat, total_cols, mod_bias_cat, (const int4*)style_chunks, styles_out);
    int rc = check_launch("styles_all_kernel");
    if (rc != GANCE_OK || n_demod_chunks == 0) return rc;
    GANCE_REQUIRE(wsq_cat && demod_chunks && demod_out, GANCE_E_NULL, "gance_styles_all: NULL demod pointer");
    demod_all_kernel<<<dim3(n_demod_chunks, sample_groups), 32 * kStyleSlices, 0, s>>>(styles_out, batch, wsq_cat,
                                                                        (const DemodChunk*)demod_chunks, demod_out);
    return check_launch("demod_all_kernel");
}

int gance_const_input(const float* konst, const float* styles, int batch, int channels, int dtype, void* out,
                      void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(konst && styles && out, GANCE_E_NULL, "gance_const_input: NULL pointer");
    GANCE_REQUIRE(batch > 0 && channels > 0 && channels % 8 == 0, GANCE_E_SHAPE, "gance_const_input: channels=%d",
                  channels);
    GANCE_REQUIRE(aligned16(out), GANCE_E_ALIGN, "gance_const_input: out not 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == GANCE_F32) {
        int grid = grid_for((size_t)batch * 16 * channels / 4, 256);
        const_input_kernel<float><<<grid, 256, 0, s>>>(konst, styles, batch, channels, (float*)out);
    } else if (dtype == GANCE_BF16) {
        int grid = grid_for((size_t)batch * 16 * channels / 8, 256);
        const_input_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(konst, styles, batch, channels, (__nv_bfloat16*)out);
    } else if (dtype == GANCE_F16) {
        int grid = grid_for((size_t)batch * 16 * channels / 8, 256);
        const_input_kernel<__half><<<grid, 256, 0, s>>>(konst, styles, batch, channels, (__half*)out);
    } else {
        return fail(GANCE_E_DTYPE, "gance_const_input: dtype %d", dtype);
    }
    return check_launch("const_input_kernel");
}

int gance_scale_channels(const void* x, int dtype, int batch, int h, int w, int channels, const float* styles,
                         void* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && styles && out, GANCE_E_NULL, "gance_scale_channels: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h > 0 && w > 0 && channels > 0 && channels % 8 == 0, GANCE_E_SHAPE,
                  "gance_scale_channels: shape [%d,%d,%d,%d]", batch, h, w, channels);
    GANCE_REQUIRE(aligned16(x) && aligned16(out), GANCE_E_ALIGN, "gance_scale_channels: pointers must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t pixels = (size_t)h * w;
    if (dtype == GANCE_F32) {
        const size_t vecs = (size_t)batch * pixels * channels / 4;
        scale_channels_kernel<float><<<grid_for(vecs, 256), 256, 0, s>>>((const float*)x, pixels, channels, vecs, styles,
                                                                          (float*)out);
    } else if (dtype == GANCE_BF16) {
        const size_t vecs = (size_t)batch * pixels * channels / 8;
        scale_channels_kernel<__nv_bfloat16><<<grid_for(vecs, 256), 256, 0, s>>>((const __nv_bfloat16*)x, pixels, channels,
                                                                                  vecs, styles, (__nv_bfloat16*)out);
    } else if (dtype == GANCE_F16) {
        const size_t vecs = (size_t)batch * pixels * channels / 8;
        scale_channels_kernel<__half><<<grid_for(vecs, 256), 256, 0, s>>>((const __half*)x, pixels, channels, vecs, styles,
                                                                           (__half*)out);
    } else {
        return fail(GANCE_E_DTYPE, "gance_scale_channels: dtype %d", dtype);
    }
    return check_launch("scale_channels_kernel");
}

int gance_upfir2x_pad(const void* x, int dtype, int batch, int h, int w, int channels, const float* styles,
                      void* u, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && u, GANCE_E_NULL, "gance_upfir2x_pad: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h > 0 && w > 0 && channels > 0 && channels % 8 == 0, GANCE_E_SHAPE,
                  "gance_upfir2x_pad: shape [%d,%d,%d,%d]", batch, h, w, channels);
    GANCE_REQUIRE(aligned16(x) && aligned16(u), GANCE_E_ALIGN, "gance_upfir2x_pad: pointers must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == GANCE_F32) {
        int grid = grid_for((size_t)batch * (h + 1) * (w + 1) * channels / 4, 256);
        upfir2x_pad_kernel<float><<<grid, 256, 0, s>>>((const float*)x, batch, h, w, channels, styles, (float*)u);
    } else if (dtype == GANCE_BF16) {
        int grid = grid_for((size_t)batch * (h + 1) * (w + 1) * channels / 8, 256);
        upfir2x_pad_kernel<__nv_bfloat16>
            <<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, batch, h, w, channels, styles, (__nv_bfloat16*)u);
    } else if (dtype == GANCE_F16) {
        int grid = grid_for((size_t)batch * (h + 1) * (w + 1) * channels / 8, 256);
        upfir2x_pad_kernel<__half><<<grid, 256, 0, s>>>((const __half*)x, batch, h, w, channels, styles, (__half*)u);
    } else {
        return fail(GANCE_E_DTYPE, "gance_upfir2x_pad: dtype %d", dtype);
    }
    return check_launch("upfir2x_pad_kernel");
}

int gance_torgb_skip(const void* x, int dtype, int batch, int h, int w, int channels, const float* styles,
                     const float* weight, const float* bias, const float* y_prev, float* y_out, uint8_t* frames,
                     void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && styles && weight && bias, GANCE_E_NULL, "gance_torgb_skip: NULL pointer");
    GANCE_REQUIRE(y_out || frames, GANCE_E_NULL, "gance_torgb_skip: neither y_out nor frames given");
    GANCE_REQUIRE(batch > 0 && h > 0 && w > 0 && channels % 8 == 0 && channels <= 4096, GANCE_E_SHAPE,
                  "gance_torgb_skip: shape [%d,%d,%d,%d]", batch, h, w, channels);
    GANCE_REQUIRE(y_prev == nullptr || (h % 2 == 0 && w % 2 == 0), GANCE_E_SHAPE,
                  "gance_torgb_skip: skip input needs even h,w");
    GANCE_REQUIRE(aligned16(x) && aligned16(y_prev) && aligned16(y_out), GANCE_E_ALIGN,
                  "gance_torgb_skip: pointers must be 16-byte aligned");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == GANCE_F32)
        return launch_torgb<float>(x, batch, h, w, channels, styles, weight, bias, y_prev, y_out, frames, s);
    if (dtype == GANCE_BF16)
        return launch_torgb<__nv_bfloat16>(x, batch, h, w, channels, styles, weight, bias, y_prev, y_out, frames, s);
    if (dtype == GANCE_F16)
        return launch_torgb<__half>(x, batch, h, w, channels, styles, weight, bias, y_prev, y_out, frames, s);
    return fail(GANCE_E_DTYPE, "gance_torgb_skip: dtype %d", dtype);
}

int gance_to_uint8_nhwc(const float* y, int batch, int h, int w, uint8_t* frames, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(y && frames, GANCE_E_NULL, "gance_to_uint8_nhwc: NULL pointer");
    GANCE_REQUIRE(batch > 0 && h > 0 && w > 0, GANCE_E_SHAPE, "gance_to_uint8_nhwc: shape");
    size_t pixels = (size_t)batch * h * w;
    to_uint8_kernel<<<grid_for(pixels, 256), 256, 0, (cudaStream_t)stream>>>((const float4*)y, pixels, frames, nonfinite_flag());
    return check_launch("to_uint8_kernel");
}

int gance_normal_noise(float* out, long long count, unsigned long long seed, unsigned long long offset, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(out != nullptr, GANCE_E_NULL, "gance_normal_noise: NULL pointer");
    GANCE_REQUIRE(count > 0 && count % 4 == 0, GANCE_E_SHAPE, "gance_normal_noise: count %lld must be a positive multiple of 4", count);
    GANCE_REQUIRE(aligned16(out), GANCE_E_ALIGN, "gance_normal_noise: out must be 16-byte aligned");
    const size_t groups = (size_t)count / 4;
    normal_noise_kernel<<<grid_for(groups, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4*>(out), groups, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)), offset, nullptr);
    return check_launch("normal_noise_kernel");
}

int gance_normal_noise_next(float* out, long long count, unsigned long long seed, unsigned long long* counter, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(out != nullptr && counter != nullptr, GANCE_E_NULL, "gance_normal_noise_next: NULL pointer");
    GANCE_REQUIRE(count > 0 && count % 4 == 0, GANCE_E_SHAPE, "gance_normal_noise_next: count %lld must be a positive multiple of 4", count);
    GANCE_REQUIRE(aligned16(out), GANCE_E_ALIGN, "gance_normal_noise_next: out must be 16-byte aligned");
    const size_t groups = (size_t)count / 4;
    normal_noise_kernel<<<grid_for(groups, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<float4*>(out), groups, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)), 0ull, counter);
    int rc = check_launch("normal_noise_kernel");
    if (rc != GANCE_OK) return rc;
    advance_counter_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(counter, (unsigned long long)groups);
    return check_launch("advance_counter_kernel");
}

}  // extern "C"
