// Audio -> per-frame latents on the device, in float64 (the min-max scale is anchored
// on the quietest dB bin, which float32 gets wrong by percents). Reference code being replaced:
//   gance/apply_spectrogram.py:20-118, gance/vector_sources/vector_sources_common.py:136-428,
//   gance/data_into_network_visualization/visualization_inputs.py:53-270.
// Operation order and rounding follow numpy/sklearn where the result is an affine map of the data
// (explicit __dmul_rn/__dadd_rn: no FMA contraction), so those stages are bit-exact given equal
// inputs; the DFT, log10 and the resampling product differ from numpy only by summation order.
#include "common.cuh"

namespace gance {


// Total order on doubles as unsigned integers, for atomicMin/atomicMax on signed values.
__device__ __forceinline__ unsigned long long encode_ordered(double v) {
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double decode_ordered(unsigned long long u) {
    u = (u & 0x8000000000000000ull) ? (u & 0x7FFFFFFFFFFFFFFFull) : ~u;
    return __longlong_as_double((long long)u);
}

__global__ void init_reduction_slots(unsigned long long* global_max, unsigned long long* minmax) {
    if (global_max) *global_max = 0ull;  // |X| >= 0: plain bit pattern is ordered
    if (minmax) {
        minmax[0] = 0xFFFFFFFFFFFFFFFFull;  // running min (encoded)
        minmax[1] = 0ull;                   // running max (encoded)
    }
}

// ------------------------------------------------------------------------------------------------
// a3: frames of `window` samples at hop `hop`, periodic Hann, DFT bins [0, window/2), |X|.
// One CTA per frame; thread k owns bin k. The windowed frame and the twiddle table
// exp(-2 pi i m / window) live in shared memory; the twiddle index (k*n mod window) is advanced
// incrementally. Per-CTA maximum of |X| goes to *global_max by atomicMax on the bit pattern.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) stft_mag_kernel(const float* __restrict__ audio, int hop, int window,
                                                       int bins, double* __restrict__ mag,
                                                       unsigned long long* __restrict__ global_max) {
    extern __shared__ double smem_d[];
    double* xw = smem_d;               // [window]
    double* tw_cos = xw + window;      // [window]
    double* tw_sin = tw_cos + window;  // [window]
    __shared__ double warp_max[8];
    const int t = blockIdx.x;
    const float* frame = audio + (size_t)t * hop;
    for (int n = threadIdx.x; n < window; n += blockDim.x) {
        double s, c;
        sincospi(2.0 * (double)n / (double)window, &s, &c);
        tw_cos[n] = c;
        tw_sin[n] = s;
        // np.hanning(window + 1)[:-1]: 0.5 - 0.5 cos(2 pi n / window); float32 sample * float64 window
        xw[n] = (double)frame[n] * (0.5 - 0.5 * c);
    }
    __syncthreads();
    double local_max = 0.0;
    for (int k = threadIdx.x; k < bins; k += blockDim.x) {
        double re = 0.0, im = 0.0;
        int idx = 0;
        for (int n = 0; n < window; ++n) {
            const double v = xw[n];
            re = fma(v, tw_cos[idx], re);
            im = fma(-v, tw_sin[idx], im);
            idx += k;
            if (idx >= window) idx -= window;
        }
        const double m = hypot(re, im);
        mag[(size_t)t * bins + k] = m;
        local_max = fmax(local_max, m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local_max = fmax(local_max, __shfl_xor_sync(0xffffffffu, local_max, o));
    if ((threadIdx.x & 31) == 0) warp_max[threadIdx.x >> 5] = local_max;
    __syncthreads();
    if (threadIdx.x == 0) {
        double m = warp_max[0];
        for (int i = 1; i < (int)(blockDim.x >> 5); ++i) m = fmax(m, warp_max[i]);
        atomicMax(global_max, (unsigned long long)__double_as_longlong(m));
    }
}

// ------------------------------------------------------------------------------------------------
// a3 (dB) + a4 (resample): D = 20 log10(mag / max); V[t,i] = sum_k D[t,k] * Rt[k,i].
// One CTA per frame, one thread per output position; Rt is read coalesced along i (L2 resident).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) db_resample_kernel(const double* __restrict__ mag,
                                                          const unsigned long long* __restrict__ global_max,
                                                          int bins, int vector_length,
                                                          const double* __restrict__ resample_t,
                                                          double* __restrict__ out,
                                                          unsigned long long* __restrict__ minmax) {
    extern __shared__ double db[];  // [bins]
    __shared__ double red_min[16], red_max[16];
    const int t = blockIdx.x;
    const double peak = __longlong_as_double((long long)*global_max);
    for (int k = threadIdx.x; k < bins; k += blockDim.x) db[k] = 20.0 * log10(mag[(size_t)t * bins + k] / peak);
    __syncthreads();
    double lo = INFINITY, hi = -INFINITY;
    for (int i = threadIdx.x; i < vector_length; i += blockDim.x) {
        double a0 = 0.0, a1 = 0.0;
        int k = 0;
        for (; k + 1 < bins; k += 2) {
            a0 = fma(db[k], __ldg(resample_t + (size_t)k * vector_length + i), a0);
            a1 = fma(db[k + 1], __ldg(resample_t + (size_t)(k + 1) * vector_length + i), a1);
        }
        if (k < bins) a0 = fma(db[k], __ldg(resample_t + (size_t)k * vector_length + i), a0);
        const double v = a0 + a1;
        out[(size_t)t * vector_length + i] = v;
        lo = fmin(lo, v);
        hi = fmax(hi, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        red_min[threadIdx.x >> 5] = lo;
        red_max[threadIdx.x >> 5] = hi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < (int)(blockDim.x >> 5); ++i) {
            lo = fmin(lo, red_min[i]);
            hi = fmax(hi, red_max[i]);
        }
        // NaN (an all-zero frame -> log10(0/..)) is propagated by the reference; fmin/fmax drop it, so
        // the scale stays finite here. Documented divergence for degenerate audio.
        atomicMin(minmax, encode_ordered(lo));
        atomicMax(minmax + 1, encode_ordered(hi));
    }
}

// ------------------------------------------------------------------------------------------------
// a4 (min-max scale, sklearn op order) fused with a5 (Savitzky-Golay along time).
// proj is the [window, window] least-squares projection matrix: row r = fitted value at window
// position r. Interior outputs use the middle row; the first/last `half` frames use rows of the
// first/last window (scipy mode='interp').
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) minmax_savgol_time_kernel(const double* __restrict__ in,
                                                                 const unsigned long long* __restrict__ minmax,
                                                                 double range_lo, double range_hi, int num_frames,
                                                                 int vector_length, const double* __restrict__ proj,
                                                                 int window, double* __restrict__ out) {
    const double data_min = decode_ordered(minmax[0]), data_max = decode_ordered(minmax[1]);
    double data_range = data_max - data_min;
    if (data_range == 0.0) data_range = 1.0;  // sklearn _handle_zeros_in_scale
    const double scale = (range_hi - range_lo) / data_range;
    const double offset = __dsub_rn(range_lo, __dmul_rn(data_min, scale));
    const int half = window / 2;
    const size_t total = (size_t)num_frames * vector_length;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e % vector_length);
        const int t = (int)(e / vector_length);
        int row, first;
        if (t < half) {
            row = t;
            first = 0;
        } else if (t >= num_frames - half) {
            row = window - (num_frames - t);
            first = num_frames - window;
        } else {
            row = half;
            first = t - half;
        }
        const double* c = proj + (size_t)row * window;
        double acc = 0.0;
        for (int j = 0; j < window; ++j) {
            const double x = in[(size_t)(first + j) * vector_length + i];
            const double scaled = __dadd_rn(__dmul_rn(x, scale), offset);
            acc = fma(__ldg(c + j), scaled, acc);
        }
        out[e] = acc;
    }
}

// ------------------------------------------------------------------------------------------------
// a8 (circular left shift by the running sum of roll values) + a6 (Savitzky-Golay within a vector).
// One CTA per frame; the (rotated) vector is staged in shared memory.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) roll_savgol_vector_kernel(const double* __restrict__ in,
                                                                 const int* __restrict__ shift, int vector_length,
                                                                 const double* __restrict__ proj, int window,
                                                                 double* __restrict__ out) {
    extern __shared__ double xs[];  // [vector_length]
    const int t = blockIdx.x;
    int s = shift ? shift[t] % vector_length : 0;
    if (s < 0) s += vector_length;
    for (int i = threadIdx.x; i < vector_length; i += blockDim.x) {
        int src = i + s;  // np.roll(v, -s)[i] = v[(i + s) mod L]
        if (src >= vector_length) src -= vector_length;
        xs[i] = in[(size_t)t * vector_length + src];
    }
    __syncthreads();
    const int half = window / 2;
    for (int i = threadIdx.x; i < vector_length; i += blockDim.x) {
        int row, first;
        if (i < half) {
            row = i;
            first = 0;
        } else if (i >= vector_length - half) {
            row = window - (vector_length - i);
            first = vector_length - window;
        } else {
            row = half;
            first = i - half;
        }
        const double* c = proj + (size_t)row * window;
        double acc = 0.0;
        for (int j = 0; j < window; ++j) acc = fma(__ldg(c + j), xs[first + j], acc);
        out[(size_t)t * vector_length + i] = acc;
    }
}

// ------------------------------------------------------------------------------------------------
// a10 / a11: alpha blends. `a * (1 - alpha)` is a float32 product in the reference (float32 array
// times a Python float), the rest float64; no FMA contraction so the result is bit-exact.
// ------------------------------------------------------------------------------------------------
__global__ void alpha_blend_noise_kernel(const double* __restrict__ spec, const float* __restrict__ noise,
                                         float one_minus_alpha, double alpha, size_t total,
                                         double* __restrict__ out64, float* __restrict__ out32) {
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const double v = __dadd_rn((double)__fmul_rn(noise[e], one_minus_alpha), __dmul_rn(spec[e], alpha));
        if (out64) out64[e] = v;
        if (out32) out32[e] = (float)v;
    }
}

__global__ void alpha_blend_wplus_kernel(const double* __restrict__ spec, const float* __restrict__ projected,
                                         int duplication, float one_minus_alpha, double alpha, int blend_depth,
                                         int depth, int num_frames, int vector_length,
                                         double* __restrict__ out64, float* __restrict__ out32) {
    const size_t per_row = (size_t)num_frames * vector_length;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < per_row; e += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(e % vector_length);
        const int t = (int)(e / vector_length);
        const float w = projected[(size_t)(t / duplication) * vector_length + i];
        const double blended = __dadd_rn((double)__fmul_rn(w, one_minus_alpha), __dmul_rn(spec[e], alpha));
        for (int r = 0; r < depth; ++r) {
            const double v = r < blend_depth ? blended : (double)w;
            if (out64) out64[(size_t)r * per_row + e] = v;                                  // (depth, T*L)
            if (out32) out32[((size_t)t * depth + r) * vector_length + i] = (float)v;       // [T, depth, L]
        }
    }
}

// ------------------------------------------------------------------------------------------------
// a1: int16/int32/uint8 PCM -> float32 in [-1,1] (music.read_wav_file, music.py:190-205, through
// remap_values_into_range = scipy interp1d, vector_sources_common.py:44-61):
// y = slope*(x - x_lo) + y_lo in float64 with slope = (y_hi-y_lo)/(x_hi-x_lo), then float32.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void pcm_to_float_kernel(const T* __restrict__ pcm, size_t n, double x_lo, double slope, double y_lo,
                                    float* __restrict__ out) {
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x)
        out[e] = (float)__dadd_rn(__dmul_rn(slope, __dsub_rn((double)pcm[e], x_lo)), y_lo);
}

// ------------------------------------------------------------------------------------------------
// a2: band-limited sinc resampling, resampy 0.2.2 interpn.resample_f with the kaiser_best filter
// (music.py:212-230). One thread per OUTPUT sample walks the left wing then the right wing of the
// interpolation window in the original order, accumulating like the numba loop does:
// y (float32) <- float32(float64(y) + weight*x) with weight = win[o] + eta*delta[o] in float64, no FMA
// contraction - so the result is bit-identical to the reference. `times` is the float64 running sum
// time_register (built on the host by sequential addition, exactly as the loop accumulates it).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) sinc_resample_kernel(const float* __restrict__ x, int n_in,
                                                            const double* __restrict__ times, int n_out,
                                                            const double* __restrict__ win,
                                                            const double* __restrict__ delta, int n_win,
                                                            double scale, int num_table, int index_step,
                                                            float* __restrict__ y) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    const double time_register = times[t];
    const int n = (int)time_register;
    double frac = __dmul_rn(scale, __dsub_rn(time_register, (double)n));
    double index_frac = __dmul_rn(frac, (double)num_table);
    int offset = (int)index_frac;
    double eta = __dsub_rn(index_frac, (double)offset);
    float acc = 0.f;
    int i_max = (n_win - offset) / index_step;
    if (n + 1 < i_max) i_max = n + 1;
    for (int i = 0; i < i_max; ++i) {
        const int o = offset + i * index_step;
        const double weight = __dadd_rn(__ldg(win + o), __dmul_rn(eta, __ldg(delta + o)));
        acc = (float)__dadd_rn((double)acc, __dmul_rn(weight, (double)__ldg(x + n - i)));
    }
    frac = __dsub_rn(scale, frac);
    index_frac = __dmul_rn(frac, (double)num_table);
    offset = (int)index_frac;
    eta = __dsub_rn(index_frac, (double)offset);
    int k_max = (n_win - offset) / index_step;
    if (n_in - n - 1 < k_max) k_max = n_in - n - 1;
    for (int k = 0; k < k_max; ++k) {
        const int o = offset + k * index_step;
        const double weight = __dadd_rn(__ldg(win + o), __dmul_rn(eta, __ldg(delta + o)));
        acc = (float)__dadd_rn((double)acc, __dmul_rn(weight, (double)__ldg(x + n + k + 1)));
    }
    y[t] = acc;
}

static int grid1d(size_t items, int block) {
    size_t blocks = (items + block - 1) / block;
    const size_t cap = 148u * 8u;
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

}  // namespace gance

using namespace gance;

// ---- a9: the wrap-around Gaussian filter ALONG TIME of the noise source -------------------------------------------------
// scipy.ndimage.gaussian_filter(x, sigma=(s,0,0), mode="wrap") of /root/reference/gance/vector_sources/primatives.py:69
// (scipy ni_filters.c, NI_Correlate1D, symmetric-kernel branch): the line is converted to double, extended periodically,
//     acc = x[t] * w[0];  for k = radius .. 1:  acc += (x[t-k] + x[t+k]) * w[k];      (separately rounded, no FMA)
// and cast back to float. One thread per (t, j), j fastest so every tap is a coalesced row read; the 22 MB input of an
// album-length track stays in L2. On the host this filter was 61 % of the whole audio -> latents time (0.68 s of 1.1 s
// for 10 800 frames, and x8 slower with eight ranks sharing the host cores).
__global__ void gaussian_filter_time_wrap_kernel(const float* __restrict__ in, int num_vectors, int vector_length,
                                                 const double* __restrict__ weights, int radius, float* __restrict__ out) {
    const size_t total = (size_t)num_vectors * vector_length;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int t = (int)(idx / vector_length), j = (int)(idx - (size_t)t * vector_length);
        double acc = __dmul_rn((double)in[idx], __ldg(weights));
        int before = (t - radius) % num_vectors, after = (t + radius) % num_vectors;  // t-k and t+k, kept in [0, T)
        if (before < 0) before += num_vectors;
        for (int k = radius; k >= 1; --k) {
            const double pair = __dadd_rn((double)__ldg(in + (size_t)before * vector_length + j),
                                          (double)__ldg(in + (size_t)after * vector_length + j));
            acc = __dadd_rn(acc, __dmul_rn(pair, __ldg(weights + k)));
            if (++before == num_vectors) before = 0;
            if (--after < 0) after = num_vectors - 1;
        }
        out[idx] = (float)acc;
    }
}

extern "C" {

int gance_pcm_to_float(const void* pcm, int bits, long long num_samples, float* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(pcm && out, GANCE_E_NULL, "gance_pcm_to_float: NULL pointer");
    GANCE_REQUIRE(num_samples > 0, GANCE_E_SHAPE, "gance_pcm_to_float: num_samples=%lld", num_samples);
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = (size_t)num_samples;
    const int grid = grid1d(n, 256);
    // input ranges of music.py:198-203 (the reference maps 8-bit PCM from (0, 255))
    if (bits == 16) {
        pcm_to_float_kernel<int16_t><<<grid, 256, 0, s>>>((const int16_t*)pcm, n, -32768.0, 2.0 / 65535.0, -1.0, out);
    } else if (bits == 32) {
        pcm_to_float_kernel<int32_t><<<grid, 256, 0, s>>>((const int32_t*)pcm, n, -2147483648.0, 2.0 / 4294967295.0,
                                                          -1.0, out);
    } else if (bits == 8) {
        pcm_to_float_kernel<int8_t><<<grid, 256, 0, s>>>((const int8_t*)pcm, n, 0.0, 2.0 / 255.0, -1.0, out);
    } else {
        return fail(GANCE_E_DTYPE, "gance_pcm_to_float: %d-bit PCM is not supported", bits);
    }
    return check_launch("pcm_to_float_kernel");
}

int gance_sinc_resample(const float* x, int n_in, const double* times, int n_out, const double* interp_win,
                        const double* interp_delta, int n_win, double scale, int num_table, float* y,
                        void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(x && times && interp_win && interp_delta && y, GANCE_E_NULL, "gance_sinc_resample: NULL pointer");
    const int index_step = (int)(scale * num_table);
    GANCE_REQUIRE(n_in > 0 && n_out > 0 && n_win > 0 && num_table > 0 && scale > 0.0 && scale <= 1.0 && index_step > 0,
                  GANCE_E_SHAPE, "gance_sinc_resample: n_in=%d n_out=%d n_win=%d scale=%g", n_in, n_out, n_win, scale);
    sinc_resample_kernel<<<(n_out + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        x, n_in, times, n_out, interp_win, interp_delta, n_win, scale, num_table, index_step, y);
    return check_launch("sinc_resample_kernel");
}

int gance_stft_mag(const float* audio, int num_frames, int vector_length, double* mag, double* global_max,
                   void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(audio && mag && global_max, GANCE_E_NULL, "gance_stft_mag: NULL pointer");
    const int window = vector_length - 2;  // apply_spectrogram.py:68  m = num_frequency_bins - 1 * 2
    GANCE_REQUIRE(num_frames > 0 && window >= 2 && window <= 4096, GANCE_E_SHAPE,
                  "gance_stft_mag: num_frames=%d vector_length=%d", num_frames, vector_length);
    cudaStream_t s = (cudaStream_t)stream;
    init_reduction_slots<<<1, 1, 0, s>>>((unsigned long long*)global_max, nullptr);
    int rc = check_launch("init_reduction_slots");
    if (rc != GANCE_OK) return rc;
    const size_t smem = (size_t)3 * window * sizeof(double);
    stft_mag_kernel<<<num_frames, 256, smem, s>>>(audio, vector_length, window, window / 2, mag,
                                                  (unsigned long long*)global_max);
    return check_launch("stft_mag_kernel");
}

int gance_db_resample(const double* mag, const double* global_max, int num_frames, int bins, int vector_length,
                      const double* resample_matrix_t, double* out, double* minmax, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(mag && global_max && resample_matrix_t && out && minmax, GANCE_E_NULL,
                  "gance_db_resample: NULL pointer");
    GANCE_REQUIRE(num_frames > 0 && bins > 0 && bins <= 4096 && vector_length > 0, GANCE_E_SHAPE,
                  "gance_db_resample: shape");
    cudaStream_t s = (cudaStream_t)stream;
    init_reduction_slots<<<1, 1, 0, s>>>(nullptr, (unsigned long long*)minmax);
    int rc = check_launch("init_reduction_slots");
    if (rc != GANCE_OK) return rc;
    db_resample_kernel<<<num_frames, 512, (size_t)bins * sizeof(double), s>>>(
        mag, (const unsigned long long*)global_max, bins, vector_length, resample_matrix_t, out,
        (unsigned long long*)minmax);
    return check_launch("db_resample_kernel");
}

int gance_minmax_savgol_time(const double* in, const double* minmax, double range_lo, double range_hi,
                             int num_frames, int vector_length, const double* coeffs, int window, double* out,
                             void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(in && minmax && coeffs && out, GANCE_E_NULL, "gance_minmax_savgol_time: NULL pointer");
    GANCE_REQUIRE(window % 2 == 1 && num_frames >= window && vector_length > 0, GANCE_E_SHAPE,
                  "gance_minmax_savgol_time: %d frames is fewer than the filter window %d (scipy raises too)",
                  num_frames, window);
    minmax_savgol_time_kernel<<<grid1d((size_t)num_frames * vector_length, 256), 256, 0, (cudaStream_t)stream>>>(
        in, (const unsigned long long*)minmax, range_lo, range_hi, num_frames, vector_length, coeffs, window, out);
    return check_launch("minmax_savgol_time_kernel");
}

int gance_roll_savgol_vector(const double* in, const int* shift, int num_frames, int vector_length,
                             const double* coeffs, int window, double* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(in && coeffs && out, GANCE_E_NULL, "gance_roll_savgol_vector: NULL pointer");
    GANCE_REQUIRE(in != out, GANCE_E_SHAPE, "gance_roll_savgol_vector: in-place operation is not supported");
    GANCE_REQUIRE(window % 2 == 1 && vector_length >= window && vector_length <= 6144 && num_frames > 0,
                  GANCE_E_SHAPE, "gance_roll_savgol_vector: vector_length=%d window=%d", vector_length, window);
    roll_savgol_vector_kernel<<<num_frames, 512, (size_t)vector_length * sizeof(double), (cudaStream_t)stream>>>(
        in, shift, vector_length, coeffs, window, out);
    return check_launch("roll_savgol_vector_kernel");
}

int gance_alpha_blend_noise(const double* spec, const float* noise, double alpha, int num_frames, int vector_length,
                            double* latents_f64, float* latents_f32, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(spec && noise && (latents_f64 || latents_f32), GANCE_E_NULL, "gance_alpha_blend_noise: NULL pointer");
    GANCE_REQUIRE(num_frames > 0 && vector_length > 0, GANCE_E_SHAPE, "gance_alpha_blend_noise: shape");
    const size_t total = (size_t)num_frames * vector_length;
    alpha_blend_noise_kernel<<<grid1d(total, 256), 256, 0, (cudaStream_t)stream>>>(
        spec, noise, (float)(1.0 - alpha), alpha, total, latents_f64, latents_f32);
    return check_launch("alpha_blend_noise_kernel");
}

int gance_alpha_blend_wplus(const double* spec, const float* projected, int duplication, double alpha,
                            int blend_depth, int depth, int num_frames, int vector_length, double* latents_f64,
                            float* latents_f32, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(spec && projected && (latents_f64 || latents_f32), GANCE_E_NULL,
                  "gance_alpha_blend_wplus: NULL pointer");
    GANCE_REQUIRE(num_frames > 0 && vector_length > 0 && duplication > 0 && num_frames % duplication == 0 &&
                      depth > 0 && blend_depth >= 0,
                  GANCE_E_SHAPE, "gance_alpha_blend_wplus: %d frames cannot be split into runs of %d", num_frames,
                  duplication);
    alpha_blend_wplus_kernel<<<grid1d((size_t)num_frames * vector_length, 256), 256, 0, (cudaStream_t)stream>>>(
        spec, projected, duplication, (float)(1.0 - alpha), alpha, blend_depth, depth, num_frames, vector_length,
        latents_f64, latents_f32);
    return check_launch("alpha_blend_wplus_kernel");
}

int gance_gaussian_filter_time_wrap(const float* in, int num_vectors, int vector_length, const double* weights, int radius,
                                    float* out, void* stream) {
    GANCE_REQUIRE(is_initialised(), GANCE_E_NOT_INIT, "gance_init() has not been called");
    GANCE_REQUIRE(in && weights && out && in != out, GANCE_E_NULL, "gance_gaussian_filter_time_wrap: NULL or aliased pointer");
    GANCE_REQUIRE(num_vectors > 0 && vector_length > 0 && radius >= 0, GANCE_E_SHAPE,
                  "gance_gaussian_filter_time_wrap: T=%d L=%d radius=%d", num_vectors, vector_length, radius);
    const size_t n = (size_t)num_vectors * vector_length;
    gaussian_filter_time_wrap_kernel<<<grid1d(n, 256), 256, 0, (cudaStream_t)stream>>>(in, num_vectors, vector_length, weights,
                                                                                      radius, out);
    return check_launch("gaussian_filter_time_wrap_kernel");
}

}  // extern "C"
