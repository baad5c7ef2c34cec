// fp32 signal-processing kernels of the front end (MelSpec) and back end (Vocos ISTFT head):
//   melspec      : reflect pad + periodic Hann + 1024-point FFT + |.| + HTK mel filterbank + log(clamp)   (modules.py:75-143)
//   istft_head   : exp / clip / cos / sin + inverse FFT + window + overlap-add + envelope normalise + centre trim
//                  (vocos ISTFTHead; in-tree corroboration runtime/triton_trtllm/scripts/export_vocoder_to_onnx.py:45-59,
//                  scripts/conv_stft.py:193-234)
// Both are HBM/launch-bound (about 1 KB in, 0.4 KB out per frame): one CTA processes 8 frames (hops) with all FFTs in
// shared memory, two REAL frames packed into one complex 1024-point FFT, so every input byte is read from HBM once
// (the 3-frame halo of the overlap-add is served by L2) and every output byte written once.
#include "host.h"

namespace f5e {

constexpr int NFFT = 1024;
constexpr int HOP = 256;
constexpr int NBIN = NFFT / 2 + 1;  // 513
constexpr int DSP_THREADS = 256;

__device__ __forceinline__ int bitrev10(int k) { return (int)(__brev((unsigned)k) >> 22); }

// In-place radix-2 decimation-in-frequency FFT of `nfft` independent 1024-point complex sequences held in shared
// memory (z[f * 1024 + n]).  Natural-order input, BIT-REVERSED output: X[k] ends up at index bitrev10(k).
// tw[j] = exp(-2*pi*i*j/1024), j < 512 (fp32, rounded from fp64); INVERSE conjugates the twiddles (no 1/N scale).
template <bool INVERSE>
__device__ __forceinline__ void fft1024_smem(float2* z, const float2* __restrict__ tw, int nfft) {
  for (int s = 0; s < 10; ++s) {
    const int half = 512 >> s;
    for (int j = threadIdx.x; j < nfft * 512; j += blockDim.x) {
      const int f = j >> 9, b = j & 511;
      const int pos = b & (half - 1);
      const int i0 = f * NFFT + ((b - pos) << 1) + pos, i1 = i0 + half;
      const float2 a = z[i0], c = z[i1];
      float2 w = __ldg(tw + (pos << s));
      if (INVERSE) w.y = -w.y;
      const float dx = a.x - c.x, dy = a.y - c.y;
      z[i0] = make_float2(a.x + c.x, a.y + c.y);
      z[i1] = make_float2(dx * w.x - dy * w.y, dx * w.y + dy * w.x);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// melspec: wav [B, nw] fp32 -> log-mel [B, n_mels, T], T = nw / 256 + 1
// grid (ceil(T / 8), B); dynamic smem: 4 complex FFT buffers (32 KB) + 8 x 513 magnitudes (16.4 KB)
// fbT: [n_mels, 513] filterbank (transposed), fb_lo/fb_hi: first / one-past-last non-zero bin of every filter.
// Algorithmic bytes: 4*nw + 4*n_mels*T per utterance.
// ---------------------------------------------------------------------------------------------
constexpr int MEL_FR = 8;
__global__ void __launch_bounds__(DSP_THREADS)
melspec_kernel(const float* __restrict__ wav, long long ld_wav, int nw, const float* __restrict__ window,
               const float2* __restrict__ tw, const float* __restrict__ fbT, const int* __restrict__ fb_lo,
               const int* __restrict__ fb_hi, int n_mels, float* __restrict__ out, int T) {
  extern __shared__ __align__(16) uint8_t dsp_smem[];
  float2* z = reinterpret_cast<float2*>(dsp_smem);                       // [4][1024]
  float* mag = reinterpret_cast<float*>(dsp_smem + 4 * NFFT * 8);       // [8][513]
  const int b = blockIdx.y, t0 = blockIdx.x * MEL_FR;
  const float* x = wav + (size_t)b * ld_wav;
  // 1. windowed frames, reflect padding of NFFT/2 at both ends (torch.stft center=True, pad_mode="reflect")
  for (int i = threadIdx.x; i < 4 * NFFT; i += blockDim.x) {
    const int f = i >> 10, n = i & (NFFT - 1);
    const float w = __ldg(window + n);
    float v[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int t = t0 + 2 * f + h;
      int p = t * HOP - NFFT / 2 + n;
      if (p < 0) p = -p;
      if (p >= nw) p = 2 * (nw - 1) - p;
      v[h] = (t < T) ? x[p] * w : 0.f;
    }
    z[i] = make_float2(v[0], v[1]);
  }
  __syncthreads();
  fft1024_smem<false>(z, tw, 4);
  // 2. split the packed pair: Xa[k] = (Z[k] + conj Z[N-k]) / 2, Xb[k] = (Z[k] - conj Z[N-k]) / (2i); magnitudes
  for (int i = threadIdx.x; i < 4 * NBIN; i += blockDim.x) {
    const int f = i / NBIN, k = i - f * NBIN;
    const float2 p = z[f * NFFT + bitrev10(k)], q = z[f * NFFT + bitrev10((NFFT - k) & (NFFT - 1))];
    const float ar = 0.5f * (p.x + q.x), ai = 0.5f * (p.y - q.y);
    const float br = 0.5f * (p.y + q.y), bi = 0.5f * (q.x - p.x);
    mag[(2 * f) * NBIN + k] = sqrtf(ar * ar + ai * ai);
    mag[(2 * f + 1) * NBIN + k] = sqrtf(br * br + bi * bi);
  }
  __syncthreads();
  // 3. mel filterbank (sparse triangular rows) + log(clamp(., 1e-5)); thread -> (mel, frame) with the frame fastest
  for (int i = threadIdx.x; i < n_mels * MEL_FR; i += blockDim.x) {
    const int m = i / MEL_FR, fr = i - m * MEL_FR;
    const int t = t0 + fr;
    if (t >= T) continue;
    const float* fr_mag = mag + fr * NBIN;
    const float* frow = fbT + (size_t)m * NBIN;
    float acc = 0.f;
    for (int k = __ldg(fb_lo + m); k < __ldg(fb_hi + m); ++k) acc = fmaf(fr_mag[k], __ldg(frow + k), acc);
    out[((size_t)b * n_mels + m) * T + t] = logf(fmaxf(acc, 1e-5f));
  }
}

int melspec(cudaStream_t st, const float* wav, long long ld_wav, int batch, int nw, const float* window, const float* tw,
            const float* fbT, const int* fb_lo, const int* fb_hi, int n_mels, float* out) {
  F5E_REQUIRE(batch > 0 && nw > NFFT / 2, "melspec: need at least %d samples for reflect padding (nw=%d)", NFFT / 2 + 1, nw);
  const int T = nw / HOP + 1;
  const int smem = 4 * NFFT * 8 + MEL_FR * NBIN * 4;
  static bool configured = false;
  if (!configured) {
    F5E_CUDA_CHECK(cudaFuncSetAttribute(melspec_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid((T + MEL_FR - 1) / MEL_FR, batch);
  LaunchScope ls(st, PROF_DSP, batch, nw, 0);
  melspec_kernel<<<grid, DSP_THREADS, smem, st>>>(wav, ld_wav, nw, window, reinterpret_cast<const float2*>(tw), fbT, fb_lo,
                                                  fb_hi, n_mels, out, T);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// istft_head: spec [B, t_pad, ld_spec] fp32 rows = (log-magnitude[513] | phase[513]) per frame; frames >= n_frames[b]
// are ignored.  out [B, ld_out] fp32, samples [0, (n_frames[b]-1)*256) written, the rest of [0, out_len) zeroed.
//   S = min(exp(mag), 100) * (cos p + i sin p);  x_t = irfft(S_t, 1024) * w;  y[p] = sum_t x_t[p - 256 t] / sum_t w^2[p - 256 t]
//   out[n] = y[n + 512]                                  (torch.istft center=True semantics)
// grid (ceil(hops / 8), B): a CTA produces 8 hops (2048 samples) from the 11 frames that overlap them (12 with
// pairing = 6 complex inverse FFTs, 48 KB smem).  Algorithmic bytes per frame: 1026*4 read + 256*4 written.
// ---------------------------------------------------------------------------------------------
constexpr int IST_HOPS = 8;
constexpr int IST_FRAMES = 12;
__global__ void __launch_bounds__(DSP_THREADS)
istft_head_kernel(const float* __restrict__ spec, long long ld_spec, int t_pad, const int* __restrict__ n_frames,
                  const float* __restrict__ window, const float2* __restrict__ tw, float* __restrict__ out,
                  long long ld_out, int out_len) {
  extern __shared__ __align__(16) uint8_t dsp_smem[];
  float2* z = reinterpret_cast<float2*>(dsp_smem);  // [6][1024]
  const int b = blockIdx.y, h0 = blockIdx.x * IST_HOPS;
  const int T = n_frames ? min(__ldg(n_frames + b), t_pad) : t_pad;
  const int len = T > 0 ? (T - 1) * HOP : 0;
  float* o = out + (size_t)b * ld_out;
  if (h0 * HOP >= len) {  // nothing but zero fill for this CTA
    for (int i = threadIdx.x; i < IST_HOPS * HOP; i += blockDim.x)
      if (h0 * HOP + i < out_len) o[h0 * HOP + i] = 0.f;
    return;
  }
  const int tbase = h0 - 1;  // first frame touching hop h0
  // 1. Hermitian-extended spectra of frame pairs: Z[k] = Sa~[k] + i Sb~[k].  Pass 0 writes the even frame of every pair
  //    (S at k, conj S at N-k), pass 1 adds i * (the odd frame): i*S = (-im, re) at k, i*conj S = (im, re) at N-k.
  for (int pass = 0; pass < 2; ++pass) {
    for (int i = threadIdx.x; i < (IST_FRAMES / 2) * NBIN; i += blockDim.x) {
      const int pr = i / NBIN, k = i - pr * NBIN;
      const int t = tbase + 2 * pr + pass;
      float re = 0.f, im = 0.f;
      if (t >= 0 && t < T) {
        const float* row = spec + ((size_t)b * t_pad + t) * ld_spec;
        const float m = fminf(expf(row[k]), 100.0f);
        float sn, cs;
        sincosf(row[NBIN + k], &sn, &cs);
        re = m * cs;
        im = (k == 0 || k == NFFT / 2) ? 0.f : m * sn;  // irfft ignores the imaginary part of DC / Nyquist
      }
      float2* zz = z + pr * NFFT;
      const bool mirror = (k != 0 && k != NFFT / 2);
      if (pass == 0) {
        zz[k] = make_float2(re, im);
        if (mirror) zz[NFFT - k] = make_float2(re, -im);
      } else {
        const float2 a = zz[k];
        zz[k] = make_float2(a.x - im, a.y + re);
        if (mirror) {
          const float2 c = zz[NFFT - k];
          zz[NFFT - k] = make_float2(c.x + im, c.y + re);
        }
      }
    }
    __syncthreads();
  }
  fft1024_smem<true>(z, tw, IST_FRAMES / 2);
  // 2. window, overlap-add (fixed frame order: deterministic), envelope, trim
  for (int i = threadIdx.x; i < IST_HOPS * HOP; i += blockDim.x) {
    const int n = h0 * HOP + i;  // output sample
    if (n >= out_len) continue;
    if (n >= len) { o[n] = 0.f; continue; }
    const int p = n + NFFT / 2;  // untrimmed coordinate
    const int h = n / HOP;
    float acc = 0.f, env = 0.f;
#pragma unroll
    for (int d = -1; d <= 2; ++d) {
      const int t = h + d;
      if (t < 0 || t >= T) continue;
      const int m = p - t * HOP;  // position inside frame t, in [0, 1024)
      const int fr = t - tbase;
      const float2 v = z[(fr >> 1) * NFFT + bitrev10(m)];
      const float w = __ldg(window + m);
      acc = fmaf(((fr & 1) ? v.y : v.x) * (1.0f / NFFT), w, acc);
      env = fmaf(w, w, env);
    }
    o[n] = acc / env;
  }
}

int istft_head(cudaStream_t st, const float* spec, long long ld_spec, int batch, int t_pad, const int* n_frames,
               const float* window, const float* tw, float* out, long long ld_out, int out_len) {
  F5E_REQUIRE(batch > 0 && t_pad > 0 && out_len >= 0, "istft_head: bad shape");
  F5E_REQUIRE(ld_spec >= 2 * NBIN, "istft_head: ld_spec=%lld must be at least %d", ld_spec, 2 * NBIN);
  if (out_len == 0) return 0;
  const int smem = (IST_FRAMES / 2) * NFFT * 8;
  static bool configured = false;
  if (!configured) {
    F5E_CUDA_CHECK(cudaFuncSetAttribute(istft_head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid((out_len + IST_HOPS * HOP - 1) / (IST_HOPS * HOP), batch);
  LaunchScope ls(st, PROF_DSP, batch, t_pad, 1);
  istft_head_kernel<<<grid, DSP_THREADS, smem, st>>>(spec, ld_spec, t_pad, n_frames, window,
                                                     reinterpret_cast<const float2*>(tw), out, ld_out, out_len);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------
// fp32 helpers of the Vocos backbone (activations stay fp32; GEMM operands are 3-term bf16 splits)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void split_store(__nv_bfloat16* o, int split, float a, float b) {
  const __nv_bfloat16 ah = __float2bfloat16(a), bh = __float2bfloat16(b);
  const uint32_t hi = pack_bf16x2(__bfloat162float(ah), __bfloat162float(bh));
  const uint32_t lo = pack_bf16x2(a - __bfloat162float(ah), b - __bfloat162float(bh));
  *reinterpret_cast<uint32_t*>(o) = hi;
  *reinterpret_cast<uint32_t*>(o + split) = hi;
  *reinterpret_cast<uint32_t*>(o + 2 * split) = lo;
}

// layernorm_f32: y = LN(x; eps) * w + b over D = 32*2*CH columns, one warp per row, fp32 in.
// Rows with (row % n_pad) >= row_len[row / n_pad] are written as zeros (sequence padding must stay exactly zero so that
// the following "same"-padded convolutions see what a per-utterance run would see).
// out_f32 (optional) receives fp32; out_split (optional) the bf16 3-term split [hi | hi | lo] with stride D.
template <int CH>
__global__ void __launch_bounds__(256)
layernorm_f32_kernel(const float* __restrict__ x, int M, const float* __restrict__ w, const float* __restrict__ b, float eps,
                     const int* __restrict__ row_len, int n_pad, float* __restrict__ out_f32,
                     __nv_bfloat16* __restrict__ out_split) {
  constexpr int D = CH * 64;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const float2* xr = reinterpret_cast<const float2*>(x + (size_t)row * D);
  float2 v[CH];
  float sum = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c) { v[c] = xr[c * 32 + lane]; sum += v[c].x + v[c].y; }
  const float mean = warp_sum(sum) * (1.0f / D);
  float sq = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c) { const float a = v[c].x - mean, d = v[c].y - mean; sq += a * a + d * d; }
  const float rstd = rsqrtf(warp_sum(sq) * (1.0f / D) + eps);
  bool masked = false;
  if (row_len) { const int s = row / n_pad; masked = (row - s * n_pad) >= __ldg(row_len + s); }
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int e = (c * 32 + lane) * 2;
    float a = (v[c].x - mean) * rstd * __ldg(w + e) + __ldg(b + e);
    float d = (v[c].y - mean) * rstd * __ldg(w + e + 1) + __ldg(b + e + 1);
    if (masked) a = d = 0.f;
    if (out_f32) *reinterpret_cast<float2*>(out_f32 + (size_t)row * D + e) = make_float2(a, d);
    if (out_split) split_store(out_split + (size_t)row * 3 * D + e, D, a, d);
  }
}

int layernorm_f32(cudaStream_t st, const float* x, int M, int D, const float* w, const float* b, float eps,
                  const int* row_len, int n_pad, float* out_f32, __nv_bfloat16* out_split) {
  F5E_REQUIRE(M > 0 && D % 64 == 0 && D >= 64 && D <= 1024, "layernorm_f32: D=%d must be a multiple of 64 in [64,1024]", D);
  if (n_pad <= 0) n_pad = M;
  const int grid = (M + 7) / 8;
  LaunchScope ls(st, PROF_NORM, M, D, 1);
#define F5E_LN(CH) \
  case CH: layernorm_f32_kernel<CH><<<grid, 256, 0, st>>>(x, M, w, b, eps, row_len, n_pad, out_f32, out_split); break;
  switch (D / 64) {
    F5E_LN(1) F5E_LN(2) F5E_LN(3) F5E_LN(4) F5E_LN(5) F5E_LN(6) F5E_LN(7) F5E_LN(8)
    F5E_LN(9) F5E_LN(10) F5E_LN(11) F5E_LN(12) F5E_LN(13) F5E_LN(14) F5E_LN(15) F5E_LN(16)
  }
#undef F5E_LN
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// dwconv7_f32: depthwise Conv1d(C, C, k=7, padding=3) + bias over frames per sample, fp32 [samples, n_pad, C].
__global__ void __launch_bounds__(256)
dwconv7_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                   float* __restrict__ out, int n_pad, int C) {
  constexpr int FR = 16;
  const int s = blockIdx.z, n0 = blockIdx.y * FR;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float wk[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) wk[k] = __ldg(w + (size_t)c * 7 + k);
  const float bias = __ldg(b + c);
  const float* xs = x + (size_t)s * n_pad * C + c;
  float win[FR + 6];
#pragma unroll
  for (int i = 0; i < FR + 6; ++i) {
    const int n = n0 + i - 3;
    win[i] = (n >= 0 && n < n_pad) ? xs[(size_t)n * C] : 0.f;
  }
#pragma unroll
  for (int i = 0; i < FR; ++i) {
    const int n = n0 + i;
    if (n >= n_pad) break;
    float a = bias;
#pragma unroll
    for (int k = 0; k < 7; ++k) a = fmaf(win[i + k], wk[k], a);
    out[((size_t)s * n_pad + n) * C + c] = a;
  }
}

int dwconv7_f32(cudaStream_t st, const float* x, const float* w, const float* b, float* out, int samples, int n_pad, int C) {
  F5E_REQUIRE(samples > 0 && n_pad > 0 && C > 0, "dwconv7_f32: bad shape");
  dim3 grid((C + 255) / 256, (n_pad + 15) / 16, samples);
  LaunchScope ls(st, PROF_ELEM, samples, n_pad, 7);
  dwconv7_f32_kernel<<<grid, 256, 0, st>>>(x, w, b, out, n_pad, C);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// unfold_mel_split3: mel [B, C, t_in] fp32 (channel-major, as vocoder.decode receives it), frames [t_off[b], t_off[b] +
// n_frames[b]) of each sample -> A [B * t_pad, 3 * kpad] bf16 with A[b, n, tap*C + c] = mel[b, c, t_off + n + tap - taps/2]
// (zero outside the span: the conv's zero padding), as the 3-term split [hi | hi | lo].  Turns Vocos' embed
// Conv1d(100, 512, k=7, padding=3) into one GEMM.
__global__ void __launch_bounds__(256)
unfold_mel_split3_kernel(const float* __restrict__ mel, long long ld_b, long long ld_c, long long ld_t,
                         const int* __restrict__ t_off, const int* __restrict__ n_frames, int t_pad, int C, int taps,
                         int kpad, __nv_bfloat16* __restrict__ out) {
  const int n = blockIdx.x, b = blockIdx.y;
  const int off = t_off ? __ldg(t_off + b) : 0;
  const int T = n_frames ? __ldg(n_frames + b) : t_pad;
  __nv_bfloat16* orow = out + ((size_t)b * t_pad + n) * 3 * kpad;
  for (int i = threadIdx.x; i < kpad; i += blockDim.x) {
    float v = 0.f;
    if (i < taps * C && n < T) {
      const int tap = i / C, c = i - tap * C;
      const int t = n + tap - taps / 2;
      if (t >= 0 && t < T) v = mel[(size_t)b * ld_b + (size_t)c * ld_c + (size_t)(off + t) * ld_t];
    }
    const __nv_bfloat16 h = __float2bfloat16(v);
    orow[i] = h;
    orow[kpad + i] = h;
    orow[2 * kpad + i] = __float2bfloat16(v - __bfloat162float(h));
  }
}

int unfold_mel_split3(cudaStream_t st, const float* mel, long long ld_b, long long ld_c, long long ld_t, const int* t_off,
                      const int* n_frames, int batch, int t_pad, int C, int taps, int kpad, __nv_bfloat16* out) {
  F5E_REQUIRE(batch > 0 && t_pad > 0 && kpad >= taps * C && kpad % 8 == 0, "unfold_mel_split3: bad shape");
  dim3 grid(t_pad, batch);
  LaunchScope ls(st, PROF_ELEM, batch, t_pad, 8);
  unfold_mel_split3_kernel<<<grid, 256, 0, st>>>(mel, ld_b, ld_c, ld_t, t_off, n_frames, t_pad, C, taps, kpad, out);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// split3: fp32 [M, K] (row stride ld) -> bf16 [M, 3K] = [hi | hi | lo]   (A operand of a bf16x3 GEMM)
__global__ void __launch_bounds__(256)
split3_kernel(const float* __restrict__ x, long long ld, int M, int K, __nv_bfloat16* __restrict__ out) {
  const long long total = (long long)M * (K / 2);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / (K / 2)), c = (int)(i - (long long)r * (K / 2)) * 2;
    const float2 v = *reinterpret_cast<const float2*>(x + (size_t)r * ld + c);
    split_store(out + (size_t)r * 3 * K + c, K, v.x, v.y);
  }
}

int split3(cudaStream_t st, const float* x, long long ld, int M, int K, __nv_bfloat16* out) {
  F5E_REQUIRE(M > 0 && K > 0 && K % 2 == 0 && ld % 2 == 0, "split3: K and ld must be even");
  const long long total = (long long)M * (K / 2);
  int grid = (int)((total + 255) / 256);
  if (grid > num_sms() * 16) grid = num_sms() * 16;
  LaunchScope ls(st, PROF_ELEM, M, K, 9);
  split3_kernel<<<grid, 256, 0, st>>>(x, ld, M, K, out);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace f5e
