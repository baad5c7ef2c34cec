// Kaldi filterbank front end of the PPG extractor (fp32, HBM-bound: 640 B of wave in, 320 B of features out per frame).
//   reference: ppg/wenet/dataset/feats.py:49-83 (kaldiFbank: wave * 2^15 -> torchaudio.compliance.kaldi.fbank with
//   num_mel_bins 80, 25 ms / 10 ms frames, dither 0, energy_floor 0, everything else default), followed by GlobalCMVN
//   (ppg/wenet/transformer/cmvn.py:9-47, applied at encoder.py:171-172).  torchaudio is a third-party dependency: the
//   algorithm restated here is its published kaldi.fbank (snip_edges, remove_dc_offset, preemphasis 0.97, povey window,
//   zero pad 400 -> 512, |rFFT|^2, kaldi mel banks 20 Hz .. Nyquist, log(max(., FLT_EPSILON))).
// One CTA produces 8 frames: two REAL frames are packed into one complex 512-point FFT held in shared memory, so the
// wave is read once from HBM (the 240-sample overlap of neighbouring frames is served by L1/L2).
#include <cfloat>

#include "../../include/f5e_b200.h"
#include "host.h"

namespace f5e {

constexpr int FB_NFFT = 512;
constexpr int FB_NBIN = FB_NFFT / 2 + 1;  // 257
constexpr int FB_FR = 8;                  // frames per CTA
constexpr int FB_THREADS = 256;
constexpr int FB_MAX_WIN = 512;

__device__ __forceinline__ int bitrev9(int k) { return (int)(__brev((unsigned)k) >> 23); }

// In-place radix-2 DIF FFT of `nfft` independent 512-point complex sequences in shared memory (z[f * 512 + n]);
// natural-order input, bit-reversed output.  tw[j] = exp(-2 pi i j / 512), j < 256.
__device__ __forceinline__ void fft512_smem(float2* z, const float2* __restrict__ tw, int nfft) {
  for (int s = 0; s < 9; ++s) {
    const int half = 256 >> s;
    for (int j = threadIdx.x; j < nfft * 256; j += blockDim.x) {
      const int f = j >> 8, b = j & 255;
      const int pos = b & (half - 1);
      const int i0 = f * FB_NFFT + ((b - pos) << 1) + pos, i1 = i0 + half;
      const float2 a = z[i0], c = z[i1];
      const float2 w = __ldg(tw + (pos << s));
      const float dx = a.x - c.x, dy = a.y - c.y;
      z[i0] = make_float2(a.x + c.x, a.y + c.y);
      z[i1] = make_float2(dx * w.x - dy * w.y, dx * w.y + dy * w.x);
    }
    __syncthreads();
  }
}

// wav [B, n_max] fp32 (row stride ld_wav; lens[b] valid samples, NULL: n_max) -> out [B, t_max, n_mels] fp32.
// Frames t >= 1 + (len - win) / shift of an utterance are written as zeros.
__global__ void __launch_bounds__(FB_THREADS)
kaldi_fbank_kernel(const float* __restrict__ wav, long long ld_wav, const int* __restrict__ lens, int n_max, float scale,
                   float preemph, int win, int shift, const float* __restrict__ window, const float2* __restrict__ tw,
                   const float* __restrict__ fbT, const int* __restrict__ fb_lo, const int* __restrict__ fb_hi, int n_mels,
                   const float* __restrict__ cmvn_mean, const float* __restrict__ cmvn_istd, float* __restrict__ out,
                   long long ld_out, int t_max) {
  __shared__ __align__(16) float2 z[(FB_FR / 2) * FB_NFFT];  // 16 KB
  __shared__ float raw[FB_FR][FB_MAX_WIN];                   // 16 KB: scaled frames, later the power spectra [8][257]
  __shared__ float mu[FB_FR];
  const int b = blockIdx.y, t0 = blockIdx.x * FB_FR;
  const int n = lens ? min(lens[b], n_max) : n_max;
  const int T = n >= win ? 1 + (n - win) / shift : 0;
  const float* x = wav + (size_t)b * ld_wav;
  // 1. frames (snip_edges: frame t = samples [t * shift, t * shift + win)), scaled like `wavform = x[i] * (1 << 15)`
  for (int i = threadIdx.x; i < FB_FR * win; i += blockDim.x) {
    const int f = i / win, k = i - f * win;
    const int t = t0 + f;
    raw[f][k] = t < T ? x[(size_t)t * shift + k] * scale : 0.f;
  }
  __syncthreads();
  // 2. DC offset per frame: one warp per frame, fixed order
  {
    const int f = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float s = 0.f;
    for (int k = lane; k < win; k += 32) s += raw[f][k];
    s = warp_sum(s);
    if (lane == 0) mu[f] = s / (float)win;
  }
  __syncthreads();
  // 3. (x - mu), pre-emphasis against the replicated left neighbour, povey window, zero pad; two frames per FFT
  for (int i = threadIdx.x; i < (FB_FR / 2) * FB_NFFT; i += blockDim.x) {
    const int f2 = i >> 9, k = i & (FB_NFFT - 1);
    float v[2] = {0.f, 0.f};
    if (k < win) {
      const float w = __ldg(window + k);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int f = 2 * f2 + h;
        const float m = mu[f];
        const float cur = __fsub_rn(raw[f][k], m), prev = __fsub_rn(raw[f][k > 0 ? k - 1 : 0], m);
        v[h] = __fmul_rn(__fsub_rn(cur, __fmul_rn(preemph, prev)), w);
      }
    }
    z[i] = make_float2(v[0], v[1]);
  }
  __syncthreads();
  fft512_smem(z, tw, FB_FR / 2);
  // 4. split the packed pairs, power spectrum into `raw` (reused as [8][257])
  float* pw = &raw[0][0];
  for (int i = threadIdx.x; i < (FB_FR / 2) * FB_NBIN; i += blockDim.x) {
    const int f2 = i / FB_NBIN, k = i - f2 * FB_NBIN;
    const float2 p = z[f2 * FB_NFFT + bitrev9(k & (FB_NFFT - 1))], q = z[f2 * FB_NFFT + bitrev9((FB_NFFT - k) & (FB_NFFT - 1))];
    const float ar = 0.5f * (p.x + q.x), ai = 0.5f * (p.y - q.y);
    const float br = 0.5f * (p.y + q.y), bi = 0.5f * (q.x - p.x);
    pw[(2 * f2) * FB_NBIN + k] = ar * ar + ai * ai;
    pw[(2 * f2 + 1) * FB_NBIN + k] = br * br + bi * bi;
  }
  __syncthreads();
  // 5. mel banks + log + optional global CMVN; thread -> (frame, mel) with the mel index fastest (coalesced rows)
  for (int i = threadIdx.x; i < FB_FR * n_mels; i += blockDim.x) {
    const int f = i / n_mels, m = i - f * n_mels;
    const int t = t0 + f;
    if (t >= t_max) continue;
    float r = 0.f;
    if (t < T) {
      const float* fp = pw + f * FB_NBIN;
      const float* frow = fbT + (size_t)m * FB_NBIN;
      float acc = 0.f;
      for (int k = __ldg(fb_lo + m); k < __ldg(fb_hi + m); ++k) acc = fmaf(fp[k], __ldg(frow + k), acc);
      r = logf(fmaxf(acc, FLT_EPSILON));
      if (cmvn_mean) r = __fmul_rn(__fsub_rn(r, __ldg(cmvn_mean + m)), __ldg(cmvn_istd + m));
    }
    out[(size_t)b * ld_out + (size_t)t * n_mels + m] = r;
  }
}

}  // namespace f5e

using namespace f5e;

extern "C" int f5e_kaldi_fbank(const float* wav, int64_t ld_wav, const int32_t* lens, int32_t batch, int32_t n_max,
                               float scale, float preemph, int32_t win, int32_t shift, const float* window,
                               const float* twiddles, const float* fb_t, const int32_t* fb_lo, const int32_t* fb_hi,
                               int32_t n_mels, const float* cmvn_mean, const float* cmvn_istd, float* out, int64_t ld_out,
                               int32_t t_max, void* stream) {
  F5E_REQUIRE(wav && window && twiddles && fb_t && fb_lo && fb_hi && out, "f5e_kaldi_fbank: null argument");
  F5E_REQUIRE(batch > 0 && n_max > 0 && t_max > 0 && n_mels > 0, "f5e_kaldi_fbank: bad sizes");
  F5E_REQUIRE(win > 0 && win <= FB_MAX_WIN && shift > 0, "f5e_kaldi_fbank: window must fit the 512-point FFT (win=%d)", win);
  F5E_REQUIRE((cmvn_mean == nullptr) == (cmvn_istd == nullptr), "f5e_kaldi_fbank: cmvn mean and istd go together");
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((t_max + FB_FR - 1) / FB_FR, batch);
  LaunchScope ls(st, PROF_DSP, batch, n_max, 4);
  kaldi_fbank_kernel<<<grid, FB_THREADS, 0, st>>>(wav, ld_wav, lens, n_max, scale, preemph, win, shift, window,
                                                  reinterpret_cast<const float2*>(twiddles), fb_t, fb_lo, fb_hi, n_mels,
                                                  cmvn_mean, cmvn_istd, out, ld_out, t_max);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}
