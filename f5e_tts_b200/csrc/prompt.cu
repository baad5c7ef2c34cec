// Prompt preparation and chunk assembly around the sampler (fp32 / fp64 streaming kernels, HBM-bound):
//   rms_normalize : reference eval/utils_eval.py:120-123, infer/utils_infer.py:441-443
//   resample      : torchaudio.transforms.Resample as called at utils_eval.py:125-127, utils_infer.py:444-446
//                   (third-party torchaudio: sinc_interp_hann, lowpass_filter_width 6, rolloff 0.99; the polyphase
//                   filter table is built by the caller exactly like torchaudio's _get_sinc_resample_kernel)
//   crossfade     : the numpy cross-fade chain of infer/utils_infer.py:525-556 (float64 arithmetic like numpy)
#include "../../include/f5e_b200.h"
#include "host.h"

namespace f5e {

// ---------------------------------------------------------------------------------------------
// RMS: deterministic two-stage sum of squares (fixed chunking, fixed order), then an optional gain
// ---------------------------------------------------------------------------------------------
constexpr int RMS_CHUNK = 16384;  // samples per partial sum

__global__ void __launch_bounds__(256)
rms_partial_kernel(const float* __restrict__ wav, long long ld, const int* __restrict__ lens, int n_max,
                   float* __restrict__ partial, int chunks) {
  const int b = blockIdx.y, c = blockIdx.x;
  const int n = lens ? min(lens[b], n_max) : n_max;
  const int lo = c * RMS_CHUNK, hi = min(lo + RMS_CHUNK, n);
  const float* x = wav + (size_t)b * ld;
  float acc = 0.f;
  for (int i = lo + threadIdx.x; i < hi; i += 256) {
    const float v = x[i];
    acc = fmaf(v, v, acc);
  }
  acc = warp_sum(acc);
  __shared__ float ws[8];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += ws[i];
    partial[(size_t)b * chunks + c] = t;
  }
}

__global__ void __launch_bounds__(256)
rms_apply_kernel(const float* __restrict__ wav, long long ld, const int* __restrict__ lens, int n_max,
                 const float* __restrict__ partial, int chunks, float target_rms, float* __restrict__ rms_out,
                 float* __restrict__ out, long long ldo) {
  const int b = blockIdx.y;
  const int n = lens ? min(lens[b], n_max) : n_max;
  float total = 0.f;
  for (int c = 0; c < chunks; ++c) total += partial[(size_t)b * chunks + c];  // same order in every block
  const float rms = sqrtf(total / (float)max(n, 1));
  if (blockIdx.x == 0 && threadIdx.x == 0) rms_out[b] = rms;
  if (!out) return;
  // reference: `if rms < target_rms: audio = audio * target_rms / rms` (left-to-right: (audio * target) / rms)
  const bool boost = rms < target_rms;
  const float* x = wav + (size_t)b * ld;
  float* y = out + (size_t)b * ldo;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n_max; i += gridDim.x * 256) {
    const float v = i < n ? x[i] : 0.f;
    y[i] = boost ? __fdiv_rn(__fmul_rn(v, target_rms), rms) : v;
  }
}

// ---------------------------------------------------------------------------------------------
// Polyphase sinc resampler: out[q * up + p] = sum_k kern[p][k] * x[q * down + k - width]  (zero outside [0, len))
// ---------------------------------------------------------------------------------------------
template <bool TABLE_IN_SMEM>
__global__ void __launch_bounds__(256)
resample_kernel(const float* __restrict__ wav, long long ld, const int* __restrict__ lens, int n_max,
                const float* __restrict__ kern, int taps, int down, int up, int width, float* __restrict__ out,
                long long ldo, int out_max) {
  extern __shared__ float sk[];  // [up][taps]; very large tables (e.g. 44.1 -> 48 kHz) stay in global memory / L2
  if (TABLE_IN_SMEM) {
    for (int i = threadIdx.x; i < up * taps; i += 256) sk[i] = kern[i];
    __syncthreads();
  }
  const int b = blockIdx.y;
  const int n = lens ? min(lens[b], n_max) : n_max;
  // torchaudio: target_length = ceil(up * length / down)
  const long long tgt = ((long long)up * n + down - 1) / down;
  const float* x = wav + (size_t)b * ld;
  float* y = out + (size_t)b * ldo;
  for (int o = blockIdx.x * 256 + threadIdx.x; o < out_max; o += gridDim.x * 256) {
    float acc = 0.f;
    if (o < tgt) {
      const int q = o / up, p = o - q * up;
      const int base = q * down - width;
      const float* kp = (TABLE_IN_SMEM ? sk : kern) + p * taps;
      for (int k = 0; k < taps; ++k) {
        const int i = base + k;
        if (i >= 0 && i < n) acc = fmaf(kp[k], x[i], acc);
      }
    }
    y[o] = acc;
  }
}

// ---------------------------------------------------------------------------------------------
// One cross-fade step: `acc` (float64, n_acc samples so far) is extended by `next` (fp32, n_next samples):
//   cf = min(cf_samples, n_acc, n_next); acc[n_acc - cf + i] = acc[..] * fade_out[i] + next[i] * fade_in[i] (i < cf),
//   acc[n_acc + i - cf] = next[i] (i >= cf).  fade = np.linspace in float64: i * step + start, last element exact.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
crossfade_kernel(double* __restrict__ acc, long long n_acc, const float* __restrict__ next, int n_next, int cf,
                 float gain_num, float gain_den, int apply_gain) {
  const double step_in = cf > 1 ? 1.0 / (double)(cf - 1) : 0.0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n_next; i += gridDim.x * 256) {
    float v = next[i];
    if (apply_gain) v = __fdiv_rn(__fmul_rn(v, gain_num), gain_den);  // `generated_wave * rms / target_rms` (utils_infer.py:495-496)
    const long long dst = n_acc - cf + i;
    if (i < cf) {
      double fin, fout;
      if (cf == 1) { fin = 0.0; fout = 1.0; }
      else if (i == cf - 1) { fin = 1.0; fout = 0.0; }
      else { fin = __dmul_rn((double)i, step_in); fout = __dadd_rn(__dmul_rn((double)i, -step_in), 1.0); }
      acc[dst] = __dadd_rn(__dmul_rn(acc[dst], fout), __dmul_rn((double)v, fin));  // no FMA contraction: numpy rounds each op
    } else {
      acc[dst] = (double)v;
    }
  }
}

}  // namespace f5e

using namespace f5e;

extern "C" {

int f5e_rms_normalize(const float* wav, int64_t ld_wav, const int32_t* lens, int32_t batch, int32_t n_max,
                      float target_rms, float* scratch, float* rms_out, float* out, int64_t ld_out, void* stream) {
  F5E_REQUIRE(wav && scratch && rms_out && batch > 0 && n_max > 0, "f5e_rms_normalize: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const int chunks = (n_max + RMS_CHUNK - 1) / RMS_CHUNK;
  {
    LaunchScope ls(st, PROF_DSP, batch, n_max, 0);
    rms_partial_kernel<<<dim3(chunks, batch), 256, 0, st>>>(wav, ld_wav, lens, n_max, scratch, chunks);
  }
  F5E_CUDA_CHECK(cudaGetLastError());
  int gx = (n_max + 255) / 256;
  const int cap = num_sms() * 8 / (batch > 0 ? batch : 1) + 1;
  if (gx > cap) gx = cap;
  {
    LaunchScope ls(st, PROF_DSP, batch, n_max, 1);
    rms_apply_kernel<<<dim3(gx, batch), 256, 0, st>>>(wav, ld_wav, lens, n_max, scratch, chunks, target_rms, rms_out, out,
                                                      ld_out);
  }
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int f5e_rms_scratch_floats(int32_t batch, int32_t n_max) { return batch * ((n_max + RMS_CHUNK - 1) / RMS_CHUNK); }

int f5e_resample(const float* wav, int64_t ld_wav, const int32_t* lens, int32_t batch, int32_t n_max, const float* kern,
                 int32_t taps, int32_t down, int32_t up, int32_t width, float* out, int64_t ld_out, int32_t out_max,
                 void* stream) {
  F5E_REQUIRE(wav && kern && out && batch > 0 && n_max > 0 && out_max > 0, "f5e_resample: bad arguments");
  F5E_REQUIRE(taps == 2 * width + down && up > 0 && down > 0, "f5e_resample: taps must be 2 * width + down");
  const size_t smem = (size_t)up * taps * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  int gx = (out_max + 255) / 256;
  const int cap = num_sms() * 8 / batch + 1;
  if (gx > cap) gx = cap;
  LaunchScope ls(st, PROF_DSP, batch, n_max, 2);
  if (smem <= 200 * 1024) {
    if (smem > 48 * 1024)
      F5E_CUDA_CHECK(cudaFuncSetAttribute(resample_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    resample_kernel<true><<<dim3(gx, batch), 256, smem, st>>>(wav, ld_wav, lens, n_max, kern, taps, down, up, width, out,
                                                              ld_out, out_max);
  } else {
    resample_kernel<false><<<dim3(gx, batch), 256, 0, st>>>(wav, ld_wav, lens, n_max, kern, taps, down, up, width, out,
                                                            ld_out, out_max);
  }
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int f5e_crossfade_append(double* acc, int64_t n_acc, const float* next, int32_t n_next, int32_t cf_samples,
                         float gain_num, float gain_den, int32_t apply_gain, int64_t* n_out, void* stream) {
  F5E_REQUIRE(acc && n_out && n_acc >= 0 && n_next >= 0, "f5e_crossfade_append: bad arguments");
  long long cf = cf_samples;
  if (cf > n_acc) cf = n_acc;
  if (cf > n_next) cf = n_next;
  if (cf < 0) cf = 0;
  *n_out = n_acc + n_next - cf;
  if (n_next == 0) return 0;
  F5E_REQUIRE(next != nullptr, "f5e_crossfade_append: null chunk");
  cudaStream_t st = (cudaStream_t)stream;
  int gx = (n_next + 255) / 256;
  if (gx > num_sms() * 8) gx = num_sms() * 8;
  LaunchScope ls(st, PROF_DSP, n_next, cf, 3);
  crossfade_kernel<<<gx, 256, 0, st>>>(acc, n_acc, next, n_next, (int)cf, gain_num, gain_den, apply_gain);
  F5E_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
