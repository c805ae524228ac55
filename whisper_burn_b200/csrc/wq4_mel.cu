// wq4_mel.cu -- batched log-mel front end on the GPU ("batched transcribe front-end (GPU log-mel)").
//
// Replaces MelSpectrogram::compute_log (src/audio/mel.rs:105-157: reflect-padded 400-point STFT with a periodic Hann
// window, hop 160, |X|^2, Slaney mel filterbank without area normalisation, log10(max(.,1e-10)), clamp to global max - 8,
// (.+4)/4) together with the framing of src/transcribe.rs:46-82 (pad / truncate the clip to 480 000 samples, transpose
// to [n_mels, frames], 3000 frames), for a batch of clips, so that the host-side single-threaded Vec<Vec<f32>> code is
// not what feeds eight GPUs.
//
// The 400-point DFT is done directly (201 bins x 400 taps per frame, 0.97 GFLOP per clip: nothing next to the encoder)
// with a 400-entry twiddle table in shared memory; one CTA transforms 8 consecutive frames so every twiddle is used 8
// times.  HBM traffic: 1.9 MB of samples in, 1.5 MB of mel out per clip (written twice: log values, then clamp+scale).
#include <cmath>
#include <vector>

#include "wq4_common.cuh"

struct wq4_logmel {
    int n_mels = 0;
    int device = 0;
    float *fb = nullptr;        // [n_mels][201]
    float *hann = nullptr;      // [400]
    float2 *tw = nullptr;       // [400]  (cos, -sin)(2 pi j / 400)
    int *maxkey = nullptr;      // [kMaxBatch] ordered-int keys of the per-clip maximum
};

namespace {

constexpr int kNfft = 400, kHop = 160, kBins = 201, kFramesPerCta = 8, kMelThreads = 256;
constexpr int kChunk = 480000, kFrames = 3000, kMaxMelBatch = 4096;

// Slaney scale, float32 arithmetic as in mel.rs:242-267
float hz_to_mel(float f)
{
    const float f_sp = 200.0f / 3.0f, min_log_hz = 1000.0f, min_log_mel = min_log_hz / f_sp, logstep = 0.06875174f;
    return f < min_log_hz ? f / f_sp : min_log_mel + logf(f / min_log_hz) / logstep;
}
float mel_to_hz(float m)
{
    const float f_sp = 200.0f / 3.0f, min_log_hz = 1000.0f, min_log_mel = min_log_hz / f_sp, logstep = 0.06875174f;
    return m < min_log_mel ? m * f_sp : min_log_hz * expf((m - min_log_mel) * logstep);
}

__device__ __forceinline__ int float_key(float f)
{
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__global__ void __launch_bounds__(kMelThreads)
logmel_frames_kernel(const float *__restrict__ pcm, long long stride, long long n_samples_all, const long long *__restrict__ lens,
                     const float *__restrict__ fb, const float *__restrict__ hann, const float2 *__restrict__ tw_g, int n_mels,
                     float *__restrict__ mel, int *__restrict__ maxkey)
{
    __shared__ float xw[kFramesPerCta][kNfft];
    __shared__ float2 tw[kNfft];
    __shared__ float pw[kFramesPerCta][kBins + 3];
    const int tid = threadIdx.x, b = blockIdx.y, frame0 = blockIdx.x * kFramesPerCta;
    const float *clip = pcm + (size_t)b * stride;
    const long long n_samples = lens ? min(lens[b], stride) : n_samples_all;     // ragged batch: clip b holds lens[b] samples
    for (int i = tid; i < kNfft; i += kMelThreads) tw[i] = tw_g[i];
    for (int i = tid; i < kFramesPerCta * kNfft; i += kMelThreads) {
        const int f = i / kNfft, j = i - f * kNfft;
        // index into the 30 s chunk with the reflect padding of mel.rs:180-193 (n_fft/2 on both sides)
        int s = (frame0 + f) * kHop + j - kNfft / 2;
        if (s < 0) s = -s;
        else if (s >= kChunk) s = 2 * kChunk - 2 - s;
        const float v = (frame0 + f < kFrames && s < n_samples) ? clip[s] : 0.0f;   // zero padding to 480 000 samples
        xw[f][j] = v * hann[j];
    }
    __syncthreads();
    if (tid < kBins) {
        float re[kFramesPerCta], im[kFramesPerCta];
#pragma unroll
        for (int f = 0; f < kFramesPerCta; f++) re[f] = im[f] = 0.0f;
        int idx = 0;
        for (int n = 0; n < kNfft; n++) {
            const float2 w = tw[idx];
#pragma unroll
            for (int f = 0; f < kFramesPerCta; f++) {
                const float x = xw[f][n];
                re[f] = fmaf(x, w.x, re[f]);
                im[f] = fmaf(x, w.y, im[f]);
            }
            idx += tid;
            if (idx >= kNfft) idx -= kNfft;
        }
#pragma unroll
        for (int f = 0; f < kFramesPerCta; f++) pw[f][tid] = re[f] * re[f] + im[f] * im[f];
    }
    __syncthreads();
    float vmax = -INFINITY;
    for (int m = tid; m < n_mels; m += kMelThreads) {
        float acc[kFramesPerCta];
#pragma unroll
        for (int f = 0; f < kFramesPerCta; f++) acc[f] = 0.0f;
        const float *row = fb + (size_t)m * kBins;
        for (int j = 0; j < kBins; j++) {                 // same summation order as mel.rs:235
            const float c = row[j];
#pragma unroll
            for (int f = 0; f < kFramesPerCta; f++) acc[f] += c * pw[f][j];
        }
        float *dst = mel + ((size_t)b * n_mels + m) * kFrames + frame0;
#pragma unroll
        for (int f = 0; f < kFramesPerCta; f++) {
            if (frame0 + f < kFrames) {
                const float v = log10f(fmaxf(acc[f], 1e-10f));
                dst[f] = v;
                vmax = fmaxf(vmax, v);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
    if ((tid & 31) == 0 && vmax > -INFINITY) atomicMax(maxkey + b, float_key(vmax));
}

__global__ void logmel_finish_kernel(float *__restrict__ mel, const int *__restrict__ maxkey, long long per_clip, long long total)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const float floor_v = key_float(maxkey[i / per_clip]) - 8.0f;
        mel[i] = (fmaxf(mel[i], floor_v) + 4.0f) / 4.0f;          // mel.rs:138-151
    }
}

}  // namespace

extern "C" WQ4_API int32_t wq4_logmel_create(int32_t n_mels, int32_t device, void *stream, wq4_logmel **out)
{
    if (!out || n_mels < 1 || n_mels > 512) { wq4_set_error("logmel: n_mels must be in 1..512"); return WQ4_ERR_INVALID; }
    *out = nullptr;
    DeviceGuard g(device);
    if (!g.ok) return wq4_cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice");
    // mel.rs:270-312: n_mels + 2 points equally spaced on the mel scale, triangles between neighbours, closed intervals
    std::vector<float> hz(n_mels + 2), fb((size_t)n_mels * kBins, 0.0f), hann(kNfft);
    const float mel_min = hz_to_mel(0.0f), mel_max = hz_to_mel(16000.0f / 2.0f);
    for (int i = 0; i < n_mels + 2; i++) hz[i] = mel_to_hz(mel_min + (mel_max - mel_min) * (float)i / (float)(n_mels + 1));
    for (int i = 0; i < n_mels; i++) {
        const float lo = hz[i], ce = hz[i + 1], up = hz[i + 2];
        for (int j = 0; j < kBins; j++) {
            const float fr = (float)j * 16000.0f / (float)kNfft;
            if (fr >= lo && fr <= ce && ce > lo) fb[(size_t)i * kBins + j] = (fr - lo) / (ce - lo);
            else if (fr > ce && fr <= up && up > ce) fb[(size_t)i * kBins + j] = (up - fr) / (up - ce);
        }
    }
    const float pi_f = 3.14159265358979323846f;
    for (int i = 0; i < kNfft; i++) hann[i] = 0.5f * (1.0f - cosf(2.0f * pi_f * (float)i / (float)kNfft));   // mel.rs:315-319
    std::vector<float2> tw(kNfft);
    for (int i = 0; i < kNfft; i++) {
        const double a = 2.0 * 3.14159265358979323846 * (double)i / (double)kNfft;
        tw[i] = make_float2((float)cos(a), (float)-sin(a));
    }
    wq4_logmel *m = new wq4_logmel();
    m->n_mels = n_mels;
    m->device = device;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    cudaError_t e = cudaMalloc(&m->fb, fb.size() * 4);
    if (e == cudaSuccess) e = cudaMalloc(&m->hann, kNfft * 4);
    if (e == cudaSuccess) e = cudaMalloc(&m->tw, kNfft * sizeof(float2));
    if (e == cudaSuccess) e = cudaMalloc(&m->maxkey, kMaxMelBatch * sizeof(int));
    if (e == cudaSuccess) e = cudaMemcpyAsync(m->fb, fb.data(), fb.size() * 4, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(m->hann, hann.data(), kNfft * 4, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(m->tw, tw.data(), kNfft * sizeof(float2), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);      // the host vectors go out of scope
    if (e != cudaSuccess) {
        cudaFree(m->fb); cudaFree(m->hann); cudaFree(m->tw); cudaFree(m->maxkey);
        delete m;
        return wq4_cuda_fail(e, "logmel_create");
    }
    *out = m;
    return WQ4_OK;
}

extern "C" WQ4_API int32_t wq4_logmel_destroy(wq4_logmel *m)
{
    if (!m) return WQ4_OK;
    DeviceGuard g(m->device);
    cudaFree(m->fb); cudaFree(m->hann); cudaFree(m->tw); cudaFree(m->maxkey);
    delete m;
    return WQ4_OK;
}

static int32_t logmel_run(const wq4_logmel *m, const float *pcm_dev, int64_t batch, int64_t n_samples, const long long *lens_dev,
                          float *mel_dev, void *stream);

extern "C" WQ4_API int32_t wq4_logmel_forward(const wq4_logmel *m, const float *pcm_dev, int64_t batch, int64_t n_samples,
                                              float *mel_dev, void *stream)
{
    return logmel_run(m, pcm_dev, batch, n_samples, nullptr, mel_dev, stream);
}

// Ragged batch: clip b occupies pcm_dev[b*stride .. b*stride + lens_dev[b]) (lens_dev: device int64 [batch], each <= stride).
extern "C" WQ4_API int32_t wq4_logmel_forward_ragged(const wq4_logmel *m, const float *pcm_dev, int64_t batch, int64_t stride,
                                                     const int64_t *lens_dev, float *mel_dev, void *stream)
{
    if (!lens_dev) { wq4_set_error("logmel_forward_ragged: lens_dev is null"); return WQ4_ERR_INVALID; }
    return logmel_run(m, pcm_dev, batch, stride, reinterpret_cast<const long long *>(lens_dev), mel_dev, stream);
}

static int32_t logmel_run(const wq4_logmel *m, const float *pcm_dev, int64_t batch, int64_t n_samples, const long long *lens_dev,
                          float *mel_dev, void *stream)
{
    if (!m || !pcm_dev || !mel_dev || batch < 1 || n_samples < 1) { wq4_set_error("logmel_forward: null pointer or empty input"); return WQ4_ERR_INVALID; }
    if (batch > kMaxMelBatch) { wq4_set_error("logmel_forward: batch %lld > %d", (long long)batch, kMaxMelBatch); return WQ4_ERR_UNSUPPORTED; }
    DeviceGuard g(m->device);
    if (!g.ok) return wq4_cuda_fail(cudaErrorInvalidDevice, "cudaSetDevice");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    WQ4_CUDA(cudaMemsetAsync(m->maxkey, 0x80, (size_t)batch * sizeof(int), st));      // 0x80808080: below every key
    dim3 grid((kFrames + kFramesPerCta - 1) / kFramesPerCta, (unsigned)batch);
    logmel_frames_kernel<<<grid, kMelThreads, 0, st>>>(pcm_dev, (long long)n_samples, (long long)n_samples, lens_dev, m->fb, m->hann, m->tw,
                                                       m->n_mels, mel_dev, m->maxkey);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    const long long per_clip = (long long)m->n_mels * kFrames, total = per_clip * batch;
    logmel_finish_kernel<<<(unsigned)std::min<long long>((total + 255) / 256, 148 * 16), 256, 0, st>>>(mel_dev, m->maxkey, per_clip, total);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}
