/* core.cu - device management, error reporting, the device-resident IF stream
 * and the synthetic front end of libsdrb200.
 *
 * The IF stream replaces the reference's host ring buffer sdrstat.buff and
 * rcvgetbuff() (src/sdrrcv.c:406-463; file front end 469-532): samples keep
 * their ABSOLUTE index, `capacity` samples are retained, and the first
 * SDRB200_STREAM_GUARD bytes of the ring are mirrored behind its end so that a
 * kernel can read any window shorter than the guard with plain linear
 * (vectorised) loads even when the window wraps.
 */
#include <stdio.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include "sdrb200_internal.h"

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

extern "C" void sdrb200_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char *sdrb200_last_error(void) { return g_err; }
extern "C" const char *sdrb200_version(void) { return "sdrb200 0.1 (sm_100a)"; }
extern "C" void sdrb200_count_launch(int n) { g_launches.fetch_add((uint64_t)n); }
extern "C" uint64_t sdrb200_launch_count(void) { return g_launches.load(); }

extern "C" int sdrb200_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
extern "C" int sdrb200_ensure_device(void)
{
    if (sdrb200_device_count() < 1) {
        sdrb200_set_error("no CUDA device: libsdrb200 has no CPU fallback");
        return SDRB200_ENODEV;
    }
    return SDRB200_OK;
}
extern "C" int sdrb200_set_device(int dev)
{
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    SDRB200_CUDA_TRY(cudaSetDevice(dev));
    return SDRB200_OK;
}
extern "C" int sdrb200_sm_count(void)
{
    int dev = 0, n = 0;
    if (sdrb200_ensure_device()) return 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return n;
}

/* ------------------------------------------------------------ IF stream */
extern "C" int sdrb200_stream_create(sdrb200_stream_t **out, int dtype, uint64_t capacity)
{
    if (!out || (dtype != 1 && dtype != 2) || capacity == 0) {
        sdrb200_set_error("stream_create: bad argument"); return SDRB200_EINVAL;
    }
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    sdrb200_stream_t *s = (sdrb200_stream_t *)calloc(1, sizeof(*s));
    if (!s) return SDRB200_ENOMEM;
    s->dtype = dtype; s->capacity = capacity;
    cudaStream_t st;
    if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) { free(s); return SDRB200_ECUDA; }
    s->cuda_stream = st;
    size_t bytes = (size_t)capacity * dtype + SDRB200_STREAM_GUARD;
    cudaError_t e = cudaMalloc((void **)&s->d_buf, bytes);
    if (e != cudaSuccess) {
        sdrb200_set_error("stream_create: cudaMalloc(%zu) -> %s", bytes, cudaGetErrorString(e));
        cudaStreamDestroy(st); free(s); return SDRB200_ENOMEM;
    }
    cudaMemsetAsync(s->d_buf, 0, bytes, st);
    cudaStreamSynchronize(st);
    *out = s;
    return SDRB200_OK;
}

extern "C" void sdrb200_stream_destroy(sdrb200_stream_t *s)
{
    if (!s) return;
    cudaFree(s->d_buf);
    cudaStreamDestroy((cudaStream_t)s->cuda_stream);
    free(s);
}

extern "C" uint64_t sdrb200_stream_count(const sdrb200_stream_t *s) { return s ? s->count : 0; }

/* after bytes [off, off+len) of the ring were written, refresh the guard mirror */
static int mirror_guard(sdrb200_stream_t *s, size_t off, size_t len)
{
    size_t ring = (size_t)s->capacity * s->dtype;
    if (off >= SDRB200_STREAM_GUARD) return SDRB200_OK;
    size_t m = len; if (off + m > SDRB200_STREAM_GUARD) m = SDRB200_STREAM_GUARD - off;
    if (off + m > ring) m = ring - off;
    SDRB200_CUDA_TRY(cudaMemcpyAsync(s->d_buf + ring + off, s->d_buf + off, m,
                                     cudaMemcpyDeviceToDevice, (cudaStream_t)s->cuda_stream));
    return SDRB200_OK;
}

/* generic chunked append; src_kind 0: host int8, 1: host rtlsdr u8, 2: device int8, 3: host packed bytes */
__global__ void k_rtlsdr_expand(const uint8_t *__restrict__ in, int8_t *__restrict__ out, size_t n)
{
    /* rtlsdr_exp: (char)(buf-127.5), src/rcv/rtlsdr/rtlsdr.c:136-142.
     * float subtraction then C truncation toward zero.                          */
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[i] = (int8_t)(int)((float)in[i] - 127.5f);
}

/* ---- ranging codes on the device (gencode, src/sdrcode.c:101-154, 426-444, 523-539) ----
 * One CTA per code.  C/A: lane 0 runs G1 (x^10+x^3+1), lane 1 runs G2 (x^10+x^9+x^8+x^6+x^3+x^2+1),
 * both from all ones; chip i = G1[i] xor G2[(i - delay) mod 1023] with the reference's polarity
 * (see csrc/codegen.c).  GLONASS ST code: x^9+x^5+1, output stage 7, 511 chips.            */
__global__ void k_gencode(const int *__restrict__ prn, const unsigned short *__restrict__ delay, int ctype,
                          int8_t *__restrict__ out, int stride)
{
    __shared__ unsigned char g[2][1024];
    int8_t *o = out + (size_t)blockIdx.x * stride;
    if (ctype == SDRB200_CTYPE_G1) {
        if (threadIdx.x == 0) {
            unsigned int reg = 0x1FF;
            for (int i = 0; i < 511; i++) {
                const unsigned int fb = ((reg >> 4) ^ (reg >> 8)) & 1u;
                g[0][i] = (unsigned char)((reg >> 6) & 1u);
                reg = ((reg << 1) | fb) & 0x1FF;
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < 511; i += blockDim.x) o[i] = g[0][i] ? 1 : -1;
        return;
    }
    if (threadIdx.x < 2) {
        const unsigned int taps = threadIdx.x == 0 ? ((1u << 2) | (1u << 9))
                                                   : ((1u << 1) | (1u << 2) | (1u << 5) | (1u << 7) | (1u << 8) | (1u << 9));
        unsigned int reg = 0x3FF;
        for (int i = 0; i < 1023; i++) {
            const unsigned int fb = (unsigned int)__popc(reg & taps) & 1u;
            g[threadIdx.x][i] = (unsigned char)((reg >> 9) & 1u);
            reg = ((reg << 1) | fb) & 0x3FF;
        }
    }
    __syncthreads();
    const int d = (int)delay[prn[blockIdx.x] - 1];
    for (int i = threadIdx.x; i < 1023; i += blockDim.x) {
        int k = i + 1023 - d; if (k >= 1023) k -= 1023;
        o[i] = (g[0][i] ^ g[1][k]) ? 1 : -1;
    }
}

extern "C" int sdrb200_gencode_device(int n, const int *prn, int ctype, int8_t *codes, int stride)
{
    const int len = (ctype == SDRB200_CTYPE_G1) ? 511 : 1023;
    if (n < 1 || !prn || !codes || stride < len ||
        (ctype != SDRB200_CTYPE_L1CA && ctype != SDRB200_CTYPE_L1SBAS && ctype != SDRB200_CTYPE_G1)) {
        sdrb200_set_error("gencode_device: bad argument"); return SDRB200_EINVAL;
    }
    if (ctype != SDRB200_CTYPE_G1)
        for (int i = 0; i < n; i++)
            if (prn[i] < 1 || prn[i] > 210) { sdrb200_set_error("gencode_device: PRN %d out of range", prn[i]); return SDRB200_EINVAL; }
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    int *d_prn = NULL; unsigned short *d_delay = NULL; int8_t *d_out = NULL;
    int rc = SDRB200_OK;
    if (cudaMalloc((void **)&d_prn, sizeof(int) * n) != cudaSuccess || cudaMalloc((void **)&d_delay, sizeof(unsigned short) * 210) != cudaSuccess ||
        cudaMalloc((void **)&d_out, (size_t)n * stride) != cudaSuccess) rc = SDRB200_ENOMEM;
    if (rc == SDRB200_OK) {
        cudaMemcpy(d_prn, prn, sizeof(int) * n, cudaMemcpyHostToDevice);
        cudaMemcpy(d_delay, sdrb200_g2_delay_table(), sizeof(unsigned short) * 210, cudaMemcpyHostToDevice);
        cudaMemset(d_out, 0, (size_t)n * stride);
        k_gencode<<<n, 128>>>(d_prn, d_delay, ctype, d_out, stride);
        sdrb200_count_launch(1);
        if (cudaGetLastError() != cudaSuccess || cudaMemcpy(codes, d_out, (size_t)n * stride, cudaMemcpyDeviceToHost) != cudaSuccess) {
            sdrb200_set_error("gencode_device: kernel failed"); rc = SDRB200_ECUDA;
        }
    } else sdrb200_set_error("gencode_device: cudaMalloc failed");
    cudaFree(d_prn); cudaFree(d_delay); cudaFree(d_out);
    return rc == SDRB200_OK ? len : rc;
}


/* packed front-end formats: every input byte holds one sample (real, or I and Q).
 * The per-byte expansion tables restate the reference's unpack routines:
 *   STEREO front end 1 (MAX2769, 2 bit real)  stereo_exp DTYPEI   src/rcv/stereo/stereo.c:163-174,192-196
 *   STEREO front end 2 (MAX2112, 3+3 bit I/Q) stereo_exp DTYPEIQ  src/rcv/stereo/stereo.c:163-174,199-204
 *   GN3S v3 2 bit real                        gn3s_exp_v3 mode 32/38  src/rcv/gn3s/gn3s.cpp:143-177
 *   GN3S v3 4 bit I/Q                         gn3s_exp_v3 mode 36/42  src/rcv/gn3s/gn3s.cpp:143-177   */
struct expand_lut_t { int8_t v[256][2]; };

static int expand_lut_build(int fmt, expand_lut_t *t, int *outs)
{
    static const int8_t base1[4] = {-3, -1, +1, +3};
    static const int8_t base2[8] = {+1, +3, +5, +7, -7, -5, -3, -1};
    static const int8_t g2[4] = {1, -1, 3, -3};
    static const int8_t gi4[16] = {1, -1, 0, 0, 3, -3, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    static const int8_t gq4[16] = {1, 0, -1, 0, 0, 0, 0, 0, 3, 0, -3, 0, 0, 0, 0, 0};
    for (int r = 0; r < 256; r++) {
        switch (fmt) {
        case SDRB200_FMT_STEREO_FE1: t->v[r][0] = base1[(r >> 6) & 3]; t->v[r][1] = 0; *outs = 1; break;
        case SDRB200_FMT_STEREO_FE2: t->v[r][0] = base2[(r >> 3) & 7]; t->v[r][1] = base2[r & 7]; *outs = 2; break;
        case SDRB200_FMT_GN3S_2BIT:  t->v[r][0] = g2[r & 0x03]; t->v[r][1] = 0; *outs = 1; break;
        case SDRB200_FMT_GN3S_4BIT:  t->v[r][0] = gi4[r & 0x05]; t->v[r][1] = gq4[r & 0x0A]; *outs = 2; break;
        default: return SDRB200_EINVAL;
        }
    }
    return SDRB200_OK;
}

/* one thread expands 16 packed bytes (one 16-byte load) into 16 or 32 samples */
__global__ void k_expand_lut(const uint8_t *__restrict__ in, int8_t *__restrict__ out, size_t n, int outs,
                             const expand_lut_t lut)
{
    __shared__ int8_t sl[256][2];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) { sl[i][0] = lut.v[i][0]; sl[i][1] = lut.v[i][1]; }
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x * 16;
    for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16; i < n; i += stride) {
        const bool vec = (i + 16 <= n) && ((((size_t)(in + i)) & 15) == 0);
        uint8_t b[16];
        if (vec) *(uint4 *)b = *(const uint4 *)(in + i);
        else for (int k = 0; k < 16; k++) b[k] = (i + k < n) ? in[i + k] : 0;
        for (int k = 0; k < 16 && i + k < n; k++) {
            if (outs == 1) out[i + k] = sl[b[k]][0];
            else { out[2 * (i + k)] = sl[b[k]][0]; out[2 * (i + k) + 1] = sl[b[k]][1]; }
        }
    }
}

static int stream_push(sdrb200_stream_t *s, int kind, const void *src, uint64_t n,
                       const expand_lut_t *lut = NULL, int outs = 0)
{
    if (!s || !src) { sdrb200_set_error("stream_push: bad argument"); return SDRB200_EINVAL; }
    cudaStream_t st = (cudaStream_t)s->cuda_stream;
    const size_t bps = (size_t)s->dtype;
    uint64_t done = 0;
    while (done < n) {
        uint64_t pos = s->count % s->capacity;
        uint64_t chunk = n - done;
        if (chunk > s->capacity - pos) chunk = s->capacity - pos;
        int8_t *dst = s->d_buf + pos * bps;
        if (kind == 0) {
            SDRB200_CUDA_TRY(cudaMemcpyAsync(dst, (const int8_t *)src + done * bps, chunk * bps,
                                             cudaMemcpyHostToDevice, st));
        } else if (kind == 1) {
            uint8_t *tmp = NULL;
            SDRB200_CUDA_TRY(cudaMallocAsync((void **)&tmp, chunk * bps, st));
            SDRB200_CUDA_TRY(cudaMemcpyAsync(tmp, (const uint8_t *)src + done * bps, chunk * bps,
                                             cudaMemcpyHostToDevice, st));
            k_rtlsdr_expand<<<148 * 4, 256, 0, st>>>(tmp, dst, chunk * bps);
            sdrb200_count_launch(1);
            SDRB200_CUDA_TRY(cudaGetLastError());
            SDRB200_CUDA_TRY(cudaFreeAsync(tmp, st));
        } else if (kind == 3) {
            uint8_t *tmp = NULL;                         /* one packed byte per sample */
            SDRB200_CUDA_TRY(cudaMallocAsync((void **)&tmp, chunk, st));
            SDRB200_CUDA_TRY(cudaMemcpyAsync(tmp, (const uint8_t *)src + done, chunk, cudaMemcpyHostToDevice, st));
            k_expand_lut<<<148 * 4, 256, 0, st>>>(tmp, dst, chunk, outs, *lut);
            sdrb200_count_launch(1);
            SDRB200_CUDA_TRY(cudaGetLastError());
            SDRB200_CUDA_TRY(cudaFreeAsync(tmp, st));
        } else {
            SDRB200_CUDA_TRY(cudaMemcpyAsync(dst, (const int8_t *)src + done * bps, chunk * bps,
                                             cudaMemcpyDeviceToDevice, st));
        }
        int rc = mirror_guard(s, pos * bps, chunk * bps);
        if (rc) return rc;
        s->count += chunk;
        done += chunk;
    }
    SDRB200_CUDA_TRY(cudaStreamSynchronize(st));
    return SDRB200_OK;
}

extern "C" int sdrb200_stream_push_host(sdrb200_stream_t *s, const int8_t *host, uint64_t n)
{
    return stream_push(s, 0, host, n);
}
extern "C" int sdrb200_stream_push_host_packed(sdrb200_stream_t *s, const uint8_t *host, uint64_t n, int format)
{
    expand_lut_t lut;
    int outs = 0;
    if (!s || !host) { sdrb200_set_error("stream_push_packed: bad argument"); return SDRB200_EINVAL; }
    if (format == SDRB200_FMT_RTLSDR) return sdrb200_stream_push_host_rtlsdr(s, host, n);
    if (expand_lut_build(format, &lut, &outs)) { sdrb200_set_error("stream_push_packed: unknown format %d", format); return SDRB200_EINVAL; }
    if (s->dtype != outs) { sdrb200_set_error("stream_push_packed: format %d yields dtype %d samples", format, outs); return SDRB200_EINVAL; }
    return stream_push(s, 3, host, n, &lut, outs);
}

extern "C" int sdrb200_stream_push_host_rtlsdr(sdrb200_stream_t *s, const uint8_t *host, uint64_t n)
{
    if (s && s->dtype != 2) { sdrb200_set_error("rtlsdr samples are I/Q (dtype 2)"); return SDRB200_EINVAL; }
    return stream_push(s, 1, host, n);
}
/* IF that is already in device memory (another kernel produced it, or a DMA engine wrote it):
 * appended with a device-to-device copy on the stream's own CUDA stream                   */
extern "C" int sdrb200_stream_push_device(sdrb200_stream_t *s, const int8_t *d_src, uint64_t n)
{
    return stream_push(s, 2, d_src, n);
}

/* advance the absolute sample index by n without storing samples (their content is undefined):
 * lets a mirror of a long-running host ring start at the first sample anybody will ask for   */
extern "C" int sdrb200_stream_skip(sdrb200_stream_t *s, uint64_t n)
{
    if (!s) return SDRB200_EINVAL;
    s->count += n;
    return SDRB200_OK;
}

extern "C" int sdrb200_stream_read(const sdrb200_stream_t *s, uint64_t buffloc, uint64_t n, int8_t *host)
{
    if (!s || !host) return SDRB200_EINVAL;
    if (buffloc + n > s->count || s->count - buffloc > s->capacity) {
        sdrb200_set_error("stream_read: [%llu,+%llu) not held (count %llu, capacity %llu)",
                          (unsigned long long)buffloc, (unsigned long long)n,
                          (unsigned long long)s->count, (unsigned long long)s->capacity);
        return SDRB200_ERANGE;
    }
    const size_t bps = (size_t)s->dtype;
    uint64_t done = 0;
    while (done < n) {
        uint64_t pos = (buffloc + done) % s->capacity;
        uint64_t chunk = n - done;
        if (chunk > s->capacity - pos) chunk = s->capacity - pos;
        SDRB200_CUDA_TRY(cudaMemcpy(host + done * bps, s->d_buf + pos * bps, chunk * bps, cudaMemcpyDeviceToHost));
        done += chunk;
    }
    return SDRB200_OK;
}

/* ------------------------------------------------- measurement helpers
 * CUDA events live in THIS library's runtime instance (cudart is linked
 * statically), so bench.py times the library's own streams through these
 * instead of torch.cuda.Event, which only sees torch's current stream.        */
extern "C" int sdrb200_evt_create(void **evt)
{
    if (!evt) return SDRB200_EINVAL;
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    cudaEvent_t e;
    SDRB200_CUDA_TRY(cudaEventCreate(&e));
    *evt = (void *)e;
    return SDRB200_OK;
}
extern "C" void sdrb200_evt_destroy(void *evt) { if (evt) cudaEventDestroy((cudaEvent_t)evt); }
extern "C" int sdrb200_evt_record(void *evt, void *cuda_stream)
{
    if (!evt) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaEventRecord((cudaEvent_t)evt, (cudaStream_t)cuda_stream));
    return SDRB200_OK;
}
extern "C" int sdrb200_evt_elapsed_ms(void *start, void *stop, float *ms)
{
    if (!start || !stop || !ms) return SDRB200_EINVAL;
    SDRB200_CUDA_TRY(cudaEventSynchronize((cudaEvent_t)stop));
    SDRB200_CUDA_TRY(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return SDRB200_OK;
}
extern "C" int sdrb200_device_sync(void)
{
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    SDRB200_CUDA_TRY(cudaDeviceSynchronize());
    return SDRB200_OK;
}

/* overwrite a buffer larger than the 126 MB L2 so that the next timed
 * iteration starts from HBM                                                   */
static void *g_flush_buf = NULL;
static const size_t k_flush_bytes = (size_t)384 << 20;
extern "C" int sdrb200_l2_flush(void *cuda_stream)
{
    if (sdrb200_ensure_device()) return SDRB200_ENODEV;
    if (!g_flush_buf) SDRB200_CUDA_TRY(cudaMalloc(&g_flush_buf, k_flush_bytes));
    static int v = 0;
    SDRB200_CUDA_TRY(cudaMemsetAsync(g_flush_buf, ++v & 0xff, k_flush_bytes, (cudaStream_t)cuda_stream));
    return SDRB200_OK;
}

