/* testkit.cu -> libsdrb200_testkit.so: TEST AND BENCH SCAFFOLDING, not product.
 *
 *   sdrb200tk_ifgen_device / sdrb200tk_stream_push_ifgen  the synthetic IF generator (ifgen.h) as a
 *       CUDA kernel, so that a 30 s / 491 MS stream is made in HBM in milliseconds and appended to a
 *       product IF stream through the public sdrb200_stream_push_device()
 *   sdrb200tk_alu_peak      issue-rate micro-benchmarks for the ALU rooflines quoted beside the HBM one
 *   sdrb200tk_nco_selftest  the device build of the NCO walker (csrc/nco_exact.h) against its host build
 * The CPU side of the generator lives in oracle/libifgen.so (oracle/ifgen_host.c).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cuda_runtime.h>
#include "ifgen.h"
#include "../../include/sdrb200.h"
#include "../../erlangnetwork-gnsslib-sdr_b200/csrc/nco_exact.h"


#define TK_TRY(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { \
    fprintf(stderr, "testkit: %s -> %s\n", #expr, cudaGetErrorString(e__)); return -4; } } while (0)

/* this library links its own copy of the CUDA runtime: its current device is set separately */
extern "C" int sdrb200tk_set_device(int dev)
{
    TK_TRY(cudaSetDevice(dev));
    return 0;
}

__global__ void k_ifgen(ifgen_cfg_t cfg, const signed char *__restrict__ codes, uint64_t k0, uint64_t n,
                        signed char *__restrict__ out)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        signed char v[2];
        ifgen_sample(&cfg, codes, k0 + i, v);
        if (cfg.dtype == 2) { out[2 * i] = v[0]; out[2 * i + 1] = v[1]; }
        else out[i] = v[0];
    }
}

static int ifgen_codes(const ifgen_cfg_t *c, signed char *codes /*[MAXSAT][MAXCLEN]*/)
{
    memset(codes, 0, (size_t)IFGEN_MAXSAT * IFGEN_MAXCLEN);
    if (c->nsat < 0 || c->nsat > IFGEN_MAXSAT || c->clen < 1 || c->clen > IFGEN_MAXCLEN) return -1;
    for (int s = 0; s < c->nsat; s++) {
        int8_t tmp[IFGEN_MAXCLEN];
        double cr;
        int len = sdrb200_gencode(c->sat[s].prn, (c->clen == 511) ? SDRB200_CTYPE_G1 : SDRB200_CTYPE_L1CA, tmp, &cr);
        if (len != c->clen) return -1;
        memcpy(codes + (size_t)s * IFGEN_MAXCLEN, tmp, (size_t)len);
    }
    return 0;
}

/* n samples from index k0 into DEVICE memory d_out (n * dtype bytes), on the legacy default stream */
extern "C" int sdrb200tk_ifgen_device(const void *vcfg, uint64_t k0, uint64_t n, void *d_out)
{
    if (!vcfg || !d_out) return -1;
    ifgen_cfg_t c = *(const ifgen_cfg_t *)vcfg;
    static signed char codes[IFGEN_MAXSAT * IFGEN_MAXCLEN];
    if (ifgen_codes(&c, codes)) return -1;
    ifgen_prepare(&c);
    signed char *d_codes = NULL;
    TK_TRY(cudaMalloc((void **)&d_codes, sizeof(codes)));
    TK_TRY(cudaMemcpy(d_codes, codes, sizeof(codes), cudaMemcpyHostToDevice));
    int blocks = (int)((n + 255) / 256); if (blocks > 148 * 16) blocks = 148 * 16; if (blocks < 1) blocks = 1;
    k_ifgen<<<blocks, 256>>>(c, d_codes, k0, n, (signed char *)d_out);
    TK_TRY(cudaGetLastError());
    TK_TRY(cudaDeviceSynchronize());
    cudaFree(d_codes);
    return 0;
}

/* append samples [k0, k0 + n) of the generator to a product IF stream */
extern "C" int sdrb200tk_stream_push_ifgen(sdrb200_stream_t *s, const void *vcfg, uint64_t k0, uint64_t n)
{
    if (!s || !vcfg) return -1;
    const int dtype = ((const ifgen_cfg_t *)vcfg)->dtype;
    const uint64_t chunk_max = 64ull << 20;
    int8_t *tmp = NULL;
    TK_TRY(cudaMalloc((void **)&tmp, (size_t)(n < chunk_max ? n : chunk_max) * dtype));
    int rc = 0;
    for (uint64_t done = 0; done < n && !rc; ) {
        const uint64_t chunk = (n - done < chunk_max) ? n - done : chunk_max;
        rc = sdrb200tk_ifgen_device(vcfg, k0 + done, chunk, tmp);
        if (!rc) rc = sdrb200_stream_push_device(s, tmp, chunk);
        done += chunk;
    }
    cudaFree(tmp);
    return rc;
}

/* issue-rate micro-benchmarks for the ALU rooflines of DESIGN.md: 8 independent
 * chains per thread of fp32 FMA (kind 0) or int32 multiply-add (kind 1), every
 * SM filled; returns operations per second counting one FMA / one MAD as 2 ops */
template <int KIND>
__global__ void __launch_bounds__(256) k_alu_peak(float *out, int iters, float a, int ia)
{
    float f[8]; int n[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { f[i] = (float)(threadIdx.x + i); n[i] = threadIdx.x + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 16; r++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (KIND == 0) f[i] = fmaf(f[i], a, 1.0f);
                else n[i] = n[i] * ia + 1;
            }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += (KIND == 0) ? f[i] : (float)n[i];
    if (s == 12345.678f) out[0] = s;                     /* keeps the chains alive */
}

extern "C" int sdrb200tk_alu_peak(int kind, double *ops_per_s)
{
    if (!ops_per_s || (kind != 0 && kind != 1)) return -1;
    cudaDeviceProp pr; int dev = 0;
    TK_TRY(cudaGetDevice(&dev));
    TK_TRY(cudaGetDeviceProperties(&pr, dev));
    float *d = NULL;
    TK_TRY(cudaMalloc(&d, 64));
    cudaEvent_t e0, e1;
    TK_TRY(cudaEventCreate(&e0)); TK_TRY(cudaEventCreate(&e1));
    const int grid = pr.multiProcessorCount * 8, iters = 1024;
    double best = 0.0;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0, 0);
        if (kind == 0) k_alu_peak<0><<<grid, 256>>>(d, iters, 0.999f, 3);
        else k_alu_peak<1><<<grid, 256>>>(d, iters, 0.999f, 3);
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
        const double ops = 2.0 * (double)grid * 256.0 * (double)iters * 16.0 * 8.0 / (ms * 1e-3);
        if (rep > 0 && ops > best) best = ops;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    TK_TRY(cudaGetLastError());
    *ops_per_s = best;
    return 0;
}
#define TRK_MAXSEG 96
/* ---- self-test of the device NCO walk (tests/test_gpu_correlator.py) ----
 * n random (start, step, count, wrap) cases, one CTA each: the device build of the walker the
 * tracking kernel's NCO lanes run (nco_walk_l = nco_walk_raw + nco_raw_to_iseg) against the host
 * walker nco_walk_i (both fuzzed against the literal loops by tests/csrc/nco_check.c):
 * every truncated value and the final accumulator must agree.                       */
__global__ void k_nco_selftest(const double *x, const double *step, const int *count, const double *wrap,
                               nco_iseg_t *segs, int *ns, double *xe)
{
    const int c = blockIdx.x;
    __shared__ nco_steptab_t tab;
    double e;
    if (threadIdx.x != 0) return;
    nco_steptab_build(step[c], &tab);
    const int r = nco_walk_l(x[c], step[c], count[c], wrap[c], &tab, segs + (size_t)c * TRK_MAXSEG, TRK_MAXSEG, &e);
    ns[c] = r; xe[c] = e;
}

extern "C" int sdrb200tk_nco_selftest(int n, unsigned long long seed, int mode, int *mismatches)
{
    if (n < 1 || !mismatches) return -1;
    double *hx = (double *)malloc(sizeof(double) * n * 3), *hs = hx + n, *hw = hs + n;
    int *hc = (int *)malloc(sizeof(int) * n);
    unsigned long long rs = seed;
    auto urand = [&]() { rs = rs * 6364136223846793005ULL + 1442695040888963407ULL; return (double)(rs >> 11) * (1.0 / 9007199254740992.0); };
    for (int i = 0; i < n; i++) {
        const int m = (mode < 0) ? i % 5 : mode;
        if (m == 0) {            /* IF carrier */
            double phi, ps; nco_carrier_init(4.092e6 + (urand() - 0.5) * 2e4, urand() * NCO_TWO_PI, 1.0 / 16.368e6, &phi, &ps);
            hx[i] = phi; hs[i] = ps; hw[i] = 0.0; hc[i] = 16360 + (int)(urand() * 20);
        } else if (m == 1) {     /* baseband carrier, both signs */
            double phi, ps; nco_carrier_init((urand() - 0.5) * 2e4, urand() * NCO_TWO_PI, 1.0 / 2.048e6, &phi, &ps);
            hx[i] = phi; hs[i] = ps; hw[i] = 0.0; hc[i] = 2040 + (int)(urand() * 20);
        } else if (m == 2) {     /* code, 16.368 MS/s */
            const double ci = (1.0 / 16.368e6) * (1.023e6 + (urand() - 0.5) * 20.0);
            hx[i] = nco_code_init(urand() * 1023.0, 18, ci, 1023); hs[i] = ci; hw[i] = 1023.0; hc[i] = 16396 + (int)(urand() * 20);
        } else if (m == 3) {     /* code, 2.048 MS/s and G1 */
            const int len = (urand() < 0.5) ? 1023 : 511;
            const double ci = (1.0 / 2.048e6) * ((len == 1023 ? 1.023e6 : 0.511e6) + (urand() - 0.5) * 20.0);
            hx[i] = nco_code_init(urand() * len, 4, ci, len); hs[i] = ci; hw[i] = (double)len; hc[i] = 2048 + (int)(urand() * 20);
        } else {                 /* short mantissas: exact ties in many binades, negative steps, zero */
            double st = ldexp(floor(urand() * 4096.0) + 0.5, -(int)(urand() * 12));
            if (urand() < 0.3) st = -st;
            if (urand() < 0.05) st = 0.0;
            hx[i] = ldexp(floor(urand() * 65536.0), -(int)(urand() * 20)); hs[i] = st; hw[i] = 0.0; hc[i] = 3000 + (int)(urand() * 3000);
        }
    }
    double *dx = NULL, *dxe = NULL; int *dc = NULL, *dns = NULL; nco_iseg_t *dseg = NULL;
    int rc = 0, bad = 0;
    if (cudaMalloc(&dx, sizeof(double) * n * 3) != cudaSuccess || cudaMalloc(&dxe, sizeof(double) * n) != cudaSuccess ||
        cudaMalloc(&dc, sizeof(int) * n) != cudaSuccess || cudaMalloc(&dns, sizeof(int) * n) != cudaSuccess ||
        cudaMalloc(&dseg, sizeof(nco_iseg_t) * TRK_MAXSEG * (size_t)n) != cudaSuccess) rc = -3;
    nco_iseg_t *hseg = (nco_iseg_t *)malloc(sizeof(nco_iseg_t) * TRK_MAXSEG * (size_t)n);
    int *hns = (int *)malloc(sizeof(int) * n);
    double *hxe = (double *)malloc(sizeof(double) * n);
    if (rc == 0) {
        cudaMemcpy(dx, hx, sizeof(double) * n * 3, cudaMemcpyHostToDevice);
        cudaMemcpy(dc, hc, sizeof(int) * n, cudaMemcpyHostToDevice);
        k_nco_selftest<<<n, 32>>>(dx, dx + n, dc, dx + 2 * n, dseg, dns, dxe);
        if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "nco_selftest: %s\n", cudaGetErrorString(cudaGetLastError())); rc = -4; }
        cudaMemcpy(hseg, dseg, sizeof(nco_iseg_t) * TRK_MAXSEG * (size_t)n, cudaMemcpyDeviceToHost);
        cudaMemcpy(hns, dns, sizeof(int) * n, cudaMemcpyDeviceToHost);
        cudaMemcpy(hxe, dxe, sizeof(double) * n, cudaMemcpyDeviceToHost);
    }
    for (int i = 0; i < n && rc == 0; i++) {
        static nco_iseg_t ref[NCO_MAXSEG];
        double xe;
        const int nr = nco_walk_i(hx[i], hs[i], hc[i], hw[i], ref, NCO_MAXSEG, &xe);
        const nco_iseg_t *g = hseg + (size_t)i * TRK_MAXSEG;
        int a = 0, b = 0, miss = 0;
        if (nr < 0 || hns[i] < 0 || hns[i] > TRK_MAXSEG) { bad += (nr < 0 && hns[i] < 0) ? 0 : 1; continue; }
        if (memcmp(&xe, &hxe[i], sizeof(double)) != 0) miss = 1;
        for (int k = 0; k < hc[i] && !miss; k++) {
            while (a + 1 < nr && ref[a + 1].k0 <= k) a++;
            while (b + 1 < hns[i] && g[b + 1].k0 <= k) b++;
            if (nco_iseg_trunc(&ref[a], k) != nco_iseg_trunc(&g[b], k)) miss = 1;
        }
        bad += miss;
    }
    cudaFree(dx); cudaFree(dxe); cudaFree(dc); cudaFree(dns); cudaFree(dseg);
    free(hx); free(hc); free(hseg); free(hns); free(hxe);
    *mismatches = bad;
    return rc;
}
