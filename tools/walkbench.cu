// microbenchmark: one lane walking the NCO tables of one code period (tracking kernel's NCO lanes),
// alone on an SM (NTHR=64) and beside busy integer warps (NTHR=1024).  mode 0: carrier at IF,
// 1: code, 2: baseband carrier with negative Doppler.  Old = nco_walk_i, lean = nco_walk_l.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../erlangnetwork-gnsslib-sdr_b200/csrc/nco_exact.h"
__device__ __noinline__ int walk_old(double x, double step, int count, double wrap, nco_iseg_t *seg, double *xe)
{ return nco_walk_i(x, step, count, wrap, seg, 96, xe); }
__device__ __noinline__ int walk_lean(double x, double step, int count, double wrap, const nco_steptab_t *t, nco_iseg_t *seg, double *xe, nco_raw_t *raw)
{ (void)seg; return nco_walk_raw(x, step, count, wrap, t, raw, 96, xe); }
__device__ __noinline__ double rem_call(double x) { return nco_carrier_rem(x); }
__global__ void k(double x0, double step, int n, double wrap, long long *out, int reps, int mode, int lean)
{
    __shared__ nco_iseg_t seg[96];
    __shared__ nco_steptab_t tab;
    __shared__ nco_raw_t raw[96];
    __shared__ volatile int stop;
    if (threadIdx.x == 0) { stop = 0; nco_steptab_build(step, &tab); }
    __syncthreads();
    if (threadIdx.x >= 32) {                       /* busy neighbours: integer ALU loop until told to stop */
        unsigned a = threadIdx.x, b = 12345u;
        if (blockDim.x > 64) while (!stop) { for (int i = 0; i < 64; i++) { a = a * 1664525u + b; b ^= a >> 3; } }
        if (a == 0x12345678u) out[7] = b;
        return;
    }
    if (threadIdx.x) return;
    double xe = 0, x = x0; int ns = 0; long long tw = 0, tr = 0;
    for (int r = 0; r < reps; r++) {
        long long t0 = clock64();
        ns += lean ? walk_lean(x, step, n, wrap, &tab, seg, &xe, raw) : walk_old(x, step, n, wrap, seg, &xe);
        long long t1 = clock64();
        tw += t1 - t0;
        if (mode != 1) { double rm = rem_call(xe); long long t2 = clock64(); tr += t2 - t1; if (rm < 0) rm += NCO_TWO_PI; x = rm * 32.0 / NCO_TWO_PI; }
        else x = xe - floor(xe / wrap) * wrap;
    }
    stop = 1;
    out[0] = tw; out[1] = tr; out[2] = ns; out[3] = (long long)(x * 1e6);
}
int main()
{
    long long *d, h[4];
    cudaMalloc(&d, 64);
    const char *names[3] = {"carrier IF", "code", "carrier baseband, negative Doppler"};
    for (int nthr = 64; nthr <= 1024; nthr += 960)
        for (int mode = 0; mode < 3; mode++)
            for (int lean = 0; lean < 2; lean++) {
                double phi, ps, wrap = 0.0; int n = 16368;
                if (mode == 0) nco_carrier_init(4.092e6 + 1234.5, 1.2345, 1.0 / 16.368e6, &phi, &ps);
                else if (mode == 1) { phi = 516.175; ps = (1.0 / 16.368e6) * (1.023e6 + 0.7); wrap = 1023.0; n = 16368 + 36; }
                else { nco_carrier_init(-3211.0, 2.2, 1.0 / 2.048e6, &phi, &ps); n = 2048; }
                for (int it = 0; it < 2; it++) { k<<<1, nthr>>>(phi, ps, n, wrap, d, 100, mode, lean); cudaDeviceSynchronize(); }
                cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
                printf("nthr %4d %-36s %-4s: walk %6lld cycles, 2pi reduction %5lld cycles, %.1f segments, chk %lld\n", nthr, names[mode],
                       lean ? "lean" : "old", h[0] / 100, h[1] / 100, h[2] / 100.0, h[3]);
            }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
