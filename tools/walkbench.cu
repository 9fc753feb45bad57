// microbenchmark: one lane walking NCO tables on an otherwise idle SM
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../erlangnetwork-gnsslib-sdr_b200/csrc/nco_exact.h"
__device__ __noinline__ int walk_call(double x, double step, int count, double wrap, nco_iseg_t *seg, double *xe)
{ return nco_walk_i(x, step, count, wrap, seg, 96, xe); }
__device__ __noinline__ double rem_call(double x) { return nco_carrier_rem(x); }
__global__ void k(double x0, double step, int n, double wrap, long long *out, int reps, int mode)
{
    __shared__ nco_iseg_t seg[96];
    __shared__ volatile int stop;
    if (threadIdx.x == 0) stop = 0;
    __syncthreads();
    if (threadIdx.x >= 32) {                       /* busy neighbours: integer ALU loop until told to stop */
        unsigned a = threadIdx.x, b = 12345u;
        while (!stop) { for (int i = 0; i < 64; i++) { a = a * 1664525u + b; b ^= a >> 3; } }
        if (a == 0x12345678u) out[7] = b;
        return;
    }
    if (threadIdx.x) return;
    double xe = 0, x = x0; int ns = 0; long long tw = 0, tr = 0;
    for (int r = 0; r < reps; r++) {
        long long t0 = clock64();
        ns += walk_call(x, step, n, wrap, seg, &xe);
        long long t1 = clock64();
        tw += t1 - t0;
        if (mode == 0) { double rm = rem_call(xe); tr += clock64() - t1; x = rm * 32.0 / NCO_TWO_PI; }
        else x = xe - floor(xe / wrap) * wrap;
    }
    stop = 1;
    out[0] = tw; out[1] = tr; out[2] = ns; out[3] = (long long)(x * 1e6);
}
int main()
{
    long long *d, h[4];
    cudaMalloc(&d, 64); int nthr = getenv("NTHR") ? atoi(getenv("NTHR")) : 32;
    for (int mode = 0; mode < 2; mode++) {
        double phi, ps;
        nco_carrier_init(4.092e6 + 1234.5, 1.2345, 1.0 / 16.368e6, &phi, &ps);
        double ci = (1.0 / 16.368e6) * (1.023e6 + 0.7);
        for (int it = 0; it < 2; it++) {
            if (mode == 0) k<<<1, nthr>>>(phi, ps, 16368, 0.0, d, 100, 0);
            else k<<<1, nthr>>>(516.175, ci, 16368 + 36, 1023.0, d, 100, 1);
            cudaDeviceSynchronize();
        }
        cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
        printf("mode %d: walk %lld cycles/call, rem %lld cycles/call, segs/call %.1f chk %lld\n", mode, h[0] / 100, h[1] / 100, h[2] / 100.0, h[3]);
    }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
