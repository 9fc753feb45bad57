/* f32x2bench.cu - issue rate of the packed fp32 instructions of sm_100a (FADD2 / FMUL2 / FFMA2,
 * PTX add/mul/fma.rn.f32x2) against their scalar forms, 8 warps per scheduler, 16 independent
 * chains per thread.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2bench f32x2bench.cu */
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c){ float2 r;
 asm volatile("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0,%1}, rd;}" : "=f"(r.x), "=f"(r.y) : "f"(a.x),"f"(a.y),"f"(b.x),"f"(b.y),"f"(c.x),"f"(c.y)); return r;}
__device__ __forceinline__ float2 add2(float2 a, float2 b){ float2 r;
 asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}" : "=f"(r.x), "=f"(r.y) : "f"(a.x),"f"(a.y),"f"(b.x),"f"(b.y)); return r;}
template <int MODE> __global__ void __launch_bounds__(256) k(float2 *p, int iters)
{
    float2 v[16]; const float2 w = p[threadIdx.x & 31], u = p[32 + (threadIdx.x & 31)];
#pragma unroll
    for (int e = 0; e < 16; e++) v[e] = p[64 + e];
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int e = 0; e < 16; e++) {
            if (MODE == 0) { v[e].x = fmaf(v[e].x, w.x, u.x); v[e].y = fmaf(v[e].y, w.y, u.y); }
            if (MODE == 1) v[e] = fma2(v[e], w, u);
            if (MODE == 2) { v[e].x = v[e].x + u.x; v[e].y = v[e].y + u.y; }
            if (MODE == 3) v[e] = add2(v[e], u);
            if (MODE == 4) { v[e] = fma2(make_float2(v[e].y, v[e].x), make_float2(-1.f, 1.f), u); }   /* swap + half negate */
            if (MODE == 5) { v[e] = fma2(make_float2(-w.y, w.x), make_float2(v[e].y, v[e].y), u); }   /* broadcast + swap */
        }
    }
    float2 s = v[0];
#pragma unroll
    for (int e = 1; e < 16; e++) { s.x += v[e].x; s.y += v[e].y; }
    if (s.x == 123.456f) p[threadIdx.x] = s;
}
template <int MODE> static void run(const char *name, float2 *d, int ops_per_elem)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 4096, grid = 148 * 8;
    k<MODE><<<grid, 256>>>(d, 16);
    cudaEventRecord(a); k<MODE><<<grid, 256>>>(d, iters); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double lane_ops = (double)grid * 256 * iters * 16 * 2;          /* scalar fp32 results */
    printf("%-28s %8.3f ms  %7.2f T lane-results/s  (%.1f T%s/s)\n", name, ms, lane_ops / ms * 1e-9,
           lane_ops * ops_per_elem / ms * 1e-9, ops_per_elem == 2 ? "FLOP" : "OP");
}
int main()
{
    float2 *d; cudaMalloc(&d, 4096 * sizeof(float2)); cudaMemset(d, 0, 4096 * sizeof(float2));
    run<0>("FFMA  (2 per float2)", d, 2); run<1>("FFMA2", d, 2);
    run<2>("FADD  (2 per float2)", d, 1); run<3>("FADD2", d, 1);
    run<4>("FFMA2 swap+NP modifiers", d, 2); run<5>("FFMA2 broadcast+swap.NP", d, 2);
    cudaError_t e = cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(e)); return e != cudaSuccess;
}
