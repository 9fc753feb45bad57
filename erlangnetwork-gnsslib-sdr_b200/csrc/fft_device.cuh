/* fft_device.cuh - CTA-level complex FFT in shared memory (hand-written
 * Stockham autosort, radix 16 with a radix 2/4/8 tail), the building block that
 * replaces the reference's FFTW calls cpxfft()/cpxifft() (src/sdrcmn.c:134-175)
 * on the acquisition path.
 *
 * One CTA of T = L/16 threads transforms L points (256 <= L <= 16384).
 * Every thread owns 16 points in registers, v[e] <-> x[tid + e*T], on entry
 * (natural order) AND on exit (natural order), so callers fuse their own
 * loads (carrier wipe-off, spectrum product) and stores (power, statistics)
 * around the transform without an extra pass over memory.
 * Between passes the data makes one round trip through shared memory; the
 * buffer is padded by one element per 16 so that the stride-16 scatter of the
 * first pass is bank-conflict free.
 */
#ifndef SDRB200_FFT_DEVICE_CUH
#define SDRB200_FFT_DEVICE_CUH

#include <cuda_runtime.h>

#define FFT_PAD(i) ((i) + ((i) >> 4))
#define FFT_SMEM_ELEMS(L) ((L) + ((L) >> 4) + 16)

/* Packed fp32 arithmetic of sm_100a: add/sub/mul/fma.rn.f32x2 work on a 64-bit register
 * pair and compile to FADD2 / FMUL2 / FFMA2, whose operand modifiers (half swap, scalar
 * broadcast, per-half negation) ptxas derives from the float2 shuffles written below - so a
 * complex add is ONE instruction and a complex multiply TWO (FMUL2 + FFMA2) instead of 2 and 4.
 * The kernel is issue/latency bound, not fp32-pipe bound, so halving the instruction count of
 * the butterflies is what counts (tools/f32x2bench.cu measures the rates).                    */
__device__ __forceinline__ float2 f2add(float2 a, float2 b)
{
    float2 r;
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 f2sub(float2 a, float2 b)
{
    float2 r;
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; sub.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 f2mul(float2 a, float2 b)
{
    float2 r;
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mul.rn.f32x2 rc, ra, rb; mov.b64 {%0,%1}, rc;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return r;
}
__device__ __forceinline__ float2 f2fma(float2 a, float2 b, float2 c)
{
    float2 r;
    asm("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; "
        "fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0,%1}, rd;}"
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return r;
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return f2add(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return f2sub(a, b); }
/* a * b = b.x * a + b.y * (i a) */
__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    return f2fma(make_float2(-a.y, a.x), make_float2(b.y, b.y), f2mul(make_float2(b.x, b.x), a));
}
/* a * conj(b) = b.x * a + b.y * (-i a) */
__device__ __forceinline__ float2 cmulc(float2 a, float2 b)
{
    return f2fma(make_float2(a.y, -a.x), make_float2(b.y, b.y), f2mul(make_float2(b.x, b.x), a));
}
/* multiply by SIGN*i : SIGN=-1 -> (y,-x) ; SIGN=+1 -> (-y,x) */
template <int SIGN> __device__ __forceinline__ float2 crot(float2 a)
{
    return SIGN < 0 ? make_float2(a.y, -a.x) : make_float2(-a.y, a.x);
}
/* c + SIGN*i*d and c - SIGN*i*d as one FFMA2 each (multiplier (1,1), swap and half negation as modifiers) */
template <int SIGN> __device__ __forceinline__ float2 caddrot(float2 c, float2 d)
{
    return f2fma(make_float2(d.y, d.x), SIGN < 0 ? make_float2(1.f, -1.f) : make_float2(-1.f, 1.f), c);
}
template <int SIGN> __device__ __forceinline__ float2 csubrot(float2 c, float2 d)
{
    return f2fma(make_float2(d.y, d.x), SIGN < 0 ? make_float2(-1.f, 1.f) : make_float2(1.f, -1.f), c);
}
/* exp(SIGN * 2*pi*i * K/16) as a compile-time constant */
template <int SIGN, int K> __device__ __forceinline__ float2 cw16(float2 a)
{
    constexpr float C[16] = {1.f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                             0.f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                             -1.f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f,
                             0.f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f};
    constexpr float S[16] = {0.f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f,
                             1.f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                             0.f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                             -1.f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f};
    constexpr int k = K & 15;
    if (k == 0) return a;
    if (k == 4) return crot<SIGN>(a);
    if (k == 8) return make_float2(-a.x, -a.y);
    if (k == 12) return crot<-SIGN>(a);
    const float c = C[k], s = (SIGN < 0) ? -S[k] : S[k];
    return f2fma(make_float2(-a.y, a.x), make_float2(s, s), f2mul(make_float2(c, c), a));
}

/* 4-point DFT in place on (a0,a1,a2,a3): b_u = sum_t a_t w4^(t*u), w4 = SIGN*i */
template <int SIGN>
__device__ __forceinline__ void fft4(float2 &a0, float2 &a1, float2 &a2, float2 &a3)
{
    float2 s02 = cadd(a0, a2), d02 = csub(a0, a2);
    float2 s13 = cadd(a1, a3), d13 = csub(a1, a3);
    a0 = cadd(s02, s13);
    a2 = csub(s02, s13);
    a1 = caddrot<SIGN>(d02, d13);
    a3 = csubrot<SIGN>(d02, d13);
}

/* 16-point DFT: v[u] = sum_t v[t] w16^(t*u), natural order in and out.
 * t = 4*t1 + t2, u = u1 + 4*u2                                              */
template <int SIGN>
__device__ __forceinline__ void fft16(float2 (&v)[16])
{
    /* step 1: over t1 for each t2 -> c[t2][u1] stored at v[4*u1 + t2] */
#pragma unroll
    for (int t2 = 0; t2 < 4; t2++) fft4<SIGN>(v[t2], v[4 + t2], v[8 + t2], v[12 + t2]);
    /* step 2: twiddle w16^(t2*u1) on element (u1,t2) = v[4*u1+t2] */
    v[5]  = cw16<SIGN, 1>(v[5]);  v[6]  = cw16<SIGN, 2>(v[6]);  v[7]  = cw16<SIGN, 3>(v[7]);
    v[9]  = cw16<SIGN, 2>(v[9]);  v[10] = cw16<SIGN, 4>(v[10]); v[11] = cw16<SIGN, 6>(v[11]);
    v[13] = cw16<SIGN, 3>(v[13]); v[14] = cw16<SIGN, 6>(v[14]); v[15] = cw16<SIGN, 9>(v[15]);
    /* step 3: over t2 for each u1 -> b[u1 + 4*u2] lands in v[4*u1 + u2] */
#pragma unroll
    for (int u1 = 0; u1 < 4; u1++) fft4<SIGN>(v[4 * u1], v[4 * u1 + 1], v[4 * u1 + 2], v[4 * u1 + 3]);
    /* transpose 4x4 so that v[u] = b[u]: currently v[4*u1+u2] = b[u1+4*u2] */
    float2 t;
#define FFT_SWAP(i, j) t = v[i]; v[i] = v[j]; v[j] = t;
    FFT_SWAP(1, 4) FFT_SWAP(2, 8) FFT_SWAP(3, 12) FFT_SWAP(6, 9) FFT_SWAP(7, 13) FFT_SWAP(11, 14)
#undef FFT_SWAP
}

/* exp(SIGN * 2*pi*i * K/32) * a, K = 0..15 */
template <int SIGN, int K> __device__ __forceinline__ float2 cw32(float2 a)
{
    constexpr float C[16] = {1.f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f,
                             0.70710678118654752f, 0.55557023301960218f, 0.38268343236508977f, 0.19509032201612825f,
                             0.f, -0.19509032201612825f, -0.38268343236508977f, -0.55557023301960218f,
                             -0.70710678118654752f, -0.83146961230254524f, -0.92387953251128674f, -0.98078528040323043f};
    constexpr float S[16] = {0.f, 0.19509032201612825f, 0.38268343236508977f, 0.55557023301960218f,
                             0.70710678118654752f, 0.83146961230254524f, 0.92387953251128674f, 0.98078528040323043f,
                             1.f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f,
                             0.70710678118654752f, 0.55557023301960218f, 0.38268343236508977f, 0.19509032201612825f};
    static_assert(K >= 0 && K < 16, "cw32: K in 0..15");
    if (K == 0) return a;
    if (K == 8) return crot<SIGN>(a);
    const float c = C[K], s = (SIGN < 0) ? -S[K] : S[K];
    return f2fma(make_float2(-a.y, a.x), make_float2(s, s), f2mul(make_float2(c, c), a));
}

/* 32-point DFT in registers, natural order in and out: even/odd 16-point transforms and one
 * radix-2 stage.  HALF: only outputs 0..15 are formed (v[16..31] are left undefined) - the
 * recombination of 32 sub-transforms needs lags below half the transform length only.       */
template <int SIGN, bool HALF>
__device__ __forceinline__ void fft32(float2 (&v)[32])
{
    float2 e[16], o[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { e[k] = v[2 * k]; o[k] = v[2 * k + 1]; }
    fft16<SIGN>(e);
    fft16<SIGN>(o);
#define FFT32_OUT(K) { const float2 t = cw32<SIGN, K>(o[K]); v[K] = cadd(e[K], t); if (!HALF) v[K + 16] = csub(e[K], t); }
    FFT32_OUT(0) FFT32_OUT(1) FFT32_OUT(2) FFT32_OUT(3) FFT32_OUT(4) FFT32_OUT(5) FFT32_OUT(6) FFT32_OUT(7)
    FFT32_OUT(8) FFT32_OUT(9) FFT32_OUT(10) FFT32_OUT(11) FFT32_OUT(12) FFT32_OUT(13) FFT32_OUT(14) FFT32_OUT(15)
#undef FFT32_OUT
}

/* v[k] *= w^k for k = 1..31: w^16 by four squarings, then one running power w^k that serves
 * v[k] and (times w^16) v[k+16]; CONJ multiplies by the conjugates                          */
template <bool CONJ>
__device__ __forceinline__ void twiddle32(float2 (&v)[32], float2 w)
{
    const float2 w2 = cmul(w, w), w4 = cmul(w2, w2), w8 = cmul(w4, w4), w16 = cmul(w8, w8);
    v[16] = CONJ ? cmulc(v[16], w16) : cmul(v[16], w16);
    float2 wk = w;
#pragma unroll
    for (int k = 1; k < 16; k++) {
        const float2 wh = cmul(wk, w16);
        v[k] = CONJ ? cmulc(v[k], wk) : cmul(v[k], wk);
        v[k + 16] = CONJ ? cmulc(v[k + 16], wh) : cmul(v[k + 16], wh);
        if (k < 15) wk = cmul(wk, w);
    }
}

/* 1024-point transform by ONE WARP: lane l holds v[e] = x[l + 32 e] on entry and X[l + 32 e] on
 * exit.  x = 32-point transform over e in registers, twiddle W_1024^(l a), transpose through the
 * warp's own 33 x 32 shared-memory tile (no CTA barrier), 32-point transform over l.
 * w1 = exp(SIGN 2 pi i l / 1024) of this lane.  before_smem() as in cta_fft_pre_hook.        */
#define WFFT_TILE 1056
template <int SIGN, typename Hook>
__device__ __forceinline__ void warp_fft1024(float2 (&v)[32], float2 *tile, float2 w1, Hook before_smem)
{
    const int lane = threadIdx.x & 31;
    fft32<SIGN, false>(v);
    twiddle32<false>(v, w1);
    before_smem();
#pragma unroll
    for (int a = 0; a < 32; a++) tile[a * 33 + lane] = v[a];
    __syncwarp();
#pragma unroll
    for (int l = 0; l < 32; l++) v[l] = tile[lane * 33 + l];
    __syncwarp();
    fft32<SIGN, false>(v);
}

/* final pass of radix RL (2, 4 or 8) over the 16 registers: 16/RL butterflies,
 * butterfly i uses elements e = i + t*(16/RL); no twiddles (m == 1).          */
template <int SIGN, int RL>
__device__ __forceinline__ void fft_tail(float2 (&v)[16])
{
    if (RL == 2) {
#pragma unroll
        for (int i = 0; i < 8; i++) { float2 a = v[i], b = v[i + 8]; v[i] = cadd(a, b); v[i + 8] = csub(a, b); }
    } else if (RL == 4) {
#pragma unroll
        for (int i = 0; i < 4; i++) fft4<SIGN>(v[i], v[i + 4], v[i + 8], v[i + 12]);
    } else if (RL == 8) {
        /* t = 4*t1 + t2 (t1<2,t2<4), u = u1 + 2*u2; element t of butterfly i is v[i + 2*t] */
#pragma unroll
        for (int i = 0; i < 2; i++) {
            float2 c0[4], c1[4];
#pragma unroll
            for (int t2 = 0; t2 < 4; t2++) {
                float2 a = v[i + 2 * t2], b = v[i + 2 * (4 + t2)];
                c0[t2] = cadd(a, b);
                c1[t2] = csub(a, b);
            }
            c1[1] = cw16<SIGN, 2>(c1[1]);                      /* w8^1 = w16^2 */
            c1[2] = cw16<SIGN, 4>(c1[2]);
            c1[3] = cw16<SIGN, 6>(c1[3]);
            fft4<SIGN>(c0[0], c0[1], c0[2], c0[3]);            /* u1 = 0 -> b[2*u2]   */
            fft4<SIGN>(c1[0], c1[1], c1[2], c1[3]);            /* u1 = 1 -> b[1+2*u2] */
#pragma unroll
            for (int u2 = 0; u2 < 4; u2++) {
                v[i + 2 * (2 * u2)] = c0[u2];
                v[i + 2 * (2 * u2 + 1)] = c1[u2];
            }
        }
    }
}

template <int L> struct fft_plan_t {
    static constexpr int T = L / 16;
    static constexpr int N16 = (L >= 4096) ? 3 : 2;
    static constexpr int RL = L / (N16 == 3 ? 4096 : 256);     /* 1, 2, 4 or 8 */
    static_assert(L == 256 || L == 512 || L == 1024 || L == 2048 || L == 4096 || L == 8192 || L == 16384,
                  "unsupported CTA FFT length");
};

/* L-point transform by the whole CTA (blockDim.x == L/16).
 * tw: forward twiddles exp(-2*pi*i*k/N) for a table of N = L*tw_stride entries
 * sm: FFT_SMEM_ELEMS(L) float2 of shared memory
 * Contains __syncthreads(); must be called by all threads of the CTA.
 * The per-pass base twiddles depend on the thread index only; cta_fft_twiddles
 * fetches them so that a caller can issue the loads early (together with its
 * own operand loads) and cta_fft_pre consumes them.                            */
template <int L>
__device__ __forceinline__ void cta_fft_twiddles(const float2 *__restrict__ tw, int tw_stride,
                                                 float2 (&w)[fft_plan_t<L>::N16])
{
    using P = fft_plan_t<L>;
    const int tid = threadIdx.x;
    int s = 1;
#pragma unroll
    for (int pass = 0; pass < P::N16; pass++) {
        const int ps = tid - (tid & (s - 1));
        if (!((pass == P::N16 - 1) && (P::RL == 1))) w[pass] = __ldg(&tw[(size_t)ps * tw_stride]);
        else w[pass] = make_float2(1.f, 0.f);
        s *= 16;
    }
}

struct fft_no_hook { __device__ __forceinline__ void operator()() const {} };

/* before_smem() runs once, after the register part of the first pass and before the first write to
 * sm: a caller whose buffer is still being read (by peer CTAs) waits there instead of up front    */
template <int L, int SIGN, typename Hook>
__device__ __forceinline__ void cta_fft_pre_hook(float2 (&v)[16], float2 *sm, const float2 (&w)[fft_plan_t<L>::N16],
                                                 Hook before_smem)
{
    using P = fft_plan_t<L>;
    constexpr int T = P::T;
    const int tid = threadIdx.x;
    /* gather address of element e: FFT_PAD(tid + e*T) = gbase + e*(T + T/16), T % 16 == 0 */
    float2 *const gbase = sm + tid + (tid >> 4);
    int s = 1;
#pragma unroll
    for (int pass = 0; pass < P::N16; pass++) {
        fft16<SIGN>(v);
        const bool last = (pass == P::N16 - 1) && (P::RL == 1);
        if (!last) {
            const int q = tid & (s - 1);
            const int ps = tid - q;                               /* p*s */
            {
                /* unconditional: threads with ps == 0 multiply by 1 (no divergent path) */
                float2 w1 = w[pass];
                if (SIGN > 0) w1.y = -w1.y;
                float2 w2 = cmul(w1, w1), w3 = cmul(w2, w1), w4 = cmul(w2, w2);
                float2 w5 = cmul(w4, w1), w6 = cmul(w3, w3), w7 = cmul(w4, w3), w8 = cmul(w4, w4);
                v[1] = cmul(v[1], w1); v[2] = cmul(v[2], w2); v[3] = cmul(v[3], w3); v[4] = cmul(v[4], w4);
                v[5] = cmul(v[5], w5); v[6] = cmul(v[6], w6); v[7] = cmul(v[7], w7); v[8] = cmul(v[8], w8);
                float2 w9 = cmul(w8, w1), w10 = cmul(w5, w5), w11 = cmul(w8, w3), w12 = cmul(w6, w6);
                v[9] = cmul(v[9], w9); v[10] = cmul(v[10], w10); v[11] = cmul(v[11], w11); v[12] = cmul(v[12], w12);
                float2 w13 = cmul(w8, w5), w14 = cmul(w7, w7), w15 = cmul(w8, w7);
                v[13] = cmul(v[13], w13); v[14] = cmul(v[14], w14); v[15] = cmul(v[15], w15);
            }
            /* scatter: y[q + s*(16p + u)]; with idx = q + 16*ps + s*u and q < s:
             * s == 1:  FFT_PAD(idx) = 17*tid + u
             * s >= 16: FFT_PAD(idx) = q + (q >> 4) + 17*ps + u*(s + s/16)        */
            float2 *const sbase = sm + ((s == 1) ? 17 * tid : q + (q >> 4) + 17 * ps);
            const int sstep = (s == 1) ? 1 : s + s / 16;
            if (pass == 0) before_smem();
#pragma unroll
            for (int u = 0; u < 16; u++) sbase[sstep * u] = v[u];
            __syncthreads();
#pragma unroll
            for (int e = 0; e < 16; e++) v[e] = gbase[e * (T + T / 16)];
            __syncthreads();
        }
        s *= 16;
    }
    if (P::RL > 1) fft_tail<SIGN, P::RL>(v);
}

template <int L, int SIGN>
__device__ __forceinline__ void cta_fft_pre(float2 (&v)[16], float2 *sm, const float2 (&w)[fft_plan_t<L>::N16])
{
    cta_fft_pre_hook<L, SIGN>(v, sm, w, fft_no_hook());
}

template <int L, int SIGN>
__device__ __forceinline__ void cta_fft(float2 (&v)[16], float2 *sm, const float2 *__restrict__ tw,
                                        int tw_stride)
{
    float2 w[fft_plan_t<L>::N16];
    cta_fft_twiddles<L>(tw, tw_stride, w);
    cta_fft_pre<L, SIGN>(v, sm, w);
}

#endif
