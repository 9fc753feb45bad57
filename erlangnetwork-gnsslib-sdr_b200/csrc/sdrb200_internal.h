/* sdrb200_internal.h - shared declarations of the libsdrb200 translation units */
#ifndef SDRB200_INTERNAL_H
#define SDRB200_INTERNAL_H

#include <stdint.h>
#include <stddef.h>
#include "../../include/sdrb200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* reference constants of the carrier table: CDIV, CMASK, CSCALE (src/sdrcmn.c:9-11) */
#define SDRB200_CDIV   32
#define SDRB200_CMASK  0x1F
#define SDRB200_CSCALE (1.0 / 32.0)

/* error plumbing (core.cu) */
void sdrb200_set_error(const char *fmt, ...);
int  sdrb200_ensure_device(void);          /* SDRB200_OK or SDRB200_ENODEV      */
void sdrb200_count_launch(int n);

/* codegen.c */
int sdrb200_gencode_l1ca(int prn, int8_t *code);
int sdrb200_gencode_g1(int8_t *code);
const uint16_t *sdrb200_g2_delay_table(void);   /* [210] G2 delays, PRN 1..210 */

/* device IF stream (core.cu) */
struct sdrb200_stream {
    int       dtype;        /* bytes per sample                                  */
    uint64_t  capacity;     /* samples held (ring)                               */
    uint64_t  count;        /* samples pushed so far                             */
    int8_t   *d_buf;        /* capacity*dtype + guard bytes                      */
    void     *cuda_stream;  /* cudaStream_t used for pushes                      */
};
#define SDRB200_STREAM_GUARD (4u << 20)     /* mirrored bytes past the ring end   */

/* per-thread workspace for the single-call primitives (ws.cu): a stream, a
 * device scratch arena and a pinned host staging arena, both grow-only.        */
typedef struct {
    void   *cuda_stream;
    char   *d_arena;  size_t d_cap;
    char   *h_arena;  size_t h_cap;
} sdrb200_ws_t;
/* returns NULL (and sets the error) if no device / allocation failure; the
 * arenas are at least d_bytes / h_bytes large and 256-byte aligned             */
sdrb200_ws_t *sdrb200_ws(size_t d_bytes, size_t h_bytes);

#ifdef __cplusplus
}
#endif

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define SDRB200_CUDA_TRY(expr)                                                   \
    do {                                                                         \
        cudaError_t e__ = (expr);                                                \
        if (e__ != cudaSuccess) {                                                \
            sdrb200_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,       \
                              cudaGetErrorString(e__));                          \
            return SDRB200_ECUDA;                                                \
        }                                                                        \
    } while (0)
#endif

#endif
