/* ws.cu - per-thread workspace of the single-call primitives.
 *
 * The reference's hot functions malloc/free their temporaries on every call
 * (src/sdrcmn.c:697-699, 747-750) and are called concurrently from one pthread
 * per channel (src/sdrmain.c:144-149).  Here each calling thread owns one CUDA
 * stream plus grow-only device and pinned-host arenas, so calls from different
 * channel threads never serialise on a lock or an allocation.
 */
#include <stdlib.h>
#include "sdrb200_internal.h"

struct ws_holder {
    sdrb200_ws_t ws{};
    ~ws_holder()
    {
        if (ws.d_arena) cudaFree(ws.d_arena);
        if (ws.h_arena) cudaFreeHost(ws.h_arena);
        if (ws.cuda_stream) cudaStreamDestroy((cudaStream_t)ws.cuda_stream);
    }
};
static thread_local ws_holder g_ws;

extern "C" sdrb200_ws_t *sdrb200_ws(size_t d_bytes, size_t h_bytes)
{
    sdrb200_ws_t *w = &g_ws.ws;
    if (sdrb200_ensure_device()) return NULL;
    if (!w->cuda_stream) {
        cudaStream_t st;
        if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) {
            sdrb200_set_error("workspace: cudaStreamCreate failed"); return NULL;
        }
        w->cuda_stream = st;
    }
    if (d_bytes > w->d_cap) {
        size_t cap = d_bytes + (d_bytes >> 2) + 4096;
        cudaStreamSynchronize((cudaStream_t)w->cuda_stream);
        if (w->d_arena) cudaFree(w->d_arena);
        w->d_arena = NULL; w->d_cap = 0;
        if (cudaMalloc((void **)&w->d_arena, cap) != cudaSuccess) {
            sdrb200_set_error("workspace: cudaMalloc(%zu) failed", cap); cudaGetLastError(); return NULL;
        }
        w->d_cap = cap;
    }
    if (h_bytes > w->h_cap) {
        size_t cap = h_bytes + (h_bytes >> 2) + 4096;
        cudaStreamSynchronize((cudaStream_t)w->cuda_stream);
        if (w->h_arena) cudaFreeHost(w->h_arena);
        w->h_arena = NULL; w->h_cap = 0;
        if (cudaMallocHost((void **)&w->h_arena, cap) != cudaSuccess) {
            sdrb200_set_error("workspace: cudaMallocHost(%zu) failed", cap); cudaGetLastError(); return NULL;
        }
        w->h_cap = cap;
    }
    return w;
}
