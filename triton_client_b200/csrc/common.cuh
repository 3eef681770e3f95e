// Shared host/device helpers for libtcb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include "../../include/tcb200.h"

#define TCB_ERRLEN 512

struct tcb_ctx {
    int device;
    int num_sms;
    char* ws;          // device workspace (one allocation, made in tcb_create)
    size_t ws_bytes;
    size_t ws_used;    // bump pointer, reset at the start of every entry point
    long long launches;
    int pdl;           // launch the hot kernels with programmatic stream serialization (TCB_PDL, default on)
    int* tickets;      // TCB_TICKETS ints behind the workspace, zero between calls: "last CTA of a group" counters that
                       // the kernel using them resets itself, so no per-call memset is needed
    float pre_div; int pre_div_known, pre_exact_recip, pre_exact2;   // tcb_preprocess_image: exact-division checks of the last divisor
    char* stage;       // TCB_STAGE_SLOTS pinned slots of TCB_STAGE_BYTES for small host tables (tcb_stage_h2d)
    cudaEvent_t stage_done[8];
    int stage_next;
    char err[TCB_ERRLEN];
};
#define TCB_TICKETS 65536
#define TCB_STAGE_SLOTS 8
#define TCB_STAGE_BYTES 16384

// Copy a small host table to the device without blocking the caller: through a pinned slot of the context (a slot is
// reused only after the copy that read it has finished); tables larger than a slot go straight from the caller's memory.
int tcb_stage_h2d(tcb_ctx* ctx, void* dev, const void* host, size_t bytes, cudaStream_t st);

// Per-file set-up run once by tcb_create on the context's device (opt-ins to > 48 KB of dynamic shared memory), so the
// entry points themselves make no driver calls besides their launches.
typedef int (*tcb_init_fn)(tcb_ctx*);
struct TcbInitReg { explicit TcbInitReg(tcb_init_fn fn); };
#define TCB_SMEM_OPT_IN(ctx, kernel, bytes) \
    TCB_CUDA((ctx), cudaFuncSetAttribute((const void*)(kernel), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)))
#define TCB_SMEM_LIMIT (227 * 1024)

int tcb_fail(tcb_ctx* ctx, int code, const char* fmt, ...);
void tcb_clear_error(tcb_ctx* ctx);

// Bump allocation out of the context workspace (256 B aligned). Returns nullptr when exhausted.
static inline void* tcb_ws_alloc(tcb_ctx* ctx, size_t bytes) {
    size_t off = (ctx->ws_used + 255) & ~(size_t)255;
    if (off + bytes > ctx->ws_bytes) return nullptr;
    ctx->ws_used = off + bytes;
    return ctx->ws + off;
}
static inline void tcb_ws_reset(tcb_ctx* ctx) { ctx->ws_used = 0; }

#define TCB_CHECK_CTX(ctx)                                                     \
    do {                                                                       \
        if (!(ctx)) return tcb_fail(nullptr, TCB_ERR_ARG, "%s: null context", __func__); \
        tcb_clear_error(ctx);                                                  \
        tcb_ws_reset(ctx);                                                     \
    } while (0)

#define TCB_REQUIRE(ctx, cond, ...)                                            \
    do {                                                                       \
        if (!(cond)) return tcb_fail((ctx), TCB_ERR_ARG, __VA_ARGS__);         \
    } while (0)

#define TCB_CUDA(ctx, call)                                                    \
    do {                                                                       \
        cudaError_t e__ = (call);                                              \
        if (e__ != cudaSuccess)                                                \
            return tcb_fail((ctx), TCB_ERR_CUDA, "%s: %s -> %s", __func__, #call, cudaGetErrorString(e__)); \
    } while (0)

// After a kernel launch: count it and surface launch-configuration errors.
#define TCB_LAUNCHED(ctx)                                                      \
    do {                                                                       \
        (ctx)->launches++;                                                     \
        cudaError_t e__ = cudaGetLastError();                                  \
        if (e__ != cudaSuccess)                                                \
            return tcb_fail((ctx), TCB_ERR_CUDA, "%s: launch failed: %s", __func__, cudaGetErrorString(e__)); \
    } while (0)

// Launch `kernel` (which must begin with pdl_wait() / pdl_enter()) with programmatic stream serialization when the context allows it.
#define TCB_KLAUNCH(ctx, kernel, grid, block, smem, st, ...)                                             \
    TCB_CUDA((ctx), tcb_launch((ctx)->pdl != 0, kernel, dim3(grid), dim3(block), (size_t)(smem), (st), __VA_ARGS__))

#define TCB_WS(ctx, ptr, type, count)                                          \
    type* ptr = (type*)tcb_ws_alloc((ctx), sizeof(type) * (size_t)(count));    \
    if (!ptr) return tcb_fail((ctx), TCB_ERR_WORKSPACE, "%s: workspace too small (need > %zu B, have %zu B)", \
                              __func__, (ctx)->ws_used + sizeof(type) * (size_t)(count), (ctx)->ws_bytes)

static inline int tcb_div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

#ifdef __CUDACC__
// 128-bit streaming load/store (read-once inputs, write-once outputs: keep them out of L1).
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_f4(float* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }

// Programmatic dependent launch. A kernel launched through tcb_launch(pdl = true) is scheduled as soon as every CTA of
// the kernel before it on the stream has exited, without waiting for that kernel's memory flush and the launch latency
// behind it; it must call pdl_wait() before it touches global memory (a no-op without the launch attribute).
// No kernel here triggers its successor early (griddepcontrol.launch_dependents): measured on the bench step and the
// voxelisation chain, the early trigger gained nothing over this form (configs[1] step with every launch edge intact:
// 121.2 us wait-only, 122.6 us with the trigger — the waiting CTAs take room from the producer's tail), and a
// successor that is resident while its producer still runs must not read the producer's output through the
// non-coherent path (ld.global.nc — what `const T* __restrict__` and __ldg compile to), which cost a stale pair
// count in the wide rotated-NMS path when it was tried.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() {
#ifdef TCB_PDL_EARLY   // A/B builds only: trigger the successor before waiting (see above for why this is not the default)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
    pdl_wait();
}
// A plain (coherent) load where one would write __ldg(p).
template <typename T>
__device__ __forceinline__ T ld_coherent(const T* p) { return *p; }
template <typename... KA, typename... A>
static inline cudaError_t tcb_launch(bool pdl, void (*kernel)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr = {};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KA(args)...);
}
#endif
