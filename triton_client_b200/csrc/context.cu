// Context, error reporting and workspace ownership for libtcb200 (SURVEY.md §8b "error convention").
#include "common.cuh"
#include <new>
#include <stdlib.h>

static thread_local char g_err[TCB_ERRLEN] = "";

int tcb_fail(tcb_ctx* ctx, int code, const char* fmt, ...) {
    char* dst = ctx ? ctx->err : g_err;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, TCB_ERRLEN, fmt, ap);
    va_end(ap);
    if (ctx) { strncpy(g_err, dst, TCB_ERRLEN - 1); g_err[TCB_ERRLEN - 1] = 0; }
    return code;
}

void tcb_clear_error(tcb_ctx* ctx) { if (ctx) ctx->err[0] = 0; }

// Filled by the static TcbInitReg objects of the other translation units while the library loads (plain zero-initialised
// storage, so the order of those constructors does not matter).
static tcb_init_fn g_init[32];
static int g_num_init;
TcbInitReg::TcbInitReg(tcb_init_fn fn) { if (g_num_init < 32) g_init[g_num_init++] = fn; }

int tcb_stage_h2d(tcb_ctx* ctx, void* dev, const void* host, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return TCB_OK;
    if (bytes > TCB_STAGE_BYTES) {
        TCB_CUDA(ctx, cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, st));
        return TCB_OK;
    }
    const int slot = ctx->stage_next;
    ctx->stage_next = (slot + 1) % TCB_STAGE_SLOTS;
    TCB_CUDA(ctx, cudaEventSynchronize(ctx->stage_done[slot]));   // the copy that last read this slot (long finished, normally)
    char* pin = ctx->stage + (size_t)slot * TCB_STAGE_BYTES;
    memcpy(pin, host, bytes);
    TCB_CUDA(ctx, cudaMemcpyAsync(dev, pin, bytes, cudaMemcpyHostToDevice, st));
    TCB_CUDA(ctx, cudaEventRecord(ctx->stage_done[slot], st));
    return TCB_OK;
}

extern "C" {

const char* tcb_version(void) { return "tcb200 0.1 (sm_100a)"; }

int tcb_create(int device, size_t workspace_bytes, tcb_ctx** out) {
    if (!out) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_create: out is null");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return tcb_fail(nullptr, TCB_ERR_CUDA, "tcb_create: cudaGetDeviceCount -> %s", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return tcb_fail(nullptr, TCB_ERR_ARG, "tcb_create: device %d out of range [0,%d)", device, ndev);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return tcb_fail(nullptr, TCB_ERR_CUDA, "tcb_create: cudaSetDevice -> %s", cudaGetErrorString(e));
    tcb_ctx* c = new (std::nothrow) tcb_ctx();
    if (!c) return tcb_fail(nullptr, TCB_ERR_ALLOC, "tcb_create: host allocation failed");
    memset(c, 0, sizeof(*c));
    c->device = device;
    { const char* e2 = getenv("TCB_PDL"); c->pdl = !(e2 && e2[0] == '0'); }
    cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, device);
    if (c->num_sms <= 0) c->num_sms = 148;
    // one allocation: the caller's workspace, then the self-resetting ticket counters
    const size_t ws_round = (workspace_bytes + 255) & ~(size_t)255;
    e = cudaMalloc((void**)&c->ws, ws_round + sizeof(int) * TCB_TICKETS);
    if (e == cudaSuccess) e = cudaMemset(c->ws + ws_round, 0, sizeof(int) * TCB_TICKETS);
    if (e != cudaSuccess) {
        if (c->ws) cudaFree(c->ws);
        delete c;
        return tcb_fail(nullptr, TCB_ERR_ALLOC, "tcb_create: cudaMalloc(%zu) -> %s", workspace_bytes, cudaGetErrorString(e));
    }
    c->tickets = reinterpret_cast<int*>(c->ws + ws_round);
    c->ws_bytes = workspace_bytes;
    e = cudaMallocHost((void**)&c->stage, (size_t)TCB_STAGE_SLOTS * TCB_STAGE_BYTES);
    for (int i = 0; i < TCB_STAGE_SLOTS && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&c->stage_done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) {
        tcb_destroy(c);
        return tcb_fail(nullptr, TCB_ERR_ALLOC, "tcb_create: pinned staging -> %s", cudaGetErrorString(e));
    }
    for (int i = 0; i < g_num_init; i++) {
        const int rc = g_init[i](c);
        if (rc != TCB_OK) {
            tcb_fail(nullptr, rc, "tcb_create: %s", c->err);
            tcb_destroy(c);
            return rc;
        }
    }
    *out = c;
    return TCB_OK;
}

int tcb_destroy(tcb_ctx* ctx) {
    if (!ctx) return TCB_OK;
    if (ctx->ws) cudaFree(ctx->ws);
    for (int i = 0; i < TCB_STAGE_SLOTS; i++) if (ctx->stage_done[i]) cudaEventDestroy(ctx->stage_done[i]);
    if (ctx->stage) cudaFreeHost(ctx->stage);
    delete ctx;
    return TCB_OK;
}

const char* tcb_last_error(const tcb_ctx* ctx) { return ctx ? ctx->err : g_err; }
size_t tcb_workspace_bytes(const tcb_ctx* ctx) { return ctx ? ctx->ws_bytes : 0; }
long long tcb_launch_count(const tcb_ctx* ctx) { return ctx ? ctx->launches : 0; }

}  // extern "C"
