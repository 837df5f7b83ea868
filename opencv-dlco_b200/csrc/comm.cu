// dlco_comm: the all-reduce of a pair-sharded run (the D x D gradient of pj-learn, the 2 x szBatch margins,
// src/pj-learn.cpp:332-418 split over GPUs that each own a row range of "Distance") as ONE kernel per GPU that
// reads its peers' partial sums directly over NVLink / NVSwitch peer memory -- no library collective.
//
// Every rank owns a staging area [2 parities][max_count floats] and a row of n "ready" words, all mapped into
// every peer's address space (cudaDeviceEnablePeerAccess inside one process, CUDA IPC handles across
// processes).  One call = one kernel on the rank's own stream:
//   1. copy the local input into the staging buffer of this call's parity;
//   2. the last block to finish publishes the call's epoch into every peer's ready word (release, system
//      scope): "rank r's partial of epoch e is in place";
//   3. every block waits until all n ready words hold the epoch (acquire), then sums the n staging buffers in
//      rank order 0..n-1 -- every rank adds the same numbers in the same order, so the result is bit-identical
//      everywhere -- and writes its own output.
// Two parities make a second barrier unnecessary: staging buffer e & 1 is overwritten by call e + 2, which a rank
// enters only after every peer signalled epoch e + 1 from its own call e + 1, i.e. after that peer's call e (the
// last reader of the buffer) has completed in stream order.  A spin that sees no progress for ~2 s sets an
// error word and gives up (a dead peer must not hang the GPU).
#include "common.cuh"
#include "kernels.h"

#define COMM_MAX_RANKS 16

struct dlco_comm {
    dlco_ctx *ctx;
    int n, rank;
    size_t max_count;
    float *stage[COMM_MAX_RANKS];        // rank h's staging area (2 * max_count floats), peer-mapped
    uint32_t *ready[COMM_MAX_RANKS];     // rank h's ready words (n words), peer-mapped
    uint32_t *d_state;                   // local: [0] block counter, [1] error word
    uint32_t epoch;
    int owns_ipc;                        // peers were opened with cudaIpcOpenMemHandle
    void *local_base;                    // one allocation: staging + ready words
};

struct CommArgs {
    float *stage[COMM_MAX_RANKS];
    uint32_t *ready[COMM_MAX_RANKS];
};

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(256)
k_peer_allreduce(const __grid_constant__ CommArgs a, int n, int rank, uint32_t epoch, size_t max_count, size_t count,
                 const float *__restrict__ in, float *__restrict__ out, uint32_t *__restrict__ state)
{
    const size_t off = (size_t)(epoch & 1) * max_count;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t)gridDim.x * blockDim.x;
    // 1. local partial -> staging (16-byte pieces where the pointers allow)
    float *mine = a.stage[rank] + off;
    const bool vec = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0 && (count & 3) == 0;
    if (vec) {
        for (size_t i = tid; i < count / 4; i += nth) reinterpret_cast<float4 *>(mine)[i] = reinterpret_cast<const float4 *>(in)[i];
    } else {
        for (size_t i = tid; i < count; i += nth) mine[i] = in[i];
    }
    // 2. the last block publishes the epoch to every peer
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&state[0], 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last) {
        if (threadIdx.x == 0) state[0] = 0;                      // for the next call (stream order)
        __threadfence_system();
        if ((int)threadIdx.x < n) st_release_sys(a.ready[threadIdx.x] + rank, epoch);
    }
    // 3. wait for every rank's partial of this epoch
    if ((int)threadIdx.x < n) {
        const uint32_t *w = a.ready[rank] + threadIdx.x;
        const long long t0 = clock64();
        // epochs only grow; a peer that is a whole call ahead has already left this epoch's buffer alone (parity)
        while ((int32_t)(ld_acquire_sys(w) - epoch) < 0) {
            if (clock64() - t0 > 4000000000LL) { atomicExch(&state[1], 1u); break; }
            __nanosleep(64);
        }
    }
    __syncthreads();
    // sum in rank order: identical bits on every rank
    if (vec) {
        for (size_t i = tid; i < count / 4; i += nth) {
            float4 s = reinterpret_cast<const float4 *>(a.stage[0] + off)[i];
            for (int h = 1; h < n; h++) {
                const float4 v = reinterpret_cast<const float4 *>(a.stage[h] + off)[i];
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            reinterpret_cast<float4 *>(out)[i] = s;
        }
    } else {
        for (size_t i = tid; i < count; i += nth) {
            float s = a.stage[0][off + i];
            for (int h = 1; h < n; h++) s += a.stage[h][off + i];
            out[i] = s;
        }
    }
}

static size_t comm_bytes(int n, size_t max_count) { return 2 * max_count * sizeof(float) + 256 + (size_t)n * sizeof(uint32_t); }
static uint32_t *comm_ready_of(void *base, size_t max_count) { return (uint32_t *)((uint8_t *)base + 2 * max_count * sizeof(float) + 256); }

static int comm_alloc(dlco_ctx *ctx, int n, int rank, size_t max_count, dlco_comm **out)
{
    DLCO_REQUIRE(ctx && out && n >= 1 && n <= COMM_MAX_RANKS && rank >= 0 && rank < n && max_count > 0,
                 "dlco_comm: bad arguments (n=%d rank=%d)", n, rank);
    DLCO_CUDA(cudaSetDevice(ctx->device));
    dlco_comm *c = (dlco_comm *)calloc(1, sizeof(dlco_comm));
    if (!c) return DLCO_ENOMEM;
    c->ctx = ctx; c->n = n; c->rank = rank;
    c->max_count = (max_count + 3) / 4 * 4;
    if (cudaMalloc(&c->local_base, comm_bytes(n, c->max_count)) != cudaSuccess ||
        cudaMalloc(&c->d_state, 2 * sizeof(uint32_t)) != cudaSuccess) {
        cudaFree(c->local_base);
        free(c);
        dlco_set_error("dlco_comm: out of device memory");
        return DLCO_ENOMEM;
    }
    DLCO_CUDA(cudaMemset(c->local_base, 0, comm_bytes(n, c->max_count)));
    DLCO_CUDA(cudaMemset(c->d_state, 0, 2 * sizeof(uint32_t)));
    c->stage[rank] = (float *)c->local_base;
    c->ready[rank] = comm_ready_of(c->local_base, c->max_count);
    *out = c;
    return DLCO_OK;
}

// ---- one process driving n GPUs (the C++ tools): peer access + direct pointers -------------------------------
extern "C" int dlco_comm_create_local(dlco_ctx **ctxs, int n, size_t max_count, dlco_comm **comms)
{
    DLCO_REQUIRE(ctxs && comms && n >= 1 && n <= COMM_MAX_RANKS, "dlco_comm_create_local: bad arguments");
    for (int g = 0; g < n; g++) comms[g] = nullptr;
    for (int g = 0; g < n; g++) {
        int r = comm_alloc(ctxs[g], n, g, max_count, &comms[g]);
        if (r != DLCO_OK) return r;
    }
    for (int g = 0; g < n; g++) {
        DLCO_CUDA(cudaSetDevice(ctxs[g]->device));
        for (int h = 0; h < n; h++) {
            if (h == g || ctxs[h]->device == ctxs[g]->device) continue;
            int can = 0;
            DLCO_CUDA(cudaDeviceCanAccessPeer(&can, ctxs[g]->device, ctxs[h]->device));
            DLCO_REQUIRE(can, "dlco_comm_create_local: GPU %d cannot access GPU %d's memory", ctxs[g]->device, ctxs[h]->device);
            cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[h]->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { dlco_set_error("cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e)); return DLCO_ECUDA; }
        }
        for (int h = 0; h < n; h++) {
            comms[g]->stage[h] = comms[h]->stage[h];
            comms[g]->ready[h] = comms[h]->ready[h];
        }
    }
    return DLCO_OK;
}

// ---- one process per GPU (torch.distributed launches): CUDA IPC ---------------------------------------------
extern "C" int dlco_comm_create(dlco_ctx *ctx, int n, int rank, size_t max_count, dlco_comm **comm)
{
    return comm_alloc(ctx, n, rank, max_count, comm);
}
// 64-byte handle of this rank's area, to be sent to every peer (e.g. torch.distributed.all_gather)
extern "C" int dlco_comm_export(dlco_comm *c, void *h_handle64)
{
    DLCO_REQUIRE(c && h_handle64, "dlco_comm_export: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DLCO_CUDA(cudaSetDevice(c->ctx->device));
    DLCO_CUDA(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)h_handle64, c->local_base));
    return DLCO_OK;
}
extern "C" int dlco_comm_import(dlco_comm *c, int peer, const void *h_handle64)
{
    DLCO_REQUIRE(c && h_handle64 && peer >= 0 && peer < c->n, "dlco_comm_import: bad arguments");
    if (peer == c->rank) return DLCO_OK;
    DLCO_CUDA(cudaSetDevice(c->ctx->device));
    void *base = nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, h_handle64, sizeof(h));
    DLCO_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    c->stage[peer] = (float *)base;
    c->ready[peer] = comm_ready_of(base, c->max_count);
    c->owns_ipc = 1;
    return DLCO_OK;
}

extern "C" int dlco_comm_destroy(dlco_comm *c)
{
    if (!c) return DLCO_OK;
    cudaSetDevice(c->ctx->device);
    cudaStreamSynchronize(c->ctx->stream);
    if (c->owns_ipc)
        for (int h = 0; h < c->n; h++)
            if (h != c->rank && c->stage[h]) cudaIpcCloseMemHandle(c->stage[h]);
    cudaFree(c->local_base);
    cudaFree(c->d_state);
    free(c);
    return DLCO_OK;
}

// out[i] = sum over ranks of in[i], i < count, on the context's stream; every rank must make the same sequence of
// calls.  d_in and d_out may be the same buffer.  Asynchronous; dlco_comm_check reports a timed-out wait.
extern "C" int dlco_comm_allreduce_f32(dlco_comm *c, const float *d_in, float *d_out, size_t count)
{
    DLCO_REQUIRE(c && d_in && d_out, "dlco_comm_allreduce_f32: NULL argument");
    DLCO_REQUIRE(count > 0 && count <= c->max_count, "dlco_comm_allreduce_f32: %zu elements, the communicator holds %zu", count,
                 c->max_count);
    for (int h = 0; h < c->n; h++) DLCO_REQUIRE(c->stage[h] && c->ready[h], "dlco_comm_allreduce_f32: peer %d was never imported", h);
    DLCO_CUDA(cudaSetDevice(c->ctx->device));
    CommArgs a;
    memcpy(a.stage, c->stage, sizeof(a.stage));
    memcpy(a.ready, c->ready, sizeof(a.ready));
    c->epoch++;
    // few blocks: the job is latency (a 1.18 MB matrix from 7 peers is ~10 us of NVLink time); all of them must be
    // co-resident for the spin-wait (they are: far fewer than one per SM)
    int blocks = (int)((count / 4 + 255) / 256);
    if (blocks < 1) blocks = 1;
    const int cap = c->ctx->sm_count < 64 ? c->ctx->sm_count : 64;
    if (blocks > cap) blocks = cap;
    k_peer_allreduce<<<blocks, 256, 0, c->ctx->stream>>>(a, c->n, c->rank, c->epoch, c->max_count, count, d_in, d_out, c->d_state);
    DLCO_CHECK_LAUNCH(c->ctx);
    return DLCO_OK;
}

// synchronises the context stream and reports whether any wait of this communicator timed out
extern "C" int dlco_comm_check(dlco_comm *c)
{
    DLCO_REQUIRE(c, "dlco_comm_check: NULL");
    DLCO_CUDA(cudaSetDevice(c->ctx->device));
    uint32_t st[2] = {0, 0};
    DLCO_CUDA(cudaMemcpyAsync(st, c->d_state, sizeof(st), cudaMemcpyDeviceToHost, c->ctx->stream));
    DLCO_CUDA(cudaStreamSynchronize(c->ctx->stream));
    DLCO_REQUIRE(st[1] == 0, "dlco_comm: rank %d timed out waiting for a peer", c->rank);
    return DLCO_OK;
}

// ---- the two index helpers a pair-sharded step needs around the exchange -----------------------------------------
__global__ void k_scatter_f32(const float *__restrict__ src, const int32_t *__restrict__ pos, int n, float *__restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[pos[i]] = src[i];
}
__global__ void k_gather_f32(const float *__restrict__ src, const int32_t *__restrict__ pos, int n, float *__restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[pos[i]];
}
// dst[0 .. n_dst) = 0, then dst[pos[i]] = src[i] for i < n  (each slot has one owner across the ranks)
extern "C" int dlco_scatter_f32(dlco_ctx *ctx, const float *d_src, const int32_t *d_pos, int n, float *d_dst, int n_dst)
{
    DLCO_REQUIRE(ctx && d_dst && n >= 0 && n_dst > 0 && (n == 0 || (d_src && d_pos)), "dlco_scatter_f32: bad arguments");
    DLCO_CUDA(cudaMemsetAsync(d_dst, 0, (size_t)n_dst * sizeof(float), ctx->stream));
    if (n == 0) return DLCO_OK;
    k_scatter_f32<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_src, d_pos, n, d_dst);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
extern "C" int dlco_gather_f32(dlco_ctx *ctx, const float *d_src, const int32_t *d_pos, int n, float *d_dst)
{
    DLCO_REQUIRE(ctx && n >= 0 && (n == 0 || (d_src && d_pos && d_dst)), "dlco_gather_f32: bad arguments");
    if (n == 0) return DLCO_OK;
    k_gather_f32<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_src, d_pos, n, d_dst);
    DLCO_CHECK_LAUNCH(ctx);
    return DLCO_OK;
}
