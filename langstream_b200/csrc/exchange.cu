// lsc_exchange: one GPU's end of the candidate exchange of a row-sharded top-k search.
//
// The corpus behind `query-vector-db` is sharded by rows, one shard per GPU; a query batch is answered by every
// shard and the per-shard top-k lists have to meet before the final merge.  The payload is tiny (B x k records
// of 16 bytes, 160 B to 160 KiB per shard), so what matters is latency: instead of a collective call between two
// kernel launches, the scan kernel itself stores its records into every peer's receive buffer over NVLink
// (peer-to-peer mapped memory, plain stores), raises a flag per query and merges the peers' records as soon as
// their flags arrive (topk_fuse.cuh).  This file owns the memory that makes this possible: a receive buffer +
// flag array per GPU, mapped into the other GPUs either by cudaDeviceEnablePeerAccess (all GPUs driven by one
// process: the sharded handle in sharded_index.cu) or by CUDA IPC handles (one process per GPU: torchrun / one
// agent pod per GPU; the handles travel over whatever channel the processes already share).
// No counterpart in the reference: HerdDB answers the query inside one JVM.
#include <cstring>
#include <mutex>

#include "common.h"
#include "exchange.h"
#include "topk.h"

using namespace lsc;

namespace lsc {

size_t exchange_block_bytes(int world, int xB) {
    const size_t rec = 2ull * world * xB * TKX_MAX_XK * sizeof(TopkRecord);
    const size_t flg = 2ull * world * xB * sizeof(uint32_t);
    return rec + flg;
}

void exchange_fill(const lsc_exchange* x, TopkFuse* f, uint32_t epoch) {
    f->world = x->world;
    f->rank = x->rank;
    f->epoch = epoch;
    f->xB = x->xB;
    f->xk = TKX_MAX_XK;
    for (int r = 0; r < x->world; ++r) {
        f->recv[r] = reinterpret_cast<TopkRecord*>(x->peer_base[r]);
        f->flags[r] = reinterpret_cast<uint32_t*>(x->peer_base[r] + 2ull * x->world * x->xB * TKX_MAX_XK * sizeof(TopkRecord));
    }
}

}  // namespace lsc

extern "C" {

int lsc_exchange_create(int device, int32_t rank, int32_t world, int32_t max_queries, lsc_exchange** out) {
    if (out == nullptr) return fail(LSC_E_INVALID, "lsc_exchange_create: out is null");
    *out = nullptr;
    if (world < 1 || world > TKX_MAX_WORLD) return fail(LSC_E_UNSUPPORTED, "lsc_exchange_create: world must be in [1,8]");
    if (rank < 0 || rank >= world) return fail(LSC_E_INVALID, "lsc_exchange_create: rank outside [0, world)");
    if (max_queries < 1 || max_queries > (1 << 16)) return fail(LSC_E_INVALID, "lsc_exchange_create: max_queries must be in [1, 65536]");
    int ndev = 0;
    LSC_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(LSC_E_INVALID, "lsc_exchange_create: no such device");
    LSC_CUDA(cudaSetDevice(device));
    LSC_TRY(topk_preload_kernels());  // nothing may be left to load once a peer's scan waits for this GPU
    lsc_exchange* x = new lsc_exchange();
    x->device = device;
    x->rank = rank;
    x->world = world;
    x->xB = max_queries;
    x->bytes = exchange_block_bytes(world, max_queries);
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&x->base), x->bytes);
    if (e == cudaSuccess) e = cudaMemset(x->base, 0, x->bytes);  // flags start at epoch 0 = "nothing yet"
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        if (x->base) cudaFree(x->base);
        delete x;
        return fail_cuda(e, "cudaMalloc/cudaMemset(exchange block)", __FILE__, __LINE__);
    }
    x->peer_base[rank] = x->base;
    x->connected = world == 1;
    *out = x;
    return LSC_OK;
}

int lsc_exchange_ipc_handle(lsc_exchange* x, void* handle_out) {
    if (x == nullptr || handle_out == nullptr) return fail(LSC_E_INVALID, "lsc_exchange_ipc_handle: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == LSC_EXCHANGE_HANDLE_BYTES, "IPC handle size");
    LSC_CUDA(cudaSetDevice(x->device));
    cudaIpcMemHandle_t h;
    LSC_CUDA(cudaIpcGetMemHandle(&h, x->base));
    memcpy(handle_out, &h, sizeof(h));
    return LSC_OK;
}

int lsc_exchange_connect_ipc(lsc_exchange* x, const void* all_handles) {
    if (x == nullptr || all_handles == nullptr) return fail(LSC_E_INVALID, "lsc_exchange_connect_ipc: null argument");
    if (x->connected) return fail(LSC_E_INVALID, "lsc_exchange_connect_ipc: already connected");
    LSC_CUDA(cudaSetDevice(x->device));
    const unsigned char* hs = static_cast<const unsigned char*>(all_handles);
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, hs + static_cast<size_t>(r) * LSC_EXCHANGE_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        LSC_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        x->peer_base[r] = static_cast<unsigned char*>(p);
        x->ipc_opened[r] = true;
    }
    x->connected = true;
    return LSC_OK;
}

int lsc_exchange_connect_local(lsc_exchange** endpoints, int32_t world) {
    if (endpoints == nullptr || world < 1 || world > TKX_MAX_WORLD) return fail(LSC_E_INVALID, "lsc_exchange_connect_local: bad argument");
    for (int r = 0; r < world; ++r)
        if (endpoints[r] == nullptr || endpoints[r]->world != world || endpoints[r]->rank != r || endpoints[r]->xB != endpoints[0]->xB)
            return fail(LSC_E_INVALID, "lsc_exchange_connect_local: endpoints must be ranks 0..world-1 of one group");
    for (int a = 0; a < world; ++a) {
        LSC_CUDA(cudaSetDevice(endpoints[a]->device));
        for (int b = 0; b < world; ++b) {
            if (a == b) continue;
            if (endpoints[a]->device != endpoints[b]->device) {
                int can = 0;
                LSC_CUDA(cudaDeviceCanAccessPeer(&can, endpoints[a]->device, endpoints[b]->device));
                if (!can) return fail(LSC_E_UNSUPPORTED, "lsc_exchange_connect_local: the GPUs cannot address each other's memory (no P2P)");
                cudaError_t e = cudaDeviceEnablePeerAccess(endpoints[b]->device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) return fail_cuda(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
            }
            endpoints[a]->peer_base[b] = endpoints[b]->base;
        }
        endpoints[a]->connected = true;
    }
    return LSC_OK;
}

void lsc_exchange_destroy(lsc_exchange* x) {
    if (x == nullptr) return;
    cudaSetDevice(x->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < x->world; ++r)
        if (x->ipc_opened[r]) cudaIpcCloseMemHandle(x->peer_base[r]);
    if (x->base) cudaFree(x->base);
    delete x;
}

}  // extern "C"
