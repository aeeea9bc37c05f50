// One GPU's end of the sharded top-k's candidate exchange (exchange.cu); used by index.cu and sharded_index.cu.
#pragma once
#include "common.h"
#include "topk_fuse.cuh"

struct lsc_exchange {
    int device = 0;
    int rank = 0, world = 1;
    int xB = 0;                      // queries per exchange (larger batches go in slices)
    size_t bytes = 0;
    unsigned char* base = nullptr;   // this GPU's block: records, then flags
    unsigned char* peer_base[lsc::TKX_MAX_WORLD] = {};  // every rank's block as addressable from this GPU
    bool ipc_opened[lsc::TKX_MAX_WORLD] = {};
    bool connected = false;
    uint32_t epoch = 0;              // searches done through this endpoint (all ranks count alike)
};

namespace lsc {
size_t exchange_block_bytes(int world, int xB);
// the peer pointers and strides of `x` for a search with sequence number `epoch`
void exchange_fill(const lsc_exchange* x, TopkFuse* f, uint32_t epoch);
}  // namespace lsc
