// kvm_peer.cuh - peer-memory exchange between the ranks of one NVSwitch node.
//
// Every rank owns one "exchange buffer" of identical layout, mapped into every other rank's
// address space (the host does the mapping; the kernels only see the base pointers).  The
// sharded analysis moves its small intermediate results with plain stores / 64-bit atomics
// into the peers' buffers over NVLink and synchronises with arrival counters - no collective
// library call on the data path.
#pragma once
#include <cstdint>

#include "../../include/kvmini_b200.h"

struct KvmPeerTable {
  unsigned char* base[KVM_PEER_MAX_WORLD];  // exchange buffer of every rank (own one included)
  int32_t world;
  int32_t rank;
};

static inline int kvm_peer_table(KvmPeerTable* pt, void* const* peers, int32_t world, int32_t rank) {
  if (!peers || world < 1 || world > KVM_PEER_MAX_WORLD || rank < 0 || rank >= world) return -1;
  for (int i = 0; i < KVM_PEER_MAX_WORLD; ++i)
    pt->base[i] = i < world ? reinterpret_cast<unsigned char*>(peers[i]) : nullptr;
  pt->world = world;
  pt->rank = rank;
  return 0;
}
