// Internals of the one-shot NVLink allreduce shared by p2p_allreduce.cu (communicator, reduce kernel)
// and gemv.cu (the GEMV epilogue that pushes its partial rows straight into the peers' buffers).
#pragma once
#include <stddef.h>
#include <vector>

namespace b200 {
constexpr int kP2PThreads = 256;
constexpr int kP2PMaxCtas = 64;
constexpr int kP2PMaxWorld = 8;
constexpr size_t kP2PHeader = 256;   // [0] call counter, [1] ticket

// what a kernel needs to reach the peers: bases of every rank's allocation as mapped in this process
struct P2PWire {
    unsigned char* peer[kP2PMaxWorld];
    int world, rank;
    size_t slot_bytes;   // capacity of one receive slot (16-byte lines carrying 8 payload bytes)
};
}  // namespace b200

struct b200_p2p {
    int world = 0, rank = 0;
    size_t slot_bytes = 0, total = 0;
    unsigned char* local = nullptr;
    std::vector<unsigned char*> peers;   // mapped bases (peers[rank] == local)
    std::vector<bool> opened;
    bool connected = false;
};

namespace b200 {
inline P2PWire p2p_wire(const b200_p2p* c) {
    P2PWire w{};
    for (int r = 0; r < c->world; ++r) w.peer[r] = c->peers[r];
    w.world = c->world;
    w.rank = c->rank;
    w.slot_bytes = c->slot_bytes;
    return w;
}
}  // namespace b200
