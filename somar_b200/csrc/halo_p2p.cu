// halo_p2p.cu -- y-halo exchange through NVLink peer memory (replaces pack -> ncclSend/ncclRecv -> unpack).
//
// The reference exchanges ghost cells with LevelData::exchange(Copier) (MPI point-to-point; GSRB.cpp:81,175,
// MappedAMRPoissonOp.cpp:2222-2238).  With NCCL that is five launches per exchange and ~65 us of latency, about sixty
// times per V-cycle; at 8 GPUs it was a quarter of the V-cycle.  Here every rank owns an "arena" (one cudaMalloc,
// exported with cudaIpcGetMemHandle, mapped by its two y-neighbours) that holds, per depth and side, two receive
// slots and two 64-bit counters:
//     push kernel:  wait until the neighbour has consumed message m-2 (ack counter in MY arena, written by the
//                   neighbour), store my boundary rows straight into the neighbour's slot m&1 over NVLink,
//                   fence, last block publishes  neighbour.ready = m;
//     pull kernel:  wait until  my.ready >= m, copy slot m&1 into my ghost rows (L1 bypassed), last block
//                   publishes  neighbour.ack = m.
// Two launches, no host involvement, no NCCL; every rank issues the same sequence of exchanges on its stream, so the
// message numbers agree without any handshake.  A wait that exceeds ~20 s traps instead of hanging the GPU.
// If the arena cannot be mapped on ANY rank (agreed by an all-reduce), all ranks stay on the NCCL path.
#include "level.hpp"
#include <cstdlib>
#include <cstring>
#include <time.h>

namespace somar {

typedef unsigned long long u64;

#define HALO_TPB 256
#define HALO_TIMEOUT_CYCLES (40ll * 1000 * 1000 * 1000)

__device__ __forceinline__ void spin_until(const volatile u64* flag, u64 want) {
    const long long t0 = clock64();
    while (*flag < want) {
        __nanosleep(32);
        if (clock64() - t0 > HALO_TIMEOUT_CYCLES) __trap();
    }
}

struct HaloSide {
    double* slot;        // push: the neighbour's receive slot; pull: my receive slot
    const u64* wait;     // push: my ack counter;               pull: my ready counter
    u64* post;           // push: the neighbour's ready counter; pull: the neighbour's ack counter
    u64 m;               // message number of this exchange on this channel
    int on;
};

// blockIdx.y = side.  Message layout = k_pack_y's: [row l][plane k][i], i over n0 + 2*GHOST, all planes.
__global__ void k_halo_push(const Geom g, const double* __restrict__ x, int w, HaloSide s0, HaloSide s1,
                            unsigned int* tickets) {
    const int side = blockIdx.y;
    const HaloSide s = side ? s1 : s0;
    if (!s.on) return;
    if (threadIdx.x == 0 && s.m > 2) spin_until(s.wait, s.m - 2);
    __syncthreads();
    const int nxx = g.n[0] + 2 * SOMAR_GHOST;
    const int nzz = g.n[2] + 2 * g.gh[2];
    const long long tot = (long long)nxx * nzz * w;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(t % nxx) - SOMAR_GHOST;
        const long long r = t / nxx;
        const int k = (int)(r % nzz) - g.gh[2];
        const int l = (int)(r / nzz);
        const int j = (side == 0) ? l : g.n[1] - w + l;   // my valid rows next to that neighbour, ascending y
        s.slot[t] = x[g.idx(i, j, k)];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(&tickets[side], 1u);
        if (t == gridDim.x - 1) {
            tickets[side] = 0;
            __threadfence_system();
            *(volatile u64*)s.post = s.m;
        }
    }
}

__global__ void k_halo_pull(const Geom g, double* __restrict__ x, int w, HaloSide s0, HaloSide s1,
                            unsigned int* tickets) {
    const int side = blockIdx.y;
    const HaloSide s = side ? s1 : s0;
    if (!s.on) return;
    if (threadIdx.x == 0) {
        spin_until(s.wait, s.m);
        __threadfence_system();
    }
    __syncthreads();
    const int nxx = g.n[0] + 2 * SOMAR_GHOST;
    const int nzz = g.n[2] + 2 * g.gh[2];
    const long long tot = (long long)nxx * nzz * w;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(t % nxx) - SOMAR_GHOST;
        const long long r = t / nxx;
        const int k = (int)(r % nzz) - g.gh[2];
        const int l = (int)(r / nzz);
        const int j = (side == 0) ? -w + l : g.n[1] + l;   // my ghost rows on that side, ascending y
        x[g.idx(i, j, k)] = __ldcg(s.slot + t);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(&tickets[2 + side], 1u);
        if (t == gridDim.x - 1) {
            tickets[2 + side] = 0;
            __threadfence_system();
            *(volatile u64*)s.post = s.m;
        }
    }
}

// ---- host side --------------------------------------------------------------------------------------------
static size_t slot_bytes(const Geom& g) {
    const size_t n = (size_t)SOMAR_GHOST * (g.n[0] + 2 * SOMAR_GHOST) * (g.n[2] + 2 * g.gh[2]) * sizeof(double);
    return (n + 255) / 256 * 256;
}

void halo_p2p_free(SomarLevel* L) {
    HaloArena& H = L->halo;
    if (H.state == 1) {
        cudaStreamSynchronize(L->cx.stream);
        for (int s = 0; s < 2; ++s)
            if (H.peer[s] && !(s == 1 && H.peer[1] == H.peer[0])) cudaIpcCloseMemHandle(H.peer[s]);
        // The last remote write into my arena is a neighbour's "ack" of my last message: wait (bounded, no
        // collective -- levels may be destroyed in any order) until every ack has caught up with what I sent.
        const int nd = (int)H.seq[0].size();
        std::vector<u64> flags((size_t)nd * 4);
        for (int tries = 0; tries < 200; ++tries) {
            if (cudaMemcpy(flags.data(), H.base + H.flagOff, flags.size() * sizeof(u64), cudaMemcpyDeviceToHost) !=
                cudaSuccess)
                break;
            bool done = true;
            for (int d = 0; d < nd; ++d)
                for (int s = 0; s < 2; ++s)
                    if (L->depths[d].g.nbr[s] >= 0 && flags[4 * d + 2 + s] < H.seq[s][d]) done = false;
            if (done) break;
            struct timespec ts = {0, 1000000};
            nanosleep(&ts, nullptr);
        }
    }
    if (H.base) cudaFree(H.base);
    H = HaloArena();
}

// Collective over the level's ranks (first exchange of level_finalize).  state: 1 enabled, -1 stay on NCCL.
static int halo_p2p_setup(SomarLevel* L) {
    HaloArena& H = L->halo;
    H.state = -1;
    const char* e = getenv("SOMAR_B200_HALO");
    const bool want = !(e && (e[0] == 'n' || e[0] == 'N' || e[0] == '0'));   // SOMAR_B200_HALO=nccl keeps NCCL
    const int nd = L->ndepths();
    const int N = L->desc.nranks;
    H.slotOff.resize(nd);
    H.slotBytes.resize(nd);
    size_t off = 0;
    for (int d = 0; d < nd; ++d) {
        H.slotBytes[d] = slot_bytes(L->depths[d].g);
        H.slotOff[d] = off;
        off += 4 * H.slotBytes[d];
    }
    H.flagOff = off;
    off += (size_t)nd * 4 * sizeof(u64);
    H.ticketOff = off;
    off += (size_t)nd * 4 * sizeof(unsigned int);
    H.bytes = (off + 255) / 256 * 256;
    H.seq[0].assign(nd, 0);
    H.seq[1].assign(nd, 0);
    int ok = want ? 1 : 0;
    cudaIpcMemHandle_t mine;
    std::memset(&mine, 0, sizeof(mine));
    if (ok && (cudaMalloc(&H.base, H.bytes) != cudaSuccess || cudaMemset(H.base, 0, H.bytes) != cudaSuccess ||
               cudaDeviceSynchronize() != cudaSuccess || cudaIpcGetMemHandle(&mine, H.base) != cudaSuccess)) {
        cudaGetLastError();
        ok = 0;
    }
    // all-gather the handles (64 bytes each) through the level's communicator
    char *dsend = nullptr, *dall = nullptr;
    std::vector<cudaIpcMemHandle_t> all(N);
    SOMAR_CUDA(cudaMalloc(&dsend, sizeof(mine)));
    SOMAR_CUDA(cudaMalloc(&dall, sizeof(mine) * N));
    SOMAR_CUDA(cudaMemcpyAsync(dsend, &mine, sizeof(mine), cudaMemcpyHostToDevice, L->cx.stream));
    SOMAR_NCCL(ncclAllGather(dsend, dall, sizeof(mine), ncclChar, L->comm, L->cx.stream));
    SOMAR_CUDA(cudaMemcpyAsync(all.data(), dall, sizeof(mine) * N, cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    cudaFree(dsend);
    cudaFree(dall);
    const Geom& g0 = L->depths[0].g;
    for (int s = 0; s < 2 && ok; ++s) {
        const int nb = g0.nbr[s];
        if (nb < 0) continue;
        if (s == 1 && nb == g0.nbr[0]) { H.peer[1] = H.peer[0]; continue; }   // two ranks, periodic y: one peer
        void* p = nullptr;
        if (cudaIpcOpenMemHandle(&p, all[nb], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            ok = 0;
        }
        H.peer[s] = static_cast<char*>(p);
    }
    // everybody or nobody
    double flag = ok ? 1.0 : 0.0;
    SOMAR_CUDA(cudaMemcpyAsync(L->d_scal + 7, &flag, sizeof(double), cudaMemcpyHostToDevice, L->cx.stream));
    SOMAR_NCCL(ncclAllReduce(L->d_scal + 7, L->d_scal + 7, 1, ncclDouble, ncclMin, L->comm, L->cx.stream));
    SOMAR_CUDA(cudaMemcpyAsync(&flag, L->d_scal + 7, sizeof(double), cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    if (flag == 1.0) {
        H.state = 1;
    } else {
        for (int s = 0; s < 2; ++s)
            if (H.peer[s] && !(s == 1 && H.peer[1] == H.peer[0])) cudaIpcCloseMemHandle(H.peer[s]);
        H.peer[0] = H.peer[1] = nullptr;
    }
    return SOMAR_OK;
}

// 1: the exchange was done here; 0: caller uses the NCCL path; < 0: error
int halo_p2p_exchange(SomarLevel* L, int depth, double* x, int width) {
    HaloArena& H = L->halo;
    if (H.state == 0) {
        int rc = halo_p2p_setup(L);
        if (rc != SOMAR_OK) return -1;
    }
    if (H.state != 1) return 0;
    const Geom& g = L->depths[depth].g;
    HaloSide push[2], pull[2];
    for (int s = 0; s < 2; ++s) {
        const u64 m = ++H.seq[s][depth];
        push[s].on = pull[s].on = g.nbr[s] >= 0;
        push[s].m = pull[s].m = m;
        if (g.nbr[s] < 0) {
            push[s].slot = pull[s].slot = nullptr;
            push[s].wait = pull[s].wait = nullptr;
            push[s].post = pull[s].post = nullptr;
            continue;
        }
        char* mine = H.base;
        char* peer = H.peer[s];
        const size_t sb = H.slotBytes[depth], so = H.slotOff[depth];
        u64* myFlags = reinterpret_cast<u64*>(mine + H.flagOff) + 4 * depth;     // ready[0], ready[1], ack[0], ack[1]
        u64* peerFlags = reinterpret_cast<u64*>(peer + H.flagOff) + 4 * depth;
        // my side-s rows are the neighbour's side-(1-s) ghosts
        push[s].slot = reinterpret_cast<double*>(peer + so + ((1 - s) * 2 + (m & 1)) * sb);
        push[s].wait = myFlags + 2 + s;
        push[s].post = peerFlags + (1 - s);
        pull[s].slot = reinterpret_cast<double*>(mine + so + (s * 2 + (m & 1)) * sb);
        pull[s].wait = myFlags + s;
        pull[s].post = peerFlags + 2 + (1 - s);
    }
    unsigned int* tickets = reinterpret_cast<unsigned int*>(H.base + H.ticketOff) + 4 * depth;
    const long long tot = (long long)(g.n[0] + 2 * SOMAR_GHOST) * (g.n[2] + 2 * g.gh[2]) * width;
    long long nb = (tot + HALO_TPB * 4 - 1) / (HALO_TPB * 4);
    if (nb > 2 * L->cx.sms) nb = 2 * L->cx.sms;
    if (nb < 1) nb = 1;
    const dim3 grid((unsigned)nb, 2);
    k_halo_push<<<grid, HALO_TPB, 0, L->cx.stream>>>(g, x, width, push[0], push[1], tickets);
    k_halo_pull<<<grid, HALO_TPB, 0, L->cx.stream>>>(g, x, width, pull[0], pull[1], tickets);
    L->cx.launches += 2;
    return 1;
}

}  // namespace somar
