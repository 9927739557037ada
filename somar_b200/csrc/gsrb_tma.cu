// gsrb_tma.cu -- the fused red+black Gauss-Seidel ITERATION of gsrb_fused.cu (same tile, same z-march, same
// arithmetic: bit-identical results) with the B200 data path: every byte comes from HBM through TMA
// (cp.async.bulk.tensor.3d) into multi-stage shared-memory rings, completion is signalled on mbarriers, and the
// 512 compute threads never issue a global LOAD.
//
// Why (profiles/r1_ncu_k_gsrb_fused_v5.txt): the register-prefetch version is latency/barrier bound -- 63 % of
// the HBM peak, issue slots 33 % busy, top stalls "barrier" and "long scoreboard": 16 warps run in lock step
// (two __syncthreads per plane) and the registers can hold only one step of loads in flight.  Here
//   * one producer thread (warp 16) keeps the loads of the next step(s) in flight at all times:
//       phi planes      ring of 5  (66 x 18 box incl. the 1-cell ring), plane k+3 is requested at the end of step k
//       coefficients    ring of 3  (J g^xx 66x16, J g^yy 64x17, 1/J, lapDiag, rhs 64x16), stage k+2
//       J g^zz planes   ring of 4  (64x16; plane k is the low z-face of cell plane k), plane k+3
//   * there is no CTA-wide barrier in the steady state.  The only cross-thread dependence of the sweep is
//     "black update of plane k reads the red values of plane k written by OTHER threads one step earlier"
//     (its z-neighbours are the thread's own red cells), so a split barrier does it: every warp ARRIVES on
//     `redDone` after its red update of step k and WAITS for it at the top of step k+1, with the black update of
//     plane k-1 and the stores in between;
//   * a ring slot is handed back to the producer through two alternating `empty` mbarriers.
// Shared memory: 5 x 9.4 KB + 3 x 40.75 KB + 4 x 8 KB = 201 KB, one CTA (544 threads) per SM.
//
// Step k of a CTA (k = k0-1 .. k1), consumers:
//   wait redDone(k-1), full(phi k+1), full(coef k), full(Jgzz k+1)
//   red update of plane k on the 64x16 extended tile (in place in the phi ring)       -> arrive redDone(k)
//   black update of plane k-1 on the 62x14 core tile -> HBM, together with the red values of plane k-1
//   fence.proxy.async; arrive empty(k)   => producer: phi plane k-2, coef stage k-1, Jgzz plane k-1 are dead
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#include "kernels.hpp"
#include "stencil_ortho.cuh"

namespace somar {

namespace {

constexpr int EX = 64, EY = 16;          // extended tile (cells)
constexpr int CX = EX - 2, CY = EY - 2;  // core tile 62 x 14
constexpr int PX = EX + 2, PY = EY + 2;  // phi plane in shared memory 66 x 18
constexpr int NPHI = 5, NCF = 3, NZF = 4;
constexpr int CONSUMERS = (EX / 2) * EY; // 512: thread = x-pair of one row
constexpr int NTHREADS = CONSUMERS + 32; // + producer warp
constexpr int NWARPS = CONSUMERS / 32;

constexpr int PHI_BYTES = PX * PY * 8;         // 9504 bytes land per plane
constexpr int PHI_SLOT = 9600;                 // slot pitch, multiple of 128
constexpr int JG0_BYTES = EY * PX * 8;         // 66 columns: column 64 is the high face of the last cell
constexpr int JG1_BYTES = (EY + 1) * EX * 8;   // 17 rows
constexpr int CC_BYTES = EY * EX * 8;
constexpr int CF_BYTES = JG0_BYTES + JG1_BYTES + 3 * CC_BYTES;   // 41728
constexpr int OFF_PHI = 0;
constexpr int OFF_CF = OFF_PHI + NPHI * PHI_SLOT;
constexpr int OFF_ZF = OFF_CF + NCF * CF_BYTES;
constexpr int OFF_BAR = OFF_ZF + NZF * CC_BYTES;
constexpr int NBAR = NPHI + NCF + NZF + 1 + 2;
constexpr int SMEM_BYTES = OFF_BAR + NBAR * 8;
static_assert(OFF_CF % 128 == 0 && OFF_ZF % 128 == 0 && OFF_BAR % 8 == 0, "TMA destinations must be 128-byte aligned");
static_assert(JG0_BYTES % 128 == 0 && JG1_BYTES % 128 == 0 && CC_BYTES % 128 == 0 && CF_BYTES % 128 == 0, "alignment");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

struct CfStage {
    double jg0[EY][PX];
    double jg1[EY + 1][EX];
    double ji[EY][EX];
    double lap[EY][EX];
    double rhs[EY][EX];
};
static_assert(sizeof(CfStage) == CF_BYTES, "stage layout");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(b)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* tm, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_u32(dst)), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
        : "memory");
}

}  // namespace

__global__ void __launch_bounds__(NTHREADS, 1)
k_gsrb_tma(const __grid_constant__ CUtensorMap tmPhi, const __grid_constant__ CUtensorMap tmJg0,
           const __grid_constant__ CUtensorMap tmJg1, const __grid_constant__ CUtensorMap tmJg2,
           const __grid_constant__ CUtensorMap tmJi, const __grid_constant__ CUtensorMap tmLap,
           const __grid_constant__ CUtensorMap tmRhs, const Geom g, double* __restrict__ phiOut, int kc) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* const fullPhi = bars;
    uint64_t* const fullCf = fullPhi + NPHI;
    uint64_t* const fullZf = fullCf + NCF;
    uint64_t* const redDone = fullZf + NZF;
    uint64_t* const empty = redDone + 1;

    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * CX - 1;      // local x of extended column 0
    const int y0 = blockIdx.y * CY - 1;
    const int k0 = blockIdx.z * kc;
    const int k1 = min(k0 + kc, g.n[2]);

    if (tid == 0) {
        for (int i = 0; i < NPHI + NCF + NZF; ++i) mbar_init(bars + i, 1);
        mbar_init(redDone, NWARPS);
        mbar_init(empty, NWARPS);
        mbar_init(empty + 1, NWARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    __syncthreads();

    if (tid >= CONSUMERS) {
        // ---------------------------------------------------------------------------------------------------
        // producer: one thread
        // ---------------------------------------------------------------------------------------------------
        if (tid != CONSUMERS) return;
        const int cxP = x0 - 1 + g.gh[0], cyP = y0 - 1 + g.gh[1];   // phi box origin (allocation coordinates)
        const int cxC = x0 + g.gh[0], cyC = y0 + g.gh[1];           // coefficient box origin
        const int cz = g.gh[2];
        auto issue_phi = [&](int p) {
            const int slot = (p - (k0 - 2)) % NPHI;
            mbar_expect_tx(fullPhi + slot, PHI_BYTES);
            tma_load_3d(smem + OFF_PHI + slot * PHI_SLOT, &tmPhi, cxP, cyP, p + cz, fullPhi + slot);
        };
        auto issue_cf = [&](int s) {
            const int slot = (s - (k0 - 1)) % NCF;
            CfStage* st = reinterpret_cast<CfStage*>(smem + OFF_CF + slot * CF_BYTES);
            mbar_expect_tx(fullCf + slot, CF_BYTES);
            tma_load_3d(st->jg0, &tmJg0, cxC, cyC, s + cz, fullCf + slot);
            tma_load_3d(st->jg1, &tmJg1, cxC, cyC, s + cz, fullCf + slot);
            tma_load_3d(st->ji, &tmJi, cxC, cyC, s + cz, fullCf + slot);
            tma_load_3d(st->lap, &tmLap, cxC, cyC, s + cz, fullCf + slot);
            tma_load_3d(st->rhs, &tmRhs, cxC, cyC, s + cz, fullCf + slot);
        };
        auto issue_zf = [&](int p) {
            const int slot = (p - (k0 - 1)) % NZF;
            mbar_expect_tx(fullZf + slot, CC_BYTES);
            tma_load_3d(smem + OFF_ZF + slot * CC_BYTES, &tmJg2, cxC, cyC, p + cz, fullZf + slot);
        };
        // prologue: everything steps k0-1 and k0 need
        issue_phi(k0 - 2);
        issue_phi(k0 - 1);
        issue_phi(k0);
        issue_zf(k0 - 1);
        issue_zf(k0);
        issue_cf(k0 - 1);
        issue_phi(k0 + 1);
        issue_zf(k0 + 1);
        if (k0 <= k1) issue_cf(k0);
        for (int k = k0 - 1, j = 0; k + 3 <= k1 + 1; ++k, ++j) {
            mbar_wait(empty + (j & 1), (j >> 1) & 1);   // all consumers are done with step k
            fence_proxy_async();
            if (k + 3 <= k1 + 1) issue_phi(k + 3);
            if (k + 2 <= k1) issue_cf(k + 2);
            if (k + 3 <= k1 + 1) issue_zf(k + 3);
        }
        return;
    }

    // -------------------------------------------------------------------------------------------------------
    // consumers: thread (t, r) owns the x-pair (2t, 2t+1) of row r of the extended tile
    // -------------------------------------------------------------------------------------------------------
    const int t = tid & 31, r = tid >> 5;
    const int isz = (int)g.sz;
    const int y = y0 + r;
    const bool yin = (y >= 0) && (y < g.n[1]);
    const bool ycore = (r >= 1) && (r <= CY);
    const int parBase = g.g0[0] + x0 + g.g0[1] + y + g.g0[2];
    const int cbase = (int)g.idx(x0 + 2 * t, y, 0);
    int colMask = 0, coreMask = 0;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int ex = 2 * t + p, x = x0 + ex;
        const bool col = yin && x >= 0 && x < g.n[0];
        if (col) colMask |= 1 << p;
        if (col && ycore && ex >= 1 && ex <= CX) coreMask |= 1 << p;
    }
    // block-uniform: does the extended tile touch the domain rim in x or y?
    const bool tileRim = (g.g0[0] + x0 <= 0) || (g.g0[0] + x0 + EX - 1 >= g.gn[0] - 1) || (g.g0[1] + y0 <= 0) ||
                         (g.g0[1] + y0 + EY - 1 >= g.gn[1] - 1);
    const int ctr = (r + 1) * PX + 1;   // + ex: this thread's cell in a phi plane
    const int cfi = r * EX;             // + ex: this thread's cell in a 64-wide coefficient plane

    mbar_wait(fullPhi + 0, 0);   // planes k0-2, k0-1
    mbar_wait(fullPhi + 1, 0);
    mbar_wait(fullZf + 0, 0);    // J g^zz plane k0-1
    int kOff = (k0 - 1) * isz;
    for (int k = k0 - 1, j = 0; k <= k1; ++k, ++j, kOff += isz) {
        if (j > 0) mbar_wait(redDone, (j - 1) & 1);             // everybody's red values of plane k-1 are in place
        mbar_wait(fullPhi + (j + 2) % NPHI, ((j + 2) / NPHI) & 1);   // phi plane k+1
        mbar_wait(fullCf + j % NCF, (j / NCF) & 1);                  // coefficients of plane k
        mbar_wait(fullZf + (j + 1) % NZF, ((j + 1) / NZF) & 1);      // J g^zz plane k+1

        double* const Pk = reinterpret_cast<double*>(smem + OFF_PHI + ((j + 1) % NPHI) * PHI_SLOT);
        const double* const Pm = reinterpret_cast<const double*>(smem + OFF_PHI + (j % NPHI) * PHI_SLOT);
        const double* const Pp = reinterpret_cast<const double*>(smem + OFF_PHI + ((j + 2) % NPHI) * PHI_SLOT);
        const double* const Pmm = reinterpret_cast<const double*>(smem + OFF_PHI + ((j + 4) % NPHI) * PHI_SLOT);
        const double* const Zk = reinterpret_cast<const double*>(smem + OFF_ZF + (j % NZF) * CC_BYTES);
        const double* const Zp = reinterpret_cast<const double*>(smem + OFF_ZF + ((j + 1) % NZF) * CC_BYTES);
        const double* const Zm = reinterpret_cast<const double*>(smem + OFF_ZF + ((j + 3) % NZF) * CC_BYTES);
        const CfStage* const C = reinterpret_cast<const CfStage*>(smem + OFF_CF + (j % NCF) * CF_BYTES);
        const CfStage* const Cm = reinterpret_cast<const CfStage*>(smem + OFF_CF + ((j + 2) % NCF) * CF_BYTES);

        const int par = (parBase + k) & 1, parn = par ^ 1;
        const int ex = 2 * t + par;
        const int x = x0 + ex;
        // red update of plane k (extended tile), in place in the ring
        if (k >= 0 && k < g.n[2] && ((colMask >> par) & 1)) {
            const int gk = g.g0[2] + k;
            const bool fast = !tileRim && gk != 0 && gk != g.gn[2] - 1;
            GsrbCoef q;
            q.glo[0] = C->jg0[r][ex];
            q.ghi[0] = C->jg0[r][ex + 1];
            q.glo[1] = C->jg1[r][ex];
            q.ghi[1] = C->jg1[r + 1][ex];
            q.glo[2] = Zk[cfi + ex];
            q.ghi[2] = Zp[cfi + ex];
            q.ji = C->ji[r][ex];
            q.lap = C->lap[r][ex];
            q.rhs = C->rhs[r][ex];
            double nb[3][2];
            nb[0][0] = Pk[ctr + ex - 1];
            nb[0][1] = Pk[ctr + ex + 1];
            nb[1][0] = Pk[ctr + ex - PX];
            nb[1][1] = Pk[ctr + ex + PX];
            nb[2][0] = Pm[ctr + ex];
            nb[2][1] = Pp[ctr + ex];
            Pk[ctr + ex] = gsrb_update_vals<3>(g, q, x, y, k, Pk[ctr + ex], nb, fast);
        }
        __syncwarp();
        if (t == 0) mbar_arrive(redDone);
        // black update of plane k-1 (core tile): x/y neighbours are red values of step k-1 (redDone(k-1) above),
        // z neighbours are this thread's own red cells of planes k-2 and k
        if (k - 1 >= k0 && k - 1 < k1) {
            const int kb = k - 1;
            const int cb0 = cbase + kOff - isz;
            if ((coreMask >> par) & 1) {
                const int gk = g.g0[2] + kb;
                const bool fast = !tileRim && gk != 0 && gk != g.gn[2] - 1;
                GsrbCoef q;
                q.glo[0] = Cm->jg0[r][ex];
                q.ghi[0] = Cm->jg0[r][ex + 1];
                q.glo[1] = Cm->jg1[r][ex];
                q.ghi[1] = Cm->jg1[r + 1][ex];
                q.glo[2] = Zm[cfi + ex];
                q.ghi[2] = Zk[cfi + ex];
                q.ji = Cm->ji[r][ex];
                q.lap = Cm->lap[r][ex];
                q.rhs = Cm->rhs[r][ex];
                double nb[3][2];
                nb[0][0] = Pm[ctr + ex - 1];
                nb[0][1] = Pm[ctr + ex + 1];
                nb[1][0] = Pm[ctr + ex - PX];
                nb[1][1] = Pm[ctr + ex + PX];
                nb[2][0] = Pmm[ctr + ex];
                nb[2][1] = Pk[ctr + ex];
                phiOut[cb0 + par] = gsrb_update_vals<3>(g, q, x, y, kb, Pm[ctr + ex], nb, fast);
            }
            // the other cell of the pair is red in plane k-1: its new value sits in the ring
            if ((coreMask >> parn) & 1) phiOut[cb0 + parn] = Pm[ctr + 2 * t + parn];
        }
        // hand phi plane k-2, coefficient stage k-1 and J g^zz plane k-1 back to the producer
        fence_proxy_async();
        __syncwarp();
        if (t == 0) mbar_arrive(empty + (j & 1));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host side: tensor maps
// ---------------------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            (void)cudaGetLastError();
    }
    return fn;
}

// 3-D map over one brick array: dims (pitch, rows, planes), box (bx, by, 1)
bool make_map(CUtensorMap* tm, const Geom& g, const double* base, int bx, int by) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)g.sy, (cuuint64_t)(g.n[1] + 2 * g.gh[1]), (cuuint64_t)(g.n[2] + 2 * g.gh[2])};
    const cuuint64_t strides[2] = {(cuuint64_t)g.sy * 8, (cuuint64_t)g.sz * 8};
    const cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    return fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base), dims, strides, box, es,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct TmaCache {
    CUtensorMap jg0, jg1, jg2, ji, lap;
    const double* key[5];
};

}  // namespace

bool gsrb_tma_usable(const Geom& g, const Metric& m) {
    if (!gsrb_fused_usable(g, m)) return false;
    if (g.gh[0] < 2 || g.gh[1] < 2 || g.gh[2] < 2) return false;
    return encode_fn() != nullptr;
}

void gsrb_tma_free(void* cache) { delete static_cast<TmaCache*>(cache); }

bool launch_gsrb_tma(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn, const double* rhs,
                     void** cache) {
    static bool attr = false;
    if (!attr) {
        if (cudaFuncSetAttribute(k_gsrb_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) {
            (void)cudaGetLastError();
            return false;
        }
        attr = true;
    }
    TmaCache* tc = static_cast<TmaCache*>(*cache);
    const double* key[5] = {m.jg[0][0], m.jg[1][1], m.jg[2][2], m.jinv, m.lapdiag};
    bool fresh = tc != nullptr;
    if (tc)
        for (int i = 0; i < 5; ++i) fresh = fresh && tc->key[i] == key[i];
    if (!fresh) {
        if (!tc) *cache = tc = new TmaCache();
        if (!make_map(&tc->jg0, g, key[0], PX, EY) || !make_map(&tc->jg1, g, key[1], EX, EY + 1) ||
            !make_map(&tc->jg2, g, key[2], EX, EY) || !make_map(&tc->ji, g, key[3], EX, EY) ||
            !make_map(&tc->lap, g, key[4], EX, EY))
            return false;
        for (int i = 0; i < 5; ++i) tc->key[i] = key[i];
    }
    CUtensorMap tmPhi, tmRhs;
    if (!make_map(&tmPhi, g, phiIn, PX, PY) || !make_map(&tmRhs, g, rhs, EX, EY)) return false;
    const int kc = (g.n[2] > 64) ? 64 : g.n[2];
    dim3 grid((g.n[0] + CX - 1) / CX, (g.n[1] + CY - 1) / CY, (g.n[2] + kc - 1) / kc);
    k_gsrb_tma<<<grid, NTHREADS, SMEM_BYTES, cx.stream>>>(tmPhi, tc->jg0, tc->jg1, tc->jg2, tc->ji, tc->lap, tmRhs, g,
                                                          phiOut, kc);
    cx.launches++;
    return true;
}

}  // namespace somar
