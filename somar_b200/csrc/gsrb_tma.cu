// gsrb_tma.cu -- the fused red+black Gauss-Seidel ITERATION of gsrb_fused.cu (same tile, same z-march, same
// arithmetic: bit-identical results) with the B200 data path: every byte comes from HBM through the TMA engine's
// bulk copies (cp.async.bulk global -> shared, SASS UBLKCP) into multi-stage shared-memory rings, completion is
// signalled on mbarriers (complete_tx), and the 512 compute threads never issue a global LOAD.
//
// (The tiled flavour, cp.async.bulk.tensor with a CUtensorMap, raises "illegal instruction" on this pool's B200
// boxes for every descriptor form tried -- tests/cuda/tma_probe.cu, modes 0-16 -- while plain bulk copies work; a
// tile row is contiguous in the brick, so one bulk copy per row does the same job: 115 copies of 528 B per step,
// issued by the 32 lanes of the producer warp.)
//
// Why (profiles/r1_ncu_k_gsrb_fused_v5.txt): the register-prefetch version is latency/barrier bound -- 63 % of
// the HBM peak, issue slots 33 % busy, top stalls "barrier" and "long scoreboard": 16 warps run in lock step
// (two __syncthreads per plane) and the registers can hold only one step of loads in flight.  Here
//   * the producer warp keeps the loads of the next step(s) in flight at all times:
//       phi planes      ring of 5  (18 rows x 66: the extended tile + 1-cell ring), plane k+3 requested at step k
//       coefficients    ring of 3  (J g^xx 16 rows, J g^yy 17 rows, 1/J, lapDiag, rhs 16 rows; all 66 wide), stage k+2
//       J g^zz planes   ring of 4  (16 rows x 66; plane k is the low z-face of cell plane k), plane k+3
//     every row starts at x0-1, which is 16-byte aligned in the brick (SOMAR_XO = 4, even tile pitch);
//   * there is no CTA-wide barrier in the steady state.  The only cross-thread dependence of the sweep is
//     "black update of plane k reads the red values of plane k written by OTHER threads one step earlier"
//     (its z-neighbours are the thread's own red cells), so a split barrier does it: every warp ARRIVES on
//     `redDone` after its red update of step k and WAITS for it at the top of step k+1, with the black update of
//     plane k-1 and the stores in between;
//   * a ring slot is handed back to the producer through two alternating `empty` mbarriers.
// Shared memory: 5 x 9.4 KB + 3 x 41.8 KB + 4 x 8.25 KB = 205 KB, one CTA (544 threads) per SM.
//
// Step k of a CTA (k = k0-1 .. k1), consumers:
//   wait redDone(k-1), full(phi k+1), full(coef k), full(Jgzz k+1)
//   red update of plane k on the 64x16 extended tile (in place in the phi ring)       -> arrive redDone(k)
//   black update of plane k-1 on the 62x14 core tile -> HBM, together with the red values of plane k-1
//   fence.proxy.async; arrive empty(k)   => producer: phi plane k-2, coef stage k-1, Jgzz plane k-1 are dead
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "kernels.hpp"
#include "stencil_ortho.cuh"

namespace somar {

namespace {

constexpr int EX = 64, EY = 16;          // extended tile (cells)
constexpr int CX = EX - 2, CY = EY - 2;  // core tile 62 x 14
constexpr int PX = EX + 2, PY = EY + 2;  // rows are 66 doubles wide: column c <-> x = x0 - 1 + c
constexpr int NPHI = 5, NCF = 3, NZF = 4;
constexpr int CONSUMERS = (EX / 2) * EY; // 512: thread = x-pair of one row
constexpr int MAXPROD = 7;               // producer warps (lane 0 of each issues copies)
constexpr int NTHREADS = CONSUMERS + 32 * MAXPROD;
constexpr int NWARPS = CONSUMERS / 32;
constexpr int ROWB = PX * 8;             // 528 bytes per row copy

constexpr int PHI_SLOT = 9600;           // 18 rows x 528 B = 9504, padded to a multiple of 128
// coefficient stage: row offsets of the five arrays
constexpr int R_JG0 = 0, R_JG1 = R_JG0 + EY, R_JI = R_JG1 + EY + 1, R_LAP = R_JI + EY, R_RHS = R_LAP + EY;
constexpr int CF_ROWS = R_RHS + EY;      // 81
constexpr int CF_BYTES = CF_ROWS * ROWB; // 42768
constexpr int ZF_BYTES = EY * ROWB;      // 8448
constexpr int OFF_PHI = 0;
constexpr int OFF_CF = OFF_PHI + NPHI * PHI_SLOT;
constexpr int OFF_ZF = OFF_CF + NCF * CF_BYTES;
constexpr int OFF_BAR = OFF_ZF + NZF * ZF_BYTES;
constexpr int NBAR = NPHI + NCF + NZF + 1 + 2;
constexpr int SMEM_BYTES = OFF_BAR + NBAR * 8;
static_assert(OFF_CF % 16 == 0 && OFF_ZF % 16 == 0 && OFF_BAR % 8 == 0 && ROWB % 16 == 0, "bulk copies need 16-byte alignment");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
static_assert(PY + CF_ROWS + EY == 115, "row copies per steady-state step");

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

// one row: global -> shared through the TMA engine, completion counted in bytes on `bar`
__device__ __forceinline__ void bulk_row(void* dst, const double* src, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(ROWB), "r"(smem_u32(bar))
                 : "memory");
}

// rows rr in [0, nrows) of a tile segment starting at local row yStart that exist in the brick
__device__ __forceinline__ int rows_lo(const Geom& g, int yStart) { return max(0, -g.gh[1] - yStart); }
__device__ __forceinline__ int rows_hi(const Geom& g, int yStart, int nrows) {
    return min(nrows, g.n[1] + g.gh[1] - yStart);
}

}  // namespace

__global__ void __launch_bounds__(NTHREADS, 1)
k_gsrb_tma(const Geom g, const Metric m, double* __restrict__ phiOut, const double* __restrict__ phi,
           const double* __restrict__ rhs, int kc, int nprod, int noload) {
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
        for (int i = 0; i < NPHI; ++i) mbar_init(fullPhi + i, nprod);
        for (int i = 0; i < NCF; ++i) mbar_init(fullCf + i, nprod);
        for (int i = 0; i < NZF; ++i) mbar_init(fullZf + i, nprod);
        mbar_init(redDone, NWARPS);
        mbar_init(empty, NWARPS);
        mbar_init(empty + 1, NWARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    __syncthreads();

    if (tid >= CONSUMERS) {
        // ---------------------------------------------------------------------------------------------------
        // producers: lane 0 of each producer warp.  The 115 row copies of a step form one flat list
        //   [phi 18][J g^xx 16][J g^yy 17][1/J 16][lapDiag 16][rhs 16][J g^zz 16]
        // cut into nprod equal ranges; every producer arrives once per phase on each full barrier with the bytes
        // it is about to issue for it (a bulk copy is a per-warp instruction, so one thread issues in a uniform
        // loop and the producers run in parallel on different schedulers).
        // ---------------------------------------------------------------------------------------------------
        if ((tid & 31) != 0) return;
        const int pw = (tid - CONSUMERS) >> 5;
        const int qa = (pw * 115) / nprod, qb = ((pw + 1) * 115) / nprod;   // my rows of the flat list
        const long long rowBase = g.idx(x0 - 1, 0, 0);   // element offset of (x0-1, y = 0, k = 0)
        auto plane_ok = [&](int p) { return p >= -g.gh[2] && p < g.n[2] + g.gh[2]; };
        // rows [lo, hi) of segment `s0` (flat start s0, nrows rows, first local y = ys) that are mine and exist
        auto my_lo = [&](int s0, int ys) { return max(max(qa - s0, 0), -g.gh[1] - ys); };
        auto my_hi = [&](int s0, int nrows, int ys) { return min(min(qb - s0, nrows), g.n[1] + g.gh[1] - ys); };
        const int loP = my_lo(0, y0 - 1), hiP = my_hi(0, PY, y0 - 1);
        const int lo0 = my_lo(18, y0), hi0 = my_hi(18, EY, y0);
        const int lo1 = my_lo(34, y0), hi1 = my_hi(34, EY + 1, y0);
        const int lo2 = my_lo(51, y0), hi2 = my_hi(51, EY, y0);
        const int lo3 = my_lo(67, y0), hi3 = my_hi(67, EY, y0);
        const int lo4 = my_lo(83, y0), hi4 = my_hi(83, EY, y0);
        const int loZ = my_lo(99, y0), hiZ = my_hi(99, EY, y0);
        const int nPhi = max(0, hiP - loP), nZf = max(0, hiZ - loZ);
        const int nCf = max(0, hi0 - lo0) + max(0, hi1 - lo1) + max(0, hi2 - lo2) + max(0, hi3 - lo3) + max(0, hi4 - lo4);
        auto arm = [&](uint64_t* bar, bool ok, int rows) {
            if (ok && rows > 0 && !noload) mbar_expect_tx(bar, rows * ROWB); else mbar_arrive(bar);
        };
        auto seg = [&](const double* base, int p, int yStart, unsigned char* dst, int lo, int hi, uint64_t* bar) {
            if (noload) return;
            const double* src = base + rowBase + (long long)p * g.sz + (long long)(yStart + lo) * g.sy;
            unsigned char* d = dst + lo * ROWB;
            for (int rr = lo; rr < hi; ++rr, src += g.sy, d += ROWB) bulk_row(d, src, bar);
        };
        auto load_phi = [&](int p) {
            const int slot = (p - (k0 - 2)) % NPHI;
            const bool ok = plane_ok(p);
            arm(fullPhi + slot, ok, nPhi);
            if (ok) seg(phi, p, y0 - 1, smem + OFF_PHI + slot * PHI_SLOT, loP, hiP, fullPhi + slot);
        };
        auto load_cf = [&](int s) {
            const int slot = (s - (k0 - 1)) % NCF;
            unsigned char* st = smem + OFF_CF + slot * CF_BYTES;
            uint64_t* bar = fullCf + slot;
            const bool ok = plane_ok(s);
            arm(bar, ok, nCf);
            if (!ok) return;
            seg(m.jg[0][0], s, y0, st + R_JG0 * ROWB, lo0, hi0, bar);
            seg(m.jg[1][1], s, y0, st + R_JG1 * ROWB, lo1, hi1, bar);
            seg(m.jinv, s, y0, st + R_JI * ROWB, lo2, hi2, bar);
            seg(m.lapdiag, s, y0, st + R_LAP * ROWB, lo3, hi3, bar);
            seg(rhs, s, y0, st + R_RHS * ROWB, lo4, hi4, bar);
        };
        auto load_zf = [&](int p) {
            const int slot = (p - (k0 - 1)) % NZF;
            const bool ok = plane_ok(p);
            arm(fullZf + slot, ok, nZf);
            if (ok) seg(m.jg[2][2], p, y0, smem + OFF_ZF + slot * ZF_BYTES, loZ, hiZ, fullZf + slot);
        };
        // prologue: everything steps k0-1 and k0 need
        load_phi(k0 - 2); load_phi(k0 - 1); load_phi(k0);
        load_zf(k0 - 1); load_zf(k0);
        load_cf(k0 - 1);
        load_phi(k0 + 1); load_zf(k0 + 1);
        load_cf(k0);
        for (int k = k0 - 1, j = 0; k + 3 <= k1 + 1; ++k, ++j) {
            mbar_wait(empty + (j & 1), (j >> 1) & 1);   // all consumers are done with step k
            fence_proxy_async();
            load_phi(k + 3);
            load_zf(k + 3);
            if (k + 2 <= k1) load_cf(k + 2);
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
    const int ctr = (r + 1) * PX + 1;   // + ex: this thread's cell in a phi plane (rows start at y0-1)
    const int cfi = r * PX + 1;         // + ex: this thread's cell in a coefficient plane (rows start at y0)

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
        const double* const Zk = reinterpret_cast<const double*>(smem + OFF_ZF + (j % NZF) * ZF_BYTES);
        const double* const Zp = reinterpret_cast<const double*>(smem + OFF_ZF + ((j + 1) % NZF) * ZF_BYTES);
        const double* const Zm = reinterpret_cast<const double*>(smem + OFF_ZF + ((j + 3) % NZF) * ZF_BYTES);
        const double* const C = reinterpret_cast<const double*>(smem + OFF_CF + (j % NCF) * CF_BYTES);
        const double* const Cm = reinterpret_cast<const double*>(smem + OFF_CF + ((j + 2) % NCF) * CF_BYTES);

        const int par = (parBase + k) & 1, parn = par ^ 1;
        const int ex = 2 * t + par;
        const int x = x0 + ex;
        // red update of plane k (extended tile), in place in the ring
        if (k >= 0 && k < g.n[2] && ((colMask >> par) & 1)) {
            const int gk = g.g0[2] + k;
            const bool fast = !tileRim && gk != 0 && gk != g.gn[2] - 1;
            const int c = cfi + ex;
            GsrbCoef q;
            q.glo[0] = C[R_JG0 * PX + c];
            q.ghi[0] = C[R_JG0 * PX + c + 1];
            q.glo[1] = C[R_JG1 * PX + c];
            q.ghi[1] = C[R_JG1 * PX + c + PX];
            q.glo[2] = Zk[c];
            q.ghi[2] = Zp[c];
            q.ji = C[R_JI * PX + c];
            q.lap = C[R_LAP * PX + c];
            q.rhs = C[R_RHS * PX + c];
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
                const int c = cfi + ex;
                GsrbCoef q;
                q.glo[0] = Cm[R_JG0 * PX + c];
                q.ghi[0] = Cm[R_JG0 * PX + c + 1];
                q.glo[1] = Cm[R_JG1 * PX + c];
                q.ghi[1] = Cm[R_JG1 * PX + c + PX];
                q.glo[2] = Zm[c];
                q.ghi[2] = Zk[c];
                q.ji = Cm[R_JI * PX + c];
                q.lap = Cm[R_LAP * PX + c];
                q.rhs = Cm[R_RHS * PX + c];
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

bool gsrb_tma_usable(const Geom& g, const Metric& m) {
    if (!gsrb_fused_usable(g, m)) return false;
    if (g.gh[0] < 2 || (g.gh[0] & 1) || g.gh[1] < 2 || g.gh[2] < 2) return false;   // rows start 16-byte aligned
    if ((g.sy & 1) || (g.off0 & 1)) return false;
    return true;
}

void gsrb_tma_free(void* /*cache*/) {}

bool launch_gsrb_tma(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn, const double* rhs,
                     void** /*cache*/) {
    static bool attr = false;
    if (!attr) {
        if (cudaFuncSetAttribute(k_gsrb_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) {
            (void)cudaGetLastError();
            return false;
        }
        attr = true;
    }
    static int nprod = 0, noload = 0;
    if (!nprod) {
        const char* e = getenv("SOMAR_B200_TMA_PROD");
        nprod = e ? atoi(e) : MAXPROD;
        if (nprod < 1 || nprod > MAXPROD) nprod = MAXPROD;
        noload = getenv("SOMAR_B200_TMA_NOLOAD") != nullptr;   // diagnostic only: wrong results, consumer-side time
    }
    const int kc = (g.n[2] > 64) ? 64 : g.n[2];
    dim3 grid((g.n[0] + CX - 1) / CX, (g.n[1] + CY - 1) / CY, (g.n[2] + kc - 1) / kc);
    k_gsrb_tma<<<grid, CONSUMERS + 32 * nprod, SMEM_BYTES, cx.stream>>>(g, m, phiOut, phiIn, rhs, kc, nprod, noload);
    cx.launches++;
    return true;
}

}  // namespace somar
