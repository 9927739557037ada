// gsrb_async.cu -- the fused red+black Gauss-Seidel ITERATION of gsrb_fused.cu (same z-march, same arithmetic:
// bit-identical results) as a software pipeline over shared-memory rings fed by asynchronous copies:
//   * every input plane goes HBM -> shared memory with cp.async.cg (16 B per lane = one 512-byte tile row per warp
//     instruction, SASS LDGSTS.BYPASS): no registers are tied up by loads in flight and the copies run two to
//     three planes ahead of the arithmetic.  Warp r copies row r of every array;
//   * the thread keeps what it owns in registers -- its two cells of planes k-2 .. k+1 (updated in place as it
//     relaxes them) and the coefficients of the partner column of plane k, needed one step later for that cell's
//     black update -- and reads shared memory only for: one conflict-free 16-byte pair per coefficient array and
//     plane, and the four y-neighbours (other warps' cells).  x-neighbours come from the adjacent lane by shuffle,
//     z-neighbours are the thread's own cells;
//   * no CTA-wide barrier in the steady state.  The only cross-thread dependence of the sweep is "the black update
//     of plane k reads red values of plane k written by OTHER warps one step earlier", so one split mbarrier does
//     it: every warp ARRIVES on `redDone` after its red update of step k -- and after its own copies for step k+1
//     have landed (cp.async.wait_group) -- and WAITS for it at the top of step k+1, with the black update of plane
//     k-1 and the stores in between.  Ring slots are recycled through two alternating `empty` mbarriers.
//
// History (profiles/README.md): the register-prefetch kernel (gsrb_fused.cu) is latency/barrier bound at 63 % of the
// HBM peak.  Feeding the rings with the TMA engine was tried first: cp.async.bulk.tensor raises "illegal
// instruction" on this pool's B200 boxes for every descriptor form (tests/cuda/tma_probe.cu), and row-wise
// cp.async.bulk copies (UBLKCP) saturate at one 528-byte copy per ~44 cycles per SM (6.0 ms per iteration,
// independent of the number of issuing warps).  Dedicated LDGSTS producer warps worked (5.1 ms) but cost the
// consumers their registers (ptxas budgets 640 threads at 96 registers), and the shared-memory pipe -- which also
// carries the LDGSTS writes -- was saturated by scalar, 2-way bank-conflicted coefficient reads.
//
// Tile: shared-memory rows are 64 doubles = 512 B, column c <-> local x = x0 + c with x0 = 60 bx - 2 (16-byte
// aligned in the brick: SOMAR_XO = 4).  Columns 0 and 63 are read-only halo, red updates cover columns 1..62 (the
// core 2..61 plus a 1-cell ring that the neighbouring CTA recomputes), rows likewise: 18 rows of phi, red updates on
// rows y0..y0+15, core rows y0+1..y0+14 (y0 = 14 by - 1).  Thread (t, r) owns the x-pair (2t, 2t+1) of row r:
// exactly one of the two cells is red in any plane, and the black cell it updates in plane k-1 is in the same
// column as its red cell of plane k.
//   phi planes      ring of 5   read at step k: k-1, k, k+1     in flight: k+2, k+3
//   coefficients    ring of 3   read at step k: k               in flight: k+1, k+2
//   J g^zz planes   ring of 4   read at step k: k+1             in flight: k+2, k+3   (plane k = low z-face of cells k)
// Shared memory: 5 x 9 KB + 3 x 32.5 KB + 4 x 8 KB = 174.5 KB, one CTA of 512 threads per SM, <= 128 registers.
//
// Step k of a warp (k = k0-1 .. k1):
//   wait redDone(k-1)        => all red values of plane k-1 are in the ring, and copy group k-2 has landed:
//                               coefficients k, phi plane k+1, J g^zz plane k+1
//   read my pairs, red update of plane k (ring + registers)
//   wait empty(k-1)          => every warp has finished step k-1: phi plane k-2, coefficient stage k-1 and
//                               J g^zz plane k-1 are dead;  issue copy group k = {coef k+2, phi k+3, J g^zz k+3}
//   wait_group 1 (my copies of group k-1 have landed);  arrive redDone(k)
//   black update of plane k-1 (core tile) from the set-aside coefficients -> HBM with the red value beside it
//   arrive empty(k)
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "kernels.hpp"
#include "stencil_ortho.cuh"

namespace somar {

namespace {

constexpr int PX = 64;                   // columns of a shared-memory row
constexpr int EY = 16, PY = EY + 2;      // red-update rows, phi rows
constexpr int CX = PX - 4, CY = EY - 2;  // core tile 60 x 14
constexpr int NPHI = 5, NCF = 3, NZF = 4;
constexpr int NTHREADS = (PX / 2) * EY;  // 512
constexpr int NWARPS = NTHREADS / 32;
constexpr int ROWB = PX * 8;             // 512 bytes per row

constexpr int PHI_BYTES = PY * ROWB;     // 9216
constexpr int R_JG0 = 0, R_JG1 = R_JG0 + EY, R_JI = R_JG1 + EY + 1, R_RHS = R_JI + EY;   // lapDiag is not streamed:
constexpr int CF_ROWS = R_RHS + EY;      // 65            gsrb_update_vals<3, true> forms it from the faces
constexpr int CF_BYTES = CF_ROWS * ROWB;
constexpr int ZF_BYTES = EY * ROWB;
constexpr int OFF_PHI = 0;
constexpr int OFF_CF = OFF_PHI + NPHI * PHI_BYTES;
constexpr int OFF_ZF = OFF_CF + NCF * CF_BYTES;
constexpr int OFF_BAR = OFF_ZF + NZF * ZF_BYTES;
constexpr int SMEM_BYTES = OFF_BAR + 3 * 8;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
static_assert(EY == NWARPS, "warp r copies tile row r");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
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
// 16 bytes global -> shared, asynchronous, L1 bypassed
__device__ __forceinline__ void copy16(uint32_t dst, const double* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void copy_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void copy_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// What a thread knows about itself (constant over the sweep).
struct Lane {
    int t, r;               // pair index in the row, row = warp
    int x0, y;              // local x of column 0, local y of my row
    int k0, k1;             // z chunk [k0, k1)
    int isz;                // plane stride of the brick (elements)
    int psz, cb;            // plane stride of the phi INPUT (0: every plane is the level's zero plane), my pair in plane 0
    int own, cown;          // my pair in a phi plane / in a coefficient plane (element index)
    int ctr;                // start of my row in a phi plane
    int redMask, coreMask;  // bit p: column 2t+p takes part in red updates / is a core cell
    bool tileRim;           // block-uniform: the red-update region touches the domain rim in x or y
    bool rowOk, rowUp;      // my row / the row above exist in the brick (copies)
    uint32_t sRow;          // shared address of my 16 bytes of row r of a plane at offset 0
};

// What changes from step to step.
struct Pipe {
    double2 Am2, Am1, A0;   // my pair in planes k-2, k-1, k
    double2 Zc;             // J g^zz pair of plane k
    GsrbCoef carry;         // coefficients of the partner column of plane k-1 (this step's black update)
    int i5, i4, i3;         // ring positions: j % 5, j % 4, j % 3
    int gk;                 // brick index of my pair in plane k
};

__device__ __forceinline__ int wrap(int i, int n) { return i >= n ? i - n : i; }

// One step of the sweep for a warp whose red cell of plane k is column 2t + PAR (PAR alternates from step to step,
// so the loop below is unrolled by two and every pick of a pair element is resolved at compile time).
template <int PAR>
__device__ __forceinline__ void sweep_step(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                           const double* __restrict__ rhs, double* __restrict__ phiOut,
                                           unsigned char* smem, uint64_t* redDone, uint64_t* empty, const Lane& L,
                                           Pipe& s, int k, int j) {
    const unsigned FULL = 0xffffffffu;
    if (j > 0) mbar_wait(redDone, (j - 1) & 1);

    const double* const Pm = reinterpret_cast<const double*>(smem + OFF_PHI + s.i5 * PHI_BYTES);
    double* const Pk = reinterpret_cast<double*>(smem + OFF_PHI + wrap(s.i5 + 1, NPHI) * PHI_BYTES);
    const double* const Pp = reinterpret_cast<const double*>(smem + OFF_PHI + wrap(s.i5 + 2, NPHI) * PHI_BYTES);
    const double* const Zp = reinterpret_cast<const double*>(smem + OFF_ZF + wrap(s.i4 + 1, NZF) * ZF_BYTES);
    const double* const C = reinterpret_cast<const double*>(smem + OFF_CF + s.i3 * CF_BYTES);
    const int col = 2 * L.t + PAR;
    const int x = L.x0 + col;

    // plane k+1 of my cells, coefficient pairs of plane k: 16-byte conflict-free loads
    const double2 Ap1 = *reinterpret_cast<const double2*>(Pp + L.own);
    const double2 J0 = *reinterpret_cast<const double2*>(C + R_JG0 * PX + L.cown);
    const double2 Y0 = *reinterpret_cast<const double2*>(C + R_JG1 * PX + L.cown);
    const double2 Y1 = *reinterpret_cast<const double2*>(C + (R_JG1 + 1) * PX + L.cown);
    const double2 JI = *reinterpret_cast<const double2*>(C + R_JI * PX + L.cown);
    const double2 RH = *reinterpret_cast<const double2*>(C + R_RHS * PX + L.cown);
    const double2 Zn = *reinterpret_cast<const double2*>(Zp + L.cown);
    const double J0n = __shfl_down_sync(FULL, J0.x, 1);       // x-face 2t+2 (lane 31: unused)
    // x-neighbour outside my pair, planes k (old black) and k-1 (new red), from the adjacent lane's registers
    const double xnK = PAR ? __shfl_down_sync(FULL, s.A0.x, 1) : __shfl_up_sync(FULL, s.A0.y, 1);
    const double xnM = PAR ? __shfl_down_sync(FULL, s.Am1.x, 1) : __shfl_up_sync(FULL, s.Am1.y, 1);

    // red update of plane k, column `col`
    if (k >= 0 && k < g.n[2] && ((L.redMask >> PAR) & 1)) {
        const int gz = g.g0[2] + k;
        const bool fast = !L.tileRim && gz != 0 && gz != g.gn[2] - 1;
        GsrbCoef q;
        q.glo[0] = PAR ? J0.y : J0.x;
        q.ghi[0] = PAR ? J0n : J0.y;
        q.glo[1] = PAR ? Y0.y : Y0.x;
        q.ghi[1] = PAR ? Y1.y : Y1.x;
        q.glo[2] = PAR ? s.Zc.y : s.Zc.x;
        q.ghi[2] = PAR ? Zn.y : Zn.x;
        q.ji = PAR ? JI.y : JI.x;
        q.lap = 0.0;
        q.rhs = PAR ? RH.y : RH.x;
        double nb[3][2];
        nb[0][0] = PAR ? s.A0.x : xnK;
        nb[0][1] = PAR ? xnK : s.A0.y;
        nb[1][0] = Pk[L.ctr + col - PX];
        nb[1][1] = Pk[L.ctr + col + PX];
        nb[2][0] = PAR ? s.Am1.y : s.Am1.x;
        nb[2][1] = PAR ? Ap1.y : Ap1.x;
        const double v = gsrb_update_vals<3, true>(g, q, x, L.y, k, PAR ? s.A0.y : s.A0.x, nb, fast);
        Pk[L.ctr + col] = v;          // for the rows above and below
        if (PAR) s.A0.y = v; else s.A0.x = v;
    }

    // recycle the slots every warp has finished with and keep the copies two to three planes ahead:
    // group k = {coefficients k+2, phi plane k+3, J g^zz plane k+3}
    if (j > 0) mbar_wait(empty + ((j - 1) & 1), ((j - 1) >> 1) & 1);
    if (k + 2 <= L.k1) {
        const uint32_t d = L.sRow + OFF_CF + wrap(s.i3 + 2, NCF) * CF_BYTES;
        const long long o = s.gk + 2 * L.isz;
        if (L.rowOk) {
            copy16(d + R_JG0 * ROWB, m.jg[0][0] + o);
            copy16(d + R_JG1 * ROWB, m.jg[1][1] + o);
            copy16(d + R_JI * ROWB, m.jinv + o);
            copy16(d + R_RHS * ROWB, rhs + o);
        }
        if (L.r == EY - 1 && L.rowUp) copy16(d + (R_JG1 + 1) * ROWB, m.jg[1][1] + o + g.sy);   // 17th row of J g^yy
    }
    if (k + 3 <= L.k1 + 1) {
        const long long o = s.gk + 3 * L.isz;
        const long long po = L.cb + (long long)(k + 3) * L.psz;
        const uint32_t d = L.sRow + OFF_PHI + wrap(s.i5 + 4, NPHI) * PHI_BYTES + ROWB;
        if (L.rowOk) {
            copy16(d, phi + po);
            copy16(L.sRow + OFF_ZF + wrap(s.i4 + 3, NZF) * ZF_BYTES, m.jg[2][2] + o);
        }
        if (L.r == 0 && L.y - 1 >= -g.gh[1]) copy16(d - ROWB, phi + po - g.sy);
        if (L.r == EY - 1 && L.rowUp) copy16(d + ROWB, phi + po + g.sy);
    }
    copy_commit();
    copy_wait<1>();            // my copies of the previous group (data of step k+1) have landed
    __syncwarp();
    if (L.t == 0) mbar_arrive(redDone);

    // black update of plane k-1 (core tile), same column: coefficients were set aside one step ago; x/y
    // neighbours are red values of step k-1 (redDone(k-1) above), z neighbours are my own red cells
    if (k - 1 >= L.k0 && k - 1 < L.k1) {
        const int kb = k - 1;
        double vb = PAR ? s.Am1.y : s.Am1.x;
        if ((L.coreMask >> PAR) & 1) {
            const int gz = g.g0[2] + kb;
            const bool fast = !L.tileRim && gz != 0 && gz != g.gn[2] - 1;
            double nb[3][2];
            nb[0][0] = PAR ? s.Am1.x : xnM;
            nb[0][1] = PAR ? xnM : s.Am1.y;
            nb[1][0] = Pm[L.ctr + col - PX];
            nb[1][1] = Pm[L.ctr + col + PX];
            nb[2][0] = PAR ? s.Am2.y : s.Am2.x;
            nb[2][1] = PAR ? s.A0.y : s.A0.x;
            vb = gsrb_update_vals<3, true>(g, s.carry, x, L.y, kb, vb, nb, fast);
        }
        // plane k-1 of my pair is final: the black result and the red value next to it, one 16-byte store
        double* const o = phiOut + (s.gk - L.isz);
        if (L.coreMask == 3) {
            *reinterpret_cast<double2*>(o) = PAR ? make_double2(s.Am1.x, vb) : make_double2(vb, s.Am1.y);
        } else {
            if ((L.coreMask >> PAR) & 1) o[PAR] = vb;
            if ((L.coreMask >> (PAR ^ 1)) & 1) o[PAR ^ 1] = PAR ? s.Am1.x : s.Am1.y;
        }
    }
    // coefficients of the partner column of plane k: black update of the next step
    s.carry.glo[0] = PAR ? J0.x : J0.y;
    s.carry.ghi[0] = PAR ? J0.y : J0n;
    s.carry.glo[1] = PAR ? Y0.x : Y0.y;
    s.carry.ghi[1] = PAR ? Y1.x : Y1.y;
    s.carry.glo[2] = PAR ? s.Zc.x : s.Zc.y;
    s.carry.ghi[2] = PAR ? Zn.x : Zn.y;
    s.carry.ji = PAR ? JI.x : JI.y;
    s.carry.lap = 0.0;
    s.carry.rhs = PAR ? RH.x : RH.y;
    s.Am2 = s.Am1;
    s.Am1 = s.A0;
    s.A0 = Ap1;
    s.Zc = Zn;
    s.i5 = wrap(s.i5 + 1, NPHI);
    s.i4 = wrap(s.i4 + 1, NZF);
    s.i3 = wrap(s.i3 + 1, NCF);
    s.gk += L.isz;
    __syncwarp();
    if (L.t == 0) mbar_arrive(empty + (j & 1));
}

}  // namespace

__global__ void __launch_bounds__(NTHREADS, 1)
k_gsrb_async(const Geom g, const Metric m, double* __restrict__ phiOut, const double* __restrict__ phi,
             const double* __restrict__ rhs, int kc, int phiPlaneStride) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* const redDone = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* const empty = redDone + 1;

    const int tid = threadIdx.x;
    Lane L;
    L.t = tid & 31;
    L.r = tid >> 5;
    L.x0 = blockIdx.x * CX - 2;              // local x of shared-memory column 0
    const int y0 = blockIdx.y * CY - 1;      // local y of the first red-update row
    L.y = y0 + L.r;
    L.k0 = blockIdx.z * kc;
    L.k1 = min(L.k0 + kc, g.n[2]);
    L.isz = (int)g.sz;
    L.psz = phiPlaneStride;
    const int k0 = L.k0;

    if (tid == 0) {
        mbar_init(redDone, NWARPS);
        mbar_init(empty, NWARPS);
        mbar_init(empty + 1, NWARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }

    // ---- copies: warp r moves row r of every array (lane t: bytes 16 t .. 16 t + 15 of the row) ----------------
    L.rowOk = (L.y >= -g.gh[1]) && (L.y < g.n[1] + g.gh[1]);
    L.rowUp = (L.y + 1 < g.n[1] + g.gh[1]);                     // row y+1 exists (y+1 >= -gh always)
    const int cbase = (int)g.idx(L.x0 + 2 * L.t, L.y, 0);       // my pair in plane 0 (brick index)
    L.cb = cbase;
    L.sRow = smem_u32(smem) + L.r * ROWB + 16 * L.t;
    // planes k0-2 .. k1+1 exist in the brick (two ghost planes: gsrb_async_usable)
    auto load_phi = [&](int p) {        // phi rows y0-1 .. y0+16: mine is r+1; warps 0 and 15 add the outer rows
        const uint32_t d = L.sRow + OFF_PHI + ((p - (k0 - 2)) % NPHI) * PHI_BYTES + ROWB;
        const double* s = phi + cbase + (long long)p * L.psz;
        if (L.rowOk) copy16(d, s);
        if (L.r == 0 && L.y - 1 >= -g.gh[1]) copy16(d - ROWB, s - g.sy);
        if (L.r == EY - 1 && L.rowUp) copy16(d + ROWB, s + g.sy);
    };
    auto load_cf = [&](int p) {
        const uint32_t d = L.sRow + OFF_CF + ((p - (k0 - 1)) % NCF) * CF_BYTES;
        const long long o = cbase + (long long)p * L.isz;
        if (L.rowOk) {
            copy16(d + R_JG0 * ROWB, m.jg[0][0] + o);
            copy16(d + R_JG1 * ROWB, m.jg[1][1] + o);
            copy16(d + R_JI * ROWB, m.jinv + o);
            copy16(d + R_RHS * ROWB, rhs + o);
        }
        if (L.r == EY - 1 && L.rowUp) copy16(d + (R_JG1 + 1) * ROWB, m.jg[1][1] + o + g.sy);
    };
    auto load_zf = [&](int p) {
        if (L.rowOk) copy16(L.sRow + OFF_ZF + ((p - (k0 - 1)) % NZF) * ZF_BYTES, m.jg[2][2] + cbase + (long long)p * L.isz);
    };
    // prologue: group A = what step k0-1 needs, group B = what step k0 adds
    load_phi(k0 - 2); load_phi(k0 - 1); load_phi(k0);
    load_zf(k0 - 1); load_zf(k0);
    load_cf(k0 - 1);
    copy_commit();
    load_phi(k0 + 1); load_zf(k0 + 1);
    load_cf(k0);
    copy_commit();

    // ---- per-thread constants of the sweep ---------------------------------------------------------------------
    // Rows that take part: my own, plus -- for the red updates only -- the first ghost row towards a neighbour
    // rank (its phi, metric and right-hand side were exchanged; relaxing its red cells here, exactly as the
    // owner does, spares the exchange between the colour passes).
    const bool yown = (L.y >= 0) && (L.y < g.n[1]);
    const bool yred = (L.y >= ((g.nbr[0] >= 0 || g.periodic[1]) ? -1 : 0)) &&
                      (L.y < g.n[1] + ((g.nbr[1] >= 0 || g.periodic[1]) ? 1 : 0));   // periodic y: the image row
    const bool ycore = (L.r >= 1) && (L.r <= CY);
    L.redMask = 0;
    L.coreMask = 0;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int col = 2 * L.t + p, x = L.x0 + col;
        const bool xin = x >= 0 && x < g.n[0];
        if (yred && xin && col >= 1 && col <= PX - 2) L.redMask |= 1 << p;
        if (yown && xin && ycore && col >= 2 && col <= PX - 3) L.coreMask |= 1 << p;
    }
    L.tileRim = (g.g0[0] + L.x0 + 1 <= 0) || (g.g0[0] + L.x0 + PX - 2 >= g.gn[0] - 1) || (g.g0[1] + y0 <= 0) ||
                (g.g0[1] + y0 + EY - 1 >= g.gn[1] - 1);
    L.ctr = (L.r + 1) * PX;          // phi rows start at y0-1
    L.own = L.ctr + 2 * L.t;
    L.cown = L.r * PX + 2 * L.t;     // coefficient rows start at y0

    copy_wait<1>();      // my part of group A has landed ...
    __syncthreads();     // ... and everybody else's; the mbarriers are initialised
    Pipe s;
    s.Am2 = make_double2(0.0, 0.0);
    s.Am1 = *reinterpret_cast<const double2*>(reinterpret_cast<const double*>(smem + OFF_PHI) + L.own);              // plane k0-2
    s.A0 = *reinterpret_cast<const double2*>(reinterpret_cast<const double*>(smem + OFF_PHI + PHI_BYTES) + L.own);   // plane k0-1
    s.Zc = *reinterpret_cast<const double2*>(reinterpret_cast<const double*>(smem + OFF_ZF) + L.cown);               // J g^zz plane k0-1
    s.carry = GsrbCoef();
    s.i5 = s.i4 = s.i3 = 0;
    s.gk = cbase + (k0 - 1) * L.isz;
    // the red column of my pair alternates with k; a warp is one row, so the start parity is warp-uniform
    const int par0 = (g.g0[0] + L.x0 + g.g0[1] + L.y + g.g0[2] + (k0 - 1)) & 1;
    int k = k0 - 1, j = 0;
    if (par0) {
        while (k <= L.k1) {
            sweep_step<1>(g, m, phi, rhs, phiOut, smem, redDone, empty, L, s, k, j);
            ++k; ++j;
            if (k > L.k1) break;
            sweep_step<0>(g, m, phi, rhs, phiOut, smem, redDone, empty, L, s, k, j);
            ++k; ++j;
        }
    } else {
        while (k <= L.k1) {
            sweep_step<0>(g, m, phi, rhs, phiOut, smem, redDone, empty, L, s, k, j);
            ++k; ++j;
            if (k > L.k1) break;
            sweep_step<1>(g, m, phi, rhs, phiOut, smem, redDone, empty, L, s, k, j);
            ++k; ++j;
        }
    }
    copy_wait<0>();
}

bool gsrb_async_usable(const Geom& g, const Metric& m) {
    if (!m.diagonal || g.nd != 3) return false;
    // periodic y is a neighbour like any other (the image rows are filled by the exchange before each iteration and
    // their red cells relaxed here); periodic x / z images would need a fill between the colour passes
    if (g.periodic[0] || g.periodic[2]) return false;
    if (g.total >= (1LL << 31)) return false;                             // 32-bit brick indices
    if (g.gh[0] < 2 || (g.gh[0] & 1) || g.gh[1] < 2 || g.gh[2] < 2) return false;   // rows start 16-byte aligned
    if ((g.sy & 1) || (g.off0 & 1)) return false;
    return true;
}

// zeroPlane != nullptr: the input is identically zero and is not read from phiIn at all -- every input plane is
// the level's (L2-resident) zero plane, plane stride 0 (the first relaxation of a V-cycle visit)
bool launch_gsrb_async(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                       const double* rhs, const double* zeroPlane) {
    static bool attr = false;
    if (!attr) {
        if (cudaFuncSetAttribute(k_gsrb_async, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) {
            (void)cudaGetLastError();
            return false;
        }
        attr = true;
    }
    // z-chunking: one CTA per SM, so the launch runs in ceil(CTAs / SMs) waves of (kc + ~4 pipeline-fill planes) steps.
    // With few tiles (a rank's slab at 8 GPUs: 180 tiles on 148 SMs) one chunk per tile wastes most of the second
    // wave; pick the chunk count that minimises waves x steps.  Many tiles: this keeps 64-plane chunks.
    const int tiles = ((g.n[0] + CX - 1) / CX) * ((g.n[1] + CY - 1) / CY);
    int kc = g.n[2];
    {
        long long best = -1;
        const int maxChunks = g.n[2] >= 8 ? g.n[2] / 4 : 1;
        for (int ch = 1; ch <= maxChunks; ++ch) {
            const int len = (g.n[2] + ch - 1) / ch;
            if (len > 64) continue;
            const int nch = (g.n[2] + len - 1) / len;
            const long long waves = ((long long)tiles * nch + cx.sms - 1) / cx.sms;
            const long long cost = waves * (len + 4);
            if (best < 0 || cost < best) { best = cost; kc = len; }
        }
        if (best < 0) kc = 64;
    }
    dim3 grid((g.n[0] + CX - 1) / CX, (g.n[1] + CY - 1) / CY, (g.n[2] + kc - 1) / kc);
    k_gsrb_async<<<grid, NTHREADS, SMEM_BYTES, cx.stream>>>(g, m, phiOut, zeroPlane ? zeroPlane : phiIn, rhs, kc,
                                                            zeroPlane ? 0 : (int)g.sz);
    cx.launches++;
    return true;
}

}  // namespace somar
