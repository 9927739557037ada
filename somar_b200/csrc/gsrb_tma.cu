// gsrb_tma.cu -- k_gsrb_async (gsrb_async.cu: same z-march, same rings, same arithmetic, bit-identical results) with the
// rings fed by the TMA engine instead of per-lane cp.async copies:
//   * ONE thread issues six tensor-map tile loads per step (cp.async.bulk.tensor.3d: J g^xx 64x16, J g^yy 64x17, 1/J
//     64x16, rhs 64x16 of plane k+2; phi 64x18 and J g^zz 64x16 of plane k+3) that complete on an mbarrier with a
//     transaction count -- against 61 LDGSTS per lane and step plus their address arithmetic and row guards in the
//     cp.async version, whose issue slots were 54 % busy with 16 warps per SM (profiles/r2_ncu_gsrb_async.txt);
//   * rows and columns of a tile that lie outside the brick are zero-filled by the TMA unit (they only ever feed
//     cells that are not updated), so the copy guards disappear;
//   * a brick is a 3-D tensor (pitch, rows incl. ghosts, planes incl. ghosts) of doubles; the metric's maps are
//     encoded once per depth, phi's and the right-hand side's per launch (host-side, ~1 us each);
//   * "the input is zero" (first relaxation of a V-cycle visit) is a tensor map over the level's zero plane whose
//     plane coordinate is always 0.
// Round 1 had concluded that tensor-map TMA "raises an illegal instruction on this pool"; the CUDA programming guide's
// own example and a step-by-step bisect (tests/cuda/tma_guide_example.cu, tma_bisect.cu) run fine on the same boxes:
// the round-1 probe was at fault, not the hardware.
// Synchronisation: full[4] (one arrival + bytes) says "group G has landed"; group G = what step G consumes first
// (coefficients of plane k, phi and J g^zz of plane k+1); redDone and empty[2] as in gsrb_async.cu.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "kernels.hpp"
#include "stencil_ortho.cuh"

namespace somar {

namespace {

#define k_gsrb_async k_gsrb_tma

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
constexpr int SMEM_BYTES = OFF_BAR + 8 * 8;   // redDone, empty[2], full[4]
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
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
// one box of a 3-D tensor -> shared memory, completion on an mbarrier
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tm, int cx, int cy, int cz, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
        "l"(tm), "r"(cx), "r"(cy), "r"(cz), "r"(smem_u32(bar))
        : "memory");
}
struct TmaMaps {
    CUtensorMap phi, rhs, jx, jy, jz, ji;
};
constexpr uint32_t TX_CF = (3 * EY + (EY + 1)) * ROWB;   // J g^xx, 1/J, rhs: 16 rows each; J g^yy: 17
constexpr uint32_t TX_PHI = PY * ROWB, TX_ZF = EY * ROWB;

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
    uint32_t sBase;         // shared address of the ring buffers
    int tcx, tcy, tcz;      // tensor coordinates of (column 0, first red-update row, plane 0); phiZ: phi plane coordinate mode
    int phiZ;               // 1: plane coordinate of the phi input follows k, 0: always plane 0 (zero input)
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
__device__ __forceinline__ void sweep_step(const Geom& g, const TmaMaps& tm, double* __restrict__ phiOut,
                                           unsigned char* smem, uint64_t* redDone, uint64_t* empty, uint64_t* full,
                                           const Lane& L, Pipe& s, int k, int j) {
    const unsigned FULL = 0xffffffffu;
    if (j > 0) mbar_wait(redDone, (j - 1) & 1);
    mbar_wait(full + (j & 3), (j >> 2) & 1);   // group j has landed: coefficients k, phi k+1, J g^zz k+1

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
    if (threadIdx.x == 0) {
        // the slots about to be refilled were read AND written (red values) through the generic proxy
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        uint64_t* const fb = full + ((j + 2) & 3);
        const bool cf = k + 2 <= L.k1, pz = k + 3 <= L.k1 + 1;
        mbar_expect_tx(fb, (cf ? TX_CF : 0u) + (pz ? TX_PHI + TX_ZF : 0u));
        if (cf) {
            const uint32_t d = L.sBase + OFF_CF + wrap(s.i3 + 2, NCF) * CF_BYTES;
            const int z = L.tcz + k + 2;
            tma_load_3d(d + R_JG0 * ROWB, &tm.jx, L.tcx, L.tcy, z, fb);
            tma_load_3d(d + R_JG1 * ROWB, &tm.jy, L.tcx, L.tcy, z, fb);
            tma_load_3d(d + R_JI * ROWB, &tm.ji, L.tcx, L.tcy, z, fb);
            tma_load_3d(d + R_RHS * ROWB, &tm.rhs, L.tcx, L.tcy, z, fb);
        }
        if (pz) {
            const int z = L.tcz + k + 3;
            tma_load_3d(L.sBase + OFF_PHI + wrap(s.i5 + 4, NPHI) * PHI_BYTES, &tm.phi, L.tcx, L.tcy - 1, L.phiZ ? z : 0, fb);
            tma_load_3d(L.sBase + OFF_ZF + wrap(s.i4 + 3, NZF) * ZF_BYTES, &tm.jz, L.tcx, L.tcy, z, fb);
        }
    }
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
k_gsrb_async(const Geom g, const __grid_constant__ TmaMaps tm, double* __restrict__ phiOut, int kc, int phiFollowsK) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* const redDone = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* const empty = redDone + 1;
    uint64_t* const full = redDone + 3;

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
    L.psz = 0;
    const int k0 = L.k0;
    L.sBase = smem_u32(smem);
    L.sRow = L.sBase + L.r * ROWB + 16 * L.t;
    L.tcx = g.gh[0] + L.x0;                  // brick coordinates: x pad, ghost rows, ghost planes
    L.tcy = g.gh[1] + y0;
    L.tcz = g.gh[2];
    L.phiZ = phiFollowsK;
    L.rowOk = L.rowUp = true;
    const int cbase = (int)g.idx(L.x0 + 2 * L.t, L.y, 0);       // my pair in plane 0 (brick index)
    L.cb = cbase;

    if (tid == 0) {
        mbar_init(redDone, NWARPS);
        mbar_init(empty, NWARPS);
        mbar_init(empty + 1, NWARPS);
        for (int q = 0; q < 4; ++q) mbar_init(full + q, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        // prologue.  Group 0 = what step k0-1 needs: phi k0-2 .. k0, J g^zz k0-1 .. k0, coefficients k0-1;
        //            group 1 = what step k0 adds:     phi k0+1, J g^zz k0+1, coefficients k0.
        // Ring positions as in gsrb_async.cu: phi plane p -> (p - (k0-2)) % 5, J g^zz and coefficients -> (p - (k0-1)) % n.
        // Planes k0-2 .. k1+1 exist in the brick (two ghost planes); rows / columns outside it are zero-filled.
        auto phi_at = [&](int p, uint64_t* fb) {
            tma_load_3d(L.sBase + OFF_PHI + ((p - (k0 - 2)) % NPHI) * PHI_BYTES, &tm.phi, L.tcx, L.tcy - 1,
                        L.phiZ ? L.tcz + p : 0, fb);
        };
        auto zf_at = [&](int p, uint64_t* fb) {
            tma_load_3d(L.sBase + OFF_ZF + ((p - (k0 - 1)) % NZF) * ZF_BYTES, &tm.jz, L.tcx, L.tcy, L.tcz + p, fb);
        };
        auto cf_at = [&](int p, uint64_t* fb) {
            const uint32_t d = L.sBase + OFF_CF + ((p - (k0 - 1)) % NCF) * CF_BYTES;
            tma_load_3d(d + R_JG0 * ROWB, &tm.jx, L.tcx, L.tcy, L.tcz + p, fb);
            tma_load_3d(d + R_JG1 * ROWB, &tm.jy, L.tcx, L.tcy, L.tcz + p, fb);
            tma_load_3d(d + R_JI * ROWB, &tm.ji, L.tcx, L.tcy, L.tcz + p, fb);
            tma_load_3d(d + R_RHS * ROWB, &tm.rhs, L.tcx, L.tcy, L.tcz + p, fb);
        };
        mbar_expect_tx(full + 0, 3 * TX_PHI + 2 * TX_ZF + TX_CF);
        phi_at(k0 - 2, full + 0); phi_at(k0 - 1, full + 0); phi_at(k0, full + 0);
        zf_at(k0 - 1, full + 0); zf_at(k0, full + 0);
        cf_at(k0 - 1, full + 0);
        mbar_expect_tx(full + 1, TX_PHI + TX_ZF + TX_CF);
        phi_at(k0 + 1, full + 1); zf_at(k0 + 1, full + 1);
        cf_at(k0, full + 1);
    }

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

    __syncthreads();                 // the mbarriers are initialised
    mbar_wait(full + 0, 0);          // group 0 has landed (step 0 waits for it again: same phase, returns at once)
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
            sweep_step<1>(g, tm, phiOut, smem, redDone, empty, full, L, s, k, j);
            ++k; ++j;
            if (k > L.k1) break;
            sweep_step<0>(g, tm, phiOut, smem, redDone, empty, full, L, s, k, j);
            ++k; ++j;
        }
    } else {
        while (k <= L.k1) {
            sweep_step<0>(g, tm, phiOut, smem, redDone, empty, full, L, s, k, j);
            ++k; ++j;
            if (k > L.k1) break;
            sweep_step<1>(g, tm, phiOut, smem, redDone, empty, full, L, s, k, j);
            ++k; ++j;
        }
    }
}
#undef k_gsrb_async

// ---- host side ------------------------------------------------------------------------------------------------
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// a brick (or one plane of it: planes == 1) as a 3-D tensor of doubles, boxes of 64 x rows x 1
bool encode_brick(CUtensorMap* tm, const Geom& g, const double* base, int boxRows, long long planes) {
    const cuuint64_t dims[3] = {(cuuint64_t)g.sy, (cuuint64_t)(g.sz / g.sy), (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)g.sy * 8, (cuuint64_t)g.sz * 8};
    const cuuint32_t box[3] = {(cuuint32_t)PX, (cuuint32_t)boxRows, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    return encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(base), dims, strides, box, es,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

bool gsrb_tma_usable(const Geom& g, const Metric& m) {
    const char* e = getenv("SOMAR_B200_TMA");   // read per call: the tests switch it within one process
    const bool on = !(e && e[0] == '0');
    if (!on || !gsrb_async_usable(g, m) || encode_fn() == nullptr) return false;
    if ((g.sy * 8) % 16 || (g.sz * 8) % 16 || g.sz % g.sy) return false;
    return true;
}

// Same contract as launch_gsrb_async.  zeroPlane != nullptr: the input is identically zero (the level's zero plane,
// one plane of off0 + sz doubles, serves every input plane).
bool launch_gsrb_tma(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn, const double* rhs,
                     const double* zeroPlane) {
    static bool attr = false;
    if (!attr) {
        if (cudaFuncSetAttribute(k_gsrb_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES) != cudaSuccess) {
            (void)cudaGetLastError();
            return false;
        }
        attr = true;
    }
    const long long planes = g.total / g.sz;
    TmaMaps tm;
    bool ok = encode_brick(&tm.jx, g, m.jg[0][0], EY, planes) && encode_brick(&tm.jy, g, m.jg[1][1], EY + 1, planes) &&
              encode_brick(&tm.jz, g, m.jg[2][2], EY, planes) && encode_brick(&tm.ji, g, m.jinv, EY, planes) &&
              encode_brick(&tm.rhs, g, rhs, EY, planes) &&
              (zeroPlane ? encode_brick(&tm.phi, g, zeroPlane, PY, 1) : encode_brick(&tm.phi, g, phiIn, PY, planes));
    if (!ok) return false;
    const int tiles = ((g.n[0] + CX - 1) / CX) * ((g.n[1] + CY - 1) / CY);
    int kc = g.n[2];
    {   // z-chunking by wave count, as launch_gsrb_async
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
    k_gsrb_tma<<<grid, NTHREADS, SMEM_BYTES, cx.stream>>>(g, tm, phiOut, kc, zeroPlane ? 0 : 1);
    cx.launches++;
    return true;
}

}  // namespace somar
