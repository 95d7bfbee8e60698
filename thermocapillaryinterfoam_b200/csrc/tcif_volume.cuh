// tcif_volume.cuh -- the cell-centric sweeps of the hot path for meshes with a compile-time slot count
// K (2: one-cell-thick hex, 3: hex, 4: tet/prism), restructured for memory-level parallelism.
//
// round-1 ncu (profiles/r01_ncu_full_n400_*.csv) showed the first versions of these kernels latency-
// bound: ~80 % of warp stalls on long-scoreboard with 0.8 eligible warps per scheduler, because
// every gather sat behind a branch on its index (`if (fd >= 0)`), which forbids hoisting the load.
// Here every thread (1) loads all 2K slot indices, (2) turns padding (-1) into a safe in-range
// address (its own slot), (3) issues ALL gathers unconditionally, and only then (4) runs the
// reference's arithmetic under the original `valid` predicates -- same operations, same order, so
// results stay bit-identical to the serial oracle.  Never-reused streams (Sf, w, Q+-, sums, outputs)
// use evict-first loads/stores (__ldcs/__stcs) so that L2 keeps the arrays neighbours gather from.
//
// round 2: the patch faces CONTINUE the volume sweep instead of repeating it.
//   volume launch (kv_*/kvp_*)  every cell over its internal faces, in tile order.  A boundary-adjacent cell stops
//       after its internal faces and writes its partial accumulators (and the few own-cell values the patch faces need)
//       into DMesh::part, a small SoA buffer indexed by the cell's rank among the boundary-adjacent cells -- a handful
//       of stores at the end of the thread, no extra loads, no longer-lived warps.
//   fix launch (kfix_*)  one thread per boundary-adjacent cell: reads the partials (coalesced: consecutive threads,
//       consecutive ranks), adds the patch faces in patch order, finishes the cell.  Internal faces come before patch
//       faces in the reference's accumulation order, so continuing from the partial sums reproduces it bit for bit.
//       Runs after the volume launch on the same stream; the processor-patch exchange it needs hides behind the
//       volume launch on the side stream (tcif.cu sweep()).
// What was measured on the way (profiles/r03_variants.md, 400^3 and 200x400x400 per rank): a generic-K twin that
// RE-COMPUTES the boundary cells concurrently on a high-priority stream (round 1) cost the latency/occupancy-bound volume
// sweeps 10-12 %; patch faces inline in the volume launch 7-17 % (a CTA stays resident until its one boundary lane is
// done); a re-computing twin after the volume launch 5.0 ms per stage for 1.5 % of the cells, because the cells next to
// an x wall are one per mesh row: every one of their ~100 loads fetches its own 32-byte sector.  Continuing from the
// partials leaves only the outputs strided.
#pragma once
#include "tcif_device.cuh"
#include "tcif_alpha.cuh"      // PredBatch (kvp_pred_jacobi)

namespace tcif {

template <int K>
__device__ __forceinline__ int owner_of_k(int stride, int fd) {   // fd mod stride, fd < K*stride
    int o = fd;
#pragma unroll
    for (int j = 1; j < K; ++j) o -= (fd >= j * stride) ? stride : 0;
    return o;
}

#ifndef TCIF_TPB_FORCE
#define TCIF_TPB_FORCE 128
#endif
// resident 128-thread CTAs per SM the force sweeps are compiled for (profiles/r03_variants.md: div at 4 = 128 registers
// instead of 168 is 12 % faster, 5 spills and loses; grad needs 120 either way)
#ifndef TCIF_MINB_GRAD
#define TCIF_MINB_GRAD 3
#endif
#ifndef TCIF_MINB_DIV
#define TCIF_MINB_DIV 4
#endif
// register caps chosen from the A/B runs in profiles/r01_variants.md, r03_variants.md
#ifndef TCIF_MINB_SETUP
#define TCIF_MINB_SETUP 4
#endif
#ifndef TCIF_MINB_ITER
#define TCIF_MINB_ITER 4
#endif
#ifndef TCIF_MINB_FINAL
#define TCIF_MINB_FINAL 3
#endif

// slot indices of cell c: neighbour-side faces (fdN, owner oN) and owner-side faces (neighbour nO)
template <int K>
struct Slots {
    int fdN[K], oN[K], nO[K];
    bool vN[K], vO[K];
    __device__ __forceinline__ void load(const DMesh &m, int c) {
#pragma unroll
        for (int k = 0; k < K; ++k) { fdN[k] = m.neiFace[k * m.stride + c]; nO[k] = m.nbrOwn[k * m.stride + c]; }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            vN[k] = fdN[k] >= 0; vO[k] = nO[k] >= 0;
            fdN[k] = vN[k] ? fdN[k] : c;          // padding -> own slot 0 (in range, already cached)
            oN[k] = owner_of_k<K>(m.stride, fdN[k]);
            nO[k] = vO[k] ? nO[k] : c;
        }
    }
};

// The volume sweeps visit the mesh in 256-cell tiles.  Within one launch cells are independent, so the ORDER in
// which CTAs visit the tiles is free: m.tileOrder (host-built, build_tile_order() in tcif.cu) walks banded
// numberings strip by strip so that the +-z neighbour planes are still in L2 when they are gathered
// (profiles/r02_variants.md: at 400^3 one plane of all streams is 70-90 MB, more than L2 keeps).
// vblock counts CTAs in units of TPB_ cells (TPB_ = 128 or 256); c = cell of this thread, cp = cell whose
// own-slot streams this thread prefetches (-1: none).
struct CellIdx { int c, cp; };
template <int TPB_> __device__ __forceinline__ CellIdx cell_of(const DMesh &m, int vblock) {
    static_assert(TPB_ == 128 || TPB_ == 256, "tiles are 256 cells");
    const int tile = TPB_ == 256 ? vblock : (vblock >> 1);
    const int sub = (TPB_ == 256 ? 0 : (vblock & 1) * 128) + threadIdx.x;
    CellIdx r;
    r.c = (m.tileOrder ? m.tileOrder[tile] : tile) * 256 + sub;
    r.cp = -1;
    // Prefetch duty: one warp in four, each lane one 32-byte sector (4 doubles / 8 indices), so that ONE instruction
    // covers 128 cells of a stream -- the same sectors as 32 lanes x 4 warps asking for their own 8 bytes each, with a
    // quarter of the prefetch and address instructions (the sweeps' issue slots are 50-70 % busy; a sweep issues ~40
    // prefetches per cell).  The other warps skip the prefetch blocks (cp = -1, warp-uniform branch).
    const int tp = tile + m.pfTiles;
    if (m.pfTiles > 0 && tp < m.nTiles && (threadIdx.x & 96) == 0) {
        const int q = sub >> 7;                  // which 128-cell half of the tile this warp covers
        const int cp = (m.tileOrder ? m.tileOrder[tp] : tp) * 256 + q * 128 + (threadIdx.x & 31) * 4;
        if (cp < m.nC) r.cp = cp;
    }
    return r;
}

// Rank of cell c among the boundary-adjacent cells (index into bndCells / bndRec / part), -1 for a cell without patch
// faces: from the bit mask (prefix count of the mask word + popcount below the bit), so interior cells read one mask
// word per warp and nothing else.  Evaluated where it is needed (the end of a thread) rather than kept in a register
// across the sweep (the volume kernels are register-starved).  That load sits at the end of every thread with nothing
// left to overlap it -- 11 % of the stall samples of the grad sweep when it went to DRAM -- so the word is prefetched
// twice: into L2 a wave ahead with the slot indices (pf_slots) and into L1 when the thread starts (pick_cell).  The
// grad sweep, which is not capped at 64 registers, simply loads the word first and keeps it (bnd_rank_of).
__device__ __forceinline__ int bnd_rank_of(const DMesh &m, int c, unsigned bw) {
    const unsigned bit = 1u << (c & 31);
    return (bw & bit) ? m.bndWordRank[c >> 5] + __popc(bw & (bit - 1u)) : -1;
}
__device__ __forceinline__ int bnd_rank(const DMesh &m, int c) {
    const int w = c >> 5;
    const unsigned bit = 1u << (c & 31), bw = m.bndMask[w];
    return (bw & bit) ? m.bndWordRank[w] + __popc(bw & (bit - 1u)) : -1;
}
// The cell of this thread in the volume launch (every cell: boundary-adjacent ones stop after their internal faces).
struct CellCtx { int c, cp; };
template <int TPB_, bool MASK_L1 = true>
__device__ __forceinline__ bool pick_cell(const DMesh &m, CellCtx &x) {
    const CellIdx ci = cell_of<TPB_>(m, blockIdx.x);
    x.c = ci.c; x.cp = ci.cp;
    if (x.c >= m.nC) return false;
    if (MASK_L1) asm volatile("prefetch.global.L1 [%0];" ::"l"(m.bndMask + (x.c >> 5)));
    return true;
}
// slot j of the partial record of boundary-adjacent cell bi (SoA: the fix launch reads it coalesced)
__device__ __forceinline__ double &part_at(const DMesh &m, int j, int bi) { return m.part[(size_t)j * m.nBndCells + bi]; }

// L2 prefetch of the contiguous (own-slot) streams a later CTA will read (cp of cell_of()).  The volume
// sweeps are bound by memory latency (ncu: long-scoreboard stalls, ~45 % occupancy), and everything a cell
// reads from HBM for the first time is one of these streams, so a prefetch issued one or two waves ahead
// turns those misses into L2 hits.  No registers are tied up (prefetch has no destination).
template <int K> __device__ __forceinline__ void pf_slots(const DMesh &m, int cp) {
#pragma unroll
    for (int k = 0; k < K; ++k) { pf_l2(m.neiFace + k * m.stride + cp); pf_l2(m.nbrOwn + k * m.stride + cp); }
    pf_l2(m.bndMask + (cp >> 5));            // the word bnd_rank() tests at the end of the thread
}
template <int K> __device__ __forceinline__ void pf_face(const double *a, int stride, int cp) {
#pragma unroll
    for (int k = 0; k < K; ++k) pf_l2(a + k * stride + cp);
}
template <int K> __device__ __forceinline__ void pf_geom(const DMesh &m, int cp) {
    pf_face<K>(m.w, m.stride, cp); pf_face<K>(m.Sfx, m.stride, cp); pf_face<K>(m.Sfy, m.stride, cp); pf_face<K>(m.Sfz, m.stride, cp);
    pf_l2(m.V + cp);
}

// tensor of all pairs from the cell values (phaseSystem.C:1074-1096); stores Tq (pair-summed when a.tsum)
template <int P>
__device__ __forceinline__ void store_pair_tensors(const DMesh &m, const ForceArgs &a, int c, const double *aS, const double (*g)[3], double Tc) {
    double Ts6[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    int q = 0;
#pragma unroll
    for (int p1 = 0; p1 < P; ++p1)
#pragma unroll
        for (int p2 = p1 + 1; p2 < P; ++p2, ++q) {
            double T6[6];
            double sg = sigma_eval(a.sig[q], Tc);
            pair_tensor(aS[p1], aS[p2], g[p1], g[p2], sg, a.deltaN, T6);
            if (a.tsum) {
#pragma unroll
                for (int k = 0; k < 6; ++k) Ts6[k] += T6[k];
            } else {
#pragma unroll
                for (int k = 0; k < 6; ++k) a.Tq[(size_t)(q * 6 + k) * m.stride + c] = T6[k];
            }
        }
    if (a.tsum) {
#pragma unroll
        for (int k = 0; k < 6; ++k) a.Tq[(size_t)k * m.stride + c] = Ts6[k];
    }
}

// ---- K1: grad(alpha_p) for all phases + capillary stress tensors of all pairs ---------------------
// partial record of a boundary-adjacent cell: g[p][d] before /V (3P), alpha_p (P), 1/V, T  ->  4P + 2 slots
template <int P, int K>
__global__ void __launch_bounds__(TCIF_TPB_FORCE, TCIF_MINB_GRAD)
kv_grad_tensor(const __grid_constant__ DMesh m, const __grid_constant__ ForceArgs a) {
    CellCtx x;
    if (!pick_cell<TCIF_TPB_FORCE, false>(m, x)) return;
    const int c = x.c;
    const unsigned bw = m.bndMask[c >> 5];
    Slots<K> s;
    s.load(m, c);
    {
        const int cp = x.cp;
        if (cp >= 0) {
            pf_slots<K>(m, cp); pf_geom<K>(m, cp);
#pragma unroll
            for (int p = 0; p < P; ++p) pf_l2(a.alpha[p] + cp);
            if (a.T) pf_l2(a.T + cp);
        }
    }
    double aS[P], aO[K][P], aN[K][P];
#pragma unroll
    for (int p = 0; p < P; ++p) aS[p] = a.alpha[p][c];
#pragma unroll
    for (int k = 0; k < K; ++k)
#pragma unroll
        for (int p = 0; p < P; ++p) { aO[k][p] = a.alpha[p][s.oN[k]]; aN[k][p] = a.alpha[p][s.nO[k]]; }
    double wN[K], sxN[K], syN[K], szN[K], wO[K], sxO[K], syO[K], szO[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int fd = s.fdN[k], fo = k * m.stride + c;
        wN[k] = m.w[fd]; sxN[k] = m.Sfx[fd]; syN[k] = m.Sfy[fd]; szN[k] = m.Sfz[fd];
        wO[k] = __ldcs(m.w + fo); sxO[k] = __ldcs(m.Sfx + fo); syO[k] = __ldcs(m.Sfy + fo); szO[k] = __ldcs(m.Sfz + fo);
    }
    const double rV = f_rcp(m.V[c]);
    const double Tc = a.T ? a.T[c] : 0.0;
    double g[P][3];
#pragma unroll
    for (int p = 0; p < P; ++p) g[p][0] = g[p][1] = g[p][2] = 0.0;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (s.vN[k]) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double af = f_mad(wN[k], aO[k][p] - aS[p], aS[p]);
                g[p][0] = f_mad(-sxN[k], af, g[p][0]); g[p][1] = f_mad(-syN[k], af, g[p][1]); g[p][2] = f_mad(-szN[k], af, g[p][2]);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (s.vO[k]) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                double af = f_mad(wO[k], aS[p] - aN[k][p], aN[k][p]);
                g[p][0] = f_mad(sxO[k], af, g[p][0]); g[p][1] = f_mad(syO[k], af, g[p][1]); g[p][2] = f_mad(szO[k], af, g[p][2]);
            }
        }
    }
    const int bi = bnd_rank_of(m, c, bw);
    if (bi >= 0) {                           // the patch faces continue from here (kfix_grad_tensor)
#pragma unroll
        for (int p = 0; p < P; ++p) {
#pragma unroll
            for (int d = 0; d < 3; ++d) part_at(m, p * 3 + d, bi) = g[p][d];
            part_at(m, 3 * P + p, bi) = aS[p];
        }
        part_at(m, 4 * P, bi) = rV; part_at(m, 4 * P + 1, bi) = Tc;
        return;
    }
#pragma unroll
    for (int p = 0; p < P; ++p) { g[p][0] = f_div(g[p][0], rV); g[p][1] = f_div(g[p][1], rV); g[p][2] = f_div(g[p][2], rV); }
    if (a.G) {
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int d = 0; d < 3; ++d) a.G[(size_t)(p * 3 + d) * m.stride + c] = g[p][d];
    }
    if (a.Tq) store_pair_tensors<P>(m, a, c, aS, g, Tc);
}

// Packed record of a boundary-adjacent cell (DMesh::bndRec): the patch values of a cell are ONE dependent load away from
// its rank (record -> values) instead of three (bndCellStart -> bndFaces -> bfType/bfPatch -> values), and the helpers
// below issue all loads of the (<= 3) patch faces before the first store, so nothing serialises on a possible alias.
struct BndRec { int n, bf[3], tp[3]; };
__device__ __forceinline__ BndRec bnd_load(const DMesh &m, int bi) {
    const int4 a = __ldg(m.bndRec + 2 * (size_t)bi), b = __ldg(m.bndRec + 2 * (size_t)bi + 1);
    BndRec r;
    r.n = a.x; r.bf[0] = a.y; r.bf[1] = a.z; r.bf[2] = a.w; r.tp[0] = b.x; r.tp[1] = b.y; r.tp[2] = b.z;
    return r;
}

// fix launch of K1: patch faces (gaussGrad: boundary loop after the internal faces), /V, gBnd, G, tensors
template <int P>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_grad_tensor(const __grid_constant__ DMesh m, const __grid_constant__ ForceArgs a) {
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    if (bi >= m.nBndCells) return;
    const int c = m.bndCells[bi];
    const BndRec r = bnd_load(m, bi);
    double g[P][3], aS[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
#pragma unroll
        for (int d = 0; d < 3; ++d) g[p][d] = part_at(m, p * 3 + d, bi);
        aS[p] = part_at(m, 3 * P + p, bi);
    }
    const double rV = part_at(m, 4 * P, bi), Tc = part_at(m, 4 * P + 1, bi);
    if (r.n >= 0) {
        double wb[3], sx[3], sy[3], sz[3], ab[3][P];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int bf = r.bf[k];
            wb[k] = m.wb[bf]; sx[k] = m.Sfbx[bf]; sy[k] = m.Sfby[bf]; sz[k] = m.Sfbz[bf];
#pragma unroll
            for (int p = 0; p < P; ++p) ab[k][p] = a.alphab[p][bf];
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (k < r.n) {
                const bool coupled = (r.tp[k] & 15) == PATCH_PROCESSOR;
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const double af = coupled ? wb[k] * aS[p] + (1.0 - wb[k]) * ab[k][p] : ab[k][p];
                    g[p][0] += sx[k] * af; g[p][1] += sy[k] * af; g[p][2] += sz[k] * af;
                }
            }
        }
    } else {
        for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) {
            const int bf = m.bndFaces[j];
            const bool coupled = m.bfType[bf] == PATCH_PROCESSOR;
            const double wb = m.wb[bf], sx = m.Sfbx[bf], sy = m.Sfby[bf], sz = m.Sfbz[bf];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const double ab = a.alphab[p][bf];
                const double af = coupled ? wb * aS[p] + (1.0 - wb) * ab : ab;
                g[p][0] += sx * af; g[p][1] += sy * af; g[p][2] += sz * af;
            }
        }
    }
#pragma unroll
    for (int p = 0; p < P; ++p) { g[p][0] = f_div(g[p][0], rV); g[p][1] = f_div(g[p][1], rV); g[p][2] = f_div(g[p][2], rV); }
#pragma unroll
    for (int p = 0; p < P; ++p)                  // kept for the patch values of the tensor (kfix_div_tensor)
#pragma unroll
        for (int d = 0; d < 3; ++d) a.gBnd[(size_t)(p * 3 + d) * m.nBndCells + bi] = g[p][d];
    if (a.G) {
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int d = 0; d < 3; ++d) a.G[(size_t)(p * 3 + d) * m.stride + c] = g[p][d];
    }
    if (a.Tq) store_pair_tensors<P>(m, a, c, aS, g, Tc);
}

// S_b . T_b over the patch faces of boundary cell bi (A.4): processor faces interpolate with the neighbour-rank
// tensor, elsewhere T_b is the tensor expression evaluated on the patch values alpha_b, (grad alpha)_b after the
// gaussGrad correction, sigma(T_b); pair q only, or (tsum) all pairs summed in pair order.
template <int P>
__device__ __forceinline__ void div_bnd_faces(const DMesh &m, const ForceArgs &a, int bi, int c, int q, const double *Ts, double *acc) {
    for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) {
        const int bf = m.bndFaces[j];
        const double sx = m.Sfbx[bf], sy = m.Sfby[bf], sz = m.Sfbz[bf];
        double t[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        if (m.bfType[bf] == PATCH_PROCESSOR) {
            const double wb = m.wb[bf];
#pragma unroll
            for (int kk = 0; kk < 6; ++kk)
                t[kk] = wb * Ts[kk] + (1.0 - wb) * a.TqHalo[(size_t)(q * 6 + kk) * m.nBF + bf];
        } else {
            const double mag = m.magSfb[bf];
            const double nx = sx / mag, ny = sy / mag, nz = sz / mag;
            int qq = 0;
            for (int p1 = 0; p1 < P; ++p1)
                for (int p2 = p1 + 1; p2 < P; ++p2, ++qq) {
                    if (!a.tsum && qq != q) continue;
                    double gb[2][3], ab[2];
                    const int pp[2] = {p1, p2};
#pragma unroll
                    for (int s = 0; s < 2; ++s) {
                        const int p = pp[s];
                        const double gc0 = a.gBnd[(size_t)(p * 3 + 0) * m.nBndCells + bi];
                        const double gc1 = a.gBnd[(size_t)(p * 3 + 1) * m.nBndCells + bi];
                        const double gc2 = a.gBnd[(size_t)(p * 3 + 2) * m.nBndCells + bi];
                        ab[s] = a.alphab[p][bf];
                        double sn = 0.0;
                        if (a.bc[p][m.bfPatch[bf]] == BC_FIXED_VALUE) sn = m.dcb[bf] * (ab[s] - a.alpha[p][c]);
                        const double ng = nx * gc0 + ny * gc1 + nz * gc2;
                        const double sc = sn - ng;
                        gb[s][0] = gc0 + nx * sc; gb[s][1] = gc1 + ny * sc; gb[s][2] = gc2 + nz * sc;
                    }
                    const double sg = sigma_eval(a.sig[qq], a.Tb ? a.Tb[bf] : 0.0);
                    double tq[6];
                    pair_tensor(ab[0], ab[1], gb[0], gb[1], sg, a.deltaN, tq);
                    if (a.tsum) {
#pragma unroll
                        for (int kk = 0; kk < 6; ++kk) t[kk] += tq[kk];
                    } else {
#pragma unroll
                        for (int kk = 0; kk < 6; ++kk) t[kk] = tq[kk];
                    }
                }
        }
        acc[0] += sx * t[0] + sy * t[1] + sz * t[2];
        acc[1] += sx * t[1] + sy * t[3] + sz * t[4];
        acc[2] += sx * t[2] + sy * t[4] + sz * t[5];
    }
}

// ---- K2: F = sum_q div(T_q) ---------------------------------------------------------------------------
// partial record: 1/V, then per tensor q: acc[3] over the internal faces, own tensor Ts[6]  ->  1 + 9 nQ slots
template <int P, int K>
__global__ void __launch_bounds__(TCIF_TPB_FORCE, TCIF_MINB_DIV)
kv_div_tensor(const __grid_constant__ DMesh m, const __grid_constant__ ForceArgs a) {
    constexpr int Q = P * (P - 1) / 2;
    CellCtx x;
    if (!pick_cell<TCIF_TPB_FORCE>(m, x)) return;
    const int c = x.c;
    Slots<K> s;
    s.load(m, c);
    {
        const int cp = x.cp;
        if (cp >= 0) {
            pf_slots<K>(m, cp); pf_geom<K>(m, cp);
#pragma unroll
            for (int kk = 0; kk < 6; ++kk) pf_l2(a.Tq + (size_t)kk * m.stride + cp);
        }
    }
    double wN[K], sxN[K], syN[K], szN[K], wO[K], sxO[K], syO[K], szO[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        const int fd = s.fdN[k], fo = k * m.stride + c;
        wN[k] = m.w[fd]; sxN[k] = m.Sfx[fd]; syN[k] = m.Sfy[fd]; szN[k] = m.Sfz[fd];
        wO[k] = __ldcs(m.w + fo); sxO[k] = __ldcs(m.Sfx + fo); syO[k] = __ldcs(m.Sfy + fo); szO[k] = __ldcs(m.Sfz + fo);
    }
    const double rV = f_rcp(m.V[c]);
    double F[3] = {0.0, 0.0, 0.0};
    const int nQ = a.tsum ? 1 : Q;
#pragma unroll 1
    for (int q = 0; q < nQ; ++q) {
        const double *Tq = a.Tq + (size_t)q * 6 * m.stride;
        double Ts[6], To[K][6], Tn[K][6];
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) Ts[kk] = Tq[(size_t)kk * m.stride + c];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int kk = 0; kk < 6; ++kk) {
                To[k][kk] = Tq[(size_t)kk * m.stride + s.oN[k]];
                Tn[k][kk] = Tq[(size_t)kk * m.stride + s.nO[k]];
            }
        double acc[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (s.vN[k]) face_div_acc(wN[k], sxN[k], syN[k], szN[k], To[k], Ts, -1.0, acc);
#pragma unroll
        for (int k = 0; k < K; ++k)
            if (s.vO[k]) face_div_acc(wO[k], sxO[k], syO[k], szO[k], Ts, Tn[k], 1.0, acc);
        const int bi = bnd_rank(m, c);
        if (bi >= 0) {                       // the patch faces continue from here (kfix_div_tensor)
            part_at(m, 0, bi) = rV;
#pragma unroll
            for (int d = 0; d < 3; ++d) part_at(m, 1 + 9 * q + d, bi) = acc[d];
#pragma unroll
            for (int kk = 0; kk < 6; ++kk) part_at(m, 1 + 9 * q + 3 + kk, bi) = Ts[kk];
        }
        F[0] += f_div(acc[0], rV); F[1] += f_div(acc[1], rV); F[2] += f_div(acc[2], rV);
    }
    __stcs(a.F + c, F[0]); __stcs(a.F + (size_t)m.stride + c, F[1]); __stcs(a.F + (size_t)2 * m.stride + c, F[2]);
}

template <int P>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_div_tensor(const __grid_constant__ DMesh m, const __grid_constant__ ForceArgs a) {
    constexpr int Q = P * (P - 1) / 2;
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    if (bi >= m.nBndCells) return;
    const int c = m.bndCells[bi];
    const double rV = part_at(m, 0, bi);
    const int nQ = a.tsum ? 1 : Q;
    double F[3] = {0.0, 0.0, 0.0};
    for (int q = 0; q < nQ; ++q) {
        double acc[3], Ts[6];
#pragma unroll
        for (int d = 0; d < 3; ++d) acc[d] = part_at(m, 1 + 9 * q + d, bi);
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) Ts[kk] = part_at(m, 1 + 9 * q + 3 + kk, bi);
        div_bnd_faces<P>(m, a, bi, c, q, Ts, acc);
        F[0] += f_div(acc[0], rV); F[1] += f_div(acc[1], rV); F[2] += f_div(acc[2], rV);
    }
    a.F[c] = F[0]; a.F[(size_t)m.stride + c] = F[1]; a.F[(size_t)2 * m.stride + c] = F[2];
}

// =====================================================================================================
// MULES sweeps, phase-looped (kvp_*): one thread per cell, ALL phases of the batch handled by that thread.
// The 2K slot indices are loaded once per cell instead of once per (cell, phase), which removes one of
// the two dependent memory round trips (indices -> gathers) for every phase after the first, and the
// per-phase pointers are compile-time offsets into the kernel parameters (NPH is a template argument).
// Arithmetic and order: SURVEY.md A.6 (see also the generic-K kernels k_mules_* in tcif_device.cuh).
// Partial records of a boundary-adjacent cell, per phase p (slot = p*PART_* + j):
//   set-up  : psiMaxn, psiMinn, sumPhiBD, sumPhip, mSumPhim, psi, psi0, Su, Sp            (V is read by the fix launch)
//   iter    : sumlPhip, mSumlPhim, lambda+_C, lambda-_C, Q+, Q-, sumPhip, mSumPhim (both over the internal faces)
// sumPhip / mSumPhim (the totals of the anti-diffusive out- and inflow of a cell) are NOT stored between the sweeps: every
// limiter sweep has phiCorr of all 2K faces in registers and re-derives them with the set-up sweep's operations in the
// set-up sweep's order, i.e. to the same bits (16 B per cell, phase and sweep less to write and to read).
//   final   : sum (faces so far), lambda+_C, lambda-_C, psi0, Su, Sp                      (MODE 2; MODE 0/1: the two lambdas)
// =====================================================================================================
enum { PART_SETUP = 9, PART_ITER = 8, PART_FINAL = 6 };

// finish of the set-up sweep for one cell and phase: extrema coefficients, Q+-, lambda+- of iteration 1
__device__ __forceinline__ void setup_finish(const MulesPh &ph, const MulesParams &prm, int c, double V, double psiC, double psi0, double Su,
                                             double Sp, double psiMaxn, double psiMinn, double sBD, double sP, double sM,
                                             double *lpDst, double *lmDst) {
    psiMaxn = dmin(psiMaxn + prm.extremaCoeff * (prm.psiMax - prm.psiMin), prm.psiMax);
    psiMinn = dmax(psiMinn - prm.extremaCoeff * (prm.psiMax - prm.psiMin), prm.psiMin);
    if (prm.smoothLimiter > TCIF_SMALL) {
        double sl = prm.smoothLimiter;
        psiMaxn = dmin(sl * psiC + (1.0 - sl) * psiMaxn, prm.psiMax);
        psiMinn = dmax(sl * psiC + (1.0 - sl) * psiMinn, prm.psiMin);
    }
    double Qp = V * ((prm.rDeltaT - Sp) * psiMaxn - Su) - (V * prm.rDeltaT) * psi0 + sBD;
    double Qm = V * (Su - (prm.rDeltaT - Sp) * psiMinn) + (V * prm.rDeltaT) * psi0 - sBD;
    __stcs(ph.Qp + c, Qp); __stcs(ph.Qm + c, Qm);
    // limiter iteration 1, cell pass: lambda == 1 so sumlPhip == sumPhip, mSumlPhim == mSumPhim
    lmDst[c] = dmax(dmin((sP + Qp) / (sM + TCIF_ROOTVSMALL), 1.0), 0.0);
    lpDst[c] = dmax(dmin((sM + Qm) / (sP + TCIF_ROOTVSMALL), 1.0), 0.0);
}

template <int K, int NPH>
__global__ void __launch_bounds__(TCIF_TPB, TCIF_MINB_SETUP)
kvp_mules_setup(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b,
                const __grid_constant__ MulesParams prm) {
    CellCtx x;
    if (!pick_cell<TCIF_TPB>(m, x)) return;
    const int c = x.c;
    Slots<K> s;
    s.load(m, c);
    double phiN[K], phiO[K];
#pragma unroll
    for (int k = 0; k < K; ++k) { phiN[k] = prm.phi[s.fdN[k]]; phiO[k] = __ldcs(prm.phi + k * m.stride + c); }
    const double V = m.V[c];
    const int cp = x.cp;
    const bool pf = cp >= 0;
    if (pf) { pf_slots<K>(m, cp); pf_face<K>(prm.phi, m.stride, cp); pf_l2(m.V + cp); }
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        if (pf) {
            pf_l2(ph.psi + cp); pf_face<K>(ph.phiIn, m.stride, cp);
            if (!prm.storeOld) pf_l2(ph.psi0 + cp);
            if (ph.Su) pf_l2(ph.Su + cp);
            if (ph.Sp) pf_l2(ph.Sp + cp);
        }
        const double psiC = ph.psi[c];
        double psiO[K], ppN[K], psiN[K], ppO[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int fo = k * m.stride + c;
            psiO[k] = ph.psi[s.oN[k]]; ppN[k] = ph.phiIn[s.fdN[k]];
            psiN[k] = ph.psi[s.nO[k]]; ppO[k] = ph.phiIn[fo];
        }
        // storeOld: this sweep opens a (sub-)time step, psi.oldTime() := psi rides on it (8 B stored instead of a
        // separate 16 B copy sweep, and psi0 need not be read)
        double psi0;
        if (prm.storeOld) { psi0 = psiC; __stcs(ph.psi0w + c, psiC); }
        else psi0 = __ldcs(ph.psi0 + c);
        const double Su = ph.Su ? __ldcs(ph.Su + c) : 0.0, Sp = ph.Sp ? __ldcs(ph.Sp + c) : 0.0;
        double psiMaxn = prm.psiMin, psiMinn = prm.psiMax, sBD = 0.0, sP = 0.0, sM = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (s.vN[k]) {
                double bd = upwind_flux(phiN[k], psiO[k], psiC);
                double pp = ppN[k];
                if (prm.addUpwind) pp = pp + bd;
                double pc = pp - bd;
                psiMaxn = dmax(psiMaxn, psiO[k]); psiMinn = dmin(psiMinn, psiO[k]);
                sBD -= bd;
                if (pc > 0) sM += pc; else sP -= pc;
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (s.vO[k]) {
                double bd = upwind_flux(phiO[k], psiC, psiN[k]);
                double pp = ppO[k];
                if (prm.addUpwind) pp = pp + bd;
                double pc = pp - bd;
                ph.phiCorr[k * m.stride + c] = pc;
                psiMaxn = dmax(psiMaxn, psiN[k]); psiMinn = dmin(psiMinn, psiN[k]);
                sBD += bd;
                if (pc > 0) sP += pc; else sM -= pc;
            }
        }
        const int bi = bnd_rank(m, c);
        if (bi >= 0) {                       // the patch faces continue from here (kfix_mules_setup)
            const int o = p * PART_SETUP;
            part_at(m, o, bi) = psiMaxn; part_at(m, o + 1, bi) = psiMinn; part_at(m, o + 2, bi) = sBD;
            part_at(m, o + 3, bi) = sP; part_at(m, o + 4, bi) = sM; part_at(m, o + 5, bi) = psiC;
            part_at(m, o + 6, bi) = psi0; part_at(m, o + 7, bi) = Su; part_at(m, o + 8, bi) = Sp;
        }
        // (a boundary-adjacent cell gets provisional values here, overwritten by the fix launch: no divergent exit)
        setup_finish(ph, prm, c, V, psiC, psi0, Su, Sp, psiMaxn, psiMinn, sBD, sP, sM, ph.lpA, ph.lmA);
    }
}

// patch face of the set-up sweep: phiBD_b, phiCorr_b, lambda_b = 1, local extrema, sums
__device__ __forceinline__ void bnd_setup_face(const MulesPh &ph, const MulesParams &prm, int bf, bool coupled, int kind, double psb,
                                               double phb, double pp, double psiC, double &psiMaxn, double &psiMinn, double &sBD,
                                               double &sP, double &sM) {
    double bd;
    if (coupled) {
        const double w = phb >= 0 ? 1.0 : 0.0;
        const double up = phb * (w * psiC + (1.0 - w) * psb);
        if (prm.addUpwind) pp = pp + up;
        bd = up;
    } else {
        if (prm.addUpwind) pp = pp + phb * psb;
        bd = pp;                       // phiBD_b := phiPsi_b on non-coupled patches
    }
    const double pc = pp - bd;
    ph.phiBDb[bf] = bd; ph.phiCorrb[bf] = pc; ph.lamb[bf] = 1.0;
    if (coupled || kind == BC_FIXED_VALUE || prm.boundaryExtrema == 1) {
        psiMaxn = dmax(psiMaxn, psb); psiMinn = dmin(psiMinn, psb);
    } else {
        psiMaxn = dmax(psiMaxn, prm.psiMax); psiMinn = dmin(psiMinn, prm.psiMin);
    }
    sBD += bd;
    if (pc > 0) sP += pc; else sM -= pc;
}

template <int NPH>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_mules_setup(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b, const __grid_constant__ MulesParams prm) {
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    if (bi >= m.nBndCells) return;
    const int c = m.bndCells[bi];
    const BndRec r = bnd_load(m, bi);
    const double V = m.V[c];
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        const int o = p * PART_SETUP;
        double psiMaxn = part_at(m, o, bi), psiMinn = part_at(m, o + 1, bi), sBD = part_at(m, o + 2, bi), sP = part_at(m, o + 3, bi),
               sM = part_at(m, o + 4, bi);
        const double psiC = part_at(m, o + 5, bi), psi0 = part_at(m, o + 6, bi), Su = part_at(m, o + 7, bi), Sp = part_at(m, o + 8, bi);
        if (r.n >= 0) {
            double psb[3], phb[3], pin[3];
            int kind[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {           // every load first (unused slots repeat face 0: in range)
                const int bf = r.bf[k];
                psb[k] = ph.psib[bf]; phb[k] = prm.phib[bf]; pin[k] = ph.phiInb[bf]; kind[k] = ph.bc[r.tp[k] >> 4];
            }
#pragma unroll
            for (int k = 0; k < 3; ++k)
                if (k < r.n) bnd_setup_face(ph, prm, r.bf[k], (r.tp[k] & 15) == PATCH_PROCESSOR, kind[k], psb[k], phb[k], pin[k], psiC,
                                            psiMaxn, psiMinn, sBD, sP, sM);
        } else {
            for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) {
                const int bf = m.bndFaces[j];
                bnd_setup_face(ph, prm, bf, m.bfType[bf] == PATCH_PROCESSOR, ph.bc[m.bfPatch[bf]], ph.psib[bf], prm.phib[bf], ph.phiInb[bf],
                               psiC, psiMaxn, psiMinn, sBD, sP, sM);
            }
        }
        setup_finish(ph, prm, c, V, psiC, psi0, Su, Sp, psiMaxn, psiMinn, sBD, sP, sM, ph.lpA, ph.lmA);
    }
}

template <int K, bool FIRST, int NPH>
__global__ void __launch_bounds__(TCIF_TPB, TCIF_MINB_ITER)
kvp_mules_iter(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b) {
    CellCtx x;
    if (!pick_cell<TCIF_TPB>(m, x)) return;
    const int c = x.c;
    Slots<K> s;
    s.load(m, c);
    const int cp = x.cp;
    const bool pf = cp >= 0;
    if (pf) pf_slots<K>(m, cp);
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        const double *lpIn = ph.lpIn, *lmIn = ph.lmIn, *lamIn = ph.lamIn;      // ping-pong resolved on the host
        double *lpOut = ph.lpOut, *lmOut = ph.lmOut, *lamOut = ph.lamOut;
        if (pf) {
            pf_face<K>(ph.phiCorr, m.stride, cp);
            if (!FIRST) pf_face<K>(lamIn, m.stride, cp);
            pf_l2_keep(lpIn + cp); pf_l2_keep(lmIn + cp); pf_l2(ph.Qp + cp); pf_l2(ph.Qm + cp);
        }
        const double lpC = ld_keep(lpIn + c), lmC = ld_keep(lmIn + c);
        double pcN[K], laN[K], lpO[K], lmO[K], pcO[K], laO[K], lpN[K], lmN[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int fo = k * m.stride + c;
            pcN[k] = ph.phiCorr[s.fdN[k]]; laN[k] = FIRST ? 1.0 : lamIn[s.fdN[k]];
            lpO[k] = lpIn[s.oN[k]]; lmO[k] = lmIn[s.oN[k]];
            pcO[k] = ld_keep(ph.phiCorr + fo); laO[k] = FIRST ? 1.0 : ld_keep(lamIn + fo);
            lpN[k] = lpIn[s.nO[k]]; lmN[k] = lmIn[s.nO[k]];
        }
        const double Qp = __ldcs(ph.Qp + c), Qm = __ldcs(ph.Qm + c);
        double slp = 0.0, slm = 0.0, sP = 0.0, sM = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (s.vN[k]) {
                double lam = lambda_face(laN[k], pcN[k], lpO[k], lmO[k], lpC, lmC);
                double l = lam * pcN[k];
                if (l > 0) slm += l; else slp -= l;
                if (pcN[k] > 0) sM += pcN[k]; else sP -= pcN[k];
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (s.vO[k]) {
                double lam = lambda_face(laO[k], pcO[k], lpC, lmC, lpN[k], lmN[k]);
                lamOut[k * m.stride + c] = lam;
                double l = lam * pcO[k];
                if (l > 0) slp += l; else slm -= l;
                if (pcO[k] > 0) sP += pcO[k]; else sM -= pcO[k];
            }
        }
        const int bi = bnd_rank(m, c);
        if (bi >= 0) {                       // the patch faces continue from here (kfix_mules_iter)
            const int o = p * PART_ITER;
            part_at(m, o, bi) = slp; part_at(m, o + 1, bi) = slm; part_at(m, o + 2, bi) = lpC; part_at(m, o + 3, bi) = lmC;
            part_at(m, o + 4, bi) = Qp; part_at(m, o + 5, bi) = Qm; part_at(m, o + 6, bi) = sP; part_at(m, o + 7, bi) = sM;
        }
        lmOut[c] = dmax(dmin((slp + Qp) / (sM + TCIF_ROOTVSMALL), 1.0), 0.0);
        lpOut[c] = dmax(dmin((slm + Qm) / (sP + TCIF_ROOTVSMALL), 1.0), 0.0);
    }
}

// patch faces of a limiter iteration: face rule + contribution to the next cell pass
template <int NPH>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_mules_iter(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b) {
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    if (bi >= m.nBndCells) return;
    const int c = m.bndCells[bi];
    const BndRec r = bnd_load(m, bi);
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        const int o = p * PART_ITER;
        double slp = part_at(m, o, bi), slm = part_at(m, o + 1, bi);
        const double lpC = part_at(m, o + 2, bi), lmC = part_at(m, o + 3, bi), Qp = part_at(m, o + 4, bi), Qm = part_at(m, o + 5, bi);
        double sP = part_at(m, o + 6, bi), sM = part_at(m, o + 7, bi);
        if (r.n >= 0) {
            double pc[3], la[3], bd[3], lpH[3], lmH[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int bf = r.bf[k];
                pc[k] = ph.phiCorrb[bf]; la[k] = ph.lamb[bf]; bd[k] = ph.phiBDb[bf]; lpH[k] = ph.lpHalo[bf]; lmH[k] = ph.lmHalo[bf];
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (k < r.n) {
                    const double lam = lambda_bface(r.tp[k] & 15, la[k], pc[k], bd[k], lpC, lmC, lpH[k], lmH[k]);
                    ph.lamb[r.bf[k]] = lam;
                    const double l = lam * pc[k];
                    if (l > 0) slp += l; else slm -= l;
                    if (pc[k] > 0) sP += pc[k]; else sM -= pc[k];
                }
            }
        } else {
            for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) {
                const int bf = m.bndFaces[j];
                const double pc = ph.phiCorrb[bf];
                const double lam = lambda_bface(m.bfType[bf], ph.lamb[bf], pc, ph.phiBDb[bf], lpC, lmC, ph.lpHalo[bf], ph.lmHalo[bf]);
                ph.lamb[bf] = lam;
                const double l = lam * pc;
                if (l > 0) slp += l; else slm -= l;
                if (pc > 0) sP += pc; else sM -= pc;
            }
        }
        ph.lmOut[c] = dmax(dmin((slp + Qp) / (sM + TCIF_ROOTVSMALL), 1.0), 0.0);
        ph.lpOut[c] = dmax(dmin((slm + Qm) / (sP + TCIF_ROOTVSMALL), 1.0), 0.0);
    }
}

// MULES::limitSum on one face, all phases in registers (same operations as k_limit_sum)
template <int NPH> __device__ __forceinline__ void limit_sum_face(double *v) {
    double sumPos = 0.0, sumNeg = 0.0;
#pragma unroll
    for (int p = 0; p < NPH; ++p) { if (v[p] > 0) sumPos += v[p]; else sumNeg += v[p]; }
    const double sum = sumPos + sumNeg;
    if (sum > 0 && sumPos > TCIF_VSMALL) {
        const double lambda = -sumNeg / sumPos;
#pragma unroll
        for (int p = 0; p < NPH; ++p) if (v[p] > 0) v[p] = v[p] * lambda;
    } else if (sum < 0 && sumNeg < -TCIF_VSMALL) {
        const double lambda = -sumPos / sumNeg;
#pragma unroll
        for (int p = 0; p < NPH; ++p) if (v[p] < 0) v[p] = v[p] * lambda;
    }
}

// MODE 0: returnCorr  -> phiPsi = lambda*phiCorr           (owner-side faces)
// MODE 1: !returnCorr -> phiPsi = phiBD + lambda*phiCorr   (owner-side faces)
// MODE 2: explicitSolve: MODE 1 + surfaceIntegrate + Euler update into psiNew
// LSUM (MODE 0 only, batch = all phases): MULES::limitSum of the owner-side faces applied in registers
// before the store, which saves the separate read-modify-write sweep over the P flux arrays.
template <int K, int MODE, bool FIRST, int NPH, bool LSUM>
__global__ void __launch_bounds__(TCIF_TPB, MODE == 2 ? TCIF_MINB_FINAL : 4)      // MODE 0/1 fit 64 registers, MODE 2 spills there
kvp_mules_final(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b,
                const __grid_constant__ MulesParams prm, int nIter) {
    static_assert(!LSUM || MODE == 0, "limitSum is fused into the returnCorr=true apply only");
    CellCtx x;
    if (!pick_cell<TCIF_TPB>(m, x)) return;
    const int c = x.c;
    Slots<K> s;
    s.load(m, c);
    double phN[K], phO[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
        if (MODE == 2) phN[k] = prm.phi[s.fdN[k]];
        if (MODE >= 1) phO[k] = __ldcs(prm.phi + k * m.stride + c);
    }
    const double V = (MODE == 2) ? m.V[c] : 1.0;
    const int cp = x.cp;
    const bool pf = cp >= 0;
    if (pf) {
        pf_slots<K>(m, cp);
        if (MODE >= 1) pf_face<K>(prm.phi, m.stride, cp);
        if (MODE == 2) pf_l2(m.V + cp);
    }
    double flAll[LSUM ? K : 1][LSUM ? NPH : 1];
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        const double *lpIn = ph.lpIn, *lmIn = ph.lmIn, *lamIn = ph.lamIn;      // ping-pong resolved on the host
        if (pf) {
            pf_face<K>(ph.phiCorr, m.stride, cp);
            if (!FIRST) pf_face<K>(lamIn, m.stride, cp);
            pf_l2_keep(lpIn + cp); pf_l2_keep(lmIn + cp);
            if (MODE >= 1) pf_l2_keep(ph.psi + cp);
            if (MODE == 2) {
                pf_l2(ph.psi0 + cp);
                if (ph.Su) pf_l2(ph.Su + cp);
                if (ph.Sp) pf_l2(ph.Sp + cp);
            }
        }
        const double lpC = ld_keep(lpIn + c), lmC = ld_keep(lmIn + c);
        const double psiC = (MODE >= 1) ? ld_keep(ph.psi + c) : 0.0;
        double pcN[K], laN[K], lpO[K], lmO[K], psO[K];
        double pcO[K], laO[K], lpN[K], lmN[K], psN[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int fo = k * m.stride + c;
            if (MODE == 2) {
                pcN[k] = ph.phiCorr[s.fdN[k]]; laN[k] = FIRST ? 1.0 : lamIn[s.fdN[k]];
                lpO[k] = lpIn[s.oN[k]]; lmO[k] = lmIn[s.oN[k]];
                psO[k] = ph.psi[s.oN[k]];
            }
            pcO[k] = ld_keep(ph.phiCorr + fo); laO[k] = FIRST ? 1.0 : ld_keep(lamIn + fo);
            lpN[k] = lpIn[s.nO[k]]; lmN[k] = lmIn[s.nO[k]];
            if (MODE >= 1) psN[k] = ph.psi[s.nO[k]];
        }
        double psi0 = 0.0, Su = 0.0, Sp = 0.0;
        if (MODE == 2) {
            psi0 = __ldcs(ph.psi0 + c);
            Su = ph.Su ? __ldcs(ph.Su + c) : 0.0; Sp = ph.Sp ? __ldcs(ph.Sp + c) : 0.0;
        }
        double sum = 0.0;
        if (MODE == 2) {
#pragma unroll
            for (int k = 0; k < K; ++k) {
                if (s.vN[k]) {
                    double lam = laN[k];
                    if (nIter > 0) lam = lambda_face(lam, pcN[k], lpO[k], lmO[k], lpC, lmC);
                    double bd = upwind_flux(phN[k], psO[k], psiC);
                    sum -= bd + lam * pcN[k];
                }
            }
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (s.vO[k]) {
                double lam = laO[k];
                if (nIter > 0) lam = lambda_face(lam, pcO[k], lpC, lmC, lpN[k], lmN[k]);
                double fl;
                if (MODE == 0) fl = pcO[k] * lam;
                else           fl = upwind_flux(phO[k], psiC, psN[k]) + lam * pcO[k];
                if (LSUM) flAll[k][p] = fl;
                else ph.phiPsi[k * m.stride + c] = fl;
                sum += fl;
            }
        }
        const int bi = bnd_rank(m, c);
        if (bi >= 0) {                       // the patch faces continue from here (kfix_mules_final)
            const int o = p * PART_FINAL;
            part_at(m, o, bi) = sum; part_at(m, o + 1, bi) = lpC; part_at(m, o + 2, bi) = lmC;
            if (MODE == 2) { part_at(m, o + 3, bi) = psi0; part_at(m, o + 4, bi) = Su; part_at(m, o + 5, bi) = Sp; }
        }
        if (MODE == 2) {
            double d = sum / V;
            ph.psiNew[c] = (psi0 * prm.rDeltaT + Su - d) / (prm.rDeltaT - Sp);
        }
    }
    if (LSUM) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (s.vO[k]) {
                limit_sum_face<NPH>(flAll[k]);
#pragma unroll
                for (int p = 0; p < NPH; ++p) b.ph[p].phiPsi[k * m.stride + c] = flAll[k][p];
            }
        }
    }
}

// patch faces of the last face pass: lambda_b, limited patch flux, contribution to surfaceIntegrate; Euler update (MODE 2)
template <int MODE, int NPH>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_mules_final(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b, const __grid_constant__ MulesParams prm, int nIter) {
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    if (bi >= m.nBndCells) return;
    const int c = m.bndCells[bi];
    const BndRec r = bnd_load(m, bi);
    const double V = (MODE == 2) ? m.V[c] : 1.0;
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        const int o = p * PART_FINAL;
        double sum = part_at(m, o, bi);
        const double lpC = part_at(m, o + 1, bi), lmC = part_at(m, o + 2, bi);
        if (r.n >= 0) {
            double pc[3], la[3], bd[3], lpH[3], lmH[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int bf = r.bf[k];
                pc[k] = ph.phiCorrb[bf]; la[k] = ph.lamb[bf]; bd[k] = ph.phiBDb[bf]; lpH[k] = ph.lpHalo[bf]; lmH[k] = ph.lmHalo[bf];
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (k < r.n) {
                    double lam = la[k];
                    if (nIter > 0) lam = lambda_bface(r.tp[k] & 15, lam, pc[k], bd[k], lpC, lmC, lpH[k], lmH[k]);
                    ph.lamb[r.bf[k]] = lam;
                    const double fl = (MODE == 0) ? pc[k] * lam : bd[k] + lam * pc[k];
                    ph.phiPsib[r.bf[k]] = fl;
                    sum += fl;
                }
            }
        } else {
            for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) {
                const int bf = m.bndFaces[j];
                const double pc = ph.phiCorrb[bf], bd = ph.phiBDb[bf];
                double lam = ph.lamb[bf];
                if (nIter > 0) lam = lambda_bface(m.bfType[bf], lam, pc, bd, lpC, lmC, ph.lpHalo[bf], ph.lmHalo[bf]);
                ph.lamb[bf] = lam;
                const double fl = (MODE == 0) ? pc * lam : bd + lam * pc;
                ph.phiPsib[bf] = fl;
                sum += fl;
            }
        }
        if (MODE == 2) {
            const double psi0 = part_at(m, o + 3, bi), Su = part_at(m, o + 4, bi), Sp = part_at(m, o + 5, bi);
            double d = sum / V;
            ph.psiNew[c] = (psi0 * prm.rDeltaT + Su - d) / (prm.rDeltaT - Sp);
        }
    }
}


// ---- explicitSolve split in two sweeps: kvp_mules_final<MODE 1> stores the limited fluxes of the owner-side faces, then
// this sweep does surfaceIntegrate + the Euler update from the STORED fluxes (one gather per neighbour-side face instead
// of re-deriving that face's lambda and upwind flux from five).  Same values in the same order as MODE 2, so the bits
// agree: the stored flux of a neighbour-side face is what MODE 2 recomputes from the same operands.
// partial record (PART_FINAL layout): slot 0 = sum over the internal faces, 3..5 = psi0, Su, Sp
template <int K, int NPH>
__global__ void __launch_bounds__(TCIF_TPB, 4)
kvp_mules_update(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b, const __grid_constant__ MulesParams prm) {
    CellCtx x;
    if (!pick_cell<TCIF_TPB>(m, x)) return;
    const int c = x.c;
    Slots<K> s;
    s.load(m, c);
    const double V = m.V[c];
    const int cp = x.cp;
    const bool pf = cp >= 0;
    if (pf) { pf_slots<K>(m, cp); pf_l2(m.V + cp); }
    const int bi = bnd_rank(m, c);
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        if (pf) {
            pf_face<K>(ph.phiPsi, m.stride, cp); pf_l2(ph.psi0 + cp);
            if (ph.Su) pf_l2(ph.Su + cp);
            if (ph.Sp) pf_l2(ph.Sp + cp);
        }
        double fN[K], fO[K];
#pragma unroll
        for (int k = 0; k < K; ++k) { fN[k] = ph.phiPsi[s.fdN[k]]; fO[k] = ph.phiPsi[k * m.stride + c]; }
        const double psi0 = __ldcs(ph.psi0 + c);
        const double Su = ph.Su ? __ldcs(ph.Su + c) : 0.0, Sp = ph.Sp ? __ldcs(ph.Sp + c) : 0.0;
        double sum = 0.0;
#pragma unroll
        for (int k = 0; k < K; ++k) if (s.vN[k]) sum -= fN[k];
#pragma unroll
        for (int k = 0; k < K; ++k) if (s.vO[k]) sum += fO[k];
        if (bi >= 0) {                       // the patch faces continue from here (kfix_mules_update)
            const int o = p * PART_FINAL;
            part_at(m, o, bi) = sum; part_at(m, o + 3, bi) = psi0; part_at(m, o + 4, bi) = Su; part_at(m, o + 5, bi) = Sp;
        }
        const double d = sum / V;
        ph.psiNew[c] = (psi0 * prm.rDeltaT + Su - d) / (prm.rDeltaT - Sp);
    }
}

template <int NPH>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_mules_update(const __grid_constant__ DMesh m, const __grid_constant__ MulesBatch b, const __grid_constant__ MulesParams prm) {
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    if (bi >= m.nBndCells) return;
    const int c = m.bndCells[bi];
    const BndRec r = bnd_load(m, bi);
    const double V = m.V[c];
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        const MulesPh &ph = b.ph[p];
        const int o = p * PART_FINAL;
        double sum = part_at(m, o, bi);
        if (r.n >= 0) {
            double fl[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) fl[k] = ph.phiPsib[r.bf[k]];
#pragma unroll
            for (int k = 0; k < 3; ++k) if (k < r.n) sum += fl[k];
        } else {
            for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) sum += ph.phiPsib[m.bndFaces[j]];
        }
        const double psi0 = part_at(m, o + 3, bi), Su = part_at(m, o + 4, bi), Sp = part_at(m, o + 5, bi);
        const double d = sum / V;
        ph.psiNew[c] = (psi0 * prm.rDeltaT + Su - d) / (prm.rDeltaT - Sp);
    }
}

// ---- Jacobi sweep of the implicit upwind predictor, phase-looped (see k_pred_jacobi, tcif_alpha.cuh) -----------
// The matrix coefficients max(phi,0) / -min(phi,0) and the slot indices are shared by the phases of the batch:
// loaded once per cell.  Same operations in the same order as k_pred_jacobi (bit-identical iterates).
// partial record of a boundary-adjacent cell, per phase: acc over the internal faces, x_C, diag (kfix_pred_jacobi adds
// the processor-patch faces -- the only patch faces with an off-diagonal coefficient -- and finishes the cell)
enum { PART_PRED = 3 };

// max |dpsi| per phase: warp shuffle, then one CTA-level value, then at most one atomic per CTA and phase -- and only if
// it can still raise the running maximum (one atomic per warp on three addresses serialised in L2 and cost more than
// the sweep itself)
template <int NPH>
__device__ __forceinline__ void pred_reduce_max(const PredBatch &b, const double *dl) {
    __shared__ double red[NPH][TCIF_TPB / 32];
#pragma unroll
    for (int p = 0; p < NPH; ++p) {
        double d = dl[p];
        for (int o = 16; o > 0; o >>= 1) d = dmax(d, __shfl_xor_sync(0xffffffffu, d, o));
        if ((threadIdx.x & 31) == 0) red[p][threadIdx.x >> 5] = d;
    }
    __syncthreads();
    if (threadIdx.x < NPH) {
        double d = 0.0;
#pragma unroll
        for (int w = 0; w < TCIF_TPB / 32; ++w) d = dmax(d, red[threadIdx.x][w]);
        if (d > 0.0 && d > *(volatile double *)(b.dmax + threadIdx.x)) atomic_max_nonneg(b.dmax + threadIdx.x, d);
    }
}

template <int K, int NPH>
__global__ void __launch_bounds__(TCIF_TPB, 4)
kvp_pred_jacobi(const __grid_constant__ DMesh m, const __grid_constant__ PredBatch b) {
    const CellIdx ci = cell_of<TCIF_TPB>(m, blockIdx.x);
    const int c = ci.c;
    double dl[NPH];
#pragma unroll
    for (int p = 0; p < NPH; ++p) dl[p] = 0.0;
    if (c < m.nC) {
        Slots<K> s;
        s.load(m, c);
        double cN[K], cO[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            cN[k] = dmax(b.phi[s.fdN[k]], 0.0);
            cO[k] = dmin(__ldcs(b.phi + k * m.stride + c), 0.0);
        }
        const int cp = ci.cp;
        if (cp >= 0) { pf_slots<K>(m, cp); pf_face<K>(b.phi, m.stride, cp); }
        const int bi = bnd_rank(m, c);
#pragma unroll
        for (int p = 0; p < NPH; ++p) {
            const PredPh &ph = b.ph[p];
            if (!ph.active) continue;
            if (cp >= 0) { pf_l2(ph.rhs + cp); pf_l2(ph.diag + cp); pf_l2(ph.x + cp); }
            double xO[K], xN[K];
#pragma unroll
            for (int k = 0; k < K; ++k) { xO[k] = ph.x[s.oN[k]]; xN[k] = ph.x[s.nO[k]]; }
            const double xC = ph.x[c], dg = __ldcs(ph.diag + c);
            double acc = __ldcs(ph.rhs + c);
#pragma unroll
            for (int k = 0; k < K; ++k) if (s.vN[k]) acc += cN[k] * xO[k];
#pragma unroll
            for (int k = 0; k < K; ++k) if (s.vO[k]) acc -= cO[k] * xN[k];
            if (bi >= 0) {                   // the processor-patch faces continue from here (kfix_pred_jacobi)
                const int o = p * PART_PRED;
                part_at(m, o, bi) = acc; part_at(m, o + 1, bi) = xC; part_at(m, o + 2, bi) = dg;
            } else {
                const double xn = acc / dg;
                ph.xn[c] = xn;
                dl[p] = fabs(xn - xC);
            }
        }
    }
    pred_reduce_max<NPH>(b, dl);
}

template <int NPH>
__global__ void __launch_bounds__(TCIF_TPB)
kfix_pred_jacobi(const __grid_constant__ DMesh m, const __grid_constant__ PredBatch b) {
    const int bi = blockIdx.x * blockDim.x + threadIdx.x;
    double dl[NPH];
#pragma unroll
    for (int p = 0; p < NPH; ++p) dl[p] = 0.0;
    if (bi < m.nBndCells) {
        const int c = m.bndCells[bi];
        const BndRec r = bnd_load(m, bi);
        double cb[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) cb[k] = (r.n >= 0) ? dmin(b.phib[r.bf[k]], 0.0) : 0.0;
#pragma unroll
        for (int p = 0; p < NPH; ++p) {
            const PredPh &ph = b.ph[p];
            if (!ph.active) continue;
            const int o = p * PART_PRED;
            double acc = part_at(m, o, bi);
            const double xC = part_at(m, o + 1, bi), dg = part_at(m, o + 2, bi);
            if (r.n >= 0) {
                double xb[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) xb[k] = ph.psib[r.bf[k]];
#pragma unroll
                for (int k = 0; k < 3; ++k) if (k < r.n && (r.tp[k] & 15) == PATCH_PROCESSOR) acc -= cb[k] * xb[k];
            } else {
                for (int j = m.bndCellStart[bi], e = m.bndCellStart[bi + 1]; j < e; ++j) {
                    const int bf = m.bndFaces[j];
                    if (m.bfType[bf] == PATCH_PROCESSOR) acc -= dmin(b.phib[bf], 0.0) * ph.psib[bf];
                }
            }
            const double xn = acc / dg;
            ph.xn[c] = xn;
            dl[p] = fabs(xn - xC);
        }
    }
    pred_reduce_max<NPH>(b, dl);
}

}  // namespace tcif
