// mesh_host.cpp -- host-side mesh model behind include/tcif_mesh.h (plain C++17, no CUDA).
// Builds what the OpenFOAM shim would read off an fvMesh: single-block hex meshes in blockMesh
// ordering, and processor sub-meshes in decomposePar conventions.
// Checked bit-exactly against the independent numpy restatement oracle/mesh_oracle.py.
#include "mesh_host.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

using namespace tcif_host;
namespace tcif_host { std::string &last_error() { thread_local std::string e; return e; } }
#define g_err (tcif_host::last_error())

namespace {

// block index range of `simple`: cells i with (i*p)/n == b
void block_range(int n, int p, int b, int &lo, int &hi) {
    lo = n; hi = 0;
    for (int i = 0; i < n; ++i)
        if ((int)(((long long)i * p) / n) == b) { lo = std::min(lo, i); hi = std::max(hi, i + 1); }
    if (lo >= hi) failf("simple decomposition: empty block %d of %d over %d cells", b, p, n);
}

tcif_mesh *box_block(int nx, int ny, int nz, const double *xs, const double *ys, const double *zs,
                     const int32_t pt[6], int px, int py, int pz, int rank) {
    if (nx < 1 || ny < 1 || nz < 1 || px < 1 || py < 1 || pz < 1) failf("box_block: bad sizes");
    if (rank < 0 || rank >= px * py * pz) failf("box_block: rank %d outside %dx%dx%d", rank, px, py, pz);
    if ((long long)nx * ny * nz > 0x7fffffffLL) failf("box_block: more than 2^31 cells");
    const int bi = rank % px, bj = (rank / px) % py, bk = rank / (px * py);
    int i0, i1, j0, j1, k0, k1;
    block_range(nx, px, bi, i0, i1); block_range(ny, py, bj, j0, j1); block_range(nz, pz, bk, k0, k1);
    const int nxl = i1 - i0, nyl = j1 - j0, nzl = k1 - k0;
    std::vector<double> dx(nx), dy(ny), dz(nz), xc(nx), yc(ny), zc(nz);
    for (int i = 0; i < nx; ++i) { dx[i] = xs[i + 1] - xs[i]; xc[i] = 0.5 * (xs[i] + xs[i + 1]); }
    for (int j = 0; j < ny; ++j) { dy[j] = ys[j + 1] - ys[j]; yc[j] = 0.5 * (ys[j] + ys[j + 1]); }
    for (int k = 0; k < nz; ++k) { dz[k] = zs[k + 1] - zs[k]; zc[k] = 0.5 * (zs[k] + zs[k + 1]); }
    auto gid = [&](int i, int j, int k) { return (int64_t)i + (int64_t)nx * (j + (int64_t)ny * k); };
    auto lid = [&](int i, int j, int k) { return (int32_t)((i - i0) + nxl * ((j - j0) + nyl * (k - k0))); };
    auto cc = [&](int i, int j, int k) { return V3{xc[i], yc[j], zc[k]}; };
    // global index of the first internal face owned by global cell (i,j,k) (upper-triangular order)
    std::vector<int64_t> planeStart(nz + 1, 0);
    for (int k = 0; k < nz; ++k)
        planeStart[k + 1] = planeStart[k] + (int64_t)(nx - 1) * ny + (int64_t)nx * (ny - 1) + (k < nz - 1 ? (int64_t)nx * ny : 0);
    auto faceBase = [&](int i, int j, int k) {
        int64_t f = planeStart[k];
        int64_t zrow = (k < nz - 1) ? nx : 0;
        f += (int64_t)j * ((nx - 1) + nx + zrow);            // full rows j' < j (all have j' < ny-1)
        f += (int64_t)i * (1 + (j < ny - 1 ? 1 : 0) + (k < nz - 1 ? 1 : 0));
        return f;
    };
    auto fx = [&](int i, int j, int k) { return faceBase(i, j, k); };
    auto fy = [&](int i, int j, int k) { return faceBase(i, j, k) + (i < nx - 1 ? 1 : 0); };
    auto fz = [&](int i, int j, int k) { return faceBase(i, j, k) + (i < nx - 1 ? 1 : 0) + (j < ny - 1 ? 1 : 0); };

    tcif_mesh *m = new tcif_mesh();
    const size_t nC = (size_t)nxl * nyl * nzl;
    m->nCells = (int32_t)nC;
    m->V.resize(nC); m->C.resize(3 * nC); m->cellGlobal.resize(nC);
    const size_t nIFest = 3 * nC;
    m->owner.reserve(nIFest); m->neighbour.reserve(nIFest); m->Sf.reserve(3 * nIFest); m->Cf.reserve(3 * nIFest);
    m->w.reserve(nIFest); m->faceAddr.reserve(nIFest);
    for (int k = k0; k < k1; ++k)
        for (int j = j0; j < j1; ++j)
            for (int i = i0; i < i1; ++i) {
                int32_t c = lid(i, j, k);
                m->V[c] = dx[i] * dy[j] * dz[k];
                m->C[3 * (size_t)c] = xc[i]; m->C[3 * (size_t)c + 1] = yc[j]; m->C[3 * (size_t)c + 2] = zc[k];
                m->cellGlobal[c] = gid(i, j, k);
                V3 cP = cc(i, j, k);
                if (i + 1 < i1) {
                    V3 S{dy[j] * dz[k], 0.0, 0.0}, cf{xs[i + 1], yc[j], zc[k]};
                    m->push_face(c, lid(i + 1, j, k), S, cf, linear_weight(S, cf, cP, cc(i + 1, j, k)), fx(i, j, k));
                }
                if (j + 1 < j1) {
                    V3 S{0.0, dx[i] * dz[k], 0.0}, cf{xc[i], ys[j + 1], zc[k]};
                    m->push_face(c, lid(i, j + 1, k), S, cf, linear_weight(S, cf, cP, cc(i, j + 1, k)), fy(i, j, k));
                }
                if (k + 1 < k1) {
                    V3 S{0.0, 0.0, dx[i] * dy[j]}, cf{xc[i], yc[j], zs[k + 1]};
                    m->push_face(c, lid(i, j, k + 1), S, cf, linear_weight(S, cf, cP, cc(i, j, k + 1)), fz(i, j, k));
                }
            }
    // global boundary-face numbering: patches in order, faces in ascending global cell order
    const int64_t gsize[6] = {(int64_t)ny * nz, (int64_t)ny * nz, (int64_t)nx * nz, (int64_t)nx * nz, (int64_t)nx * ny, (int64_t)nx * ny};
    int64_t gstart[6];
    {
        int64_t pos = 0;
        for (int p = 0; p < 6; ++p) { gstart[p] = pos; if (pt[p] != TCIF_PATCH_EMPTY) pos += gsize[p]; }
    }
    // original patches (possibly empty on this rank)
    for (int p = 0; p < 6; ++p) {
        const int d = p / 2, hi = p % 2;
        m->patchType.push_back(pt[p]); m->patchStart.push_back((int32_t)m->faceCells.size()); m->patchNbr.push_back(-1);
        m->bndAddr.push_back(p);
        bool touches = (d == 0 && (hi ? i1 == nx : i0 == 0)) || (d == 1 && (hi ? j1 == ny : j0 == 0)) || (d == 2 && (hi ? k1 == nz : k0 == 0));
        if (pt[p] == TCIF_PATCH_EMPTY || !touches) { m->patchSize.push_back(0); continue; }
        int cnt = 0;
        for (int k = k0; k < k1; ++k)
            for (int j = j0; j < j1; ++j)
                for (int i = i0; i < i1; ++i) {
                    bool on = d == 0 ? (i == (hi ? nx - 1 : 0)) : d == 1 ? (j == (hi ? ny - 1 : 0)) : (k == (hi ? nz - 1 : 0));
                    if (!on) continue;
                    double a = d == 0 ? dy[j] * dz[k] : d == 1 ? dx[i] * dz[k] : dx[i] * dy[j];
                    V3 S{0.0, 0.0, 0.0}, cf{xc[i], yc[j], zc[k]};
                    double sa = hi ? a : -a;
                    if (d == 0) { S.x = sa; cf.x = hi ? xs[nx] : xs[0]; }
                    else if (d == 1) { S.y = sa; cf.y = hi ? ys[ny] : ys[0]; }
                    else { S.z = sa; cf.z = hi ? zs[nz] : zs[0]; }
                    V3 dl = cf - cc(i, j, k);
                    int64_t rankInPatch = d == 0 ? (int64_t)j + (int64_t)ny * k : d == 1 ? (int64_t)i + (int64_t)nx * k : (int64_t)i + (int64_t)nx * j;
                    m->push_bface(lid(i, j, k), S, cf, 1.0 / mag(dl), 1.0, V3{0, 0, 0}, -1, gstart[p] + rankInPatch + 1);
                    ++cnt;
                }
        m->patchSize.push_back(cnt);
    }
    // processor patches, ascending neighbour rank: -z, -y, -x, +x, +y, +z
    struct Nb { int d, hi; };
    const Nb order[6] = {{2, 0}, {1, 0}, {0, 0}, {0, 1}, {1, 1}, {2, 1}};
    for (const Nb &nb : order) {
        int nbi = bi + (nb.d == 0 ? (nb.hi ? 1 : -1) : 0), nbj = bj + (nb.d == 1 ? (nb.hi ? 1 : -1) : 0), nbk = bk + (nb.d == 2 ? (nb.hi ? 1 : -1) : 0);
        if (nbi < 0 || nbi >= px || nbj < 0 || nbj >= py || nbk < 0 || nbk >= pz) continue;
        int e = nbi + px * (nbj + py * nbk);
        m->patchType.push_back(TCIF_PATCH_PROCESSOR); m->patchStart.push_back((int32_t)m->faceCells.size());
        m->patchNbr.push_back(e); m->bndAddr.push_back(-1);
        int cnt = 0;
        for (int k = k0; k < k1; ++k)
            for (int j = j0; j < j1; ++j)
                for (int i = i0; i < i1; ++i) {
                    bool on = nb.d == 0 ? (i == (nb.hi ? i1 - 1 : i0)) : nb.d == 1 ? (j == (nb.hi ? j1 - 1 : j0)) : (k == (nb.hi ? k1 - 1 : k0));
                    if (!on) continue;
                    int ni = i + (nb.d == 0 ? (nb.hi ? 1 : -1) : 0), nj = j + (nb.d == 1 ? (nb.hi ? 1 : -1) : 0), nk = k + (nb.d == 2 ? (nb.hi ? 1 : -1) : 0);
                    // the global face is owned by the lower cell
                    int oi = nb.hi ? i : ni, oj = nb.hi ? j : nj, ok = nb.hi ? k : nk;
                    double a = nb.d == 0 ? dy[j] * dz[k] : nb.d == 1 ? dx[i] * dz[k] : dx[i] * dy[j];
                    V3 Sg{0.0, 0.0, 0.0}, cf{xc[i], yc[j], zc[k]};
                    int64_t gf;
                    if (nb.d == 0) { Sg.x = a; cf.x = xs[oi + 1]; gf = fx(oi, oj, ok); }
                    else if (nb.d == 1) { Sg.y = a; cf.y = ys[oj + 1]; gf = fy(oi, oj, ok); }
                    else { Sg.z = a; cf.z = zs[ok + 1]; gf = fz(oi, oj, ok); }
                    bool flip = !nb.hi;                          // local cell is the global neighbour
                    V3 S = flip ? neg(Sg) : Sg;
                    V3 cl = cc(i, j, k), cn = cc(ni, nj, nk);
                    V3 dl = cn - cl;
                    m->push_bface(lid(i, j, k), S, cf, 1.0 / mag(dl), processor_weight(S, cf, cl, cn), cn, gid(ni, nj, nk),
                                  flip ? -(gf + 1) : (gf + 1));
                    ++cnt;
                }
        m->patchSize.push_back(cnt);
    }
    return m;
}

tcif_mesh *decompose(const tcif_mesh *g, const int32_t *cellProc, int nProcs, int r) {
    if (!g || !cellProc || nProcs < 1 || r < 0 || r >= nProcs) failf("decompose: bad arguments");
    const int nC = g->nCells, nIF = (int)g->owner.size(), nP = (int)g->patchType.size();
    for (int c = 0; c < nC; ++c) if (cellProc[c] < 0 || cellProc[c] >= nProcs) failf("decompose: cellProc[%d] = %d out of range", c, cellProc[c]);
    for (int p = 0; p < nP; ++p) if (g->patchType[p] == TCIF_PATCH_PROCESSOR) failf("decompose: the global mesh already has processor patches");
    tcif_mesh *m = new tcif_mesh();
    std::vector<int32_t> g2l(nC, -1);
    for (int c = 0; c < nC; ++c)
        if (cellProc[c] == r) {
            g2l[c] = m->nCells++;
            m->V.push_back(g->V[c]);
            m->C.insert(m->C.end(), {g->C[3 * (size_t)c], g->C[3 * (size_t)c + 1], g->C[3 * (size_t)c + 2]});
            m->cellGlobal.push_back(c);
        }
    if (m->nCells == 0) { delete m; failf("decompose: rank %d owns no cells", r); }
    auto S3 = [&](const std::vector<double> &v, size_t f) { return V3{v[3 * f], v[3 * f + 1], v[3 * f + 2]}; };
    for (int f = 0; f < nIF; ++f) {
        int o = g->owner[f], n = g->neighbour[f];
        if (cellProc[o] == r && cellProc[n] == r) m->push_face(g2l[o], g2l[n], S3(g->Sf, f), S3(g->Cf, f), g->w[f], f);
    }
    for (int p = 0; p < nP; ++p) {
        m->patchType.push_back(g->patchType[p]); m->patchStart.push_back((int32_t)m->faceCells.size());
        m->patchNbr.push_back(-1); m->bndAddr.push_back(p);
        int cnt = 0;
        for (int i = 0; i < g->patchSize[p]; ++i) {
            size_t b = (size_t)g->patchStart[p] + i;
            int fc = g->faceCells[b];
            if (cellProc[fc] != r) continue;
            m->push_bface(g2l[fc], S3(g->Sfb, b), S3(g->Cfb, b), g->dcb[b], g->wb[b], V3{0, 0, 0}, -1, (int64_t)b + 1);
            ++cnt;
        }
        m->patchSize.push_back(cnt);
    }
    // processor patches: neighbour ranks ascending, faces in global face order
    std::vector<std::vector<int>> byNbr(nProcs);
    for (int f = 0; f < nIF; ++f) {
        int po = cellProc[g->owner[f]], pn = cellProc[g->neighbour[f]];
        if (po == pn) continue;
        if (po == r) byNbr[pn].push_back(f);
        else if (pn == r) byNbr[po].push_back(f);
    }
    for (int e = 0; e < nProcs; ++e) {
        if (e == r || byNbr[e].empty()) continue;
        m->patchType.push_back(TCIF_PATCH_PROCESSOR); m->patchStart.push_back((int32_t)m->faceCells.size());
        m->patchNbr.push_back(e); m->bndAddr.push_back(-1);
        for (int f : byNbr[e]) {
            bool flip = cellProc[g->neighbour[f]] == r;
            int lc = flip ? g->neighbour[f] : g->owner[f], nc = flip ? g->owner[f] : g->neighbour[f];
            V3 Sg = S3(g->Sf, f), cf = S3(g->Cf, f);
            V3 S = flip ? neg(Sg) : Sg;
            V3 cl = g->centre(lc), cn = g->centre(nc);
            m->push_bface(g2l[lc], S, cf, 1.0 / mag(cn - cl), processor_weight(S, cf, cl, cn), cn, nc,
                          flip ? -((int64_t)f + 1) : ((int64_t)f + 1));
        }
        m->patchSize.push_back((int32_t)byNbr[e].size());
    }
    return m;
}

}  // namespace

extern "C" {

const char *tcif_mesh_last_error(void) { return g_err.c_str(); }

int tcif_mesh_box_block(int32_t nx, int32_t ny, int32_t nz, const double *xs, const double *ys, const double *zs,
                        const int32_t patchTypes[6], int32_t px, int32_t py, int32_t pz, int32_t rank, tcif_mesh **out) {
    try {
        if (!xs || !ys || !zs || !patchTypes || !out) failf("tcif_mesh_box_block: null argument");
        *out = box_block(nx, ny, nz, xs, ys, zs, patchTypes, px, py, pz, rank);
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return 1; }
}

int tcif_mesh_decompose(const tcif_mesh *global, const int32_t *cellProc, int32_t nProcs, int32_t rank, tcif_mesh **out) {
    try {
        if (!out) failf("tcif_mesh_decompose: null argument");
        *out = decompose(global, cellProc, nProcs, rank);
        return 0;
    } catch (const std::exception &e) { g_err = e.what(); return 1; }
}

int tcif_mesh_simple_map(int32_t nx, int32_t ny, int32_t nz, int32_t px, int32_t py, int32_t pz, int32_t *cellProc) {
    if (!cellProc || nx < 1 || ny < 1 || nz < 1 || px < 1 || py < 1 || pz < 1) { g_err = "tcif_mesh_simple_map: bad arguments"; return 1; }
    for (int k = 0; k < nz; ++k)
        for (int j = 0; j < ny; ++j)
            for (int i = 0; i < nx; ++i)
                cellProc[(size_t)i + (size_t)nx * (j + (size_t)ny * k)] =
                    (int32_t)(((long long)i * px) / nx + px * (((long long)j * py) / ny + py * (((long long)k * pz) / nz)));
    return 0;
}

void tcif_mesh_free(tcif_mesh *m) { delete m; }

int tcif_mesh_get_desc(const tcif_mesh *m, tcif_mesh_desc *d) {
    if (!m || !d) { g_err = "tcif_mesh_get_desc: null argument"; return 1; }
    d->nCells = m->nCells; d->nInternalFaces = (int32_t)m->owner.size();
    d->nPatches = (int32_t)m->patchType.size(); d->nBoundaryFaces = (int32_t)m->faceCells.size();
    d->owner = m->owner.data(); d->neighbour = m->neighbour.data(); d->Sf = m->Sf.data(); d->weights = m->w.data();
    d->V = m->V.data(); d->patchType = m->patchType.data(); d->patchStart = m->patchStart.data();
    d->patchSize = m->patchSize.data(); d->patchNeighbProcNo = m->patchNbr.data(); d->faceCells = m->faceCells.data();
    d->SfBoundary = m->Sfb.data(); d->magSfBoundary = m->magSfb.data(); d->deltaCoeffsBoundary = m->dcb.data();
    d->weightsBoundary = m->wb.data();
    return 0;
}

int tcif_mesh_array_f64(const tcif_mesh *m, const char *name, const double **ptr, int64_t *count) {
    if (!m || !name || !ptr || !count) return 1;
    const std::vector<double> *v = nullptr;
    if (!strcmp(name, "C")) v = &m->C; else if (!strcmp(name, "Cf")) v = &m->Cf; else if (!strcmp(name, "Cfb")) v = &m->Cfb;
    else if (!strcmp(name, "nbrC")) v = &m->nbrC;
    if (!v) { g_err = std::string("unknown f64 array ") + name; return 1; }
    *ptr = v->data(); *count = (int64_t)v->size();
    return 0;
}

int tcif_mesh_array_i64(const tcif_mesh *m, const char *name, const int64_t **ptr, int64_t *count) {
    if (!m || !name || !ptr || !count) return 1;
    const std::vector<int64_t> *v = nullptr;
    if (!strcmp(name, "cellGlobal")) v = &m->cellGlobal; else if (!strcmp(name, "nbrGlobal")) v = &m->nbrGlobal;
    else if (!strcmp(name, "faceProcAddressing")) v = &m->faceAddr; else if (!strcmp(name, "bfaceProcAddressing")) v = &m->bfaceAddr;
    else if (!strcmp(name, "boundaryProcAddressing")) v = &m->bndAddr;
    if (!v) { g_err = std::string("unknown i64 array ") + name; return 1; }
    *ptr = v->data(); *count = (int64_t)v->size();
    return 0;
}

}  // extern "C"
