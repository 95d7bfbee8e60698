// mesh_host.h -- private to libtcif_host.so: the mesh object behind include/tcif_mesh.h and the small vector helpers
// shared by mesh_host.cpp (block generator, decomposition) and polymesh_host.cpp (polyMesh reader).
#pragma once
#include "../../include/tcif_mesh.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

namespace tcif_host {

const double VSMALL = 1e-300;

[[noreturn]] inline void failf(const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    throw std::runtime_error(buf);
}
std::string &last_error();          // thread-local message behind tcif_mesh_last_error()

struct V3 { double x, y, z; };
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline double mag(V3 a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }

// surfaceInterpolation::makeWeights (linear)
inline double linear_weight(V3 S, V3 Cf, V3 CP, V3 CN) {
    double a = std::fabs(dot(S, CN - Cf)), b = std::fabs(dot(S, Cf - CP));
    return a / (b + a);
}
// processorFvPatch::makeWeights: unit-normal projections of both sides
inline double processor_weight(V3 S, V3 Cf, V3 Clocal, V3 Cnbr) {
    double m = mag(S);
    V3 nf{S.x / m, S.y / m, S.z / m};
    double dP = dot(nf, Cf - Clocal);
    V3 SN = neg(S);
    double mn = m + VSMALL;
    V3 nN{SN.x / mn, SN.y / mn, SN.z / mn};
    double dN = dot(nN, Cf - Cnbr);
    return dN / (dP + dN);
}

}  // namespace tcif_host

struct tcif_mesh {
    int32_t nCells = 0;
    std::vector<int32_t> owner, neighbour;
    std::vector<double> Sf, w, V, C, Cf;
    std::vector<int32_t> patchType, patchStart, patchSize, patchNbr;
    std::vector<int32_t> faceCells;
    std::vector<double> Sfb, magSfb, dcb, wb, Cfb, nbrC;
    std::vector<int64_t> cellGlobal, nbrGlobal, faceAddr, bfaceAddr, bndAddr;

    void push_face(int32_t o, int32_t n, tcif_host::V3 S, tcif_host::V3 cf, double wt, int64_t gface) {
        owner.push_back(o); neighbour.push_back(n);
        Sf.insert(Sf.end(), {S.x, S.y, S.z}); Cf.insert(Cf.end(), {cf.x, cf.y, cf.z});
        w.push_back(wt); faceAddr.push_back(gface + 1);
    }
    void push_bface(int32_t fc, tcif_host::V3 S, tcif_host::V3 cf, double dc, double wt, tcif_host::V3 nc, int64_t ng, int64_t addr) {
        faceCells.push_back(fc);
        Sfb.insert(Sfb.end(), {S.x, S.y, S.z}); Cfb.insert(Cfb.end(), {cf.x, cf.y, cf.z});
        magSfb.push_back(tcif_host::mag(S)); dcb.push_back(dc); wb.push_back(wt);
        nbrC.insert(nbrC.end(), {nc.x, nc.y, nc.z}); nbrGlobal.push_back(ng); bfaceAddr.push_back(addr);
    }
    tcif_host::V3 centre(int32_t c) const { return {C[3 * (size_t)c], C[3 * (size_t)c + 1], C[3 * (size_t)c + 2]}; }
};
