// polymesh_host.cpp -- reader of an OpenFOAM ASCII polyMesh directory (constant/polyMesh/{points,faces,owner,
// neighbour,boundary}) and of labelList files (a scotch `cellDecomposition` as decomposePar -cellDist writes it) into
// the host mesh model of include/tcif_mesh.h.  This is how a case that was meshed and decomposed elsewhere reaches
// the device path without OpenFOAM being installed: tcif_mesh_read_polymesh + tcif_mesh_read_label_list +
// tcif_mesh_decompose give every rank its processor sub-mesh in decomposePar conventions.
//
// Geometry follows the [UPSTREAM] OpenFOAM-v2106 formulas (the source is not in /root/reference; restated from
// src/OpenFOAM/meshes/primitiveMesh/primitiveMeshFaceCentresAndAreas.C, primitiveMeshCellCentresAndVols.C and
// src/finiteVolume/interpolation/surfaceInterpolation/surfaceInterpolation/surfaceInterpolation.C::makeWeights):
//   face   : triangles directly; otherwise a fan about the vertex average fC: n_k = (p_k+1 - p_k) x (fC - p_k),
//            Sf = 1/2 sum n_k, Cf = 1/3 sum(|n_k| (p_k + p_k+1 + fC)) / sum |n_k|
//   cell   : cEst = mean of its face centres; pyramids about cEst: 3V_pyr = Sf.(Cf - cEst) (sign by ownership),
//            C = sum(3V_pyr (3/4 Cf + 1/4 cEst)) / sum 3V_pyr, V = 1/3 sum 3V_pyr; faces in ascending label, the
//            owner loop before the neighbour loop
//   weights: w = |Sf.(C_N - Cf)| / (|Sf.(Cf - C_P)| + |Sf.(C_N - Cf)|); boundary w = 1, deltaCoeffs = 1/|Cf - C|
// Checked bit-exactly against the independent numpy restatement oracle/polymesh_oracle.py (tests/test_polymesh.py).
#include "mesh_host.h"

#include <cctype>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

using namespace tcif_host;

namespace {

// file content with the FoamFile header and C/C++ comments removed
std::string foam_body(const std::string &path) {
    std::ifstream in(path);
    if (!in) failf("cannot open %s", path.c_str());
    std::stringstream ss;
    ss << in.rdbuf();
    std::string t = ss.str(), out;
    out.reserve(t.size());
    for (size_t i = 0; i < t.size();) {
        if (t.compare(i, 2, "/*") == 0) { size_t e = t.find("*/", i + 2); i = e == std::string::npos ? t.size() : e + 2; }
        else if (t.compare(i, 2, "//") == 0) { size_t e = t.find('\n', i); i = e == std::string::npos ? t.size() : e; }
        else out.push_back(t[i++]);
    }
    size_t h = out.find("FoamFile");
    if (h != std::string::npos) {
        size_t e = out.find('}', h);
        if (e == std::string::npos) failf("%s: unterminated FoamFile header", path.c_str());
        out.erase(0, e + 1);
    }
    return out;
}

struct Cursor {
    const std::string &s; size_t i; const std::string &path;
    void ws() { while (i < s.size() && std::isspace((unsigned char)s[i])) ++i; }
    bool eof() { ws(); return i >= s.size(); }
    char peek() { ws(); return i < s.size() ? s[i] : '\0'; }
    void expect(char c) { ws(); if (i >= s.size() || s[i] != c) failf("%s: expected '%c' at offset %zu", path.c_str(), c, i); ++i; }
    long long integer() {
        ws();
        char *e = nullptr;
        long long v = std::strtoll(s.c_str() + i, &e, 10);
        if (e == s.c_str() + i) failf("%s: expected an integer at offset %zu", path.c_str(), i);
        i = (size_t)(e - s.c_str());
        return v;
    }
    double real() {
        ws();
        char *e = nullptr;
        double v = std::strtod(s.c_str() + i, &e);
        if (e == s.c_str() + i) failf("%s: expected a number at offset %zu", path.c_str(), i);
        i = (size_t)(e - s.c_str());
        return v;
    }
    std::string word() {
        ws();
        size_t b = i;
        while (i < s.size() && !std::isspace((unsigned char)s[i]) && s[i] != ';' && s[i] != '{' && s[i] != '}' && s[i] != '(' && s[i] != ')') ++i;
        if (b == i) failf("%s: expected a word at offset %zu", path.c_str(), i);
        return s.substr(b, i - b);
    }
};

// labelList: `N ( l0 l1 ... )` or the uniform form `N{l}`
std::vector<int32_t> read_labels(const std::string &path) {
    const std::string body = foam_body(path);
    Cursor c{body, 0, path};
    const long long n = c.integer();
    if (n < 0 || n > 0x7fffffffLL) failf("%s: bad list size %lld", path.c_str(), n);
    std::vector<int32_t> v((size_t)n);
    if (c.peek() == '{') { c.expect('{'); const long long x = c.integer(); c.expect('}'); std::fill(v.begin(), v.end(), (int32_t)x); return v; }
    c.expect('(');
    for (long long k = 0; k < n; ++k) v[(size_t)k] = (int32_t)c.integer();
    c.expect(')');
    return v;
}

V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 scale(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }

int patch_type(const std::string &t, const std::string &path) {
    if (t == "patch" || t == "wall" || t == "symmetryPlane" || t == "symmetry") return TCIF_PATCH_GENERIC;
    if (t == "empty") return TCIF_PATCH_EMPTY;
    if (t == "wedge") return TCIF_PATCH_WEDGE;
    if (t == "processor") return TCIF_PATCH_PROCESSOR;
    failf("%s: patch type %s is not supported by the device path (patch | wall | symmetry(Plane) | empty | wedge | processor)", path.c_str(), t.c_str());
}

tcif_mesh *read_polymesh(const std::string &dir) {
    // points
    std::vector<V3> pts;
    {
        const std::string path = dir + "/points", body = foam_body(path);
        Cursor c{body, 0, path};
        const long long n = c.integer();
        if (n < 4) failf("%s: %lld points", path.c_str(), n);
        pts.resize((size_t)n);
        c.expect('(');
        for (auto &p : pts) { c.expect('('); p.x = c.real(); p.y = c.real(); p.z = c.real(); c.expect(')'); }
        c.expect(')');
    }
    // faces (compact: vertex labels + offsets)
    std::vector<int32_t> fStart(1, 0), fPts;
    {
        const std::string path = dir + "/faces", body = foam_body(path);
        Cursor c{body, 0, path};
        const long long n = c.integer();
        c.expect('(');
        for (long long f = 0; f < n; ++f) {
            const long long nv = c.integer();
            if (nv < 3) failf("%s: face %lld has %lld vertices", path.c_str(), f, nv);
            c.expect('(');
            for (long long k = 0; k < nv; ++k) {
                const long long v = c.integer();
                if (v < 0 || v >= (long long)pts.size()) failf("%s: face %lld: vertex label %lld out of range", path.c_str(), f, v);
                fPts.push_back((int32_t)v);
            }
            c.expect(')');
            fStart.push_back((int32_t)fPts.size());
        }
        c.expect(')');
    }
    const size_t nF = fStart.size() - 1;
    std::vector<int32_t> owner = read_labels(dir + "/owner"), neighbour = read_labels(dir + "/neighbour");
    if (owner.size() != nF) failf("%s/owner: %zu labels for %zu faces", dir.c_str(), owner.size(), nF);
    const size_t nIF = neighbour.size();
    if (nIF > nF) failf("%s/neighbour: more labels than faces", dir.c_str());
    int32_t nCells = 0;
    for (size_t f = 0; f < nF; ++f) { if (owner[f] < 0) failf("owner[%zu] < 0", f); nCells = std::max(nCells, owner[f] + 1); }
    for (size_t f = 0; f < nIF; ++f) {
        if (neighbour[f] <= owner[f]) failf("face %zu: neighbour %d <= owner %d (internal faces must be upper-triangular)", f, neighbour[f], owner[f]);
        nCells = std::max(nCells, neighbour[f] + 1);
    }
    // face centres and area vectors
    std::vector<V3> Sf(nF), Cf(nF);
    for (size_t f = 0; f < nF; ++f) {
        const int32_t *v = &fPts[fStart[f]];
        const int n = fStart[f + 1] - fStart[f];
        if (n == 3) {
            const V3 p0 = pts[v[0]], p1 = pts[v[1]], p2 = pts[v[2]];
            Cf[f] = scale(1.0 / 3.0, add(add(p0, p1), p2));
            Sf[f] = scale(0.5, cross(p1 - p0, p2 - p0));
            continue;
        }
        V3 fC = pts[v[0]];
        for (int k = 1; k < n; ++k) fC = add(fC, pts[v[k]]);
        fC = {fC.x / n, fC.y / n, fC.z / n};
        V3 sumN{0, 0, 0}, sumAc{0, 0, 0};
        double sumA = 0.0;
        for (int k = 0; k < n; ++k) {
            const V3 p = pts[v[k]], pn = pts[v[(k + 1) % n]];
            const V3 c = add(add(p, pn), fC);
            const V3 nn = cross(pn - p, fC - p);
            const double a = mag(nn);
            sumN = add(sumN, nn); sumA = sumA + a; sumAc = add(sumAc, scale(a, c));
        }
        if (sumA > VSMALL) { const V3 t = scale(1.0 / 3.0, sumAc); Cf[f] = {t.x / sumA, t.y / sumA, t.z / sumA}; }
        else Cf[f] = fC;
        Sf[f] = scale(0.5, sumN);
    }
    // cell centres and volumes
    std::vector<V3> cEst((size_t)nCells, V3{0, 0, 0}), C((size_t)nCells, V3{0, 0, 0});
    std::vector<double> nCellFaces((size_t)nCells, 0.0), V((size_t)nCells, 0.0);
    for (size_t f = 0; f < nF; ++f) { cEst[owner[f]] = add(cEst[owner[f]], Cf[f]); nCellFaces[owner[f]] += 1.0; }
    for (size_t f = 0; f < nIF; ++f) { cEst[neighbour[f]] = add(cEst[neighbour[f]], Cf[f]); nCellFaces[neighbour[f]] += 1.0; }
    for (int32_t c = 0; c < nCells; ++c) {
        if (nCellFaces[c] < 4) failf("cell %d has %d faces", c, (int)nCellFaces[c]);
        cEst[c] = {cEst[c].x / nCellFaces[c], cEst[c].y / nCellFaces[c], cEst[c].z / nCellFaces[c]};
    }
    for (size_t f = 0; f < nF; ++f) {
        const int32_t o = owner[f];
        const double pyr = dot(Sf[f], Cf[f] - cEst[o]);
        const V3 pc = add(scale(0.75, Cf[f]), scale(0.25, cEst[o]));
        C[o] = add(C[o], scale(pyr, pc)); V[o] += pyr;
    }
    for (size_t f = 0; f < nIF; ++f) {
        const int32_t n = neighbour[f];
        const double pyr = dot(Sf[f], cEst[n] - Cf[f]);
        const V3 pc = add(scale(0.75, Cf[f]), scale(0.25, cEst[n]));
        C[n] = add(C[n], scale(pyr, pc)); V[n] += pyr;
    }
    tcif_mesh *m = new tcif_mesh();
    try {
        m->nCells = nCells;
        m->C.resize(3 * (size_t)nCells); m->V.resize((size_t)nCells);
        for (int32_t c = 0; c < nCells; ++c) {
            if (!(V[c] > 0)) failf("cell %d has non-positive volume", c);
            m->C[3 * (size_t)c] = C[c].x / V[c]; m->C[3 * (size_t)c + 1] = C[c].y / V[c]; m->C[3 * (size_t)c + 2] = C[c].z / V[c];
            m->V[c] = V[c] * (1.0 / 3.0);
            m->cellGlobal.push_back(c);
        }
        for (size_t f = 0; f < nIF; ++f)
            m->push_face(owner[f], neighbour[f], Sf[f], Cf[f], linear_weight(Sf[f], Cf[f], m->centre(owner[f]), m->centre(neighbour[f])), (int64_t)f);
        // boundary: `N ( name { type T; nFaces n; startFace s; [neighbProcNo e;] ... } ... )`
        const std::string path = dir + "/boundary", body = foam_body(path);
        Cursor c{body, 0, path};
        const long long nP = c.integer();
        c.expect('(');
        size_t covered = nIF;
        int64_t bfGlobal = 0;
        for (long long p = 0; p < nP; ++p) {
            const std::string name = c.word();
            c.expect('{');
            std::string type;
            long long nFaces = -1, startFace = -1, nbr = -1;
            while (c.peek() != '}') {
                const std::string key = c.word();
                if (key == "type") type = c.word();
                else if (key == "nFaces") nFaces = c.integer();
                else if (key == "startFace") startFace = c.integer();
                else if (key == "neighbProcNo") nbr = c.integer();
                else {                                   // any other entry (inGroups, physicalType, myProcNo, ...): skip to ';'
                    size_t e = body.find(';', c.i);
                    if (e == std::string::npos) failf("%s: patch %s: unterminated entry %s", path.c_str(), name.c_str(), key.c_str());
                    c.i = e;
                }
                c.expect(';');
            }
            c.expect('}');
            if (type.empty() || nFaces < 0 || startFace < 0) failf("%s: patch %s lacks type/nFaces/startFace", path.c_str(), name.c_str());
            if ((size_t)startFace != covered || (size_t)(startFace + nFaces) > nF) failf("%s: patch %s does not continue the face list", path.c_str(), name.c_str());
            covered += (size_t)nFaces;
            const int t = patch_type(type, path);
            m->patchType.push_back(t); m->patchStart.push_back((int32_t)m->faceCells.size()); m->patchNbr.push_back((int32_t)nbr);
            m->bndAddr.push_back(p);
            if (t == TCIF_PATCH_EMPTY) { m->patchSize.push_back(0); continue; }      // fields carry no values on it
            m->patchSize.push_back((int32_t)nFaces);
            for (long long i = 0; i < nFaces; ++i) {
                const size_t f = (size_t)(startFace + i);
                const V3 cc = m->centre(owner[f]);
                const V3 d = Cf[f] - cc;
                m->push_bface(owner[f], Sf[f], Cf[f], 1.0 / mag(d), 1.0, V3{0, 0, 0}, -1, ++bfGlobal);
            }
        }
        c.expect(')');
        if (covered != nF) failf("%s: patches cover %zu of %zu faces", path.c_str(), covered, nF);
    } catch (...) { delete m; throw; }
    return m;
}

}  // namespace

extern "C" {

int tcif_mesh_read_polymesh(const char *polyMeshDir, tcif_mesh **out) {
    if (out) *out = nullptr;
    try {
        if (!polyMeshDir || !out) failf("tcif_mesh_read_polymesh: null argument");
        *out = read_polymesh(polyMeshDir);
        return 0;
    } catch (const std::exception &e) { last_error() = e.what(); return 1; }
}

int tcif_mesh_read_label_list(const char *path, int32_t capacity, int32_t *labels, int32_t *count) {
    if (count) *count = 0;
    try {
        if (!path || !count || capacity < 0 || (capacity && !labels)) failf("tcif_mesh_read_label_list: bad arguments");
        const std::vector<int32_t> v = read_labels(path);
        *count = (int32_t)v.size();
        if ((size_t)capacity < v.size()) failf("%s holds %zu labels, capacity is %d", path, v.size(), capacity);
        if (!v.empty()) std::memcpy(labels, v.data(), sizeof(int32_t) * v.size());
        return 0;
    } catch (const std::exception &e) { last_error() = e.what(); return 1; }
}

}  // extern "C"
