// tcif.cu -- C ABI (include/tcif.h) of the B200-native interface hot path: device-resident
// mesh + fields, kernel orchestration, processor-patch exchange, timers.
//
// Mirrors (does not copy) the reference call surface:
//   phaseSystem::divCapillaryStress()      /root/reference/phasesSystem/phaseSystem/phaseSystem.C:1038-1105
//   multiphaseSystem::solveAlphas() MULES  /root/reference/phasesSystem/multiphaseSystem/multiphaseSystem.C:296-375
// There is no CPU fallback: every entry point needs a CUDA device.
#include "tcif_device.cuh"
#include "tcif_sweeps.h"
#include "tcif_alpha.cuh"
#include "tcif_mass.cuh"
#include "tcif_species.cuh"
#include "tcif_laser.cuh"
#include "../../include/tcif.h"

#include <dlfcn.h>
#include <unistd.h>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <type_traits>
#include <unordered_map>
#include <vector>
#include <map>

using namespace tcif;

namespace {

thread_local std::string g_createError;

[[noreturn]] void failf(const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    throw std::runtime_error(buf);
}

#define CUDA_OK(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) failf("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

// ---- timers ---------------------------------------------------------------------------
enum TimerId {
    T_GRAD_TENSOR = 0, T_DIV_TENSOR, T_MULES_SETUP, T_MULES_ITER, T_MULES_FINAL, T_LIMIT_SUM,
    T_BOUNDARY, T_HALO, T_CONVERT, T_STATS, T_FLUX_ASSEMBLY, T_PREDICTOR, T_RHOPHI, T_MASS, T_LASER, T_COUNT
};
const char *kTimerNames[T_COUNT] = {
    "grad_tensor", "div_tensor", "mules_setup", "mules_iter", "mules_final", "limit_sum",
    "boundary", "halo", "convert", "stats", "flux_assembly", "predictor", "rho_phi", "alpha_transfer", "laser_dtrm"};

struct EvPair { cudaEvent_t a, b; int id; double bytes; };

// ---- NCCL, bound at run time (torch ships libnccl.so.2; the system has one too) --------
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
struct NcclApi {
    void *h = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
const int kNcclDouble = 8;   // ncclFloat64
NcclApi g_nccl;
std::mutex g_ncclMutex;

void nccl_load() {
    std::lock_guard<std::mutex> lk(g_ncclMutex);
    if (g_nccl.h) return;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
    if (!h) failf("cannot load libnccl.so.2: %s", dlerror());
#define NSYM(field, sym)                                                   \
    *(void **)(&g_nccl.field) = dlsym(h, sym);                             \
    if (!g_nccl.field) failf("libnccl: missing symbol %s", sym);
    NSYM(GetUniqueId, "ncclGetUniqueId") NSYM(CommInitRank, "ncclCommInitRank") NSYM(CommDestroy, "ncclCommDestroy")
    NSYM(Send, "ncclSend") NSYM(Recv, "ncclRecv") NSYM(AllGather, "ncclAllGather")
    NSYM(GroupStart, "ncclGroupStart") NSYM(GroupEnd, "ncclGroupEnd") NSYM(GetErrorString, "ncclGetErrorString")
#undef NSYM
    g_nccl.h = h;
}
#define NCCL_OK(call)                                                                        \
    do {                                                                                     \
        int r_ = (call);                                                                     \
        if (r_ != 0) failf("%s failed: %s", #call, g_nccl.GetErrorString(r_));               \
    } while (0)

}  // namespace

// ---- in-process group: several contexts on one GPU, one host thread each ----------------
struct tcif_group {
    int n = 0;
    std::mutex mu;
    std::condition_variable cv;
    int waiting = 0;
    long generation = 0;
    std::vector<tcif_ctx *> members;
    std::vector<double> red;     // n * 16 doubles
    std::vector<const double *> gatherPtr;   // n pointers (allgather_host)
    // set by a member that found an error inside a collective section; every member checks it after the closing
    // barrier of that section and fails too, so that nobody is left waiting in the next barrier
    std::atomic<int> failed{0};
    void barrier() {
        std::unique_lock<std::mutex> lk(mu);
        long gen = generation;
        if (++waiting == n) { waiting = 0; ++generation; cv.notify_all(); }
        else cv.wait(lk, [&] { return generation != gen; });
    }
};

struct tcif_ctx {
    std::string err;
    int device = 0;
    cudaStream_t ownStream = nullptr, stream = nullptr;
    // O(surface) work (halo exchange + boundary-cell kernels) runs on `side`, concurrently with the
    // O(volume) sweep on `stream`; evV/evB order the two (see sweep()).
    cudaStream_t side = nullptr;
    // host-buffer entry points: uploads on sIn, downloads on sOut, overlapped with compute (PCIe is full duplex)
    cudaStream_t sIn = nullptr, sOut = nullptr;
    cudaEvent_t evIn[2]{}, evC[2]{};
    double *stagingOut = nullptr;
    cudaEvent_t evV = nullptr, evB = nullptr;
    bool noOverlap = false;
    // host copy of the small mesh tables
    int nC = 0, nIF = 0, nBF = 0, nPatches = 0, stride = 0, Ko = 0, Kn = 0, nBndCells = 0, nCoupled = 0;
    size_t nFd = 0;
    std::vector<int> patchType, patchStart, patchSize, patchNbr;
    std::vector<int> cpPatch, cpOff;    // coupled patches: patch id, offset in coupled-face units
    double sumV = 0.0;
    DMesh dm{};
    int tileStrips = 0;      // 1: volume sweeps walk the tiles strip by strip (build_tile_order)
    const int *forceOrder = nullptr;   // tile order of the two force sweeps (TCIF_FORCE_STRIP_CELLS)
    int tsum = 0;            // force kernels carry the pair-summed stress tensor (6 comps) instead of Q tensors
    int *d_faceMap = nullptr;           // host face f -> device slot
    std::vector<void *> allocs;
    int64_t devBytes = 0;
    // config
    tcif_config cfg{};
    int P = 0, Q = 0;
    // fields
    double *alpha[4]{}, *alpha0[4]{}, *alphaNew[4]{}, *alphab[4]{};
    int *bc[4]{};
    double *Tc = nullptr, *Tb = nullptr;
    bool haveT = false;
    SigmaModel hSig[6];
    bool sigSet[6]{};
    SigmaModel *dSig = nullptr;
    double *phi = nullptr, *phib = nullptr;
    double *phiHO[4]{}, *phiHOb[4]{}, *phiPsi[4]{}, *phiPsib[4]{}, *phiCorr[4]{}, *phiCorrb[4]{}, *phiBDb[4]{};
    double *Qp[4]{}, *Qm[4]{}, *sP[4]{}, *sM[4]{}, *lpA[4]{}, *lmA[4]{}, *lpB[4]{}, *lmB[4]{};
    double *lamA[4]{}, *lamB[4]{}, *lamb[4]{}, *lpHalo[4]{}, *lmHalo[4]{};
    double *Su[4]{}, *Sp[4]{};
    bool haveSu[4]{}, haveSp[4]{};
    double *Tq = nullptr, *gBnd = nullptr, *TqHalo = nullptr, *G = nullptr, *Gb = nullptr, *F = nullptr, *Fb = nullptr;
    // rest of solveAlphas (flux assembly, predictor, rhoPhi)
    double *Cx = nullptr, *Cy = nullptr, *Cz = nullptr, *dbx = nullptr, *dby = nullptr, *dbz = nullptr;
    double cAlpha[16];
    double rhoU[4]{};
    double *rho[4]{}, *rhob[4]{};
    bool haveRho[4]{}, rhoSet[4]{};
    double *rhoPhi = nullptr, *rhoPhib = nullptr;
    double *predDiag[4]{}, *predRhs[4]{};
    double *dmaxDev = nullptr;
    int predictorSweeps[4]{};
    // Su/Sp from alphaTransfer
    struct MassModel { int from, to, var, divU; double C, Tact, aMin; double *dmdt; bool haveDmdt; };
    std::vector<MassModel> massModels;
    std::vector<int> massPairs;              // totalPhasePairs order; empty = all (i<j) in phase order
    bool massInSolve = false;                // calculateSuSp() inside every solveAlphas corrector
    long long *ddtMaxDev = nullptr;
    // species MULES (solveYi): phase gradients kept aside + one species' working set swapped into a phase slot
    double *GA = nullptr, *GAb = nullptr;
    double *spY = nullptr, *spY0 = nullptr, *spYNew = nullptr, *spYb = nullptr, *spSu = nullptr, *spSp = nullptr;
    double *spPhiPsi = nullptr, *spPhiPsib = nullptr, *spPhiHO = nullptr, *spPhiHOb = nullptr;
    int *spBc = nullptr;
    double ddtAlphaMax = 0.0;
    // laserDTRM
    LaserGeo lg{};
    bool laserGeoSet = false;
    std::vector<LaserRefl> lzModels;
    double lzAlphaCut = 0.5;
    int lzNAbs = 0, lzAbsPhase[4]{};
    double lzA[4]{}, lzE[4]{}, lzEE[4]{};
    bool lzAbsFields = false;
    double *lzAF = nullptr;                 // a, e, E fields [3*stride] when the host supplies them
    int *lzRefl = nullptr, *lzCounters = nullptr;
    double *lzNHat = nullptr, *lzNHatPts = nullptr, *lzQ = nullptr;
    LaserPart *lzCloud = nullptr;
    int lzCap = 0, lzParticles = 0, lzGenerations = 0;
    bool lzLocalFind = false;             // reflected particles inserted behind a processor patch are dropped (reference behaviour)
    // decomposed runs: point weights kept on the host for the global normalisation, shared points, mailboxes
    std::vector<double> lzPwRaw;          // per point: LOCAL sum of the raw inverse-distance weights (never rescaled)
    std::vector<double> lzPwDenom;        // per point: the sum the device weights are currently normalised with
    std::vector<int> lzPcStart;
    std::vector<unsigned char> lzIsPatch;
    double *lzPw = nullptr;
    unsigned char *lzIsPatchDev = nullptr;
    int lzNShared = 0, lzMaxShared = 0;
    bool lzSharedSet = false;
    int *lzSharedPts = nullptr;
    double *lzSharedBuf = nullptr;
    std::vector<std::vector<int>> lzSharedSrc;      // per shared point: indices rank*maxShared + j of its copies (all ranks)
    LaserPart *lzOutbox = nullptr, *lzInbox = nullptr;
    int lzMailCap = 0;
    int *lzPatchInfo = nullptr;
    double lzDeltaN = 0.0;
    long long lzCreated = 0;
    int *lzSpCells = nullptr;
    double *lzSpVals = nullptr;
    int lzSpCap = 0;
    // F is a pure function of (alpha, T, sigma): the solver evaluates it 2-4x per time step on unchanged
    // fields (UEqn.H:29, pEqn.H:35 per corrector, thermocapillaryInterFoam.C:150); fieldsVersion counts
    // every change of those inputs and forceVersion remembers which one F was computed from
    uint64_t fieldsVersion = 1, forceVersion = 0;
    bool cacheForce = true;
    // residency (tcif_set_alpha_stamped & co.): the caller's modification stamp of the host field each device copy equals
    uint64_t alphaStamp[4]{}, tStamp = 0;
    int64_t xfer[4]{};                  // alpha uploaded / skipped, T uploaded / skipped
    std::vector<int> bcHost[4];         // last alpha BC kinds uploaded (tcif_set_alpha_bc re-uploads only on change)
    // host-buffer entry points (tcif_stage_host / tcif_time_step_host): one-shot stamps of the caller's alpha arrays and
    // delta read-back of alpha and F (tcif_host_alpha_stamps, tcif_host_sparse_returns)
    bool hsArmed = false;
    uint64_t hsIn[4]{}, hsOut[4]{};
    bool sparseReturns = false;
    double *shAlpha[4]{}, *shF = nullptr;      // device record of what the caller's arrays hold
    double *shFace[5]{};                       // same for alphaPhi[p] (host face order)
    const void *shFaceHost[5]{};
    const void *shFHost = nullptr;             // the host F array shF mirrors
    // (two sets of index/value buffers: the host scatters read-back i while the GPU compacts and copies read-back i+1)
    int *dIdx2[2] = {nullptr, nullptr}, *dCount = nullptr;
    double *dVal2[2] = {nullptr, nullptr};
    int dCap = 0, dPar = 0;
    int *hIdx2[2] = {nullptr, nullptr}, *hCount = nullptr;    // pinned
    double *hVal2[2] = {nullptr, nullptr};
    std::thread scatterTh;                     // the one host scatter in flight (joined before the API call returns)
    int64_t d2hBytes = 0, h2dBytes = 0;        // bytes actually moved by the last *_host call
    int64_t forceEvaluations = 0, forceCacheHits = 0;
    double deltaN = 0.0;
    bool deltaNValid = false;
    double *staging = nullptr;
    size_t stagingDoubles = 0;
    double *sendBuf = nullptr, *recvBuf = nullptr;
    size_t haloDoubles = 0;
    double *statsPart = nullptr, *statsOut = nullptr;
    double *h_small = nullptr;          // pinned, 64 doubles
    // comm
    int rank = 0, nRanks = 1;
    tcif_group *group = nullptr;
    ncclComm_t nccl = nullptr;
    double *d_gather = nullptr;         // nRanks*16 doubles (nccl scalar reductions)
    // processor-patch exchange over peer memory (k_halo_push / k_halo_wait / k_halo_pull), NCCL ranks on one node
    bool peerOn = false;
    double *peerBuf = nullptr;          // own receive buffer: 2 parities x haloDoubles, then nRanks flag words
    unsigned *peerFlags = nullptr;      // = (unsigned *)(peerBuf + 2*haloDoubles): one word per source rank
    std::vector<void *> peerMapped;     // per rank: that rank's peerBuf mapped here (null: not a neighbour)
    HaloPeer hp{};
    unsigned exSeq = 0;
    int *peerCounter = nullptr, *peerStatus = nullptr;
    cudaEvent_t evX = nullptr;          // end of the previous exchange (exchanges are totally ordered on a rank)
    double *agBuf = nullptr;            // allgather_host staging
    int agCap = 0;
    // timers
    bool timersOn = false;
    std::vector<EvPair> evs;
    std::vector<EvPair> evPool;
    double tMs[T_COUNT]{}, tBytes[T_COUNT]{};
    int64_t tLaunch[T_COUNT]{};
    int64_t launches = 0;

    // TCIF_GUARD=1 (debug): every device allocation sits between two 512 B guard zones filled with a pattern;
    // tcif_debug_guard_check() counts the zones that no longer hold it (an out-of-bounds WRITE next to an array)
    static constexpr size_t kGuard = 512;
    static constexpr int kGuardByte = 0xA5;
    struct GuardRec { char *base; size_t bytes; };
    std::vector<GuardRec> guards;
    bool guardOn = [] { const char *e = getenv("TCIF_GUARD"); return e && atoi(e) != 0; }();
    template <class T> T *dalloc(size_t n, bool zero = true) {
        void *p = nullptr;
        size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
        if (guardOn) {
            const size_t padded = (bytes + 255) / 256 * 256;
            CUDA_OK(cudaMalloc(&p, padded + 2 * kGuard));
            allocs.push_back(p);
            CUDA_OK(cudaMemsetAsync(p, kGuardByte, kGuard, stream));
            CUDA_OK(cudaMemsetAsync((char *)p + kGuard + bytes, kGuardByte, padded - bytes + kGuard, stream));
            guards.push_back({(char *)p, bytes});
            p = (char *)p + kGuard;
        } else {
            CUDA_OK(cudaMalloc(&p, bytes));
            allocs.push_back(p);
        }
        devBytes += (int64_t)bytes;
        if (zero) CUDA_OK(cudaMemsetAsync(p, 0, bytes, stream));
        return (T *)p;
    }
    template <class T> T *upload(const std::vector<T> &v) {
        T *p = dalloc<T>(v.size(), false);
        if (!v.empty()) CUDA_OK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        return p;
    }
};

namespace {

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <class F> void timed_on(tcif_ctx *c, cudaStream_t st, int id, double bytes, F &&launch) {
    if (c->timersOn) {
        EvPair e;
        if (!c->evPool.empty()) { e = c->evPool.back(); c->evPool.pop_back(); }
        else { CUDA_OK(cudaEventCreate(&e.a)); CUDA_OK(cudaEventCreate(&e.b)); }
        e.id = id; e.bytes = bytes;
        CUDA_OK(cudaEventRecord(e.a, st));
        launch();
        CUDA_OK(cudaEventRecord(e.b, st));
        c->evs.push_back(e);
    } else launch();
    CUDA_OK(cudaGetLastError());
}
template <class F> void timed(tcif_ctx *c, int id, double bytes, F &&launch) { timed_on(c, c->stream, id, bytes, launch); }

// One sweep of the path = the O(volume) kernel over the interior cells, the processor-patch exchange it does NOT need
// and the O(surface) kernel over the deferred (boundary-adjacent) cells, which consumes the halo:
//     main stream :  volume_k ------------------> [wait halo_k] deferred_k   volume_k+1 ...
//     side stream :  halo_k (push / wait / pull)
// The exchange hides behind the volume kernel; the deferred kernel runs AFTER it on the same stream.  (Rounds 1 ran
// the deferred kernel concurrently on the high-priority side stream: the volume sweeps are latency/occupancy-bound and
// lost 10-12 % to the co-runner -- 42.5 ms of volume kernels in isolation under ncu, 48.4 ms in the bench -- and
// handling patch faces inline in the volume kernel cost the same through longer-lived divergent warps; see
// profiles/r03_variants.md.)
template <class FV, class FH, class FB> void sweep(tcif_ctx *c, FV &&volume, FH &&halo, FB &&boundary) {
    const bool ex = c->nCoupled > 0;
    if (ex) {
        CUDA_OK(cudaEventRecord(c->evV, c->stream));
        CUDA_OK(cudaStreamWaitEvent(c->side, c->evV, 0));
        halo();
        CUDA_OK(cudaEventRecord(c->evB, c->side));
    }
    volume();
    if (ex) CUDA_OK(cudaStreamWaitEvent(c->stream, c->evB, 0));
    boundary();
}
template <class FV, class FB> void sweep(tcif_ctx *c, FV &&volume, FB &&boundary) { volume(); boundary(); }
void join(tcif_ctx *c) { CUDA_OK(cudaStreamWaitEvent(c->stream, c->evB, 0)); }     // (kept: harmless, the sweeps already waited)

void timers_collect(tcif_ctx *c) {
    join(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    for (auto &e : c->evs) {
        float ms = 0.f;
        CUDA_OK(cudaEventElapsedTime(&ms, e.a, e.b));
        c->tMs[e.id] += ms; c->tBytes[e.id] += e.bytes; c->tLaunch[e.id] += 1;
        c->evPool.push_back(e);
    }
    c->evs.clear();
}

// ---- mesh upload: OpenFOAM ldu addressing -> owner-slot-major padded ELL ----------------
// Traversal order of the 256-cell tiles for the volume sweeps (DMesh::tileOrder).  Any permutation is valid
// (cells of one launch are independent); this one targets numberings that are banded like a structured
// block (blockMesh / simple decomposition output): the modal neighbour offset of each owner slot gives the
// strides (1, a, b) = (cell, row, plane).  Ascending order has a reuse distance of one plane (b cells) for the
// +-z neighbours, which at 400^3 is more than L2 holds across all streams of a sweep; walking the mesh in
// strips of H rows (all planes of strip 0, then strip 1, ...) shortens it to a*H cells at the price of one
// extra row fetch per strip edge.  Returns empty (= ascending) when no such structure is found, when one
// plane is small enough for L2 anyway (< 64 K cells), or with TCIF_TILE_ORDER=0.  stripCells = a*H (TCIF_STRIP_CELLS, default 16384, for the MULES sweeps; TCIF_FORCE_STRIP_CELLS for the force sweeps).
std::vector<int> build_tile_order(int nC, int stride, int Ko, const std::vector<int> &nbrOwn, long long stripCells) {
    const char *e = getenv("TCIF_TILE_ORDER");
    const int mode = e ? atoi(e) : 1;                  // 0 off, 1 auto, 2 force (also on small meshes)
    if (mode == 0 || Ko < 3) return {};
    std::vector<long long> offs;
    for (int k = 0; k < Ko; ++k) {
        std::map<long long, long long> hist;
        long long valid = 0;
        for (int c = 0; c < nC; c += (nC > (1 << 20) ? 7 : 1)) {      // a sample is enough for the mode
            int n = nbrOwn[(size_t)k * stride + c];
            if (n >= 0) { ++hist[(long long)n - c]; ++valid; }
        }
        long long best = 0, cnt = 0;
        for (auto &kv : hist) if (kv.second > cnt) { cnt = kv.second; best = kv.first; }
        if (valid == 0 || cnt < 0.8 * valid || best <= 0) return {};
        offs.push_back(best);
    }
    std::sort(offs.begin(), offs.end());
    const long long a = offs[offs.size() - 2], b = offs[offs.size() - 1];
    if (a < 2 || b <= a || b % a != 0) return {};
    if (mode == 1 && b * 512 < (32 << 20)) return {};  // a plane of all streams of a sweep (~0.5 KB/cell) fits in L2 anyway
    const long long R = b / a;                         // rows per plane
    if (stripCells <= 0) return {};
    long long H = std::max<long long>(2, stripCells / a);
    if (H >= R) return {};
    const long long nStrips = (R + H - 1) / H;
    H = (R + nStrips - 1) / nStrips;                   // even strips
    const int nTiles = cdiv(nC, 256);
    std::vector<int> order(nTiles);
    for (int t = 0; t < nTiles; ++t) order[t] = t;
    auto strip_of = [&](int t) { long long c0 = (long long)t * 256; return ((c0 % b) / a) / H; };
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return strip_of(x) < strip_of(y); });
    return order;
}

void build_mesh(tcif_ctx *c, const tcif_mesh_desc *m) {
    const int nC = m->nCells, nIF = m->nInternalFaces, nBF = m->nBoundaryFaces, nP = m->nPatches;
    if (nC <= 0 || nIF < 0 || nBF < 0 || nP < 0) failf("tcif_create: bad mesh sizes");
    if (nIF && (!m->owner || !m->neighbour || !m->Sf || !m->weights)) failf("tcif_create: null internal-face arrays");
    if (!m->V) failf("tcif_create: null V");
    c->nC = nC; c->nIF = nIF; c->nBF = nBF; c->nPatches = nP;
    c->stride = (nC + 31) / 32 * 32;
    const int stride = c->stride;
    // rank of each face among its owner's / neighbour's faces in ascending face index
    std::vector<int> cntO(nC, 0), cntN(nC, 0), kOwn(nIF), kNei(nIF);
    for (int f = 0; f < nIF; ++f) {
        int o = m->owner[f], n = m->neighbour[f];
        if (o < 0 || o >= nC || n < 0 || n >= nC || o == n) failf("tcif_create: face %d has bad owner/neighbour (%d,%d)", f, o, n);
        kOwn[f] = cntO[o]++; kNei[f] = cntN[n]++;
    }
    int Ko = 0, Kn = 0;
    for (int i = 0; i < nC; ++i) { Ko = std::max(Ko, cntO[i]); Kn = std::max(Kn, cntN[i]); }
    Ko = std::max(Ko, 1); Kn = std::max(Kn, 1);
    // fixed-K kernels (tcif_volume.cuh) want Ko == Kn == K in {2,3,4}: pad the smaller side
    { int K = std::max(Ko, Kn); if (K <= 4) { K = std::max(K, 2); Ko = Kn = K; } }
    c->Ko = Ko; c->Kn = Kn;
    c->nFd = (size_t)Ko * stride;
    if ((size_t)std::max(Ko, Kn) * stride > 0x7fffffffULL) failf("tcif_create: mesh too large for int32 face slots");
    std::vector<int> faceMap(nIF), nbrOwn((size_t)Ko * stride, -1), neiFace((size_t)Kn * stride, -1);
    std::vector<double> sx(c->nFd, 0.0), sy(c->nFd, 0.0), sz(c->nFd, 0.0), w(c->nFd, 0.0);
    for (int f = 0; f < nIF; ++f) {
        int o = m->owner[f], n = m->neighbour[f];
        size_t fd = (size_t)kOwn[f] * stride + o;
        faceMap[f] = (int)fd;
        nbrOwn[fd] = n;
        neiFace[(size_t)kNei[f] * stride + n] = (int)fd;
        sx[fd] = m->Sf[3 * (size_t)f]; sy[fd] = m->Sf[3 * (size_t)f + 1]; sz[fd] = m->Sf[3 * (size_t)f + 2];
        w[fd] = m->weights[f];
    }
    DMesh &d = c->dm;
    { const char *e = getenv("TCIF_PF_AHEAD"); d.pfTiles = (e ? atoi(e) : 148 * 256) / 256; }   // one CTA per SM ahead (profiles/r02_variants.md)
    d.nTiles = cdiv(nC, 256);
    {
        const char *sc = getenv("TCIF_STRIP_CELLS"), *fc = getenv("TCIF_FORCE_STRIP_CELLS");
        std::vector<int> order = build_tile_order(nC, stride, Ko, nbrOwn, sc ? atoll(sc) : 16384);
        d.tileOrder = order.empty() ? nullptr : c->upload(order);
        c->tileStrips = order.empty() ? 0 : 1;
        // the two force sweeps gather fewer arrays per cell: their own strip size (0 = ascending order)
        std::vector<int> forder = build_tile_order(nC, stride, Ko, nbrOwn, fc ? atoll(fc) : 0);
        c->forceOrder = forder.empty() ? nullptr : c->upload(forder);
    }
    d.nC = nC; d.stride = stride; d.Ko = Ko; d.Kn = Kn; d.nFd = (int)c->nFd; d.nBF = nBF; d.nPatches = nP;
    d.nbrOwn = c->upload(nbrOwn); d.neiFace = c->upload(neiFace);
    d.Sfx = c->upload(sx); d.Sfy = c->upload(sy); d.Sfz = c->upload(sz); d.w = c->upload(w);
    c->d_faceMap = c->upload(faceMap);
    std::vector<double> V(m->V, m->V + nC);
    double sumV = 0.0;
    for (int i = 0; i < nC; ++i) { if (!(V[i] > 0)) failf("tcif_create: cell %d has non-positive volume", i); sumV += V[i]; }
    c->sumV = sumV;
    d.V = c->upload(V);
    // patches
    c->patchType.assign(m->patchType, m->patchType + nP);
    c->patchStart.assign(m->patchStart, m->patchStart + nP);
    c->patchSize.assign(m->patchSize, m->patchSize + nP);
    c->patchNbr.assign(nP, -1);
    if (m->patchNeighbProcNo) c->patchNbr.assign(m->patchNeighbProcNo, m->patchNeighbProcNo + nP);
    std::vector<int> bfPatch(nBF, 0), bfType(nBF, 0), coupled, cOff, cSize, cPatchIdx;
    int covered = 0;
    for (int p = 0; p < nP; ++p) {
        int s = c->patchStart[p], n = c->patchSize[p], t = c->patchType[p];
        if (t < 0 || t > 3) failf("tcif_create: patch %d has unknown type %d", p, t);
        if (t == TCIF_PATCH_EMPTY && n != 0) failf("tcif_create: empty patch %d must have size 0 (fields carry no values on it)", p);
        if (n < 0 || s < 0 || s + n > nBF || s != covered) failf("tcif_create: patch %d range [%d,%d) is not contiguous in patch order", p, s, s + n);
        covered += n;
        if (t == TCIF_PATCH_PROCESSOR) {
            if (c->patchNbr[p] < 0) failf("tcif_create: processor patch %d without neighbour rank", p);
            c->cpPatch.push_back(p); c->cpOff.push_back((int)coupled.size());
            int off = (int)coupled.size();
            for (int i = 0; i < n; ++i) { coupled.push_back(s + i); cOff.push_back(off); cSize.push_back(n); cPatchIdx.push_back((int)c->cpPatch.size() - 1); }
        }
        for (int i = 0; i < n; ++i) { bfPatch[s + i] = p; bfType[s + i] = t; }
    }
    if (covered != nBF) failf("tcif_create: patches cover %d of %d boundary faces", covered, nBF);
    c->nCoupled = (int)coupled.size();
    // boundary faces grouped by cell (ascending cell, ascending boundary-face index = patch order)
    std::vector<int> fc(nBF);
    for (int b = 0; b < nBF; ++b) { fc[b] = m->faceCells[b]; if (fc[b] < 0 || fc[b] >= nC) failf("tcif_create: faceCells[%d] out of range", b); }
    std::vector<int> order(nBF);
    for (int b = 0; b < nBF; ++b) order[b] = b;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return fc[a] < fc[b]; });
    std::vector<int> bndCells, bndStart;
    for (int k = 0; k < nBF; ++k) {
        if (k == 0 || fc[order[k]] != fc[order[k - 1]]) { bndCells.push_back(fc[order[k]]); bndStart.push_back(k); }
    }
    bndStart.push_back(nBF);
    c->nBndCells = (int)bndCells.size();
    d.nBndCells = c->nBndCells; d.nCoupled = c->nCoupled;
    std::vector<unsigned> mask((size_t)stride / 32, 0u);
    for (int bc : bndCells) mask[bc >> 5] |= 1u << (bc & 31);
    d.bndMask = c->upload(mask);
    {
        std::vector<int> rank((size_t)stride / 32, 0);
        int acc = 0;
        for (size_t wd = 0; wd < rank.size(); ++wd) { rank[wd] = acc; acc += __builtin_popcount(mask[wd]); }
        d.bndWordRank = c->upload(rank);
    }
    d.bndCells = c->upload(bndCells); d.bndCellStart = c->upload(bndStart); d.bndFaces = c->upload(order);
    {
        // one 32-byte record per boundary-adjacent cell: its (<= 3) patch faces with their patch type and id, so that a
        // thread reaches the patch values with ONE dependent load after its cell index instead of three (start, face, type)
        std::vector<int> rec(8 * std::max<size_t>(bndCells.size(), 1), 0);
        for (size_t k = 0; k < bndCells.size(); ++k) {
            const int n = bndStart[k + 1] - bndStart[k];
            int *r = &rec[8 * k];
            r[0] = n <= 3 ? n : -n;
            for (int j = 0; j < 3; ++j) {
                const int bf = order[bndStart[k] + (j < n ? j : 0)];
                r[1 + j] = bf; r[4 + j] = bfType[bf] | (bfPatch[bf] << 4);
            }
        }
        d.bndRec = (const int4 *)c->upload(rec);
    }
    d.faceCells = c->upload(fc); d.bfPatch = c->upload(bfPatch); d.bfType = c->upload(bfType);
    std::vector<double> bx(nBF), by(nBF), bz(nBF), mag(nBF), dc(nBF), wb(nBF);
    for (int b = 0; b < nBF; ++b) {
        bx[b] = m->SfBoundary[3 * (size_t)b]; by[b] = m->SfBoundary[3 * (size_t)b + 1]; bz[b] = m->SfBoundary[3 * (size_t)b + 2];
        mag[b] = m->magSfBoundary ? m->magSfBoundary[b] : std::sqrt(bx[b] * bx[b] + by[b] * by[b] + bz[b] * bz[b]);
        dc[b] = m->deltaCoeffsBoundary ? m->deltaCoeffsBoundary[b] : 0.0;
        wb[b] = m->weightsBoundary ? m->weightsBoundary[b] : 1.0;
    }
    d.Sfbx = c->upload(bx); d.Sfby = c->upload(by); d.Sfbz = c->upload(bz);
    d.magSfb = c->upload(mag); d.dcb = c->upload(dc); d.wb = c->upload(wb);
    d.coupledFaces = c->upload(coupled); d.coupledOff = c->upload(cOff); d.coupledSize = c->upload(cSize);
    d.coupledPatchIdx = c->upload(cPatchIdx);
}

void alloc_fields(tcif_ctx *c) {
    const int P = c->P, Q = c->Q, nBF = c->nBF;
    const size_t cs = c->stride, fs = c->nFd, bs = std::max(nBF, 1);
    for (int p = 0; p < P; ++p) {
        c->alpha[p] = c->dalloc<double>(cs); c->alpha0[p] = c->dalloc<double>(cs); c->alphaNew[p] = c->dalloc<double>(cs);
        c->alphab[p] = c->dalloc<double>(bs);
        c->bc[p] = c->dalloc<int>(std::max(c->nPatches, 1));
        c->phiHO[p] = c->dalloc<double>(fs); c->phiHOb[p] = c->dalloc<double>(bs);
        c->phiPsi[p] = c->dalloc<double>(fs); c->phiPsib[p] = c->dalloc<double>(bs);
        c->phiCorr[p] = c->dalloc<double>(fs); c->phiCorrb[p] = c->dalloc<double>(bs); c->phiBDb[p] = c->dalloc<double>(bs);
        c->Qp[p] = c->dalloc<double>(cs); c->Qm[p] = c->dalloc<double>(cs);
        if (c->Ko != c->Kn || c->Ko < 2 || c->Ko > 4) { c->sP[p] = c->dalloc<double>(cs); c->sM[p] = c->dalloc<double>(cs); }   // generic-K kernels only
        c->lpA[p] = c->dalloc<double>(cs); c->lmA[p] = c->dalloc<double>(cs); c->lpB[p] = c->dalloc<double>(cs); c->lmB[p] = c->dalloc<double>(cs);
        c->lamA[p] = c->dalloc<double>(fs); c->lamB[p] = c->dalloc<double>(fs); c->lamb[p] = c->dalloc<double>(bs);
        c->lpHalo[p] = c->dalloc<double>(bs); c->lmHalo[p] = c->dalloc<double>(bs);
    }
    c->Tc = c->dalloc<double>(cs); c->Tb = c->dalloc<double>(bs);
    c->phi = c->dalloc<double>(fs); c->phib = c->dalloc<double>(bs);
    c->dSig = c->dalloc<SigmaModel>(std::max(Q, 1));
    c->Tq = c->dalloc<double>((size_t)std::max(Q, 1) * 6 * cs);
    c->gBnd = c->dalloc<double>((size_t)P * 3 * std::max(c->nBndCells, 1));
    c->TqHalo = c->dalloc<double>((size_t)std::max(Q, 1) * 6 * bs);
    c->F = c->dalloc<double>(3 * cs); c->Fb = c->dalloc<double>(3 * bs);
    {   // partial records of the boundary-adjacent cells: the largest of the sweeps' layouts (tcif_volume.cuh)
        const int slots = std::max(std::max(4 * P + 2, 1 + 9 * std::max(Q, 1)), 9 * P);
        c->dm.part = c->dalloc<double>((size_t)slots * std::max(c->nBndCells, 1));
    }
    c->stagingDoubles = std::max<size_t>(3 * cs, std::max<size_t>((size_t)std::max(c->nIF, 1), 3 * bs));
    c->staging = c->dalloc<double>(c->stagingDoubles);
    int maxComp = std::max(std::max(Q * 6, 3 * P), 3);
    c->haloDoubles = (size_t)std::max(c->nCoupled, 1) * maxComp;
    c->sendBuf = c->dalloc<double>(c->haloDoubles); c->recvBuf = c->dalloc<double>(c->haloDoubles);
    c->statsPart = c->dalloc<double>(1024 * 3); c->statsOut = c->dalloc<double>(16);
    CUDA_OK(cudaHostAlloc((void **)&c->h_small, 64 * sizeof(double), cudaHostAllocDefault));
    for (int i = 0; i < 16; ++i) c->cAlpha[i] = 1.0;
    c->dmaxDev = c->dalloc<double>(8);
    for (int q = 0; q < 6; ++q) { c->hSig[q].kind = SIGMA_CONSTANT; c->hSig[q].n = 1; memset(c->hSig[q].c, 0, sizeof c->hSig[q].c); }
}

// ---- scalar all-reduce over ranks (deterministic: gathered, combined in rank order) -----
enum { RED_SUM = 0, RED_MIN = 1, RED_MAX = 2 };
void allreduce_host(tcif_ctx *c, double *v, const int *ops, int n) {
    if (n > 16) failf("allreduce_host: n > 16");
    if (c->group) {
        tcif_group *g = c->group;
        memcpy(&g->red[(size_t)c->rank * 16], v, sizeof(double) * n);
        g->barrier();
        for (int i = 0; i < n; ++i) {
            double r = g->red[i];
            for (int k = 1; k < g->n; ++k) {
                double x = g->red[(size_t)k * 16 + i];
                r = ops[i] == RED_SUM ? r + x : ops[i] == RED_MIN ? std::min(r, x) : std::max(r, x);
            }
            v[i] = r;
        }
        g->barrier();
    } else if (c->nccl) {
        memcpy(c->h_small, v, sizeof(double) * n);
        CUDA_OK(cudaMemcpyAsync(c->statsOut, c->h_small, sizeof(double) * 16, cudaMemcpyHostToDevice, c->stream));
        NCCL_OK(g_nccl.AllGather(c->statsOut, c->d_gather, 16, kNcclDouble, c->nccl, c->stream));
        std::vector<double> all((size_t)c->nRanks * 16);
        CUDA_OK(cudaMemcpyAsync(all.data(), c->d_gather, all.size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
        for (int i = 0; i < n; ++i) {
            double r = all[i];
            for (int k = 1; k < c->nRanks; ++k) {
                double x = all[(size_t)k * 16 + i];
                r = ops[i] == RED_SUM ? r + x : ops[i] == RED_MIN ? std::min(r, x) : std::max(r, x);
            }
            v[i] = r;
        }
    }
}

// all-gather of n doubles per rank (same n everywhere) into all[rank*n + i]; host buffers
void allgather_host(tcif_ctx *c, const double *send, int n, std::vector<double> &all) {
    all.resize((size_t)c->nRanks * std::max(n, 0));
    if (n <= 0) return;
    if (c->group) {
        tcif_group *g = c->group;
        g->barrier();
        g->gatherPtr[c->rank] = send;
        g->barrier();
        for (int r = 0; r < g->n; ++r) memcpy(&all[(size_t)r * n], g->gatherPtr[r], sizeof(double) * n);
        g->barrier();
    } else if (c->nccl) {
        if (n > c->agCap) {             // grow-only device buffers: [n] to send, [n*nRanks] to receive
            if (c->agBuf) CUDA_OK(cudaFree(c->agBuf));
            c->agBuf = nullptr; c->agCap = 0;
            CUDA_OK(cudaMalloc((void **)&c->agBuf, sizeof(double) * (size_t)n * (c->nRanks + 1)));
            c->agCap = n;
        }
        double *ds = c->agBuf, *dr = c->agBuf + n;
        CUDA_OK(cudaMemcpyAsync(ds, send, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
        NCCL_OK(g_nccl.AllGather(ds, dr, n, kNcclDouble, c->nccl, c->stream));
        CUDA_OK(cudaMemcpyAsync(all.data(), dr, sizeof(double) * n * c->nRanks, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
    } else {
        memcpy(all.data(), send, sizeof(double) * n);
    }
}

void require_comm(tcif_ctx *c) {
    if (c->nCoupled > 0 && !c->group && !c->nccl)
        failf("mesh has processor patches but no communicator is attached (tcif_comm_init_nccl / tcif_comm_init_local)");
}

// deltaN = 1e-8/cbrt(average(mesh.V()))   phaseSystem.C:1287-1291 (global average)
void ensure_deltaN(tcif_ctx *c) {
    if (c->deltaNValid) return;
    double v[2] = {c->sumV, (double)c->nC};
    int ops[2] = {RED_SUM, RED_SUM};
    allreduce_host(c, v, ops, 2);
    c->deltaN = 1e-8 / cbrt(v[0] / v[1]);
    c->deltaNValid = true;
}

// ---- processor-patch exchange -----------------------------------------------------------
// send buffer (patch-major) -> neighbour ranks -> recv buffer.
void exchange(tcif_ctx *c, int ncomp, cudaStream_t st) {
    if (c->group) {
        tcif_group *g = c->group;
        CUDA_OK(cudaStreamSynchronize(st));
        g->barrier();
        std::string bad;                 // a rank-local throw between the barriers would leave the other ranks waiting
        for (size_t i = 0; i < c->cpPatch.size() && bad.empty(); ++i) {
            int p = c->cpPatch[i], e = c->patchNbr[p], n = c->patchSize[p];
            if (e < 0 || e >= g->n || !g->members[e]) { bad = "processor patch " + std::to_string(p) + ": neighbour rank " + std::to_string(e) + " not in group"; break; }
            tcif_ctx *o = g->members[e];
            int peerOff = -1;
            for (size_t j = 0; j < o->cpPatch.size(); ++j)
                if (o->patchNbr[o->cpPatch[j]] == c->rank) { peerOff = o->cpOff[j]; if (o->patchSize[o->cpPatch[j]] != n) bad = "processor patch size mismatch"; break; }
            if (peerOff < 0) bad = "rank " + std::to_string(e) + " has no patch facing rank " + std::to_string(c->rank);
            if (!bad.empty()) break;
            if (cudaMemcpyAsync(c->recvBuf + (size_t)c->cpOff[i] * ncomp, o->sendBuf + (size_t)peerOff * ncomp,
                                sizeof(double) * (size_t)n * ncomp, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
                bad = std::string("halo copy failed: ") + cudaGetErrorString(cudaGetLastError());
        }
        if (cudaStreamSynchronize(st) != cudaSuccess && bad.empty()) bad = std::string("halo copy failed: ") + cudaGetErrorString(cudaGetLastError());
        if (!bad.empty()) g->failed.store(1);
        g->barrier();
        if (!bad.empty()) failf("%s", bad.c_str());
        if (g->failed.load()) failf("processor-patch exchange failed on another rank of the group");
    } else if (c->nccl) {
        NCCL_OK(g_nccl.GroupStart());
        for (size_t i = 0; i < c->cpPatch.size(); ++i) {
            int p = c->cpPatch[i], e = c->patchNbr[p], n = c->patchSize[p];
            NCCL_OK(g_nccl.Send(c->sendBuf + (size_t)c->cpOff[i] * ncomp, (size_t)n * ncomp, kNcclDouble, e, c->nccl, st));
            NCCL_OK(g_nccl.Recv(c->recvBuf + (size_t)c->cpOff[i] * ncomp, (size_t)n * ncomp, kNcclDouble, e, c->nccl, st));
        }
        NCCL_OK(g_nccl.GroupEnd());
    } else failf("exchange without communicator");
}

// Map every neighbour's receive buffer (collective over the NCCL ranks).  Falls back to ncclSend/ncclRecv -- on ALL
// ranks together -- when the ranks are not on one node, a rank has too many processor patches, or the mapping fails.
// TCIF_PEER_HALO=0 keeps the NCCL path (the checked alternative).
void peer_setup(tcif_ctx *c) {
    const char *e = getenv("TCIF_PEER_HALO");
    const bool want = !(e && atoi(e) == 0);
    const int nR = c->nRanks;
    const size_t bufDoubles = 2 * c->haloDoubles + (size_t)(nR + 1) / 2 + 8;
    bool ok = want && (int)c->cpPatch.size() <= TCIF_MAX_PEER_PATCHES;
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof mine);
    if (ok) {
        ok = cudaMalloc((void **)&c->peerBuf, sizeof(double) * bufDoubles) == cudaSuccess;
        if (ok) {
            c->devBytes += (int64_t)(sizeof(double) * bufDoubles);
            c->peerFlags = (unsigned *)(c->peerBuf + 2 * c->haloDoubles);
            ok = cudaMemset(c->peerBuf, 0, sizeof(double) * bufDoubles) == cudaSuccess && cudaIpcGetMemHandle(&mine, c->peerBuf) == cudaSuccess;
        }
        if (!ok) cudaGetLastError();
    }
    // one all-gather: [ok, host id, handle (8 doubles), haloDoubles, cpOff of the patch facing each rank (nR)]
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t travels as 8 doubles");
    char host[256] = {0};
    gethostname(host, sizeof host - 1);
    unsigned long long hh = 1469598103934665603ULL;
    for (const char *q = host; *q; ++q) hh = (hh ^ (unsigned char)*q) * 1099511628211ULL;
    const int W = 11 + nR;
    std::vector<double> me(W, -1.0), all;
    me[0] = ok ? 1.0 : 0.0; me[1] = (double)(hh & 0xffffffffffffULL);
    memcpy(&me[2], &mine, 64);
    me[10] = (double)c->haloDoubles;
    for (size_t i = 0; i < c->cpPatch.size(); ++i) me[11 + c->patchNbr[c->cpPatch[i]]] = (double)c->cpOff[i];
    allgather_host(c, me.data(), W, all);
    for (int r = 0; r < nR; ++r) ok = ok && all[(size_t)r * W] == 1.0 && all[(size_t)r * W + 1] == me[1];
    c->peerMapped.assign(nR, nullptr);
    if (ok) {
        c->hp = HaloPeer{};
        c->hp.nPatches = (int)c->cpPatch.size();
        for (size_t i = 0; i < c->cpPatch.size() && ok; ++i) {
            const int nb = c->patchNbr[c->cpPatch[i]];
            const double peerOff = all[(size_t)nb * W + 11 + c->rank];
            if (peerOff < 0) { ok = false; break; }              // the neighbour has no patch facing this rank
            if (!c->peerMapped[nb]) {
                cudaIpcMemHandle_t h;
                memcpy(&h, &all[(size_t)nb * W + 2], 64);
                void *ptr = nullptr;
                if (cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
                c->peerMapped[nb] = ptr;
            }
            const size_t nbHalo = (size_t)all[(size_t)nb * W + 10];
            c->hp.base[i] = (double *)c->peerMapped[nb];
            c->hp.peerOff[i] = (int)peerOff;
            c->hp.remoteFlag[i] = (unsigned *)((double *)c->peerMapped[nb] + 2 * nbHalo) + c->rank;
            c->hp.localFlag[i] = c->peerFlags + nb;
            c->hp.stride[i] = nbHalo;
        }
    }
    double v[1] = {ok ? 0.0 : 1.0};
    const int ops[1] = {RED_MAX};
    allreduce_host(c, v, ops, 1);                                    // everybody uses the same transport
    if (v[0] != 0.0) {
        for (void *ptr : c->peerMapped) if (ptr) cudaIpcCloseMemHandle(ptr);
        c->peerMapped.assign(nR, nullptr);
        if (c->peerBuf) { cudaFree(c->peerBuf); c->peerBuf = nullptr; }
        cudaGetLastError();
        c->peerOn = false;
        return;
    }
    c->peerCounter = c->dalloc<int>(2);
    c->peerStatus = c->peerCounter + 1;
    CUDA_OK(cudaEventCreateWithFlags(&c->evX, cudaEventDisableTiming));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->peerOn = true;
}

void peer_check(tcif_ctx *c) {           // called where the host synchronises anyway
    if (!c->peerOn) return;
    int st = 0;
    CUDA_OK(cudaMemcpy(&st, c->peerStatus, sizeof(int), cudaMemcpyDeviceToHost));
    if (st) failf("processor-patch exchange over peer memory timed out waiting for a neighbour rank");
}

// processorFvPatchField::evaluate for ncomp cell arrays at once: dst[d][bf] := neighbour-rank src[d][faceCell]
void halo_cells(tcif_ctx *c, int ncomp, const double *const *src, double *const *dst, cudaStream_t st = nullptr) {
    if (!st) st = c->stream;
    if (c->nCoupled == 0) return;
    require_comm(c);
    if ((size_t)ncomp * c->nCoupled > c->haloDoubles || ncomp > 36) failf("halo_cells: too many components");
    HaloArgs h;
    h.ncomp = ncomp;
    for (int d = 0; d < ncomp; ++d) { h.src[d] = src[d]; h.dst[d] = dst[d]; }
    double bytes = 16.0 * ncomp * c->nCoupled;
    const int grid = cdiv(c->nCoupled, TCIF_TPB);
    if (c->peerOn) {
        timed_on(c, st, T_HALO, bytes, [&] {
            if (c->exSeq) CUDA_OK(cudaStreamWaitEvent(st, c->evX, 0));
            const unsigned seq = ++c->exSeq;
            const unsigned parity = seq & 1u;
            k_halo_push<<<grid, TCIF_TPB, 0, st>>>(c->dm, h, c->hp, parity, seq, c->peerCounter);
            k_halo_wait<<<1, 32, 0, st>>>(c->hp, seq, c->peerStatus);
            k_halo_pull<<<grid, TCIF_TPB, 0, st>>>(c->dm, h, c->peerBuf + (size_t)parity * c->haloDoubles);
            c->launches += 3;
            CUDA_OK(cudaEventRecord(c->evX, st));
        });
        return;
    }
    timed_on(c, st, T_HALO, bytes, [&] {
        k_halo_pack<<<grid, TCIF_TPB, 0, st>>>(c->dm, h, c->sendBuf);
        c->launches++;
        exchange(c, ncomp, st);
        k_halo_unpack<<<grid, TCIF_TPB, 0, st>>>(c->dm, h, c->recvBuf);
        c->launches++;
    });
}

// ---- kernel dispatch ----------------------------------------------------------------------
int kshape(const tcif_ctx *c) { return (c->Ko == c->Kn && c->Ko >= 2 && c->Ko <= 4) ? c->Ko : 0; }

// fixed-K meshes: the sweeps of tcif_volume.cuh (one body for the volume launch and the deferred launch);
// other meshes: the generic-K kernels k_* of tcif_device.cuh (volume launch + BND twin over all boundary cells)
const SweepFns *fns(const tcif_ctx *c) {
    switch (kshape(c)) {
        case 2: return sweep_fns_k2();
        case 3: return sweep_fns_k3();
        case 4: return sweep_fns_k4();
        default: return nullptr;
    }
}

#define DISPATCH_P(...)                                      \
    switch (c->P) {                                          \
        case 2: { constexpr int PP = 2; __VA_ARGS__; } break; \
        case 3: { constexpr int PP = 3; __VA_ARGS__; } break; \
        case 4: { constexpr int PP = 4; __VA_ARGS__; } break; \
        default: failf("nPhases must be 2..4");              \
    }

ForceArgs force_args(tcif_ctx *c, bool withTensor, bool withG) {
    ForceArgs a{};
    for (int p = 0; p < c->P; ++p) { a.alpha[p] = c->alpha[p]; a.alphab[p] = c->alphab[p]; a.bc[p] = c->bc[p]; }
    a.T = c->Tc; a.Tb = c->Tb; a.sig = c->dSig; a.deltaN = c->deltaN;
    a.Tq = withTensor ? c->Tq : nullptr; a.gBnd = c->gBnd; a.TqHalo = c->TqHalo;
    a.G = withG ? c->G : nullptr; a.F = c->F;
    a.tsum = c->tsum;
    return a;
}

// the force sweeps gather few arrays (their +-z planes fit in L2): ascending tile order is faster for them
// (profiles/r02_variants.md), the strip order is for the MULES sweeps
DMesh dm_force(const tcif_ctx *c) { DMesh d = c->dm; d.tileOrder = c->forceOrder; return d; }
// BND: the O(surface) launch over the boundary-adjacent cells (fixed-K: same body as the volume launch; generic K: twin)
template <bool BND, bool WG> void launch_grad_tensor(tcif_ctx *c, const ForceArgs &a) {
    if (const SweepFns *f = fns(c)) {
        if (BND && !c->nBndCells) return;
        f->grad_tensor(c->P, BND, c->stream, BND ? c->dm : dm_force(c), a);
        c->launches++;
        return;
    }
    int n = BND ? c->nBndCells : c->nC;
    if (n == 0) return;
    DISPATCH_P(k_grad_tensor<PP, 0, 0, BND, WG><<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a));
    c->launches++;
}
template <bool BND> void launch_div_tensor(tcif_ctx *c, const ForceArgs &a) {
    if (const SweepFns *f = fns(c)) {
        if (BND && !c->nBndCells) return;
        f->div_tensor(c->P, BND, c->stream, dm_force(c), a);
        c->launches++;
        return;
    }
    int n = BND ? c->nBndCells : c->nC;
    if (n == 0) return;
    DISPATCH_P(k_div_tensor<PP, 0, 0, BND><<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a));
    c->launches++;
}

// compulsory HBM bytes of the volume sweeps (every array touched once; DESIGN.md "Kernels")
double bytes_grad_tensor(const tcif_ctx *c, bool tensor, bool g) {
    double perCell = 8.0 * c->P + 8 /*V*/ + 4.0 * (c->Ko + c->Kn) + 32.0 * c->Ko;
    if (tensor) perCell += 48.0 * (c->tsum ? 1 : c->Q) + (c->haveT ? 8 : 0);
    if (g) perCell += 24.0 * c->P;
    return perCell * c->nC;
}
double bytes_div_tensor(const tcif_ctx *c) {
    return (48.0 * (c->tsum ? 1 : c->Q) + 8 + 4.0 * (c->Ko + c->Kn) + 32.0 * c->Ko + 24) * c->nC;
}
// MULES sweeps: arrays shared by the phases of a batch (slot indices, V, phi) are counted once per launch
double bytes_mules_setup(const tcif_ctx *c, int nPh, bool storeOld = false) {
    double shared = 4.0 * (c->Ko + c->Kn) + 8 /*V*/ + 8.0 * c->Ko /*phi*/;
    (void)storeOld;      // psi0 is one 8 B stream either way: read, or (storeOld) written
    const bool sums = kshape(c) == 0;        // only the generic-K kernels store sumPhip / mSumPhim
    double perPh = 8 + 8 /*psi psi0*/ + 8.0 * c->Ko * 2 /*phiIn phiCorr*/ + 32 /*Q+-,l+-*/ + (sums ? 16 : 0);
    return (double)c->nC * (shared + nPh * perPh);
}
double bytes_mules_iter(const tcif_ctx *c, int nPh, bool first) {
    double shared = 4.0 * (c->Ko + c->Kn);
    double perPh = 8.0 * c->Ko * (first ? 2 : 3) + 16 + 16 /*l+- in, out*/ + 16 /*Q+-*/ + (kshape(c) == 0 ? 16 : 0) /*s+-*/;
    return (double)c->nC * (shared + nPh * perPh);
}
double bytes_mules_final(const tcif_ctx *c, int nPh, int mode, bool first) {
    double shared = 4.0 * (c->Ko + c->Kn);
    double perPh = 8.0 * c->Ko * (first ? 2 : 3) + 16;
    if (mode >= 1) { shared += 8.0 * c->Ko; perPh += 8; }      // phi; psi
    if (mode == 2) { shared += 8; perPh += 16; }               // V; psi0, psiNew
    return (double)c->nC * (shared + nPh * perPh);
}

inline void touch(tcif_ctx *c) { ++c->fieldsVersion; }   // alpha / T / sigma changed: F is stale
// the device changed alpha_p: F is stale and no host field equals the device copy any more
inline void alpha_changed(tcif_ctx *c, int p) { ++c->fieldsVersion; c->alphaStamp[p] = 0; }
inline bool sigma_needs_T(const tcif_ctx *c) {
    for (int q = 0; q < c->Q; ++q) if (!c->sigSet[q] || c->hSig[q].kind != SIGMA_CONSTANT) return true;
    return false;
}

void do_force(tcif_ctx *c) {
    if (c->Q < 1) failf("divCapillaryStress needs at least two phases");
    for (int q = 0; q < c->Q; ++q) {
        if (!c->sigSet[q]) failf("surfaceTension entry missing for pair %d (all pairs need one; phaseSystem.C:1135)", q);
        if (c->hSig[q].kind != SIGMA_CONSTANT && !c->haveT) failf("temperatureDependent sigma without a temperature field");
    }
    require_comm(c);
    if (c->cacheForce && c->forceVersion == c->fieldsVersion) { c->forceCacheHits++; return; }
    c->forceEvaluations++;
    ensure_deltaN(c);
    ForceArgs a = force_args(c, true, false);
    sweep(c, [&] { timed(c, T_GRAD_TENSOR, bytes_grad_tensor(c, true, false), [&] { launch_grad_tensor<false, false>(c, a); }); },
             [&] { timed(c, T_BOUNDARY, 0, [&] { launch_grad_tensor<true, false>(c, a); }); });
    sweep(c, [&] { timed(c, T_DIV_TENSOR, bytes_div_tensor(c), [&] { launch_div_tensor<false>(c, a); }); },
             [&] {                   // neighbour-rank tensors of the processor faces: hidden behind the volume sweep
                 const double *src[36]; double *dst[36];
                 const int nT = (c->tsum ? 1 : c->Q) * 6;
                 for (int k = 0; k < nT; ++k) { src[k] = c->Tq + (size_t)k * c->stride; dst[k] = c->TqHalo + (size_t)k * std::max(c->nBF, 1); }
                 halo_cells(c, nT, src, dst, c->side);
             },
             [&] { timed(c, T_BOUNDARY, 0, [&] { launch_div_tensor<true>(c, a); }); });
    join(c);
    if (c->nBF) {
        timed(c, T_BOUNDARY, 0, [&] {
            k_extrapolate_bnd<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, 3, (size_t)c->stride, c->F, c->Fb);
            c->launches++;
        });
        if (c->nCoupled) {
            const double *src[3]; double *dst[3];
            for (int d = 0; d < 3; ++d) { src[d] = c->F + (size_t)d * c->stride; dst[d] = c->Fb + (size_t)d * c->nBF; }
            halo_cells(c, 3, src, dst);
        }
    }
    c->forceVersion = c->fieldsVersion;
}

// fvc::grad(alpha_p) for every phase into G / Gb (Gauss linear + gaussGrad boundary correction)
void do_grad(tcif_ctx *c) {
    require_comm(c);
    if (!c->G) { c->G = c->dalloc<double>((size_t)c->P * 3 * c->stride); c->Gb = c->dalloc<double>((size_t)c->P * 3 * std::max(c->nBF, 1)); }
    ForceArgs a = force_args(c, false, true);
    sweep(c, [&] { timed(c, T_GRAD_TENSOR, bytes_grad_tensor(c, false, true), [&] { launch_grad_tensor<false, true>(c, a); }); },
             [&] { timed(c, T_BOUNDARY, 0, [&] { launch_grad_tensor<true, true>(c, a); }); });
    join(c);
    if (!c->nBF) return;
    for (int p = 0; p < c->P; ++p) {
        double *gi = c->G + (size_t)p * 3 * c->stride, *gb = c->Gb + (size_t)p * 3 * c->nBF;
        timed(c, T_BOUNDARY, 0, [&] {
            k_extrapolate_bnd<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, 3, (size_t)c->stride, gi, gb);
            c->launches++;
        });
        if (c->nCoupled) {
            const double *src[3]; double *dst[3];
            for (int d = 0; d < 3; ++d) { src[d] = gi + (size_t)d * c->stride; dst[d] = gb + (size_t)d * c->nBF; }
            halo_cells(c, 3, src, dst);
        }
        timed(c, T_BOUNDARY, 0, [&] {
            k_grad_bnd_correct<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->alpha[p], c->alphab[p], c->bc[p], gb);
            c->launches++;
        });
    }
}

// psi.correctBoundaryConditions() for a set of phases: zeroGradient patches + processor halo
void correct_alpha_bcs(tcif_ctx *c, const int *phases, int nPh) {
    for (int k = 0; k < nPh; ++k) alpha_changed(c, phases[k]);
    if (!c->nBF) return;
    BcArgs ba{};
    ba.n = nPh;
    for (int k = 0; k < nPh; ++k) { const int p = phases[k]; ba.vi[k] = c->alpha[p]; ba.vb[k] = c->alphab[p]; ba.bc[k] = c->bc[p]; }
    timed(c, T_BOUNDARY, 0, [&] {
        k_correct_bc_scalar<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, ba);
        c->launches++;
    });
    if (c->nCoupled) {
        const double *src[4]; double *dst[4];
        for (int k = 0; k < nPh; ++k) { src[k] = c->alpha[phases[k]]; dst[k] = c->alphab[phases[k]]; }
        halo_cells(c, nPh, src, dst);
    }
}

MulesBatch mules_batch(tcif_ctx *c, const int *phases, int nPh, bool fromHO) {
    MulesBatch b{};
    b.nPh = nPh;
    for (int k = 0; k < nPh; ++k) {
        int p = phases[k];
        MulesPh &ph = b.ph[k];
        ph.psi = c->alpha[p]; ph.psi0 = c->alpha0[p]; ph.psi0w = c->alpha0[p]; ph.psib = c->alphab[p];
        ph.Su = c->haveSu[p] ? c->Su[p] : nullptr; ph.Sp = c->haveSp[p] ? c->Sp[p] : nullptr;
        ph.bc = c->bc[p];
        ph.phiIn = fromHO ? c->phiHO[p] : c->phiPsi[p]; ph.phiInb = fromHO ? c->phiHOb[p] : c->phiPsib[p];
        ph.phiPsi = c->phiPsi[p]; ph.phiPsib = c->phiPsib[p];
        ph.phiCorr = c->phiCorr[p]; ph.phiCorrb = c->phiCorrb[p]; ph.phiBDb = c->phiBDb[p];
        ph.Qp = c->Qp[p]; ph.Qm = c->Qm[p]; ph.sP = c->sP[p]; ph.sM = c->sM[p];
        ph.lpA = c->lpA[p]; ph.lmA = c->lmA[p]; ph.lpB = c->lpB[p]; ph.lmB = c->lmB[p];
        ph.lamA = c->lamA[p]; ph.lamB = c->lamB[p]; ph.lamb = c->lamb[p];
        ph.lpHalo = c->lpHalo[p]; ph.lmHalo = c->lmHalo[p];
        ph.psiNew = c->alphaNew[p];
    }
    return b;
}

// ping-pong buffers of the limiter resolved for one sweep
MulesBatch resolve_flip(const MulesBatch &b0, int flip) {
    MulesBatch b = b0;
    for (int k = 0; k < b.nPh; ++k) {
        MulesPh &ph = b.ph[k];
        ph.lpIn = flip ? ph.lpB : ph.lpA; ph.lmIn = flip ? ph.lmB : ph.lmA; ph.lamIn = flip ? ph.lamB : ph.lamA;
        ph.lpOut = flip ? ph.lpA : ph.lpB; ph.lmOut = flip ? ph.lmA : ph.lmB; ph.lamOut = flip ? ph.lamA : ph.lamB;
    }
    return b;
}

template <bool BND> void launch_setup(tcif_ctx *c, const MulesBatch &b, const MulesParams &prm) {
    if (const SweepFns *f = fns(c)) {
        if (BND && !c->nBndCells) return;
        f->setup(b.nPh, BND, c->stream, c->dm, b, prm);
        c->launches++;
        return;
    }
    int n = BND ? c->nBndCells : c->nC;
    if (n == 0) return;
    k_mules_setup<0, 0, BND><<<cdiv(n, TCIF_TPB) * b.nPh, TCIF_TPB, 0, c->stream>>>(c->dm, b, prm);
    c->launches++;
}
template <bool BND> void launch_iter(tcif_ctx *c, const MulesBatch &b, int first, int flip) {
    if (const SweepFns *f = fns(c)) {
        if (BND && !c->nBndCells) return;
        f->iter(b.nPh, first != 0, BND, c->stream, c->dm, resolve_flip(b, flip));
        c->launches++;
        return;
    }
    int n = BND ? c->nBndCells : c->nC;
    if (n == 0) return;
    k_mules_iter<0, 0, BND><<<cdiv(n, TCIF_TPB) * b.nPh, TCIF_TPB, 0, c->stream>>>(c->dm, b, first, flip);
    c->launches++;
}
template <bool BND, int MODE>
void launch_final(tcif_ctx *c, const MulesBatch &b, const MulesParams &prm, int first, int flip, int nIter, bool lsum = false) {
    if (const SweepFns *f = fns(c)) {
        if (BND && !c->nBndCells) return;
        f->fin(b.nPh, MODE, first != 0, lsum, BND, c->stream, c->dm, resolve_flip(b, flip), prm, nIter);
        c->launches++;
        return;
    }
    if (lsum) failf("launch_final: fused limitSum needs the fixed-K kernels");
    int n = BND ? c->nBndCells : c->nC;
    if (n == 0) return;
    k_mules_final<0, 0, BND, MODE><<<cdiv(n, TCIF_TPB) * b.nPh, TCIF_TPB, 0, c->stream>>>(c->dm, b, prm, first, flip, nIter);
    c->launches++;
}

void halo_lambda(tcif_ctx *c, const MulesBatch &b, int flip) {
    if (!c->nCoupled) return;
    const double *src[8]; double *dst[8];
    for (int k = 0; k < b.nPh; ++k) {
        src[2 * k] = flip ? b.ph[k].lpB : b.ph[k].lpA; dst[2 * k] = b.ph[k].lpHalo;
        src[2 * k + 1] = flip ? b.ph[k].lmB : b.ph[k].lmA; dst[2 * k + 1] = b.ph[k].lmHalo;
    }
    halo_cells(c, 2 * b.nPh, src, dst, c->side);
}

// MULES::limit for a batch of phases.  mode 0: returnCorr=true; 1: returnCorr=false;
// 2: explicitSolve's limit fused with surfaceIntegrate + Euler update (psi -> alphaNew).
// fuseLimitSum (mode 0, batch = all phases): limitSum of the owner faces rides on the apply sweep;
// the caller finishes with do_limit_sum_rest().
// storeOld: psi.oldTime() := psi for the phases of the batch before anything else (rides on the set-up sweep).
void do_limit(tcif_ctx *c, const int *phases, int nPh, double rDeltaT, int mode, bool addUpwind, bool fromHO,
              bool fuseLimitSum = false, bool storeOld = false) {
    require_comm(c);
    MulesBatch b = mules_batch(c, phases, nPh, fromHO);
    MulesParams prm{};
    prm.rDeltaT = rDeltaT; prm.psiMax = 1.0; prm.psiMin = 0.0;
    prm.extremaCoeff = c->cfg.extremaCoeff; prm.smoothLimiter = c->cfg.smoothLimiter;
    prm.boundaryExtrema = c->cfg.boundaryExtrema; prm.addUpwind = addUpwind ? 1 : 0;
    prm.phi = c->phi; prm.phib = c->phib;
    if (storeOld) {
        if (fns(c)) prm.storeOld = 1;
        else for (int k = 0; k < nPh; ++k)
            CUDA_OK(cudaMemcpyAsync(c->alpha0[phases[k]], c->alpha[phases[k]], sizeof(double) * c->nC, cudaMemcpyDeviceToDevice, c->stream));
    }
    const int L = c->cfg.nLimiterIter;
    sweep(c, [&] { timed(c, T_MULES_SETUP, bytes_mules_setup(c, nPh, prm.storeOld != 0), [&] { launch_setup<false>(c, b, prm); }); },
             [&] { timed(c, T_BOUNDARY, 0, [&] { launch_setup<true>(c, b, prm); }); });
    int first = 1, flip = 0;
    for (int j = 1; j < L; ++j) {
        sweep(c, [&] { timed(c, T_MULES_ITER, bytes_mules_iter(c, nPh, first), [&] { launch_iter<false>(c, b, first, flip); }); },
                 [&] { halo_lambda(c, b, flip); },      // syncTools::syncFaceList(min) of the previous iteration
                 [&] { timed(c, T_BOUNDARY, 0, [&] { launch_iter<true>(c, b, first, flip); }); });
        flip ^= 1; first = 0;
    }
    double fb = bytes_mules_final(c, nPh, mode, first);
    if (fuseLimitSum && (mode != 0 || !fns(c))) failf("do_limit: fused limitSum needs the fixed-K kernels and returnCorr=true");
    auto fin = [&](auto modeTag) {
        constexpr int MODE = decltype(modeTag)::value;
        sweep(c, [&] { timed(c, T_MULES_FINAL, fb, [&] { launch_final<false, MODE>(c, b, prm, first, flip, L, fuseLimitSum); }); },
                 [&] { if (L > 0) halo_lambda(c, b, flip); },
                 [&] { timed(c, T_BOUNDARY, 0, [&] { launch_final<true, MODE>(c, b, prm, first, flip, L, fuseLimitSum); }); });
    };
    if (mode == 0) fin(std::integral_constant<int, 0>{});
    else if (mode == 1) fin(std::integral_constant<int, 1>{});
    else {
        // explicitSolve on fixed-K meshes: the limited fluxes first (MODE 1), then surfaceIntegrate + Euler update from the
        // STORED fluxes in a second, lighter sweep; TCIF_SPLIT_FINAL=0 does both in one sweep (MODE 2), which re-derives the
        // flux of every neighbour-side face from five gathers and is 1 ms per stage slower at 64 M cells although it
        // moves fewer bytes (profiles/r03_variants.md).  Bit-identical either way.
        static const int split = [] { const char *e = getenv("TCIF_SPLIT_FINAL"); return e ? atoi(e) : 1; }();
        const SweepFns *f = fns(c);
        if (split && f) {
            fb = bytes_mules_final(c, nPh, 1, first);
            fin(std::integral_constant<int, 1>{});
            const double ub = (double)c->nC * (4.0 * (c->Ko + c->Kn) + 8 + nPh * (8.0 * c->Ko + 16));
            sweep(c, [&] { timed(c, T_MULES_FINAL, ub, [&] { f->update(nPh, false, c->stream, c->dm, b, prm); c->launches++; }); },
                     [&] { timed(c, T_BOUNDARY, 0, [&] { if (c->nBndCells) { f->update(nPh, true, c->stream, c->dm, b, prm); c->launches++; } }); });
        } else fin(std::integral_constant<int, 2>{});
        for (int k = 0; k < nPh; ++k) { std::swap(c->alpha[phases[k]], c->alphaNew[phases[k]]); alpha_changed(c, phases[k]); }
    }
    join(c);
}

void do_limit_sum_coupled(tcif_ctx *c) {
    if (!c->nCoupled) return;
    LimitSumArgs ab{};
    ab.nPh = c->P;
    for (int p = 0; p < c->P; ++p) ab.f[p] = c->phiPsib[p];
    timed(c, T_BOUNDARY, 0, [&] {
        k_limit_sum_coupled<<<cdiv(c->nCoupled, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nCoupled, c->dm.coupledFaces, ab);
        c->launches++;
    });
}

void do_limit_sum(tcif_ctx *c) {
    LimitSumArgs a{};
    a.nPh = c->P;
    for (int p = 0; p < c->P; ++p) a.f[p] = c->phiPsi[p];
    timed(c, T_LIMIT_SUM, 16.0 * c->P * c->nFd, [&] {
        k_limit_sum<<<cdiv((long long)c->nFd, TCIF_TPB), TCIF_TPB, 0, c->stream>>>((int)c->nFd, a);
        c->launches++;
    });
    do_limit_sum_coupled(c);
}

// whether limit(all phases, returnCorr=true) + limitSum can run fused (do_limit(..., fuseLimitSum=true) + do_limit_sum_rest)
bool can_fuse_limit_sum(tcif_ctx *c) { return fns(c) != nullptr; }
// the part of limitSum the fused sweep did not cover: the coupled patch faces (every internal face is owned by a cell
// of the volume or the deferred launch)
void do_limit_sum_rest(tcif_ctx *c) { do_limit_sum_coupled(c); }

// MULES::explicitSolve for a batch of phases (optionally preceded by phiAlpha += upwind flux)
void do_explicit(tcif_ctx *c, const int *phases, int nPh, double deltaT, bool addUpwind) {
    if (!(deltaT > 0)) failf("explicitSolve: deltaT must be positive");
    correct_alpha_bcs(c, phases, nPh);
    do_limit(c, phases, nPh, 1.0 / deltaT, 2, addUpwind, false);
    correct_alpha_bcs(c, phases, nPh);
}

// phiAlpha[p] += upwind<scalar>(mesh, phi).flux(alpha_p)
void do_add_upwind(tcif_ctx *c, int phase) {
    timed(c, T_MULES_SETUP, 0, [&] {
        int grid = cdiv(c->nC, TCIF_TPB);
        k_add_upwind<0><<<grid, TCIF_TPB, 0, c->stream>>>(c->dm, c->phi, c->alpha[phase], c->phiPsi[phase]);
        c->launches++;
        if (c->nBF) {
            k_add_upwind_bnd<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->phib, c->alpha[phase], c->alphab[phase], c->phiPsib[phase]);
            c->launches++;
        }
    });
}

void ensure_G(tcif_ctx *c) {
    if (!c->G) { c->G = c->dalloc<double>((size_t)c->P * 3 * c->stride); c->Gb = c->dalloc<double>((size_t)c->P * 3 * std::max(c->nBF, 1)); }
}

// phic_k = min(pa[k]*phic0, pb[k]): closed form of the cumulative `phic = min(cAlpha*phic, max(phic))`
// (multiphaseSystem.C:237) over the (phase1, phase2) visits of ALL correctors of one solveAlphas() call
struct PhicState { double a = 1.0, b = 1e300, M0 = 0.0; bool trivial = true; };

void phic_begin(tcif_ctx *c, PhicState &st) {
    st = PhicState();
    for (int p1 = 0; p1 < c->P; ++p1) for (int p2 = 0; p2 < c->P; ++p2) if (p1 != p2 && c->cAlpha[p1 * c->P + p2] != 1.0) st.trivial = false;
    if (st.trivial) return;                       // cAlpha == 1 everywhere: min(phic, max(phic)) == phic
    CUDA_OK(cudaMemsetAsync(c->dmaxDev, 0, sizeof(double) * 8, c->stream));
    long long n = (long long)c->nFd + c->nBF;
    k_phic_max<<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->phi, c->phib, c->dmaxDev);
    c->launches++;
    CUDA_OK(cudaMemcpyAsync(c->h_small, c->dmaxDev, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    double v[1] = {c->h_small[0]};
    int ops[1] = {RED_MAX};
    allreduce_host(c, v, ops, 1);
    st.M0 = v[0];
}

// fvc::flux(phi, alpha_p, scheme) + interface compression for every phase -> phiHO[p] (multiphaseSystem.C:199-277)
void do_flux_assembly(tcif_ctx *c, int schemePhi, int schemePhir, PhicState &st) {
    for (int sc : {schemePhi, schemePhir}) {
        if (sc < 0 || sc > 2) failf("unsupported convection scheme %d (upwind | linear | vanLeer)", sc);
        if (sc == SCHEME_VANLEER && !c->Cx) failf("vanLeer needs cell centres: call tcif_set_cell_centres first");
    }
    require_comm(c);
    ensure_deltaN(c);
    ensure_G(c);
    ForceArgs ga = force_args(c, false, true);
    sweep(c, [&] { timed(c, T_GRAD_TENSOR, bytes_grad_tensor(c, false, true), [&] { launch_grad_tensor<false, true>(c, ga); }); },
             [&] { timed(c, T_BOUNDARY, 0, [&] { launch_grad_tensor<true, true>(c, ga); }); });
    join(c);
    if (c->nCoupled) {                            // neighbour-rank gradients on processor faces
        const double *src[36]; double *dst[36];
        for (int k = 0; k < c->P * 3; ++k) { src[k] = c->G + (size_t)k * c->stride; dst[k] = c->Gb + (size_t)k * c->nBF; }
        halo_cells(c, c->P * 3, src, dst);
    }
    FluxArgs a{};
    a.P = c->P; a.schemePhi = schemePhi; a.schemePhir = schemePhir;
    for (int p = 0; p < c->P; ++p) { a.alpha[p] = c->alpha[p]; a.alphab[p] = c->alphab[p]; a.out[p] = c->phiHO[p]; a.outb[p] = c->phiHOb[p]; }
    a.G = c->G; a.Gb = c->Gb; a.Cx = c->Cx; a.Cy = c->Cy; a.Cz = c->Cz; a.dbx = c->dbx; a.dby = c->dby; a.dbz = c->dbz;
    a.phi = c->phi; a.phib = c->phib; a.deltaN = c->deltaN;
    int step = 0;
    for (int p1 = 0; p1 < c->P; ++p1)
        for (int p2 = 0; p2 < c->P; ++p2) {
            if (p1 == p2) continue;
            if (!st.trivial) {
                double cA = c->cAlpha[p1 * c->P + p2];
                double M = std::min(st.a * st.M0, st.b);
                st.b = std::min(cA * st.b, M);
                st.a = cA * st.a;
            }
            a.pa[step] = st.a; a.pb[step] = st.b;
            ++step;
        }
    double bytes = (double)c->nC * (8.0 * c->P + 24.0 * c->P + 4.0 * c->Ko + 8.0 * c->Ko * (5 + c->P) + (c->Cx ? 24 : 0));
    timed(c, T_FLUX_ASSEMBLY, bytes, [&] {
        int grid = cdiv(c->nC, TCIF_TPB_FLUX);
        DISPATCH_P(k_flux_assembly<PP><<<grid, TCIF_TPB_FLUX, 0, c->stream>>>(c->dm, a));
        c->launches++;
        if (c->nBF) {
            DISPATCH_P(k_flux_assembly_bnd<PP><<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a));
            c->launches++;
        }
    });
}

// implicit upwind predictor for a batch of phases (multiphaseSystem.C:348-363): Jacobi on the M-matrix
// EulerDdt + upwind fvmDiv == Su + Sp until max|dpsi| <= tol per phase; then psi.correctBoundaryConditions()
// and phiAlpha += alpha1Eqn.flux().
// addFlux=false leaves `phiAlpha += alpha1Eqn.flux()` to the caller (solve() folds it into explicitSolve's set-up
// sweep, which evaluates the same upwind flux of the same psi anyway).
void do_predictor(tcif_ctx *c, const int *phases, int nPh, double deltaT, double tol, int maxIter, bool addFlux = true) {
    if (!(deltaT > 0)) failf("predictor: deltaT must be positive");
    require_comm(c);
    for (int k = 0; k < nPh; ++k) {
        int p = phases[k];
        if (!c->predDiag[p]) { c->predDiag[p] = c->dalloc<double>(c->stride); c->predRhs[p] = c->dalloc<double>(c->stride); }
        c->predictorSweeps[p] = 0;
    }
    auto batch = [&](const bool *active) {
        PredBatch b{};
        b.nPh = nPh; b.phi = c->phi; b.phib = c->phib; b.rDeltaT = 1.0 / deltaT; b.dmax = c->dmaxDev;
        for (int k = 0; k < nPh; ++k) {
            int p = phases[k];
            PredPh &ph = b.ph[k];
            ph.psi0 = c->alpha0[p]; ph.Su = c->haveSu[p] ? c->Su[p] : nullptr; ph.Sp = c->haveSp[p] ? c->Sp[p] : nullptr;
            ph.bc = c->bc[p]; ph.psib = c->alphab[p]; ph.diag = c->predDiag[p]; ph.rhs = c->predRhs[p];
            ph.x = c->alpha[p]; ph.xn = c->alphaNew[p]; ph.active = active ? (active[k] ? 1 : 0) : 1;
        }
        return b;
    };
    int grid = cdiv(c->nC, TCIF_TPB) * nPh, gridB = cdiv(std::max(c->nBndCells, 1), TCIF_TPB) * nPh;
    {
        PredBatch b = batch(nullptr);
        sweep(c, [&] { timed(c, T_PREDICTOR, (double)nPh * c->nC * (40.0 + 4.0 * (c->Ko + c->Kn) + 8.0 * c->Ko), [&] {
                           k_pred_setup<false><<<grid, TCIF_TPB, 0, c->stream>>>(c->dm, b); c->launches++; }); },
                 [&] { if (c->nBndCells) timed(c, T_BOUNDARY, 0, [&] {
                           k_pred_setup<true><<<gridB, TCIF_TPB, 0, c->stream>>>(c->dm, b); c->launches++; }); });
    }
    bool active[4] = {true, true, true, true};
    const SweepFns *kvpPred = fns(c);
    for (int it = 0; it < maxIter; ++it) {
        bool any = false;
        for (int k = 0; k < nPh; ++k) any = any || active[k];
        if (!any) break;
        CUDA_OK(cudaMemsetAsync(c->dmaxDev, 0, sizeof(double) * 8, c->stream));
        PredBatch b = batch(active);
        sweep(c, [&] { timed(c, T_PREDICTOR, (double)nPh * c->nC * (32.0 + 4.0 * (c->Ko + c->Kn) + 8.0 * c->Ko), [&] {
                           if (kvpPred) kvpPred->pred_jacobi(nPh, false, c->stream, c->dm, b);
                           else k_pred_jacobi<false><<<grid, TCIF_TPB, 0, c->stream>>>(c->dm, b);
                           c->launches++; }); },
                 [&] {                            // patchNeighbourField of the iterate
                     const double *src[4]; double *dst[4]; int n = 0;
                     for (int k = 0; k < nPh; ++k) if (active[k]) { src[n] = c->alpha[phases[k]]; dst[n] = c->alphab[phases[k]]; ++n; }
                     halo_cells(c, n, src, dst, c->side);
                 },
                 [&] {
                     if (c->nBndCells) timed(c, T_BOUNDARY, 0, [&] {
                         if (kvpPred) kvpPred->pred_jacobi(nPh, true, c->stream, c->dm, b);
                         else k_pred_jacobi<true><<<gridB, TCIF_TPB, 0, c->stream>>>(c->dm, b);
                         c->launches++; });
                 });
        join(c);
        CUDA_OK(cudaMemcpyAsync(c->h_small, c->dmaxDev, sizeof(double) * 4, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
        double v[4]; int ops[4] = {RED_MAX, RED_MAX, RED_MAX, RED_MAX};
        for (int k = 0; k < nPh; ++k) v[k] = c->h_small[k];
        allreduce_host(c, v, ops, nPh);
        for (int k = 0; k < nPh; ++k) {
            if (!active[k]) continue;
            std::swap(c->alpha[phases[k]], c->alphaNew[phases[k]]);
            alpha_changed(c, phases[k]);
            c->predictorSweeps[phases[k]]++;
            if (v[k] <= tol) active[k] = false;
        }
    }
    correct_alpha_bcs(c, phases, nPh);
    if (addFlux) for (int k = 0; k < nPh; ++k) do_add_upwind(c, phases[k]);
}

void store_old(tcif_ctx *c) {
    for (int p = 0; p < c->P; ++p)
        CUDA_OK(cudaMemcpyAsync(c->alpha0[p], c->alpha[p], sizeof(double) * c->nC, cudaMemcpyDeviceToDevice, c->stream));
}

void do_mules_stage(tcif_ctx *c, double deltaT);
void do_stage(tcif_ctx *c, double deltaT) {
    if (!(deltaT > 0)) failf("tcif_stage: deltaT must be positive");
    do_force(c);
    do_mules_stage(c, deltaT);
}
// One solveAlphas corrector restricted to MULES (the metric stage): limit(returnCorr) for all phases, limitSum,
// explicitSolve for all phases.
void do_mules_stage(tcif_ctx *c, double deltaT) {
    int all[4] = {0, 1, 2, 3};
    const bool fuse = can_fuse_limit_sum(c);
    do_limit(c, all, c->P, 1.0 / deltaT, 0, false, true, fuse, true); // alpha.oldTime() := alpha, then multiphaseSystem.C:296-319
    if (fuse) do_limit_sum_rest(c); else do_limit_sum(c);           // :321
    do_explicit(c, all, c->P, deltaT, true);                        // :346 (bounded flux) + :365-375
}

// rhoPhi_ (+)= sum_p interpolate(rho_p)*alphaPhi_p   (multiphaseSystem.C:397-406, 158)
void do_rho_phi(tcif_ctx *c, double wgt, bool accumulate) {
    for (int p = 0; p < c->P; ++p) if (!c->rhoSet[p]) failf("phase %d has no density: call tcif_set_rho first", p);
    if (!c->rhoPhi) { c->rhoPhi = c->dalloc<double>(c->nFd); c->rhoPhib = c->dalloc<double>(std::max(c->nBF, 1)); }
    RhoPhiArgs a{};
    a.P = c->P; a.out = c->rhoPhi; a.outb = c->rhoPhib; a.wgt = wgt; a.accumulate = accumulate ? 1 : 0;
    for (int p = 0; p < c->P; ++p) {
        a.rhoU[p] = c->rhoU[p]; a.rho[p] = c->haveRho[p] ? c->rho[p] : nullptr; a.rhob[p] = c->haveRho[p] ? c->rhob[p] : nullptr;
        a.aPhi[p] = c->phiPsi[p]; a.aPhib[p] = c->phiPsib[p];
    }
    timed(c, T_RHOPHI, (double)c->nC * (4.0 * c->Ko + 8.0 * c->Ko * (2 + c->P)), [&] {
        k_rho_phi<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a);
        c->launches++;
        if (c->nBF) { k_rho_phi_bnd<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a); c->launches++; }
    });
}

// alpha += alpha*(1 - sumAlpha), cells and patch values   (multiphaseSystem.C:414-426)
void do_renormalise(tcif_ctx *c) {
    for (int p = 0; p < c->P; ++p) alpha_changed(c, p);
    RenormArgs r{};
    r.P = c->P;
    for (int p = 0; p < c->P; ++p) r.a[p] = c->alpha[p];
    timed(c, T_RHOPHI, 16.0 * c->P * c->nC, [&] {
        k_renormalise<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, r);
        c->launches++;
        if (c->nBF) {
            RenormArgs rb{};
            rb.P = c->P;
            for (int p = 0; p < c->P; ++p) rb.a[p] = c->alphab[p];
            k_renormalise<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nBF, rb);
            c->launches++;
        }
    });
}

// MassTransferPhaseSystem::correctMassSources(T) (:447-497): dmdt_[key] = Lee::Kexp(T) for every model
void do_correct_mass_sources(tcif_ctx *c) {
    if (!c->haveT) failf("tcif_correct_mass_sources: no temperature field (tcif_set_temperature)");
    for (auto &mm : c->massModels) {
        if (!c->rhoSet[mm.from]) failf("tcif_correct_mass_sources: phase %d has no density (tcif_set_rho)", mm.from);
        if (!mm.dmdt) mm.dmdt = c->dalloc<double>(c->stride);
        timed(c, T_MASS, 32.0 * c->nC, [&] {
            k_lee_kexp<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, mm.C, mm.Tact, mm.aMin, c->alpha[mm.from],
                c->haveRho[mm.from] ? c->rho[mm.from] : nullptr, c->rhoU[mm.from], c->Tc, mm.dmdt);
            c->launches++;
        });
        mm.haveDmdt = true;
    }
}

// Su_/Sp_ := 0; alphaTransfer(Su_, Sp_)   (multiphaseSystem.C:282-292, MassTransferPhaseSystem.C:503-704)
void do_alpha_transfer(tcif_ctx *c) {
    require_comm(c);
    MassArgs a{};
    a.P = c->P;
    std::vector<int> pairs = c->massPairs;
    if (pairs.empty()) for (int i = 0; i < c->P; ++i) for (int j = i + 1; j < c->P; ++j) { pairs.push_back(i); pairs.push_back(j); }
    a.nPairs = (int)pairs.size() / 2;
    if (a.nPairs > TCIF_MAX_PAIRS) failf("tcif_alpha_transfer: more than %d phase pairs", TCIF_MAX_PAIRS);
    a.nModels = (int)c->massModels.size();
    for (int i = 0; i < a.nModels; ++i) {
        const auto &mm = c->massModels[i];
        a.mFrom[i] = mm.from; a.mVar[i] = mm.var; a.mC[i] = mm.C; a.mTact[i] = mm.Tact; a.mAmin[i] = mm.aMin;
        a.dmdt[i] = mm.haveDmdt ? mm.dmdt : nullptr;
    }
    int includeDivU = 1;                                                    // :512
    for (int q = 0; q < a.nPairs; ++q) {
        int p1 = pairs[2 * q], p2 = pairs[2 * q + 1];
        if (p1 < 0 || p1 >= c->P || p2 < 0 || p2 >= c->P || p1 == p2) failf("tcif_alpha_transfer: bad phase pair (%d,%d)", p1, p2);
        a.pairs[2 * q] = p1; a.pairs[2 * q + 1] = p2; a.m12[q] = -1; a.m21[q] = -1;
        for (int i = 0; i < a.nModels; ++i) {
            if (c->massModels[i].from == p1 && c->massModels[i].to == p2) a.m12[q] = i;
            if (c->massModels[i].from == p2 && c->massModels[i].to == p1) a.m21[q] = i;
        }
        if (a.m12[q] >= 0) includeDivU = c->massModels[a.m12[q]].divU;    // :555
        if (a.m21[q] >= 0) includeDivU = c->massModels[a.m21[q]].divU;    // :583
        a.divU[q] = includeDivU;
        a.anyDivU |= includeDivU;
    }
    for (int p = 0; p < c->P; ++p) {
        if (!c->rhoSet[p]) failf("tcif_alpha_transfer: phase %d has no density (tcif_set_rho)", p);
        a.alpha[p] = c->alpha[p]; a.rhoU[p] = c->rhoU[p]; a.rho[p] = c->haveRho[p] ? c->rho[p] : nullptr;
        if (!c->Su[p]) c->Su[p] = c->dalloc<double>(c->stride);
        if (!c->Sp[p]) c->Sp[p] = c->dalloc<double>(c->stride);
        a.Su[p] = c->Su[p]; a.Sp[p] = c->Sp[p];
        c->haveSu[p] = c->haveSp[p] = true;
    }
    a.phi = c->phi; a.phib = c->phib;
    if (!c->ddtMaxDev) c->ddtMaxDev = c->dalloc<long long>(2);
    const long long init[2] = {dbl_ord(-1e300), dbl_ord(-1e300)};
    CUDA_OK(cudaMemcpyAsync(c->ddtMaxDev, init, sizeof init, cudaMemcpyHostToDevice, c->stream));
    a.ddtMax = c->ddtMaxDev;
    double bytes = (double)c->nC * (4.0 * (c->Ko + c->Kn) + 8.0 * c->Ko + 8 + 8.0 * c->P + 8.0 * a.nModels + 16.0 * c->P);
    timed(c, T_MASS, bytes, [&] {
        k_alpha_transfer<false><<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a);
        c->launches++;
        if (c->nBndCells) {
            k_alpha_transfer<true><<<cdiv(c->nBndCells, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a);
            c->launches++;
        }
    });
    long long h[2];
    CUDA_OK(cudaMemcpyAsync(h, c->ddtMaxDev, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    double v[2] = {ord_dbl(h[0]), ord_dbl(h[1])};
    const int ops[2] = {RED_MAX, RED_MAX};
    allreduce_host(c, v, ops, 2);                                           // gMax
    c->ddtAlphaMax = std::max(v[0], v[1]);
}

void check_phase(tcif_ctx *c, int p);
void up_cells(tcif_ctx *c, double *dst, const double *src);
void up_bnd(tcif_ctx *c, double *dst, const double *src);

// MultiComponentPhaseModel::solveYi (MultiComponentPhaseModel.C:182-388) for the species of phase `ph`.
// Host buffers in/out: Yint[i] [nCells], Ybnd[i] [nBF]; the inert species receives max(1 - sum Yi, 0).
// The species' working set (Yi, Yi.oldTime, patch values, BC kinds, Su/Sp, flux) is swapped into phase slot `ph`
// so that the limiter / predictor / explicitSolve sweeps of the phase fractions run on it unchanged.
void do_solve_yi(tcif_ctx *c, int ph, int nSpecies, int inertIndex, double *const *Yint, double *const *Ybnd,
                 const int32_t *Ybc, const double *const *Su, const double *const *Sp, double cYi,
                 int schemePhi, int schemePhir, int nYiSubCycles, double deltaT, double tol, int maxIter) {
    check_phase(c, ph);
    if (nSpecies < 1 || !Yint || (c->nBF && !Ybnd) || !Ybc) failf("tcif_solve_yi: bad arguments");
    if (!(deltaT > 0)) failf("tcif_solve_yi: deltaT must be positive");
    for (int sc : {schemePhi, schemePhir}) {
        if (sc < 0 || sc > 2) failf("unsupported convection scheme %d (upwind | linear | vanLeer)", sc);
        if (sc == SCHEME_VANLEER && !c->Cx) failf("vanLeer needs cell centres: call tcif_set_cell_centres first");
    }
    require_comm(c);
    ensure_deltaN(c);
    const size_t cs = c->stride, bs = std::max(c->nBF, 1);
    if (!c->GA) {
        c->GA = c->dalloc<double>((size_t)c->P * 3 * cs); c->GAb = c->dalloc<double>((size_t)c->P * 3 * bs);
        c->spY = c->dalloc<double>(cs); c->spY0 = c->dalloc<double>(cs); c->spYNew = c->dalloc<double>(cs);
        c->spYb = c->dalloc<double>(bs); c->spSu = c->dalloc<double>(cs); c->spSp = c->dalloc<double>(cs);
        c->spPhiPsi = c->dalloc<double>(c->nFd); c->spPhiPsib = c->dalloc<double>(bs);
        c->spPhiHO = c->dalloc<double>(c->nFd); c->spPhiHOb = c->dalloc<double>(bs);
        c->spBc = c->dalloc<int>(std::max(c->nPatches, 1));
    }
    // grad(alpha_p) of the phases (nHatf, vanLeer of alpha2), kept aside: G is re-used for grad(Yi)
    do_grad(c);
    CUDA_OK(cudaMemcpyAsync(c->GA, c->G, sizeof(double) * c->P * 3 * cs, cudaMemcpyDeviceToDevice, c->stream));
    CUDA_OK(cudaMemcpyAsync(c->GAb, c->Gb, sizeof(double) * c->P * 3 * bs, cudaMemcpyDeviceToDevice, c->stream));
    // max(phic) over all faces (:199-201)
    CUDA_OK(cudaMemsetAsync(c->dmaxDev, 0, sizeof(double) * 8, c->stream));
    {
        long long n = (long long)c->nFd + c->nBF;
        k_phic_max<<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->phi, c->phib, c->dmaxDev, 1);
        c->launches++;
    }
    CUDA_OK(cudaMemcpyAsync(c->h_small, c->dmaxDev, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    double mx[1] = {c->h_small[0]};
    { int ops[1] = {RED_MAX}; allreduce_host(c, mx, ops, 1); }

    // the phase's own state, restored at the end
    struct Saved { double *alpha, *alpha0, *alphaNew, *alphab, *Su, *Sp, *phiPsi, *phiPsib, *phiHO, *phiHOb; int *bc; bool haveSu, haveSp; } sv =
        {c->alpha[ph], c->alpha0[ph], c->alphaNew[ph], c->alphab[ph], c->Su[ph], c->Sp[ph], c->phiPsi[ph], c->phiPsib[ph],
         c->phiHO[ph], c->phiHOb[ph], c->bc[ph], c->haveSu[ph], c->haveSp[ph]};
    const double *alphaPh = sv.alpha, *alphabPh = sv.alphab;
    std::vector<double> Yt((size_t)c->nC, 0.0), Ytb((size_t)c->nBF, 0.0), tmp((size_t)std::max(c->nC, c->nBF));
    double *yCur = c->spY, *yNew = c->spYNew;
    auto restore = [&] {
        c->alpha[ph] = sv.alpha; c->alpha0[ph] = sv.alpha0; c->alphaNew[ph] = sv.alphaNew; c->alphab[ph] = sv.alphab;
        c->Su[ph] = sv.Su; c->Sp[ph] = sv.Sp; c->phiPsi[ph] = sv.phiPsi; c->phiPsib[ph] = sv.phiPsib; c->bc[ph] = sv.bc;
        c->phiHO[ph] = sv.phiHO; c->phiHOb[ph] = sv.phiHOb;
        c->haveSu[ph] = sv.haveSu; c->haveSp[ph] = sv.haveSp;
        touch(c);
    };
    try {
        for (int i = 0; i < nSpecies; ++i) {
            if (i == inertIndex) continue;
            if (!Yint[i] || (c->nBF && !Ybnd[i])) failf("tcif_solve_yi: species %d has no field", i);
            up_cells(c, yCur, Yint[i]);
            if (c->nBF) up_bnd(c, c->spYb, Ybnd[i]);
            CUDA_OK(cudaMemcpyAsync(c->spBc, Ybc + (size_t)i * c->nPatches, sizeof(int) * c->nPatches, cudaMemcpyHostToDevice, c->stream));
            const bool hsu = Su && Su[i], hsp = Sp && Sp[i];
            if (hsu) up_cells(c, c->spSu, Su[i]);
            if (hsp) up_cells(c, c->spSp, Sp[i]);
            c->alpha[ph] = yCur; c->alpha0[ph] = c->spY0; c->alphaNew[ph] = yNew; c->alphab[ph] = c->spYb;
            c->Su[ph] = c->spSu; c->Sp[ph] = c->spSp; c->haveSu[ph] = hsu; c->haveSp[ph] = hsp;
            c->phiPsi[ph] = c->spPhiPsi; c->phiPsib[ph] = c->spPhiPsib; c->bc[ph] = c->spBc;
            c->phiHO[ph] = c->spPhiHO; c->phiHOb[ph] = c->spPhiHOb;
            correct_alpha_bcs(c, &ph, 1);
            CUDA_OK(cudaMemcpyAsync(c->spY0, yCur, sizeof(double) * c->nC, cudaMemcpyDeviceToDevice, c->stream));   // Yi.oldTime()
            do_grad(c);                                   // slot ph of G/Gb now holds grad(Yi)
            YiFluxArgs a{};
            a.P = c->P; a.ph = ph; a.schemePhi = schemePhi; a.schemePhir = schemePhir;
            for (int p = 0; p < c->P; ++p) { a.alpha[p] = p == ph ? alphaPh : c->alpha[p]; a.alphab[p] = p == ph ? alphabPh : c->alphab[p]; }
            a.GA = c->GA; a.GAb = c->GAb; a.Y = yCur; a.Yb = c->spYb;
            a.GY = c->G + (size_t)ph * 3 * cs; a.GYb = c->Gb + (size_t)ph * 3 * c->nBF;
            a.Cx = c->Cx; a.Cy = c->Cy; a.Cz = c->Cz; a.dbx = c->dbx; a.dby = c->dby; a.dbz = c->dbz;
            a.phi = c->phi; a.phib = c->phib; a.deltaN = c->deltaN; a.cYi = cYi; a.maxPhic = mx[0];
            a.out = c->phiHO[ph]; a.outb = c->phiHOb[ph];
            timed(c, T_FLUX_ASSEMBLY, 0, [&] {
                DISPATCH_P(k_flux_yi<PP><<<cdiv(c->nC, 128), 128, 0, c->stream>>>(c->dm, a));
                c->launches++;
                if (c->nBF) {
                    DISPATCH_P(k_flux_yi_bnd<PP><<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a));
                    c->launches++;
                }
            });
            do_limit(c, &ph, 1, 1.0 / deltaT, 0, false, true);                        // :299-311
            do_predictor(c, &ph, 1, deltaT, tol, maxIter, false);                      // :323-338
            if (nYiSubCycles > 1) {
                // subCycle<volScalarField> (:345-364): the first sub-cycle keeps the real old-time field (the
                // constructor advances the field's time index), from the second one on Yi.oldTime() is re-stored
                // from the current Yi; deltaT/n; the flux left in place by the previous explicitSolve
                for (int k = 0; k < nYiSubCycles; ++k) {
                    if (k) CUDA_OK(cudaMemcpyAsync(c->spY0, c->alpha[ph], sizeof(double) * c->nC, cudaMemcpyDeviceToDevice, c->stream));
                    do_explicit(c, &ph, 1, deltaT / nYiSubCycles, k == 0);
                }
            } else
                do_explicit(c, &ph, 1, deltaT, true);                                  // :343 (+= YiEqn.flux()) and :361-378
            yCur = c->alpha[ph]; yNew = c->alphaNew[ph];                               // explicitSolve swapped the buffers
            CUDA_OK(cudaMemcpyAsync(Yint[i], yCur, sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
            if (c->nBF) CUDA_OK(cudaMemcpyAsync(Ybnd[i], c->spYb, sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
            CUDA_OK(cudaStreamSynchronize(c->stream));
            for (int k = 0; k < c->nC; ++k) Yt[k] += Yint[i][k];                       // Yt += Yi (:380)
            for (int b = 0; b < c->nBF; ++b) Ytb[b] += Ybnd[i][b];
        }
    } catch (...) { restore(); throw; }
    restore();
    c->spY = yCur; c->spYNew = yNew;
    if (inertIndex >= 0 && inertIndex < nSpecies) {                                    // :383-384
        if (!Yint[inertIndex] || (c->nBF && !Ybnd[inertIndex])) failf("tcif_solve_yi: inert species has no field");
        for (int k = 0; k < c->nC; ++k) Yint[inertIndex][k] = std::max(1.0 - Yt[k], 0.0);
        for (int b = 0; b < c->nBF; ++b) Ybnd[inertIndex][b] = std::max(1.0 - Ytb[b], 0.0);
    }
}

// multiphaseSystem::solve() (multiphaseSystem.C:128-168) = nAlphaSubCycles x solveAlphas() (:170-429)
void do_solve_alphas(tcif_ctx *c, const tcif_alpha_controls *ctl, double deltaT) {
    if (!ctl) failf("tcif_solve_alphas: null controls");
    if (!(deltaT > 0)) failf("tcif_solve_alphas: deltaT must be positive");
    if (ctl->nAlphaCorr < 1) failf("tcif_solve_alphas: nAlphaCorr must be >= 1");
    int all[4] = {0, 1, 2, 3};
    const int nSub = std::max(1, ctl->nAlphaSubCycles);
    const double dt = deltaT / nSub;
    for (int sub = 0; sub < nSub; ++sub) {
        store_old(c);                                                     // subCycle refreshes oldTime()
        PhicState st;
        phic_begin(c, st);
        for (int acorr = 0; acorr < ctl->nAlphaCorr; ++acorr) {
            do_flux_assembly(c, ctl->schemePhi, ctl->schemePhir, st);     // :199-277
            if (c->massInSolve) do_alpha_transfer(c);                     // :280-292 calculateSuSp()
            const bool fuse = can_fuse_limit_sum(c);
            do_limit(c, all, c->P, 1.0 / dt, 0, false, true, fuse);       // :296-319
            if (fuse) do_limit_sum_rest(c); else do_limit_sum(c);         // :321
            if (ctl->implicitPredictor) {
                do_predictor(c, all, c->P, dt, ctl->predictorTol, ctl->predictorMaxIter, false);   // :348-361
                do_explicit(c, all, c->P, dt, true);                      // :363 (+= alpha1Eqn.flux()) and :365-375
            } else {
                do_explicit(c, all, c->P, dt, true);                      // :346 (the commented-out bounded flux)
            }
            if (acorr == ctl->nAlphaCorr - 1) {                           // :382-427
                do_rho_phi(c, dt / deltaT, nSub > 1 && sub > 0);
                do_renormalise(c);
            }
        }
    }
}

// ---- host <-> device field transfers ----------------------------------------------------
void up_cells(tcif_ctx *c, double *dst, const double *src) {
    CUDA_OK(cudaMemcpyAsync(dst, src, sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
}
void up_bnd(tcif_ctx *c, double *dst, const double *src) {
    if (c->nBF && src) CUDA_OK(cudaMemcpyAsync(dst, src, sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
}
void up_faces(tcif_ctx *c, double *dst, const double *src) {
    if (!c->nIF) return;
    CUDA_OK(cudaMemcpyAsync(c->staging, src, sizeof(double) * c->nIF, cudaMemcpyDefault, c->stream));
    timed(c, T_CONVERT, 20.0 * c->nIF, [&] {
        k_face_scatter<<<cdiv(c->nIF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nIF, c->d_faceMap, c->staging, dst);
        c->launches++;
    });
}
void down_faces(tcif_ctx *c, double *dst, const double *src) {
    if (!c->nIF) return;
    timed(c, T_CONVERT, 20.0 * c->nIF, [&] {
        k_face_gather<<<cdiv(c->nIF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nIF, c->d_faceMap, src, c->staging);
        c->launches++;
    });
    CUDA_OK(cudaMemcpyAsync(dst, c->staging, sizeof(double) * c->nIF, cudaMemcpyDefault, c->stream));
}
// SoA (3 comps, given strides) -> host AoS
void down_vec(tcif_ctx *c, double *dst, const double *src, int n, size_t strideSrc) {
    if (!n || !dst) return;
    timed(c, T_CONVERT, 48.0 * n, [&] {
        k_soa_to_aos3<<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(n, strideSrc, src, c->staging);
        c->launches++;
    });
    CUDA_OK(cudaMemcpyAsync(dst, c->staging, sizeof(double) * 3 * (size_t)n, cudaMemcpyDefault, c->stream));
}

void check_phase(tcif_ctx *c, int p) { if (p < 0 || p >= c->P) failf("phase index %d out of range [0,%d)", p, c->P); }
int pair_index(int P, int p1, int p2) {
    int q = 0;
    for (int a = 0; a < P; ++a) for (int b = a + 1; b < P; ++b, ++q) if (a == p1 && b == p2) return q;
    return -1;
}

// ---- laserDTRM, kernels in tcif_laser.cuh ------------------------------------------------
// polyMesh data of the cell walk: face centres in the slot layout, mesh.points()/faces()/tetBasePtIs and the
// point -> cell weights of volPointInterpolation (built once on the host from the faces' vertices).
void do_laser_set_polymesh(tcif_ctx *c, const double *Cf, const double *Cfb, int nPoints, const double *points,
                           const int32_t *fpStart, const int32_t *fpPts, const int32_t *tetBasePt, double boundsMag) {
    const int nIF = c->nIF, nBF = c->nBF, nC = c->nC, nF = nIF + nBF;
    if ((nIF && !Cf) || (nBF && !Cfb) || nPoints <= 0 || !points || !fpStart || !fpPts || !tetBasePt || !(boundsMag > 0))
        failf("tcif_laser_set_polymesh: bad arguments");
    if (!c->Cx) failf("tcif_laser_set_polymesh: cell centres needed, call tcif_set_cell_centres first");
    // 2-D cases: the reference subtracts particle::deviationFromMeshCentre() from every displacement (DTRMParticle.C:110)
    // and the cells are open in the empty direction here (empty patches carry no faces): not supported, not guessed
    for (int pa = 0; pa < c->nPatches; ++pa)
        if (c->patchType[pa] == TCIF_PATCH_EMPTY || c->patchType[pa] == TCIF_PATCH_WEDGE)
            failf("tcif_laser_set_polymesh: laserDTRM on meshes with empty or wedge patches (fewer than 3 solution directions) is not supported");
    if (fpStart[0] != 0) failf("tcif_laser_set_polymesh: fpStart[0] must be 0");
    for (int f = 0; f < nF; ++f) {
        if (fpStart[f + 1] - fpStart[f] < 3) failf("tcif_laser_set_polymesh: face %d has fewer than 3 vertices", f);
        if (tetBasePt[f] >= fpStart[f + 1] - fpStart[f]) failf("tcif_laser_set_polymesh: tetBasePt[%d] out of range", f);
    }
    const int nFP = fpStart[nF];
    for (int j = 0; j < nFP; ++j) if (fpPts[j] < 0 || fpPts[j] >= nPoints) failf("tcif_laser_set_polymesh: vertex label out of range");
    LaserGeo &g = c->lg;
    g = LaserGeo{};
    const size_t bs = std::max(nBF, 1);
    // face centres: internal faces scattered into the slots of Sf, boundary faces as SoA
    {
        double *cf = c->dalloc<double>(3 * c->nFd);
        g.Cfx = cf; g.Cfy = cf + c->nFd; g.Cfz = cf + 2 * c->nFd;
        double *cb = c->dalloc<double>(3 * bs);
        g.Cfbx = cb; g.Cfby = cb + bs; g.Cfbz = cb + 2 * bs;
        double *tmp = nullptr;
        CUDA_OK(cudaMalloc(&tmp, sizeof(double) * 3 * (size_t)std::max(std::max(nIF, nBF), 1)));
        if (nIF) {
            CUDA_OK(cudaMemcpyAsync(tmp, Cf, sizeof(double) * 3 * nIF, cudaMemcpyDefault, c->stream));
            k_face_scatter_vec<<<cdiv(nIF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(nIF, c->d_faceMap, tmp, cf, cf + c->nFd, cf + 2 * c->nFd);
            CUDA_OK(cudaStreamSynchronize(c->stream));
        }
        if (nBF) {
            CUDA_OK(cudaMemcpyAsync(tmp, Cfb, sizeof(double) * 3 * nBF, cudaMemcpyDefault, c->stream));
            k_aos3_to_soa<<<cdiv(nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(nBF, bs, tmp, cb);
            CUDA_OK(cudaStreamSynchronize(c->stream));
        }
        CUDA_OK(cudaFree(tmp));
        c->launches += 2;
    }
    g.Cx = c->Cx; g.Cy = c->Cy; g.Cz = c->Cz;
    {
        int *bi = c->dalloc<int>(nC, false);
        CUDA_OK(cudaMemsetAsync(bi, 0xff, sizeof(int) * nC, c->stream));
        if (c->nBndCells) k_laser_bnd_index<<<cdiv(c->nBndCells, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nBndCells, c->dm.bndCells, bi);
        int *ff = c->dalloc<int>(c->nFd, false);
        CUDA_OK(cudaMemsetAsync(ff, 0xff, sizeof(int) * c->nFd, c->stream));
        if (nIF) k_laser_fd_to_face<<<cdiv(nIF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(nIF, c->d_faceMap, ff);
        g.bndIdx = bi; g.fdToFace = ff;
        c->launches += 2;
    }
    g.nIF = nIF; g.nPoints = nPoints; g.boundsMag = boundsMag;
    g.points = c->upload(std::vector<double>(points, points + 3 * (size_t)nPoints));
    g.fpStart = c->upload(std::vector<int>(fpStart, fpStart + nF + 1));
    g.fpPts = c->upload(std::vector<int>(fpPts, fpPts + nFP));
    g.tetBasePt = c->upload(std::vector<int>(tetBasePt, tetBasePt + nF));
    // owner / neighbour / faceCells back from the device tables
    std::vector<int> faceMap(nIF), nbrOwn(c->nFd), fc(nBF);
    if (nIF) CUDA_OK(cudaMemcpy(faceMap.data(), c->d_faceMap, sizeof(int) * nIF, cudaMemcpyDeviceToHost));
    CUDA_OK(cudaMemcpy(nbrOwn.data(), c->dm.nbrOwn, sizeof(int) * c->nFd, cudaMemcpyDeviceToHost));
    if (nBF) CUDA_OK(cudaMemcpy(fc.data(), c->dm.faceCells, sizeof(int) * nBF, cudaMemcpyDeviceToHost));
    std::vector<double> Cc(3 * (size_t)nC);
    {
        std::vector<double> soa(3 * (size_t)c->stride);
        CUDA_OK(cudaMemcpy(soa.data(), c->Cx, sizeof(double) * 3 * c->stride, cudaMemcpyDeviceToHost));
        for (int i = 0; i < nC; ++i) for (int d = 0; d < 3; ++d) Cc[3 * (size_t)i + d] = soa[(size_t)d * c->stride + i];
    }
    // point -> cells (ascending, unique) with weights (1/|x_p - C_c|)/sum   [volPointInterpolation::makeWeights]
    std::vector<int> cnt(nPoints + 1, 0);
    for (int f = 0; f < nF; ++f)
        for (int j = fpStart[f]; j < fpStart[f + 1]; ++j) cnt[fpPts[j] + 1] += f < nIF ? 2 : 1;
    for (int p = 0; p < nPoints; ++p) cnt[p + 1] += cnt[p];
    std::vector<int> tmp(cnt[nPoints]), fill(cnt.begin(), cnt.end() - 1);
    for (int f = 0; f < nF; ++f)
        for (int j = fpStart[f]; j < fpStart[f + 1]; ++j) {
            const int p = fpPts[j];
            if (f < nIF) { tmp[fill[p]++] = faceMap[f] % c->stride; tmp[fill[p]++] = nbrOwn[faceMap[f]]; }
            else tmp[fill[p]++] = fc[f - nIF];
        }
    std::vector<int> pcStart(nPoints + 1), pc;
    c->lzPwRaw.clear();
    std::vector<double> pw;
    pc.reserve(tmp.size() / 2); pw.reserve(tmp.size() / 2);
    for (int p = 0; p < nPoints; ++p) {
        pcStart[p] = (int)pc.size();
        std::sort(tmp.begin() + cnt[p], tmp.begin() + cnt[p + 1]);
        double sum = 0.0;
        const size_t first = pc.size();
        for (int i = cnt[p]; i < cnt[p + 1]; ++i) {
            if (i > cnt[p] && tmp[i] == tmp[i - 1]) continue;
            const double *C = &Cc[3 * (size_t)tmp[i]], *x = points + 3 * (size_t)p;
            const double d[3] = {x[0] - C[0], x[1] - C[1], x[2] - C[2]};
            const double w = 1.0 / std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            pc.push_back(tmp[i]); pw.push_back(w);
            sum += w;
        }
        for (size_t i = first; i < pc.size(); ++i) pw[i] /= sum;
        c->lzPwRaw.push_back(sum);           // local sum of the raw weights of point p
    }
    pcStart[nPoints] = (int)pc.size();
    std::vector<unsigned char> isPatch(nPoints, 0);
    for (int pa = 0; pa < c->nPatches; ++pa) {
        if (c->patchType[pa] == TCIF_PATCH_EMPTY || c->patchType[pa] == TCIF_PATCH_PROCESSOR) continue;
        for (int i = 0; i < c->patchSize[pa]; ++i) {
            const int f = nIF + c->patchStart[pa] + i;
            for (int j = fpStart[f]; j < fpStart[f + 1]; ++j) isPatch[fpPts[j]] = 1;
        }
    }
    g.pcStart = c->upload(pcStart); g.pc = c->upload(pc);
    c->lzPw = c->upload(pw); c->lzIsPatchDev = c->upload(isPatch);
    g.pw = c->lzPw; g.isPatchPoint = c->lzIsPatchDev;
    c->lzPcStart = pcStart; c->lzIsPatch = isPatch;
    {
        std::vector<int> info(2 * (size_t)std::max(c->nPatches, 1), -1);
        for (int pa = 0; pa < c->nPatches; ++pa) info[2 * pa + 1] = c->patchStart[pa];
        for (size_t i = 0; i < c->cpPatch.size(); ++i) info[2 * c->cpPatch[i]] = (int)i;
        c->lzPatchInfo = c->upload(info);
    }
    c->lzPwDenom = c->lzPwRaw;
    c->lzSharedSet = false;
    // walk table: one contiguous record of all face planes per cell (TCIF_LASER_TABLE=0 keeps the ELL gathers)
    {
        std::vector<int> nf(nC, 0);
        for (int f = 0; f < nIF; ++f) { nf[faceMap[f] % c->stride]++; nf[nbrOwn[faceMap[f]]]++; }
        for (int b = 0; b < nBF; ++b) nf[fc[b]]++;
        int W = 0;
        for (int i = 0; i < nC; ++i) W = std::max(W, nf[i]);
        const char *e = getenv("TCIF_LASER_TABLE");
        if (W > 0 && W <= 8 && nC < 0x40000000 && !(e && atoi(e) == 0)) {
            double *wd = c->dalloc<double>((size_t)nC * W * 6, false);
            int *wi = c->dalloc<int>((size_t)nC * W * 2, false);
            k_laser_build_walk<<<cdiv(nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, g, W, wd, wi);
            c->launches++;
            CUDA_OK(cudaStreamSynchronize(c->stream));
            g.W = W; g.wd = wd; g.wi = wi;
        }
    }
    c->lzRefl = c->dalloc<int>(nC);
    c->lzNHat = c->dalloc<double>(3 * (size_t)c->stride);
    c->lzNHatPts = c->dalloc<double>(3 * (size_t)nPoints);
    c->lzQ = c->dalloc<double>(nC);
    c->lzCounters = c->dalloc<int>(4 + 64);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->laserGeoSet = true;
}

// Points on processor patches are shared by ranks: volPointInterpolation sums weights and weighted values over
// all ranks holding the point (syncUntransformedData with plusEqOp).  The solver hands over the global label of
// every such local point (globalMeshData); collective over the ranks.
void do_laser_set_shared_points(tcif_ctx *c, int nShared, const int32_t *localPoint, const int64_t *globalPoint) {
    if (!c->laserGeoSet) failf("tcif_laser_set_shared_points: call tcif_laser_set_polymesh first");
    if (nShared < 0 || (nShared && (!localPoint || !globalPoint))) failf("tcif_laser_set_shared_points: bad arguments");
    require_comm(c);
    const int nPoints = c->lg.nPoints;
    for (int i = 0; i < nShared; ++i) if (localPoint[i] < 0 || localPoint[i] >= nPoints) failf("tcif_laser_set_shared_points: local point out of range");
    double mx[1] = {(double)nShared};
    { int ops[1] = {RED_MAX}; allreduce_host(c, mx, ops, 1); }
    const int maxS = (int)mx[0];
    c->lzNShared = nShared; c->lzMaxShared = maxS;
    c->lzSharedSrc.assign(nShared, {});
    if (maxS > 0) {
        // (global label, local sum of raw weights, on-a-patch flag) of every shared point of every rank
        std::vector<double> mine(3 * (size_t)maxS, -1.0), all;
        for (int i = 0; i < nShared; ++i) {
            mine[3 * i] = (double)globalPoint[i];
            mine[3 * i + 1] = c->lzPwRaw[localPoint[i]];
            mine[3 * i + 2] = c->lzIsPatch[localPoint[i]] ? 1.0 : 0.0;
        }
        allgather_host(c, mine.data(), 3 * maxS, all);
        std::unordered_map<long long, int> idx;
        for (int i = 0; i < nShared; ++i) idx[(long long)globalPoint[i]] = i;
        std::vector<double> sumW(nShared, 0.0);
        std::vector<int> flag(nShared, 0);
        for (int r = 0; r < c->nRanks; ++r)
            for (int j = 0; j < maxS; ++j) {
                const double id = all[((size_t)r * maxS + j) * 3];
                if (id < 0) continue;
                auto it = idx.find((long long)id);
                if (it == idx.end()) continue;
                c->lzSharedSrc[it->second].push_back(r * maxS + j);
                sumW[it->second] += all[((size_t)r * maxS + j) * 3 + 1];
                if (all[((size_t)r * maxS + j) * 3 + 2] > 0) flag[it->second] = 1;
            }
        // renormalise the weights of the shared points by the global sum, OR the patch flags
        std::vector<double> pw(c->lzPcStart[nPoints]);
        CUDA_OK(cudaMemcpy(pw.data(), c->lzPw, sizeof(double) * pw.size(), cudaMemcpyDeviceToHost));
        for (int i = 0; i < nShared; ++i) {
            const int p = localPoint[i];
            const double f = c->lzPwDenom[p] / sumW[i];      // idempotent: a second call finds the weights already global
            for (int k = c->lzPcStart[p]; k < c->lzPcStart[p + 1]; ++k) pw[k] *= f;
            c->lzPwDenom[p] = sumW[i];
            if (flag[i]) c->lzIsPatch[p] = 1;
        }
        CUDA_OK(cudaMemcpy(c->lzPw, pw.data(), sizeof(double) * pw.size(), cudaMemcpyHostToDevice));
        CUDA_OK(cudaMemcpy(c->lzIsPatchDev, c->lzIsPatch.data(), nPoints, cudaMemcpyHostToDevice));
        c->lzSharedPts = c->upload(std::vector<int>(localPoint, localPoint + nShared));
        c->lzSharedBuf = c->dalloc<double>(3 * (size_t)std::max(nShared, 1));
    }
    c->lzSharedSet = true;
}

// sum the partial point values of the shared points over the ranks, in rank order on every rank
void laser_sync_points(tcif_ctx *c) {
    const int nS = c->lzNShared, maxS = c->lzMaxShared;
    if (maxS == 0) return;
    std::vector<double> mine(3 * (size_t)maxS, 0.0), all;
    if (nS) {
        k_laser_points_gather<<<cdiv(nS, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(nS, c->lzSharedPts, c->lzNHatPts, c->lzSharedBuf);
        c->launches++;
        CUDA_OK(cudaMemcpyAsync(mine.data(), c->lzSharedBuf, sizeof(double) * 3 * nS, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
    }
    allgather_host(c, mine.data(), 3 * maxS, all);
    if (!nS) return;
    for (int i = 0; i < nS; ++i)
        for (int d = 0; d < 3; ++d) {
            double v = 0.0;
            for (int src : c->lzSharedSrc[i]) v += all[(size_t)src * 3 + d];
            mine[3 * i + d] = v;
        }
    CUDA_OK(cudaMemcpyAsync(c->lzSharedBuf, mine.data(), sizeof(double) * 3 * nS, cudaMemcpyHostToDevice, c->stream));
    k_laser_points_scatter<<<cdiv(nS, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(nS, c->lzSharedPts, c->lzSharedBuf, c->lzNHatPts);
    c->launches++;
    CUDA_OK(cudaStreamSynchronize(c->stream));
}

// laserDTRM::calculate(): mark the reflecting cells, interpolate nHat to the points, walk the cloud generation by
// generation (reflected particles are appended by the walk, particles reaching a processor patch are handed to the
// neighbour rank between generations), Q /= V
int do_laser_calculate(tcif_ctx *c, int nParticles, const double *p0, const double *p1, const double *I, const double *dA,
                       const int32_t *cell, double sigmaSB, int maxParticles) {
    // On a decomposed mesh every check below is made by all ranks and agreed on BEFORE anybody throws: a rank-local
    // throw would leave the others waiting in the next collective.
    const int nCp = (int)c->cpPatch.size();
    const bool par = c->nCoupled > 0;
    std::string argErr;
    if (!c->laserGeoSet) argErr = "call tcif_laser_set_polymesh first";
    else if (!c->haveT) argErr = "temperature not set";
    else if (!c->lzAbsFields && c->lzNAbs == 0) argErr = "no absorption model set";
    else if (nParticles < 0 || (nParticles && (!p0 || !p1 || !I || !dA || !cell))) argErr = "bad arguments";
    else if (par && !c->lzSharedSet) argErr = "mesh has processor patches, call tcif_laser_set_shared_points first";
    else if (par && nCp > 64) argErr = "more than 64 processor patches";
    int nSeeds = 0;
    if (argErr.empty())
        for (int k = 0; k < nParticles; ++k) {
            if (cell[k] < 0) continue;
            if (cell[k] >= c->nC) { argErr = "cell[" + std::to_string(k) + "] out of range"; break; }
            ++nSeeds;
        }
    // seeds of the whole cloud (each seed is owned by the one rank that found its cell, laserDTRM.C:287-296) and the
    // largest maxParticles: cloud and mailbox capacities must be the same on every rank, because a rank with few or no
    // seeds of its own can receive the whole cloud from its neighbours
    long long globalSeeds = nSeeds;
    int maxPartAll = maxParticles;
    if (par) {
        require_comm(c);
        double v[3] = {argErr.empty() ? 0.0 : 1.0, (double)nSeeds, (double)maxParticles};
        const int ops[3] = {RED_MAX, RED_SUM, RED_MAX};
        allreduce_host(c, v, ops, 3);
        if (v[0] != 0.0) failf("tcif_laser_calculate: %s", argErr.empty() ? "another rank rejected its arguments" : argErr.c_str());
        globalSeeds = (long long)v[1]; maxPartAll = (int)v[2];
    } else if (!argErr.empty()) failf("tcif_laser_calculate: %s", argErr.c_str());
    if (c->lzDeltaN == 0.0) {
        double v[2] = {c->sumV, (double)c->nC};
        int ops[2] = {RED_SUM, RED_SUM};
        if (c->nRanks > 1) allreduce_host(c, v, ops, 2);
        c->lzDeltaN = 1e-7 / cbrt(v[0] / v[1]);                                        // laserDTRM.C:121-125
    }
    LaserArgs a{};
    a.nModels = (int)c->lzModels.size(); a.P = c->P;
    for (int i = 0; i < a.nModels; ++i) a.models[i] = c->lzModels[i];
    a.alphaCut = c->lzAlphaCut; a.sigmaSB = sigmaSB; a.deltaN = c->lzDeltaN;
    a.localFindCell = c->lzLocalFind ? 1 : 0;
    a.maxSteps = 64 * (int)cbrt((double)std::max(c->nC, 1)) + 4096;      // a straight ray crosses O(nCells^(1/3)) cells
    for (int p = 0; p < c->P; ++p) a.alpha[p] = c->alpha[p];
    a.refl = c->lzRefl; a.nHat = c->lzNHat; a.nHatPts = c->lzNHatPts; a.Q = c->lzQ; a.T = c->Tc;
    a.nAbs = c->lzNAbs;
    for (int i = 0; i < c->lzNAbs; ++i) { a.absPhase[i] = c->lzAbsPhase[i]; a.aCoeff[i] = c->lzA[i]; a.eCoeff[i] = c->lzE[i]; a.ECoeff[i] = c->lzEE[i]; }
    if (c->lzAbsFields) { a.aF = c->lzAF; a.eF = c->lzAF + c->stride; a.EF = c->lzAF + 2 * (size_t)c->stride; }
    // the seeds that found a cell on this rank (laserDTRM.C:287-296)
    std::vector<LaserPart> seeds;
    seeds.reserve(nParticles);
    for (int k = 0; k < nParticles; ++k) {
        if (cell[k] < 0) continue;
        LaserPart q{};
        for (int d = 0; d < 3; ++d) { q.pos[d] = q.p0[d] = p0[3 * (size_t)k + d]; q.p1[d] = p1[3 * (size_t)k + d]; }
        q.I0 = q.I = I[k]; q.dA = dA[k]; q.frac = 0; q.cell = cell[k]; q.tid = -1; q.face = -1; q.keep = 1;
        seeds.push_back(q);
    }
    // on a decomposed mesh any rank may end up walking every ray: capacity from the global seed count
    const long long base = std::max<long long>(globalSeeds, 1);
    const long long cap64 = std::max<long long>(maxPartAll > 0 ? maxPartAll : 64 * base, nSeeds);
    if (cap64 > 0x7fffffffLL / 2) failf("tcif_laser_calculate: cloud capacity %lld too large", cap64);
    const int cap = (int)cap64;
    if (cap > c->lzCap) {
        if (c->lzCloud) CUDA_OK(cudaFree(c->lzCloud));
        c->lzCloud = nullptr; c->lzCap = 0;
        CUDA_OK(cudaMalloc((void **)&c->lzCloud, sizeof(LaserPart) * (size_t)cap));
        c->lzCap = cap;
    }
    const int mailCap = par ? (int)std::min<long long>(std::max<long long>(4 * base, 1024), cap64) : 0;
    if (par && mailCap > c->lzMailCap) {
        if (c->lzOutbox) CUDA_OK(cudaFree(c->lzOutbox));
        if (c->lzInbox) CUDA_OK(cudaFree(c->lzInbox));
        c->lzOutbox = c->lzInbox = nullptr; c->lzMailCap = 0;
        CUDA_OK(cudaMalloc((void **)&c->lzOutbox, sizeof(LaserPart) * (size_t)mailCap * nCp));
        CUDA_OK(cudaMalloc((void **)&c->lzInbox, sizeof(LaserPart) * (size_t)mailCap * nCp));
        c->lzMailCap = mailCap;
    }
    a.cloud = c->lzCloud; a.cap = cap; a.counters = c->lzCounters;
    a.outbox = par ? c->lzOutbox : nullptr; a.mailCap = c->lzMailCap; a.patchInfo = c->lzPatchInfo;
    if (nSeeds) CUDA_OK(cudaMemcpyAsync(c->lzCloud, seeds.data(), sizeof(LaserPart) * nSeeds, cudaMemcpyHostToDevice, c->stream));
    int cnt[4 + 64] = {0};
    cnt[0] = nSeeds;
    CUDA_OK(cudaMemcpyAsync(c->lzCounters, cnt, sizeof(cnt), cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaMemsetAsync(c->lzQ, 0, sizeof(double) * c->nC, c->stream));
    if (a.nModels) { do_grad(c); a.G = c->G; }
    timed(c, T_LASER, (16.0 * a.nModels * 4 + 28.0) * c->nC, [&] {
        k_laser_mark<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, a);
        k_laser_points<<<cdiv(c->lg.nPoints, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->lg, c->lzNHat, c->lzNHatPts);
        c->launches += 2;
    });
    if (par) laser_sync_points(c);
    int done = 0, total = nSeeds, generations = 0;
    long long created = nSeeds;
    std::string error;
    for (;;) {
        if (total > done) {
            timed(c, T_LASER, 0.0, [&] {
                k_laser_move<<<cdiv(total - done, 128), 128, 0, c->stream>>>(c->dm, c->lg, a, done, total);
                c->launches++;
            });
        }
        CUDA_OK(cudaMemcpyAsync(cnt, c->lzCounters, sizeof(int) * (4 + nCp), cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
        if (cnt[1] == LASER_ERR_PROCESSOR) error = "a particle reached a processor patch of a context without mailboxes";
        if (cnt[1] == LASER_ERR_OVERFLOW || cnt[0] > cap) error = "cloud overflow, raise maxParticles";
        if (cnt[1] == LASER_ERR_STUCK) error = "a particle crossed more cells than a straight ray can (warped or non-convex cells?)";
        created += cnt[0] - total;
        done = total; total = std::min(cnt[0], cap);
        if (++generations > 10000) error = "reflection cascade does not end";
        if (!par) {
            if (!error.empty()) failf("tcif_laser_calculate: %s (capacity %d)", error.c_str(), cap);
            if (done == total) break;
            continue;
        }
        // hand-over: every rank learns how many particles each rank sends to each rank, who failed, who still has work
        std::vector<double> mine(c->nRanks + 3, 0.0), all;
        for (int i = 0; i < nCp; ++i) mine[c->patchNbr[c->cpPatch[i]]] += cnt[4 + i];
        mine[c->nRanks] = error.empty() ? 0.0 : 1.0;
        mine[c->nRanks + 1] = total - done;
        mine[c->nRanks + 2] = cap - total;                 // free slots of this rank's cloud
        allgather_host(c, mine.data(), c->nRanks + 3, all);
        const int W = c->nRanks + 3;
        bool anyErr = false;
        double pending = 0.0;
        for (int r = 0; r < c->nRanks; ++r) {
            anyErr = anyErr || all[(size_t)r * W + c->nRanks] != 0.0;
            pending += all[(size_t)r * W + c->nRanks + 1];
            for (int e = 0; e < c->nRanks; ++e) pending += all[(size_t)r * W + e];
        }
        if (anyErr) failf("tcif_laser_calculate: %s", error.empty() ? "another rank failed" : error.c_str());
        // receiver-side capacity, evaluated by every rank for every rank (the counts and the free slots are global
        // knowledge now), so that all ranks fail together
        for (int r = 0; r < c->nRanks; ++r) {
            double got = 0.0;
            for (int e = 0; e < c->nRanks; ++e) {
                const double n = all[(size_t)e * W + r];
                if (n > c->lzMailCap) failf("tcif_laser_calculate: %d particles handed from rank %d to rank %d exceed the mailbox (%d), raise maxParticles", (int)n, e, r, c->lzMailCap);
                got += n;
            }
            if (got > all[(size_t)r * W + c->nRanks + 2])
                failf("tcif_laser_calculate: cloud overflow on hand-over to rank %d, raise maxParticles (capacity %d)", r, cap);
        }
        if (pending == 0.0) break;
        std::vector<int> inCount(nCp, 0);
        for (int i = 0; i < nCp; ++i) inCount[i] = (int)all[(size_t)c->patchNbr[c->cpPatch[i]] * W + c->rank];
        if (c->group) {
            tcif_group *g = c->group;
            g->barrier();
            std::string bad;
            for (int i = 0; i < nCp && bad.empty(); ++i) {
                if (!inCount[i]) continue;
                tcif_ctx *o = g->members[c->patchNbr[c->cpPatch[i]]];
                int peer = -1;
                for (size_t j = 0; o && j < o->cpPatch.size(); ++j) if (o->patchNbr[o->cpPatch[j]] == c->rank) { peer = (int)j; break; }
                if (peer < 0) { bad = "neighbour rank has no patch facing this rank"; break; }
                if (cudaMemcpyAsync(c->lzInbox + (size_t)i * c->lzMailCap, o->lzOutbox + (size_t)peer * o->lzMailCap,
                                    sizeof(LaserPart) * inCount[i], cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess)
                    bad = std::string("hand-over copy failed: ") + cudaGetErrorString(cudaGetLastError());
            }
            if (cudaStreamSynchronize(c->stream) != cudaSuccess && bad.empty()) bad = std::string("hand-over copy failed: ") + cudaGetErrorString(cudaGetLastError());
            if (!bad.empty()) g->failed.store(1);
            g->barrier();
            if (!bad.empty()) failf("tcif_laser_calculate: %s", bad.c_str());
            if (g->failed.load()) failf("tcif_laser_calculate: hand-over failed on another rank of the group");
        } else {
            static_assert(sizeof(LaserPart) % sizeof(double) == 0, "LaserPart is sent as doubles");
            const size_t dpp = sizeof(LaserPart) / sizeof(double);
            NCCL_OK(g_nccl.GroupStart());
            for (int i = 0; i < nCp; ++i) {
                const int e = c->patchNbr[c->cpPatch[i]];
                if (cnt[4 + i]) NCCL_OK(g_nccl.Send(c->lzOutbox + (size_t)i * c->lzMailCap, cnt[4 + i] * dpp, kNcclDouble, e, c->nccl, c->stream));
                if (inCount[i]) NCCL_OK(g_nccl.Recv(c->lzInbox + (size_t)i * c->lzMailCap, inCount[i] * dpp, kNcclDouble, e, c->nccl, c->stream));
            }
            NCCL_OK(g_nccl.GroupEnd());
        }
        int got = 0;
        for (int i = 0; i < nCp; ++i) got += inCount[i];
        if (total + got > cap) failf("tcif_laser_calculate: internal error: hand-over exceeds the agreed capacity %d", cap);
        for (int i = 0; i < nCp; ++i) {
            if (!inCount[i]) continue;
            k_laser_receive<<<cdiv(inCount[i], TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->lg, inCount[i], c->lzInbox + (size_t)i * c->lzMailCap,
                                                                                c->patchStart[c->cpPatch[i]], c->lzCloud + total, c->lzCounters);
            c->launches++;
            total += inCount[i];
        }
        // reflected particles located here on behalf of a neighbour count as created on this rank
        CUDA_OK(cudaMemcpyAsync(cnt, c->lzCounters, sizeof(int) * 4, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
        created += cnt[2];
        cnt[0] = total; cnt[1] = 0; cnt[2] = 0;
        for (int i = 0; i < nCp; ++i) cnt[4 + i] = 0;
        CUDA_OK(cudaMemcpyAsync(c->lzCounters, cnt, sizeof(int) * (4 + nCp), cudaMemcpyHostToDevice, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
    }
    k_laser_finish<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, c->dm.V, c->lzQ);
    c->launches++;
    c->lzParticles = total; c->lzGenerations = generations; c->lzCreated = created;
    return (int)created;
}

// ---- host-buffer entry points: what crosses PCIe ---------------------------------------------------------------
void host_pipeline_init(tcif_ctx *c) {
    if (c->sIn) return;
    CUDA_OK(cudaStreamCreateWithFlags(&c->sIn, cudaStreamNonBlocking));
    CUDA_OK(cudaStreamCreateWithFlags(&c->sOut, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        CUDA_OK(cudaEventCreateWithFlags(&c->evIn[i], cudaEventDisableTiming));
        CUDA_OK(cudaEventCreateWithFlags(&c->evC[i], cudaEventDisableTiming));
    }
    c->stagingOut = c->dalloc<double>(c->stagingDoubles);
}
void delta_init(tcif_ctx *c) {
    if (c->dIdx2[0]) return;
    // (index, value) pairs cost 1.5x (scalar) / 1.17x (vector) the dense bytes per entry: beyond 1/4 of the entries the
    // dense copy is preferred
    c->dCap = std::max(std::max(c->nC, c->nIF) / 4, 1024);
    const size_t nVal = std::max<size_t>((size_t)c->dCap, 3 * (size_t)std::max(c->nC / 4, 1024));
    c->dCount = c->dalloc<int>(8);
    for (int b = 0; b < 2; ++b) {
        c->dIdx2[b] = c->dalloc<int>(c->dCap); c->dVal2[b] = c->dalloc<double>(nVal);
        CUDA_OK(cudaHostAlloc((void **)&c->hIdx2[b], sizeof(int) * (size_t)c->dCap, cudaHostAllocDefault));
        CUDA_OK(cudaHostAlloc((void **)&c->hVal2[b], sizeof(double) * nVal, cudaHostAllocDefault));
    }
    CUDA_OK(cudaHostAlloc((void **)&c->hCount, sizeof(int) * 8, cudaHostAllocDefault));
}
// dst[idx[k]] = val[k] (ncomp values per entry) on a few host threads: the indices of one list are distinct
void host_scatter(double *dst, const int *ix, const double *v, int n, int ncomp, int nRanks) {
    auto work = [=](int k0, int k1) {
        if (ncomp == 1) for (int k = k0; k < k1; ++k) dst[ix[k]] = v[k];
        else for (int k = k0; k < k1; ++k) { double *d = dst + 3 * (size_t)ix[k]; d[0] = v[3 * (size_t)k]; d[1] = v[3 * (size_t)k + 1]; d[2] = v[3 * (size_t)k + 2]; }
    };
    // (the ranks of one node share its cores: 8 threads per rank on a 32-core host with 8 ranks only fight each other)
    const unsigned share = std::max(1u, std::thread::hardware_concurrency() / (unsigned)std::max(nRanks, 1));
    const int nT = n < (1 << 18) ? 1 : (int)std::min<unsigned>(8, share);
    if (nT == 1) { work(0, n); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nT; ++t) th.emplace_back(work, (int)((long long)n * t / nT), (int)((long long)n * (t + 1) / nT));
    for (auto &t : th) t.join();
}
// Bring the entries of a device array `cur` (n entries of ncomp SoA components `stride` apart) whose value differs from
// `shadow` to the host array `dst` (AoS when ncomp == 3) on stream st, and update the shadow.  Returns false when too
// many differ (nothing was written: copy densely).  Synchronises st twice; every kernel the result depends on must
// have been enqueued before.
inline void scatter_join(tcif_ctx *c) { if (c->scatterTh.joinable()) c->scatterTh.join(); }
// joins the scatter in flight when an API call that reads back deltas returns, normally or by exception
struct ScatterJoin { tcif_ctx *c; explicit ScatterJoin(tcif_ctx *c_) : c(c_) {} ~ScatterJoin() { scatter_join(c); } };
// TCIF_HOST_TRACE=1: one line per delta read-back on stderr (entries, ms waiting for the GPU, ms scattering on the host)
static const bool g_hostTrace = [] { const char *e = getenv("TCIF_HOST_TRACE"); return e && atoi(e) != 0; }();
static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
bool down_delta(tcif_ctx *c, cudaStream_t st, double *dst, const double *cur, double *shadow, int n, int ncomp, size_t stride) {
    delta_init(c);
    const double t0 = g_hostTrace ? now_ms() : 0.0;
    int *dIdx = c->dIdx2[c->dPar], *hIdx = c->hIdx2[c->dPar];
    double *dVal = c->dVal2[c->dPar], *hVal = c->hVal2[c->dPar];
    const int cap = ncomp == 1 ? c->dCap : std::min(c->dCap, std::max(c->nC / 4, 1024));
    CUDA_OK(cudaMemsetAsync(c->dCount, 0, sizeof(int), st));
    if (ncomp == 1) k_delta_compact<<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, st>>>(n, cur, shadow, cap, dIdx, dVal, c->dCount);
    else k_delta_compact3<<<cdiv(n, TCIF_TPB), TCIF_TPB, 0, st>>>(n, stride, cur, shadow, cap, dIdx, dVal, c->dCount);
    c->launches++;
    CUDA_OK(cudaMemcpyAsync(c->hCount, c->dCount, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaStreamSynchronize(st));
    const int m = c->hCount[0];
    c->d2hBytes += sizeof(int);
    if (m > cap) return false;
    if (m == 0) return true;
    CUDA_OK(cudaMemcpyAsync(hIdx, dIdx, sizeof(int) * (size_t)m, cudaMemcpyDeviceToHost, st));
    CUDA_OK(cudaMemcpyAsync(hVal, dVal, sizeof(double) * (size_t)m * ncomp, cudaMemcpyDeviceToHost, st));
    k_delta_apply<<<cdiv(m, TCIF_TPB), TCIF_TPB, 0, st>>>(m, ncomp, stride, dIdx, dVal, shadow);
    c->launches++;
    CUDA_OK(cudaStreamSynchronize(st));
    c->d2hBytes += (int64_t)m * (sizeof(int) + sizeof(double) * ncomp);
    const double t1 = g_hostTrace ? now_ms() : 0.0;
    // the scatter runs beside the GPU work of the NEXT read-back (other buffer set); at most one is in flight
    scatter_join(c);
    const int nRanks = c->nRanks;
    c->scatterTh = std::thread([=] {
        host_scatter(dst, hIdx, hVal, m, ncomp, nRanks);
        if (g_hostTrace) fprintf(stderr, "[tcif trace] delta n=%d ncomp=%d entries=%d gpu+copy %.2f ms, host scatter %.2f ms\n", n, ncomp, m, t1 - t0, now_ms() - t1);
    });
    c->dPar ^= 1;
    return true;
}
// a face field (device slots) -> caller's array in OpenFOAM face order: delta against what the SAME host array received
// from the same slot last time, or the dense copy.  slot: 0..P-1 alphaPhi, P rhoPhi
void down_faces_host(tcif_ctx *c, cudaStream_t st, int slot, double *dst, const double *src) {
    if (!c->nIF || !dst) return;
    const size_t fb = sizeof(double) * c->nIF;
    k_face_gather<<<cdiv(c->nIF, TCIF_TPB), TCIF_TPB, 0, st>>>(c->nIF, c->d_faceMap, src, c->stagingOut);
    c->launches++;
    bool done = false;
    if (c->sparseReturns && slot < c->P) {          // rhoPhi (slot P) changes wherever phi does: always dense
        if (!c->shFace[slot]) c->shFace[slot] = c->dalloc<double>(c->nIF);
        if (c->shFaceHost[slot] == (const void *)dst) done = down_delta(c, st, dst, c->stagingOut, c->shFace[slot], c->nIF, 1, 0);
        if (!done) CUDA_OK(cudaMemcpyAsync(c->shFace[slot], c->stagingOut, fb, cudaMemcpyDeviceToDevice, st));
        c->shFaceHost[slot] = dst;
    }
    if (!done) { CUDA_OK(cudaMemcpyAsync(dst, c->stagingOut, fb, cudaMemcpyDefault, st)); c->d2hBytes += fb; }
}
// alpha[p] -> caller's array (in/out): delta against the snapshot taken at entry, or the dense copy
void down_alpha(tcif_ctx *c, cudaStream_t st, int p, double *ai, double *ab) {
    const size_t cb = sizeof(double) * c->nC, bb = sizeof(double) * c->nBF;
    if (!(c->sparseReturns && c->shAlpha[p] && down_delta(c, st, ai, c->alpha[p], c->shAlpha[p], c->nC, 1, 0))) {
        CUDA_OK(cudaMemcpyAsync(ai, c->alpha[p], cb, cudaMemcpyDefault, st));
        c->d2hBytes += cb;
    }
    if (c->nBF && ab) { CUDA_OK(cudaMemcpyAsync(ab, c->alphab[p], bb, cudaMemcpyDefault, st)); c->d2hBytes += bb; }
}
// F -> caller's AoS arrays: delta against what the SAME host array received last time, or the dense copy
void down_force(tcif_ctx *c, cudaStream_t st, double *Fi, double *Fb) {
    const size_t cb = sizeof(double) * c->nC, bb = sizeof(double) * c->nBF;
    if (Fi) {
        bool done = false;
        if (c->sparseReturns) {
            if (!c->shF) c->shF = c->dalloc<double>(3 * (size_t)c->stride);
            if (c->shFHost == (const void *)Fi) done = down_delta(c, st, Fi, c->F, c->shF, c->nC, 3, (size_t)c->stride);
        }
        if (!done) {
            k_soa_to_aos3<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, st>>>(c->nC, (size_t)c->stride, c->F, c->stagingOut);
            c->launches++;
            CUDA_OK(cudaMemcpyAsync(Fi, c->stagingOut, 3 * cb, cudaMemcpyDefault, st));
            c->d2hBytes += 3 * cb;
            if (c->sparseReturns) CUDA_OK(cudaMemcpyAsync(c->shF, c->F, sizeof(double) * 3 * (size_t)c->stride, cudaMemcpyDeviceToDevice, st));
        }
        if (c->sparseReturns) c->shFHost = Fi;       // the host array now equals F
    }
    if (Fb && c->nBF) {
        k_soa_to_aos3<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, st>>>(c->nBF, (size_t)c->nBF, c->Fb, c->stagingOut);
        c->launches++;
        CUDA_OK(cudaMemcpyAsync(Fb, c->stagingOut, 3 * bb, cudaMemcpyDefault, st));
        c->d2hBytes += 3 * bb;
    }
}
// uploads of the caller's alpha arrays on stream `in`, skipped where the one-shot stamps say the device copy is that
// array; then (sparse returns) the snapshot of alpha the deltas are taken against, on the same stream
void up_alphas_host(tcif_ctx *c, cudaStream_t in, double *const *ai, double *const *ab) {
    const size_t cb = sizeof(double) * c->nC, bb = sizeof(double) * c->nBF;
    for (int p = 0; p < c->P; ++p) {
        const bool resident = c->hsArmed && c->hsIn[p] != 0 && c->hsIn[p] == c->alphaStamp[p];
        if (!resident) {
            alpha_changed(c, p);
            CUDA_OK(cudaMemcpyAsync(c->alpha[p], ai[p], cb, cudaMemcpyDefault, in));
            c->h2dBytes += cb;
            if (c->nBF) { CUDA_OK(cudaMemcpyAsync(c->alphab[p], ab[p], bb, cudaMemcpyDefault, in)); c->h2dBytes += bb; }
            c->xfer[0]++;
        } else c->xfer[1]++;
        if (c->sparseReturns) {
            if (!c->shAlpha[p]) c->shAlpha[p] = c->dalloc<double>(c->stride);
            CUDA_OK(cudaMemcpyAsync(c->shAlpha[p], c->alpha[p], cb, cudaMemcpyDeviceToDevice, in));
        }
    }
}
void finish_host_stamps(tcif_ctx *c) {
    if (c->hsArmed) for (int p = 0; p < c->P; ++p) c->alphaStamp[p] = c->hsOut[p];
    c->hsArmed = false;
}

}  // namespace

// ========================================================================================
// C ABI
// ========================================================================================
#define API_BEGIN(ctx)                                        \
    if (!(ctx)) return 1;                                     \
    try {                                                     \
        DeviceGuard guard_((ctx)->device);
#define API_END(ctx)                                          \
        return 0;                                             \
    } catch (const std::exception &e) {                       \
        (ctx)->err = e.what();                                \
        return 1;                                             \
    }

extern "C" {

const char *tcif_version(void) {
#if TCIF_FORCE_EXACT
    return "tcif 0.2 (sm_100a, fp64, fmad=off, force kernels exact)";
#else
    return "tcif 0.2 (sm_100a, fp64, MULES fmad=off, force kernels reciprocal+fma)";
#endif
}

int tcif_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char *tcif_last_error(const tcif_ctx *ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }

int tcif_create(const tcif_mesh_desc *mesh, const tcif_config *cfg, tcif_ctx **out) {
    if (out) *out = nullptr;
    tcif_ctx *c = nullptr;
    int prevDev = -1;
    try {
        if (!mesh || !cfg || !out) failf("tcif_create: null argument");
        int nDev = tcif_device_count();
        if (nDev <= 0) failf("tcif_create: no CUDA device available (this library has no CPU fallback)");
        if (cfg->nPhases < 2 || cfg->nPhases > TCIF_MAX_PHASES) failf("tcif_create: nPhases must be 2..%d", TCIF_MAX_PHASES);
        if (cfg->nLimiterIter < 0 || cfg->nLimiterIter > 64) failf("tcif_create: nLimiterIter out of range");
        if (cfg->boundaryExtrema != TCIF_BEXT_GLOBAL && cfg->boundaryExtrema != TCIF_BEXT_PATCHVALUE) failf("tcif_create: bad boundaryExtrema");
        c = new tcif_ctx();
        c->cfg = *cfg;
        c->tsum = TCIF_FORCE_EXACT ? 0 : 1;
        if (const char *e = getenv("TCIF_TSUM")) c->tsum = atoi(e) ? 1 : 0;
        cudaGetDevice(&prevDev);
        c->device = cfg->device >= 0 ? cfg->device : prevDev;
        if (c->device >= nDev) failf("tcif_create: device %d not present (%d visible)", c->device, nDev);
        CUDA_OK(cudaSetDevice(c->device));
        cudaDeviceProp prop;
        CUDA_OK(cudaGetDeviceProperties(&prop, c->device));
        if (prop.major < 10) failf("tcif_create: device %d is sm_%d%d; this build targets sm_100a only", c->device, prop.major, prop.minor);
        CUDA_OK(cudaStreamCreateWithFlags(&c->ownStream, cudaStreamNonBlocking));
        c->stream = c->ownStream;
        {
            int lo = 0, hi = 0;
            CUDA_OK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            // TCIF_NO_OVERLAP=1 (profiling aid): surface work runs on the main stream, in sequence
            const char *no = getenv("TCIF_NO_OVERLAP");
            if (no && no[0] == '1') { c->side = c->ownStream; c->noOverlap = true; }
            else CUDA_OK(cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, hi));   // surface work first
            CUDA_OK(cudaEventCreateWithFlags(&c->evV, cudaEventDisableTiming));
            CUDA_OK(cudaEventCreateWithFlags(&c->evB, cudaEventDisableTiming));
        }
        c->P = cfg->nPhases; c->Q = c->P * (c->P - 1) / 2;
        build_mesh(c, mesh);
        alloc_fields(c);
        CUDA_OK(cudaStreamSynchronize(c->stream));
        if (prevDev >= 0) cudaSetDevice(prevDev);
        *out = c;
        return 0;
    } catch (const std::exception &e) {
        g_createError = e.what();
        if (c) { tcif_destroy(c); }
        if (prevDev >= 0) cudaSetDevice(prevDev);
        return 1;
    }
}

void tcif_destroy(tcif_ctx *c) {
    if (!c) return;
    DeviceGuard guard(c->device);
    if (c->side) cudaStreamSynchronize(c->side);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (void *ptr : c->peerMapped) if (ptr) cudaIpcCloseMemHandle(ptr);
    if (c->peerBuf) cudaFree(c->peerBuf);
    if (c->evX) cudaEventDestroy(c->evX);
    if (c->nccl && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl);
    if (c->group) { std::lock_guard<std::mutex> lk(c->group->mu); if (c->rank < (int)c->group->members.size()) c->group->members[c->rank] = nullptr; }
    for (auto &e : c->evs) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    for (auto &e : c->evPool) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    for (void *p : c->allocs) cudaFree(p);
    if (c->agBuf) cudaFree(c->agBuf);
    if (c->lzCloud) cudaFree(c->lzCloud);
    if (c->lzOutbox) cudaFree(c->lzOutbox);
    if (c->lzSpCells) cudaFree(c->lzSpCells);
    if (c->lzSpVals) cudaFree(c->lzSpVals);
    if (c->lzInbox) cudaFree(c->lzInbox);
    if (c->h_small) cudaFreeHost(c->h_small);
    scatter_join(c);
    for (int b = 0; b < 2; ++b) { if (c->hIdx2[b]) cudaFreeHost(c->hIdx2[b]); if (c->hVal2[b]) cudaFreeHost(c->hVal2[b]); }
    if (c->hCount) cudaFreeHost(c->hCount);
    if (c->sIn) cudaStreamDestroy(c->sIn);
    if (c->sOut) cudaStreamDestroy(c->sOut);
    for (int i = 0; i < 2; ++i) { if (c->evIn[i]) cudaEventDestroy(c->evIn[i]); if (c->evC[i]) cudaEventDestroy(c->evC[i]); }
    if (c->evV) cudaEventDestroy(c->evV);
    if (c->evB) cudaEventDestroy(c->evB);
    if (c->side && !c->noOverlap) cudaStreamDestroy(c->side);
    if (c->ownStream) cudaStreamDestroy(c->ownStream);
    delete c;
}

int tcif_set_stream(tcif_ctx *c, void *s) {
    API_BEGIN(c)
    join(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->stream = s ? (cudaStream_t)s : c->ownStream;
    if (c->noOverlap) c->side = c->stream;
    API_END(c)
}

int tcif_synchronize(tcif_ctx *c) {
    API_BEGIN(c)
    join(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    peer_check(c);
    API_END(c)
}

void *tcif_host_alloc(uint64_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void tcif_host_free(void *p) { if (p) cudaFreeHost(p); }

int tcif_get_ldu_csr(tcif_ctx *c, int32_t *ownerStart, int32_t *losort, int32_t *losortStart) {
    API_BEGIN(c)
    const int nC = c->nC, nIF = c->nIF;
    std::vector<int> faceMap(nIF), nbrOwn(c->nFd);
    CUDA_OK(cudaMemcpy(faceMap.data(), c->d_faceMap, sizeof(int) * nIF, cudaMemcpyDeviceToHost));
    CUDA_OK(cudaMemcpy(nbrOwn.data(), c->dm.nbrOwn, sizeof(int) * c->nFd, cudaMemcpyDeviceToHost));
    std::vector<int> os(nC + 1, 0), ls(nC + 1, 0);
    for (int f = 0; f < nIF; ++f) { os[faceMap[f] % c->stride + 1]++; ls[nbrOwn[faceMap[f]] + 1]++; }
    for (int i = 0; i < nC; ++i) { os[i + 1] += os[i]; ls[i + 1] += ls[i]; }
    if (ownerStart) memcpy(ownerStart, os.data(), sizeof(int) * (nC + 1));
    if (losortStart) memcpy(losortStart, ls.data(), sizeof(int) * (nC + 1));
    if (losort) {
        std::vector<int> pos(ls.begin(), ls.end() - 1);
        for (int f = 0; f < nIF; ++f) losort[pos[nbrOwn[faceMap[f]]]++] = f;
    }
    API_END(c)
}

// ---- communication -----------------------------------------------------------------------
int tcif_comm_unique_id(void *id128) {
    try {
        if (!id128) failf("null id buffer");
        nccl_load();
        ncclUniqueId id;
        NCCL_OK(g_nccl.GetUniqueId(&id));
        memcpy(id128, &id, 128);
        return 0;
    } catch (const std::exception &e) { g_createError = e.what(); return 1; }
}

int tcif_comm_init_nccl(tcif_ctx *c, int rank, int nRanks, const void *id128) {
    API_BEGIN(c)
    if (c->group || c->nccl) failf("communicator already attached");
    if (rank < 0 || rank >= nRanks || !id128) failf("tcif_comm_init_nccl: bad arguments");
    nccl_load();
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    NCCL_OK(g_nccl.CommInitRank(&c->nccl, nRanks, id, rank));
    c->rank = rank; c->nRanks = nRanks;
    for (int p : c->cpPatch) if (c->patchNbr[p] >= nRanks || c->patchNbr[p] == rank) failf("processor patch %d: neighbour rank %d invalid", p, c->patchNbr[p]);
    c->d_gather = c->dalloc<double>((size_t)nRanks * 16);
    c->deltaNValid = false;
    peer_setup(c);
    API_END(c)
}

int tcif_group_create(int nRanks, tcif_group **out) {
    if (!out || nRanks < 1) return 1;
    tcif_group *g = new tcif_group();
    g->n = nRanks; g->members.assign(nRanks, nullptr); g->red.assign((size_t)nRanks * 16, 0.0);
    g->gatherPtr.assign(nRanks, nullptr);
    *out = g;
    return 0;
}
void tcif_group_destroy(tcif_group *g) { delete g; }

int tcif_comm_init_local(tcif_ctx *c, int rank, tcif_group *g) {
    API_BEGIN(c)
    if (c->group || c->nccl) failf("communicator already attached");
    if (!g || rank < 0 || rank >= g->n) failf("tcif_comm_init_local: bad arguments");
    {
        std::lock_guard<std::mutex> lk(g->mu);
        if (g->members[rank]) failf("rank %d already joined", rank);
        g->members[rank] = c;
    }
    c->group = g; c->rank = rank; c->nRanks = g->n;
    c->deltaNValid = false;
    API_END(c)
}

// ---- fields ---------------------------------------------------------------------------------
int tcif_set_alpha_bc(tcif_ctx *c, int phase, const int32_t *bcKind) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (c->nPatches) {
        if (!bcKind) failf("tcif_set_alpha_bc: null kinds");
        for (int p = 0; p < c->nPatches; ++p) if (bcKind[p] != TCIF_BC_ZERO_GRADIENT && bcKind[p] != TCIF_BC_FIXED_VALUE) failf("unsupported alpha boundary condition %d on patch %d", bcKind[p], p);
        std::vector<int> &h = c->bcHost[phase];
        if (h.size() == (size_t)c->nPatches && std::equal(h.begin(), h.end(), bcKind)) return 0;    // unchanged: F stays valid
        h.assign(bcKind, bcKind + c->nPatches);
        touch(c);
        CUDA_OK(cudaMemcpyAsync(c->bc[phase], bcKind, sizeof(int) * c->nPatches, cudaMemcpyDefault, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
    }
    API_END(c)
}

int tcif_set_alpha(tcif_ctx *c, int phase, const double *internal, const double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (!internal || (c->nBF && !boundary)) failf("tcif_set_alpha: null field");
    alpha_changed(c, phase);
    c->xfer[0]++;
    up_cells(c, c->alpha[phase], internal);
    up_bnd(c, c->alphab[phase], boundary);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_alpha_stamped(tcif_ctx *c, int phase, uint64_t stamp, const double *internal, const double *boundary, int32_t *uploaded) {
    if (uploaded) *uploaded = 0;
    API_BEGIN(c)
    check_phase(c, phase);
    if (!internal || (c->nBF && !boundary)) failf("tcif_set_alpha_stamped: null field");
    if (stamp != 0 && stamp == c->alphaStamp[phase]) { c->xfer[1]++; return 0; }      // the device copy IS this field
    alpha_changed(c, phase);
    c->xfer[0]++;
    up_cells(c, c->alpha[phase], internal);
    up_bnd(c, c->alphab[phase], boundary);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->alphaStamp[phase] = stamp;
    if (uploaded) *uploaded = 1;
    API_END(c)
}

int tcif_mark_alpha(tcif_ctx *c, int phase, uint64_t stamp) {
    API_BEGIN(c)
    check_phase(c, phase);
    c->alphaStamp[phase] = stamp;
    API_END(c)
}

int tcif_transfer_counters(const tcif_ctx *c, int64_t uploads[4]) {
    if (!c || !uploads) return 1;
    for (int i = 0; i < 4; ++i) uploads[i] = c->xfer[i];
    return 0;
}

int tcif_get_alpha(tcif_ctx *c, int phase, double *internal, double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (internal) CUDA_OK(cudaMemcpyAsync(internal, c->alpha[phase], sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
    if (boundary && c->nBF) CUDA_OK(cudaMemcpyAsync(boundary, c->alphab[phase], sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    peer_check(c);
    API_END(c)
}

int tcif_store_old_alphas(tcif_ctx *c) {
    API_BEGIN(c)
    store_old(c);
    API_END(c)
}

int tcif_set_alpha_old(tcif_ctx *c, int phase, const double *internal) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (!internal) failf("tcif_set_alpha_old: null field");
    up_cells(c, c->alpha0[phase], internal);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_temperature(tcif_ctx *c, const double *internal, const double *boundary) {
    API_BEGIN(c)
    touch(c);
    if (!internal || (c->nBF && !boundary)) failf("tcif_set_temperature: null field");
    c->tStamp = 0;
    c->xfer[2]++;
    up_cells(c, c->Tc, internal);
    up_bnd(c, c->Tb, boundary);
    c->haveT = true;
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_temperature_stamped(tcif_ctx *c, uint64_t stamp, const double *internal, const double *boundary, int32_t *uploaded) {
    if (uploaded) *uploaded = 0;
    API_BEGIN(c)
    if (!internal || (c->nBF && !boundary)) failf("tcif_set_temperature_stamped: null field");
    if (stamp != 0 && stamp == c->tStamp && c->haveT) { c->xfer[3]++; return 0; }
    if (sigma_needs_T(c)) touch(c);          // F depends on T only through a temperatureDependent sigma
    c->xfer[2]++;
    up_cells(c, c->Tc, internal);
    up_bnd(c, c->Tb, boundary);
    c->haveT = true;
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->tStamp = stamp;
    if (uploaded) *uploaded = 1;
    API_END(c)
}

int tcif_set_sigma(tcif_ctx *c, int phase1, int phase2, int kind, int n, const double *coeffs) {
    API_BEGIN(c)
    touch(c);
    check_phase(c, phase1); check_phase(c, phase2);
    if (phase1 == phase2) failf("tcif_set_sigma: a pair needs two different phases");
    int q = pair_index(c->P, std::min(phase1, phase2), std::max(phase1, phase2));
    SigmaModel s;
    memset(&s, 0, sizeof s);
    if (kind == TCIF_SIGMA_CONSTANT) { if (n < 1 || !coeffs) failf("constant sigma needs one coefficient"); s.kind = SIGMA_CONSTANT; s.n = 1; s.c[0] = coeffs[0]; }
    else if (kind == TCIF_SIGMA_POLYNOMIAL || kind == TCIF_SIGMA_TABLE) {
        if (n < 1 || 2 * n > TCIF_SIGMA_MAX_COEFFS || !coeffs) failf("sigma Function1 needs 1..%d (x,y) pairs", TCIF_SIGMA_MAX_COEFFS / 2);
        s.kind = kind == TCIF_SIGMA_POLYNOMIAL ? SIGMA_POLYNOMIAL : SIGMA_TABLE; s.n = n;
        for (int i = 0; i < 2 * n; ++i) s.c[i] = coeffs[i];
        if (kind == TCIF_SIGMA_TABLE) for (int i = 1; i < n; ++i) if (!(s.c[2 * i] > s.c[2 * (i - 1)])) failf("sigma table abscissae must increase");
    } else failf("unsupported surfaceTensionModel kind %d (constant | temperatureDependent polynomial/table)", kind);
    c->hSig[q] = s; c->sigSet[q] = true;
    CUDA_OK(cudaMemcpyAsync(c->dSig + q, &c->hSig[q], sizeof(SigmaModel), cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_phi(tcif_ctx *c, const double *internal, const double *boundary) {
    API_BEGIN(c)
    if ((c->nIF && !internal) || (c->nBF && !boundary)) failf("tcif_set_phi: null field");
    up_faces(c, c->phi, internal);
    up_bnd(c, c->phib, boundary);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_phi_alpha(tcif_ctx *c, int phase, const double *internal, const double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    if ((c->nIF && !internal) || (c->nBF && !boundary)) failf("tcif_set_phi_alpha: null field");
    up_faces(c, c->phiHO[phase], internal);
    up_bnd(c, c->phiHOb[phase], boundary);
    CUDA_OK(cudaMemcpyAsync(c->phiPsi[phase], c->phiHO[phase], sizeof(double) * c->nFd, cudaMemcpyDeviceToDevice, c->stream));
    if (c->nBF) CUDA_OK(cudaMemcpyAsync(c->phiPsib[phase], c->phiHOb[phase], sizeof(double) * c->nBF, cudaMemcpyDeviceToDevice, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_get_phi_alpha(tcif_ctx *c, int phase, double *internal, double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (internal) down_faces(c, internal, c->phiPsi[phase]);
    if (boundary && c->nBF) CUDA_OK(cudaMemcpyAsync(boundary, c->phiPsib[phase], sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_su_sp(tcif_ctx *c, int phase, const double *Su, const double *Sp) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (Su) { if (!c->Su[phase]) c->Su[phase] = c->dalloc<double>(c->stride); up_cells(c, c->Su[phase], Su); }
    if (Sp) { if (!c->Sp[phase]) c->Sp[phase] = c->dalloc<double>(c->stride); up_cells(c, c->Sp[phase], Sp); }
    c->haveSu[phase] = Su != nullptr; c->haveSp[phase] = Sp != nullptr;
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

// ---- Su/Sp from the mass-transfer models -----------------------------------------
int tcif_mass_transfer_clear(tcif_ctx *c) {
    API_BEGIN(c)
    c->massModels.clear(); c->massPairs.clear(); c->massInSolve = false;
    API_END(c)
}

int tcif_add_mass_transfer_lee(tcif_ctx *c, const tcif_lee_model *mdl) {
    API_BEGIN(c)
    if (!mdl) failf("tcif_add_mass_transfer_lee: null model");
    check_phase(c, mdl->from); check_phase(c, mdl->to);
    if (mdl->from == mdl->to) failf("tcif_add_mass_transfer_lee: from == to");
    if (mdl->modelVariable != TCIF_VAR_T && mdl->modelVariable != TCIF_VAR_ALPHA) failf("tcif_add_mass_transfer_lee: bad modelVariable");
    if (!(mdl->Tactivate != 0)) failf("tcif_add_mass_transfer_lee: Tactivate must be non-zero");
    for (const auto &mm : c->massModels)
        if (mm.from == mdl->from && mm.to == mdl->to) failf("massTransferModel (%d to %d) given twice", mdl->from, mdl->to);
    if ((int)c->massModels.size() >= TCIF_MAX_MASS_MODELS) failf("too many mass transfer models");
    c->massModels.push_back({mdl->from, mdl->to, mdl->modelVariable, mdl->includeDivU ? 1 : 0, mdl->C, mdl->Tactivate, mdl->alphaMin, nullptr, false});
    API_END(c)
}

int tcif_set_phase_pairs(tcif_ctx *c, int nPairs, const int32_t *pairs) {
    API_BEGIN(c)
    if (nPairs < 0 || nPairs > TCIF_MAX_PAIRS || (nPairs && !pairs)) failf("tcif_set_phase_pairs: bad arguments");
    for (int q = 0; q < nPairs; ++q) { check_phase(c, pairs[2 * q]); check_phase(c, pairs[2 * q + 1]); if (pairs[2 * q] == pairs[2 * q + 1]) failf("tcif_set_phase_pairs: pair of a phase with itself"); }
    c->massPairs.assign(pairs, pairs + 2 * nPairs);
    API_END(c)
}

static tcif_ctx::MassModel &find_mass_model(tcif_ctx *c, int from, int to) {
    for (auto &mm : c->massModels) if (mm.from == from && mm.to == to) return mm;
    failf("no massTransferModel (%d to %d)", from, to);
    throw std::logic_error("unreachable");
}

int tcif_set_dmdt(tcif_ctx *c, int from, int to, const double *internal) {
    API_BEGIN(c)
    auto &mm = find_mass_model(c, from, to);
    if (!internal) { mm.haveDmdt = false; }
    else {
        if (!mm.dmdt) mm.dmdt = c->dalloc<double>(c->stride);
        up_cells(c, mm.dmdt, internal);
        CUDA_OK(cudaStreamSynchronize(c->stream));
        mm.haveDmdt = true;
    }
    API_END(c)
}

int tcif_get_dmdt(tcif_ctx *c, int from, int to, double *internal) {
    API_BEGIN(c)
    auto &mm = find_mass_model(c, from, to);
    if (!internal) failf("tcif_get_dmdt: null output");
    if (mm.haveDmdt) CUDA_OK(cudaMemcpyAsync(internal, mm.dmdt, sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
    else memset(internal, 0, sizeof(double) * c->nC);          // dmdt(key) of an empty entry is a zero field (:127-151)
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_correct_mass_sources(tcif_ctx *c) {
    API_BEGIN(c)
    do_correct_mass_sources(c);
    API_END(c)
}

int tcif_alpha_transfer(tcif_ctx *c, double *ddtAlphaMax) {
    API_BEGIN(c)
    do_alpha_transfer(c);
    if (ddtAlphaMax) *ddtAlphaMax = c->ddtAlphaMax;
    API_END(c)
}

int tcif_set_mass_transfer_in_solve(tcif_ctx *c, int on) {
    API_BEGIN(c)
    c->massInSolve = on != 0;
    API_END(c)
}

int tcif_get_su_sp(tcif_ctx *c, int phase, double *Su, double *Sp) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (Su) { if (c->haveSu[phase]) CUDA_OK(cudaMemcpyAsync(Su, c->Su[phase], sizeof(double) * c->nC, cudaMemcpyDefault, c->stream)); else memset(Su, 0, sizeof(double) * c->nC); }
    if (Sp) { if (c->haveSp[phase]) CUDA_OK(cudaMemcpyAsync(Sp, c->Sp[phase], sizeof(double) * c->nC, cudaMemcpyDefault, c->stream)); else memset(Sp, 0, sizeof(double) * c->nC); }
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_interface_porosity(tcif_ctx *c, int nModels, const int32_t *solidPhase, const double *Cu, double *Udiag) {
    API_BEGIN(c)
    if (nModels < 0 || nModels > 8 || (nModels && (!solidPhase || !Cu)) || !Udiag) failf("tcif_interface_porosity: bad arguments");
    PorosityArgs a{};
    a.n = nModels;
    for (int i = 0; i < nModels; ++i) { check_phase(c, solidPhase[i]); a.alpha[i] = c->alpha[solidPhase[i]]; a.Cu[i] = Cu[i]; }
    if (c->stagingDoubles < (size_t)c->nC) failf("tcif_interface_porosity: staging buffer too small");
    up_cells(c, c->staging, Udiag);
    timed(c, T_MASS, (16.0 + 8.0 * nModels + 8) * c->nC, [&] {
        k_interface_porosity<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, c->dm.V, a, c->staging);
        c->launches++;
    });
    CUDA_OK(cudaMemcpyAsync(Udiag, c->staging, sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_solve_yi(tcif_ctx *c, int phase, int nSpecies, int inertIndex, double *const *Yinternal, double *const *Yboundary,
                  const int32_t *Ybc, const double *const *Su, const double *const *Sp, double cYi,
                  int schemePhi, int schemePhir, int nYiSubCycles, double deltaT, double tol, int maxIter) {
    API_BEGIN(c)
    do_solve_yi(c, phase, nSpecies, inertIndex, Yinternal, Yboundary, Ybc, Su, Sp, cYi, schemePhi, schemePhir, nYiSubCycles,
                deltaT, tol, maxIter);
    API_END(c)
}

// ---- hot path ---------------------------------------------------------------------------------
// ---- laserDTRM ----------------------------------------------------------------------------
int tcif_laser_set_polymesh(tcif_ctx *c, const double *Cf, const double *Cfb, int32_t nPoints, const double *points,
                            const int32_t *fpStart, const int32_t *fpPts, const int32_t *tetBasePt, double boundsMag) {
    API_BEGIN(c)
    do_laser_set_polymesh(c, Cf, Cfb, nPoints, points, fpStart, fpPts, tetBasePt, boundsMag);
    API_END(c)
}

int tcif_laser_set_reflection(tcif_ctx *c, int32_t nModels, const tcif_reflection_model *models, double alphaCut) {
    API_BEGIN(c)
    if (nModels < 0 || nModels > 8 || (nModels && !models)) failf("tcif_laser_set_reflection: 0..8 models");
    std::vector<LaserRefl> v;
    for (int i = 0; i < nModels; ++i) {
        const tcif_reflection_model &m = models[i];
        if (m.kind < 0 || m.kind > 2) failf("tcif_laser_set_reflection: unknown reflection model %d", m.kind);
        check_phase(c, m.from); check_phase(c, m.to);
        if (m.from == m.to) failf("tcif_laser_set_reflection: pair (%d %d)", m.from, m.to);
        LaserRefl r{m.kind, m.from, m.to, {m.nk1[0], m.nk1[1]}, {m.nk2[0], m.nk2[1]}, m.epsilon};
        v.push_back(r);
    }
    c->lzModels = v; c->lzAlphaCut = alphaCut;
    API_END(c)
}

int tcif_laser_set_local_density_absorption(tcif_ctx *c, int32_t nNames, const int32_t *phases, const double *aCoeff,
                                            const double *eCoeff, const double *ECoeff) {
    API_BEGIN(c)
    if (nNames < 1 || nNames > 4 || !phases || !aCoeff || !eCoeff || !ECoeff) failf("tcif_laser_set_local_density_absorption: bad arguments");
    for (int i = 0; i < nNames; ++i) {
        check_phase(c, phases[i]);
        c->lzAbsPhase[i] = phases[i]; c->lzA[i] = aCoeff[i]; c->lzE[i] = eCoeff[i]; c->lzEE[i] = ECoeff[i];
    }
    c->lzNAbs = nNames; c->lzAbsFields = false;
    API_END(c)
}

int tcif_laser_set_absorption_fields(tcif_ctx *c, const double *a, const double *e, const double *E) {
    API_BEGIN(c)
    if (!a || !e || !E) failf("tcif_laser_set_absorption_fields: null array");
    if (!c->lzAF) c->lzAF = c->dalloc<double>(3 * (size_t)c->stride);
    up_cells(c, c->lzAF, a); up_cells(c, c->lzAF + c->stride, e); up_cells(c, c->lzAF + 2 * (size_t)c->stride, E);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->lzAbsFields = true;
    API_END(c)
}

int tcif_laser_calculate(tcif_ctx *c, int32_t nParticles, const double *p0, const double *p1, const double *I, const double *dA,
                         const int32_t *cell, double sigmaSB, int32_t maxParticles, double *Q, int32_t *nParticlesFinal) {
    API_BEGIN(c)
    const int n = do_laser_calculate(c, nParticles, p0, p1, I, dA, cell, sigmaSB, maxParticles);
    if (Q) CUDA_OK(cudaMemcpyAsync(Q, c->lzQ, sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    if (nParticlesFinal) *nParticlesFinal = n;
    API_END(c)
}

int tcif_laser_get(tcif_ctx *c, double *Q, int32_t *reflectingCells, double *nHat, double *nHatPoints) {
    API_BEGIN(c)
    if (!c->laserGeoSet) failf("tcif_laser_get: call tcif_laser_set_polymesh first");
    if (Q) CUDA_OK(cudaMemcpyAsync(Q, c->lzQ, sizeof(double) * c->nC, cudaMemcpyDefault, c->stream));
    if (reflectingCells) CUDA_OK(cudaMemcpyAsync(reflectingCells, c->lzRefl, sizeof(int) * c->nC, cudaMemcpyDefault, c->stream));
    if (nHat) {
        k_soa_to_aos3<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, (size_t)c->stride, c->lzNHat, c->staging);
        c->launches++;
        CUDA_OK(cudaMemcpyAsync(nHat, c->staging, sizeof(double) * 3 * c->nC, cudaMemcpyDefault, c->stream));
    }
    if (nHatPoints) CUDA_OK(cudaMemcpyAsync(nHatPoints, c->lzNHatPts, sizeof(double) * 3 * c->lg.nPoints, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_laser_get_sparse(tcif_ctx *c, int32_t capacity, int32_t *cells, double *values, int32_t *n) {
    if (n) *n = 0;               // callers tell "capacity too small" (rc != 0, *n > capacity) from any other failure (*n == 0)
    API_BEGIN(c)
    if (!c->laserGeoSet) failf("tcif_laser_get_sparse: call tcif_laser_set_polymesh first");
    if (capacity < 0 || (capacity && (!cells || !values)) || !n) failf("tcif_laser_get_sparse: bad arguments");
    if (capacity > c->lzSpCap) {
        if (c->lzSpCells) CUDA_OK(cudaFree(c->lzSpCells));
        if (c->lzSpVals) CUDA_OK(cudaFree(c->lzSpVals));
        c->lzSpCells = nullptr; c->lzSpVals = nullptr; c->lzSpCap = 0;
        CUDA_OK(cudaMalloc((void **)&c->lzSpCells, sizeof(int) * (size_t)capacity));
        CUDA_OK(cudaMalloc((void **)&c->lzSpVals, sizeof(double) * (size_t)capacity));
        c->lzSpCap = capacity;
    }
    CUDA_OK(cudaMemsetAsync(c->lzCounters + 3, 0, sizeof(int), c->stream));
    k_laser_compact<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, c->lzQ, capacity, c->lzSpCells, c->lzSpVals, c->lzCounters + 3);
    c->launches++;
    int found = 0;
    CUDA_OK(cudaMemcpyAsync(&found, c->lzCounters + 3, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    *n = found;
    if (found > capacity) failf("tcif_laser_get_sparse: %d cells hold energy, capacity is %d", found, capacity);
    if (found) {
        CUDA_OK(cudaMemcpyAsync(cells, c->lzSpCells, sizeof(int) * found, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaMemcpyAsync(values, c->lzSpVals, sizeof(double) * found, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
    }
    API_END(c)
}

int tcif_laser_generations(const tcif_ctx *c) { return c ? c->lzGenerations : -1; }

int tcif_laser_set_local_find_cell(tcif_ctx *c, int on) {
    API_BEGIN(c)
    c->lzLocalFind = on != 0;
    API_END(c)
}

int tcif_laser_set_shared_points(tcif_ctx *c, int32_t nShared, const int32_t *localPoint, const int64_t *globalPoint) {
    API_BEGIN(c)
    do_laser_set_shared_points(c, nShared, localPoint, globalPoint);
    API_END(c)
}

int tcif_div_capillary_stress(tcif_ctx *c) {
    API_BEGIN(c)
    do_force(c);
    API_END(c)
}

int tcif_get_force(tcif_ctx *c, double *internal, double *boundary) {
    API_BEGIN(c)
    down_vec(c, internal, c->F, c->nC, c->stride);
    if (boundary && c->nBF) down_vec(c, boundary, c->Fb, c->nBF, c->nBF);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    peer_check(c);
    API_END(c)
}

int tcif_div_capillary_stress_host(tcif_ctx *c, const double *const *ai, const double *const *ab, const double *Ti,
                                   const double *Tb, double *Fi, double *Fb) {
    API_BEGIN(c)
    if (!ai || (c->nBF && !ab)) failf("tcif_div_capillary_stress_host: null alpha lists");
    for (int p = 0; p < c->P; ++p) alpha_changed(c, p);
    c->tStamp = 0;
    for (int p = 0; p < c->P; ++p) { up_cells(c, c->alpha[p], ai[p]); if (c->nBF) up_bnd(c, c->alphab[p], ab[p]); }
    if (Ti) { up_cells(c, c->Tc, Ti); up_bnd(c, c->Tb, Tb); c->haveT = true; }
    do_force(c);
    down_vec(c, Fi, c->F, c->nC, c->stride);
    if (Fb && c->nBF) down_vec(c, Fb, c->Fb, c->nBF, c->nBF);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

// fvc::flux(fluid.divCapillaryStress())  (pEqn.H:31-38): F is evaluated only if alpha/T/sigma changed since
int tcif_flux_of_force(tcif_ctx *c, double *internal, double *boundary) {
    API_BEGIN(c)
    do_force(c);
    double *fl = c->phiCorr[0], *flb = c->phiCorrb[0];       // scratch surface field (MULES work array, free between limits)
    k_flux_of_force<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->F, fl);
    c->launches++;
    if (c->nBF) { k_flux_of_force_bnd<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->dm, c->F, c->Fb, flb); c->launches++; }
    if (internal) down_faces(c, internal, fl);
    if (boundary && c->nBF) CUDA_OK(cudaMemcpyAsync(boundary, flb, sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_force_cache(tcif_ctx *c, int on) {
    API_BEGIN(c)
    c->cacheForce = on != 0;
    API_END(c)
}

int tcif_force_counters(const tcif_ctx *c, int64_t *evaluations, int64_t *cacheHits) {
    if (!c) return 1;
    if (evaluations) *evaluations = c->forceEvaluations;
    if (cacheHits) *cacheHits = c->forceCacheHits;
    return 0;
}

int tcif_grad_alpha(tcif_ctx *c, int phase, double *internal, double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    do_grad(c);
    down_vec(c, internal, c->G + (size_t)phase * 3 * c->stride, c->nC, c->stride);
    if (boundary && c->nBF) down_vec(c, boundary, c->Gb + (size_t)phase * 3 * c->nBF, c->nBF, c->nBF);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_mules_limit(tcif_ctx *c, int phase, double rDeltaT, int returnCorr) {
    API_BEGIN(c)
    check_phase(c, phase);
    do_limit(c, &phase, 1, rDeltaT, returnCorr ? 0 : 1, false, false);
    API_END(c)
}

int tcif_mules_limit_sum(tcif_ctx *c) {
    API_BEGIN(c)
    do_limit_sum(c);
    API_END(c)
}

int tcif_add_upwind_flux(tcif_ctx *c, int phase) {
    API_BEGIN(c)
    check_phase(c, phase);
    do_add_upwind(c, phase);
    API_END(c)
}

int tcif_mules_explicit_solve(tcif_ctx *c, int phase, double deltaT) {
    API_BEGIN(c)
    check_phase(c, phase);
    do_explicit(c, &phase, 1, deltaT, false);
    API_END(c)
}

int tcif_stage(tcif_ctx *c, double deltaT) {
    API_BEGIN(c)
    do_stage(c, deltaT);
    API_END(c)
}

// Pipelined host-buffer stage: [H2D alpha,T] -> force -> [D2H F || H2D phi, phiAlpha] -> MULES -> [D2H alpha, alphaPhi]
// All compute is enqueued before the first download, because the delta read-back synchronises the download stream.
int tcif_stage_host(tcif_ctx *c, double deltaT, double *const *ai, double *const *ab, const double *Ti, const double *Tb,
                    const double *phii, const double *phib, const double *const *pai, const double *const *pab,
                    double *const *apo, double *const *apob, double *Fi, double *Fb) {
    API_BEGIN(c)
    ScatterJoin scatterGuard(c);
    if (!ai || !pai || !apo || (c->nBF && (!ab || !pab || !apob))) failf("tcif_stage_host: null field lists");
    if (!(deltaT > 0)) failf("tcif_stage_host: deltaT must be positive");
    host_pipeline_init(c);
    join(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    const double tr0 = g_hostTrace ? now_ms() : 0.0;
    c->h2dBytes = c->d2hBytes = 0;
    const size_t cb = sizeof(double) * c->nC, bb = sizeof(double) * c->nBF, fb = sizeof(double) * c->nIF;
    cudaStream_t in = c->sIn, out = c->sOut;
    auto upFaces = [&](double *dst, const double *src) {
        if (!c->nIF) return;
        CUDA_OK(cudaMemcpyAsync(c->staging, src, fb, cudaMemcpyDefault, in));
        k_face_scatter<<<cdiv(c->nIF, TCIF_TPB), TCIF_TPB, 0, in>>>(c->nIF, c->d_faceMap, c->staging, dst);
        c->launches++;
        c->h2dBytes += fb;
    };
    // group A: what the force needs
    up_alphas_host(c, in, ai, ab);
    if (Ti) {
        touch(c); c->tStamp = 0; c->xfer[2]++;
        CUDA_OK(cudaMemcpyAsync(c->Tc, Ti, cb, cudaMemcpyDefault, in));
        if (c->nBF && Tb) CUDA_OK(cudaMemcpyAsync(c->Tb, Tb, bb, cudaMemcpyDefault, in));
        c->h2dBytes += cb + (Tb ? bb : 0);
        c->haveT = true;
    }
    CUDA_OK(cudaEventRecord(c->evIn[0], in));
    // group B: what MULES needs, streaming in while the force runs
    if (phii || !c->nIF) { upFaces(c->phi, phii); if (c->nBF && phib) { CUDA_OK(cudaMemcpyAsync(c->phib, phib, bb, cudaMemcpyDefault, in)); c->h2dBytes += bb; } }
    for (int p = 0; p < c->P; ++p) {
        upFaces(c->phiHO[p], pai[p]);
        if (c->nBF) { CUDA_OK(cudaMemcpyAsync(c->phiHOb[p], pab[p], bb, cudaMemcpyDefault, in)); c->h2dBytes += bb; }
    }
    CUDA_OK(cudaEventRecord(c->evIn[1], in));
    // compute: force, then the MULES corrector
    CUDA_OK(cudaStreamWaitEvent(c->stream, c->evIn[0], 0));
    do_force(c);
    CUDA_OK(cudaEventRecord(c->evC[0], c->stream));
    CUDA_OK(cudaStreamWaitEvent(c->stream, c->evIn[1], 0));
    do_mules_stage(c, deltaT);
    CUDA_OK(cudaEventRecord(c->evC[1], c->stream));
    // downloads
    const double tr1 = g_hostTrace ? now_ms() : 0.0;
    CUDA_OK(cudaStreamWaitEvent(out, c->evC[0], 0));
    down_force(c, out, Fi, Fb);
    const double tr2 = g_hostTrace ? now_ms() : 0.0;
    CUDA_OK(cudaStreamWaitEvent(out, c->evC[1], 0));
    if (g_hostTrace) { CUDA_OK(cudaStreamSynchronize(c->stream)); fprintf(stderr, "[tcif trace] stage_host: enqueue %.2f ms, F back at %.2f ms, compute done at %.2f ms\n", tr1 - tr0, tr2 - tr0, now_ms() - tr0); }
    for (int p = 0; p < c->P; ++p) {
        down_faces_host(c, out, p, apo[p], c->phiPsi[p]);
        if (c->nBF) { CUDA_OK(cudaMemcpyAsync(apob[p], c->phiPsib[p], bb, cudaMemcpyDefault, out)); c->d2hBytes += bb; }
    }
    for (int p = 0; p < c->P; ++p) down_alpha(c, out, p, ai[p], c->nBF ? ab[p] : nullptr);
    CUDA_OK(cudaStreamSynchronize(out));
    CUDA_OK(cudaStreamSynchronize(in));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    CUDA_OK(cudaGetLastError());
    finish_host_stamps(c);
    scatter_join(c);
    if (g_hostTrace) fprintf(stderr, "[tcif trace] stage_host: done at %.2f ms\n", now_ms() - tr0);
    API_END(c)
}

// The reference's own call surface with host buffers: fluid.solve() (thermocapillaryInterFoam.C:105) then
// fluid.divCapillaryStress() on the new alpha (UEqn.H:29 / thermocapillaryInterFoam.C:150), pipelined.
int tcif_time_step_host(tcif_ctx *c, const tcif_alpha_controls *ctl, double deltaT, double *const *ai, double *const *ab,
                        const double *Ti, const double *Tb, const double *phii, const double *phib,
                        double *const *apo, double *const *apob, double *rpi, double *rpb, double *Fi, double *Fb) {
    API_BEGIN(c)
    ScatterJoin scatterGuard(c);
    if (!ai || (c->nBF && !ab) || !phii) failf("tcif_time_step_host: null field lists");
    host_pipeline_init(c);
    join(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->h2dBytes = c->d2hBytes = 0;
    const size_t cb = sizeof(double) * c->nC, bb = sizeof(double) * c->nBF, fb = sizeof(double) * c->nIF;
    cudaStream_t in = c->sIn, out = c->sOut;
    up_alphas_host(c, in, ai, ab);
    if (c->nIF) {
        CUDA_OK(cudaMemcpyAsync(c->staging, phii, fb, cudaMemcpyDefault, in));
        k_face_scatter<<<cdiv(c->nIF, TCIF_TPB), TCIF_TPB, 0, in>>>(c->nIF, c->d_faceMap, c->staging, c->phi);
        c->launches++;
        c->h2dBytes += fb;
    }
    if (c->nBF && phib) { CUDA_OK(cudaMemcpyAsync(c->phib, phib, bb, cudaMemcpyDefault, in)); c->h2dBytes += bb; }
    CUDA_OK(cudaEventRecord(c->evIn[0], in));
    if (Ti) {
        touch(c); c->tStamp = 0; c->xfer[2]++;
        CUDA_OK(cudaMemcpyAsync(c->Tc, Ti, cb, cudaMemcpyDefault, in));
        if (c->nBF && Tb) CUDA_OK(cudaMemcpyAsync(c->Tb, Tb, bb, cudaMemcpyDefault, in));
        c->h2dBytes += cb + (Tb ? bb : 0);
        c->haveT = true;
    }
    CUDA_OK(cudaEventRecord(c->evIn[1], in));
    // compute: solve(), then the force on the new alpha
    CUDA_OK(cudaStreamWaitEvent(c->stream, c->evIn[0], 0));
    do_solve_alphas(c, ctl, deltaT);
    CUDA_OK(cudaEventRecord(c->evC[0], c->stream));
    CUDA_OK(cudaStreamWaitEvent(c->stream, c->evIn[1], 0));
    do_force(c);
    CUDA_OK(cudaEventRecord(c->evC[1], c->stream));
    // downloads: the dense flux arrays first, then the deltas
    CUDA_OK(cudaStreamWaitEvent(out, c->evC[0], 0));
    down_faces_host(c, out, c->P, rpi, c->rhoPhi);
    if (rpb && c->nBF) { CUDA_OK(cudaMemcpyAsync(rpb, c->rhoPhib, bb, cudaMemcpyDefault, out)); c->d2hBytes += bb; }
    if (apo) for (int p = 0; p < c->P; ++p) {
        down_faces_host(c, out, p, apo[p], c->phiPsi[p]);
        if (apob && c->nBF) { CUDA_OK(cudaMemcpyAsync(apob[p], c->phiPsib[p], bb, cudaMemcpyDefault, out)); c->d2hBytes += bb; }
    }
    for (int p = 0; p < c->P; ++p) down_alpha(c, out, p, ai[p], c->nBF ? ab[p] : nullptr);
    CUDA_OK(cudaStreamWaitEvent(out, c->evC[1], 0));
    down_force(c, out, Fi, Fb);
    CUDA_OK(cudaStreamSynchronize(out));
    CUDA_OK(cudaStreamSynchronize(in));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    CUDA_OK(cudaGetLastError());
    finish_host_stamps(c);
    API_END(c)
}

// one-shot stamps of the alpha arrays the NEXT tcif_stage_host / tcif_time_step_host call receives (in) and of the
// state those arrays hold when it returns (out); see include/tcif.h
int tcif_host_alpha_stamps(tcif_ctx *c, const uint64_t *in, const uint64_t *out) {
    API_BEGIN(c)
    for (int p = 0; p < c->P; ++p) { c->hsIn[p] = in ? in[p] : 0; c->hsOut[p] = out ? out[p] : 0; }
    c->hsArmed = true;
    API_END(c)
}
int tcif_host_sparse_returns(tcif_ctx *c, int on) {
    API_BEGIN(c)
    c->sparseReturns = on != 0;
    c->shFHost = nullptr;
    for (auto &h : c->shFaceHost) h = nullptr;
    API_END(c)
}
int tcif_host_bytes_moved(const tcif_ctx *c, int64_t *h2d, int64_t *d2h) {
    if (!c) return 1;
    if (h2d) *h2d = c->h2dBytes;
    if (d2h) *d2h = c->d2hBytes;
    return 0;
}

int tcif_set_cell_centres(tcif_ctx *c, const double *C, const double *deltaB) {
    API_BEGIN(c)
    if (!C || (c->nBF && !deltaB)) failf("tcif_set_cell_centres: null array");
    if (!c->Cx) {
        c->Cx = c->dalloc<double>(3 * (size_t)c->stride); c->Cy = c->Cx + c->stride; c->Cz = c->Cy + c->stride;
        c->dbx = c->dalloc<double>(3 * (size_t)std::max(c->nBF, 1)); c->dby = c->dbx + std::max(c->nBF, 1); c->dbz = c->dby + std::max(c->nBF, 1);
    }
    CUDA_OK(cudaMemcpyAsync(c->staging, C, sizeof(double) * 3 * c->nC, cudaMemcpyDefault, c->stream));
    k_aos3_to_soa<<<cdiv(c->nC, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nC, (size_t)c->stride, c->staging, c->Cx);
    c->launches++;
    if (c->nBF) {
        CUDA_OK(cudaStreamSynchronize(c->stream));
        CUDA_OK(cudaMemcpyAsync(c->staging, deltaB, sizeof(double) * 3 * c->nBF, cudaMemcpyDefault, c->stream));
        k_aos3_to_soa<<<cdiv(c->nBF, TCIF_TPB), TCIF_TPB, 0, c->stream>>>(c->nBF, (size_t)std::max(c->nBF, 1), c->staging, c->dbx);
        c->launches++;
    }
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_set_calpha(tcif_ctx *c, const double *cAlpha) {
    API_BEGIN(c)
    if (!cAlpha) failf("tcif_set_calpha: null table");
    for (int i = 0; i < c->P * c->P; ++i) {
        if (!(cAlpha[i] >= 0)) failf("tcif_set_calpha: negative or NaN entry %d", i);
        c->cAlpha[i] = cAlpha[i];
    }
    API_END(c)
}

int tcif_set_rho(tcif_ctx *c, int phase, double uniform, const double *internal, const double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    c->rhoU[phase] = uniform;
    if (internal) {
        if (c->nBF && !boundary) failf("tcif_set_rho: field without boundary values");
        if (!c->rho[phase]) { c->rho[phase] = c->dalloc<double>(c->stride); c->rhob[phase] = c->dalloc<double>(std::max(c->nBF, 1)); }
        up_cells(c, c->rho[phase], internal);
        up_bnd(c, c->rhob[phase], boundary);
        CUDA_OK(cudaStreamSynchronize(c->stream));
    }
    c->haveRho[phase] = internal != nullptr;
    c->rhoSet[phase] = true;
    API_END(c)
}

int tcif_flux_assembly(tcif_ctx *c, int schemePhi, int schemePhir) {
    API_BEGIN(c)
    PhicState st;
    phic_begin(c, st);
    do_flux_assembly(c, schemePhi, schemePhir, st);
    for (int p = 0; p < c->P; ++p) {      // becomes the input of tcif_mules_limit like tcif_set_phi_alpha does
        CUDA_OK(cudaMemcpyAsync(c->phiPsi[p], c->phiHO[p], sizeof(double) * c->nFd, cudaMemcpyDeviceToDevice, c->stream));
        if (c->nBF) CUDA_OK(cudaMemcpyAsync(c->phiPsib[p], c->phiHOb[p], sizeof(double) * c->nBF, cudaMemcpyDeviceToDevice, c->stream));
    }
    API_END(c)
}

int tcif_get_flux_assembly(tcif_ctx *c, int phase, double *internal, double *boundary) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (internal) down_faces(c, internal, c->phiHO[phase]);
    if (boundary && c->nBF) CUDA_OK(cudaMemcpyAsync(boundary, c->phiHOb[phase], sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_upwind_predictor(tcif_ctx *c, int phase, double deltaT, double tol, int maxIter) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (!(tol > 0) || maxIter < 1) failf("tcif_upwind_predictor: tol must be > 0 and maxIter >= 1");
    do_predictor(c, &phase, 1, deltaT, tol, maxIter);
    API_END(c)
}

int tcif_predictor_sweeps(const tcif_ctx *c, int phase) { return (c && phase >= 0 && phase < c->P) ? c->predictorSweeps[phase] : -1; }

int tcif_solve_alphas(tcif_ctx *c, const tcif_alpha_controls *ctl, double deltaT) {
    API_BEGIN(c)
    do_solve_alphas(c, ctl, deltaT);
    API_END(c)
}

int tcif_get_rho_phi(tcif_ctx *c, double *internal, double *boundary) {
    API_BEGIN(c)
    if (!c->rhoPhi) failf("tcif_get_rho_phi: rhoPhi has not been computed (tcif_solve_alphas)");
    if (internal) down_faces(c, internal, c->rhoPhi);
    if (boundary && c->nBF) CUDA_OK(cudaMemcpyAsync(boundary, c->rhoPhib, sizeof(double) * c->nBF, cudaMemcpyDefault, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END(c)
}

int tcif_alpha_stats(tcif_ctx *c, int phase, double out3[3]) {
    API_BEGIN(c)
    check_phase(c, phase);
    int nPart = std::min(1024, cdiv(c->nC, TCIF_TPB));
    timed(c, T_STATS, 16.0 * c->nC, [&] {
        k_stats_partial<<<nPart, TCIF_TPB, 0, c->stream>>>(c->nC, c->dm.V, c->alpha[phase], c->statsPart);
        k_stats_final<<<1, 32, 0, c->stream>>>(nPart, c->statsPart, c->statsOut);
        c->launches += 2;
    });
    CUDA_OK(cudaMemcpyAsync(c->h_small, c->statsOut, sizeof(double) * 3, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    double v[3] = {c->h_small[0], c->h_small[1], c->h_small[2]};
    int ops[3] = {RED_SUM, RED_MIN, RED_MAX};
    allreduce_host(c, v, ops, 3);
    out3[0] = v[0]; out3[1] = v[1]; out3[2] = v[2];
    peer_check(c);
    API_END(c)
}

// ---- tcif_alpha_stats without any allocation (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The scratch of the two-level reduction (one (sum, min, max) triple per CTA of k_stats_partial, then the result
// triple) lives in this file-scope device array instead of the context's cudaMalloc'ed buffers: 24.6 KB that every GPU
// which loads the library holds until exit.  In host code `g_statsStaticScratch` is a host-side stand-in for the
// symbol, NOT a device address: the address a kernel can use comes from cudaGetSymbolAddress, resolved once per
// device with that device current (stats_static_scratch()).
//   * every call on a device shares the array and finds the previous call's contents in it: nothing has to start at
//     zero (the first kernel overwrites all nPart triples the second one reads, the second overwrites the result),
//     so nothing is cleared;
//   * for the same reason two calls that overlap on one GPU (different streams / contexts / host threads) RACE on it.
constexpr int kStatsStaticMaxPart = 1024;                       // CTAs of k_stats_partial = triples of scratch
constexpr int64_t kStatsStaticMaxCells = TCIF_ALPHA_STATS_STATIC_MAX_CELLS;
static __device__ __align__(256) double g_statsStaticScratch[3 * kStatsStaticMaxPart + 32];

static double *stats_static_scratch(int device) {               // caller has made `device` current
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static double *addr[kMaxDev] = {};
    if (device < 0 || device >= kMaxDev) failf("tcif_alpha_stats_static: device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        double *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_statsStaticScratch);
        if (e != cudaSuccess || !p) failf("tcif_alpha_stats_static: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
        addr[device] = p;
    }
    return addr[device];
}

int tcif_alpha_stats_static(tcif_ctx *c, int phase, double out3[3]) {
    API_BEGIN(c)
    check_phase(c, phase);
    if (!out3) failf("tcif_alpha_stats_static: null output");
    if ((int64_t)c->nC > kStatsStaticMaxCells)
        failf("tcif_alpha_stats_static: %d cells exceed the maximum of %lld this entry point is built for "
              "(TCIF_ALPHA_STATS_STATIC_MAX_CELLS); use tcif_alpha_stats", c->nC, (long long)kStatsStaticMaxCells);
    const int nPart = std::min(kStatsStaticMaxPart, cdiv(c->nC, TCIF_TPB));     // same tree as tcif_alpha_stats: same bits
    double *scr = stats_static_scratch(c->device);
    double *res = scr + 3 * kStatsStaticMaxPart;
    k_stats_partial<<<nPart, TCIF_TPB, 0, c->stream>>>(c->nC, c->dm.V, c->alpha[phase], scr);
    k_stats_final<<<1, 32, 0, c->stream>>>(nPart, scr, res);
    c->launches += 2;
    CUDA_OK(cudaMemcpyAsync(c->h_small, res, sizeof(double) * 3, cudaMemcpyDeviceToHost, c->stream));   // h_small: pinned at tcif_create
    CUDA_OK(cudaStreamSynchronize(c->stream));
    CUDA_OK(cudaGetLastError());
    out3[0] = c->h_small[0]; out3[1] = c->h_small[1]; out3[2] = c->h_small[2];
    API_END(c)
}

// ---- measurement ------------------------------------------------------------------------------
int tcif_timers_enable(tcif_ctx *c, int on) {
    API_BEGIN(c)
    if (!on && c->timersOn) timers_collect(c);
    c->timersOn = on != 0;
    API_END(c)
}
int tcif_timers_reset(tcif_ctx *c) {
    API_BEGIN(c)
    timers_collect(c);
    for (int i = 0; i < T_COUNT; ++i) { c->tMs[i] = 0; c->tBytes[i] = 0; c->tLaunch[i] = 0; }
    API_END(c)
}
int tcif_timer_count(void) { return T_COUNT; }
const char *tcif_timer_name(int id) { return (id >= 0 && id < T_COUNT) ? kTimerNames[id] : ""; }
int tcif_timers_read(tcif_ctx *c, double *ms, int64_t *launches, double *bytes) {
    API_BEGIN(c)
    timers_collect(c);
    for (int i = 0; i < T_COUNT; ++i) {
        if (ms) ms[i] = c->tMs[i];
        if (launches) launches[i] = c->tLaunch[i];
        if (bytes) bytes[i] = c->tBytes[i];
    }
    API_END(c)
}
int64_t tcif_launch_count(const tcif_ctx *c) { return c ? c->launches : 0; }
int64_t tcif_device_bytes(const tcif_ctx *c) { return c ? c->devBytes : 0; }
int tcif_debug_guard_check(tcif_ctx *c, int64_t *nZones, int64_t *nCorrupt) {
    API_BEGIN(c)
    if (!nZones || !nCorrupt) failf("tcif_debug_guard_check: null output");
    *nZones = 0; *nCorrupt = 0;
    if (!c->guardOn) failf("tcif_debug_guard_check: the context was not created under TCIF_GUARD=1");
    CUDA_OK(cudaDeviceSynchronize());
    std::vector<unsigned char> h;
    for (const auto &g : c->guards) {
        const size_t padded = (g.bytes + 255) / 256 * 256;
        const size_t tail = padded - g.bytes + tcif_ctx::kGuard;
        h.resize(tcif_ctx::kGuard + tail);
        CUDA_OK(cudaMemcpy(h.data(), g.base, tcif_ctx::kGuard, cudaMemcpyDeviceToHost));
        CUDA_OK(cudaMemcpy(h.data() + tcif_ctx::kGuard, g.base + tcif_ctx::kGuard + g.bytes, tail, cudaMemcpyDeviceToHost));
        bool lo = true, hi = true;
        for (size_t i = 0; i < tcif_ctx::kGuard; ++i) lo = lo && h[i] == tcif_ctx::kGuardByte;
        for (size_t i = tcif_ctx::kGuard; i < h.size(); ++i) hi = hi && h[i] == tcif_ctx::kGuardByte;
        *nZones += 2; *nCorrupt += (lo ? 0 : 1) + (hi ? 0 : 1);
    }
    API_END(c)
}
int tcif_comm_transport(const tcif_ctx *c) { return !c ? 0 : c->group ? 1 : c->nccl ? (c->peerOn ? 3 : 2) : 0; }
int tcif_mesh_info(const tcif_ctx *c, int32_t out[8]) {
    if (!c || !out) return 1;
    out[0] = c->Ko; out[1] = c->Kn; out[2] = kshape(c); out[3] = c->tileStrips; out[4] = c->dm.pfTiles;
    out[5] = c->nBndCells; out[6] = c->nBndCells; out[7] = c->nCoupled;
    return 0;
}

}  // extern "C"
