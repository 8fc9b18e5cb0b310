// engine.cu -- host side of the B200 bundle-adjustment engine and its C ABI
// (include/cugo_b200.h).  One Solver = one device, one stream.  The product path
// has no CPU fallback: every numeric stage is a CUDA kernel in kernels_*.cuh.
#include "../../include/cugo_b200.h"

#include <cub/cub.cuh>
#include <cuda_runtime.h>
#include <dlfcn.h>
// NCCL is loaded with dlopen at run time (several ranks only); a single-GPU build does not need its headers: the
// handful of types / enumerators used below are declared here when <nccl.h> is absent (values as in nccl.h 2.x).
#if __has_include(<nccl.h>) && !defined(CUGO_B200_NO_NCCL_HEADER)
#include <nccl.h>
#else
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclSuccess = 0 } ncclResult_t;
typedef enum { ncclUint8 = 1, ncclInt32 = 2, ncclUint64 = 5, ncclFloat64 = 8 } ncclDataType_t;
typedef enum { ncclSum = 0 } ncclRedOp_t;
#endif

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "engine.h"
#include "kernels_chol3.cuh"
#include "kernels_edge.cuh"
#include "kernels_lm.cuh"
#include "kernels_peer.cuh"
#include "kernels_poseonly.cuh"
#include "kernels_schur.cuh"
#include "kernels_struct.cuh"
#include "kernels_sym.cuh"
#include "symbolic.h"

namespace cugo_b200 {

struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct ArgError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct StateError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct UnsupportedError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            throw ::cugo_b200::CudaError(std::string(#call) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + \
                            ":" + std::to_string(__LINE__) + ")");                                 \
    } while (0)

static inline int divUp(long long a, int b) { return (int)((a + b - 1) / b); }

// interleaved measurements [n x dim] / informations of one edge set -> the raw SoA arrays
__global__ void k_scatter_meas(const double* meas, const double* omega, int n, int dim, int nomega, double* m0,
                               double* m1, double* m2, double* w) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    m0[i] = meas[(size_t)i * dim]; m1[i] = meas[(size_t)i * dim + 1];
    m2[i] = dim == 3 ? meas[(size_t)i * 3 + 2] : 0.0;
    w[i] = nomega > 1 ? omega[i] : omega[0];
}
__global__ void k_fill_index(int* a, int n, int base, int step) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = base + step * i;
}
__global__ void k_int_to_flag(const int* in, int n, uint8_t* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i] ? 1 : 0;
}

// NCCL is bound at run time (dlopen) the first time a communicator is asked for: a single-GPU process
// never loads it, and inside a torch.distributed process the already-loaded libnccl.so.2 is reused.
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string& err) {
        if (lib) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) { lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
        if (!lib) { err = std::string("cannot load libnccl: ") + dlerror(); return false; }
        bool ok = true;
        auto sym = [&](const char* n) { void* p = dlsym(lib, n); if (!p) ok = false; return p; };
        GetUniqueId = (decltype(GetUniqueId))sym("ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))sym("ncclCommInitRank");
        CommDestroy = (decltype(CommDestroy))sym("ncclCommDestroy");
        AllReduce = (decltype(AllReduce))sym("ncclAllReduce");
        AllGather = (decltype(AllGather))sym("ncclAllGather");
        Broadcast = (decltype(Broadcast))sym("ncclBroadcast");
        GroupStart = (decltype(GroupStart))sym("ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))sym("ncclGroupEnd");
        GetErrorString = (decltype(GetErrorString))sym("ncclGetErrorString");
        if (!ok) { err = "libnccl lacks an expected symbol"; lib = nullptr; }
        return ok;
    }
};
static NcclApi g_nccl;
#define NK(call)                                                                                   \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess)                                                                     \
            throw ::cugo_b200::CudaError(std::string(#call) + ": " + g_nccl.GetErrorString(r_));   \
    } while (0)

// Grow-only device buffer.  Allocations are stream-ordered on the stream of the solver in use (Solver::use() sets
// t_allocStream) out of the device's default memory pool, whose release threshold cugo_b200_create raises: the ~100
// buffers of a structure build then cost microseconds each, and a solver created after another one was destroyed reuses
// its memory (a cold build_structure used to spend a third of its time in cudaMalloc).
static thread_local cudaStream_t t_allocStream = nullptr;
template <typename T>
struct DBuf {
    T* p = nullptr;
    size_t cap = 0;
    ~DBuf() { if (p) cudaFree(p); }          // synchronising free: also valid for stream-ordered allocations
    T* alloc(size_t n) {
        if (n > cap) {
            if (p) { if (t_allocStream) CK(cudaFreeAsync(p, t_allocStream)); else CK(cudaFree(p)); }
            p = nullptr;
            cap = 0;
            const size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
            if (t_allocStream) CK(cudaMallocAsync(&p, bytes, t_allocStream)); else CK(cudaMalloc(&p, bytes));
            cap = std::max<size_t>(n, 1);
        }
        return p;
    }
    operator T*() const { return p; }
};

struct HostSet {
    int dim = 0, n = 0, nomega = 0, ncam = 0, rk = 0;   // dim: the edge kind (device_math.cuh: 2 mono, 3 stereo, 4 depth)
    int cols() const { return dim == 2 ? 2 : 3; }       // measurement columns
    double delta = 1.0;
    std::vector<int> iP, iL, out;
    std::vector<double> meas, omega, cam;
    bool stale = false;      // measurements / informations were last updated on the device only (refreshHostSets)
    double outThr = 0.0;     // EdgeSet::setOutlierThreshold (0 = no gating)
    int outCount = 0;        // EdgeSet::getOutlierCount: edges of the set flagged so far
};

// one PlaneEdgeSet / LineEdgeSet of the reference (include/icp_types.h): pose-only edges, kept apart from the BA sets
// (own id range, cugo_b200.h: CUGO_B200_ICP_SET_BASE) because they never enter the landmark-major structure
constexpr int ICP_SET_BASE = 1024;
struct IcpHostSet {
    int kind = EDGE_PLANE, n = 0, nomega = 1, rk = 0;
    double delta = 1.0;
    int cols() const { return kind == EDGE_PLANE ? 7 : 10; }
    // The caller's arrays go straight to the device when the set is added (no host copy: an ICP caller re-matches and
    // re-adds its sets before every solve, so this IS its hot path -- 1.5 M matches are 100 MB).
    std::unique_ptr<DBuf<int>> iP;
    std::unique_ptr<DBuf<double>> meas, omega;
};

struct Solver {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr, side2 = nullptr;   // low-priority branches of the trial graph: k_chol_zinv beside the factorisation
    cudaEvent_t evFork = nullptr, evJoin = nullptr, evJoin2 = nullptr;   // (two branches, levels alternate: one chain of 13 x 32 us ended after the forward pass)
    int panelRB = 0;               // option "panel_rb": block rows of L21 per panel CTA; 0 = per tree level (planPanelSlabs), else 1..CHOL_RB
    bool bwdSplit = true;          // option "bwd_split" = 0: one CTA per supernode in the backward solve
    bool zinvOn = true;            // option "zinv" = 0: backward solves by blocked substitution on every level
    std::string err;
    // host copies of the problem
    int nP_all = 0, nP = 0, nL_all = 0, nL = 0;
    bool haveVertices = false;
    std::vector<HostSet> sets;
    bool structDirty = true, systemBuilt = false, solved = false;
    // SURVEY.md 8(f) row 4 (kernels_poseonly.cuh): ICP edge sets, and graphs without free landmarks
    std::vector<IcpHostSet> icpSets;
    bool poseOnly = false;         // nL == 0: no Schur complement, Hpp is block diagonal
    long nIcp = 0;
    IcpDev ic{};
    DBuf<int> i_rIP, i_src, i_iP, i_posePtr, i_sliceCnt, i_slicePtr, i_keyOut;
    DBuf<int2> i_units;
    DBuf<uint8_t> i_rset, i_set;
    DBuf<double> i_rw, i_w, i_rpay, i_pay, i_part;
    int nHpl_ = 0, nChunks_ = 0;   // sizes set by buildEdgeOrders()
    int addIcpEdgeSet(int kind, int n, const int* iP, const double* meas, const double* omega, int nomega, int rk, double delta);
    void refreshIcpParams();
    void buildEdgeOrders();
    void buildSchurStructure(int TB, int gE, int gE1, int nHpl, int nChunks, bool perEdgeCam, bool anyOut);
    bool lapOn = false;
    std::chrono::steady_clock::time_point lapPrev;
    void lap(const char* what);
    void buildIcp();
    void finishPoseOnly();
    void enqueueErrors(int which, int guarded, bool viaLaunch);
    // device state
    Dev d{};
    Chol c{};
    SetParams sp{};
    long E = 0;
    int nmulRef = 0, nnzSymm = 0, maxRt = 0;
    SymCounts counts{};
    std::vector<int> h_levelPtr, h_trsmPtr, h_updPtr, h_levMaxNt;
    int h_nchild = 0;              // non-root supernodes
    // buffers
    DBuf<double> pose0, pose1, poseBak, lm0, lm1, lmBak;
    DBuf<int> r_iP, r_iL, r_cam;
    DBuf<double> r_m0, r_m1, r_m2, r_w, cams;
    DBuf<uint8_t> r_set, r_out;
    DBuf<int> s_iP, s_iL, s_cam, s_gid, s_hplPre, lmEdgePtr, chunkLm, flag, wgt, wexcl;
    DBuf<double2> s_uv, p_uv;
    DBuf<double> s_r, s_w, p_r, p_w;
    DBuf<uint8_t> s_set, s_out, p_set, p_out;
    DBuf<int> posePtr, p_iL, p_cam, p_src, pkeyIn, pkeyOut, pvalIn;
    DBuf<int> colPtr, rowInd, hplLm, edge2Hpl;
    DBuf<unsigned long long> keyA, keyB, valA, valB;
    DBuf<int> ivalA, prodCnt, prodOff, head, blkOf, bRowPtr, bColInd, bRowOf, prodPtr;
    DBuf<int2> prodIJ;
    DBuf<int> poseSliceCnt, poseSlicePtr, poseTicket;
    DBuf<int2> poseUnits;
    DBuf<double> posePart;
    int nPoseUnits = 0;
    DBuf<int> pmFlag, pmPre, poseSlotPtr, poseSlots, slotRank, prodRank, poseDeg, poseLm, rowWork, rowWorkS, rowIds, rowOrder, unitCnt, unitPtr;
    DBuf<int2> rowUnits, blkPass;
    int nRowUnits = 0, schurPasses = 1, schurUnit = 2048;
    int maxPoseDeg = 0, schurCap = 1;
    bool f32 = false;              // USE_FLOAT32 mode: Hpl blocks stored in single precision (arithmetic stays FP64)
    DBuf<float> HplF;
    DBuf<double> Hpp, Hll, Hpl, b, x, invHll, bsc, Hsc, chiPartial, scalePartial, maxDiagL, maxDiagP, chiEdge;
    DBuf<double> scalars;          // small device scratch for reductions
    DBuf<LMState> state;
    DBuf<uint8_t> cubTmp;
    // symbolic
    DBuf<int> y_adjPtr, y_adj, y_verts, y_iperm, y_mark, y_level, y_queue, y_tmp, y_stack, y_snStart, y_snParent,
        y_colToSn, y_snRowPtr, y_snLevel, y_childHead, y_childNext, y_snRows;
    DBuf<int> z_levelPtr, z_levelSn, z_childPtr, z_childIdx, z_invPtr, z_inv, z_snBlkPtr, z_asmBlk, z_asmRow,
        z_asmCol, z_trsmPtr, z_updPtr;
    DBuf<long long> z_Lptr, z_Uptr;
    DBuf<SymCounts> y_counts;
    // parallel symbolic analysis (kernels_sym.cuh)
    DBuf<int> y_vOf, y_isStart, y_symCnt, y_rcount, y_ak1, y_ak2, y_av1, y_av2;
    DBuf<int2> y_regA, y_regB;
    DBuf<int4> y_emits;
    DBuf<unsigned> y_bits;
    bool symParallel = true;       // option "parallel_symbolic" = 0: the single-thread run of symbolic.h (the twin) on the device
    void symbolicParallel(const int* gRowOfP, const int* gRowPtrP, const int* gColIndP, int P);
    DBuf<double> cL, cU, cX, cW, cZ, cPart;
    DBuf<int> cCnt, cFlag;
    DBuf<int> z_levMaxNt;
    // multi-GPU
    int nranks = 1, rank = 0;
    ncclComm_t comm = nullptr;
    DBuf<unsigned long long> lkeys, gkeysAll, gkeysSorted, gkeys;
    DBuf<int> gHead, gIdx, gRowOf, gColInd, gRowPtr, hscSlot, d_cnt;
    DBuf<double> redA, redC, HscPartBuf;
    // exchange over NVLink peer memory instead of NCCL (kernels_peer.cuh): one cudaMalloc'ed block per rank, opened by the
    // others through CUDA IPC.  Option "peer_reduce" = 0 keeps ncclAllReduce (also the fallback when IPC is not available).
    int peerMode = 1;
    bool peerOn = false;
    PeerCtx peer{};
    char* peerLocal = nullptr;
    size_t peerNred = 0;
    void peerSetup(size_t nred);
    void peerRelease();
    void smallReduce(double* buf, int n) {
        if (peerOn && n <= PEER_SMALL_MAX) k_peer_small<<<1, PEER_TPB, 0, stream>>>(peer, buf, n, d.st, 0);
        else allReduceSum(buf, (size_t)n);
    }
    // factorisation distributed by assembly-tree subtree (several ranks): supernodes on tree levels
    // >= distLc are factorised by every rank, the subtrees below are dealt to the ranks
    int distMode = 1;              // option "dist_factor": 0 never, 1 when the factor flops make it pay, 2 always
    int distLc = 0;                // 0: not distributed
    std::vector<int> h_own;        // [nsn] owner rank, -1 = every rank
    struct Bcast { size_t off, count; int root; };
    std::vector<Bcast> distBcast;  // update matrices of the subtree roots, sent by their owners before level distLc
    DBuf<int> zr_levelPtr, zr_levelSn, zr_trsmPtr, zr_updPtr, zr_colOwn;
    void planDistFactor();
    void planPanelSlabs();
    DBuf<int> zs_trsmPtr, zs_rbLev, zs_updPtr;
    std::vector<int> h_tbLev;        // update-tile height per level: CHOL_TB, or UPD_TBS where the level has few tiles
    bool schurMma = false;          // option "schur_mma" = 1: the Schur products as DMMAs (measured slower: 0.184 vs 0.165 ms)
    int updSmall = 1;               // option "upd_small" = 0: 48x48 update tiles on every level    // panel-slab task table / slab height per level when chosen per level (planPanelSlabs)
    int nblkG = 0;
    void commInit(int n, int r, const char* id);
    void allReduceSum(double* buf, size_t n) {
        NK(g_nccl.AllReduce(buf, buf, n, ncclFloat64, ncclSum, comm, stream));
    }
    void requireSingle(const char* what) {
        if (nranks > 1) throw StateError(std::string(what) + ": the staged API is single-GPU; with several ranks use optimize()");
    }
    DBuf<double> z_factStats;
    double factStats[4] = {0, 0, 0, 0};
    // graph of one trial slot
    cudaGraph_t slotGraph = nullptr;
    cudaGraphExec_t slotExec = nullptr;
    // several ranks: the FIRST trial of an optimize() also exchanges the diagonal maxima (lambda0 = tau * max diag of the
    // global system, iteration 0 only); every later trial runs the graph without that collective
    cudaGraph_t slotGraph0 = nullptr;
    cudaGraphExec_t slotExec0 = nullptr;
    bool withA = true;
    void dropGraphs() {
        if (slotExec) { cudaGraphExecDestroy(slotExec); slotExec = nullptr; }
        if (slotGraph) { cudaGraphDestroy(slotGraph); slotGraph = nullptr; }
        if (slotExec0) { cudaGraphExecDestroy(slotExec0); slotExec0 = nullptr; }
        if (slotGraph0) { cudaGraphDestroy(slotGraph0); slotGraph0 = nullptr; }
    }
    int kernelsPerSlot = 0;
    // per-kernel profiling of one trial slot (events on the launching stream)
    bool profiling = false;
    std::vector<cudaEvent_t> profEvents;
    std::vector<std::string> profNames;
    void mark(const char* name) {
        kernelsPerSlot++;
        if (!profiling) return;
        cudaEvent_t e;
        CK(cudaEventCreate(&e));
        CK(cudaEventRecord(e, stream));
        profEvents.push_back(e);
        profNames.push_back(name);
    }
    // Launch of a trial kernel with programmatic stream serialisation (engine.h: pdl_wait / pdl_trigger);
    // option "pdl" = 0 falls back to plain stream order.  Captured into the trial graph as a programmatic edge.
    bool pdl = true;
    bool panelDC = true;           // panel kernel with the decoupled diagonal chain (option "panel_dc" = 0: CTA-wide barriers)
    bool fuseBuild = true;         // both passes of the system build in one launch (option "fuse_build" = 0: two)
    template <typename... KA, typename... A>
    void launch(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, A... args) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = (pdl && !profiling) ? 1 : 0;
        CK(cudaLaunchKernelEx(&cfg, kern, KA(args)...));
    }
    int profileSlot(int cap, char* names, int nameLen, double* ms);
    void updateMeasurements(int setId, const double* meas, const double* omega, int nomega);
    void refreshHostSets();
    DBuf<double> u_meas, u_w;      // staging of one set's interleaved measurements / informations
    bool perEdgeCam_ = false, anyOut_ = false;
    LMState* h_state = nullptr;    // pinned
    double* h_scalars = nullptr;   // pinned
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double lastOptimizeMs = 0.0;
    long lastKernels = 0;

    ~Solver() {
        cudaSetDevice(device);
        dropGraphs();
        if (h_state) cudaFreeHost(h_state);
        if (h_scalars) cudaFreeHost(h_scalars);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        peerRelease();
        if (comm && g_nccl.CommDestroy) g_nccl.CommDestroy(comm);
        if (evFork) cudaEventDestroy(evFork);
        if (evJoin) cudaEventDestroy(evJoin);
        if (evJoin2) cudaEventDestroy(evJoin2);
        if (side) cudaStreamDestroy(side);
        if (side2) cudaStreamDestroy(side2);
        if (stream) cudaStreamDestroy(stream);
    }

    void use() { CK(cudaSetDevice(device)); t_allocStream = stream; }
    void uploadPoses(const double* poses7);
    void setVertices(int nPa, int nPact, const double* poses7, int nLa, int nLact, const double* lms);
    void setEstimates(const double* poses7, const double* lms);
    int addEdgeSet(int dim, int n, const int* iP, const int* iL, const double* meas, const double* omega,
                   int nomega, const double* cam, int ncam, int rk, double delta);
    void refreshSetParams();
    void buildStructure();
    void ensureStructure() { if (structDirty) buildStructure(); }
    void captureSlot();
    void enqueueBuild();
    void enqueueSolve();
    void enqueueSlot();
    double computeErrors();
    void buildSystem();
    double maxDiagonal();
    bool solve();
    void update();
    double computeScale(double lambda);
    // outlier gating on the device (reference: computeOutliersKernel + EdgeSet::updateEdges)
    bool gating() const { for (auto& hs : sets) if (hs.outThr > 0) return true; return false; }
    bool maskLive = false;         // the device mask holds flags the host sets do not (syncOutliersToHost)
    DBuf<int> outCnt;
    void setOutlierThreshold(int setId, double thr);
    void updateEdges();
    void syncOutliersToHost();
    void getOutliers(int setId, int* mask, int* outliers, int* inliers);
    int optimize(int niter, int cap, double* chi2, double* lambda, int* trials);
    void getEstimates(double* poses7, double* lms);
    long getStructure(int which, int* out, long cap);
    long getValues(int which, double* out, long cap);
    double stagedLambda = 0.0;
};

void Solver::uploadPoses(const double* poses7) {
    std::vector<double> p8((size_t)nP_all * POSE_STRIDE, 0.0);
    for (int i = 0; i < nP_all; i++)
        for (int k = 0; k < 7; k++) p8[(size_t)i * POSE_STRIDE + k] = poses7[(size_t)i * 7 + k];
    CK(cudaMemcpyAsync(pose0.p, p8.data(), p8.size() * sizeof(double), cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(pose1.p, pose0.p, p8.size() * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    CK(cudaStreamSynchronize(stream));
}

void Solver::setVertices(int nPa, int nPact, const double* poses7, int nLa, int nLact, const double* lms) {
    if (nPa < 0 || nLa < 0 || nPact < 0 || nLact < 0 || nPact > nPa || nLact > nLa || !poses7 || (!lms && nLa > 0))
        throw ArgError("set_vertices: bad sizes or null pointers");
    use();
    if (nPa != nP_all || nPact != nP || nLa != nL_all || nLact != nL) structDirty = true;
    nP_all = nPa; nP = nPact; nL_all = nLa; nL = nLact;
    pose0.alloc((size_t)nP_all * POSE_STRIDE); pose1.alloc((size_t)nP_all * POSE_STRIDE);
    poseBak.alloc((size_t)nP_all * POSE_STRIDE);
    lm0.alloc((size_t)nL_all * 3); lm1.alloc((size_t)nL_all * 3); lmBak.alloc((size_t)nL_all * 3);
    haveVertices = true;
    setEstimates(poses7, lms);
}

void Solver::setEstimates(const double* poses7, const double* lms) {
    if (!haveVertices) throw StateError("set_estimates before set_vertices");
    use();
    uploadPoses(poses7);
    if (nL_all > 0) {       // one trip over PCIe; the trial buffer is filled on the device
        CK(cudaMemcpyAsync(lm0.p, lms, (size_t)nL_all * 3 * sizeof(double), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(lm1.p, lm0.p, (size_t)nL_all * 3 * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    }
    state.alloc(1);
    // estimates live in buffer 0 again
    LMState zero{};
    zero.nu = 2.0; zero.needBuild = 1;
    CK(cudaMemcpyAsync(state.p, &zero, sizeof(LMState), cudaMemcpyHostToDevice, stream));
    CK(cudaStreamSynchronize(stream));
    d.pose[0] = pose0; d.pose[1] = pose1; d.lm[0] = lm0; d.lm[1] = lm1; d.st = state;
    systemBuilt = false; solved = false;
}

int Solver::addEdgeSet(int dim, int n, const int* iP, const int* iL, const double* meas, const double* omega,
                       int nomega, const double* cam, int ncam, int rk, double delta) {
    if (!haveVertices) throw StateError("add_edge_set before set_vertices");
    if ((dim != EDGE_MONO && dim != EDGE_STEREO && dim != EDGE_DEPTH) || n < 0)
        throw ArgError("add_edge_set: dim must be 2 (mono), 3 (stereo) or 4 (depth), n >= 0");
    if (n > 0 && (!iP || !iL || !meas || !omega || !cam)) throw ArgError("add_edge_set: null pointer");
    if ((nomega != 1 && nomega != n) || (ncam != 1 && ncam != n)) throw ArgError("add_edge_set: nomega/ncam must be 1 or n");
    if (rk < 0 || rk > 3) throw ArgError("add_edge_set: unknown robust kernel");
    if ((int)sets.size() >= MAX_SETS) throw ArgError("add_edge_set: too many edge sets");
    for (int i = 0; i < n; i++) {
        if (iP[i] < 0 || iP[i] >= nP_all || iL[i] < 0 || iL[i] >= nL_all) throw ArgError("add_edge_set: vertex index out of range");
        if (iP[i] >= nP && iL[i] >= nL) throw ArgError("add_edge_set: edge with both vertices fixed is not active");
    }
    HostSet s;
    s.dim = dim; s.n = n; s.nomega = nomega; s.ncam = ncam; s.rk = rk; s.delta = delta;
    s.iP.assign(iP, iP + n); s.iL.assign(iL, iL + n);
    s.meas.assign(meas, meas + (size_t)n * s.cols());
    if (n > 0) {
        s.omega.assign(omega, omega + nomega);
        s.cam.assign(cam, cam + (size_t)ncam * 5);
    } else {
        s.omega.assign(1, 1.0);
        s.cam.assign(5, 0.0);
        s.nomega = s.ncam = 1;
    }
    sets.push_back(std::move(s));
    structDirty = true;
    return (int)sets.size() - 1;
}

void Solver::refreshSetParams() {
    memset(&sp, 0, sizeof(sp));
    for (size_t i = 0; i < sets.size(); i++) {
        sp.dim[i] = sets[i].dim; sp.rkType[i] = sets[i].rk; sp.rkDelta[i] = sets[i].delta;
        for (int k = 0; k < 5; k++) sp.cam[i][k] = sets[i].cam[k];
    }
}

// ---------------------------------------------------------------------------
// structure build: everything below is device work except size read-backs
// ---------------------------------------------------------------------------
template <typename K, typename V>
static void sortPairs(Solver& S, const K* kin, K* kout, const V* vin, V* vout, int n, int endBit) {
    size_t bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, n, 0, endBit, S.stream));
    S.cubTmp.alloc(bytes);
    CK(cub::DeviceRadixSort::SortPairs(S.cubTmp.p, bytes, kin, kout, vin, vout, n, 0, endBit, S.stream));
}
static void exclusiveScan(Solver& S, const int* in, int* out, int n) {
    size_t bytes = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, n, S.stream));
    S.cubTmp.alloc(bytes);
    CK(cub::DeviceScan::ExclusiveSum(S.cubTmp.p, bytes, in, out, n, S.stream));
}
static int bitsFor(unsigned long long maxv) {
    int b = 1;
    while (b < 64 && (maxv >> b)) b++;
    return b;
}
static int readInt(Solver& S, const int* dptr) {
    int v = 0;
    CK(cudaMemcpyAsync(&v, dptr, sizeof(int), cudaMemcpyDeviceToHost, S.stream));
    CK(cudaStreamSynchronize(S.stream));
    return v;
}

void Solver::lap(const char* what) {
    if (!lapOn) return;
    cudaStreamSynchronize(stream);
    const auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "[cugo_b200] build_structure: %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t - lapPrev).count());
    lapPrev = t;
}

void Solver::buildStructure() {
    if (!haveVertices) throw StateError("build_structure before set_vertices");
    use();
    lapOn = getenv("CUGO_B200_DEBUG") != nullptr;
    lapPrev = std::chrono::steady_clock::now();
    refreshHostSets();
    syncOutliersToHost();
    if (nP == 0) throw UnsupportedError("graph has no free poses");
    E = 0; nIcp = 0;
    bool perEdgeCam = false, anyOut = false;
    for (auto& s : sets) {
        E += s.n;
        if (s.ncam > 1) perEdgeCam = true;
        if (!s.out.empty() || s.outThr > 0) anyOut = true;
    }
    for (auto& s : icpSets) nIcp += s.n;
    // no free landmarks: the reference's doSchur_ == false path (src/block_solver.cpp:80,226-234) -- kernels_poseonly.cuh
    poseOnly = (nL == 0);
    if (E == 0 && nIcp == 0) throw StateError("build_structure: no edges");
    if (!poseOnly && E == 0) throw StateError("build_structure: free landmarks but no edge observes them");
    if ((poseOnly || nIcp > 0) && nranks > 1)
        throw UnsupportedError("pose-only graphs and ICP edge sets run on one GPU (nothing to shard: they have no landmarks)");
    if (E > 0x7fffffff / 18 || nIcp > 0x7fffffff / ICP_COLS) throw UnsupportedError("edge count exceeds 32-bit block indexing");
    perEdgeCam_ = perEdgeCam; anyOut_ = anyOut;
    refreshSetParams();
    dropGraphs();
    nHpl_ = 0; nChunks_ = 0; nPoseUnits = 0;
    if (E > 0) buildEdgeOrders();
    if (nIcp > 0) buildIcp();
    if (poseOnly) { finishPoseOnly(); return; }
    const int TB = 256;
    const int gE = divUp(E, TB), gE1 = divUp(E + 1, TB);
    const int nHpl = nHpl_, nChunks = nChunks_;
    buildSchurStructure(TB, gE, gE1, nHpl, nChunks, perEdgeCam, anyOut);
}

// raw edge arrays -> device, landmark-major order + Hpl CSC + CTA chunks, pose-major copy + slices (E > 0)
void Solver::buildEdgeOrders() {
    const bool perEdgeCam = perEdgeCam_, anyOut = anyOut_;
    // ---- raw edge arrays (input order) -> device.  The sets' arrays go up as the caller gave them (indices straight
    // into place, interleaved measurements through a staging buffer) and are split into the SoA arrays on the device.
    {
        r_iP.alloc(E); r_iL.alloc(E); r_m0.alloc(E); r_m1.alloc(E); r_m2.alloc(E); r_w.alloc(E); r_set.alloc(E);
        size_t nMeas = 0, nOm = 0, nCamRows = sets.size();
        for (auto& s : sets) { nMeas += (size_t)s.n * s.cols(); nOm += s.nomega; if (s.ncam > 1) nCamRows += s.n; }
        u_meas.alloc(std::max<size_t>(nMeas, 1)); u_w.alloc(std::max<size_t>(nOm, 1));
        if (perEdgeCam) { r_cam.alloc(E); cams.alloc(nCamRows * 5); }
        if (anyOut) { r_out.alloc(E); CK(cudaMemsetAsync(r_out.p, 0, E, stream)); }
        size_t o = 0, om = 0, oo = 0, camRow = sets.size();
        for (size_t si = 0; si < sets.size(); si++) {
            const HostSet& s = sets[si];
            if (perEdgeCam) CK(cudaMemcpyAsync(cams.p + si * 5, s.cam.data(), 5 * sizeof(double), cudaMemcpyHostToDevice, stream));
            if (s.n == 0) continue;
            const int g = divUp(s.n, 256);
            CK(cudaMemcpyAsync(r_iP.p + o, s.iP.data(), (size_t)s.n * sizeof(int), cudaMemcpyHostToDevice, stream));
            CK(cudaMemcpyAsync(r_iL.p + o, s.iL.data(), (size_t)s.n * sizeof(int), cudaMemcpyHostToDevice, stream));
            CK(cudaMemcpyAsync(u_meas.p + om, s.meas.data(), (size_t)s.n * s.cols() * sizeof(double), cudaMemcpyHostToDevice, stream));
            CK(cudaMemcpyAsync(u_w.p + oo, s.omega.data(), (size_t)s.nomega * sizeof(double), cudaMemcpyHostToDevice, stream));
            k_scatter_meas<<<g, 256, 0, stream>>>(u_meas.p + om, u_w.p + oo, s.n, s.cols(), s.nomega, r_m0.p + o, r_m1.p + o,
                                                  r_m2.p + o, r_w.p + o);
            CK(cudaMemsetAsync(r_set.p + o, (int)si, s.n, stream));
            if (perEdgeCam) {
                if (s.ncam > 1) {
                    CK(cudaMemcpyAsync(cams.p + camRow * 5, s.cam.data(), (size_t)s.n * 5 * sizeof(double), cudaMemcpyHostToDevice, stream));
                    k_fill_index<<<g, 256, 0, stream>>>(r_cam.p + o, s.n, (int)camRow, 1);
                    camRow += s.n;
                } else {
                    k_fill_index<<<g, 256, 0, stream>>>(r_cam.p + o, s.n, (int)si, 0);
                }
            }
            if (anyOut && !s.out.empty()) {
                ivalA.alloc(std::max<size_t>(E, 1));
                CK(cudaMemcpyAsync(ivalA.p, s.out.data(), (size_t)s.n * sizeof(int), cudaMemcpyHostToDevice, stream));
                k_int_to_flag<<<g, 256, 0, stream>>>(ivalA.p, s.n, r_out.p + o);
            }
            o += s.n; om += (size_t)s.n * s.cols(); oo += s.nomega;
        }
    }
    lap("raw edge arrays -> device");
    RawEdges raw{(int)E, r_iP, r_iL, r_m0, r_m1, r_m2, r_w, r_set, perEdgeCam ? r_cam.p : nullptr,
                 anyOut ? r_out.p : nullptr};
    const int TB = 256;
    const int gE = divUp(E, TB), gE1 = divUp(E + 1, TB);

    // ---- landmark-major order
    keyA.alloc(E); keyB.alloc(E); ivalA.alloc(E); s_gid.alloc(E);
    k_make_lm_keys<<<gE, TB, 0, stream>>>(raw, nP_all, keyA, ivalA);
    sortPairs(*this, keyA.p, keyB.p, ivalA.p, s_gid.p, (int)E,
              bitsFor((unsigned long long)nL_all * (unsigned long long)nP_all));
    s_iP.alloc(E); s_iL.alloc(E); s_uv.alloc(E); s_r.alloc(E); s_w.alloc(E); s_set.alloc(E);
    if (perEdgeCam) s_cam.alloc(E);
    if (anyOut) s_out.alloc(E);
    flag.alloc(E + 1); s_hplPre.alloc(E + 1);
    k_gather_lm<<<gE1, TB, 0, stream>>>(raw, s_gid, nP, nL, s_iP, s_iL, s_uv, s_r, s_w, s_set,
                                        perEdgeCam ? s_cam.p : nullptr, anyOut ? s_out.p : nullptr, flag);
    exclusiveScan(*this, flag.p, s_hplPre.p, (int)E + 1);
    lmEdgePtr.alloc(nL_all + 1);
    k_lower_bounds<<<divUp(nL_all + 1, TB), TB, 0, stream>>>(s_iL, (int)E, nL_all, lmEdgePtr);
    const int nHpl = readInt(*this, s_hplPre.p + E);
    colPtr.alloc(nL + 1); rowInd.alloc(nHpl); hplLm.alloc(nHpl); edge2Hpl.alloc(E);
    k_fill_hpl<<<std::max(gE, divUp(nL + 1, TB)), TB, 0, stream>>>((int)E, s_iP, s_iL, s_gid, flag, s_hplPre,
                                                                     lmEdgePtr, nL, colPtr, rowInd, hplLm, edge2Hpl);
    // ---- CTA chunks of the landmark-major kernels
    wgt.alloc(nL_all + 1); wexcl.alloc(nL_all + 1);
    k_lm_weights<<<divUp(nL_all + 1, TB), TB, 0, stream>>>(lmEdgePtr, nL_all, wgt);
    exclusiveScan(*this, wgt.p, wexcl.p, nL_all + 1);
    const int lastW = readInt(*this, wexcl.p + nL_all - 1);
    const int nChunks = lastW / LM_CHUNK + 1;
    chunkLm.alloc(nChunks + 1);
    k_chunk_bounds<<<divUp(nL_all + 1, TB), TB, 0, stream>>>(wexcl, nL_all, nChunks, chunkLm);

    lap("landmark-major order, Hpl CSC, chunks");
    // ---- pose-major copy
    pkeyIn.alloc(E); pkeyOut.alloc(E); pvalIn.alloc(E); p_src.alloc(E);
    k_make_pose_keys<<<gE, TB, 0, stream>>>(s_iP, (int)E, pkeyIn, pvalIn);
    sortPairs(*this, pkeyIn.p, pkeyOut.p, pvalIn.p, p_src.p, (int)E, bitsFor((unsigned long long)nP_all));
    p_iL.alloc(E); p_uv.alloc(E); p_r.alloc(E); p_w.alloc(E); p_set.alloc(E);
    if (perEdgeCam) p_cam.alloc(E);
    if (anyOut) p_out.alloc(E);
    k_gather_pose<<<gE, TB, 0, stream>>>((int)E, p_src, s_iL, s_uv, s_r, s_w, s_set, perEdgeCam ? s_cam.p : nullptr,
                                         anyOut ? s_out.p : nullptr, p_iL, p_uv, p_r, p_w, p_set,
                                         perEdgeCam ? p_cam.p : nullptr, anyOut ? p_out.p : nullptr);
    posePtr.alloc(nP_all + 1);
    k_lower_bounds<<<divUp(nP_all + 1, TB), TB, 0, stream>>>(pkeyOut, (int)E, nP_all, posePtr);
    // pose-major pass in slices (engine.h: POSE_SLICE)
    poseSliceCnt.alloc(nP + 1); poseSlicePtr.alloc(nP + 1); poseTicket.alloc(std::max(nP, 1));
    k_pose_slice_count<<<divUp(nP + 1, TB), TB, 0, stream>>>(posePtr, nP, poseSliceCnt);
    exclusiveScan(*this, poseSliceCnt.p, poseSlicePtr.p, nP + 1);
    nPoseUnits = readInt(*this, poseSlicePtr.p + nP);
    poseUnits.alloc(std::max(nPoseUnits, 1)); posePart.alloc((size_t)std::max(nPoseUnits, 1) * 27);
    k_pose_units_fill<<<divUp(std::max(nP, 1), TB), TB, 0, stream>>>(poseSlicePtr, nP, poseUnits);
    CK(cudaMemsetAsync(poseTicket.p, 0, sizeof(int) * std::max(nP, 1), stream));

    lap("pose-major order, slices");
    nHpl_ = nHpl; nChunks_ = nChunks;
}

// Hsc pattern, product lists, symbolic analysis, system buffers, trial graph (graphs with free landmarks)
void Solver::buildSchurStructure(const int TB, const int gE, const int gE1, const int nHpl, const int nChunks,
                                 const bool perEdgeCam, const bool anyOut) {
    // ---- Hsc block pattern + product lists
    prodCnt.alloc(nL + 1); prodOff.alloc(nL + 1);
    k_prod_counts<<<divUp(nL + 1, TB), TB, 0, stream>>>(colPtr, nL, prodCnt);
    exclusiveScan(*this, prodCnt.p, prodOff.p, nL + 1);
    const int nmul = readInt(*this, prodOff.p + nL);
    nmulRef = nmul;
    const long long nmulAllLL = (long long)nmul + nP;
    if (nmulAllLL > 0x7fffffff) throw UnsupportedError("product list exceeds 32-bit indexing");
    const int nmulAll = (int)nmulAllLL;
    keyA.alloc(nmulAll); keyB.alloc(nmulAll); valA.alloc(nmulAll); valB.alloc(nmulAll);
    k_gen_products<<<divUp(std::max(nL, nP), TB), TB, 0, stream>>>(colPtr, rowInd, prodOff, nL, nP, nmul, keyA, valA);
    sortPairs(*this, keyA.p, keyB.p, valA.p, valB.p, nmulAll, bitsFor((unsigned long long)nP * (unsigned long long)nP));
    head.alloc(nmulAll + 1); blkOf.alloc(nmulAll + 1);
    k_prod_heads<<<divUp(nmulAll + 1, TB), TB, 0, stream>>>(keyB, nmulAll, head);
    exclusiveScan(*this, head.p, blkOf.p, nmulAll + 1);
    const int nblk = readInt(*this, blkOf.p + nmulAll);
    bRowOf.alloc(nblk); bColInd.alloc(nblk); prodPtr.alloc(nblk + 1); prodIJ.alloc(nmulAll); bRowPtr.alloc(nP + 1);
    k_fill_blocks<<<divUp(nmulAll, TB), TB, 0, stream>>>(keyB, valB, head, blkOf, nmulAll, nP, nblk, bRowOf, bColInd,
                                                         prodPtr, prodIJ);
    k_lower_bounds<<<divUp(nP + 1, TB), TB, 0, stream>>>(bRowOf, nblk, nP, bRowPtr);
    nnzSymm = (2 * nblk - nP) * 36;

    lap("Hsc pattern, product lists");
    // ---- pose-major view of the Hpl blocks + product ranks (k_schur_rows)
    pmFlag.alloc(E + 1); pmPre.alloc(E + 1); poseSlotPtr.alloc(nP + 1); poseSlots.alloc(std::max(nHpl, 1));
    slotRank.alloc(std::max(nHpl, 1)); prodRank.alloc(nmulAll); poseDeg.alloc(nP + 1);
    k_pm_flags<<<gE1, TB, 0, stream>>>(p_src, flag, (int)E, pmFlag);
    exclusiveScan(*this, pmFlag.p, pmPre.p, (int)E + 1);
    k_pm_slots<<<gE, TB, 0, stream>>>(p_src, flag, s_hplPre, pmPre, pkeyOut, posePtr, (int)E, poseSlots, slotRank);
    k_pose_slot_ptr<<<divUp(nP + 1, TB), TB, 0, stream>>>(pmPre, posePtr, nP, poseSlotPtr, poseDeg);
    k_prod_rank<<<divUp(nmulAll, TB), TB, 0, stream>>>(prodIJ, slotRank, nmulAll, prodRank);
    poseLm.alloc(std::max(nHpl, 1));
    k_pose_lm<<<divUp(std::max(nHpl, 1), TB), TB, 0, stream>>>(poseSlots, hplLm, nHpl, poseLm);
    {
        size_t bytes = 0;
        CK(cub::DeviceReduce::Max(nullptr, bytes, poseDeg.p, poseDeg.p + nP, nP, stream));
        cubTmp.alloc(bytes);
        CK(cub::DeviceReduce::Max(cubTmp.p, bytes, poseDeg.p, poseDeg.p + nP, nP, stream));
        maxPoseDeg = readInt(*this, poseDeg.p + nP);
        rowWork.alloc(nP); rowWorkS.alloc(nP); rowIds.alloc(nP); rowOrder.alloc(nP);
        k_row_work<<<divUp(nP, TB), TB, 0, stream>>>(bRowPtr, prodPtr, nP, rowWork, rowIds);
        CK(cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, rowWork.p, rowWorkS.p, rowIds.p, rowOrder.p, nP, 0, 32, stream));
        cubTmp.alloc(bytes);
        CK(cub::DeviceRadixSort::SortPairsDescending(cubTmp.p, bytes, rowWork.p, rowWorkS.p, rowIds.p, rowOrder.p, nP, 0, 32, stream));
        schurCap = std::max(1, std::min(maxPoseDeg, SROW_MAX_CAP));
        schurPasses = std::max(1, (maxPoseDeg + schurCap - 1) / schurCap);
        schurUnit = std::max(SROW_UNIT, nmulAll / 1200);
        unitCnt.alloc(nP + 1); unitPtr.alloc(nP + 1);
        k_row_unit_count<<<divUp(nP + 1, TB), TB, 0, stream>>>(rowWorkS, nP, schurUnit, unitCnt);
        exclusiveScan(*this, unitCnt.p, unitPtr.p, nP + 1);
        nRowUnits = readInt(*this, unitPtr.p + nP);
        rowUnits.alloc(std::max(nRowUnits, 1));
        k_row_unit_fill<<<divUp(nP, TB), TB, 0, stream>>>(rowOrder, unitPtr, nP, rowUnits);
        if (schurPasses > 1) {
            blkPass.alloc((size_t)nblk * schurPasses);
            k_blk_pass_init<<<divUp(nblk * schurPasses, TB), TB, 0, stream>>>(nblk * schurPasses, blkPass);
            k_blk_pass<<<divUp(nmulAll, TB), TB, 0, stream>>>(prodRank, head, blkOf, nmulAll, schurCap, schurPasses, blkPass);
        }
        if (getenv("CUGO_B200_DEBUG")) fprintf(stderr, "[cugo_b200] max pose degree %d, schur stage %d blocks, %d row units\n", maxPoseDeg, schurCap, nRowUnits);
        CK(cudaFuncSetAttribute(k_schur_rows<double, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)srow_smem_bytes(schurCap)));
        CK(cudaFuncSetAttribute(k_schur_rows<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)srow_smem_bytes(schurCap)));
        CK(cudaFuncSetAttribute(k_schur_rows<double, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)srow_smem_bytes(schurCap)));
        CK(cudaFuncSetAttribute(k_schur_rows<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)srow_smem_bytes(schurCap)));
    }

    // ---- several ranks: the factorisation works on the union of the ranks' block patterns
    const int* gRowPtrP = bRowPtr;
    const int* gColIndP = bColInd;
    nblkG = nblk;
    if (nranks > 1) {
        d_cnt.alloc(nranks + 1);
        CK(cudaMemcpyAsync(d_cnt.p + nranks, &nblk, sizeof(int), cudaMemcpyHostToDevice, stream));
        NK(g_nccl.AllGather(d_cnt.p + nranks, d_cnt.p, 1, ncclInt32, comm, stream));
        std::vector<int> cnt(nranks);
        CK(cudaMemcpyAsync(cnt.data(), d_cnt.p, sizeof(int) * nranks, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        const int padded = *std::max_element(cnt.begin(), cnt.end());
        const long long nAllLL = (long long)padded * nranks;
        if (nAllLL > 0x7fffffff) throw UnsupportedError("global block pattern exceeds 32-bit indexing");
        const int nAll = (int)nAllLL;
        lkeys.alloc(padded); gkeysAll.alloc(nAll); gkeysSorted.alloc(nAll); gHead.alloc(nAll + 1); gIdx.alloc(nAll + 1);
        k_block_keys<<<divUp(padded, TB), TB, 0, stream>>>(bRowOf, bColInd, nblk, padded, (unsigned long long)nP, lkeys);
        NK(g_nccl.AllGather(lkeys.p, gkeysAll.p, padded, ncclUint64, comm, stream));
        {
            size_t bytes = 0;
            CK(cub::DeviceRadixSort::SortKeys(nullptr, bytes, gkeysAll.p, gkeysSorted.p, nAll, 0, 64, stream));
            cubTmp.alloc(bytes);
            CK(cub::DeviceRadixSort::SortKeys(cubTmp.p, bytes, gkeysAll.p, gkeysSorted.p, nAll, 0, 64, stream));
        }
        k_key_heads<<<divUp(nAll + 1, TB), TB, 0, stream>>>(gkeysSorted, nAll, gHead);
        exclusiveScan(*this, gHead.p, gIdx.p, nAll + 1);
        nblkG = readInt(*this, gIdx.p + nAll);
        gkeys.alloc(nblkG); gRowOf.alloc(nblkG); gColInd.alloc(nblkG); gRowPtr.alloc(nP + 1); hscSlot.alloc(nblk);
        k_compact_keys<<<divUp(nAll, TB), TB, 0, stream>>>(gkeysSorted, gHead, gIdx, nAll, (unsigned long long)nP, gkeys,
                                                           gRowOf, gColInd);
        k_lower_bounds<<<divUp(nP + 1, TB), TB, 0, stream>>>(gRowOf, nblkG, nP, gRowPtr);
        k_map_slots<<<divUp(nblk, TB), TB, 0, stream>>>(lkeys, nblk, gkeys, nblkG, hscSlot);
        gRowPtrP = gRowPtr; gColIndP = gColInd;
        redA.alloc((size_t)6 * nP + nranks); redC.alloc(4);
    }

    lap("pose-major Hpl view, row units");
    // ---- symbolic factorisation
    const int P = nP;
    int nsn = 0;
    if (symParallel) {
        symbolicParallel(nranks > 1 ? gRowOf.p : bRowOf.p, gRowPtrP, gColIndP, P);
        nsn = counts.nsn;
        lap("symbolic analysis + maps (parallel)");
    } else {
    // the sequential statement of the algorithm (symbolic.h) run by one device thread
    y_adjPtr.alloc(P + 1); y_adj.alloc(2 * (size_t)nblkG + 1); y_verts.alloc(P); y_iperm.alloc(P); y_mark.alloc(P);
    y_level.alloc(P); y_queue.alloc(P); y_tmp.alloc(P + 1); y_stack.alloc(4 * (size_t)P + 8); y_snStart.alloc(P + 1);
    y_snParent.alloc(P); y_colToSn.alloc(P); y_snRowPtr.alloc(P + 1); y_snLevel.alloc(P); y_childHead.alloc(P);
    y_childNext.alloc(P);
    const long long rowsCap = sym_rows_cap(P);
    y_snRows.alloc((size_t)rowsCap); y_counts.alloc(1);
    SymWork w{};
    w.P = P; w.bRowPtr = gRowPtrP; w.bColInd = gColIndP; w.leaf = CHOL_MAXW; w.maxw = CHOL_MAXW;
    w.adjPtr = y_adjPtr; w.adj = y_adj; w.verts = y_verts; w.iperm = y_iperm; w.mark = y_mark; w.level = y_level;
    w.queue = y_queue; w.tmp = y_tmp; w.stack = y_stack; w.snStart = y_snStart; w.snParent = y_snParent;
    w.colToSn = y_colToSn; w.snRowPtr = y_snRowPtr; w.snLevel = y_snLevel; w.childHead = y_childHead;
    w.childNext = y_childNext; w.snRows = y_snRows; w.rowsCap = rowsCap;
    {
        const long long adjCap = 2LL * (nblkG - P) + 2;
        const long long symBytes = (adjCap + (P + 1) + 3LL * P) * (long long)sizeof(int);
        const bool inSmem = symBytes <= 200 * 1024;
        if (inSmem) CK(cudaFuncSetAttribute(k_sym_analyze, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)symBytes));
        k_sym_analyze<<<1, 1, inSmem ? (size_t)symBytes : 0, stream>>>(w, y_counts, inSmem ? (int)adjCap : 0);
    }
    CK(cudaMemcpyAsync(&counts, y_counts.p, sizeof(SymCounts), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    lap("symbolic analysis");
    if (counts.err) throw CudaError("symbolic analysis overflowed its row-structure workspace");
    nsn = counts.nsn;
    z_levelPtr.alloc(counts.nlev + 1); z_levelSn.alloc(nsn); z_Lptr.alloc(nsn + 1); z_Uptr.alloc(nsn + 1);
    z_childPtr.alloc(nsn + 1); z_childIdx.alloc(nsn + 1); z_invPtr.alloc(nsn); z_inv.alloc((size_t)counts.ninv + 1);
    z_snBlkPtr.alloc(nsn + 1); z_asmBlk.alloc(nblkG); z_asmRow.alloc(nblkG); z_asmCol.alloc(nblkG);
    z_trsmPtr.alloc(nsn + 1); z_updPtr.alloc(nsn + 1);
    SymMaps m{};
    m.levelPtr = z_levelPtr; m.levelSn = z_levelSn; m.Lptr = z_Lptr; m.Uptr = z_Uptr; m.childPtr = z_childPtr;
    m.childIdx = z_childIdx; m.invPtr = z_invPtr; m.inv = z_inv; m.snBlkPtr = z_snBlkPtr; m.asmBlk = z_asmBlk;
    m.asmRow = z_asmRow; m.asmCol = z_asmCol;
    z_levMaxNt.alloc(counts.nlev + 1);
    k_sym_fill<<<1, 1, 0, stream>>>(w, y_counts, m, z_trsmPtr, z_updPtr, 0, 0, panelRB > 0 ? panelRB : CHOL_RB, CHOL_TB,
                                    z_levMaxNt, z_factStats.alloc(4));
    lap("symbolic maps + task tables");
    }
    CK(cudaMemcpyAsync(factStats, z_factStats.p, sizeof(factStats), cudaMemcpyDeviceToHost, stream));
    h_levelPtr.resize(counts.nlev + 1); h_trsmPtr.resize(nsn + 1); h_updPtr.resize(nsn + 1); h_levMaxNt.resize(counts.nlev + 1);
    CK(cudaMemcpyAsync(h_levMaxNt.data(), z_levMaxNt.p, sizeof(int) * counts.nlev, cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(h_levelPtr.data(), z_levelPtr.p, sizeof(int) * (counts.nlev + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(h_trsmPtr.data(), z_trsmPtr.p, sizeof(int) * (nsn + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(h_updPtr.data(), z_updPtr.p, sizeof(int) * (nsn + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(&h_nchild, z_childPtr.p + nsn, sizeof(int), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    c.levelPtr = z_levelPtr; c.levelSn = z_levelSn; c.trsmPtr = z_trsmPtr; c.updPtr = z_updPtr;
    distLc = 0; distBcast.clear(); h_own.clear();
    if (nranks > 1 && distMode > 0) planDistFactor();
    planPanelSlabs();
    if (getenv("CUGO_B200_DEBUG")) {
        fprintf(stderr, "[cugo_b200] elimination tree: %d supernodes, %d levels, max front %d block rows; per level "
                        "(supernodes / panel CTAs / update tiles / widest):", nsn, counts.nlev, counts.maxFront);
        for (int lev = 0; lev < counts.nlev; lev++) {
            const int q0 = h_levelPtr[lev], q1 = h_levelPtr[lev + 1];
            fprintf(stderr, " %d/%d/%d/%d", q1 - q0, h_trsmPtr[q1] - h_trsmPtr[q0], h_updPtr[q1] - h_updPtr[q0], h_levMaxNt[lev]);
        }
        fprintf(stderr, "\n");
    }
    if (counts.maxCols > CHOL_MAXW) throw CudaError("internal: supernode wider than CHOL_MAXW");
    maxRt = counts.maxFront;   // upper bound of r_t
    cL.alloc((size_t)counts.lsize); cU.alloc((size_t)counts.usize); cX.alloc((size_t)P * 6); cW.alloc((size_t)P * 21);
    cZ.alloc((size_t)P * 6 * 6 * CHOL_MAXW);
    cPart.alloc((size_t)nsn * BWD_MAXG * 6 * CHOL_MAXW); cCnt.alloc(nsn);
    CK(cudaMemsetAsync(cCnt.p, 0, sizeof(int) * nsn, stream));
    cFlag.alloc(nsn + 1);
    CK(cudaMemsetAsync(cFlag.p, 0, sizeof(int) * (nsn + 1), stream));

    // ---- system buffers
    Hpp.alloc((size_t)nP * 36); Hll.alloc((size_t)nL * 9);
    if (f32) HplF.alloc((size_t)nHpl * 18); else Hpl.alloc((size_t)nHpl * 18);
    b.alloc((size_t)nP * 6 + (size_t)nL * 3); x.alloc((size_t)nP * 6 + (size_t)nL * 3);
    invHll.alloc((size_t)nL * 9);
    Hsc.alloc((size_t)nblkG * 36 + (size_t)nP * 6);     // [Hsc | bsc] contiguous: one all-reduce with several ranks
    const int errGrid = std::min(divUp(E, 256), 148 * 8);
    chiPartial.alloc((size_t)errGrid + (nIcp > 0 ? ic.grid : 0)); scalePartial.alloc(nChunks); maxDiagL.alloc(nChunks); maxDiagP.alloc(nP);
    scalars.alloc(8);
    CK(cudaMemsetAsync(x.p, 0, ((size_t)nP * 6 + (size_t)nL * 3) * sizeof(double), stream));

    d.nP_all = nP_all; d.nP = nP; d.nL_all = nL_all; d.nL = nL; d.E = (int)E; d.nHpl = nHpl; d.nblk = nblk;
    d.nmul = nmulAll; d.nChunks = nChunks;
    d.s_iP = s_iP; d.s_iL = s_iL; d.s_uv = s_uv; d.s_r = s_r; d.s_w = s_w; d.s_set = s_set;
    d.s_cam = perEdgeCam ? s_cam.p : nullptr; d.s_out = anyOut ? s_out.p : nullptr; d.s_gid = s_gid;
    d.s_hplPre = s_hplPre; d.lmEdgePtr = lmEdgePtr; d.chunkLm = chunkLm;
    d.posePtr = posePtr; d.p_iL = p_iL; d.p_uv = p_uv; d.p_r = p_r; d.p_w = p_w; d.p_set = p_set;
    d.p_cam = perEdgeCam ? p_cam.p : nullptr; d.p_out = anyOut ? p_out.p : nullptr;
    d.cams = perEdgeCam ? cams.p : nullptr;
    d.nPoseUnits = nPoseUnits; d.poseSlicePtr = poseSlicePtr; d.poseUnits = poseUnits; d.posePart = posePart; d.poseTicket = poseTicket;
    d.colPtr = colPtr; d.rowInd = rowInd; d.hplLm = hplLm;
    d.bRowPtr = bRowPtr; d.bColInd = bColInd; d.bRowOf = bRowOf; d.prodPtr = prodPtr; d.prodIJ = prodIJ;
    d.poseSlotPtr = poseSlotPtr; d.poseSlots = poseSlots; d.prodRank = prodRank; d.poseLm = poseLm; d.rowUnits = rowUnits;
    d.Hpp = Hpp; d.Hll = Hll; d.Hpl = f32 ? nullptr : Hpl.p; d.HplF = f32 ? HplF.p : nullptr; d.b = b; d.x = x; d.invHll = invHll; d.Hsc = Hsc;
    d.bsc = Hsc.p + (size_t)nblkG * 36;
    d.HscPart = d.Hsc; d.bscPart = d.bsc;
    if (nranks > 1) {       // partial system of this rank: zeroed once, the blocks it does not own are never written
        const size_t nred = (size_t)nblkG * 36 + (size_t)nP * 6;
        HscPartBuf.alloc(nred);
        CK(cudaMemsetAsync(HscPartBuf.p, 0, nred * sizeof(double), stream));
        d.HscPart = HscPartBuf.p; d.bscPart = HscPartBuf.p + (size_t)nblkG * 36;
        peerSetup(nred);
        if (peerOn) {           // partial and summed system live in the exchange block
            d.HscPart = reinterpret_cast<double*>(peerLocal + peer.offPart); d.bscPart = d.HscPart + (size_t)nblkG * 36;
            d.Hsc = reinterpret_cast<double*>(peerLocal + peer.offDst); d.bsc = d.Hsc + (size_t)nblkG * 36;
        }
    }
    d.nranks = nranks; d.rank = rank; d.nblkG = nblkG; d.hscSlot = nranks > 1 ? hscSlot.p : nullptr;
    d.redA = nranks > 1 ? redA.p : nullptr; d.redC = nranks > 1 ? redC.p : nullptr;
    d.chiPartial = chiPartial; d.scalePartial = scalePartial; d.maxDiagL = maxDiagL; d.maxDiagP = maxDiagP;
    d.chiEdge = nullptr; d.errGrid = errGrid; d.nChi = errGrid + (nIcp > 0 ? ic.grid : 0);
    d.pose[0] = pose0; d.pose[1] = pose1; d.lm[0] = lm0; d.lm[1] = lm1; d.st = state;

    c.P = P; c.nsn = nsn; c.nlev = counts.nlev;
    c.perm = y_verts; c.snStart = y_snStart; c.snRowPtr = y_snRowPtr; c.snRows = y_snRows;
    c.Lptr = z_Lptr; c.Uptr = z_Uptr; c.childPtr = z_childPtr;     // (levelPtr / levelSn / trsmPtr / updPtr: set with the schedule above)
    c.childIdx = z_childIdx; c.invPtr = z_invPtr; c.inv = z_inv; c.snBlkPtr = z_snBlkPtr; c.asmBlk = z_asmBlk;
    c.asmRow = z_asmRow; c.asmCol = z_asmCol;
    c.L = cL; c.U = cU; c.xperm = cX; c.Winv = cW; c.Zinv = cZ; c.bwdPart = cPart; c.bwdCnt = cCnt;
    c.xFlag = cFlag; c.colToSn = y_colToSn;
    { const char* e = getenv("CUGO_B200_DEBUG"); c.dbg = e ? atoi(e) : 0; }   // diagnostic builds (make DIAG=1): 1 phase stamps, 3 also a worker warp's view

    // dynamic shared memory opt-in for the factorisation kernels
    const int ncMax = 6 * CHOL_MAXW;
    const size_t bwd2Bytes = ((size_t)ncMax * (ncMax + 1) + ncMax + 21 * CHOL_MAXW + 6 * (size_t)maxRt) * 8;
    if (bwd2Bytes > 227 * 1024) throw UnsupportedError("front too large for the shared-memory backward solve of this version");
    CK(cudaFuncSetAttribute(k_chol_bwd3<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd2Bytes));
    CK(cudaFuncSetAttribute(k_chol_zinv<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (int)((2 * (size_t)ncMax * (ncMax + 1) + 21 * CHOL_MAXW + 36 * 8) * 8)));
    if (!side) {
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));     // lo = lowest priority
        CK(cudaStreamCreateWithPriority(&side, cudaStreamNonBlocking, lo));
        CK(cudaStreamCreateWithPriority(&side2, cudaStreamNonBlocking, lo));
        CK(cudaEventCreateWithFlags(&evJoin2, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&evFork, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&evJoin, cudaEventDisableTiming));
    }
    CK(cudaFuncSetAttribute(k_chol_panel3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, PANEL_LD * ncMax * 8));
    CK(cudaFuncSetAttribute(k_chol_panel3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, PANEL_LD * ncMax * 8));
    CK(cudaFuncSetAttribute(k_chol_update3, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (2 * ncMax * UPD_LD + 2 * 6 * CHOL_TB * UPD_CLD) * 8));
    CK(cudaFuncSetAttribute(k_chol_update_small, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (2 * ncMax * UPD_LD + (UPD3_TPB / 32) * 6 * UPD_TBS * UPD_SCLD) * 8));
    CK(cudaStreamSynchronize(stream));
    systemBuilt = false; solved = false;
    lap("buffers");
    structDirty = true;
    captureSlot();
    structDirty = false;            // only once the trial graph exists
    lap("graph capture");
}

// ---------------------------------------------------------------------------
// ICP edge sets and graphs without free landmarks (kernels_poseonly.cuh)
// ---------------------------------------------------------------------------
int Solver::addIcpEdgeSet(int kind, int n, const int* iP, const double* meas, const double* omega, int nomega, int rk,
                          double delta) {
    if (!haveVertices) throw StateError("add_icp_edge_set before set_vertices");
    if ((kind != EDGE_PLANE && kind != EDGE_LINE) || n < 0) throw ArgError("add_icp_edge_set: kind must be plane or line, n >= 0");
    if (n > 0 && (!iP || !meas || !omega)) throw ArgError("add_icp_edge_set: null pointer");
    if (nomega != 1 && nomega != n) throw ArgError("add_icp_edge_set: nomega must be 1 or n");
    if (rk < 0 || rk > 3) throw ArgError("add_icp_edge_set: unknown robust kernel");
    if ((int)icpSets.size() >= MAX_ICP_SETS) throw ArgError("add_icp_edge_set: too many ICP edge sets");
    IcpHostSet s;
    s.kind = kind; s.n = n; s.nomega = n > 0 ? nomega : 1; s.rk = rk; s.delta = delta;
    for (int i = 0; i < n; i++) {
        if (iP[i] < 0 || iP[i] >= nP_all) throw ArgError("add_icp_edge_set: pose index out of range");
        if (iP[i] >= nP) throw ArgError("add_icp_edge_set: edge on a fixed pose is not active");
        if (kind == EDGE_LINE && !(meas[(size_t)i * 10 + 6] > 0)) throw ArgError("add_icp_edge_set: line length must be > 0");
    }
    if (n > 0) {
        use();
        s.iP.reset(new DBuf<int>); s.meas.reset(new DBuf<double>); s.omega.reset(new DBuf<double>);
        s.iP->alloc(n); s.meas->alloc((size_t)n * s.cols()); s.omega->alloc(nomega);
        CK(cudaMemcpyAsync(s.iP->p, iP, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(s.meas->p, meas, (size_t)n * s.cols() * sizeof(double), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(s.omega->p, omega, (size_t)nomega * sizeof(double), cudaMemcpyHostToDevice, stream));
        CK(cudaStreamSynchronize(stream));      // the caller's arrays are free again on return
    }
    icpSets.push_back(std::move(s));
    structDirty = true;
    return ICP_SET_BASE + (int)icpSets.size() - 1;
}

void Solver::refreshIcpParams() {
    for (size_t i = 0; i < icpSets.size(); i++) {
        ic.kind[i] = icpSets[i].kind; ic.rkType[i] = icpSets[i].rk; ic.rkDelta[i] = icpSets[i].delta;
    }
}

// the sets' arrays (on the device since add_*_edge_set) are split into the column-major payload and sorted by pose
void Solver::buildIcp() {
    const size_t nI = (size_t)nIcp;
    const int TB = 256, gI = divUp(nIcp, TB);
    i_rIP.alloc(nI); i_rset.alloc(nI); i_rw.alloc(nI); i_rpay.alloc(nI * ICP_COLS);
    size_t off = 0;
    for (size_t si = 0; si < icpSets.size(); si++) {
        const IcpHostSet& s = icpSets[si];
        if (s.n == 0) continue;
        CK(cudaMemcpyAsync(i_rIP.p + off, s.iP->p, (size_t)s.n * sizeof(int), cudaMemcpyDeviceToDevice, stream));
        k_icp_scatter<<<divUp(s.n, TB), TB, 0, stream>>>(s.meas->p, s.omega->p, s.n, s.cols(), s.nomega, nI, off, i_rpay, i_rw);
        CK(cudaMemsetAsync(i_rset.p + off, (int)si, s.n, stream));
        off += s.n;
    }
    i_src.alloc(nI); i_iP.alloc(nI); i_keyOut.alloc(nI);
    k_fill_index<<<gI, TB, 0, stream>>>(i_iP.p, (int)nI, 0, 1);          // (i_iP: the identity, values of the sort)
    sortPairs(*this, i_rIP.p, i_keyOut.p, i_iP.p, i_src.p, (int)nI, bitsFor((unsigned long long)nP_all));
    i_set.alloc(nI); i_w.alloc(nI); i_pay.alloc(nI * ICP_COLS);
    k_icp_gather<<<gI, TB, 0, stream>>>(i_src, nI, i_rset, i_rw, i_rpay, i_set, i_w, i_pay);
    i_posePtr.alloc(nP_all + 1);
    k_lower_bounds<<<divUp(nP_all + 1, TB), TB, 0, stream>>>(i_keyOut, (int)nI, nP_all, i_posePtr);
    i_sliceCnt.alloc(nP + 1); i_slicePtr.alloc(nP + 1);
    k_icp_slice_count<<<divUp(nP + 1, TB), TB, 0, stream>>>(i_posePtr, nP, i_sliceCnt);
    exclusiveScan(*this, i_sliceCnt.p, i_slicePtr.p, nP + 1);
    const int nUnits = readInt(*this, i_slicePtr.p + nP);
    i_units.alloc(std::max(nUnits, 1)); i_part.alloc((size_t)std::max(nUnits, 1) * 27);
    k_pose_units_fill<<<divUp(std::max(nP, 1), TB), TB, 0, stream>>>(i_slicePtr, nP, i_units);
    ic.nI = (int)nI; ic.nUnits = nUnits; ic.grid = std::max(1, std::min(divUp(nIcp, 256), 148 * 8));
    ic.iP = i_keyOut; ic.set = i_set; ic.w = i_w; ic.pay = i_pay; ic.posePtr = i_posePtr; ic.slicePtr = i_slicePtr;
    ic.units = i_units; ic.part = i_part;
    refreshIcpParams();
    lap("ICP edges: pose order, slices");
}

// graph without free landmarks: no Hpl, no Schur complement, no symbolic analysis -- Hpp is block diagonal
void Solver::finishPoseOnly() {
    Hpp.alloc((size_t)nP * 36); b.alloc((size_t)nP * 6); x.alloc((size_t)nP * 6); maxDiagP.alloc(nP); scalars.alloc(8);
    const int errGrid = std::max(1, std::min(divUp(E, 256), 148 * 8));
    chiPartial.alloc((size_t)errGrid + (nIcp > 0 ? ic.grid : 0));
    CK(cudaMemsetAsync(x.p, 0, (size_t)nP * 6 * sizeof(double), stream));
    Dev nd{};
    nd.nP_all = nP_all; nd.nP = nP; nd.nL_all = nL_all; nd.nL = 0; nd.E = (int)E;
    if (E > 0) {
        nd.s_iP = s_iP; nd.s_iL = s_iL; nd.s_uv = s_uv; nd.s_r = s_r; nd.s_w = s_w; nd.s_set = s_set;
        nd.s_cam = perEdgeCam_ ? s_cam.p : nullptr; nd.s_out = anyOut_ ? s_out.p : nullptr; nd.s_gid = s_gid;
        nd.s_hplPre = s_hplPre; nd.lmEdgePtr = lmEdgePtr; nd.chunkLm = chunkLm;
        nd.posePtr = posePtr; nd.p_iL = p_iL; nd.p_uv = p_uv; nd.p_r = p_r; nd.p_w = p_w; nd.p_set = p_set;
        nd.p_cam = perEdgeCam_ ? p_cam.p : nullptr; nd.p_out = anyOut_ ? p_out.p : nullptr;
        nd.cams = perEdgeCam_ ? cams.p : nullptr;
        nd.nPoseUnits = nPoseUnits; nd.poseSlicePtr = poseSlicePtr; nd.poseUnits = poseUnits; nd.posePart = posePart;
        nd.poseTicket = poseTicket;
    }
    nd.Hpp = Hpp; nd.b = b; nd.x = x; nd.chiPartial = chiPartial; nd.maxDiagP = maxDiagP;
    nd.errGrid = errGrid; nd.nChi = errGrid + (nIcp > 0 ? ic.grid : 0);
    nd.pose[0] = pose0; nd.pose[1] = pose1; nd.lm[0] = lm0; nd.lm[1] = lm1; nd.st = state;
    nd.nranks = 1; nd.rank = 0;
    d = nd;
    c = Chol{};
    counts = SymCounts{};
    memset(factStats, 0, sizeof(factStats));
    nmulRef = 0; nnzSymm = 0; nblkG = 0;
    if (!h_state) CK(cudaMallocHost(&h_state, sizeof(LMState)));
    CK(cudaStreamSynchronize(stream));
    systemBuilt = false; solved = false;
    structDirty = true;
    captureSlot();
    structDirty = false;
    lap("pose-only: buffers, graph capture");
}

void Solver::enqueueErrors(int which, int guarded, bool viaLaunch) {
    if (viaLaunch) {
        launch(k_errors<256>, d.errGrid, 256, 0, d, sp, which, guarded); mark("errors");
        if (nIcp > 0) { launch(k_icp_errors<256>, ic.grid, 256, 0, d, ic, which, guarded); mark("icp_errors"); }
    } else {
        k_errors<256><<<d.errGrid, 256, 0, stream>>>(d, sp, which, guarded);
        if (nIcp > 0) k_icp_errors<256><<<ic.grid, 256, 0, stream>>>(d, ic, which, guarded);
    }
}

// ---------------------------------------------------------------------------
// Factorisation distributed by assembly-tree subtree (several ranks; no counterpart in the reference, which is
// single-GPU: src/cuda_device.cpp:264-282).  Every rank holds the same symbolic analysis.  A cut level Lc splits the
// assembly tree: supernodes on levels >= Lc are factorised by EVERY rank (the replicated top: its fronts are what all
// subtrees feed, and solving it everywhere saves a broadcast of the top of x), the subtrees hanging below the cut are
// dealt to the ranks by estimated factor flops (longest-processing-time first).  A rank runs the panel / update /
// backward kernels of its own and of the replicated supernodes only (filtered level schedule).  Two exchanges per
// trial carry the coupling: the update matrices of the subtree roots are broadcast by their owners before level Lc,
// and after the backward solves the owners' pieces of x are summed (every other rank contributes zeros: exact).
// Lc is chosen to minimise replicated + max-per-rank flops; 0 (= today's fully replicated factorisation) when that does
// not pay, e.g. at the KITTI-00 size where the tree's critical path, not the flops, is the bound.
// ---------------------------------------------------------------------------
__global__ void k_col_owner(const int* colToSn, const int* own, int P, int* colOwn) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < P) colOwn[j] = own[colToSn[j]];
}
// zero the pieces of the permuted solution this rank does not answer for (replicated columns: rank 0 answers)
__global__ void k_x_mask(const int* colOwn, int P, int rank, double* xperm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 6 * P) return;
    const int o = colOwn[i / 6];
    if (!(o == rank || (o < 0 && rank == 0))) xperm[i] = 0.0;
}
__global__ void k_x_scatter(const double* xperm, const int* perm, int P, double* x) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 6 * P) x[(size_t)perm[i / 6] * 6 + i % 6] = xperm[i];
}

void Solver::planDistFactor() {
    const int nsn = counts.nsn, nlev = counts.nlev;
    std::vector<int> par(nsn), lev(nsn), st(nsn + 1), rp(nsn + 1), lsn(nsn);
    std::vector<long long> uptr(nsn + 1);
    CK(cudaMemcpyAsync(par.data(), y_snParent.p, sizeof(int) * nsn, cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(lev.data(), y_snLevel.p, sizeof(int) * nsn, cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(st.data(), y_snStart.p, sizeof(int) * (nsn + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(rp.data(), y_snRowPtr.p, sizeof(int) * (nsn + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(lsn.data(), z_levelSn.p, sizeof(int) * nsn, cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(uptr.data(), z_Uptr.p, sizeof(long long) * (nsn + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    std::vector<double> fl(nsn);
    double total = 0.0;
    for (int t = 0; t < nsn; t++) {
        const double ncd = 6.0 * (st[t + 1] - st[t]), rd = 6.0 * (rp[t + 1] - rp[t]) + 1.0;
        fl[t] = ncd * ncd * ncd / 3.0 + rd * ncd * ncd + rd * rd * ncd;
        total += fl[t];
    }
    // component (subtree below the cut) of every supernode for a cut level L: follow parents while the parent's level < L
    auto evaluate = [&](int L, std::vector<int>* ownOut) {
        std::vector<int> root(nsn, -1);
        std::vector<double> w(nsn, 0.0);
        double repl = 0.0;
        for (int t = nsn - 1; t >= 0; t--) {          // parents have larger ids than their children
            if (lev[t] >= L) { repl += fl[t]; continue; }
            const int p = par[t];
            root[t] = (p >= 0 && lev[p] < L) ? root[p] : t;
        }
        for (int t = 0; t < nsn; t++) if (root[t] >= 0) w[root[t]] += fl[t];
        std::vector<int> roots;
        for (int t = 0; t < nsn; t++) if (root[t] == t) roots.push_back(t);
        std::sort(roots.begin(), roots.end(), [&](int a, int b) { return w[a] != w[b] ? w[a] > w[b] : a < b; });
        std::vector<double> load(nranks, 0.0);
        std::vector<int> rootOwner(nsn, -1);
        for (int r : roots) {
            int best = 0;
            for (int k = 1; k < nranks; k++) if (load[k] < load[best]) best = k;
            load[best] += w[r];
            rootOwner[r] = best;
        }
        if (ownOut) {
            ownOut->assign(nsn, -1);
            for (int t = 0; t < nsn; t++) if (root[t] >= 0) (*ownOut)[t] = rootOwner[root[t]];
        }
        return repl + *std::max_element(load.begin(), load.end());
    };
    int bestL = 0;
    double bestT = total;
    for (int L = 1; L < nlev; L++) {
        const double T = evaluate(L, nullptr);
        if (T < bestT) { bestT = T; bestL = L; }
    }
    // below ~5 GFLOP per factorisation the tree's critical path is the bound, not the flops: keep it replicated
    const bool pays = bestL > 0 && bestT < 0.7 * total && (distMode >= 2 || total > 5e9);
    if (getenv("CUGO_B200_DEBUG"))
        fprintf(stderr, "[cugo_b200] distributed factorisation: %.2f GFLOP per factorisation, best cut level %d of %d -> %.2f GFLOP "
                        "per rank (%s)\n", total * 1e-9, bestL, nlev, bestT * 1e-9, pays ? "on" : "off");
    if (!pays) return;
    distLc = bestL;
    evaluate(bestL, &h_own);
    // filtered level schedule of this rank: its own and the replicated supernodes, in the order of the full schedule
    std::vector<int> fPtr(nlev + 1, 0), fSn, fTrsm, fUpd;
    int a = 0, b = 0;
    for (int l = 0; l < nlev; l++) {
        for (int q = h_levelPtr[l]; q < h_levelPtr[l + 1]; q++) {
            const int t = lsn[q];
            if (h_own[t] >= 0 && h_own[t] != rank) continue;
            fSn.push_back(t); fTrsm.push_back(a); fUpd.push_back(b);
            const int rt = rp[t + 1] - rp[t];
            const int rbU = panelRB > 0 ? panelRB : CHOL_RB;
            const int ns = (rt + rbU - 1) / rbU;
            a += ns > 0 ? ns : 1;
            if (rt > 0) { const int n2 = (rt + CHOL_TB) / CHOL_TB; b += n2 * (n2 + 1) / 2; }
        }
        fPtr[l + 1] = (int)fSn.size();
    }
    fTrsm.push_back(a); fUpd.push_back(b);
    const int nf = (int)fSn.size();
    zr_levelPtr.alloc(nlev + 1); zr_levelSn.alloc(std::max(nf, 1)); zr_trsmPtr.alloc(nf + 1); zr_updPtr.alloc(nf + 1);
    zr_colOwn.alloc(nP); y_ak1.alloc(std::max(nsn, nblkG));
    CK(cudaMemcpyAsync(zr_levelPtr.p, fPtr.data(), sizeof(int) * (nlev + 1), cudaMemcpyHostToDevice, stream));
    if (nf) CK(cudaMemcpyAsync(zr_levelSn.p, fSn.data(), sizeof(int) * nf, cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(zr_trsmPtr.p, fTrsm.data(), sizeof(int) * (nf + 1), cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(zr_updPtr.p, fUpd.data(), sizeof(int) * (nf + 1), cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(y_ak1.p, h_own.data(), sizeof(int) * nsn, cudaMemcpyHostToDevice, stream));
    k_col_owner<<<divUp(nP, 256), 256, 0, stream>>>(y_colToSn, y_ak1, nP, zr_colOwn);
    CK(cudaStreamSynchronize(stream));
    h_levelPtr = fPtr; h_trsmPtr = fTrsm; h_updPtr = fUpd;
    c.levelPtr = zr_levelPtr; c.levelSn = zr_levelSn; c.trsmPtr = zr_trsmPtr; c.updPtr = zr_updPtr;
    // what crosses the cut upwards: the update matrix of every subtree root that has a (replicated) parent
    for (int t = 0; t < nsn; t++) {
        if (h_own[t] < 0 || par[t] < 0 || lev[par[t]] < distLc) continue;
        const long long u = 6LL * (rp[t + 1] - rp[t]) + 1;
        distBcast.push_back({(size_t)uptr[t], (size_t)(u * u), h_own[t]});
    }
}

// Exchange block of this rank + the peers' blocks through CUDA IPC (collective: every rank calls it from build_structure).
void Solver::peerSetup(size_t nred) {
    peerRelease();
    if (nranks < 2 || nranks > PEER_MAXR || !peerMode) return;
    PeerCtx x{};
    x.nranks = nranks; x.rank = rank;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    x.offPart = 0;
    x.offDst = up(nred * sizeof(double));
    x.offSmall = x.offDst + up(nred * sizeof(double));
    x.offFlag = x.offSmall + up((size_t)PEER_MAXR * PEER_SMALL_MAX * sizeof(double));
    x.offEpoch = x.offFlag + up((size_t)3 * PEER_MAXR * PEER_CTAS * sizeof(int));
    const size_t bytes = x.offEpoch + up((size_t)2 * PEER_CTAS * sizeof(int));
    int ok = 1;
    cudaIpcMemHandle_t mine{};
    if (cudaMalloc(&peerLocal, bytes) != cudaSuccess) { peerLocal = nullptr; ok = 0; cudaGetLastError(); }
    if (ok) {
        CK(cudaMemsetAsync(peerLocal, 0, bytes, stream));
        if (cudaIpcGetMemHandle(&mine, peerLocal) != cudaSuccess) { ok = 0; cudaGetLastError(); }
    }
    // handles of everybody (64 bytes each) + the flag that every rank got this far
    DBuf<unsigned char> hbuf;
    const size_t hsz = sizeof(cudaIpcMemHandle_t) + 8;
    hbuf.alloc(hsz * (nranks + 1));
    std::vector<unsigned char> host(hsz * (nranks + 1), 0);
    memcpy(host.data() + hsz * nranks, &mine, sizeof(mine));
    host[hsz * nranks + sizeof(mine)] = (unsigned char)ok;
    CK(cudaMemcpyAsync(hbuf.p + hsz * nranks, host.data() + hsz * nranks, hsz, cudaMemcpyHostToDevice, stream));
    NK(g_nccl.AllGather(hbuf.p + hsz * nranks, hbuf.p, hsz, ncclUint8, comm, stream));
    CK(cudaMemcpyAsync(host.data(), hbuf.p, hsz * nranks, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    for (int r = 0; r < nranks; r++) ok &= host[hsz * r + sizeof(mine)];
    if (ok) {
        for (int r = 0; r < nranks && ok; r++) {
            if (r == rank) { x.base[r] = peerLocal; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, host.data() + hsz * r, sizeof(h));
            void* pp = nullptr;
            if (cudaIpcOpenMemHandle(&pp, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cudaGetLastError(); break; }
            x.base[r] = static_cast<char*>(pp);
        }
    }
    // every rank must take the same path: agree once more
    redC.alloc(4);
    const double mineOk = ok ? 0.0 : 1.0;
    CK(cudaMemcpyAsync(redC.p, &mineOk, sizeof(double), cudaMemcpyHostToDevice, stream));
    allReduceSum(redC.p, 1);
    double bad = 0.0;
    CK(cudaMemcpyAsync(&bad, redC.p, sizeof(double), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    peer = x;
    if (bad != 0.0) {
        if (getenv("CUGO_B200_DEBUG")) fprintf(stderr, "[cugo_b200] peer memory not available on every rank: the exchange stays on NCCL\n");
        peerRelease();
        return;
    }
    peerOn = true;
    peerNred = nred;
    if (getenv("CUGO_B200_DEBUG") && rank == 0)
        fprintf(stderr, "[cugo_b200] exchange over peer memory: %d ranks, %.1f MB per rank\n", nranks, bytes / 1048576.0);
}
void Solver::peerRelease() {
    if (peerLocal || peerOn) {
        cudaStreamSynchronize(stream);
        for (int r = 0; r < peer.nranks; r++)
            if (r != peer.rank && peer.base[r]) cudaIpcCloseMemHandle(peer.base[r]);
        if (peerLocal) cudaFree(peerLocal);
    }
    peerLocal = nullptr; peerOn = false; peer = PeerCtx{};
}

// Slab height of the panel kernel per tree level.  A panel CTA factors the diagonal part of its supernode and one slab
// of the rows below; the rows below are independent of each other, so on levels with few supernodes (the top of the
// tree, where most SMs idle) short slabs spread them over more SMs, while the leaf levels keep tall slabs so that
// the level still fits one wave.  panel_rb > 0 fixes the height for every level.
void Solver::planPanelSlabs() {
    int sms = 148;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int nlev = counts.nlev, cap = std::max(16, (sms * 9) / 10);
    const int nq = h_levelPtr[nlev];
    const int nsn = counts.nsn;
    std::vector<int> lsn(std::max(nq, 1)), rp(nsn + 1), rbLev(nlev + 1, CHOL_RB), tr(nq + 1, 0);
    if (nq) CK(cudaMemcpyAsync(lsn.data(), c.levelSn, sizeof(int) * nq, cudaMemcpyDeviceToHost, stream));
    CK(cudaMemcpyAsync(rp.data(), y_snRowPtr.p, sizeof(int) * (nsn + 1), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    static const int cand[] = {2, 3, 4, 5, 6, 8, 10, 12, 14, 16};
    int a = 0;
    for (int l = 0; l < nlev; l++) {
        int rb = panelRB > 0 ? panelRB : 10;
        if (panelRB == 0) {
            // cost of the level ~ waves of CTAs (one CTA per SM: the panel takes most of its shared memory) x time of a CTA:
            // the chain of diagonal blocks (fixed) against the slab rows its slab warps carry (grows with the slab height)
            double best = 1e300;
            for (int cnd : cand) {
                int ctas = 0;
                for (int q = h_levelPtr[l]; q < h_levelPtr[l + 1]; q++) {
                    const int rt = rp[lsn[q] + 1] - rp[lsn[q]];
                    ctas += std::max(1, (rt + cnd - 1) / cnd);
                }
                const int waves = (ctas + sms - 1) / sms;
                const double cost = waves * std::max(27.0, 12.0 + 2.2 * cnd) + (ctas > cap ? 0.5 : 0.0);
                if (cost < best - 1e-9) { best = cost; rb = cnd; }
            }
        }
        rbLev[l] = rb;
        for (int q = h_levelPtr[l]; q < h_levelPtr[l + 1]; q++) {
            const int rt = rp[lsn[q] + 1] - rp[lsn[q]];
            tr[q] = a;
            a += std::max(1, (rt + rb - 1) / rb);
        }
    }
    tr[nq] = a;
    // update tiles: 24x24 instead of 48x48 on the levels where that still fits the GPU in one go (two CTAs per SM)
    std::vector<int> up(nq + 1, 0);
    h_tbLev.assign(nlev + 1, CHOL_TB);
    int b = 0;
    for (int l = 0; l < nlev; l++) {
        auto tiles = [&](int tb) {
            int t = 0;
            for (int q = h_levelPtr[l]; q < h_levelPtr[l + 1]; q++) {
                const int rt = rp[lsn[q] + 1] - rp[lsn[q]];
                if (rt > 0) { const int n2 = (rt + tb) / tb; t += n2 * (n2 + 1) / 2; }
            }
            return t;
        };
        const int tb = (updSmall && tiles(UPD_TBS) <= (updSmall + 1) * sms) ? UPD_TBS : CHOL_TB;     // option value k: levels with up to (k + 1) x SMs small tiles
        h_tbLev[l] = tb;
        for (int q = h_levelPtr[l]; q < h_levelPtr[l + 1]; q++) {
            const int rt = rp[lsn[q] + 1] - rp[lsn[q]];
            up[q] = b;
            if (rt > 0) { const int n2 = (rt + tb) / tb; b += n2 * (n2 + 1) / 2; }
        }
    }
    up[nq] = b;
    zs_updPtr.alloc(nq + 1);
    CK(cudaMemcpyAsync(zs_updPtr.p, up.data(), sizeof(int) * (nq + 1), cudaMemcpyHostToDevice, stream));
    h_updPtr = up;
    c.updPtr = zs_updPtr;
    zs_trsmPtr.alloc(nq + 1); zs_rbLev.alloc(nlev + 1);
    CK(cudaMemcpyAsync(zs_trsmPtr.p, tr.data(), sizeof(int) * (nq + 1), cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(zs_rbLev.p, rbLev.data(), sizeof(int) * (nlev + 1), cudaMemcpyHostToDevice, stream));
    CK(cudaStreamSynchronize(stream));
    h_trsmPtr = tr;
    c.trsmPtr = zs_trsmPtr; c.rbLev = zs_rbLev;
}

// Symbolic analysis as parallel device work (kernels_sym.cuh); the arrays are those of symbolic.h, bit for bit.
void Solver::symbolicParallel(const int* gRowOfP, const int* gRowPtrP, const int* gColIndP, int P) {
    (void)gRowPtrP;
    const int TB = 256;
    const int nadj = 2 * (nblkG - P);
    y_adjPtr.alloc(P + 1); y_adj.alloc((size_t)std::max(nadj, 1)); y_vOf.alloc((size_t)std::max(nadj, 1));
    y_verts.alloc(P); y_iperm.alloc(P); y_mark.alloc(P); y_level.alloc(P); y_queue.alloc(P); y_tmp.alloc(P + 1);
    y_isStart.alloc(P); y_snStart.alloc(P + 1); y_colToSn.alloc(P); y_snParent.alloc(P); y_snLevel.alloc(P);
    y_childHead.alloc(P); y_childNext.alloc(P); y_rcount.alloc(P); y_snRowPtr.alloc(P + 1); y_counts.alloc(1);
    y_regA.alloc(P); y_regB.alloc(P); y_emits.alloc(P); y_symCnt.alloc(4);
    // adjacency: sorted neighbour lists of the symmetric block graph
    if (nadj > 0) {
        keyA.alloc((size_t)nadj); keyB.alloc((size_t)nadj);
        k_sym_pairs<<<divUp(nblkG, TB), TB, 0, stream>>>(gRowOfP, gColIndP, nblkG, (unsigned long long)P, keyA);
        size_t bytes = 0;
        const int bits = bitsFor((unsigned long long)P * (unsigned long long)P);
        CK(cub::DeviceRadixSort::SortKeys(nullptr, bytes, keyA.p, keyB.p, nadj, 0, bits, stream));
        cubTmp.alloc(bytes);
        CK(cub::DeviceRadixSort::SortKeys(cubTmp.p, bytes, keyA.p, keyB.p, nadj, 0, bits, stream));
        k_sym_adj<<<divUp(nadj, TB), TB, 0, stream>>>(keyB, nadj, (unsigned long long)P, y_adj, y_vOf);
    }
    k_lower_bounds<<<divUp(P + 1, TB), TB, 0, stream>>>(y_vOf, nadj, P, y_adjPtr);
    // nested dissection, supernode column ranges
    SymPar sp2{};
    sp2.P = P; sp2.leaf = CHOL_MAXW; sp2.maxw = CHOL_MAXW; sp2.nadj = nadj;
    sp2.adjPtr = y_adjPtr; sp2.adj = y_adj; sp2.verts = y_verts; sp2.iperm = y_iperm; sp2.isStart = y_isStart;
    sp2.snStart = y_snStart; sp2.colToSn = y_colToSn; sp2.mark = y_mark; sp2.level = y_level; sp2.queue = y_queue; sp2.tmp = y_tmp;
    sp2.regA = y_regA; sp2.regB = y_regB; sp2.emits = y_emits; sp2.counters = y_symCnt;
    const size_t ndBytes = ((size_t)nadj + 6 * (size_t)P + 2) * sizeof(int);
    sp2.useSmem = ndBytes <= 200 * 1024 ? 1 : 0;
    if (sp2.useSmem) CK(cudaFuncSetAttribute(k_sym_nd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ndBytes));
    k_sym_nd<<<1, SYM_TPB, sp2.useSmem ? ndBytes : 0, stream>>>(sp2);
    int hc[4] = {0, 0, 0, 0};
    CK(cudaMemcpyAsync(hc, y_symCnt.p, 3 * sizeof(int), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    const int nsn = hc[1];
    // row structures (a bitmap per supernode), parents, levels, sizes
    const int W = (P + 31) / 32;
    y_bits.alloc((size_t)nsn * W);
    CK(cudaMemsetAsync(y_bits.p, 0, (size_t)nsn * W * sizeof(unsigned), stream));
    SymRowsPar rp{};
    rp.P = P; rp.maxw = CHOL_MAXW; rp.W = W; rp.adjPtr = y_adjPtr; rp.adj = y_adj; rp.verts = y_verts; rp.iperm = y_iperm;
    rp.snStart = y_snStart; rp.colToSn = y_colToSn; rp.emits = y_emits; rp.counters = y_symCnt; rp.bits = y_bits;
    rp.snParent = y_snParent; rp.snLevel = y_snLevel; rp.childHead = y_childHead; rp.childNext = y_childNext;
    rp.rcount = y_rcount; rp.snRowPtr = y_snRowPtr; rp.out = y_counts;
    k_sym_rows<<<1, SYM_TPB, 0, stream>>>(rp);
    CK(cudaMemcpyAsync(&counts, y_counts.p, sizeof(SymCounts), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    y_snRows.alloc((size_t)std::max(counts.nrows, 1));
    k_sym_extract<<<divUp((long long)nsn * 32, TB), TB, 0, stream>>>(y_bits, W, nsn, y_snRowPtr, y_snRows);
    // schedule, offsets, children, task tables; inverse maps; assembly lists
    z_levelPtr.alloc(counts.nlev + 1); z_levelSn.alloc(nsn); z_Lptr.alloc(nsn + 1); z_Uptr.alloc(nsn + 1);
    z_childPtr.alloc(nsn + 1); z_childIdx.alloc(nsn + 1); z_invPtr.alloc(nsn); z_inv.alloc((size_t)counts.ninv + 1);
    z_snBlkPtr.alloc(nsn + 1); z_asmBlk.alloc(nblkG); z_asmRow.alloc(nblkG); z_asmCol.alloc(nblkG);
    z_trsmPtr.alloc(nsn + 1); z_updPtr.alloc(nsn + 1); z_levMaxNt.alloc(counts.nlev + 1);
    SymMaps m{};
    m.levelPtr = z_levelPtr; m.levelSn = z_levelSn; m.Lptr = z_Lptr; m.Uptr = z_Uptr; m.childPtr = z_childPtr;
    m.childIdx = z_childIdx; m.invPtr = z_invPtr; m.inv = z_inv; m.snBlkPtr = z_snBlkPtr; m.asmBlk = z_asmBlk;
    m.asmRow = z_asmRow; m.asmCol = z_asmCol;
    const size_t tabBytes = (5 * (size_t)nsn + counts.nlev + 2) * sizeof(int);
    if (tabBytes > 200 * 1024) throw UnsupportedError("too many supernodes for the shared-memory schedule tables of this version");
    CK(cudaFuncSetAttribute(k_sym_tables, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tabBytes));
    k_sym_tables<<<1, SYM_TPB, tabBytes, stream>>>(nsn, counts.nlev, y_snStart, y_snRowPtr, y_snParent, y_snLevel, m, z_trsmPtr,
                                                   z_updPtr, panelRB > 0 ? panelRB : CHOL_RB, CHOL_TB, z_levMaxNt, z_factStats.alloc(4), counts.lsize);
    k_sym_inv<<<divUp((long long)nsn * 32, TB), TB, 0, stream>>>(nsn, y_snStart, y_snRowPtr, y_snRows, y_snParent, z_invPtr, z_inv);
    y_ak1.alloc(nblkG); y_ak2.alloc(nblkG); y_av1.alloc(nblkG); y_av2.alloc(nblkG);
    k_sym_asm_keys<<<divUp(nblkG, TB), TB, 0, stream>>>(gRowOfP, gColIndP, nblkG, y_iperm, y_colToSn, y_ak1, y_av1);
    sortPairs(*this, y_ak1.p, y_ak2.p, y_av1.p, y_av2.p, nblkG, bitsFor((unsigned long long)std::max(nsn, 1)));
    k_lower_bounds<<<divUp(nsn + 1, TB), TB, 0, stream>>>(y_ak2, nblkG, nsn, z_snBlkPtr);
    k_sym_asm_fill<<<divUp(nblkG, TB), TB, 0, stream>>>(y_ak2, y_av2, nblkG, gRowOfP, gColIndP, y_iperm, y_snStart, y_snRowPtr,
                                                       y_snRows, z_asmBlk, z_asmRow, z_asmCol);
}

// ---------------------------------------------------------------------------
// kernel sequences
// ---------------------------------------------------------------------------
void Solver::enqueueBuild() {
    static_assert(LM_CHUNK == 128, "k_build runs both passes with 128 threads");
    if (poseOnly) {       // no free landmarks: the pose-major pass alone
        if (E > 0) { launch(k_build_pose<128>, d.nPoseUnits, 128, 0, d, sp); mark("build_pose"); }
    } else if (fuseBuild && !profiling) {
        if (f32) launch(k_build<LM_CHUNK, float>, d.nPoseUnits + d.nChunks, LM_CHUNK, 0, d, sp);
        else launch(k_build<LM_CHUNK, double>, d.nPoseUnits + d.nChunks, LM_CHUNK, 0, d, sp);
        mark("build");
    } else {
        if (f32) launch(k_build_lm<LM_CHUNK, float>, d.nChunks, LM_CHUNK, 0, d, sp);
        else launch(k_build_lm<LM_CHUNK, double>, d.nChunks, LM_CHUNK, 0, d, sp);
        mark("build_lm");
        launch(k_build_pose<128>, d.nPoseUnits, 128, 0, d, sp); mark("build_pose");
    }
    if (nIcp > 0) {       // ICP edges add to Hpp / bp / the max diagonal of their poses
        launch(k_icp_build<256>, ic.nUnits, 256, 0, d, ic); mark("icp_build");
        launch(k_icp_add<256>, d.nP, 256, 0, d, ic, E == 0 ? 1 : 0); mark("icp_add");
    }
    if (nranks > 1 && withA) {   // max diagonal of the *global* system for lambda0
        launch(k_pack_A<256>, std::min(divUp(6 * d.nP, 256), 148), 256, 0, d); mark("pack_A");
        smallReduce(d.redA, 6 * d.nP + nranks);
    }
}

// Schur complement -> multifrontal Cholesky (levels up) -> backward solve (levels
// down) -> back-substitution (+ landmark part of the update)
void Solver::enqueueSolve() {
    if (poseOnly) {       // block-diagonal system: one 6x6 solve per pose (reference: dense LU of the whole 6P x 6P matrix)
        launch(k_po_solve<128>, divUp(d.nP, 128), 128, 0, d); mark("pose_solve");
        return;
    }
    launch(k_schur_inv<256>, divUp(d.nL, 256), 256, 0, d); mark("schur_inv");
    const int2* bp = schurPasses > 1 ? blkPass.p : nullptr;
    if (schurMma) {
        if (f32) launch(k_schur_rows<float, true>, nRowUnits, SROW_TPB, srow_smem_bytes(schurCap), d, schurCap, c.dbg, schurUnit, schurPasses, bp);
        else launch(k_schur_rows<double, true>, nRowUnits, SROW_TPB, srow_smem_bytes(schurCap), d, schurCap, c.dbg, schurUnit, schurPasses, bp);
    } else {
        if (f32) launch(k_schur_rows<float, false>, nRowUnits, SROW_TPB, srow_smem_bytes(schurCap), d, schurCap, c.dbg, schurUnit, schurPasses, bp);
        else launch(k_schur_rows<double, false>, nRowUnits, SROW_TPB, srow_smem_bytes(schurCap), d, schurCap, c.dbg, schurUnit, schurPasses, bp);
    }
    mark("schur_rows");
    // the one exchange step of a trial: partial reduced camera systems [Hsc | bsc] summed over the ranks
    if (nranks > 1) {
        const size_t nred = (size_t)d.nblkG * 36 + (size_t)d.nP * 6;
        if (peerOn) k_peer_allreduce<<<PEER_CTAS, PEER_TPB, 0, stream>>>(peer, nred, d.st);
        else NK(g_nccl.AllReduce(d.HscPart, d.Hsc, nred, ncclFloat64, ncclSum, comm, stream));
        mark("exchange");
    }
    // levels [0, zTop] solve backward with the inverse of L11 (k_chol_zinv has the whole rest of the factorisation
    // to finish); the top three keep the blocked substitution
    const bool zOn = zinvOn && pdl && !profiling;
    const int zTop = c.nlev - 4;
    bool forked = false;             // the side branches carry at least one k_chol_zinv
    int nz = 0;
    for (int lev = 0; lev < c.nlev; lev++) {
        const int q0 = h_levelPtr[lev], q1 = h_levelPtr[lev + 1];
        const int ncLev = 6 * h_levMaxNt[lev];
        if (distLc > 0 && lev == distLc && !distBcast.empty()) {
            // the subtrees below the cut are complete on their owners: their roots' update matrices go to every rank
            NK(g_nccl.GroupStart());
            for (const Bcast& bc : distBcast)
                NK(g_nccl.Broadcast(c.U + bc.off, c.U + bc.off, bc.count, ncclFloat64, bc.root, comm, stream));
            NK(g_nccl.GroupEnd());
        }
        if (q1 == q0) continue;          // (distributed: none of this level's supernodes is this rank's)
        launch(panelDC ? k_chol_panel3<true> : k_chol_panel3<false>, h_trsmPtr[q1] - h_trsmPtr[q0], PANEL3_TPB,
               (size_t)PANEL_LD * ncLev * 8, d, c, lev, (pdl && !profiling && lev > 0 && lev != distLc) ? 1 : 0);
        mark("chol_panel");
        if (zOn && lev <= zTop) {      // L11^-1 of the level on the low-priority branch, beside the rest of the factorisation
            cudaStream_t sd = (nz & 1) ? side2 : side;
            CK(cudaEventRecord(evFork, stream));
            CK(cudaStreamWaitEvent(sd, evFork, 0));
            const size_t zb = (2 * (size_t)ncLev * (ncLev + 1) + 21 * CHOL_MAXW + 36 * 8) * 8;
            k_chol_zinv<256><<<q1 - q0, 256, zb, sd>>>(d, c, lev);
            kernelsPerSlot++;
            nz++;
            forked = true;
        }
        const int nUpd = h_updPtr[q1] - h_updPtr[q0];
        if (nUpd > 0) {
            const int kpad = (ncLev + 15) & ~15;
            if (h_tbLev[lev] == UPD_TBS)
                launch(k_chol_update_small, nUpd, UPD3_TPB,
                       (size_t)std::max(2 * kpad * UPD_LD, (UPD3_TPB / 32) * 6 * UPD_TBS * UPD_SCLD) * 8, d, c, lev, kpad);
            else
                launch(k_chol_update3, nUpd, UPD3_TPB, (size_t)std::max(2 * kpad * UPD_LD, 2 * 6 * CHOL_TB * UPD_CLD) * 8, d, c, lev, kpad);
            mark("chol_update");
        }
    }
    for (int lev = c.nlev - 1; lev >= 0; lev--) {
        const int ncLev = 6 * h_levMaxNt[lev];
        const size_t bytes = ((size_t)ncLev * (ncLev + 1) + ncLev + 21 * CHOL_MAXW + 6 * (size_t)maxRt) * 8;
        if (zOn && forked && lev == zTop) {
            CK(cudaEventRecord(evJoin, side));
            CK(cudaStreamWaitEvent(stream, evJoin, 0));
            if (nz > 1) {
                CK(cudaEventRecord(evJoin2, side2));
                CK(cudaStreamWaitEvent(stream, evJoin2, 0));
            }
        }
        // few supernodes on a level: several CTAs share the L21^T x_R product of each
        const int nsnLev = h_levelPtr[lev + 1] - h_levelPtr[lev];
        if (nsnLev == 0) continue;
        const int G = bwdSplit ? std::max(1, std::min(BWD_MAXG, 148 / nsnLev)) : 1;
        launch(k_chol_bwd3<256>, nsnLev * G, 256, bytes, d, c, lev, (pdl && !profiling && lev < c.nlev - 1) ? 1 : 0,
               (zOn && lev <= zTop) ? 1 : 0, G);
        mark("chol_bwd");
    }
    if (distLc > 0) {      // every rank gets the whole x: each piece from the rank that solved it
        k_x_mask<<<divUp(6 * d.nP, 256), 256, 0, stream>>>(zr_colOwn, d.nP, rank, c.xperm);
        allReduceSum(c.xperm, (size_t)6 * d.nP);
        k_x_scatter<<<divUp(6 * d.nP, 256), 256, 0, stream>>>(c.xperm, c.perm, d.nP, d.x);
        kernelsPerSlot += 2;
    }
    if (f32) launch(k_backsub<LM_CHUNK, float>, d.nChunks, LM_CHUNK, 0, d);
    else launch(k_backsub<LM_CHUNK, double>, d.nChunks, LM_CHUNK, 0, d);
    mark("backsub");
}

// one LM trial: [build] prepare solve update errors decide
void Solver::enqueueSlot() {
    kernelsPerSlot = 0;
    enqueueBuild();
    launch(k_lm_prepare<1024>, 1, 1024, 0, d); mark("lm_prepare");
    enqueueSolve();
    launch(k_update_poses<128>, divUp(d.nP, 128), 128, 0, d); mark("update_poses");
    enqueueErrors(1, 1, true);
    if (nranks > 1 && peerOn) {
        launch(k_lm_decide_peer<1024>, 1, 1024, 0, d, peer); mark("lm_decide");       // sums, exchange and decision in one kernel
        return;
    }
    if (nranks > 1) {
        launch(k_lm_sums<256>, 1, 256, 0, d); mark("lm_sums");
        smallReduce(d.redC, 3);
    }
    launch(k_lm_decide<1024>, 1, 1024, 0, d); mark("lm_decide");
}

// Runs ONE trial slot outside the graph with an event after every kernel and returns
// the per-kernel device times (ms).  The LM state must allow a trial (call after
// set_estimates): it performs the first trial of a 1-iteration optimisation.
int Solver::profileSlot(int cap, char* names, int nameLen, double* ms) {
    ensureStructure();
    use();
    enqueueErrors(0, 0, false);
    if (nranks > 1) { k_lm_sums<256><<<1, 256, 0, stream>>>(d); smallReduce(d.redC, 3); }
    k_lm_begin<256><<<1, 256, 0, stream>>>(d, 1);
    const int saved = kernelsPerSlot;
    profiling = true;
    profEvents.clear(); profNames.clear();
    cudaEvent_t e0;
    CK(cudaEventCreate(&e0));
    CK(cudaEventRecord(e0, stream));
    enqueueSlot();
    profiling = false;
    kernelsPerSlot = saved;
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    int n = 0;
    cudaEvent_t prev = e0;
    for (size_t i = 0; i < profEvents.size(); i++) {
        float t = 0;
        CK(cudaEventElapsedTime(&t, prev, profEvents[i]));
        if (n < cap) {
            if (names) { strncpy(names + (size_t)n * nameLen, profNames[i].c_str(), nameLen - 1); names[(size_t)n * nameLen + nameLen - 1] = 0; }
            if (ms) ms[n] = t;
            n++;
        }
        prev = profEvents[i];
    }
    CK(cudaEventDestroy(e0));
    for (auto e : profEvents) cudaEventDestroy(e);
    profEvents.clear();
    systemBuilt = false; solved = false;
    return n;
}

// New measurements / information values for an unchanged graph (a second
// initialize() of the reference re-uploads the whole edge arena but keeps the
// structure: src/block_solver.cpp:161-170).  H2D into the raw arrays, then the two
// gather kernels refresh the landmark-major and pose-major copies.
__global__ void k_gather_meas(const double* m0, const double* m1, const double* m2, const double* w, int n, int dim,
                              double* meas, double* omega) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    meas[(size_t)i * dim] = m0[i]; meas[(size_t)i * dim + 1] = m1[i];
    if (dim == 3) meas[(size_t)i * 3 + 2] = m2[i];
    omega[i] = w[i];
}

// Host copies of sets whose measurements were replaced on the device only (the fast path of
// updateMeasurements) are brought up to date before the raw arrays are rebuilt.
void Solver::refreshHostSets() {
    size_t off = 0;
    for (HostSet& hs : sets) {
        if (hs.stale && hs.n > 0 && off + hs.n <= (size_t)r_m0.cap) {
            use();
            u_meas.alloc((size_t)hs.n * hs.cols()); u_w.alloc(hs.n);
            k_gather_meas<<<divUp(hs.n, 256), 256, 0, stream>>>(r_m0.p + off, r_m1.p + off, r_m2.p + off, r_w.p + off, hs.n,
                                                               hs.cols(), u_meas, u_w);
            hs.meas.resize((size_t)hs.n * hs.cols()); hs.omega.resize(hs.n); hs.nomega = hs.n;
            CK(cudaMemcpyAsync(hs.meas.data(), u_meas.p, hs.meas.size() * sizeof(double), cudaMemcpyDeviceToHost, stream));
            CK(cudaMemcpyAsync(hs.omega.data(), u_w.p, hs.omega.size() * sizeof(double), cudaMemcpyDeviceToHost, stream));
            CK(cudaStreamSynchronize(stream));
        }
        hs.stale = false;
        off += hs.n;
    }
}

void Solver::updateMeasurements(int setId, const double* meas, const double* omega, int nomega) {
    if (setId < 0 || setId >= (int)sets.size()) throw ArgError("update_measurements: bad set id");
    HostSet& hs = sets[setId];
    if (hs.n == 0) return;        // the reference skips empty sets
    if (!meas || !omega || (nomega != 1 && nomega != hs.n)) throw ArgError("update_measurements: bad arguments");
    if (structDirty) {            // no device arrays yet: keep the host copy, the next build_structure uploads it
        refreshHostSets();
        hs.meas.assign(meas, meas + (size_t)hs.n * hs.cols());
        hs.omega.assign(omega, omega + nomega);
        hs.nomega = nomega;
        return;
    }
    // structure unchanged: the caller's buffers go straight to the device (asynchronous DMA when they are
    // pinned), are split into the SoA arrays there, and the host copy is only marked stale
    use();
    size_t off = 0;
    for (int i = 0; i < setId; i++) off += sets[i].n;
    const int n = hs.n;
    u_meas.alloc((size_t)n * hs.cols()); u_w.alloc(std::max(nomega, 1));
    CK(cudaMemcpyAsync(u_meas.p, meas, (size_t)n * hs.cols() * sizeof(double), cudaMemcpyHostToDevice, stream));
    CK(cudaMemcpyAsync(u_w.p, omega, (size_t)nomega * sizeof(double), cudaMemcpyHostToDevice, stream));
    k_scatter_meas<<<divUp(n, 256), 256, 0, stream>>>(u_meas, u_w, n, hs.cols(), nomega, r_m0.p + off, r_m1.p + off,
                                                      r_m2.p + off, r_w.p + off);
    hs.stale = true;
    RawEdges raw{(int)E, r_iP, r_iL, r_m0, r_m1, r_m2, r_w, r_set, perEdgeCam_ ? r_cam.p : nullptr,
                 anyOut_ ? r_out.p : nullptr};
    const int TB = 256;
    k_gather_lm<<<divUp(E + 1, TB), TB, 0, stream>>>(raw, s_gid, nP, nL, s_iP, s_iL, s_uv, s_r, s_w, s_set,
                                                     perEdgeCam_ ? s_cam.p : nullptr, anyOut_ ? s_out.p : nullptr, flag);
    k_gather_pose<<<divUp(E, TB), TB, 0, stream>>>((int)E, p_src, s_iL, s_uv, s_r, s_w, s_set,
                                                   perEdgeCam_ ? s_cam.p : nullptr, anyOut_ ? s_out.p : nullptr, p_iL,
                                                   p_uv, p_r, p_w, p_set, perEdgeCam_ ? p_cam.p : nullptr,
                                                   anyOut_ ? p_out.p : nullptr);
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    systemBuilt = false; solved = false;
}

void Solver::captureSlot() {
    use();
    if (!h_state) CK(cudaMallocHost(&h_state, sizeof(LMState)));
    if (!h_scalars) CK(cudaMallocHost(&h_scalars, 8 * sizeof(double)));
    if (!ev0) { CK(cudaEventCreate(&ev0)); CK(cudaEventCreate(&ev1)); }
    auto capture = [&](bool a, cudaGraph_t* g, cudaGraphExec_t* ex) {
        withA = a;
        CK(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
        try {
            enqueueSlot();
        } catch (...) {
            // leave neither the stream in capture mode nor a half-recorded graph behind; the structure stays dirty
            cudaGraph_t partial = nullptr;
            cudaStreamEndCapture(stream, &partial);
            if (partial) cudaGraphDestroy(partial);
            cudaGetLastError();
            structDirty = true;
            withA = true;
            throw;
        }
        withA = true;
        CK(cudaStreamEndCapture(stream, g));
        CK(cudaGetLastError());
        CK(cudaGraphInstantiate(ex, *g, 0));
    };
    if (nranks > 1) {
        capture(true, &slotGraph0, &slotExec0);
        capture(false, &slotGraph, &slotExec);
    } else {
        capture(true, &slotGraph, &slotExec);
    }
}

// ---------------------------------------------------------------------------
// staged (host-driven) API: same kernels, one BlockSolver method per call
// ---------------------------------------------------------------------------
static double readScalar(Solver& S, int slot) {
    CK(cudaMemcpyAsync(S.h_scalars + slot, S.scalars.p + slot, sizeof(double), cudaMemcpyDeviceToHost, S.stream));
    CK(cudaStreamSynchronize(S.stream));
    CK(cudaGetLastError());
    return S.h_scalars[slot];
}

void Solver::commInit(int n, int r, const char* id) {
    if (n < 1 || r < 0 || r >= n || !id) throw ArgError("comm_init: bad rank / size");
    if (comm) throw StateError("comm_init: this solver already has a communicator");
    use();
    if (n == 1) return;
    if (!g_nccl.load(err)) throw CudaError(err);
    ncclUniqueId uid;
    static_assert(sizeof(uid) == 128, "ncclUniqueId is 128 bytes");
    memcpy(&uid, id, sizeof(uid));
    NK(g_nccl.CommInitRank(&comm, n, uid, r));
    nranks = n; rank = r;
    // one collective outside any capture: connections are set up before the trial graph is recorded
    redC.alloc(4);
    CK(cudaMemsetAsync(redC.p, 0, 4 * sizeof(double), stream));
    allReduceSum(redC.p, 2);
    CK(cudaStreamSynchronize(stream));
    structDirty = true;
}

double Solver::computeErrors() {
    requireSingle("compute_errors");
    ensureStructure();
    use();
    enqueueErrors(0, 0, false);
    k_reduce<256><<<1, 256, 0, stream>>>(d.chiPartial, d.nChi, nullptr, 0, 0, scalars.p);
    return readScalar(*this, 0);
}

void Solver::buildSystem() {
    requireSingle("build_system");
    ensureStructure();
    use();
    k_state_set<<<1, 1, 0, stream>>>(d.st, stagedLambda, 1);
    enqueueBuild();
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    systemBuilt = true; solved = false;
}

double Solver::maxDiagonal() {
    if (!systemBuilt) throw StateError("max_diagonal before build_system");
    use();
    k_reduce<256><<<1, 256, 0, stream>>>(d.maxDiagL, d.nChunks, d.maxDiagP, d.nP, 1, scalars.p + 1);
    return readScalar(*this, 1);
}

bool Solver::solve() {
    requireSingle("solve");
    if (!systemBuilt) throw StateError("solve before build_system");
    use();
    k_state_set<<<1, 1, 0, stream>>>(d.st, stagedLambda, 0);
    enqueueSolve();
    CK(cudaMemcpyAsync(h_state, d.st, sizeof(LMState), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    solved = true;
    return h_state->fail == 0;
}

void Solver::update() {
    if (!solved) throw StateError("update before solve");
    use();
    // k_backsub already wrote the trial landmarks; poses follow, then the trial buffer becomes current
    k_update_poses<128><<<divUp(d.nP, 128), 128, 0, stream>>>(d);
    k_state_flip<<<1, 1, 0, stream>>>(d.st);
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    solved = false;
}

double Solver::computeScale(double lambda) {
    if (!systemBuilt) throw StateError("compute_scale before build_system");
    use();
    if (lambda == stagedLambda) {   // the landmark partials of k_backsub were formed with the lambda of the last solve()
        k_scale_poses<256><<<1, 256, 0, stream>>>(d, stagedLambda, scalars.p + 2);
    } else {                        // any other lambda: the whole sum over [xp | xl] with the caller's value
        k_scale_full<1024><<<1, 1024, 0, stream>>>(d, lambda, scalars.p + 2);
    }
    return readScalar(*this, 2);
}

// ---------------------------------------------------------------------------
// device-resident LM
// ---------------------------------------------------------------------------
int Solver::optimize(int niter, int cap, double* chi2, double* lambda, int* trials) {
    ensureStructure();
    use();
    // any iteration count is accepted, as in the reference; statistics are kept for the first MAX_STATS iterations
    CK(cudaEventRecord(ev0, stream));
    long launched = 0;
    enqueueErrors(0, 0, false);
    if (nranks > 1) { k_lm_sums<256><<<1, 256, 0, stream>>>(d); smallReduce(d.redC, 3); launched++; }
    k_lm_begin<256><<<1, 256, 0, stream>>>(d, niter);
    launched += 2 + (nIcp > 0 ? 1 : 0);
    int slots = 0;
    const int maxSlots = niter * 10 + 1;
    int batch = niter;
    while (niter > 0) {
        for (int i = 0; i < batch; i++) CK(cudaGraphLaunch((slots == 0 && i == 0 && slotExec0) ? slotExec0 : slotExec, stream));
        slots += batch;
        launched += (long)batch * kernelsPerSlot;
        CK(cudaMemcpyAsync(h_state, d.st, sizeof(LMState), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        if (h_state->done || slots >= maxSlots) break;
        batch = std::max(1, std::min(2, maxSlots - slots));   // rejected trials: top up
    }
    if (niter <= 0) {
        CK(cudaMemcpyAsync(h_state, d.st, sizeof(LMState), cudaMemcpyDeviceToHost, stream));
    }
    if (gating()) { updateEdges(); launched += 3; }     // src/cuda_graph_optimisation.cpp:150-151
    CK(cudaEventRecord(ev1, stream));
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, ev0, ev1));
    lastOptimizeMs = ms; lastKernels = launched;
    const int n = std::min(h_state->nstats, MAX_STATS);
    for (int i = 0; i < n && i < cap; i++) {
        if (chi2) chi2[i] = h_state->statChi2[i];
        if (lambda) lambda[i] = h_state->statLambda[i];
        if (trials) trials[i] = h_state->statTrials[i];
    }
    systemBuilt = false; solved = false;
    return n;
}

// ---------------------------------------------------------------------------
// outlier gating.  Reference: EdgeSet::setOutlierThreshold src/optimisable_graph.hpp:415-420,
// BlockSolver::updateEdges src/block_solver.cpp:423-429 -> EdgeSet::updateEdges src/optimisable_graph.hpp:603-640 ->
// computeOutliersKernel cuda_block_solver.cu:1135-1149, called at the end of every optimize()
// (src/cuda_graph_optimisation.cpp:150-151).  There the mask goes to the host, the edges are inactivated and the
// structure is rebuilt; here the flags stay on the device (raw order r_out + the two sorted copies), flagged edges
// weigh nothing in every kernel, their Hpl blocks are written as zeros and the structure, its symbolic analysis and
// the captured trial graph are kept.
// ---------------------------------------------------------------------------
struct OutThr { double thr[MAX_SETS]; };

// chi2 of every unflagged edge at the estimates the last k_errors saw (st->lastEval: the reference keeps the
// chiValues of its last computeErrors, a rejected trial included) against its set's threshold
template <int TPB>
__global__ void __launch_bounds__(TPB) k_compute_outliers(Dev d, SetParams sp, OutThr ot, uint8_t* s_out, uint8_t* r_out) {
    const LMState* st = d.st;
    const int buf = st->lastEval;
    const double* __restrict__ pose = d.pose[buf];
    const double* __restrict__ lm = d.lm[buf];
    for (int k = blockIdx.x * TPB + threadIdx.x; k < d.E; k += gridDim.x * TPB) {
        if (s_out[k]) continue;
        const int set = d.s_set[k];
        const double thr = ot.thr[set];
        if (!(thr > 0)) continue;
        const int iP = __ldg(d.s_iP + k), iL = __ldg(d.s_iL + k);
        const int dim = sp.dim[set];
        double q[4], t[3], cam[5], Xc[3], e[3];
        load_pose(pose, iP, q, t);
        const double X[3] = {__ldg(lm + 3 * (size_t)iL), __ldg(lm + 3 * (size_t)iL + 1), __ldg(lm + 3 * (size_t)iL + 2)};
        load_cam(d, sp, d.s_cam, k, set, cam);
        const double2 uv = __ldg(d.s_uv + k);
        const double mr = (dim != 2) ? __ldg(d.s_r + k) : 0.0;
        world_to_camera(q, t, X, Xc);
        residual(dim, Xc, cam, uv.x, uv.y, mr, e);
        const double chi = rk_rho(sp.rkType[set], sp.rkDelta[set], __ldg(d.s_w + k) * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]));
        if (chi > thr) { s_out[k] = 1; r_out[d.s_gid[k]] = 1; }
    }
}
__global__ void k_mask_to_pose(const int* p_src, const uint8_t* s_out, int E, uint8_t* p_out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < E) p_out[k] = s_out[p_src[k]];
}
// flagged edges per set (sets are contiguous ranges of the raw order); one CTA, integer sums
template <int TPB>
__global__ void __launch_bounds__(TPB) k_count_mask(const uint8_t* r_out, OutThr bounds /* prefix of set sizes as doubles */, int nsets, int* cnt) {
    __shared__ int sh[MAX_SETS];
    if (threadIdx.x < MAX_SETS) sh[threadIdx.x] = 0;
    __syncthreads();
    for (int s = 0; s < nsets; s++) {
        const int b0 = s ? (int)bounds.thr[s - 1] : 0, b1 = (int)bounds.thr[s];
        int c = 0;
        for (int i = b0 + threadIdx.x; i < b1; i += TPB) c += r_out[i] ? 1 : 0;
        if (c) atomicAdd(&sh[s], c);
    }
    __syncthreads();
    if (threadIdx.x < MAX_SETS) cnt[threadIdx.x] = sh[threadIdx.x];
}

void Solver::setOutlierThreshold(int setId, double thr) {
    if (setId < 0 || setId >= (int)sets.size()) throw ArgError("set_outlier_threshold: bad set id");
    if (!(thr >= 0)) throw ArgError("set_outlier_threshold: threshold must be >= 0");
    sets[setId].outThr = thr;
    if (structDirty || !(thr > 0) || d.s_out) return;
    // structure is current but was built without mask arrays: add them (all clear) and re-record the trial graph
    use();
    r_out.alloc(E); s_out.alloc(E); p_out.alloc(E);
    CK(cudaMemsetAsync(r_out.p, 0, E, stream));
    CK(cudaMemsetAsync(s_out.p, 0, E, stream));
    CK(cudaMemsetAsync(p_out.p, 0, E, stream));
    anyOut_ = true;
    d.s_out = s_out; d.p_out = p_out;
    dropGraphs();
    captureSlot();
}

void Solver::updateEdges() {
    if (!gating()) return;
    ensureStructure();
    use();
    OutThr ot{}, bounds{};
    size_t off = 0;
    for (size_t i = 0; i < sets.size(); i++) { ot.thr[i] = sets[i].outThr; off += sets[i].n; bounds.thr[i] = (double)off; }
    outCnt.alloc(MAX_SETS);
    k_compute_outliers<256><<<d.errGrid, 256, 0, stream>>>(d, sp, ot, s_out.p, r_out.p);
    k_mask_to_pose<<<divUp(E, 256), 256, 0, stream>>>(p_src, s_out, (int)E, p_out);
    k_count_mask<1024><<<1, 1024, 0, stream>>>(r_out, bounds, (int)sets.size(), outCnt);
    int cnt[MAX_SETS];
    CK(cudaMemcpyAsync(cnt, outCnt.p, sizeof(cnt), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    CK(cudaGetLastError());
    for (size_t i = 0; i < sets.size(); i++) {
        if (cnt[i] != sets[i].outCount) maskLive = true;
        sets[i].outCount = cnt[i];
    }
    systemBuilt = false; solved = false;
}

// before the raw arrays are rebuilt: flags made on the device become part of the host sets' masks, so a rebuilt
// structure leaves those edges' Hpl blocks out (the reference's rebuild after updateEdges)
void Solver::syncOutliersToHost() {
    if (!maskLive || !r_out.p) { maskLive = false; return; }
    use();
    std::vector<uint8_t> m(r_out.cap);
    size_t tot = 0;
    for (auto& hs : sets) tot += hs.n;
    if (tot > r_out.cap) { maskLive = false; return; }
    CK(cudaMemcpyAsync(m.data(), r_out.p, tot, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    size_t off = 0;
    for (auto& hs : sets) {
        if (hs.outCount > 0) {
            hs.out.assign(hs.n, 0);
            for (int i = 0; i < hs.n; i++) hs.out[i] = m[off + i];
        }
        off += hs.n;
    }
    maskLive = false;
}

void Solver::getOutliers(int setId, int* mask, int* outliers, int* inliers) {
    if (setId < 0 || setId >= (int)sets.size()) throw ArgError("get_outliers: bad set id");
    HostSet& hs = sets[setId];
    if (outliers) *outliers = hs.outCount;
    if (inliers) *inliers = hs.n - hs.outCount;
    if (!mask || hs.n == 0) return;
    if (structDirty || !d.s_out) {          // nothing on the device yet: the host mask (or all clear)
        for (int i = 0; i < hs.n; i++) mask[i] = hs.out.empty() ? 0 : (hs.out[i] ? 1 : 0);
        return;
    }
    use();
    size_t off = 0;
    for (int i = 0; i < setId; i++) off += sets[i].n;
    std::vector<uint8_t> m(hs.n);
    CK(cudaMemcpyAsync(m.data(), r_out.p + off, hs.n, cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    for (int i = 0; i < hs.n; i++) mask[i] = m[i];
}

void Solver::getEstimates(double* poses7, double* lms) {
    if (!haveVertices) throw StateError("get_estimates before set_vertices");
    use();
    LMState st;
    CK(cudaMemcpyAsync(&st, d.st, sizeof(LMState), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    const int cur = st.cur;
    std::vector<double> p8((size_t)nP_all * POSE_STRIDE);
    CK(cudaMemcpyAsync(p8.data(), d.pose[cur], p8.size() * sizeof(double), cudaMemcpyDeviceToHost, stream));
    if (lms && nL_all > 0)
        CK(cudaMemcpyAsync(lms, d.lm[cur], (size_t)nL_all * 3 * sizeof(double), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    if (poses7)
        for (int i = 0; i < nP_all; i++)
            for (int k = 0; k < 7; k++) poses7[(size_t)i * 7 + k] = p8[(size_t)i * POSE_STRIDE + k];
}

static int readCur(Solver& S) {
    LMState st;
    CK(cudaMemcpyAsync(&st, S.d.st, sizeof(LMState), cudaMemcpyDeviceToHost, S.stream));
    CK(cudaStreamSynchronize(S.stream));
    return st.cur;
}

// ---------------------------------------------------------------------------
// parity-only structure outputs in the reference's formats (built on demand)
// ---------------------------------------------------------------------------
// scalar CSR of the full symmetric Hsc + BSR2CSR gather map, restating
// HschurSparseBlockMatrix::convertBSRToCSR (src/sparse_block_matrix.cpp:158-262):
// scalar row 6R+dr = [transposes of blocks (R',R), R'<R ascending] then [row blocks
// ascending]; BSR2CSR = 36*blk + 6*dc + dr for an entry and its mirror.
__global__ void k_csr_counts(const int* bRowPtr, const int* bColInd, int nP, int* upCnt) {
    // upCnt[C] = number of blocks (R, C) with R < C
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= bRowPtr[nP]) return;
    int lo = 0, hi = nP;       // row of block k
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (bRowPtr[mid] <= k) lo = mid; else hi = mid; }
    if (bColInd[k] != lo) atomicAdd(&upCnt[bColInd[k]], 1);
}
__global__ void k_csr_rowptr(const int* bRowPtr, const int* upCnt, int nP, int* blkRowCnt) {
    const int R = blockIdx.x * blockDim.x + threadIdx.x;
    if (R > nP) return;
    blkRowCnt[R] = (R < nP) ? (bRowPtr[R + 1] - bRowPtr[R] + upCnt[R]) : 0;
}
__global__ void k_csr_fill(const int* bRowPtr, const int* bColInd, const int* bRowOf, const int* blkRowOff,
                           const int* upCnt, int nP, int nblk, int* rowPtr, int* colInd, int* bsr2csr) {
    // one thread per (block k, dc, dr)
    const long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (id <= 6LL * nP) {
        const int r = (int)id;
        if (r < 6 * nP) rowPtr[r] = 36 * blkRowOff[r / 6] + (r % 6) * 6 * (blkRowOff[r / 6 + 1] - blkRowOff[r / 6]);
        else rowPtr[r] = 36 * blkRowOff[nP];
    }
    if (id >= 36LL * nblk) return;
    const int k = (int)(id / 36), e = (int)(id % 36), dc = e / 6, dr = e % 6;
    const int R = bRowOf[k], C = bColInd[k];
    const int src = 36 * k + 6 * dc + dr;
    {   // entry (6R+dr, 6C+dc): position among row R's own blocks, after its upCnt[R] mirrored blocks
        const int nb = blkRowOff[R + 1] - blkRowOff[R];
        const int base = 36 * blkRowOff[R] + dr * 6 * nb;
        const int pos = upCnt[R] + (k - bRowPtr[R]);
        colInd[base + 6 * pos + dc] = 6 * C + dc;
        bsr2csr[base + 6 * pos + dc] = src;
    }
    if (R != C) {   // mirror (6C+dc, 6R+dr): rank of R among the blocks (R', C), R' < C
        int rank = 0;
        for (int Rp = 0; Rp < R; Rp++) {   // count rows R' < R that hold a block in column C
            int lo = bRowPtr[Rp], hi = bRowPtr[Rp + 1];
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (bColInd[mid] < C) lo = mid + 1; else hi = mid; }
            if (lo < bRowPtr[Rp + 1] && bColInd[lo] == C) rank++;
        }
        const int nb = blkRowOff[C + 1] - blkRowOff[C];
        const int base = 36 * blkRowOff[C] + dc * 6 * nb;
        colInd[base + 6 * rank + dr] = 6 * R + dr;
        bsr2csr[base + 6 * rank + dr] = src;
    }
}
// product list in the reference's order: (i, j, k) sorted by (i, j)
__global__ void k_widen(const float* a, int n, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)a[i];
}
__global__ void k_mul_keys(const int2* prodIJ, const int* blkOf, const int* head, int n, unsigned long long nHpl,
                           unsigned long long* keys, int* vals) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int2 ij = prodIJ[p];
    keys[p] = ij.x < 0 ? 0xffffffffffffffffULL : (unsigned long long)ij.x * nHpl + (unsigned long long)ij.y;
    vals[p] = blkOf[p] + head[p] - 1;   // blkOf is the exclusive scan of the head flags
}
__global__ void k_mul_out(const unsigned long long* keys, const int* vals, int nmul, unsigned long long nHpl, int* out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= nmul) return;
    out[3 * p] = (int)(keys[p] / nHpl);
    out[3 * p + 1] = (int)(keys[p] % nHpl);
    out[3 * p + 2] = vals[p];
}

long Solver::getStructure(int which, int* out, long cap) {
    ensureStructure();
    use();
    if (poseOnly) {       // a graph without free landmarks has no Hpl / Hsc pattern and no elimination tree
        if (which == CUGO_B200_HPL_COLPTR || which == CUGO_B200_HPL_ROWIND) return 0;
        if (which != CUGO_B200_EDGE2HPL || E == 0) throw UnsupportedError("get_structure: pose-only graph (block-diagonal Hpp, no Schur structure)");
    }
    auto copyOut = [&](const int* dptr, long n) {
        if (out && cap >= n && n > 0) {
            CK(cudaMemcpyAsync(out, dptr, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
            CK(cudaStreamSynchronize(stream));
        }
        return n;
    };
    switch (which) {
    case CUGO_B200_HPL_COLPTR: return copyOut(colPtr, nL + 1);
    case CUGO_B200_HPL_ROWIND: return copyOut(rowInd, d.nHpl);
    case CUGO_B200_HSC_ROWPTR: return copyOut(bRowPtr, nP + 1);
    case CUGO_B200_HSC_COLIND: return copyOut(bColInd, d.nblk);
    case CUGO_B200_EDGE2HPL: return copyOut(edge2Hpl, E);
    case CUGO_B200_PERM: return copyOut(y_verts, nP);
    case CUGO_B200_SN_START: return copyOut(y_snStart, counts.nsn + 1);
    case CUGO_B200_SN_ROWPTR: return copyOut(y_snRowPtr, counts.nsn + 1);
    case CUGO_B200_SN_ROWS: return copyOut(y_snRows, counts.nrows);
    case CUGO_B200_SN_PARENT: return copyOut(y_snParent, counts.nsn);
    case CUGO_B200_SN_LEVEL: return copyOut(y_snLevel, counts.nsn);
    case CUGO_B200_LEVEL_SN: return copyOut(z_levelSn, counts.nsn);
    case CUGO_B200_CHILD_IDX: return copyOut(z_childIdx, h_nchild);
    case CUGO_B200_INV_MAP: return copyOut(z_inv, counts.ninv);
    case CUGO_B200_ASM_BLK: return copyOut(z_asmBlk, nblkG);
    case CUGO_B200_ASM_ROW: return copyOut(z_asmRow, nblkG);
    case CUGO_B200_ASM_COL: return copyOut(z_asmCol, nblkG);
    case CUGO_B200_CSR_ROWPTR:
    case CUGO_B200_CSR_COLIND:
    case CUGO_B200_BSR2CSR: {
        const long n = (which == CUGO_B200_CSR_ROWPTR) ? 6L * nP + 1 : nnzSymm;
        if (!out || cap < n) return n;
        DBuf<int> upCnt, rowCnt, rowOff, rp, ci, map;
        upCnt.alloc(nP + 1); rowCnt.alloc(nP + 1); rowOff.alloc(nP + 1);
        rp.alloc(6 * (size_t)nP + 1); ci.alloc(nnzSymm); map.alloc(nnzSymm);
        CK(cudaMemsetAsync(upCnt.p, 0, (nP + 1) * sizeof(int), stream));
        k_csr_counts<<<divUp(d.nblk, 256), 256, 0, stream>>>(bRowPtr, bColInd, nP, upCnt);
        k_csr_rowptr<<<divUp(nP + 1, 256), 256, 0, stream>>>(bRowPtr, upCnt, nP, rowCnt);
        exclusiveScan(*this, rowCnt.p, rowOff.p, nP + 1);
        const long long work = std::max(36LL * d.nblk, 6LL * nP + 1);
        k_csr_fill<<<divUp(work, 256), 256, 0, stream>>>(bRowPtr, bColInd, bRowOf, rowOff, upCnt, nP, d.nblk, rp, ci, map);
        const int* src = which == CUGO_B200_CSR_ROWPTR ? rp.p : (which == CUGO_B200_CSR_COLIND ? ci.p : map.p);
        CK(cudaMemcpyAsync(out, src, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        CK(cudaGetLastError());
        return n;
    }
    case CUGO_B200_MULIDS: {
        const long n = 3L * nmulRef;
        if (!out || cap < n) return n;
        DBuf<unsigned long long> k1, k2;
        DBuf<int> v1, v2, o;
        k1.alloc(d.nmul); k2.alloc(d.nmul); v1.alloc(d.nmul); v2.alloc(d.nmul); o.alloc(std::max<long>(n, 1));
        k_mul_keys<<<divUp(d.nmul, 256), 256, 0, stream>>>(prodIJ, blkOf, head, d.nmul, (unsigned long long)d.nHpl, k1, v1);
        sortPairs(*this, k1.p, k2.p, v1.p, v2.p, d.nmul, 64);
        k_mul_out<<<divUp(std::max(nmulRef, 1), 256), 256, 0, stream>>>(k2, v2, nmulRef, (unsigned long long)d.nHpl, o);
        CK(cudaMemcpyAsync(out, o.p, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        CK(cudaGetLastError());
        return n;
    }
    default: throw ArgError("get_structure: unknown array id");
    }
}

long Solver::getValues(int which, double* out, long cap) {
    ensureStructure();
    use();
    const double* src = nullptr;
    long n = 0;
    switch (which) {
    case CUGO_B200_HPP: src = Hpp; n = 36L * nP; break;
    case CUGO_B200_HLL: src = Hll; n = 9L * nL; break;
    case CUGO_B200_HPL:
        n = 18L * d.nHpl;
        if (f32) {                       // widened copy of the single-precision blocks
            if (!out || cap < n) return n;
            chiEdge.alloc((size_t)std::max(n, 1L));
            k_widen<<<divUp((int)n, 256), 256, 0, stream>>>(HplF.p, (int)n, chiEdge.p);
            src = chiEdge;
        } else {
            src = Hpl;
        }
        break;
    case CUGO_B200_B: src = b; n = 6L * nP + 3L * nL; break;
    case CUGO_B200_X: src = x; n = 6L * nP + 3L * nL; break;
    case CUGO_B200_BSC: src = d.bsc; n = poseOnly ? 0 : 6L * nP; break;
    case CUGO_B200_HSC: src = d.Hsc; n = poseOnly ? 0 : 36L * d.nblkG; break;
    case CUGO_B200_INVHLL: src = invHll; n = 9L * nL; break;
    case CUGO_B200_CHI_PER_EDGE: {
        n = E;
        if (!out || cap < n) return n;
        chiEdge.alloc(E);
        CK(cudaMemsetAsync(chiEdge.p, 0, E * sizeof(double), stream));
        Dev dd = d;
        dd.chiEdge = chiEdge;
        k_errors<256><<<d.errGrid, 256, 0, stream>>>(dd, sp, 0, 0);
        src = chiEdge;
        break;
    }
    default: throw ArgError("get_values: unknown array id");
    }
    if (out && cap >= n && n > 0) {
        CK(cudaMemcpyAsync(out, src, n * sizeof(double), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        CK(cudaGetLastError());
    }
    return n;
}

}  // namespace cugo_b200

// ===========================================================================
// C ABI (include/cugo_b200.h)
// ===========================================================================
using cugo_b200::Solver;
struct cugo_b200_solver {
    Solver impl;
};

template <typename F>
static int guarded(cugo_b200_solver* s, F&& f) {
    if (!s) return CUGO_B200_EINVAL;
    try {
        f(s->impl);
        return CUGO_B200_OK;
    } catch (const cugo_b200::ArgError& e) {
        s->impl.err = e.what(); return CUGO_B200_EINVAL;
    } catch (const cugo_b200::StateError& e) {
        s->impl.err = e.what(); return CUGO_B200_ESTATE;
    } catch (const cugo_b200::UnsupportedError& e) {
        s->impl.err = e.what(); return CUGO_B200_EUNSUPPORTED;
    } catch (const cugo_b200::CudaError& e) {
        s->impl.err = e.what(); return CUGO_B200_ECUDA;
    } catch (const std::exception& e) {
        s->impl.err = e.what(); return CUGO_B200_ECUDA;
    } catch (...) {
        s->impl.err = "unknown error"; return CUGO_B200_ECUDA;
    }
}

extern "C" {

int cugo_b200_create(int device, int use_float32, cugo_b200_solver** out) {
    if (!out) return CUGO_B200_EINVAL;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return CUGO_B200_ENODEVICE; }
    if (device < 0 || device >= ndev) return CUGO_B200_EINVAL;
    cugo_b200_solver* s = nullptr;
    try {
        s = new cugo_b200_solver();
    } catch (...) {
        return CUGO_B200_ECUDA;
    }
    s->impl.device = device;
    s->impl.f32 = use_float32 != 0;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&s->impl.stream, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError();
        delete s;
        return CUGO_B200_ECUDA;
    }
    {   // keep freed buffers in the default pool instead of returning them to the driver at every synchronisation
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess && pool) {
            unsigned long long keep = ~0ULL;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        cudaGetLastError();
    }
    *out = s;
    return CUGO_B200_OK;
}

int cugo_b200_destroy(cugo_b200_solver* s) {
    if (!s) return CUGO_B200_EINVAL;
    cudaSetDevice(s->impl.device);
    cugo_b200::t_allocStream = nullptr;      // the solver's stream dies with it
    delete s;
    return CUGO_B200_OK;
}

const char* cugo_b200_last_error(const cugo_b200_solver* s) { return s ? s->impl.err.c_str() : "null solver"; }

int cugo_b200_set_vertices(cugo_b200_solver* s, int nP_all, int nP_active, const double* poses7, int nL_all,
                           int nL_active, const double* landmarks3) {
    return guarded(s, [&](Solver& S) { S.setVertices(nP_all, nP_active, poses7, nL_all, nL_active, landmarks3); });
}

int cugo_b200_clear_edge_sets(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) { S.sets.clear(); S.icpSets.clear(); S.structDirty = true; });
}

int cugo_b200_add_edge_set(cugo_b200_solver* s, int dim, int n, const int* iP, const int* iL,
                           const double* measurements, const double* omegas, int nomega, const double* cameras5,
                           int ncam, int robust_type, double robust_delta, int* set_id) {
    return guarded(s, [&](Solver& S) {
        const int id = S.addEdgeSet(dim, n, iP, iL, measurements, omegas, nomega, cameras5, ncam, robust_type, robust_delta);
        if (set_id) *set_id = id;
    });
}

int cugo_b200_add_depth_edge_set(cugo_b200_solver* s, int n, const int* iP, const int* iL, const double* measurements,
                                 const double* omegas, int nomega, const double* cameras5, int ncam, int robust_type,
                                 double robust_delta, int* set_id) {
    return guarded(s, [&](Solver& S) {
        const int id = S.addEdgeSet(cugo_b200::EDGE_DEPTH, n, iP, iL, measurements, omegas, nomega, cameras5, ncam, robust_type, robust_delta);
        if (set_id) *set_id = id;
    });
}
int cugo_b200_add_plane_edge_set(cugo_b200_solver* s, int n, const int* iP, const double* matches7, const double* omegas,
                                 int nomega, int robust_type, double robust_delta, int* set_id) {
    return guarded(s, [&](Solver& S) {
        const int id = S.addIcpEdgeSet(cugo_b200::EDGE_PLANE, n, iP, matches7, omegas, nomega, robust_type, robust_delta);
        if (set_id) *set_id = id;
    });
}
int cugo_b200_add_line_edge_set(cugo_b200_solver* s, int n, const int* iP, const double* matches10, const double* omegas,
                                int nomega, int robust_type, double robust_delta, int* set_id) {
    return guarded(s, [&](Solver& S) {
        const int id = S.addIcpEdgeSet(cugo_b200::EDGE_LINE, n, iP, matches10, omegas, nomega, robust_type, robust_delta);
        if (set_id) *set_id = id;
    });
}

int cugo_b200_set_robust_kernel(cugo_b200_solver* s, int set_id, int robust_type, double robust_delta) {
    return guarded(s, [&](Solver& S) {
        if (robust_type < 0 || robust_type > 3) throw cugo_b200::ArgError("set_robust_kernel: unknown kernel");
        if (set_id >= cugo_b200::ICP_SET_BASE && set_id < cugo_b200::ICP_SET_BASE + (int)S.icpSets.size()) {
            auto& is = S.icpSets[set_id - cugo_b200::ICP_SET_BASE];
            is.rk = robust_type; is.delta = robust_delta;
            if (!S.structDirty) { S.refreshIcpParams(); S.dropGraphs(); S.captureSlot(); }
            return;
        }
        if (set_id < 0 || set_id >= (int)S.sets.size()) throw cugo_b200::ArgError("set_robust_kernel: bad set id");
        if (robust_type < 0 || robust_type > 3) throw cugo_b200::ArgError("set_robust_kernel: unknown kernel");
        S.sets[set_id].rk = robust_type; S.sets[set_id].delta = robust_delta;
        // kernel parameters are launch arguments: refresh them and re-capture the trial graph
        if (!S.structDirty) {
            S.refreshSetParams();
            S.dropGraphs();
            S.captureSlot();
        }
    });
}

int cugo_b200_set_outliers(cugo_b200_solver* s, int set_id, const int* mask) {
    return guarded(s, [&](Solver& S) {
        if (set_id < 0 || set_id >= (int)S.sets.size()) throw cugo_b200::ArgError("set_outliers: bad set id");
        auto& hs = S.sets[set_id];
        S.syncOutliersToHost();
        if (mask) hs.out.assign(mask, mask + hs.n); else hs.out.clear();
        hs.outCount = 0;
        for (int v : hs.out) hs.outCount += v ? 1 : 0;
        S.structDirty = true;
    });
}

int cugo_b200_set_outlier_threshold(cugo_b200_solver* s, int set_id, double threshold) {
    return guarded(s, [&](Solver& S) { S.setOutlierThreshold(set_id, threshold); });
}
int cugo_b200_update_edges(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) { S.requireSingle("update_edges"); S.updateEdges(); });
}
int cugo_b200_get_outliers(cugo_b200_solver* s, int set_id, int* mask, int* n_outliers, int* n_inliers) {
    return guarded(s, [&](Solver& S) { S.getOutliers(set_id, mask, n_outliers, n_inliers); });
}

int cugo_b200_set_estimates(cugo_b200_solver* s, const double* poses7, const double* landmarks3) {
    return guarded(s, [&](Solver& S) {
        if (!poses7) throw cugo_b200::ArgError("set_estimates: null poses");
        S.setEstimates(poses7, landmarks3);
    });
}

int cugo_b200_build_structure(cugo_b200_solver* s) {
    // like the reference (src/block_solver.cpp:161-170) the heavy part runs only when the graph changed
    return guarded(s, [&](Solver& S) { S.ensureStructure(); });
}

int cugo_b200_compute_errors(cugo_b200_solver* s, double* chi2) {
    return guarded(s, [&](Solver& S) { const double v = S.computeErrors(); if (chi2) *chi2 = v; });
}
int cugo_b200_build_system(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) { S.stagedLambda = 0.0; S.buildSystem(); });
}
int cugo_b200_max_diagonal(cugo_b200_solver* s, double* max_diag) {
    return guarded(s, [&](Solver& S) { const double v = S.maxDiagonal(); if (max_diag) *max_diag = v; });
}
int cugo_b200_set_lambda(cugo_b200_solver* s, double lambda) {
    return guarded(s, [&](Solver& S) { S.stagedLambda = lambda; });
}
int cugo_b200_restore_diagonal(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) { S.stagedLambda = 0.0; });
}
int cugo_b200_solve(cugo_b200_solver* s, int* ok) {
    return guarded(s, [&](Solver& S) { const bool r = S.solve(); if (ok) *ok = r ? 1 : 0; });
}
int cugo_b200_update(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) { S.update(); });
}
int cugo_b200_compute_scale(cugo_b200_solver* s, double lambda, double* scale) {
    return guarded(s, [&](Solver& S) { const double v = S.computeScale(lambda); if (scale) *scale = v; });
}
int cugo_b200_push(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) {
        S.use();
        const int cur = cugo_b200::readCur(S);
        CK(cudaMemcpyAsync(S.poseBak.p, S.d.pose[cur], (size_t)S.nP_all * cugo_b200::POSE_STRIDE * sizeof(double),
                           cudaMemcpyDeviceToDevice, S.stream));
        CK(cudaMemcpyAsync(S.lmBak.p, S.d.lm[cur], (size_t)S.nL_all * 3 * sizeof(double), cudaMemcpyDeviceToDevice, S.stream));
        CK(cudaStreamSynchronize(S.stream));
    });
}
int cugo_b200_pop(cugo_b200_solver* s) {
    return guarded(s, [&](Solver& S) {
        S.use();
        const int cur = cugo_b200::readCur(S);
        CK(cudaMemcpyAsync(S.d.pose[cur], S.poseBak.p, (size_t)S.nP_all * cugo_b200::POSE_STRIDE * sizeof(double),
                           cudaMemcpyDeviceToDevice, S.stream));
        CK(cudaMemcpyAsync(S.d.lm[cur], S.lmBak.p, (size_t)S.nL_all * 3 * sizeof(double), cudaMemcpyDeviceToDevice, S.stream));
        CK(cudaStreamSynchronize(S.stream));
    });
}

int cugo_b200_optimize(cugo_b200_solver* s, int niterations, int capacity, double* chi2, double* lambda, int* trials,
                       int* nstats) {
    return guarded(s, [&](Solver& S) {
        const int n = S.optimize(niterations, capacity, chi2, lambda, trials);
        if (nstats) *nstats = n;
    });
}

int cugo_b200_get_estimates(cugo_b200_solver* s, double* poses7, double* landmarks3) {
    return guarded(s, [&](Solver& S) { S.getEstimates(poses7, landmarks3); });
}

int cugo_b200_get_structure(cugo_b200_solver* s, int which, int* out, long capacity, long* n) {
    return guarded(s, [&](Solver& S) { const long r = S.getStructure(which, out, capacity); if (n) *n = r; });
}
int cugo_b200_get_values(cugo_b200_solver* s, int which, double* out, long capacity, long* n) {
    return guarded(s, [&](Solver& S) { const long r = S.getValues(which, out, capacity); if (n) *n = r; });
}
int cugo_b200_get_sizes(cugo_b200_solver* s, long sizes[10]) {
    return guarded(s, [&](Solver& S) {
        if (!sizes) throw cugo_b200::ArgError("get_sizes: null output");
        S.ensureStructure();
        sizes[0] = S.nP; sizes[1] = S.nL; sizes[2] = S.E; sizes[3] = S.d.nHpl; sizes[4] = S.d.nblk;
        sizes[5] = S.nmulRef; sizes[6] = S.nnzSymm; sizes[7] = S.counts.nsn; sizes[8] = S.counts.nlev;
        sizes[9] = S.kernelsPerSlot;
    });
}
int cugo_b200_get_last_timing(cugo_b200_solver* s, double* optimize_ms, long* kernels_launched) {
    return guarded(s, [&](Solver& S) {
        if (optimize_ms) *optimize_ms = S.lastOptimizeMs;
        if (kernels_launched) *kernels_launched = S.lastKernels;
    });
}
int cugo_b200_profile_trial(cugo_b200_solver* s, int capacity, char* names, int name_len, double* ms, int* n) {
    return guarded(s, [&](Solver& S) { const int r = S.profileSlot(capacity, names, name_len, ms); if (n) *n = r; });
}
int cugo_b200_update_measurements(cugo_b200_solver* s, int set_id, const double* measurements, const double* omegas,
                                  int nomega) {
    return guarded(s, [&](Solver& S) { S.updateMeasurements(set_id, measurements, omegas, nomega); });
}
int cugo_b200_comm_unique_id(char id[128]) {
    if (!id) return CUGO_B200_EINVAL;
    std::string err;
    if (!cugo_b200::g_nccl.load(err)) return CUGO_B200_ECUDA;
    ncclUniqueId uid;
    if (cugo_b200::g_nccl.GetUniqueId(&uid) != ncclSuccess) return CUGO_B200_ECUDA;
    memcpy(id, &uid, 128);
    return CUGO_B200_OK;
}
int cugo_b200_comm_init(cugo_b200_solver* s, int nranks, int rank, const char id[128]) {
    return guarded(s, [&](Solver& S) { S.commInit(nranks, rank, id); });
}
int cugo_b200_get_factor_stats(cugo_b200_solver* s, double stats[4]) {
    return guarded(s, [&](Solver& S) {
        if (!stats) throw cugo_b200::ArgError("get_factor_stats: null output");
        S.ensureStructure();
        for (int i = 0; i < 4; i++) stats[i] = S.factStats[i];
    });
}
int cugo_b200_debug_counters(long long* out, int n) {
#ifdef CUGO_B200_DIAG
    if (out && n == 8192)   // per-row cycles of the Schur kernel
        return cudaMemcpyFromSymbol(out, cugo_b200::g_schurDbg, sizeof(long long) * n) == cudaSuccess ? CUGO_B200_OK : CUGO_B200_ECUDA;
    if (!out || n < 0 || n > 64 * 24) return CUGO_B200_EINVAL;
    if (n == 512)
        return cudaMemcpyFromSymbol(out, cugo_b200::g_cholTime, sizeof(long long) * n) == cudaSuccess ? CUGO_B200_OK : CUGO_B200_ECUDA;
    if (n == 176)
        return cudaMemcpyFromSymbol(out, cugo_b200::g_cholTrace, sizeof(long long) * n) == cudaSuccess ? CUGO_B200_OK : CUGO_B200_ECUDA;
    return cudaMemcpyFromSymbol(out, cugo_b200::g_cholDbg, sizeof(long long) * n) == cudaSuccess ? CUGO_B200_OK : CUGO_B200_ECUDA;
#else
    (void)out; (void)n;
    return CUGO_B200_EUNSUPPORTED;      // the phase clocks are compiled in by `make DIAG=1` only
#endif
}
int cugo_b200_set_option(cugo_b200_solver* s, const char* name, int value) {
    return guarded(s, [&](Solver& S) {
        if (!name) throw cugo_b200::ArgError("set_option: null name");
        const std::string n(name);
        if (n == "dist_factor") { S.distMode = value; S.structDirty = true; return; }
        if (n == "peer_reduce") { S.peerMode = value; S.structDirty = true; return; }
        if (n == "upd_small") { S.updSmall = value; S.structDirty = true; return; }
        if (n == "panel_rb") {
            if (value < 0 || value > cugo_b200::CHOL_RB) throw cugo_b200::ArgError("set_option: panel_rb must be in [0, 16] (0: chosen per tree level)");
            S.panelRB = value; S.structDirty = true; return;
        }
        bool* f = n == "pdl" ? &S.pdl : n == "fuse_build" ? &S.fuseBuild : n == "panel_dc" ? &S.panelDC : n == "zinv" ? &S.zinvOn
                : n == "schur_mma" ? &S.schurMma : n == "bwd_split" ? &S.bwdSplit : n == "parallel_symbolic" ? &S.symParallel : nullptr;
        if (!f) throw cugo_b200::ArgError("set_option: unknown option '" + n + "' (pdl, fuse_build, panel_dc, zinv, bwd_split, parallel_symbolic, dist_factor, panel_rb, upd_small, schur_mma, peer_reduce)");
        *f = value != 0;
        S.structDirty = true;           // the options shape the structure and the trial graph
    });
}
void* cugo_b200_stream(cugo_b200_solver* s) { return s ? (void*)s->impl.stream : nullptr; }

}  // extern "C"

// ===========================================================================
// Allocation-free chi2 of one edge set (include/cugo_b200.h: cugo_b200_chi2_noalloc).
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops).  The per-CTA partial sums that k_errors keeps
// in a cudaMalloc'ed buffer (Dev::chiPartial) live here in a file-scope __device__ array that every GPU
// loading the library holds until exit.  In host code `g_naScratch` is only a stand-in for the symbol: its
// device address is taken once per device with cudaGetSymbolAddress (that device current) and handed to the
// kernel as an ordinary pointer.
// ===========================================================================
namespace cugo_b200 {

constexpr int NA_TPB = 256;
constexpr int NA_EDGES_PER_CTA = 4 * NA_TPB;                 // 1024 edges per CTA, one partial each
constexpr int NA_MAX_CTAS = CUGO_B200_NOALLOC_MAX_EDGES / NA_EDGES_PER_CTA;   // 8192
constexpr int NA_HEAD = 32;                                   // doubles in front of the partials: [0] = arrival ticket
static_assert(CUGO_B200_NOALLOC_MAX_EDGES % NA_EDGES_PER_CTA == 0, "maximum is a whole number of CTAs");
static __device__ __align__(256) double g_naScratch[NA_HEAD + NA_MAX_CTAS];   // 64.25 KB per GPU

struct NaArgs {
    int dim, n, nomega, rkType;
    double rkDelta;
    double cam[5];
    const int* iP; const int* iL;
    const double* meas; const double* omega;
    const double* poses7; const double* lms;
    double* out;
};

// same arithmetic per edge as k_errors (world_to_camera, residual, rk_rho), caller-layout inputs: poses 7 doubles
// AoS, measurements interleaved [n x dim].  CTA partials -> scratch; the last CTA to arrive adds them in CTA order
// (bit-reproducible), writes the total and leaves the ticket at zero for the next call.
__global__ void __launch_bounds__(NA_TPB) k_errors_noalloc(NaArgs a, double* scratch) {
    __shared__ double sh[NA_TPB / 32];
    __shared__ unsigned last;
    double sum = 0;
    const int base = blockIdx.x * NA_EDGES_PER_CTA;
#pragma unroll
    for (int j = 0; j < NA_EDGES_PER_CTA / NA_TPB; j++) {
        const int k = base + j * NA_TPB + threadIdx.x;
        if (k >= a.n) break;
        const int iP = __ldg(a.iP + k), iL = __ldg(a.iL + k);
        const double* p = a.poses7 + 7 * (size_t)iP;
        const double q[4] = {__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3)};
        const double t[3] = {__ldg(p + 4), __ldg(p + 5), __ldg(p + 6)};
        const double X[3] = {__ldg(a.lms + 3 * (size_t)iL), __ldg(a.lms + 3 * (size_t)iL + 1), __ldg(a.lms + 3 * (size_t)iL + 2)};
        const double* m = a.meas + (size_t)k * a.dim;
        const double mr = (a.dim == 3) ? __ldg(m + 2) : 0.0;
        double Xc[3], e[3];
        world_to_camera(q, t, X, Xc);
        residual(a.dim, Xc, a.cam, __ldg(m), __ldg(m + 1), mr, e);
        const double w = __ldg(a.omega + (a.nomega > 1 ? k : 0));
        sum += rk_rho(a.rkType, a.rkDelta, w * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]));
    }
    const double tot = block_sum<NA_TPB>(sum, sh);
    double* part = scratch + NA_HEAD;
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    if (threadIdx.x == 0) {
        part[blockIdx.x] = tot;
        __threadfence();
        last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double s = 0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += NA_TPB) s += __ldcg(part + i);   // fixed assignment, fixed order
    const double all = block_sum<NA_TPB>(s, sh);
    if (threadIdx.x == 0) { *a.out = all; *ticket = 0u; }
}

static double* g_naPtr[64];          // device address of g_naScratch per device ordinal (set once)
static std::mutex g_naMutex;

static int na_fail(char* errbuf, int len, int code, const std::string& msg) {
    if (errbuf && len > 0) { strncpy(errbuf, msg.c_str(), (size_t)len - 1); errbuf[len - 1] = 0; }
    return code;
}

}  // namespace cugo_b200

extern "C" int cugo_b200_chi2_noalloc(int device, void* stream, int dim, int n, const int* d_iP, const int* d_iL,
                                      const double* d_measurements, const double* d_omegas, int nomega,
                                      const double camera5[5], int robust_type, double robust_delta,
                                      const double* d_poses7, const double* d_landmarks3, double* d_chi2,
                                      char* errbuf, int errbuf_len) {
    using namespace cugo_b200;
    if ((dim != 2 && dim != 3) || n < 0 || (nomega != 1 && nomega != n && n > 0) || robust_type < 0 || robust_type > 3 ||
        !camera5 || !d_chi2 || (n > 0 && (!d_iP || !d_iL || !d_measurements || !d_omegas || !d_poses7 || !d_landmarks3)))
        return na_fail(errbuf, errbuf_len, CUGO_B200_EINVAL, "chi2_noalloc: bad argument (dim 2|3, nomega 1|n, non-null device pointers)");
    if (n > CUGO_B200_NOALLOC_MAX_EDGES)
        return na_fail(errbuf, errbuf_len, CUGO_B200_EINVAL,
                       "chi2_noalloc: " + std::to_string(n) + " edges exceed CUGO_B200_NOALLOC_MAX_EDGES = " +
                           std::to_string(CUGO_B200_NOALLOC_MAX_EDGES) + " (size of the static scratch array); split the set");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return na_fail(errbuf, errbuf_len, CUGO_B200_ENODEVICE, "chi2_noalloc: no CUDA device (there is no CPU path)"); }
    if (device < 0 || device >= ndev || device >= 64) return na_fail(errbuf, errbuf_len, CUGO_B200_EINVAL, "chi2_noalloc: bad device ordinal");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, CUGO_B200_ECUDA, std::string("chi2_noalloc: cudaSetDevice: ") + cudaGetErrorString(e));
    double* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_naMutex);
        if (!g_naPtr[device]) {           // once per device, with that device current
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_naScratch);
            if (e != cudaSuccess || !p)
                return na_fail(errbuf, errbuf_len, CUGO_B200_ECUDA, std::string("chi2_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            g_naPtr[device] = static_cast<double*>(p);
        }
        scratch = g_naPtr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the array keeps the last call's contents: the ticket must start at zero (an aborted call may have left it set)
    e = cudaMemsetAsync(scratch, 0, NA_HEAD * sizeof(double), st);
    if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, CUGO_B200_ECUDA, std::string("chi2_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(e));
    if (n == 0) {                         // empty set: chi2 = 0
        e = cudaMemsetAsync(d_chi2, 0, sizeof(double), st);
        return e == cudaSuccess ? CUGO_B200_OK : na_fail(errbuf, errbuf_len, CUGO_B200_ECUDA, std::string("chi2_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(e));
    }
    NaArgs a{};
    a.dim = dim; a.n = n; a.nomega = nomega; a.rkType = robust_type; a.rkDelta = robust_delta;
    for (int i = 0; i < 5; i++) a.cam[i] = camera5[i];
    a.iP = d_iP; a.iL = d_iL; a.meas = d_measurements; a.omega = d_omegas; a.poses7 = d_poses7; a.lms = d_landmarks3;
    a.out = d_chi2;
    const int grid = (n + NA_EDGES_PER_CTA - 1) / NA_EDGES_PER_CTA;
    k_errors_noalloc<<<grid, NA_TPB, 0, st>>>(a, scratch);
    e = cudaGetLastError();
    if (e != cudaSuccess) return na_fail(errbuf, errbuf_len, CUGO_B200_ECUDA, std::string("chi2_noalloc: launch: ") + cudaGetErrorString(e));
    return CUGO_B200_OK;
}
