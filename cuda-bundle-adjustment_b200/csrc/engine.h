// engine.h -- device-side data layout of the B200 bundle-adjustment engine.
// All arrays live in HBM for the lifetime of a structure; kernels take one
// `Dev` struct of raw pointers by value.  See DESIGN.md section 3 for the layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cugo_b200 {

constexpr int MAX_SETS = 8;
constexpr int MAX_STATS = 256;     // outer LM iterations recorded per optimize()
constexpr int LM_CHUNK = 128;      // landmark-major kernels: edges per CTA (nominal)
constexpr int POSE_SLICE = 512;    // edges per CTA of the pose-major pass
constexpr int POSE_STRIDE = 8;     // poses are padded 7 -> 8 doubles (16 B vector loads)

// per edge set parameters (reference: one EdgeSet object each, include/ba_types.h)
struct SetParams {
    int dim[MAX_SETS];        // edge kind: 2 mono / 3 stereo / 4 depth (device_math.cuh)
    int rkType[MAX_SETS];
    double rkDelta[MAX_SETS];
    double cam[MAX_SETS][5];  // shared camera of the set (used when the set has no per-edge cameras)
};

// Device-resident Levenberg-Marquardt state (reference keeps these on the host:
// src/cuda_graph_optimisation.cpp:48-154)
struct LMState {
    double lambda, nu, F, Fhat, rho, scale, maxDiag;
    int iter;        // outer iteration index
    int q;           // trial index inside the outer iteration
    int cur;         // which estimate buffer holds the accepted estimates
    int done;        // optimisation finished
    int needBuild;   // linear system must be rebuilt from est[cur]
    int fail;        // factorisation hit a non-positive pivot in this trial
    int niter;       // requested outer iterations
    int nstats;
    int slots;       // trial slots executed (diagnostics)
    int lastEval;    // estimate buffer the last k_errors evaluated (the reference's chiValues belong to it)
    double statChi2[MAX_STATS];
    double statLambda[MAX_STATS];
    int statTrials[MAX_STATS];
};

struct Dev {
    // sizes
    int nP_all, nP, nL_all, nL, E, nHpl, nblk, nmul, nChunks;
    // estimates, double buffered: pose [nP_all*8], landmark [nL_all*3]
    double* pose[2];
    double* lm[2];
    // edges in landmark-major order (sorted by (iL, iP)): SoA
    const int* s_iP;
    const int* s_iL;
    const double2* s_uv;
    const double* s_r;          // stereo third measurement (0 for mono)
    const double* s_w;          // information (omega) per edge
    const uint8_t* s_set;       // edge set id
    const int* s_cam;           // per-edge camera index or nullptr
    const uint8_t* s_out;       // outlier mask or nullptr
    const int* s_gid;           // global edge id (set-major input order)
    const int* s_hplPre;        // [E+1] exclusive prefix of "has Hpl block" flags
    const int* lmEdgePtr;       // [nL_all+1]
    const int* chunkLm;         // [nChunks+1] landmark ranges of the landmark-major CTAs
    // the same edges in pose-major order (sorted by iP, then landmark)
    const int* posePtr;         // [nP_all+1]
    const int* p_iL;
    const double2* p_uv;
    const double* p_r;
    const double* p_w;
    const uint8_t* p_set;
    const int* p_cam;
    const uint8_t* p_out;
    const double* cams;         // [ncams][5] per-edge camera table (or nullptr)
    // pose-major pass in slices of POSE_SLICE edges (a loop-closure pose, or any pose of a landmark shard on several
    // GPUs, has thousands of edges): work unit = (pose, slice); partial sums meet through a ticket per pose
    int nPoseUnits;
    const int* poseSlicePtr;    // [nP+1]
    const int2* poseUnits;      // [nPoseUnits]
    double* posePart;           // [nPoseUnits][27]
    int* poseTicket;            // [nP], zero between launches
    // Hpl CSC (reference: gpu::buildHplStructure)
    const int* colPtr;          // [nL+1]
    const int* rowInd;          // [nHpl]
    const int* hplLm;           // [nHpl] landmark of each block
    // Hsc upper BSR + product lists grouped by destination block
    const int* bRowPtr;         // [nP+1]
    const int* bColInd;         // [nblk]
    const int* bRowOf;          // [nblk]
    const int* prodPtr;         // [nblk+1]
    const int2* prodIJ;         // [nmul] (Hpl slot of row pose, Hpl slot of col pose); x<0 = none
    // pose-major view of the Hpl blocks (k_schur_rows)
    const int* poseSlotPtr;     // [nP+1]
    const int* poseSlots;       // [nHpl] Hpl slots of every free pose in landmark order
    const int* poseLm;          // [nHpl] landmark of poseSlots[k]
    const int* prodRank;        // [nmul] rank of the product's row block among its pose's slots, -1 = none
    const int2* rowUnits;       // [nRowUnits] (block row, unit of the row): work list of k_schur_rows, heavy rows first
    // linear system
    double* Hpp;                // [36 nP]
    double* Hll;                // [9 nL]
    double* Hpl;                // [18 nHpl] (FP64 mode)
    float* HplF;                // [18 nHpl] single-precision storage of the same blocks (USE_FLOAT32 mode), else null
    double* b;                  // [6 nP + 3 nL]
    double* x;                  // [6 nP + 3 nL]
    double* invHll;             // [9 nL]
    double* bsc;                // [6 nP]
    double* Hsc;                // [36 nblk]
    // what the Schur kernels write: the same arrays on one GPU; with several ranks the rank's PARTIAL [Hsc | bsc] in the
    // global block layout (blocks the rank does not own stay zero), summed into Hsc / bsc by the all-reduce
    double* HscPart;
    double* bscPart;
    // reductions
    double* chiPartial;         // [nChi]: one per CTA of k_errors, then one per CTA of k_icp_errors
    double* scalePartial;       // [nChunks]
    double* maxDiagL;           // [nChunks]
    double* maxDiagP;           // [nP]
    double* chiEdge;            // [E] per-edge rho(chi2) in global edge order, or nullptr
    int errGrid;                // CTAs of k_errors
    int nChi;                   // entries of chiPartial the LM kernels sum (errGrid + the ICP kernel's CTAs)
    LMState* st;
    // multi-GPU (one process per GPU, landmarks sharded).  Every rank holds all poses
    // and the *global* Hsc pattern; bRowPtr/bColInd/bRowOf/prodPtr/nblk above describe the LOCAL blocks.
    int nranks, rank;
    int nblkG;                  // blocks of the global pattern (== nblk on one GPU)
    const int* hscSlot;         // [nblk] local block -> global block id, nullptr = identity
    double* redA;               // [6 nP + nranks] diag(Hpp) partials + one-hot landmark max-diagonal (all-reduced)
    double* redC;               // [2] chi2, gain denominator (all-reduced)
};

// Pose-only ("unary") ICP edges: the reference's PlaneEdgeSet / LineEdgeSet (include/icp_types.h:19-107).  They touch
// Hpp and bp only, so they live beside the BA edge arrays: sorted by pose, payload column-major [10][nI] in the field
// order of the reference's match structs (src/measurements.h:13-101) --
//   plane: normal[3], originDistance, pointP[3];   line: a[3], b[3], length, pointP[3].
constexpr int MAX_ICP_SETS = 8;
constexpr int ICP_COLS = 10;
constexpr int ICP_SLICE = 2048;    // edges per CTA of k_icp_build
struct IcpDev {
    int nI;                     // edges of all ICP sets
    int nUnits, grid;           // CTAs of k_icp_build / k_icp_errors
    const int* iP;              // [nI] pose of the edge (sorted)
    const uint8_t* set;         // [nI] ICP set (index into kind / rk below)
    const double* w;            // [nI] information
    const double* pay;          // [ICP_COLS][nI]
    const int* posePtr;         // [nP_all+1]
    const int* slicePtr;        // [nP+1]
    const int2* units;          // [nUnits] (pose, slice)
    double* part;               // [nUnits][27] partial JtJ (21) and Jte (6) of every unit
    int kind[MAX_ICP_SETS];     // EDGE_PLANE / EDGE_LINE
    int rkType[MAX_ICP_SETS];
    double rkDelta[MAX_ICP_SETS];
};

// multifrontal factorisation data (see symbolic.h)
struct Chol {
    int P, nsn, nlev;
    const int* perm;       // new -> old
    const int* snStart;
    const int* snRowPtr;
    const int* snRows;
    const int* levelPtr;
    const int* levelSn;
    const long long* Lptr;
    const long long* Uptr;
    const int* childPtr;
    const int* childIdx;
    const int* invPtr;
    const int* inv;
    const int* snBlkPtr;
    const int* asmBlk;
    const int* asmRow;
    const int* asmCol;
    const int* trsmPtr;    // [nsn+1] prefix of panel-slab tasks in levelSn order
    const int* updPtr;     // [nsn+1] prefix of update-tile tasks in levelSn order
    double* L;
    double* U;
    double* Zinv;          // [6 P x 6 MAXW] inverse of every supernode's diagonal block L11 (k_chol_zinv), or unused
    double* Winv;          // [21 P] inverse of each 6x6 diagonal Cholesky block (packed lower), by permuted column block
    double* xperm;         // [6P] solution in permuted order
    double* bwdPart;       // [nsn x 8 x 6 MAXW] partial L21^T x_R vectors of the CTAs sharing a supernode (k_chol_bwd3)
    int* bwdCnt;           // [nsn] arrival counters of those CTAs (zero between trials)
    int* xFlag;            // [nsn + 1] backward solve: epoch at which x of supernode t was stored; [nsn] = the current epoch
    const int* colToSn;    // [P] supernode of every (permuted) block column
    const int* rbLev;      // [nlev] block rows of L21 per panel CTA on every tree level (<= CHOL_RB; Solver::planPanelSlabs)
    int dbg;               // 1: kernels record phase timestamps (diagnostics)
};

// Programmatic dependent launch (sm_90+).  The kernels of a trial are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization: a kernel may become resident while its predecessor still
// runs and blocks in pdl_wait() until the predecessor has COMPLETED (memory flushed).  Rules kept by every
// kernel of the trial: pdl_wait() before the first read of mutable global data and before any global write, so
// kernels still complete in launch order; what runs before it touches only shared memory and structure arrays.
// Both are no-ops for a launch without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// storage type of the Hpl blocks: double, or float in the USE_FLOAT32 mode (arithmetic stays FP64)
template <typename HT> struct HplOf;
template <> struct HplOf<double> { static __device__ __forceinline__ double* get(const Dev& d) { return d.Hpl; } };
template <> struct HplOf<float> { static __device__ __forceinline__ float* get(const Dev& d) { return d.HplF; } };

}  // namespace cugo_b200
