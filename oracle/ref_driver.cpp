// ref_driver.cpp -- TEST INFRASTRUCTURE.  Drives the UNMODIFIED reference library
// (/root/reference, compiled by oracle/Makefile into oracle/_ref/) through its own
// public API (CudaGraphOptimisationImpl / vertex & edge sets, as
// samples/sample_ba_from_file/main.cpp:78-201 does) on a graph written by
// cugo_b200/graphio.py, and dumps what the parity tests pin the oracle and the
// CUDA engine against: chi2 per LM iteration, final estimates, the structure
// arrays and one stage-by-stage linearisation.  Never shipped, never linked by
// the product.  BlockSolver internals are reached by compiling THIS file with
// -Dprivate=public (the reference sources themselves are compiled untouched).
#include <ba_types.h>
#include <cuda_graph_optimisation.h>
#include <optimisable_graph.h>
#include <robust_kernel.h>
#include <block_solver.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct GraphFile {
    int P, L, Em, Es;
    double cam[5];
    std::vector<double> pose, lm, mmeas, minfo, smeas, sinfo;
    std::vector<int> pfixed, lfixed, mp, ml, sp, sl;
};

template <typename T>
static void rd(FILE* f, std::vector<T>& v, size_t n) {
    v.resize(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
}

static GraphFile readGraph(const char* path) {
    GraphFile g;
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    int hdr[6];
    if (fread(hdr, sizeof(int), 6, f) != 6 || hdr[0] != 0x31474142) { fprintf(stderr, "bad header\n"); exit(2); }
    g.P = hdr[1]; g.L = hdr[2]; g.Em = hdr[3]; g.Es = hdr[4];
    if (fread(g.cam, sizeof(double), 5, f) != 5) exit(2);
    rd(f, g.pose, (size_t)g.P * 7); rd(f, g.lm, (size_t)g.L * 3);
    rd(f, g.mmeas, (size_t)g.Em * 2); rd(f, g.minfo, g.Em);
    rd(f, g.smeas, (size_t)g.Es * 3); rd(f, g.sinfo, g.Es);
    rd(f, g.pfixed, g.P); rd(f, g.lfixed, g.L);
    rd(f, g.mp, g.Em); rd(f, g.ml, g.Em); rd(f, g.sp, g.Es); rd(f, g.sl, g.Es);
    fclose(f);
    return g;
}

template <typename T>
static void dumpHost(const std::string& prefix, const char* name, const T* p, size_t n) {
    std::string fn = prefix + "." + name + ".bin";
    FILE* f = fopen(fn.c_str(), "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", fn.c_str()); exit(2); }
    if (n) fwrite(p, sizeof(T), n, f);
    fclose(f);
}
template <typename T>
static void dumpDev(const std::string& prefix, const char* name, const T* d, size_t n) {
    std::vector<T> h(n);
    if (n) cudaMemcpy(h.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost);
    dumpHost(prefix, name, h.data(), n);
}

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: ref_driver graph.bag out_prefix niter [rk_type(0 none,1 cauchy,2 tukey) delta] [reps] [outlier_threshold niter2]\n");
        return 1;
    }
    const GraphFile g = readGraph(argv[1]);
    const std::string prefix = argv[2];
    const int niter = atoi(argv[3]);
    const int rkType = argc > 4 ? atoi(argv[4]) : 0;
    const double rkDelta = argc > 5 ? atof(argv[5]) : 1.0;
    const int reps = argc > 6 ? atoi(argv[6]) : 1;
    // outlier scenario (EdgeSet::setOutlierThreshold): initialize(); optimize(niter); optimize(niter2)
    const double outThr = argc > 7 ? atof(argv[7]) : 0.0;
    const int niter2 = argc > 8 ? atoi(argv[8]) : 0;

    cugo::GraphOptimisationOptions options;
    options.perEdgeInformation = true;
    options.perEdgeCamera = false;
    auto optimizer = std::make_unique<cugo::CudaGraphOptimisationImpl>(options);

    cugo::PoseVertexSet poseSet(false);
    cugo::LandmarkVertexSet lmSet(true);
    cugo::MonoEdgeSet monoSet;
    cugo::StereoEdgeSet stereoSet;
    std::vector<cugo::PoseVertex*> pv(g.P);
    std::vector<cugo::LandmarkVertex*> lv(g.L);
    for (int i = 0; i < g.P; i++) {
        cugo::QuatD q; cugo::Vec3d t;
        for (int k = 0; k < 4; k++) q[k] = g.pose[7 * (size_t)i + k];
        for (int k = 0; k < 3; k++) t[k] = g.pose[7 * (size_t)i + 4 + k];
        pv[i] = new cugo::PoseVertex(i, cugo::Se3D(q, t), g.pfixed[i] != 0);
        poseSet.addVertex(pv[i]);
    }
    for (int i = 0; i < g.L; i++) {
        cugo::Vec3d X;
        for (int k = 0; k < 3; k++) X[k] = g.lm[3 * (size_t)i + k];
        lv[i] = new cugo::LandmarkVertex(i, X, g.lfixed[i] != 0);
        lmSet.addVertex(lv[i]);
    }
    optimizer->addVertexSet(&poseSet);
    optimizer->addVertexSet(&lmSet);
    cugo::Camera camera(g.cam[0], g.cam[1], g.cam[2], g.cam[3], g.cam[4]);
    for (int i = 0; i < g.Em; i++) {
        auto* e = new cugo::MonoEdge();
        e->setVertex(pv[g.mp[i]], 0);
        e->setVertex(lv[g.ml[i]], 1);
        cugo::Vec2d m; m[0] = g.mmeas[2 * (size_t)i]; m[1] = g.mmeas[2 * (size_t)i + 1];
        e->setMeasurement(m);
        e->setInformation(g.minfo[i]);
        e->setCamera(camera);
        monoSet.addEdge(e);
    }
    for (int i = 0; i < g.Es; i++) {
        auto* e = new cugo::StereoEdge();
        e->setVertex(pv[g.sp[i]], 0);
        e->setVertex(lv[g.sl[i]], 1);
        cugo::Vec3d m;
        for (int k = 0; k < 3; k++) m[k] = g.smeas[3 * (size_t)i + k];
        e->setMeasurement(m);
        e->setInformation(g.sinfo[i]);
        e->setCamera(camera);
        stereoSet.addEdge(e);
    }
    monoSet.setCamera(camera);
    stereoSet.setCamera(camera);
    optimizer->addEdgeSet<cugo::MonoEdgeSet>(&monoSet);
    optimizer->addEdgeSet<cugo::StereoEdgeSet>(&stereoSet);
    // one process-global kernel in the reference (cuda_block_solver.cu:1030-1031)
    stereoSet.setRobustKernel(static_cast<cugo::RobustKernelType>(rkType), rkDelta);

    auto resetEstimates = [&]() {
        for (int i = 0; i < g.P; i++) {
            cugo::QuatD q; cugo::Vec3d t;
            for (int k = 0; k < 4; k++) q[k] = g.pose[7 * (size_t)i + k];
            for (int k = 0; k < 3; k++) t[k] = g.pose[7 * (size_t)i + 4 + k];
            pv[i]->setEstimate(cugo::Se3D(q, t));
        }
        for (int i = 0; i < g.L; i++) {
            cugo::Vec3d X;
            for (int k = 0; k < 3; k++) X[k] = g.lm[3 * (size_t)i + k];
            lv[i]->setEstimate(X);
        }
    };

    // ---- timed runs: initialize() + optimize(niter), as the sample times them
    // (samples/sample_ba_from_file/main.cpp:185-190).  rep 0 is cold (structure,
    // METIS ordering and symbolic analysis included); later reps reuse them.
    std::string json = "{\n";
    json += "  \"P\": " + std::to_string(g.P) + ", \"L\": " + std::to_string(g.L) +
            ", \"Em\": " + std::to_string(g.Em) + ", \"Es\": " + std::to_string(g.Es) +
            ", \"niter\": " + std::to_string(niter) + ", \"rk_type\": " + std::to_string(rkType) +
            ", \"rk_delta\": " + std::to_string(rkDelta) + ",\n";
    std::vector<double> wallInit, wallOpt;
    std::vector<double> chi2;
    auto listd0 = [](const std::vector<double>& v) {
        std::string s = "[";
        char buf[64];
        for (size_t i = 0; i < v.size(); i++) { snprintf(buf, sizeof(buf), "%s%.17g", i ? ", " : "", v[i]); s += buf; }
        return s + "]";
    };
    if (outThr > 0) {
        // optimize -> updateEdges (src/cuda_graph_optimisation.cpp:150-151) -> optimize again WITHOUT initialize():
        // the device mask persists, flagged edges are inactive, the structure is rebuilt without their Hpl blocks
        monoSet.setOutlierThreshold(outThr);
        stereoSet.setOutlierThreshold(outThr);
        optimizer->initialize();
        // the packed edge order of this run (unordered_set iteration order) is dumped to key the masks by (pose, landmark)
        optimizer->setVerbose(true);
        optimizer->optimize(niter);
        cudaDeviceSynchronize();
        dumpDev(prefix, "mono_edge2pl", (const int*)monoSet.d_edge2PL.data(), monoSet.nActiveEdges() * 2);
        dumpDev(prefix, "stereo_edge2pl", (const int*)stereoSet.d_edge2PL.data(), stereoSet.nActiveEdges() * 2);
        std::vector<double> chiA, chiB;
        for (const auto& st : optimizer->batchStatistics().get()) chiA.push_back(st.chi2);
        const int cntA[2] = {(int)monoSet.getOutlierCount(), (int)stereoSet.getOutlierCount()};
        dumpDev(prefix, "mono_outliers_a", monoSet.d_outliers.data(), monoSet.nActiveEdges());
        dumpDev(prefix, "stereo_outliers_a", stereoSet.d_outliers.data(), stereoSet.nActiveEdges());
        const size_t nA = optimizer->batchStatistics().get().size();
        optimizer->optimize(niter2);
        cudaDeviceSynchronize();
        const auto& all = optimizer->batchStatistics().get();     // optimize() appends to the same statistics object
        for (size_t i = (all.size() > nA ? nA : 0); i < all.size(); i++) chiB.push_back(all[i].chi2);
        const int cntB[2] = {(int)monoSet.getOutlierCount(), (int)stereoSet.getOutlierCount()};
        dumpDev(prefix, "mono_outliers_b", monoSet.d_outliers.data(), monoSet.nActiveEdges());
        dumpDev(prefix, "stereo_outliers_b", stereoSet.d_outliers.data(), stereoSet.nActiveEdges());
        std::vector<double> est((size_t)g.P * 7 + (size_t)g.L * 3);
        for (int i = 0; i < g.P; i++) {
            const cugo::Se3D& e = pv[i]->getEstimate();
            for (int k = 0; k < 4; k++) est[7 * (size_t)i + k] = e.r[k];
            for (int k = 0; k < 3; k++) est[7 * (size_t)i + 4 + k] = e.t[k];
        }
        for (int i = 0; i < g.L; i++) {
            const cugo::Vec3d& e = lv[i]->getEstimate();
            for (int k = 0; k < 3; k++) est[(size_t)g.P * 7 + 3 * (size_t)i + k] = e[k];
        }
        dumpHost(prefix, "final_estimates", est.data(), est.size());
        char buf[512];
        snprintf(buf, sizeof(buf), "  \"outlier_threshold\": %.17g, \"niter2\": %d, \"outliers_a\": [%d, %d], \"outliers_b\": [%d, %d],\n",
                 outThr, niter2, cntA[0], cntA[1], cntB[0], cntB[1]);
        json += buf;
        json += "  \"chi2_a\": " + listd0(chiA) + ",\n  \"chi2_b\": " + listd0(chiB) + "\n}\n";
        FILE* f = fopen((prefix + ".json").c_str(), "w");
        fputs(json.c_str(), f);
        fclose(f);
        printf("%s", json.c_str());
        return 0;
    }
    for (int rep = 0; rep < reps; rep++) {
        resetEstimates();
        optimizer->setVerbose(rep == 0);     // prints "levenberg iterations = q" per outer iteration (trial counts)
        cudaDeviceSynchronize();
        const auto t0 = std::chrono::steady_clock::now();
        optimizer->initialize();
        cudaDeviceSynchronize();
        const auto t1 = std::chrono::steady_clock::now();
        optimizer->optimize(niter);
        cudaDeviceSynchronize();
        const auto t2 = std::chrono::steady_clock::now();
        wallInit.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        wallOpt.push_back(std::chrono::duration<double, std::milli>(t2 - t1).count());
        if (rep == 0) {
            for (const auto& s : optimizer->batchStatistics().get()) chi2.push_back(s.chi2);
            std::vector<double> est((size_t)g.P * 7 + (size_t)g.L * 3);
            for (int i = 0; i < g.P; i++) {
                const cugo::Se3D& e = pv[i]->getEstimate();
                for (int k = 0; k < 4; k++) est[7 * (size_t)i + k] = e.r[k];
                for (int k = 0; k < 3; k++) est[7 * (size_t)i + 4 + k] = e.t[k];
            }
            for (int i = 0; i < g.L; i++) {
                const cugo::Vec3d& e = lv[i]->getEstimate();
                for (int k = 0; k < 3; k++) est[(size_t)g.P * 7 + 3 * (size_t)i + k] = e[k];
            }
            dumpHost(prefix, "final_estimates", est.data(), est.size());
        }
    }
    auto listd = [](const std::vector<double>& v) {
        std::string s = "[";
        char buf[64];
        for (size_t i = 0; i < v.size(); i++) {
            snprintf(buf, sizeof(buf), "%s%.17g", i ? ", " : "", v[i]);
            s += buf;
        }
        return s + "]";
    };
    json += "  \"chi2\": " + listd(chi2) + ",\n";
    json += "  \"wall_ms_initialize\": " + listd(wallInit) + ",\n";
    json += "  \"wall_ms_optimize\": " + listd(wallOpt) + ",\n";

    // ---- one linearisation, stage by stage (BlockSolver methods in the order of
    // src/cuda_graph_optimisation.cpp:57-87), from the ORIGINAL estimates
    resetEstimates();
    optimizer->initialize();
    cugo::BlockSolver& S = *optimizer->solver_;
    auto& edgeSets = optimizer->edgeSets;
    auto& vertexSets = optimizer->vertexSets;
    S.buildStructure(edgeSets, vertexSets);
    cudaDeviceSynchronize();
    const int numP = (int)S.numP_, numL = (int)S.numL_;
    const int nHpl = S.d_Hpl_.nnz(), nblk = S.d_Hsc_.nnz();
    const int nmul = (int)S.d_HscMulBlockIds_.size(), nnzSymm = S.Hsc_.nnzSymm();
    dumpDev(prefix, "hpl_colptr", S.d_Hpl_.outerIndices(), (size_t)numL + 1);
    dumpDev(prefix, "hpl_rowind", S.d_Hpl_.innerIndices(), (size_t)nHpl);
    dumpDev(prefix, "hsc_rowptr", S.d_Hsc_.outerIndices(), (size_t)numP + 1);
    dumpDev(prefix, "hsc_colind", S.d_Hsc_.innerIndices(), (size_t)nblk);
    dumpHost(prefix, "csr_rowptr", S.Hsc_.rowPtr(), (size_t)numP * 6 + 1);
    dumpHost(prefix, "csr_colind", S.Hsc_.colInd(), (size_t)nnzSymm);
    dumpHost(prefix, "bsr2csr", S.Hsc_.BSR2CSR(), (size_t)nnzSymm);
    dumpDev(prefix, "mulids", (const int*)S.d_HscMulBlockIds_.data(), (size_t)nmul * 3);
    dumpDev(prefix, "edge2hpl", S.d_edge2Hpl_.data(), S.d_edge2Hpl_.size());
    dumpDev(prefix, "mono_edge2pl", (const int*)monoSet.d_edge2PL.data(), monoSet.nActiveEdges() * 2);
    dumpDev(prefix, "stereo_edge2pl", (const int*)stereoSet.d_edge2PL.data(), stereoSet.nActiveEdges() * 2);

    const double F0 = S.computeErrors(edgeSets, vertexSets);
    S.buildSystem(edgeSets, vertexSets);
    cudaDeviceSynchronize();
    const double maxDiag = S.maxDiagonal();
    const double lambda = 1e-5 * maxDiag;
    dumpDev(prefix, "Hpp", S.d_Hpp_.values(), (size_t)numP * 36);
    dumpDev(prefix, "Hll", S.d_Hll_.values(), (size_t)numL * 9);
    dumpDev(prefix, "Hpl", S.d_Hpl_.values(), (size_t)nHpl * 18);
    dumpDev(prefix, "b", S.d_b_.data(), (size_t)numP * 6 + (size_t)numL * 3);
    S.push();
    S.setLambda(lambda);
    const bool ok = S.solve();
    cudaDeviceSynchronize();
    dumpDev(prefix, "bsc", S.d_bsc_.values(), (size_t)numP * 6);
    dumpDev(prefix, "Hsc", S.d_Hsc_.values(), (size_t)nblk * 36);
    dumpDev(prefix, "x", S.d_x_.data(), (size_t)numP * 6 + (size_t)numL * 3);
    S.update(vertexSets);
    const double F1 = S.computeErrors(edgeSets, vertexSets);
    const double scale = S.computeScale(lambda);
    cudaDeviceSynchronize();
    dumpDev(prefix, "updated_estimates", S.d_solution_.data(), S.d_solution_.size());
    char buf[512];
    snprintf(buf, sizeof(buf),
             "  \"numP\": %d, \"numL\": %d, \"nHpl\": %d, \"nblk\": %d, \"nmul\": %d, \"nnzSymm\": %d,\n"
             "  \"stage\": {\"F0\": %.17g, \"maxDiag\": %.17g, \"lambda\": %.17g, \"solve_ok\": %d, "
             "\"F1\": %.17g, \"scale\": %.17g}\n}\n",
             numP, numL, nHpl, nblk, nmul, nnzSymm, F0, maxDiag, lambda, ok ? 1 : 0, F1, scale);
    json += buf;
    FILE* f = fopen((prefix + ".json").c_str(), "w");
    fputs(json.c_str(), f);
    fclose(f);
    printf("%s", json.c_str());
    return 0;
}
