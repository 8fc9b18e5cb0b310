// ref_driver_f4.cpp -- TEST INFRASTRUCTURE.  What of SURVEY.md 8(f) row 4 CAN be pinned to the UNMODIFIED reference
// (/root/reference, compiled by oracle/Makefile into oracle/_ref/): the RESIDUALS of its DepthEdgeSet, PlaneEdgeSet and
// LineEdgeSet (its Jacobians for these edges are not a model, see oracle/ba_oracle.c), and its LM behaviour on a depth graph.
// Drives the reference through its own API, as oracle/ref_driver.cpp does; compiled with -Dprivate=public.
//   depth: ref_driver_f4 depth graph.bag out_prefix niter
//          mono edges -> MonoEdgeSet, stereo edges (u, v, u_r) -> DepthEdgeSet with (u, v, (u - u_r) / bf);
//          dumps computeErrors() at the start, the per-edge errors keyed by (pose, landmark), chi2 per iteration of optimize(niter)
//   icp:   ref_driver_f4 icp problem.icp out_prefix
//          PoseVertexSet alone + PlaneEdgeSet + LineEdgeSet (ICP1 file, cugo_b200/graphio.py:write_icp); dumps computeErrors()
//          of each set alone and the per-edge errors with the measurements they belong to
#include <ba_types.h>
#include <icp_types.h>
#include <cuda_graph_optimisation.h>
#include <optimisable_graph.h>
#include <robust_kernel.h>
#include <block_solver.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

template <typename T>
static void rd(FILE* f, std::vector<T>& v, size_t n) {
    v.resize(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
}
template <typename T>
static void dumpHost(const std::string& prefix, const char* name, const T* p, size_t n) {
    const std::string fn = prefix + "." + name + ".bin";
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
static std::string listd(const std::vector<double>& v) {
    std::string s = "[";
    char buf[64];
    for (size_t i = 0; i < v.size(); i++) { snprintf(buf, sizeof(buf), "%s%.17g", i ? ", " : "", v[i]); s += buf; }
    return s + "]";
}

static int runDepth(const char* path, const std::string& prefix, int niter) {
    FILE* f = fopen(path, "rb");
    int hdr[6];
    if (!f || fread(hdr, sizeof(int), 6, f) != 6 || hdr[0] != 0x31474142) { fprintf(stderr, "bad graph file\n"); return 2; }
    const int P = hdr[1], L = hdr[2], Em = hdr[3], Es = hdr[4];
    double cam[5];
    if (fread(cam, sizeof(double), 5, f) != 5) return 2;
    std::vector<double> pose, lm, mmeas, minfo, smeas, sinfo;
    std::vector<int> pfixed, lfixed, mp, ml, sp, sl;
    rd(f, pose, (size_t)P * 7); rd(f, lm, (size_t)L * 3);
    rd(f, mmeas, (size_t)Em * 2); rd(f, minfo, Em); rd(f, smeas, (size_t)Es * 3); rd(f, sinfo, Es);
    rd(f, pfixed, P); rd(f, lfixed, L); rd(f, mp, Em); rd(f, ml, Em); rd(f, sp, Es); rd(f, sl, Es);
    fclose(f);

    cugo::GraphOptimisationOptions options;
    options.perEdgeInformation = true;
    options.perEdgeCamera = false;
    auto optimizer = std::make_unique<cugo::CudaGraphOptimisationImpl>(options);
    cugo::PoseVertexSet poseSet(false);
    cugo::LandmarkVertexSet lmSet(true);
    cugo::MonoEdgeSet monoSet;
    cugo::DepthEdgeSet depthSet;
    std::vector<cugo::PoseVertex*> pv(P);
    std::vector<cugo::LandmarkVertex*> lv(L);
    auto poseOf = [&](int i) {
        cugo::QuatD q; cugo::Vec3d t;
        for (int k = 0; k < 4; k++) q[k] = pose[7 * (size_t)i + k];
        for (int k = 0; k < 3; k++) t[k] = pose[7 * (size_t)i + 4 + k];
        return cugo::Se3D(q, t);
    };
    for (int i = 0; i < P; i++) { pv[i] = new cugo::PoseVertex(i, poseOf(i), pfixed[i] != 0); poseSet.addVertex(pv[i]); }
    for (int i = 0; i < L; i++) {
        cugo::Vec3d X;
        for (int k = 0; k < 3; k++) X[k] = lm[3 * (size_t)i + k];
        lv[i] = new cugo::LandmarkVertex(i, X, lfixed[i] != 0);
        lmSet.addVertex(lv[i]);
    }
    optimizer->addVertexSet(&poseSet);
    optimizer->addVertexSet(&lmSet);
    cugo::Camera camera(cam[0], cam[1], cam[2], cam[3], cam[4]);
    for (int i = 0; i < Em; i++) {
        auto* e = new cugo::MonoEdge();
        e->setVertex(pv[mp[i]], 0); e->setVertex(lv[ml[i]], 1);
        cugo::Vec2d m; m[0] = mmeas[2 * (size_t)i]; m[1] = mmeas[2 * (size_t)i + 1];
        e->setMeasurement(m); e->setInformation(minfo[i]); e->setCamera(camera);
        monoSet.addEdge(e);
    }
    for (int i = 0; i < Es; i++) {
        auto* e = new cugo::DepthEdge();
        e->setVertex(pv[sp[i]], 0); e->setVertex(lv[sl[i]], 1);
        cugo::Vec3d m;
        m[0] = smeas[3 * (size_t)i]; m[1] = smeas[3 * (size_t)i + 1];
        m[2] = (smeas[3 * (size_t)i] - smeas[3 * (size_t)i + 2]) / cam[4];
        e->setMeasurement(m); e->setInformation(sinfo[i]); e->setCamera(camera);
        depthSet.addEdge(e);
    }
    monoSet.setCamera(camera);
    depthSet.setCamera(camera);
    optimizer->addEdgeSet<cugo::MonoEdgeSet>(&monoSet);
    optimizer->addEdgeSet<cugo::DepthEdgeSet>(&depthSet);
    depthSet.setRobustKernel(cugo::RobustKernelType::None, 1.0);    // (process-global; the kernels dereference it)

    optimizer->initialize();
    cugo::BlockSolver& S = *optimizer->solver_;
    S.buildStructure(optimizer->edgeSets, optimizer->vertexSets);
    cudaDeviceSynchronize();
    const double F0 = S.computeErrors(optimizer->edgeSets, optimizer->vertexSets);
    cudaDeviceSynchronize();
    dumpDev(prefix, "depth_edge2pl", (const int*)depthSet.d_edge2PL.data(), depthSet.nActiveEdges() * 2);
    dumpDev(prefix, "depth_errors", (const double*)depthSet.d_errors.data(), depthSet.nActiveEdges() * 3);

    for (int i = 0; i < P; i++) pv[i]->setEstimate(poseOf(i));
    optimizer->setVerbose(true);
    optimizer->initialize();
    optimizer->optimize(niter);
    cudaDeviceSynchronize();
    std::vector<double> chi2;
    for (const auto& st : optimizer->batchStatistics().get()) chi2.push_back(st.chi2);
    char buf[256];
    snprintf(buf, sizeof(buf), "{\"P\": %d, \"L\": %d, \"Em\": %d, \"Es\": %d, \"niter\": %d, \"F0\": %.17g, \"chi2\": ", P, L, Em, Es, niter, F0);
    const std::string json = std::string(buf) + listd(chi2) + "}\n";
    FILE* o = fopen((prefix + ".json").c_str(), "w");
    fputs(json.c_str(), o);
    fclose(o);
    printf("%s", json.c_str());
    return 0;
}

static int runIcp(const char* path, const std::string& prefix) {
    FILE* f = fopen(path, "rb");
    int hdr[4];
    if (!f || fread(hdr, sizeof(int), 4, f) != 4 || hdr[0] != 0x31504349) { fprintf(stderr, "bad icp file\n"); return 2; }
    const int P = hdr[1], nPlane = hdr[2], nLine = hdr[3];
    std::vector<double> pose, pm, pw, lm, lw;
    std::vector<int> fixed, pp, lp;
    rd(f, pose, (size_t)P * 7); rd(f, fixed, P);
    rd(f, pp, nPlane); rd(f, pm, (size_t)nPlane * 7); rd(f, pw, nPlane);
    rd(f, lp, nLine); rd(f, lm, (size_t)nLine * 10); rd(f, lw, nLine);
    fclose(f);

    double F[2] = {0, 0};
    for (int which = 0; which < 2; which++) {        // each set alone: computeErrors() sums the sets
        cugo::GraphOptimisationOptions options;
        options.perEdgeInformation = true;
        options.perEdgeCamera = false;
        auto optimizer = std::make_unique<cugo::CudaGraphOptimisationImpl>(options);
        cugo::PoseVertexSet poseSet(false);
        std::vector<cugo::PoseVertex*> pv(P);
        for (int i = 0; i < P; i++) {
            cugo::QuatD q; cugo::Vec3d t;
            for (int k = 0; k < 4; k++) q[k] = pose[7 * (size_t)i + k];
            for (int k = 0; k < 3; k++) t[k] = pose[7 * (size_t)i + 4 + k];
            pv[i] = new cugo::PoseVertex(i, cugo::Se3D(q, t), fixed[i] != 0);
            poseSet.addVertex(pv[i]);
        }
        optimizer->addVertexSet(&poseSet);
        cugo::PlaneEdgeSet planeSet;
        cugo::LineEdgeSet lineSet;
        if (which == 0) {
            for (int i = 0; i < nPlane; i++) {
                const double* m = &pm[(size_t)i * 7];
                cugo::Vec3d n, p;
                for (int k = 0; k < 3; k++) { n[k] = m[k]; p[k] = m[4 + k]; }
                auto* e = new cugo::PlaneEdge();
                e->setVertex(pv[pp[i]], 0);
                e->setMeasurement(cugo::PointToPlaneMatch<double>(n, m[3], p));
                e->setInformation(pw[i]);
                planeSet.addEdge(e);
            }
            optimizer->addEdgeSet<cugo::PlaneEdgeSet>(&planeSet);
        } else {
            for (int i = 0; i < nLine; i++) {
                const double* m = &lm[(size_t)i * 10];
                cugo::Vec3d a, b, p;
                for (int k = 0; k < 3; k++) { a[k] = m[k]; b[k] = m[3 + k]; p[k] = m[7 + k]; }
                cugo::PointToLineMatch<double> match(a, b);
                match.pointP = p;
                auto* e = new cugo::LineEdge();
                e->setVertex(pv[lp[i]], 0);
                e->setMeasurement(match);
                e->setInformation(lw[i]);
                lineSet.addEdge(e);
            }
            optimizer->addEdgeSet<cugo::LineEdgeSet>(&lineSet);
        }
        optimizer->initialize();
        cugo::BlockSolver& S = *optimizer->solver_;
        S.buildStructure(optimizer->edgeSets, optimizer->vertexSets);
        cudaDeviceSynchronize();
        F[which] = S.computeErrors(optimizer->edgeSets, optimizer->vertexSets);
        cudaDeviceSynchronize();
        if (which == 0) {
            dumpDev(prefix, "plane_meas", (const double*)planeSet.d_measurements.data(), planeSet.nActiveEdges() * 7);
            dumpDev(prefix, "plane_errors", (const double*)planeSet.d_errors.data(), planeSet.nActiveEdges());
            dumpDev(prefix, "plane_edge2pl", (const int*)planeSet.d_edge2PL.data(), planeSet.nActiveEdges() * 2);
        } else {
            dumpDev(prefix, "line_meas", (const double*)lineSet.d_measurements.data(), lineSet.nActiveEdges() * 10);
            dumpDev(prefix, "line_errors", (const double*)lineSet.d_errors.data(), lineSet.nActiveEdges());
            dumpDev(prefix, "line_edge2pl", (const int*)lineSet.d_edge2PL.data(), lineSet.nActiveEdges() * 2);
        }
    }
    char buf[256];
    snprintf(buf, sizeof(buf), "{\"P\": %d, \"nPlane\": %d, \"nLine\": %d, \"F_plane\": %.17g, \"F_line\": %.17g}\n", P, nPlane, nLine, F[0], F[1]);
    FILE* o = fopen((prefix + ".json").c_str(), "w");
    fputs(buf, o);
    fclose(o);
    printf("%s", buf);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 5 && !strcmp(argv[1], "depth")) return runDepth(argv[2], argv[3], atoi(argv[4]));
    if (argc >= 4 && !strcmp(argv[1], "icp")) return runIcp(argv[2], argv[3]);
    fprintf(stderr, "usage: ref_driver_f4 depth graph.bag out_prefix niter | icp problem.icp out_prefix\n");
    return 1;
}
