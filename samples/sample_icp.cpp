// sample_icp.cpp -- a pose-only graph through include/cugo_b200.hpp: the reference's PlaneEdgeSet / LineEdgeSet
// (include/icp_types.h:19-125) with the pose vertex set alone, the calls in the order of its BA sample
// (samples/sample_ba_from_file/main.cpp:78-201).  The matches come from an ICP1 file written by the test
// (tests/test_poseonly.py: header {magic, P, nPlane, nLine}, poses [P][7], fixed [P] int32, plane pose ids [nPlane] int32,
// plane matches [nPlane][7], plane informations [nPlane], line pose ids, line matches [nLine][10], line informations).
// Prints one JSON object with the chi2 of every iteration; optionally dumps the final poses by id.
//
//   usage: sample_icp problem.icp niter [robust(0 none,1 cauchy,2 tukey,3 huber) delta] [poses.bin]
#include <cugo_b200.hpp>

#include <cstdio>
#include <cstdlib>
#include <vector>

template <typename T>
static void rd(FILE* f, std::vector<T>& v, size_t n) {
    v.resize(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
}

static int run(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s problem.icp niter [robust delta] [poses.bin]\n", argv[0]); return 1; }
    FILE* f = fopen(argv[1], "rb");
    int hdr[4];
    if (!f || fread(hdr, sizeof(int), 4, f) != 4 || hdr[0] != 0x31504349) { fprintf(stderr, "cannot read %s\n", argv[1]); return 2; }
    const int P = hdr[1], nPlane = hdr[2], nLine = hdr[3];
    std::vector<double> pose, pm, pw, lm, lw;
    std::vector<int> fixed, pp, lp;
    rd(f, pose, (size_t)P * 7); rd(f, fixed, P);
    rd(f, pp, nPlane); rd(f, pm, (size_t)nPlane * 7); rd(f, pw, nPlane);
    rd(f, lp, nLine); rd(f, lm, (size_t)nLine * 10); rd(f, lw, nLine);
    fclose(f);
    const int niter = atoi(argv[2]);
    const int rk = argc > 3 ? atoi(argv[3]) : 0;
    const double delta = argc > 4 ? atof(argv[4]) : 1.0;
    const char* dump = argc > 5 ? argv[5] : nullptr;

    cugo::GraphOptimisationOptions options;
    options.perEdgeInformation = true;
    auto optimizer = std::make_unique<cugo::CudaGraphOptimisationImpl>(options);
    auto* poseVertexSet = new cugo::PoseVertexSet(false);
    auto* planeEdgeSet = new cugo::PlaneEdgeSet();
    auto* lineEdgeSet = new cugo::LineEdgeSet();
    for (int i = 0; i < P; i++)
        poseVertexSet->addVertex(new cugo::PoseVertex(i, cugo::Se3D(cugo::QuatD(&pose[(size_t)i * 7]), cugo::Vec3d(&pose[(size_t)i * 7 + 4])),
                                                      fixed[i] != 0));
    optimizer->addVertexSet(poseVertexSet);
    for (int i = 0; i < nPlane; i++) {
        const double* m = &pm[(size_t)i * 7];
        auto* e = new cugo::PlaneEdge();
        e->setVertex(poseVertexSet->getVertex(pp[i]), 0);
        e->setMeasurement(cugo::PointToPlaneMatch(cugo::Vec3d(m), m[3], cugo::Vec3d(m + 4)));
        e->setInformation(pw[i]);
        planeEdgeSet->addEdge(e);
    }
    for (int i = 0; i < nLine; i++) {
        const double* m = &lm[(size_t)i * 10];
        auto* e = new cugo::LineEdge();
        e->setVertex(poseVertexSet->getVertex(lp[i]), 0);
        e->setMeasurement(cugo::PointToLineMatch(cugo::Vec3d(m), cugo::Vec3d(m + 3), cugo::Vec3d(m + 7)));
        e->setInformation(lw[i]);
        lineEdgeSet->addEdge(e);
    }
    optimizer->addEdgeSet<cugo::PlaneEdgeSet>(planeEdgeSet);
    optimizer->addEdgeSet<cugo::LineEdgeSet>(lineEdgeSet);
    planeEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);
    lineEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);

    optimizer->initialize();
    optimizer->optimize(niter);

    printf("{\"num_poses\": %zu, \"num_edges\": [%zu, %zu], \"chi2\": [", optimizer->nVertices(0), planeEdgeSet->nedges(),
           lineEdgeSet->nedges());
    const auto& stats = optimizer->batchStatistics().get();
    for (size_t i = 0; i < stats.size(); i++) printf("%s%.17g", i ? ", " : "", stats[i].chi2);
    printf("]}\n");
    if (dump) {
        FILE* o = fopen(dump, "wb");
        if (!o) return 3;
        for (int i = 0; i < P; i++) {
            double v[7];
            poseVertexSet->getVertex(i)->getEstimate().copyTo(v, v + 4);
            fwrite(v, sizeof(double), 7, o);
        }
        fclose(o);
    }
    for (auto* e : planeEdgeSet->get()) delete e;
    for (auto* e : lineEdgeSet->get()) delete e;
    for (auto& kv : poseVertexSet->get()) delete kv.second;
    delete planeEdgeSet; delete lineEdgeSet; delete poseVertexSet;
    return 0;
}

int main(int argc, char** argv) {
    try {
        return run(argc, argv);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 4;
    }
}
