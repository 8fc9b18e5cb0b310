// sample_ba_from_bag.cpp -- the reference's sample (samples/sample_ba_from_file/main.cpp:78-201)
// rewritten against include/cugo_b200.hpp: same calls in the same order, the graph comes from a BAG1
// file (cugo_b200/graphio.py; same fields as the reference's JSON schema) instead of OpenCV
// FileStorage.  Prints one JSON object: chi2 per iteration of the timed run, timings, final estimates'
// checksum; optionally dumps the final estimates (poses by id [P][7], landmarks by id [L][3]).
//
//   g++ -std=c++17 -O2 -Iinclude samples/sample_ba_from_bag.cpp -Lcuda-bundle-adjustment_b200/lib -lcugo_b200
//   usage: sample_ba_from_bag graph.bag niter [robust(0 none,1 cauchy,2 tukey,3 huber) delta] [estimates.bin] [warmup 0|1] [depth 0|1]
//   depth 1: the stereo edges (u, v, u_r) go into a DepthEdgeSet as (u, v, (u - u_r) / bf) -- include/ba_types.h:171-234
#include <cugo_b200.hpp>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

struct GraphFile {
    int P = 0, L = 0, Em = 0, Es = 0;
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
    int hdr[6];
    if (!f || fread(hdr, sizeof(int), 6, f) != 6 || hdr[0] != 0x31474142) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
    g.P = hdr[1]; g.L = hdr[2]; g.Em = hdr[3]; g.Es = hdr[4];
    if (fread(g.cam, sizeof(double), 5, f) != 5) exit(2);
    rd(f, g.pose, (size_t)g.P * 7); rd(f, g.lm, (size_t)g.L * 3);
    rd(f, g.mmeas, (size_t)g.Em * 2); rd(f, g.minfo, g.Em); rd(f, g.smeas, (size_t)g.Es * 3); rd(f, g.sinfo, g.Es);
    rd(f, g.pfixed, g.P); rd(f, g.lfixed, g.L); rd(f, g.mp, g.Em); rd(f, g.ml, g.Em); rd(f, g.sp, g.Es); rd(f, g.sl, g.Es);
    fclose(f);
    return g;
}

static int run(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s graph.bag niter [robust delta] [estimates.bin] [warmup]\n", argv[0]); return 1; }
    const GraphFile g = readGraph(argv[1]);
    const int niter = atoi(argv[2]);
    const int rk = argc > 3 ? atoi(argv[3]) : 0;
    const double delta = argc > 4 ? atof(argv[4]) : 1.0;
    const char* dump = argc > 5 ? argv[5] : nullptr;
    const bool warmup = argc > 6 ? atoi(argv[6]) != 0 : false;
    const bool depth = argc > 7 ? atoi(argv[7]) != 0 : false;

    cugo::GraphOptimisationOptions options;
    options.perEdgeInformation = true;
    options.perEdgeCamera = false;
    auto optimizer = std::make_unique<cugo::CudaGraphOptimisationImpl>(options);

    auto* poseVertexSet = new cugo::PoseVertexSet(false);
    auto* landmarkVertexSet = new cugo::LandmarkVertexSet(true);
    auto* monoEdgeSet = new cugo::MonoEdgeSet();
    auto* stereoEdgeSet = new cugo::StereoEdgeSet();
    auto* depthEdgeSet = new cugo::DepthEdgeSet();

    for (int i = 0; i < g.P; i++) {
        const cugo::QuatD q(&g.pose[(size_t)i * 7]);
        const cugo::Vec3d t(&g.pose[(size_t)i * 7 + 4]);
        poseVertexSet->addVertex(new cugo::PoseVertex(i, cugo::Se3D(q, t), g.pfixed[i] != 0));
    }
    optimizer->addVertexSet(poseVertexSet);
    for (int i = 0; i < g.L; i++)
        landmarkVertexSet->addVertex(new cugo::LandmarkVertex(i, cugo::Vec3d(&g.lm[(size_t)i * 3]), g.lfixed[i] != 0));
    optimizer->addVertexSet(landmarkVertexSet);

    const cugo::Camera camera(g.cam[0], g.cam[1], g.cam[2], g.cam[3], g.cam[4]);
    monoEdgeSet->setCamera(camera);
    stereoEdgeSet->setCamera(camera);
    for (int i = 0; i < g.Em; i++) {
        auto* e = new cugo::MonoEdge();
        e->setVertex(poseVertexSet->getVertex(g.mp[i]), 0);
        e->setVertex(landmarkVertexSet->getVertex(g.ml[i]), 1);
        e->setMeasurement(cugo::Vec2d(&g.mmeas[(size_t)i * 2]));
        e->setInformation(g.minfo[i]);
        e->setCamera(camera);
        monoEdgeSet->addEdge(e);
    }
    depthEdgeSet->setCamera(camera);
    for (int i = 0; i < g.Es; i++) {
        auto* e = new cugo::StereoEdge();       // (DepthEdge is the same 3-row edge type)
        e->setVertex(poseVertexSet->getVertex(g.sp[i]), 0);
        e->setVertex(landmarkVertexSet->getVertex(g.sl[i]), 1);
        cugo::Vec3d m(&g.smeas[(size_t)i * 3]);
        if (depth) m[2] = (m[0] - m[2]) / g.cam[4];
        e->setMeasurement(m);
        e->setInformation(g.sinfo[i]);
        e->setCamera(camera);
        if (depth) depthEdgeSet->addEdge(e); else stereoEdgeSet->addEdge(e);
    }
    optimizer->addEdgeSet<cugo::MonoEdgeSet>(monoEdgeSet);
    if (depth) optimizer->addEdgeSet<cugo::DepthEdgeSet>(depthEdgeSet);
    else optimizer->addEdgeSet<cugo::StereoEdgeSet>(stereoEdgeSet);
    monoEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);
    stereoEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);
    depthEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);

    if (warmup) {   // as the reference's sample: a one-iteration run whose result is written back
        optimizer->initialize();
        optimizer->optimize(1);
    }
    const auto t0 = std::chrono::steady_clock::now();
    optimizer->initialize();
    const auto t1 = std::chrono::steady_clock::now();
    optimizer->optimize(niter);
    const auto t2 = std::chrono::steady_clock::now();

    printf("{\"num_poses\": %zu, \"num_landmarks\": %zu, \"num_edges\": [%zu, %zu], \"chi2\": [", optimizer->nVertices(0),
           optimizer->nVertices(1), optimizer->getEdgeSets()[0]->nedges(), optimizer->getEdgeSets()[1]->nedges());
    const auto& stats = optimizer->batchStatistics().get();
    for (size_t i = 0; i < stats.size(); i++) printf("%s%.17g", i ? ", " : "", stats[i].chi2);
    printf("], \"ms_initialize\": %.3f, \"ms_optimize\": %.3f, \"profile\": {",
           std::chrono::duration<double, std::milli>(t1 - t0).count(), std::chrono::duration<double, std::milli>(t2 - t1).count());
    bool first = true;
    for (auto& kv : optimizer->timeProfile()) { printf("%s\"%s\": %.4f", first ? "" : ", ", kv.first.c_str(), kv.second); first = false; }
    printf("}}\n");

    if (dump) {
        FILE* f = fopen(dump, "wb");
        if (!f) return 3;
        for (int i = 0; i < g.P; i++) {
            double v[7];
            poseVertexSet->getVertex(i)->getEstimate().copyTo(v, v + 4);
            fwrite(v, sizeof(double), 7, f);
        }
        for (int i = 0; i < g.L; i++) fwrite(landmarkVertexSet->getVertex(i)->getEstimate().data, sizeof(double), 3, f);
        fclose(f);
    }
    // the optimiser owns nothing: the client frees vertices, edges and sets (include/cuda_graph_optimisation.h:129-130)
    for (auto* e : monoEdgeSet->get()) delete e;
    for (auto* e : stereoEdgeSet->get()) delete e;
    for (auto* e : depthEdgeSet->get()) delete e;
    for (auto& kv : poseVertexSet->get()) delete kv.second;
    for (auto& kv : landmarkVertexSet->get()) delete kv.second;
    delete monoEdgeSet; delete stereoEdgeSet; delete depthEdgeSet; delete poseVertexSet; delete landmarkVertexSet;
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
