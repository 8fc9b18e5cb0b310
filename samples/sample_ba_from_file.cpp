// sample_ba_from_file.cpp -- the reference's sample (samples/sample_ba_from_file/main.cpp:78-201) and its
// comparison sample (samples/sample_comparison_with_g2o/main.cpp:101-148) against include/cugo_b200.hpp.
// The graph is the reference's own JSON schema, parsed by include/cugo_b200_json.hpp (OpenCV's FileStorage is
// not needed).  Same calls in the same order as the reference: per-edge information and per-edge camera, pose
// set first, a warm-up initialize() + optimize(1), then the timed initialize() + optimize(niter).
//
//   sample_ba_from_file graph.json [niter=10] [--no-warmup] [--robust none|cauchy|tukey|huber delta]
//                       [--dump result.json]      final estimates + "chi2": [...] in the same schema
//                       [--compare other.json]    chi2 table + rotation / translation / landmark RMSE against the
//                                                 run stored in other.json (a --dump of another engine: the CPU
//                                                 oracle, the reference library, g2o)
//                       [--convert out.json]      parse and re-write the graph only (no GPU needed)
#include <cugo_b200.hpp>
#include <cugo_b200_json.hpp>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

// "chi2": [...] of a result file (the graph reader skips unknown keys)
static std::vector<double> readChi2(const std::string& path) {
    std::vector<double> out;
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return out;
    std::string s;
    char buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof(buf), f)) > 0) s.append(buf, n);
    fclose(f);
    const size_t k = s.find("\"chi2\"");
    if (k == std::string::npos) return out;
    size_t p = s.find('[', k);
    if (p == std::string::npos) return out;
    p++;
    while (p < s.size() && s[p] != ']') {
        char* end = nullptr;
        const double v = strtod(s.c_str() + p, &end);
        if (end == s.c_str() + p) { p++; continue; }
        out.push_back(v);
        p = (size_t)(end - s.c_str());
    }
    return out;
}

static int run(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: %s graph.json [niter] [--no-warmup] [--robust kind delta] [--dump out.json] [--compare other.json] [--convert out.json]\n", argv[0]);
        return 1;
    }
    const std::string path = argv[1];
    int niter = 10, rk = 0;
    double delta = 1.0;
    bool warmup = true;
    const char *dump = nullptr, *compare = nullptr, *convert = nullptr;
    for (int i = 2; i < argc; i++) {
        if (!strcmp(argv[i], "--no-warmup")) warmup = false;
        else if (!strcmp(argv[i], "--dump") && i + 1 < argc) dump = argv[++i];
        else if (!strcmp(argv[i], "--compare") && i + 1 < argc) compare = argv[++i];
        else if (!strcmp(argv[i], "--convert") && i + 1 < argc) convert = argv[++i];
        else if (!strcmp(argv[i], "--robust") && i + 2 < argc) {
            const char* k = argv[++i];
            rk = !strcmp(k, "cauchy") ? 1 : !strcmp(k, "tukey") ? 2 : !strcmp(k, "huber") ? 3 : 0;
            delta = atof(argv[++i]);
        } else niter = atoi(argv[i]);
    }

    printf("Reading Graph... ");
    fflush(stdout);
    cugo::io::GraphFile g = cugo::io::readJson(path);
    printf("Done.\n\n");
    if (convert) { cugo::io::writeJson(g, convert); return 0; }

    cugo::GraphOptimisationOptions options;
    options.perEdgeInformation = true;
    options.perEdgeCamera = true;
    auto optimizer = std::make_unique<cugo::CudaGraphOptimisationImpl>(options);
    auto* poseVertexSet = new cugo::PoseVertexSet(false);
    auto* landmarkVertexSet = new cugo::LandmarkVertexSet(true);
    auto* monoEdgeSet = new cugo::MonoEdgeSet();
    auto* stereoEdgeSet = new cugo::StereoEdgeSet();

    for (size_t i = 0; i < g.nPoses(); i++) {
        const cugo::QuatD q(&g.poseQT[i * 7]);
        const cugo::Vec3d t(&g.poseQT[i * 7 + 4]);
        poseVertexSet->addVertex(new cugo::PoseVertex(g.poseId[i], cugo::Se3D(q, t), g.poseFixed[i] != 0));
    }
    optimizer->addVertexSet(poseVertexSet);
    for (size_t i = 0; i < g.nLandmarks(); i++)
        landmarkVertexSet->addVertex(new cugo::LandmarkVertex(g.lmId[i], cugo::Vec3d(&g.lmXw[i * 3]), g.lmFixed[i] != 0));
    optimizer->addVertexSet(landmarkVertexSet);

    const cugo::Camera camera(g.cam[0], g.cam[1], g.cam[2], g.cam[3], g.cam[4]);
    auto vertexOrDie = [](auto* set, int id, const char* what) {
        auto* v = set->getVertex(id);
        if (!v) throw std::runtime_error(std::string("edge refers to an unknown ") + what + " id " + std::to_string(id));
        return v;
    };
    for (size_t i = 0; i < g.monoP.size(); i++) {
        auto* e = new cugo::MonoEdge();
        e->setVertex(vertexOrDie(poseVertexSet, g.monoP[i], "pose"), 0);
        e->setVertex(vertexOrDie(landmarkVertexSet, g.monoL[i], "landmark"), 1);
        e->setMeasurement(cugo::Vec2d(&g.monoMeas[i * 2]));
        e->setInformation(g.monoInfo[i]);
        e->setCamera(camera);
        monoEdgeSet->addEdge(e);
    }
    for (size_t i = 0; i < g.stereoP.size(); i++) {
        auto* e = new cugo::StereoEdge();
        e->setVertex(vertexOrDie(poseVertexSet, g.stereoP[i], "pose"), 0);
        e->setVertex(vertexOrDie(landmarkVertexSet, g.stereoL[i], "landmark"), 1);
        e->setMeasurement(cugo::Vec3d(&g.stereoMeas[i * 3]));
        e->setInformation(g.stereoInfo[i]);
        e->setCamera(camera);
        stereoEdgeSet->addEdge(e);
    }
    optimizer->addEdgeSet<cugo::MonoEdgeSet>(monoEdgeSet);
    optimizer->addEdgeSet<cugo::StereoEdgeSet>(stereoEdgeSet);
    monoEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);
    stereoEdgeSet->setRobustKernel((cugo::RobustKernelType)rk, delta);

    if (warmup) {   // "warm-up" of the reference's sample; its result is written back into the vertices
        optimizer->initialize();
        optimizer->optimize(1);
    }
    printf("=== Graph size : \nnum poses      : %zu\nnum landmarks  : %zu\n", optimizer->nVertices(0), optimizer->nVertices(1));
    for (const auto* es : optimizer->getEdgeSets()) printf("num edges      : %zu\n\n", es->nedges());

    printf("Running BA... ");
    fflush(stdout);
    const auto t0 = std::chrono::steady_clock::now();
    optimizer->initialize();
    optimizer->optimize(niter);
    const auto t1 = std::chrono::steady_clock::now();
    printf("Done.\n\n=== Processing time : \nBA total : %.2f [msec]\n\n", std::chrono::duration<double, std::milli>(t1 - t0).count());
    for (auto& kv : optimizer->timeProfile()) printf("%s : %.3f\n", kv.first.c_str(), kv.second);

    printf("\n=== Objective function value : \n");
    std::vector<double> chi2;
    for (const auto& stat : optimizer->batchStatistics().get()) {
        printf("iter: %2d, chi2: %.1f\n", stat.iteration + 1, stat.chi2);
        chi2.push_back(stat.chi2);
    }

    // final estimates in file order, under the file's ids
    cugo::io::GraphFile res = g;
    for (size_t i = 0; i < g.nPoses(); i++) poseVertexSet->getVertex(g.poseId[i])->getEstimate().copyTo(&res.poseQT[i * 7], &res.poseQT[i * 7 + 4]);
    for (size_t i = 0; i < g.nLandmarks(); i++) landmarkVertexSet->getVertex(g.lmId[i])->getEstimate().copyTo(&res.lmXw[i * 3]);
    if (dump) {
        cugo::io::writeJson(res, dump);
        // append "chi2" by re-opening: replace the closing brace
        FILE* f = fopen(dump, "r+");
        if (f) {
            fseek(f, -2, SEEK_END);
            fprintf(f, ",\n\"chi2\": [");
            for (size_t i = 0; i < chi2.size(); i++) fprintf(f, "%s%.17g", i ? ", " : "", chi2[i]);
            fprintf(f, "]\n}\n");
            fclose(f);
        }
    }
    if (compare) {
        const cugo::io::GraphFile other = cugo::io::readJson(compare);
        cugo::io::Estimates a{res.poseId, res.lmId, res.poseQT, res.lmXw}, b{other.poseId, other.lmId, other.poseQT, other.lmXw};
        printf("\n");
        cugo::io::printComparison(readChi2(compare), chi2, b, a, "ref", "B200");
    }

    // the optimiser owns nothing: the client frees vertices, edges and sets (include/cuda_graph_optimisation.h:129-130)
    for (auto* e : monoEdgeSet->get()) delete e;
    for (auto* e : stereoEdgeSet->get()) delete e;
    for (auto& kv : poseVertexSet->get()) delete kv.second;
    for (auto& kv : landmarkVertexSet->get()) delete kv.second;
    delete monoEdgeSet; delete stereoEdgeSet; delete poseVertexSet; delete landmarkVertexSet;
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
