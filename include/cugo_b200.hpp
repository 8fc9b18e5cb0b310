// cugo_b200.hpp -- C++ host facade of the B200 bundle-adjustment engine.
//
// Header-only, depends only on the C ABI (cugo_b200.h / libcugo_b200.so).  It mirrors the public
// interface a client of the reference programs against, with the same names and argument meaning
// in `namespace cugo`, so that code written like the reference's sample
// (samples/sample_ba_from_file/main.cpp:78-201) compiles against this header unchanged:
//
//   reference interface                                   here
//   include/cuda_graph_optimisation.h:132-239             CudaGraphOptimisation, CudaGraphOptimisationImpl
//   include/cuda_graph_optimisation.h:49-107              BatchInfo, BatchStatistics, TimeProfile
//   src/graph_optimisation_options.h                      GraphOptimisationOptions
//   src/optimisable_graph.h:40-151   (vertices)           BaseVertex, Vertex<T, Marg>, PoseVertex, LandmarkVertex
//   src/optimisable_graph.h:161-330  (vertex sets)        BaseVertexSet, VertexSet<T>, PoseVertexSet, LandmarkVertexSet
//   src/optimisable_graph.h:341-505  (edges)              BaseEdge, Edge<DIM>, MonoEdge, StereoEdge
//   src/optimisable_graph.h:510-680  (edge sets)          BaseEdgeSet, EdgeSet<DIM>, MonoEdgeSet, StereoEdgeSet
//   src/fixed_vector.h:29-474, src/camera.h               Vec<T,N>, Vec2d, Vec3d, QuatD, Se3D, Camera
//   src/robust_kernel.h:12-17                             RobustKernelType (+ Huber)
//
// Differences that are deliberate: the robust kernel belongs to its edge set
// (process-global in the reference); edges are packed in insertion order (pointer-hash order
// there); the optimiser is bound to a caller-chosen device; errors are std::runtime_error carrying
// the engine's message (raw CUDA enums thrown through noexcept there).  Ownership is the
// reference's: the optimiser owns neither vertices, edges nor sets.
#pragma once
#include "cugo_b200.h"

#include <array>
#include <chrono>
#include <cstddef>
#include <cstdio>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_set>
#include <vector>

namespace cugo {

using Scalar = double;
using Information = double;

// ---- small fixed-size types (src/fixed_vector.h) --------------------------------------------
template <typename T, int N>
struct Vec {
    T data[N];
    Vec() : data{} {}
    explicit Vec(const T* v) { for (int i = 0; i < N; i++) data[i] = v[i]; }
    Vec(const std::array<T, N>& a) { for (int i = 0; i < N; i++) data[i] = a[i]; }
    template <typename... U, typename = typename std::enable_if<sizeof...(U) == N && (N > 1)>::type>
    Vec(U... v) : data{static_cast<T>(v)...} {}
    T& operator[](int i) noexcept { return data[i]; }
    const T& operator[](int i) const noexcept { return data[i]; }
    void copyTo(T* out) const noexcept { for (int i = 0; i < N; i++) out[i] = data[i]; }
};
using Vec2d = Vec<double, 2>;
using Vec3d = Vec<double, 3>;
using Vec4d = Vec<double, 4>;
using Vec5d = Vec<double, 5>;

// quaternion stored (x, y, z, w), as the reference's Quat<T>
template <typename T>
struct Quat : Vec<T, 4> {
    using Vec<T, 4>::Vec;
    Quat() = default;
    Quat(const Vec<T, 4>& v) : Vec<T, 4>(v) {}
};
using QuatD = Quat<double>;

// world -> camera transform: rotation r (quaternion) and translation t
template <typename T>
struct Se3 {
    Quat<T> r;
    Vec<T, 3> t;
    Se3() = default;
    Se3(const Quat<T>& q, const Vec<T, 3>& v) : r(q), t(v) {}
    Se3(const T* rv, const T* tv) : r(rv), t(tv) {}
    void copyTo(T* rOut, T* tOut) const noexcept { r.copyTo(rOut); t.copyTo(tOut); }
};
using Se3D = Se3<double>;

struct Camera {
    double fx = 0, fy = 0, cx = 0, cy = 0, bf = 0;
    Camera() = default;
    Camera(double fx_, double fy_, double cx_, double cy_, double bf_ = 1.0) : fx(fx_), fy(fy_), cx(cx_), cy(cy_), bf(bf_) {}
};

enum class RobustKernelType { None = 0, Cauchy = 1, Tukey = 2, Huber = 3 };

struct GraphOptimisationOptions {
    bool perEdgeInformation = false;   // false: one information value per edge set (EdgeSet::setInformation)
    bool perEdgeCamera = false;        // false: one camera per edge set (EdgeSet::setCamera)
};

struct BatchInfo {
    int iteration;
    double chi2;
};
class BatchStatistics {
public:
    BatchInfo getStartStats() const { return stats_.at(0); }
    BatchInfo getLastStats() const { return stats_.at(stats_.size() - 1); }
    BatchInfo getStatEntry(int idx) const { return stats_.at(idx); }
    void addStat(const BatchInfo& b) { stats_.push_back(b); }
    const std::vector<BatchInfo>& get() { return stats_; }
    void clear() { stats_.clear(); }
private:
    std::vector<BatchInfo> stats_;
};
using TimeProfile = std::map<std::string, double>;

class BaseEdge;
using EdgeContainer = std::unordered_set<BaseEdge*>;

// ---- vertices (src/optimisable_graph.h:40-151) ------------------------------------------------
class BaseVertex {
public:
    virtual ~BaseVertex() {}
    virtual int getId() const noexcept = 0;
    virtual void setId(int id) noexcept = 0;
    virtual EdgeContainer& getEdges() noexcept = 0;
    virtual void addEdge(BaseEdge* e) = 0;
    virtual void removeEdge(BaseEdge* e) = 0;
    virtual void setFixed(bool status) noexcept = 0;
    virtual bool isFixed() const noexcept = 0;
    virtual int getIndex() const noexcept = 0;      // position in the solver's ordering (set by initialize)
    virtual void setIndex(int idx) noexcept = 0;
    virtual bool isMarginilised() const noexcept = 0;
    virtual void clearEdges() noexcept = 0;
};

template <typename T, bool Marginilised>
class Vertex : public BaseVertex {
public:
    using EstimateType = T;
    Vertex(int id, const EstimateType& est, bool fixed = false) : est_(est), fixed_(fixed), id_(id) {}
    explicit Vertex(int id, bool fixed = false) : est_(), fixed_(fixed), id_(id) {}
    EstimateType& getEstimate() noexcept { return est_; }
    const EstimateType& getEstimate() const noexcept { return est_; }
    void setEstimate(const EstimateType& e) noexcept { est_ = e; }
    int getId() const noexcept override { return id_; }
    void setId(int id) noexcept override { id_ = id; }
    EdgeContainer& getEdges() noexcept override { return edges_; }
    void addEdge(BaseEdge* e) override { edges_.insert(e); }
    void removeEdge(BaseEdge* e) override { edges_.erase(e); }
    void setFixed(bool s) noexcept override { fixed_ = s; }
    bool isFixed() const noexcept override { return fixed_; }
    int getIndex() const noexcept override { return idx_; }
    void setIndex(int i) noexcept override { idx_ = i; }
    bool isMarginilised() const noexcept override { return Marginilised; }
    void clearEdges() noexcept override { edges_.clear(); }
private:
    EstimateType est_;
    bool fixed_;
    int id_;
    int idx_ = -1;
    EdgeContainer edges_;
};
using PoseVertex = Vertex<Se3D, false>;
using LandmarkVertex = Vertex<Vec3d, true>;

// ---- vertex sets (src/optimisable_graph.h:161-330) -------------------------------------------------
class BaseVertexSet {
public:
    virtual ~BaseVertexSet() {}
    virtual size_t size() const noexcept = 0;
    virtual bool isMarginilised() const noexcept = 0;
    virtual int getActiveSize() const noexcept = 0;
    virtual void clearVertices() noexcept = 0;
    // facade <-> engine packing (the reference's generateEstimateData / finalise,
    // src/optimisable_graph.hpp:84-154): non-fixed vertices first in ascending id, then fixed ones
    virtual void assignIndices() = 0;
    virtual void packEstimates(std::vector<double>& out) const = 0;
    virtual void unpackEstimates(const double* in) = 0;
    virtual int estimateDim() const noexcept = 0;
};

namespace detail {
inline void packOne(const Se3D& e, double* o) { e.copyTo(o, o + 4); }
inline void packOne(const Vec3d& e, double* o) { e.copyTo(o); }
inline void unpackOne(Se3D& e, const double* i) { e = Se3D(i, i + 4); }
inline void unpackOne(Vec3d& e, const double* i) { e = Vec3d(i); }
template <typename E> struct EstDim;
template <> struct EstDim<Se3D> { static constexpr int value = 7; };
template <> struct EstDim<Vec3d> { static constexpr int value = 3; };
}  // namespace detail

template <typename T>
class VertexSet : public BaseVertexSet {
public:
    explicit VertexSet(bool marginilised) : marg_(marginilised) {}
    void addVertex(T* v) { vertices_[v->getId()] = v; dirty_ = true; }
    T* getVertex(int id) const {
        auto it = vertices_.find(id);
        return it == vertices_.end() ? nullptr : it->second;
    }
    bool removeVertex(T* v) {
        dirty_ = true;
        return vertices_.erase(v->getId()) > 0;
    }
    size_t size() const noexcept override { return vertices_.size(); }
    bool isMarginilised() const noexcept override { return marg_; }
    int getActiveSize() const noexcept override { return active_; }
    void clearVertices() noexcept override { vertices_.clear(); order_.clear(); active_ = 0; dirty_ = true; }
    int estimateDim() const noexcept override { return detail::EstDim<typename T::EstimateType>::value; }
    void assignIndices() override {
        order_.clear();
        order_.reserve(vertices_.size());
        for (auto& kv : vertices_) if (!kv.second->isFixed()) order_.push_back(kv.second);
        active_ = (int)order_.size();
        for (auto& kv : vertices_) if (kv.second->isFixed()) order_.push_back(kv.second);
        for (size_t i = 0; i < order_.size(); i++) order_[i]->setIndex((int)i);
    }
    void packEstimates(std::vector<double>& out) const override {
        const int dim = estimateDim();
        out.resize(order_.size() * dim);
        for (size_t i = 0; i < order_.size(); i++) detail::packOne(order_[i]->getEstimate(), out.data() + i * dim);
    }
    void unpackEstimates(const double* in) override {
        const int dim = estimateDim();
        for (size_t i = 0; i < order_.size(); i++) detail::unpackOne(order_[i]->getEstimate(), in + i * dim);
    }
    const std::map<int, T*>& get() const noexcept { return vertices_; }
private:
    bool marg_;
    bool dirty_ = true;
    int active_ = 0;
    std::map<int, T*> vertices_;     // ascending id, as the reference's VertexMap
    std::vector<T*> order_;          // index order
};
using PoseVertexSet = VertexSet<PoseVertex>;
using LandmarkVertexSet = VertexSet<LandmarkVertex>;

// ---- edges (src/optimisable_graph.h:341-505) ------------------------------------------------------
class BaseEdge {
public:
    virtual ~BaseEdge() {}
    virtual BaseVertex* getVertex(int index) = 0;            // 0: pose, 1: landmark
    virtual void setVertex(BaseVertex* v, int index) = 0;
    virtual bool allVerticesFixed() const noexcept = 0;
    virtual bool anyVerticesNotFixed() const noexcept = 0;
    virtual bool allVerticesNotFixed() const noexcept = 0;
    virtual const double* measurementData() const noexcept = 0;
    virtual int dim() const noexcept = 0;
    virtual void setInformation(Information info) noexcept = 0;
    virtual Information getInformation() noexcept = 0;
    virtual void setCamera(const Camera& c) noexcept = 0;
    virtual Camera& getCamera() noexcept = 0;
    virtual void inactivate() noexcept = 0;
    virtual void setActive() noexcept = 0;
    virtual bool isActive() const noexcept = 0;
};

template <int DIM>
class Edge : public BaseEdge {
public:
    using Measurement = Vec<double, DIM>;
    BaseVertex* getVertex(int i) override { return v_[i]; }
    void setVertex(BaseVertex* v, int i) override { v_[i] = v; v->addEdge(this); }
    bool allVerticesFixed() const noexcept override { return v_[0]->isFixed() && v_[1]->isFixed(); }
    bool anyVerticesNotFixed() const noexcept override { return !allVerticesFixed(); }
    bool allVerticesNotFixed() const noexcept override { return !v_[0]->isFixed() && !v_[1]->isFixed(); }
    const double* measurementData() const noexcept override { return m_.data; }
    int dim() const noexcept override { return DIM; }
    void setMeasurement(const Measurement& m) noexcept { m_ = m; }
    const Measurement& getMeasurement() const noexcept { return m_; }
    void setInformation(Information info) noexcept override { info_ = info; }
    Information getInformation() noexcept override { return info_; }
    void setCamera(const Camera& c) noexcept override { cam_ = c; }
    Camera& getCamera() noexcept override { return cam_; }
    void inactivate() noexcept override { active_ = false; }
    void setActive() noexcept override { active_ = true; }
    bool isActive() const noexcept override { return active_; }
private:
    BaseVertex* v_[2] = {nullptr, nullptr};
    Measurement m_;
    Information info_ = 1.0;
    Camera cam_;
    bool active_ = true;
};
using MonoEdge = Edge<2>;      // g2o EdgeSE3ProjectXYZ
using StereoEdge = Edge<3>;    // g2o EdgeStereoSE3ProjectXYZ

// ---- edge sets (src/optimisable_graph.h:510-680) ------------------------------------------------------
class BaseEdgeSet {
public:
    virtual ~BaseEdgeSet() {}
    virtual void addEdge(BaseEdge* e) = 0;
    virtual void removeEdge(BaseEdge* e) = 0;
    virtual size_t nedges() const noexcept = 0;
    virtual size_t nActiveEdges() const noexcept = 0;
    virtual int dim() const noexcept = 0;
    virtual void clearEdges() noexcept = 0;
    virtual void setRobustKernel(RobustKernelType type, Scalar delta) noexcept = 0;
    virtual void setInformation(Information info) noexcept = 0;
    virtual void setCamera(const Camera& c) noexcept = 0;
    virtual Camera& getCamera() noexcept = 0;
    // flat arrays handed to cugo_b200_add_edge_set (the reference's EdgeSet::init,
    // src/optimisable_graph.hpp:474-572): active edges = flagged active and not both vertices fixed
    struct Packed {
        int dim = 0, n = 0, nomega = 0, ncam = 0, rk = 0;
        double delta = 1.0;
        std::vector<int> iP, iL;
        std::vector<double> meas, omega, cam;
    };
    virtual void pack(const GraphOptimisationOptions& opt, Packed& out) = 0;
};

template <int DIM>
class EdgeSet : public BaseEdgeSet {
public:
    using EdgeType = Edge<DIM>;
    void addEdge(BaseEdge* e) override { if (index_.insert(e).second) edges_.push_back(e); }
    void removeEdge(BaseEdge* e) override {
        if (!index_.erase(e)) return;
        for (size_t i = 0; i < edges_.size(); i++)
            if (edges_[i] == e) { edges_.erase(edges_.begin() + i); break; }
        for (int k = 0; k < 2; k++) if (e->getVertex(k)) e->getVertex(k)->removeEdge(e);
    }
    size_t nedges() const noexcept override { return edges_.size(); }
    size_t nActiveEdges() const noexcept override { return nActive_; }
    int dim() const noexcept override { return DIM; }
    void clearEdges() noexcept override { edges_.clear(); index_.clear(); nActive_ = 0; }
    void setRobustKernel(RobustKernelType type, Scalar delta) noexcept override { rk_ = type; delta_ = delta; }
    RobustKernelType robustKernelType() const noexcept { return rk_; }
    Scalar robustKernelDelta() const noexcept { return delta_; }
    void setInformation(Information info) noexcept override { info_ = info; }
    void setCamera(const Camera& c) noexcept override { cam_ = c; }
    Camera& getCamera() noexcept override { return cam_; }
    const std::vector<BaseEdge*>& get() const noexcept { return edges_; }
    void pack(const GraphOptimisationOptions& opt, Packed& out) override {
        out = Packed();
        out.dim = DIM; out.rk = (int)rk_; out.delta = delta_;
        for (BaseEdge* e : edges_) {
            if (!e->isActive() || e->allVerticesFixed()) continue;
            out.iP.push_back(e->getVertex(0)->getIndex());
            out.iL.push_back(e->getVertex(1)->getIndex());
            const double* m = e->measurementData();
            out.meas.insert(out.meas.end(), m, m + DIM);
            if (opt.perEdgeInformation) out.omega.push_back(e->getInformation());
            if (opt.perEdgeCamera) {
                const Camera& c = e->getCamera();
                const double v[5] = {c.fx, c.fy, c.cx, c.cy, c.bf};
                out.cam.insert(out.cam.end(), v, v + 5);
            }
        }
        out.n = (int)out.iP.size();
        nActive_ = out.n;
        if (!opt.perEdgeInformation) out.omega.assign(1, info_);
        if (!opt.perEdgeCamera) out.cam = {cam_.fx, cam_.fy, cam_.cx, cam_.cy, cam_.bf};
        out.nomega = opt.perEdgeInformation ? out.n : 1;
        out.ncam = opt.perEdgeCamera ? out.n : 1;
        if (out.n == 0) { out.omega.assign(1, info_); out.cam.assign(5, 0.0); out.nomega = out.ncam = 1; }
    }
private:
    std::vector<BaseEdge*> edges_;            // insertion order (deterministic packing)
    std::unordered_set<BaseEdge*> index_;
    size_t nActive_ = 0;
    RobustKernelType rk_ = RobustKernelType::None;
    Scalar delta_ = 1.0;
    Information info_ = 1.0;
    Camera cam_;
};
using MonoEdgeSet = EdgeSet<2>;
using StereoEdgeSet = EdgeSet<3>;

using EdgeSetVec = std::vector<BaseEdgeSet*>;
using VertexSetVec = std::vector<BaseVertexSet*>;

// ---- the optimiser (include/cuda_graph_optimisation.h:132-239) ----------------------------------------
class CudaGraphOptimisationImpl;

class CudaGraphOptimisation {
public:
    using Ptr = std::unique_ptr<CudaGraphOptimisationImpl>;
    virtual ~CudaGraphOptimisation() {}
    static Ptr create();
    virtual void initialize() = 0;
    virtual void optimize(int niterations) = 0;
    virtual BatchStatistics& batchStatistics() = 0;
    virtual const TimeProfile& timeProfile() = 0;
    virtual EdgeSetVec& getEdgeSets() = 0;
    virtual size_t nVertices(int id) = 0;
    virtual void clearEdgeSets() = 0;
    virtual void clearVertexSets() = 0;
    virtual void setVerbose(bool status) = 0;
    virtual void setProfile(bool status) = 0;
};

class CudaGraphOptimisationImpl : public CudaGraphOptimisation {
public:
    // `device`: CUDA ordinal this optimiser is bound to (the reference picks "the fastest" GPU itself,
    // src/cuda_device.cpp:15-42).  The options object is held by reference, as in the reference.
    explicit CudaGraphOptimisationImpl(GraphOptimisationOptions& options, int device = 0) : options_(options) { open(device); }
    explicit CudaGraphOptimisationImpl(int device = 0) : options_(ownOptions_) { open(device); }
    ~CudaGraphOptimisationImpl() override { if (h_) cugo_b200_destroy(h_); }
    CudaGraphOptimisationImpl(const CudaGraphOptimisationImpl&) = delete;
    CudaGraphOptimisationImpl& operator=(const CudaGraphOptimisationImpl&) = delete;

    template <typename T>
    bool addEdgeSet(T* edgeSet) { edgeSets_.push_back(edgeSet); return true; }
    template <typename T>
    bool addVertexSet(T* vertexSet) { vertexSets_.push_back(vertexSet); return true; }

    EdgeSetVec& getEdgeSets() override { return edgeSets_; }
    size_t nVertices(int id) override { return vertexSets_.at(id)->size(); }
    void clearEdgeSets() override { edgeSets_.clear(); }
    void clearVertexSets() override { vertexSets_.clear(); }
    void setVerbose(bool s) override { verbose_ = s; }
    void setProfile(bool s) override { profile_ = s; }
    BatchStatistics& batchStatistics() override { return stats_; }
    const TimeProfile& timeProfile() override { return timeProfile_; }

    // BlockSolver::initialize (src/block_solver.cpp:21-137): vertex set 0 = poses, 1 = landmarks
    // (include/ba_types.h:46-49).  An unchanged graph keeps its structure and symbolic analysis and only
    // re-uploads estimates, measurements and informations (src/block_solver.cpp:161-170).
    void initialize() override {
        const auto t0 = std::chrono::steady_clock::now();
        if (vertexSets_.size() != 2 || vertexSets_[0]->estimateDim() != 7 || vertexSets_[1]->estimateDim() != 3)
            throw std::runtime_error("cugo: add the pose vertex set first, then the landmark vertex set");
        for (auto* vs : vertexSets_) vs->assignIndices();
        vertexSets_[0]->packEstimates(poses_);
        vertexSets_[1]->packEstimates(landmarks_);
        std::vector<BaseEdgeSet::Packed> packed(edgeSets_.size());
        for (size_t i = 0; i < edgeSets_.size(); i++) edgeSets_[i]->pack(options_, packed[i]);
        const int nPa = (int)vertexSets_[0]->size(), nP = vertexSets_[0]->getActiveSize();
        const int nLa = (int)vertexSets_[1]->size(), nL = vertexSets_[1]->getActiveSize();
        bool same = uploaded_ && nPa == nPa_ && nP == nP_ && nLa == nLa_ && nL == nL_ && packed.size() == last_.size();
        for (size_t i = 0; same && i < packed.size(); i++)
            same = packed[i].dim == last_[i].dim && packed[i].iP == last_[i].iP && packed[i].iL == last_[i].iL &&
                   packed[i].nomega == last_[i].nomega && packed[i].cam == last_[i].cam;
        if (same) {
            ck(cugo_b200_set_estimates(h_, poses_.data(), landmarks_.data()));
            for (size_t i = 0; i < packed.size(); i++) {
                if (packed[i].meas != last_[i].meas || packed[i].omega != last_[i].omega)
                    ck(cugo_b200_update_measurements(h_, (int)i, packed[i].meas.data(), packed[i].omega.data(), packed[i].nomega));
                if (packed[i].rk != last_[i].rk || packed[i].delta != last_[i].delta)
                    ck(cugo_b200_set_robust_kernel(h_, (int)i, packed[i].rk, packed[i].delta));
            }
        } else {
            ck(cugo_b200_set_vertices(h_, nPa, nP, poses_.data(), nLa, nL, landmarks_.data()));
            ck(cugo_b200_clear_edge_sets(h_));
            for (auto& p : packed) {
                int id = -1;
                ck(cugo_b200_add_edge_set(h_, p.dim, p.n, p.iP.data(), p.iL.data(), p.meas.data(), p.omega.data(), p.nomega,
                                          p.cam.data(), p.ncam, p.rk, p.delta, &id));
            }
            nPa_ = nPa; nP_ = nP; nLa_ = nLa; nL_ = nL;
        }
        last_.swap(packed);
        uploaded_ = true;
        stats_.clear();
        timeProfile_["0: Initialize (host ms)"] = ms(t0);
    }

    // CudaGraphOptimisationImpl::optimize (src/cuda_graph_optimisation.cpp:48-154): the whole LM loop
    // runs on the device; the estimates are written back into the vertices afterwards (finalize).
    void optimize(int niterations) override {
        if (!uploaded_) throw std::runtime_error("cugo: optimize() before initialize()");
        const auto t0 = std::chrono::steady_clock::now();
        ck(cugo_b200_build_structure(h_));      // no-op when the structure is current
        timeProfile_["1: Build Structure + Symbolic Decomposition (host ms)"] = ms(t0);
        const auto t1 = std::chrono::steady_clock::now();
        const int cap = niterations > 0 ? niterations : 1;
        std::vector<double> chi2(cap), lambda(cap);
        std::vector<int> trials(cap);
        int n = 0;
        ck(cugo_b200_optimize(h_, niterations, cap, chi2.data(), lambda.data(), trials.data(), &n));
        for (int i = 0; i < n && i < cap; i++) {
            stats_.addStat({i, chi2[i]});
            if (verbose_) std::printf("iter: %2d, chi2: %.1f, lambda: %.3e, rejected trials: %d\n", i + 1, chi2[i], lambda[i], trials[i]);
        }
        ck(cugo_b200_get_estimates(h_, poses_.data(), landmarks_.data()));
        vertexSets_[0]->unpackEstimates(poses_.data());
        vertexSets_[1]->unpackEstimates(landmarks_.data());
        double devMs = 0; long kernels = 0;
        ck(cugo_b200_get_last_timing(h_, &devMs, &kernels));
        timeProfile_["2: LM iterations (device ms)"] = devMs;
        timeProfile_["3: LM iterations + write-back (host ms)"] = ms(t1);
        timeProfile_["4: kernels launched"] = (double)kernels;
        if (profile_) for (auto& kv : timeProfile_) std::printf("%s:  %f\n", kv.first.c_str(), kv.second);
    }

    cugo_b200_solver* handle() noexcept { return h_; }   // the C ABI handle, for stage-level calls

private:
    void open(int device) {
#ifdef USE_FLOAT32            // the reference's compile-time switch (include/scalar.h): single-precision block storage
        const int rc = cugo_b200_create(device, 1, &h_);
#else
        const int rc = cugo_b200_create(device, 0, &h_);
#endif
        if (rc != CUGO_B200_OK)
            throw std::runtime_error(rc == CUGO_B200_ENODEVICE ? "cugo: no CUDA device (this engine has no CPU path)"
                                                               : "cugo: cannot create the solver");
    }
    void ck(int rc) { if (rc != CUGO_B200_OK) throw std::runtime_error(std::string("cugo: ") + cugo_b200_last_error(h_)); }
    static double ms(std::chrono::steady_clock::time_point t0) {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    GraphOptimisationOptions ownOptions_;
    GraphOptimisationOptions& options_;
    cugo_b200_solver* h_ = nullptr;
    EdgeSetVec edgeSets_;
    VertexSetVec vertexSets_;
    BatchStatistics stats_;
    TimeProfile timeProfile_;
    bool verbose_ = false, profile_ = false, uploaded_ = false;
    int nPa_ = 0, nP_ = 0, nLa_ = 0, nL_ = 0;
    std::vector<double> poses_, landmarks_;
    std::vector<BaseEdgeSet::Packed> last_;
};

inline CudaGraphOptimisation::Ptr CudaGraphOptimisation::create() { return Ptr(new CudaGraphOptimisationImpl()); }

}  // namespace cugo
