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
//   include/ba_types.h:171-234, include/icp_types.h       DepthEdgeSet; PlaneEdgeSet / LineEdgeSet (pose-only ICP edges)
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
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
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
// Host packing walks hundreds of thousands of heap objects (one or two cache misses each): a few threads overlap those
// misses.  fn(chunk, begin, end) is called once per chunk; chunks are contiguous and ascending.
inline int packThreads(size_t n) {
    const unsigned hc = std::thread::hardware_concurrency();
    const size_t t = n / 16384;
    return (int)(t < 1 ? 1 : (t > 8 ? 8 : t)) > (int)(hc ? hc : 1) ? (int)(hc ? hc : 1) : (int)(t < 1 ? 1 : (t > 8 ? 8 : t));
}
template <typename F>
inline void parallelChunks(size_t n, int nthreads, F&& fn) {
    if (nthreads <= 1) { fn(0, (size_t)0, n); return; }
    std::vector<std::thread> th;
    th.reserve(nthreads);
    for (int c = 0; c < nthreads; c++) th.emplace_back([&, c]() { fn(c, n * c / nthreads, n * (c + 1) / nthreads); });
    for (auto& t : th) t.join();
}
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
        // the order of the last call stands while no vertex was added / removed and no fixed flag moved a vertex across
        // the free | fixed boundary (checked on the contiguous pointer array, not by walking the map again)
        if (!dirty_ && order_.size() == vertices_.size()) {
            bool ok = true;
            for (size_t i = 0; i < order_.size() && ok; i++) ok = order_[i]->T::isFixed() == (i >= (size_t)active_);
            if (ok) return;
        }
        order_.clear();
        order_.reserve(vertices_.size());
        for (auto& kv : vertices_) if (!kv.second->isFixed()) order_.push_back(kv.second);
        active_ = (int)order_.size();
        for (auto& kv : vertices_) if (kv.second->isFixed()) order_.push_back(kv.second);
        for (size_t i = 0; i < order_.size(); i++) order_[i]->setIndex((int)i);
        dirty_ = false;
    }
    void packEstimates(std::vector<double>& out) const override {
        const int dim = estimateDim();
        out.resize(order_.size() * dim);
        double* o = out.data();
        detail::parallelChunks(order_.size(), detail::packThreads(order_.size()), [&](int, size_t b, size_t e) {
            for (size_t i = b; i < e; i++) detail::packOne(order_[i]->T::getEstimate(), o + i * dim);
        });
    }
    void unpackEstimates(const double* in) override {
        const int dim = estimateDim();
        detail::parallelChunks(order_.size(), detail::packThreads(order_.size()), [&](int, size_t b, size_t e) {
            for (size_t i = b; i < e; i++) detail::unpackOne(order_[i]->T::getEstimate(), in + i * dim);
        });
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
using DepthEdge = Edge<3>;     // include/ba_types.h:258-264: measurement (u, v, 1/z)

// ---- pose-only ICP edges (include/icp_types.h:110-125, src/measurements.h:13-101): one vertex, one scalar residual.
// The match structs keep the reference's member order: they go to the engine as they lie in memory.
struct PointToLineMatch {
    Vec3d a, b;                 // two points of the line
    double length = 0;          // |a - b|
    Vec3d pointP;               // the matched point, platform coordinates
    PointToLineMatch() {}
    PointToLineMatch(const Vec3d& start, const Vec3d& finish, const Vec3d& p = Vec3d()) : a(start), b(finish), pointP(p) {
        double s = 0;
        for (int i = 0; i < 3; i++) s += (a[i] - b[i]) * (a[i] - b[i]);
        length = std::sqrt(s);
    }
};
struct PointToPlaneMatch {
    Vec3d normal;               // unit normal
    double originDistance = 0;  // signed distance of the plane from the origin
    Vec3d pointP;
    PointToPlaneMatch() {}
    PointToPlaneMatch(const Vec3d& n, double offset, const Vec3d& p) : normal(n), originDistance(offset), pointP(p) {}
};
static_assert(sizeof(PointToLineMatch) == 10 * sizeof(double) && sizeof(PointToPlaneMatch) == 7 * sizeof(double),
              "match structs are plain runs of doubles");

template <typename M>
class UnaryEdge : public BaseEdge {
public:
    using Measurement = M;
    BaseVertex* getVertex(int i) override { return i == 0 ? v_ : nullptr; }
    void setVertex(BaseVertex* v, int i) override {
        if (i != 0) throw std::runtime_error("cugo: a pose-only edge has one vertex");
        v_ = v; v->addEdge(this);
    }
    bool allVerticesFixed() const noexcept override { return v_->isFixed(); }
    bool anyVerticesNotFixed() const noexcept override { return !v_->isFixed(); }
    bool allVerticesNotFixed() const noexcept override { return !v_->isFixed(); }
    const double* measurementData() const noexcept override { return reinterpret_cast<const double*>(&m_); }
    int dim() const noexcept override { return 1; }
    void setMeasurement(const Measurement& m) noexcept { m_ = m; }
    const Measurement& getMeasurement() const noexcept { return m_; }
    void setInformation(Information info) noexcept override { info_ = info; }
    Information getInformation() noexcept override { return info_; }
    void setCamera(const Camera&) noexcept override {}
    Camera& getCamera() noexcept override { return cam_; }
    void inactivate() noexcept override { active_ = false; }
    void setActive() noexcept override { active_ = true; }
    bool isActive() const noexcept override { return active_; }
private:
    BaseVertex* v_ = nullptr;
    Measurement m_;
    Information info_ = 1.0;
    Camera cam_;
    bool active_ = true;
};
using PlaneEdge = UnaryEdge<PointToPlaneMatch>;
using LineEdge = UnaryEdge<PointToLineMatch>;

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
    // outlier gating (src/optimisable_graph.h:631-633, .hpp:415-426, 603-640, 664-675)
    virtual void setOutlierThreshold(Scalar errorThreshold) noexcept = 0;
    virtual Scalar getOutlierThreshold() const noexcept = 0;
    virtual uint32_t getOutlierCount() const noexcept = 0;
    virtual uint32_t getInlierCount() const noexcept = 0;
    // mask[nActiveEdges] in packing order from the engine: flagged edges are inactivated (EdgeSet::updateEdges)
    virtual void applyOutlierMask(const int* mask, size_t n) = 0;
    // flat arrays handed to cugo_b200_add_edge_set (the reference's EdgeSet::init,
    // src/optimisable_graph.hpp:474-572): active edges = flagged active and not both vertices fixed
    struct Packed {
        int dim = 0, n = 0, nomega = 0, ncam = 0, rk = 0;     // dim: the engine's edge kind (2 mono, 3 stereo, 4 depth, 5 plane, 6 line)
        double delta = 1.0;
        std::vector<int> iP, iL;
        std::vector<double> meas, omega, cam;
    };
    virtual void pack(const GraphOptimisationOptions& opt, Packed& out) = 0;
};

// KIND: what the engine is told (cugo_b200.h): DIM for mono / stereo, 4 depth, 5 plane, 6 line (the last two pose-only)
template <int DIM, int KIND = DIM>
class EdgeSet : public BaseEdgeSet {
public:
    static constexpr bool UNARY = KIND >= 5;
    static constexpr int COLS = KIND == 5 ? 7 : (KIND == 6 ? 10 : DIM);     // doubles per measurement
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
    void clearEdges() noexcept override { edges_.clear(); packed_.clear(); index_.clear(); nActive_ = 0; outCount_ = 0; }
    void setRobustKernel(RobustKernelType type, Scalar delta) noexcept override { rk_ = type; delta_ = delta; }
    RobustKernelType robustKernelType() const noexcept { return rk_; }
    Scalar robustKernelDelta() const noexcept { return delta_; }
    void setInformation(Information info) noexcept override { info_ = info; }
    void setCamera(const Camera& c) noexcept override { cam_ = c; }
    Camera& getCamera() noexcept override { return cam_; }
    const std::vector<BaseEdge*>& get() const noexcept { return edges_; }
    void setOutlierThreshold(Scalar t) noexcept override { outThr_ = t; }
    Scalar getOutlierThreshold() const noexcept override { return outThr_; }
    uint32_t getOutlierCount() const noexcept override { return outCount_; }
    uint32_t getInlierCount() const noexcept override { return (uint32_t)nActive_ - outCount_; }
    void applyOutlierMask(const int* mask, size_t n) override {
        uint32_t c = 0;
        for (size_t i = 0; i < n && i < packed_.size(); i++)
            if (mask[i]) { packed_[i]->inactivate(); c++; }
        outCount_ = c;
    }
    void pack(const GraphOptimisationOptions& opt, Packed& out) override {
        out = Packed();
        out.dim = KIND; out.rk = (int)rk_; out.delta = delta_;
        outCount_ = 0;
        // pass 1: which edges go to the engine (active flag set, not both vertices fixed), per chunk; pass 2: fill
        const size_t ne = edges_.size();
        const int nth = detail::packThreads(ne);
        std::vector<size_t> cnt(nth + 1, 0);
        std::vector<unsigned char> keep(ne);
        detail::parallelChunks(ne, nth, [&](int c, size_t b, size_t e) {
            size_t k = 0;
            for (size_t i = b; i < e; i++) {
                BaseEdge* ed = edges_[i];
                const bool on = ed->isActive() && !ed->allVerticesFixed();
                keep[i] = on;
                k += on;
            }
            cnt[c + 1] = k;
        });
        for (int c = 0; c < nth; c++) cnt[c + 1] += cnt[c];
        const size_t n = cnt[nth];
        packed_.resize(n);
        out.iP.resize(n); out.meas.resize(n * COLS);
        if (!UNARY) out.iL.resize(n);
        const bool perCam = opt.perEdgeCamera && !UNARY;
        if (opt.perEdgeInformation) out.omega.resize(n);
        if (perCam) out.cam.resize(n * 5);
        detail::parallelChunks(ne, nth, [&](int c, size_t b, size_t e) {
            size_t o = cnt[c];
            for (size_t i = b; i < e; i++) {
                if (!keep[i]) continue;
                BaseEdge* ed = edges_[i];
                packed_[o] = ed;
                out.iP[o] = ed->getVertex(0)->getIndex();
                if (!UNARY) out.iL[o] = ed->getVertex(1)->getIndex();
                const double* m = ed->measurementData();
                for (int k = 0; k < COLS; k++) out.meas[o * COLS + k] = m[k];
                if (opt.perEdgeInformation) out.omega[o] = ed->getInformation();
                if (perCam) {
                    const Camera& cm = ed->getCamera();
                    double* v = &out.cam[o * 5];
                    v[0] = cm.fx; v[1] = cm.fy; v[2] = cm.cx; v[3] = cm.cy; v[4] = cm.bf;
                }
                o++;
            }
        });
        out.n = (int)n;
        nActive_ = n;
        if (!opt.perEdgeInformation) out.omega.assign(1, info_);
        if (!perCam) out.cam = {cam_.fx, cam_.fy, cam_.cx, cam_.cy, cam_.bf};
        out.nomega = opt.perEdgeInformation ? out.n : 1;
        out.ncam = perCam ? out.n : 1;
        if (out.n == 0) { out.omega.assign(1, info_); out.cam.assign(5, 0.0); out.nomega = out.ncam = 1; }
    }
private:
    std::vector<BaseEdge*> edges_;            // insertion order (deterministic packing)
    std::vector<BaseEdge*> packed_;           // the edges handed to the engine, in packing order
    std::unordered_set<BaseEdge*> index_;
    size_t nActive_ = 0;
    Scalar outThr_ = 0.0;
    uint32_t outCount_ = 0;
    RobustKernelType rk_ = RobustKernelType::None;
    Scalar delta_ = 1.0;
    Information info_ = 1.0;
    Camera cam_;
};
using MonoEdgeSet = EdgeSet<2>;
using StereoEdgeSet = EdgeSet<3>;
using DepthEdgeSet = EdgeSet<3, 4>;     // include/ba_types.h:171-234
using PlaneEdgeSet = EdgeSet<1, 5>;     // include/icp_types.h:64-107
using LineEdgeSet = EdgeSet<1, 6>;      // include/icp_types.h:19-61

using EdgeSetVec = std::vector<BaseEdgeSet*>;
using VertexSetVec = std::vector<BaseVertexSet*>;

// ---- BlockSolver (src/block_solver.h:24-250): one method per stage, same names and argument meaning --------
// The reference's BlockSolver owns the device system and is driven stage by stage by the LM loop of
// CudaGraphOptimisationImpl::optimize; this one forwards every stage to the C ABI.  `optimizeOnDevice` is the one
// addition: the whole LM loop of src/cuda_graph_optimisation.cpp:48-154 run on the device in one call.
class BlockSolver {
public:
    enum ProfileItem {
        PROF_ITEM_INITIALIZE, PROF_ITEM_BUILD_STRUCTURE, PROF_ITEM_COMPUTE_ERROR, PROF_ITEM_BUILD_SYSTEM,
        PROF_ITEM_SCHUR_COMPLEMENT, PROF_ITEM_DECOMP_SYMBOLIC, PROF_ITEM_DECOMP_NUMERICAL, PROF_ITEM_UPDATE,
        PROF_ITEM_SOLVE_HPP, PROF_ITEM_COUNT
    };
    BlockSolver() = delete;
    // `device`: CUDA ordinal (the reference passes its CudaDevice, which picks "the fastest" GPU: src/cuda_device.cpp:15-42)
    explicit BlockSolver(GraphOptimisationOptions& options, int device = 0) : options(options) {
#ifdef USE_FLOAT32            // the reference's compile-time switch (include/scalar.h): single-precision block storage
        const int rc = cugo_b200_create(device, 1, &h_);
#else
        const int rc = cugo_b200_create(device, 0, &h_);
#endif
        if (rc != CUGO_B200_OK)
            throw std::runtime_error(rc == CUGO_B200_ENODEVICE ? "cugo: no CUDA device (this engine has no CPU path)"
                                                               : "cugo: cannot create the solver");
        for (double& v : profItems_) v = 0.0;
    }
    ~BlockSolver() { if (h_) cugo_b200_destroy(h_); }
    BlockSolver(const BlockSolver&) = delete;
    BlockSolver& operator=(const BlockSolver&) = delete;

    // src/block_solver.cpp:21-137: vertex set 0 = poses, 1 = landmarks (include/ba_types.h:46-49).  An unchanged
    // graph keeps its structure and symbolic analysis and only re-uploads estimates, measurements and informations
    // (src/block_solver.cpp:161-170).
    void initialize(const EdgeSetVec& edgeSets, const VertexSetVec& vertexSets) {
        const auto t0 = now();
        // (a pose-only graph -- ICP edge sets alone -- may come with the pose vertex set only, include/icp_types.h:28)
        if (vertexSets.empty() || vertexSets.size() > 2 || vertexSets[0]->estimateDim() != 7 ||
            (vertexSets.size() == 2 && vertexSets[1]->estimateDim() != 3))
            throw std::runtime_error("cugo: add the pose vertex set first, then the landmark vertex set");
        const bool haveL = vertexSets.size() == 2;
        for (auto* vs : vertexSets) vs->assignIndices();
        vertexSets[0]->packEstimates(poses_);
        if (haveL) vertexSets[1]->packEstimates(landmarks_); else landmarks_.clear();
        std::vector<BaseEdgeSet::Packed> packed(edgeSets.size());
        nedges_ = 0;
        for (size_t i = 0; i < edgeSets.size(); i++) { edgeSets[i]->pack(options, packed[i]); nedges_ += packed[i].n; }
        const int nPa = (int)vertexSets[0]->size(), nP = vertexSets[0]->getActiveSize();
        const int nLa = haveL ? (int)vertexSets[1]->size() : 0, nL = haveL ? vertexSets[1]->getActiveSize() : 0;
        bool same = uploaded_ && nPa == nPa_ && nP == nP_ && nLa == nLa_ && nL == nL_ && packed.size() == last_.size();
        for (size_t i = 0; same && i < packed.size(); i++) {
            same = packed[i].dim == last_[i].dim && packed[i].iP == last_[i].iP && packed[i].iL == last_[i].iL &&
                   packed[i].nomega == last_[i].nomega && packed[i].cam == last_[i].cam;
            // the engine refreshes measurements in place for the BA sets only: changed ICP matches go up as a new graph
            if (same && packed[i].dim >= 5) same = packed[i].meas == last_[i].meas && packed[i].omega == last_[i].omega;
        }
        if (same) {
            ck(cugo_b200_set_estimates(h_, poses_.data(), landmarks_.data()));
            for (size_t i = 0; i < packed.size(); i++) {
                if (packed[i].meas != last_[i].meas || packed[i].omega != last_[i].omega)
                    ck(cugo_b200_update_measurements(h_, ids_[i], packed[i].meas.data(), packed[i].omega.data(), packed[i].nomega));
                if (packed[i].rk != last_[i].rk || packed[i].delta != last_[i].delta)
                    ck(cugo_b200_set_robust_kernel(h_, ids_[i], packed[i].rk, packed[i].delta));
            }
        } else {
            ck(cugo_b200_set_vertices(h_, nPa, nP, poses_.data(), nLa, nL, landmarks_.data()));
            ck(cugo_b200_clear_edge_sets(h_));
            ids_.assign(packed.size(), -1);
            for (size_t i = 0; i < packed.size(); i++) {
                auto& p = packed[i];
                if (p.dim == 5)
                    ck(cugo_b200_add_plane_edge_set(h_, p.n, p.iP.data(), p.meas.data(), p.omega.data(), p.nomega, p.rk, p.delta, &ids_[i]));
                else if (p.dim == 6)
                    ck(cugo_b200_add_line_edge_set(h_, p.n, p.iP.data(), p.meas.data(), p.omega.data(), p.nomega, p.rk, p.delta, &ids_[i]));
                else
                    ck(cugo_b200_add_edge_set(h_, p.dim, p.n, p.iP.data(), p.iL.data(), p.meas.data(), p.omega.data(), p.nomega,
                                              p.cam.data(), p.ncam, p.rk, p.delta, &ids_[i]));
            }
            nPa_ = nPa; nP_ = nP; nLa_ = nLa; nL_ = nL;
        }
        for (size_t i = 0; i < edgeSets.size(); i++) {
            const bool icp = packed[i].dim >= 5;
            if (icp && edgeSets[i]->getOutlierThreshold() > 0.0)
                throw std::runtime_error("cugo: outlier gating is not available for the pose-only ICP edge sets");
            if (!icp) ck(cugo_b200_set_outlier_threshold(h_, ids_[i], edgeSets[i]->getOutlierThreshold()));
        }
        last_.swap(packed);
        uploaded_ = true;
        profItems_[PROF_ITEM_INITIALIZE] = ms(t0);
    }
    void buildStructure(const EdgeSetVec&, const VertexSetVec&) {      // src/block_solver.cpp:139-248 (no-op when current)
        need();
        const auto t0 = now();
        ck(cugo_b200_build_structure(h_));
        profItems_[PROF_ITEM_BUILD_STRUCTURE] = ms(t0);
    }
    double computeErrors(const EdgeSetVec&, const VertexSetVec&) {     // :250-272
        need();
        const auto t0 = now();
        double chi2 = 0;
        ck(cugo_b200_compute_errors(h_, &chi2));
        profItems_[PROF_ITEM_COMPUTE_ERROR] = ms(t0);
        return chi2;
    }
    void buildSystem(const EdgeSetVec&, const VertexSetVec&) {         // :274-307
        need();
        const auto t0 = now();
        ck(cugo_b200_build_system(h_));
        profItems_[PROF_ITEM_BUILD_SYSTEM] = ms(t0);
    }
    double maxDiagonal() { double v = 0; ck(cugo_b200_max_diagonal(h_, &v)); return v; }     // :309-320
    void setLambda(double lambda) { ck(cugo_b200_set_lambda(h_, lambda)); }                   // :322-329
    void restoreDiagonal() { ck(cugo_b200_restore_diagonal(h_)); }                            // :331-338
    bool solve() {                                                                            // :340-388
        const auto t0 = now();
        int ok = 0;
        ck(cugo_b200_solve(h_, &ok));
        // Schur complement, numerical decomposition and back-substitution are one enqueue here
        profItems_[PROF_ITEM_DECOMP_NUMERICAL] = ms(t0);
        return ok != 0;
    }
    void update(const VertexSetVec&) {                                                        // :390-413
        const auto t0 = now();
        ck(cugo_b200_update(h_));
        profItems_[PROF_ITEM_UPDATE] = ms(t0);
    }
    // :423-429 -> EdgeSet::updateEdges (src/optimisable_graph.hpp:603-640).  flagOnDevice = false when the device
    // LM loop has already flagged (cugo_b200_optimize does it at its end, as the reference's optimize()).
    void updateEdges(const EdgeSetVec& edgeSets, bool flagOnDevice = true) {
        bool any = false;
        for (auto* es : edgeSets) any = any || es->getOutlierThreshold() > 0.0;
        if (!any) return;
        if (flagOnDevice) ck(cugo_b200_update_edges(h_));
        std::vector<int> mask;
        for (size_t i = 0; i < edgeSets.size(); i++) {
            if (!(edgeSets[i]->getOutlierThreshold() > 0.0) || edgeSets[i]->nActiveEdges() == 0) continue;
            mask.assign(edgeSets[i]->nActiveEdges(), 0);
            int no = 0, ni = 0;
            ck(cugo_b200_get_outliers(h_, ids_[i], mask.data(), &no, &ni));
            edgeSets[i]->applyOutlierMask(mask.data(), mask.size());
        }
    }
    double computeScale(double lambda) { double v = 0; ck(cugo_b200_compute_scale(h_, lambda, &v)); return v; }   // :415-421
    void push() { ck(cugo_b200_push(h_)); }                                                   // :431-434
    void pop() { ck(cugo_b200_pop(h_)); }                                                     // :436-439
    void clear() noexcept { uploaded_ = false; last_.clear(); nedges_ = 0; }                  // :460-468
    void finalize(const VertexSetVec& vertexSets) {                                           // :441-458
        need();
        ck(cugo_b200_get_estimates(h_, poses_.data(), landmarks_.data()));
        vertexSets[0]->unpackEstimates(poses_.data());
        if (vertexSets.size() > 1) vertexSets[1]->unpackEstimates(landmarks_.data());
    }
    // the reference's nine profile strings (src/block_solver.cpp:470-488); host wall-clock ms of the last call of each
    // stage (every staged call of the C ABI synchronises).  Items 4 and 5 are inside "1" / "6" here.
    void getTimeProfile(TimeProfile& prof) const {
        static const char* names[PROF_ITEM_COUNT] = {
            "0: Initialize Optimizer", "1: Build Structure", "2: Compute Error", "3: Build System", "4: Schur Complement",
            "5: Symbolic Decomposition", "6: Numerical Decomposition", "7: Update Solution", "8: Hpp linear solver"};
        prof.clear();
        for (int i = 0; i < PROF_ITEM_COUNT; i++) prof[names[i]] = profItems_[i];
    }
    int nedges() const { return nedges_; }
    static inline uint8_t makeEdgeFlag(bool fixedP, bool fixedL) { return (uint8_t)((fixedP ? 2 : 0) | (fixedL ? 1 : 0)); }

    // the LM loop of CudaGraphOptimisationImpl::optimize on the device: one entry per outer iteration
    int optimizeOnDevice(int niterations, std::vector<double>& chi2, std::vector<double>& lambda, std::vector<int>& trials) {
        need();
        const int cap = niterations > 0 ? niterations : 1;
        chi2.assign(cap, 0.0); lambda.assign(cap, 0.0); trials.assign(cap, 0);
        int n = 0;
        ck(cugo_b200_optimize(h_, niterations, cap, chi2.data(), lambda.data(), trials.data(), &n));
        n = n < cap ? n : cap;
        chi2.resize(n); lambda.resize(n); trials.resize(n);
        return n;
    }
    void lastTiming(double& deviceMs, long& kernels) { ck(cugo_b200_get_last_timing(h_, &deviceMs, &kernels)); }
    cugo_b200_solver* handle() noexcept { return h_; }

private:
    using Clock = std::chrono::steady_clock;
    static Clock::time_point now() { return Clock::now(); }
    static double ms(Clock::time_point t0) { return std::chrono::duration<double, std::milli>(Clock::now() - t0).count(); }
    void ck(int rc) { if (rc != CUGO_B200_OK) throw std::runtime_error(std::string("cugo: ") + cugo_b200_last_error(h_)); }
    void need() const { if (!uploaded_) throw std::runtime_error("cugo: BlockSolver used before initialize()"); }
    GraphOptimisationOptions& options;      // held by reference, as in the reference (src/block_solver.h:177)
    cugo_b200_solver* h_ = nullptr;
    bool uploaded_ = false;
    int nedges_ = 0;
    int nPa_ = 0, nP_ = 0, nLa_ = 0, nL_ = 0;
    std::vector<double> poses_, landmarks_;
    std::vector<BaseEdgeSet::Packed> last_;
    std::vector<int> ids_;                  // engine id of every edge set (the ICP sets have their own range)
    double profItems_[PROF_ITEM_COUNT];
};

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
    // `device`: CUDA ordinal this optimiser is bound to.  The options object is held by reference, as in the reference.
    explicit CudaGraphOptimisationImpl(GraphOptimisationOptions& options, int device = 0)
        : options_(options), solver_(new BlockSolver(options, device)) {}
    explicit CudaGraphOptimisationImpl(int device = 0) : options_(ownOptions_), solver_(new BlockSolver(ownOptions_, device)) {}
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

    void initialize() override {                                   // src/cuda_graph_optimisation.cpp:42-46
        solver_->initialize(edgeSets_, vertexSets_);
        stats_.clear();
        initialized_ = true;
    }

    // CudaGraphOptimisationImpl::optimize (src/cuda_graph_optimisation.cpp:48-154): buildStructure, the LM loop (on
    // the device here), updateEdges, finalize (estimates written back into the vertices).
    void optimize(int niterations) override {
        if (!initialized_) throw std::runtime_error("cugo: optimize() before initialize()");
        solver_->buildStructure(edgeSets_, vertexSets_);
        const auto t1 = std::chrono::steady_clock::now();
        std::vector<double> chi2, lambda;
        std::vector<int> trials;
        const int n = solver_->optimizeOnDevice(niterations, chi2, lambda, trials);
        for (int i = 0; i < n; i++) {
            stats_.addStat({i, chi2[i]});
            if (verbose_) std::printf("iter: %2d, chi2: %.1f, lambda: %.3e, rejected trials: %d\n", i + 1, chi2[i], lambda[i], trials[i]);
        }
        solver_->updateEdges(edgeSets_, false);     // flagged at the end of the device loop; here: mask -> edges
        solver_->finalize(vertexSets_);
        solver_->getTimeProfile(timeProfile_);
        double devMs = 0; long kernels = 0;
        solver_->lastTiming(devMs, kernels);
        timeProfile_["LM iterations (device ms)"] = devMs;
        timeProfile_["LM iterations + write-back (host ms)"] =
            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
        timeProfile_["kernels launched"] = (double)kernels;
        if (profile_) for (auto& kv : timeProfile_) std::printf("%s:  %f\n", kv.first.c_str(), kv.second);
    }

    BlockSolver& solver() noexcept { return *solver_; }             // stage-level access (the reference's solver_)
    cugo_b200_solver* handle() noexcept { return solver_->handle(); }   // the C ABI handle

private:
    GraphOptimisationOptions ownOptions_;
    GraphOptimisationOptions& options_;
    std::unique_ptr<BlockSolver> solver_;
    EdgeSetVec edgeSets_;
    VertexSetVec vertexSets_;
    BatchStatistics stats_;
    TimeProfile timeProfile_;
    bool verbose_ = false, profile_ = false, initialized_ = false;
};

inline CudaGraphOptimisation::Ptr CudaGraphOptimisation::create() { return Ptr(new CudaGraphOptimisationImpl()); }

}  // namespace cugo
