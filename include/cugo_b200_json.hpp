// cugo_b200_json.hpp -- dependency-free reader / writer of the reference's JSON graph files, plus the
// comparison report of its g2o sample.  Header-only, host-only, no CUDA.
//
// The reference parses its sample inputs with OpenCV's cv::FileStorage
// (samples/sample_ba_from_file/main.cpp:78-160).  OpenCV is not a dependency here: the same schema
//
//   fx fy cx cy bf                                               top-level numbers (one shared camera)
//   pose_vertices      [{id, fixed, q: [x,y,z,w], t: [x,y,z]}]
//   landmark_vertices  [{id, fixed, Xw: [x,y,z]}]
//   monocular_edges    [{vertexP, vertexL, measurement: [u,v],    information}]
//   stereo_edges       [{vertexP, vertexL, measurement: [u,v,ur], information}]
//
// is read by a small recursive-descent JSON parser straight into flat arrays (`GraphFile`), vertex ids kept as
// written (arbitrary integers; ORB-SLAM keyframe / map-point ids).  `buildGraph` turns a GraphFile into the
// vertex / edge objects of include/cugo_b200.hpp exactly as the reference's readGraph does, so a real
// ba_kitti_*.json can be dropped in.  `printComparison` is the chi2 table + rotation / translation / landmark RMSE
// of samples/sample_comparison_with_g2o/main.cpp:101-148.
#pragma once
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace cugo {
namespace io {

struct GraphFile {
    double cam[5] = {0, 0, 0, 0, 0};               // fx fy cx cy bf
    std::vector<int> poseId, poseFixed;            // file order
    std::vector<double> poseQT;                    // [P][7] qx qy qz qw tx ty tz
    std::vector<int> lmId, lmFixed;
    std::vector<double> lmXw;                      // [L][3]
    std::vector<int> monoP, monoL, stereoP, stereoL;   // vertex IDS (not positions)
    std::vector<double> monoMeas, monoInfo;        // [Em][2], [Em]
    std::vector<double> stereoMeas, stereoInfo;    // [Es][3], [Es]
    size_t nPoses() const { return poseId.size(); }
    size_t nLandmarks() const { return lmId.size(); }
};

namespace detail {

// Minimal JSON tokenizer over an in-memory buffer.  Only what the schema needs: objects, arrays, numbers, strings
// (keys), true/false/null.  Also skips cv::FileStorage's optional first line ("%YAML:1.0" never appears in .json
// output, but a UTF-8 byte-order mark may).
class Parser {
public:
    Parser(const char* b, size_t n) : p_(b), e_(b + n) {
        if (n >= 3 && (unsigned char)b[0] == 0xEF && (unsigned char)b[1] == 0xBB && (unsigned char)b[2] == 0xBF) p_ += 3;
    }
    void ws() { while (p_ < e_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\r' || *p_ == '\t')) p_++; }
    bool peek(char c) { ws(); return p_ < e_ && *p_ == c; }
    bool take(char c) { ws(); if (p_ < e_ && *p_ == c) { p_++; return true; } return false; }
    void expect(char c) { if (!take(c)) fail(std::string("expected '") + c + "'"); }
    std::string str() {
        expect('"');
        std::string s;
        while (p_ < e_ && *p_ != '"') {
            if (*p_ == '\\' && p_ + 1 < e_) p_++;
            s.push_back(*p_++);
        }
        expect('"');
        return s;
    }
    double num() {
        ws();
        if (p_ < e_ && (std::isalpha((unsigned char)*p_))) {      // true / false / null -> 1 / 0 / 0
            const char* q = p_;
            while (p_ < e_ && std::isalpha((unsigned char)*p_)) p_++;
            return (p_ - q == 4 && !std::strncmp(q, "true", 4)) ? 1.0 : 0.0;
        }
        char* end = nullptr;
        const double v = std::strtod(p_, &end);
        if (end == p_) fail("expected a number");
        p_ = end;
        return v;
    }
    // reads up to n numbers of an array (or a single number) into out, skipping the rest
    void numbers(double* out, int n) {
        if (!take('[')) { out[0] = num(); return; }
        int k = 0;
        if (take(']')) return;
        do { const double v = num(); if (k < n) out[k++] = v; } while (take(','));
        expect(']');
    }
    void skipValue() {
        ws();
        if (peek('{')) { expect('{'); if (take('}')) return; do { str(); expect(':'); skipValue(); } while (take(',')); expect('}'); }
        else if (peek('[')) { expect('['); if (take(']')) return; do { skipValue(); } while (take(',')); expect(']'); }
        else if (peek('"')) str();
        else num();
    }
    [[noreturn]] void fail(const std::string& what) {
        throw std::runtime_error("cugo json: " + what + " at byte " + std::to_string((long long)(p_ - (e_ - size()))));
    }
    size_t size() const { return size_; }
    void setSize(size_t n) { size_ = n; }
private:
    const char* p_;
    const char* e_;
    size_t size_ = 0;
};

}  // namespace detail

// Parses a graph file in the reference's JSON schema.  Throws std::runtime_error on malformed input.
inline GraphFile readJson(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cugo json: cannot open " + path);
    std::fseek(f, 0, SEEK_END);
    const long len = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<char> buf((size_t)len + 1, 0);
    if (len > 0 && std::fread(buf.data(), 1, (size_t)len, f) != (size_t)len) { std::fclose(f); throw std::runtime_error("cugo json: short read of " + path); }
    std::fclose(f);
    detail::Parser ps(buf.data(), (size_t)len);
    ps.setSize((size_t)len);
    GraphFile g;
    bool haveCam[5] = {false, false, false, false, false};
    static const char* camKeys[5] = {"fx", "fy", "cx", "cy", "bf"};
    ps.expect('{');
    if (!ps.take('}')) {
        do {
            const std::string key = ps.str();
            ps.expect(':');
            int camIdx = -1;
            for (int i = 0; i < 5; i++) if (key == camKeys[i]) camIdx = i;
            if (camIdx >= 0) { g.cam[camIdx] = ps.num(); haveCam[camIdx] = true; continue; }
            const bool isPose = key == "pose_vertices", isLm = key == "landmark_vertices";
            const bool isMono = key == "monocular_edges", isStereo = key == "stereo_edges";
            if (!isPose && !isLm && !isMono && !isStereo) { ps.skipValue(); continue; }
            ps.expect('[');
            if (ps.take(']')) continue;
            do {      // one record
                ps.expect('{');
                int id = 0, fixed = 0, vP = 0, vL = 0;
                double q[4] = {0, 0, 0, 1}, t[3] = {0, 0, 0}, xw[3] = {0, 0, 0}, m[3] = {0, 0, 0}, info = 1.0;
                if (!ps.take('}')) {
                    do {
                        const std::string k = ps.str();
                        ps.expect(':');
                        if (k == "id") id = (int)ps.num();
                        else if (k == "fixed") fixed = ps.num() != 0.0;
                        else if (k == "q") ps.numbers(q, 4);
                        else if (k == "t") ps.numbers(t, 3);
                        else if (k == "Xw") ps.numbers(xw, 3);
                        else if (k == "vertexP") vP = (int)ps.num();
                        else if (k == "vertexL") vL = (int)ps.num();
                        else if (k == "measurement") ps.numbers(m, 3);
                        else if (k == "information") info = ps.num();
                        else ps.skipValue();
                    } while (ps.take(','));
                    ps.expect('}');
                }
                if (isPose) {
                    g.poseId.push_back(id); g.poseFixed.push_back(fixed);
                    g.poseQT.insert(g.poseQT.end(), q, q + 4); g.poseQT.insert(g.poseQT.end(), t, t + 3);
                } else if (isLm) {
                    g.lmId.push_back(id); g.lmFixed.push_back(fixed); g.lmXw.insert(g.lmXw.end(), xw, xw + 3);
                } else if (isMono) {
                    g.monoP.push_back(vP); g.monoL.push_back(vL); g.monoMeas.insert(g.monoMeas.end(), m, m + 2); g.monoInfo.push_back(info);
                } else {
                    g.stereoP.push_back(vP); g.stereoL.push_back(vL); g.stereoMeas.insert(g.stereoMeas.end(), m, m + 3); g.stereoInfo.push_back(info);
                }
            } while (ps.take(','));
            ps.expect(']');
        } while (ps.take(','));
        ps.expect('}');
    }
    for (int i = 0; i < 4; i++) if (!haveCam[i]) throw std::runtime_error(std::string("cugo json: missing camera field ") + camKeys[i]);
    return g;
}

// Writes a GraphFile in the same schema (17 significant digits: a read-back is bit-exact).
inline void writeJson(const GraphFile& g, const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "w");
    if (!f) throw std::runtime_error("cugo json: cannot write " + path);
    static const char* camKeys[5] = {"fx", "fy", "cx", "cy", "bf"};
    std::fprintf(f, "{\n");
    for (int i = 0; i < 5; i++) std::fprintf(f, "\"%s\": %.17g,\n", camKeys[i], g.cam[i]);
    std::fprintf(f, "\"pose_vertices\": [\n");
    for (size_t i = 0; i < g.nPoses(); i++) {
        const double* v = &g.poseQT[i * 7];
        std::fprintf(f, "{\"id\": %d, \"fixed\": %d, \"q\": [%.17g, %.17g, %.17g, %.17g], \"t\": [%.17g, %.17g, %.17g]}%s\n", g.poseId[i],
                     g.poseFixed[i], v[0], v[1], v[2], v[3], v[4], v[5], v[6], i + 1 < g.nPoses() ? "," : "");
    }
    std::fprintf(f, "],\n\"landmark_vertices\": [\n");
    for (size_t i = 0; i < g.nLandmarks(); i++) {
        const double* v = &g.lmXw[i * 3];
        std::fprintf(f, "{\"id\": %d, \"fixed\": %d, \"Xw\": [%.17g, %.17g, %.17g]}%s\n", g.lmId[i], g.lmFixed[i], v[0], v[1], v[2],
                     i + 1 < g.nLandmarks() ? "," : "");
    }
    std::fprintf(f, "],\n\"monocular_edges\": [\n");
    for (size_t i = 0; i < g.monoP.size(); i++)
        std::fprintf(f, "{\"vertexP\": %d, \"vertexL\": %d, \"measurement\": [%.17g, %.17g], \"information\": %.17g}%s\n", g.monoP[i],
                     g.monoL[i], g.monoMeas[2 * i], g.monoMeas[2 * i + 1], g.monoInfo[i], i + 1 < g.monoP.size() ? "," : "");
    std::fprintf(f, "],\n\"stereo_edges\": [\n");
    for (size_t i = 0; i < g.stereoP.size(); i++)
        std::fprintf(f, "{\"vertexP\": %d, \"vertexL\": %d, \"measurement\": [%.17g, %.17g, %.17g], \"information\": %.17g}%s\n",
                     g.stereoP[i], g.stereoL[i], g.stereoMeas[3 * i], g.stereoMeas[3 * i + 1], g.stereoMeas[3 * i + 2], g.stereoInfo[i],
                     i + 1 < g.stereoP.size() ? "," : "");
    std::fprintf(f, "]\n}\n");
    std::fclose(f);
}

// Final estimates keyed by vertex id, and the reference comparison sample's report over two runs.
struct Estimates {
    std::vector<int> poseId, lmId;
    std::vector<double> poseQT, lmXw;      // [P][7], [L][3], same order as the id arrays
};

struct Rmse { double rotation = 0, translation = 0, landmark = 0; };

// RMSE over quaternion coefficients / translations / landmark positions (sample_comparison_with_g2o/main.cpp:124-148).
// Both estimate sets must list the same ids in the same order.
inline Rmse rmse(const Estimates& a, const Estimates& b) {
    if (a.poseId != b.poseId || a.lmId != b.lmId) throw std::runtime_error("cugo compare: the two runs list different vertices");
    Rmse r;
    double sR = 0, sT = 0, sP = 0;
    for (size_t i = 0; i < a.poseId.size(); i++) {
        for (int k = 0; k < 4; k++) { const double d = a.poseQT[i * 7 + k] - b.poseQT[i * 7 + k]; sR += d * d; }
        for (int k = 4; k < 7; k++) { const double d = a.poseQT[i * 7 + k] - b.poseQT[i * 7 + k]; sT += d * d; }
    }
    for (size_t i = 0; i < a.lmXw.size(); i++) { const double d = a.lmXw[i] - b.lmXw[i]; sP += d * d; }
    const double nP = a.poseId.empty() ? 1.0 : (double)a.poseId.size(), nL = a.lmId.empty() ? 1.0 : (double)a.lmId.size();
    r.rotation = std::sqrt(sR / nP); r.translation = std::sqrt(sT / nP); r.landmark = std::sqrt(sP / nL);
    return r;
}

inline void printComparison(const std::vector<double>& chi2A, const std::vector<double>& chi2B, const Estimates& a,
                            const Estimates& b, const char* nameA, const char* nameB, FILE* out = stdout) {
    std::fprintf(out, "=== Objective function value : \n%10s|%16s %-4s|%16s %-4s\n", "iteration", "chi2", nameA, "chi2", nameB);
    const size_t n = chi2A.size() > chi2B.size() ? chi2A.size() : chi2B.size();
    for (size_t i = 0; i < n; i++) {
        std::fprintf(out, "%10zu|", i + 1);
        if (i < chi2A.size()) std::fprintf(out, "%21.6f|", chi2A[i]); else std::fprintf(out, "%21s|", "N/A");
        if (i < chi2B.size()) std::fprintf(out, "%21.6f\n", chi2B[i]); else std::fprintf(out, "%21s\n", "N/A");
    }
    const Rmse r = rmse(a, b);
    std::fprintf(out, "\n=== RMSE between %s estimates and %s estimates : \nRotation    : %.2e\nTranslation : %.2e\nLandmark    : %.2e\n",
                 nameA, nameB, r.rotation, r.translation, r.landmark);
}

}  // namespace io
}  // namespace cugo
