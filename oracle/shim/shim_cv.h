// TEST INFRASTRUCTURE ONLY (oracle).  Minimal cv::Mat with value semantics plus the five numeric entry
// points the reference calls; numerics delegate to oracle/thirdparty/tp_smallmat.h (see the provenance note
// there: OpenCV is not vendored, parity for these is unpinned and tolerance-level).
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>
#include "../thirdparty/tp_smallmat.h"

#define CV_8S 1
#define CV_32S 4
#define CV_32F 5

namespace cv {

struct Scalar {
    double v;
    Scalar(double a = 0) : v(a) {}
    static Scalar all(double a) { return Scalar(a); }
};

enum { DECOMP_LU = 0, DECOMP_QR = 4 };

class Mat {
public:
    int rows = 0, cols = 0, type_ = CV_32F;
    std::vector<unsigned char> buf;
    Mat() {}
    Mat(int r, int c, int type, const Scalar& s = Scalar(0)) : rows(r), cols(c), type_(type) {
        buf.resize(static_cast<size_t>(r) * c * esz());
        size_t n = static_cast<size_t>(r) * c;
        if (type == CV_32F) { float f = static_cast<float>(s.v); float* p = reinterpret_cast<float*>(buf.data()); for (size_t i = 0; i < n; ++i) p[i] = f; }
        else if (type == CV_32S) { int32_t f = static_cast<int32_t>(s.v); int32_t* p = reinterpret_cast<int32_t*>(buf.data()); for (size_t i = 0; i < n; ++i) p[i] = f; }
        else { std::memset(buf.data(), static_cast<int>(static_cast<int8_t>(s.v)), n); }
    }
    size_t esz() const { return type_ == CV_8S ? 1 : 4; }
    template <class T> T& at(int r, int c) { return reinterpret_cast<T*>(buf.data())[static_cast<size_t>(r) * cols + c]; }
    template <class T> const T& at(int r, int c) const { return reinterpret_cast<const T*>(buf.data())[static_cast<size_t>(r) * cols + c]; }
    float* f() { return reinterpret_cast<float*>(buf.data()); }
    const float* f() const { return reinterpret_cast<const float*>(buf.data()); }
    void copyTo(Mat& o) const { o = *this; }
    Mat inv(int = DECOMP_LU) const {
        Mat r(rows, cols, CV_32F);
        tp::inverse(f(), r.f(), rows);
        return r;
    }
};

inline Mat operator*(const Mat& a, const Mat& b) {
    Mat c(a.rows, b.cols, CV_32F);
    tp::matmul(a.f(), b.f(), c.f(), a.rows, a.cols, b.cols);
    return c;
}

inline void transpose(const Mat& a, Mat& t) {
    Mat r(a.cols, a.rows, CV_32F);
    for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) r.at<float>(j, i) = a.at<float>(i, j);
    t = r;
}

inline bool solve(const Mat& A, const Mat& B, Mat& X, int /*flags*/) {
    Mat r(A.cols, 1, CV_32F);
    bool ok = tp::solve_qr(A.f(), B.f(), r.f(), A.rows, A.cols);
    X = r;
    return ok;
}

// Eigenvalues descending; written into E keeping its element count (the reference pre-sizes E as 1 x n and reads
// at<float>(0,i)); eigenvectors as rows of V.
inline bool eigen(const Mat& A, Mat& E, Mat& V) {
    int n = A.rows;
    if (E.rows * E.cols != n) E = Mat(1, n, CV_32F);
    if (V.rows != n || V.cols != n) V = Mat(n, n, CV_32F);
    tp::eigen_sym(A.f(), E.f(), V.f(), n);
    return true;
}

}  // namespace cv
