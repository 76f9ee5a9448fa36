// serialize.hpp — host mirror of ilqr_planner::EigenSerialize and the std::vector converters of utils/utils.h, utils.cpp:20-80
// (record & replay helpers of the reference: CSV with 4 significant digits, a vector per line, matrices row by row, a list of
// matrices separated by a line of '=').  File I/O only: nothing here touches the device.
#pragma once
#include <array>
#include <fstream>
#include <string>
#include <vector>

#include "dense.hpp"

namespace ilqr_planner {
namespace EigenSerialize {
namespace detail_ser {
inline void put(std::ofstream& f, double v) {
    f.precision(4);  // Eigen::IOFormat(4, ...): StreamPrecision 4
    f << v;
}
inline void row_of(std::ofstream& f, const MatrixXd& A, long i) {
    for (long j = 0; j < A.cols(); j++) {
        if (j) f << ",";
        put(f, A(i, j));
    }
}
}  // namespace detail_ser

// utils.h:19-46: every line of the file is a vector of nb comma-separated values
template <size_t nb>
inline std::vector<VectorXd> load(const std::string& filename) {
    std::ifstream f(filename);
    std::vector<VectorXd> states;
    std::string line;
    while (std::getline(f, line)) {
        std::array<double, nb> vec{};
        size_t i = 0, at = 0;
        while (i < nb) {
            const size_t pos = line.find(',', at);
            vec[i++] = std::stod(line.substr(at, pos == std::string::npos ? std::string::npos : pos - at));
            if (pos == std::string::npos) break;
            at = pos + 1;
        }
        VectorXd st(static_cast<long>(nb));
        for (size_t j = 0; j < nb; j++) st(static_cast<long>(j)) = vec[j];
        states.push_back(st);
    }
    return states;
}
// utils.cpp:20-30: one vector per line
inline bool save(const std::vector<VectorXd>& states, std::string filename) {
    std::ofstream f(filename);
    for (auto const& st : states) {
        for (long i = 0; i < st.size(); i++) {
            if (i) f << ",";
            detail_ser::put(f, st(i));
        }
        f << "\n";
    }
    return true;
}
// utils.cpp:32-43: matrices row by row, a separator line after each
inline bool save(const std::vector<MatrixXd>& states, std::string filename) {
    std::ofstream f(filename);
    for (auto const& st : states) {
        for (long i = 0; i < st.rows(); i++) {
            detail_ser::row_of(f, st, i);
            if (i + 1 < st.rows()) f << "\n";
        }
        f << "\n";
        f << "=================================== \n";
    }
    return true;
}
// utils.cpp:45-52
inline bool save(const MatrixXd& A, std::string filename) {
    std::ofstream f(filename);
    for (long i = 0; i < A.rows(); i++) {
        detail_ser::row_of(f, A, i);
        if (i + 1 < A.rows()) f << "\n";
    }
    f << "\n";
    return true;
}
// utils.cpp:53-60
inline bool save(const VectorXd& b, std::string filename) {
    std::ofstream f(filename);
    for (long i = 0; i < b.size(); i++) {
        if (i) f << ",";
        detail_ser::put(f, b(i));
    }
    f << "\n";
    return true;
}
}  // namespace EigenSerialize

// utils.cpp:71-79 (column-major, as Eigen stores)
inline std::vector<double> eigenVectorToStdVector(const VectorXd& u) { return std::vector<double>(u.data(), u.data() + u.rows()); }
inline std::vector<double> eigenMatrixToStdVector(const MatrixXd& u) { return std::vector<double>(u.data(), u.data() + u.rows() * u.cols()); }

}  // namespace ilqr_planner
