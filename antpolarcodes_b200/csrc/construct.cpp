#include "construct.h"
#include "reliability_5g.h"

#include <algorithm>
#include <cmath>
#include <numeric>

namespace polar {

bool construct_frozen_bits(int N, int K, float designSnr, const std::string& type, std::vector<uint32_t>& out,
                           std::string& err)
{
    if (N < 1 || (N & (N - 1))) { err = "new blockLength is not a power of 2!"; return false; }
    if (K < 0 || K > N) { err = "Invalid polar code(" + std::to_string(N) + ", " + std::to_string(K) + ")"; return false; }
    std::string t(type);
    std::transform(t.begin(), t.end(), t.begin(), [](unsigned char c) { return (char)std::tolower(c); });
    const int nFrozen = N - K;
    int lgN = 0;
    while ((1 << lgN) < N) ++lgN;
    out.clear();
    if (t.find("be") != std::string::npos) {
        // beta expansion, beta = 2^(1/4): betaexpansion.cpp:46-79 (weight_i = sum_j beta^j * bit_j(i))
        const double beta = std::pow(2.0, 1.0 / 4.0);
        std::vector<double> w(lgN), par(N);
        for (int j = 0; j < lgN; ++j) w[j] = std::pow(beta, (double)j);
        for (int i = 0; i < N; ++i) {
            double s = 0.0;
            for (int j = 0; j < lgN; ++j) s += w[j] * ((i >> j) & 1);
            par[i] = s;
        }
        std::vector<size_t> idx(N);
        std::iota(idx.begin(), idx.end(), 0);
        std::sort(idx.begin(), idx.end(), [&par](int l, int r) { return par[l] < par[r]; });
        for (int i = 0; i < nFrozen; ++i) out.push_back((uint32_t)idx[i]);
    } else if (t.find("5g") != std::string::npos) {
        // fiveGList.cpp:17-41: the first N-K table entries, whatever N is (throws for N > 1024)
        if (N > 1024) { err = "5G standard does not allow for block size N > 1024!"; return false; }
        for (int i = 0; i < nFrozen; ++i) out.push_back(RELIABILITY_5G[i]);
    } else {
        // Bhattacharyya parameters: bhattacharrya.cpp:40-90
        const float linear = (float)std::pow(10.0, (double)designSnr / 10.0);
        const float z0 = (float)std::exp(-2.0 * linear * (double)K / (double)N);
        std::vector<double> z(N, 0.0);
        z[0] = z0;
        for (int stage = lgN - 1; stage >= 0; --stage) {
            const int B = 1 << stage;
            for (int j = 0; j < N; j += (B << 1)) {
                const double T = z[j];
                z[j + B] = T * T;
                z[j] = 2 * T - z[j + B];
            }
        }
        // stable insertion sort, descending (arrayfuncs.cpp:90-106): worst channels first
        std::vector<int> perm(N);
        std::iota(perm.begin(), perm.end(), 0);
        std::stable_sort(perm.begin(), perm.end(), [&z](int l, int r) { return z[l] > z[r]; });
        for (int i = 0; i < nFrozen; ++i) out.push_back((uint32_t)perm[i]);
    }
    std::sort(out.begin(), out.end());
    return true;
}

} // namespace polar
