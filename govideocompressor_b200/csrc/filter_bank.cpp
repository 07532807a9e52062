// filter_bank.cpp -- see filter_bank.h.  Follows libswscale/utils.c initFilter() (FFmpeg, LGPL; algorithm
// restated, not copied): 64-bit fixed-point kernel evaluation, near-zero tap trimming, edge folding,
// error-feedback normalisation to int16.
#include "filter_bank.h"

#include <algorithm>
#include <cstdlib>

namespace gvc {
namespace {

constexpr double kTrimFraction = 0.002;  // SWS_MAX_REDUCE_CUTOFF

inline int floor_log2(unsigned v) {
    int n = 0;
    while (v >>= 1) ++n;
    return n;
}

// Cubic / triangle kernel in 2.30 fixed point; d = |distance| already divided by the shrink ratio.
int64_t kernel_weight(int algo, int64_t d, int64_t unit) {
    if (algo == kBilinear) {
        const int64_t w = (int64_t(1) << 30) - d;
        return w <= 0 ? 0 : w * (unit >> 30);
    }
    if (d >= (int64_t(1) << 31)) return 0;
    const int64_t one24 = int64_t(1) << 24;
    const int64_t b = 0, c = int64_t(0.6 * double(one24));
    const int64_t d2 = (d * d) >> 30, d3 = (d2 * d) >> 30;
    int64_t       w;
    if (d < (int64_t(1) << 30))
        w = (12 * one24 - 9 * b - 6 * c) * d3 + (-18 * one24 + 12 * b + 6 * c) * d2 + (6 * one24 - 2 * b) * (int64_t(1) << 30);
    else
        w = (-b - 6 * c) * d3 + (6 * b + 30 * c) * d2 + (-12 * b - 48 * c) * d + (8 * b + 24 * c) * (int64_t(1) << 30);
    return w / ((int64_t(1) << 54) / unit);
}

}  // namespace

bool build_filter_bank(FilterBank &out, int src_len, int dst_len, int algo, int one, int phase128) {
    if (src_len < 1 || dst_len < 1 || (algo != kBilinear && algo != kBicubic)) return false;
    const int     step = int(((int64_t(src_len) << 16) + (dst_len >> 1)) / dst_len);  // 16.16 source step
    const int64_t unit = int64_t(1) << (54 - std::min(floor_log2(unsigned(src_len / dst_len)), 8));

    int                  n;
    std::vector<int64_t> w;
    std::vector<int32_t> pos(dst_len);

    if (std::abs(step - 0x10000) < 10) {  // axis not scaled: single unit tap
        n = 1;
        w.assign(dst_len, unit);
        for (int i = 0; i < dst_len; ++i) pos[i] = i;
    } else {
        const int support = algo == kBicubic ? 4 : 2;
        n = step <= 0x10000 ? 1 + support : 1 + (support * src_len + dst_len - 1) / dst_len;
        n = std::max(1, std::min(n, src_len - 2));
        w.resize(size_t(dst_len) * n);
        int64_t centre = ((phase128 * int64_t(step)) >> 7) - ((phase128 * int64_t(0x10000)) >> 7);
        for (int i = 0; i < dst_len; ++i, centre += 2 * int64_t(step)) {
            int first = int((centre - (n - 2) * (int64_t(1) << 16)) / (1 << 17));
            pos[i]    = first;
            for (int j = 0; j < n; ++j) {
                int64_t d = std::llabs(int64_t(first + j) * (1 << 17) - centre) << 13;
                if (step > 0x10000) d = d * dst_len / src_len;
                w[size_t(i) * n + j] = kernel_weight(algo, d, unit);
            }
        }
    }

    // Trim negligible taps: rotate them off the front (advancing pos) and count them off the back.
    int keep_max = 0;
    for (int i = dst_len - 1; i >= 0; --i) {
        int64_t *row = &w[size_t(i) * n];
        int64_t  acc = 0;
        for (int t = 0; t < n; ++t) {
            acc += std::llabs(row[0]);
            if (double(acc) > kTrimFraction * double(unit)) break;
            if (i < dst_len - 1 && pos[i] >= pos[i + 1]) break;
            std::rotate(row, row + 1, row + n);
            row[n - 1] = 0;
            ++pos[i];
        }
        int keep = n;
        acc      = 0;
        for (int j = n - 1; j > 0; --j) {
            acc += std::llabs(row[j]);
            if (double(acc) > kTrimFraction * double(unit)) break;
            --keep;
        }
        keep_max = std::max(keep_max, keep);
    }

    const int            m = keep_max;
    std::vector<int64_t> f(size_t(dst_len) * m);
    for (int i = 0; i < dst_len; ++i)
        for (int j = 0; j < m; ++j) f[size_t(i) * m + j] = w[size_t(i) * n + j];

    // Fold taps that would read outside [0, src_len) onto the border samples.
    for (int i = 0; i < dst_len; ++i) {
        int64_t *row = &f[size_t(i) * m];
        if (pos[i] < 0) {
            for (int j = 1; j < m; ++j) {
                const int to = std::max(j + pos[i], 0);
                row[to] += row[j];
                row[j] = 0;
            }
            pos[i] = 0;
        }
        if (pos[i] + m > src_len) {
            const int shift = pos[i] + std::min(m - src_len, 0);
            int64_t   spill = 0;
            for (int j = m - 1; j >= 0; --j)
                if (pos[i] + j >= src_len) { spill += row[j]; row[j] = 0; }
            for (int j = m - 1; j >= 0; --j) row[j] = j < shift ? 0 : row[j - shift];
            pos[i] -= shift;
            row[src_len - 1 - pos[i]] += spill;
        }
    }

    out.taps    = m;
    out.dst_len = dst_len;
    out.pos     = pos;
    out.coef.assign(size_t(dst_len) * m, 0);
    for (int i = 0; i < dst_len; ++i) {
        const int64_t *row = &f[size_t(i) * m];
        int64_t        sum = 0, carry = 0;
        for (int j = 0; j < m; ++j) sum += row[j];
        sum = (sum + one / 2) / one;
        if (sum == 0) sum = 1;
        for (int j = 0; j < m; ++j) {
            const int64_t v = row[j] + carry;
            const int64_t q = (v >= 0 ? v + (sum >> 1) : v - (sum >> 1)) / sum;
            out.coef[size_t(i) * m + j] = int16_t(q);
            carry = v - q * sum;
        }
    }
    return true;
}

}  // namespace gvc
